/* C ABI of the runtime core (libtclb_core.so).  Used by the Python package (ctypes), by the
 * `tclb` command line driver and by foreign callers.  Every function returns 0 / a valid handle
 * on success; on failure it returns non-zero / NULL and tclb_last_error() describes why.
 *
 * Handles:  tclb_model  - a loaded model library          (reference: one CLB/<model>/ build)
 *           tclb_domain - all subdomains owned by this process + their halo wiring
 *                         (reference: Lattice + MPI topology, src/Lattice.cu.Rt, src/Solver.cpp.Rt:288-370)
 *           tclb_solver - XML case runner (units, geometry, handlers)  (reference: Solver + Handlers, src/main.cpp:173-425)
 */
#ifndef TCLB_B200_H
#define TCLB_B200_H
#include <stdint.h>
#include "tclb_model_abi.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct tclb_model_h tclb_model_h;
typedef struct tclb_domain_h tclb_domain_h;
typedef struct tclb_solver_h tclb_solver_h;

const char *tclb_last_error(void);
void tclb_log_init(int rank, int level);

/* ---- models (reference src/Lists.cpp.Rt metadata) ---- */
tclb_model_h *tclb_model_load(const char *path);
const tclb_model_t *tclb_model_desc(tclb_model_h *);

/* ---- domain / lattice (reference Lattice::*, src/Lattice.h.Rt:39-191) ---- */
tclb_domain_h *tclb_domain_create(tclb_model_h *, int backend, int nx, int ny, int nz, int nranks,
                                  const int *local_ranks, int n_local, const int *devices, int n_devices);
void tclb_domain_destroy(tclb_domain_h *);
int tclb_domain_decomposition(tclb_domain_h *, int *divy, int *divz, int *boxes6 /* [nranks][6] x0,y0,z0,nx,ny,nz */);
int tclb_domain_set_halo_mode(tclb_domain_h *, int direct);              /* 1 fused in-kernel (default), 0 copy/NCCL baseline */
int tclb_domain_set_setting(tclb_domain_h *, const char *name, double value);
double tclb_domain_get_setting(tclb_domain_h *, const char *name);
int tclb_domain_set_zone_setting(tclb_domain_h *, const char *name, int zone, double value);
int tclb_domain_set_zone_series(tclb_domain_h *, const char *name, int zone, const double *v, int n);
int tclb_domain_set_flags(tclb_domain_h *, int rank, const uint32_t *flags_local);   /* [nz][ny][nx] of that rank */
int tclb_domain_get_flags(tclb_domain_h *, int rank, uint32_t *flags_local);
int tclb_domain_set_cuts(tclb_domain_h *, int rank, const uint16_t *cuts26);
int tclb_domain_init(tclb_domain_h *);                                   /* run action Init */
int tclb_domain_iterate(tclb_domain_h *, int n, int globals_on_last, int globals_every);
int tclb_domain_run_action(tclb_domain_h *, const char *action, int globals_mode);
int tclb_domain_sync(tclb_domain_h *);
int tclb_domain_iter(tclb_domain_h *);
int tclb_domain_get_globals(tclb_domain_h *, double *out);              /* n_globals values, reduced over local ranks */
int tclb_domain_get_field(tclb_domain_h *, int rank, const char *field, void *dst_native);
int tclb_domain_set_field(tclb_domain_h *, int rank, const char *field, const void *src_native);
int tclb_domain_get_quantity(tclb_domain_h *, int rank, const char *quantity, const int *box6_local, void *dst_native, double scale);
int tclb_domain_color(tclb_domain_h *, int rank, int zplane, float scale, unsigned char *rgba);
int tclb_domain_reset_average(tclb_domain_h *);
int tclb_domain_set_block(tclb_domain_h *, int bx, int by);
/* iterate() runs launch-only stretches of the loop as a replayed CUDA graph (default on for CUDA); statistics */
int tclb_domain_use_graphs(tclb_domain_h *, int on);
long long tclb_domain_graph_iterations(tclb_domain_h *);
int tclb_domain_save_pri(tclb_domain_h *, int rank, const char *file);
/* elements of the 27 margin blocks of a rank's checkpoint image, in file order (reference: src/conf.R:897-999 MarginSizes) */
int tclb_domain_pri_blocks(tclb_domain_h *, int rank, long long *out27);
int tclb_domain_load_pri(tclb_domain_h *, int rank, const char *file);
/* raw device views for DLPack/torch coupling: base pointer of the CURRENT snapshot of a rank and its shape
 * [n_fields][nz+2gz][ny+2gy][nx]; `snap`<0 means current */
void *tclb_domain_snapshot_ptr(tclb_domain_h *, int rank, int snap, int64_t shape4[4], int *ghost_y, int *ghost_z, int *current);
void *tclb_domain_flags_ptr(tclb_domain_h *, int rank);
void *tclb_domain_stream(tclb_domain_h *, int rank);
/* multi-process wiring (one process per GPU): export my receive side, import a neighbour's */
int tclb_domain_export_ipc(tclb_domain_h *, int rank, void *handle64, int64_t offsets4[4] /* snap0, snap1, halo flags, globals ring */);
int tclb_domain_import_peer(tclb_domain_h *, int my_rank, int peer_rank, const void *handle64, const int64_t offsets4[4],
                            int peer_ny, int peer_nz);
int tclb_domain_finalize_wiring(tclb_domain_h *);
/* bytes moved by the last settings upload (host->device) and the last Globals read-back (device->host) of a rank */
int tclb_domain_io_bytes(tclb_domain_h *, int rank, int64_t *h2d_bytes, int64_t *d2h_bytes);
/* hooks: external halo exchange (HALO_NONE) and cross-process globals reduction */
typedef void (*tclb_exchange_fn)(void *user, int snap_index);
typedef void (*tclb_reduce_fn)(void *user, double *values, int count, int op);
int tclb_domain_set_hooks(tclb_domain_h *, tclb_exchange_fn ex, tclb_reduce_fn red, void *user);


/* ---- device scan (reference: host loop in src/Handlers/cbFailcheck.cpp:66-72) ----
 * result3_dev[0..2] (DEVICE memory) = number of NaNs, minimum and maximum of the non-NaN values of the device array
 * dev_ptr[0..n) of floats (real_size 4) or doubles (8).  Asynchronous on `stream` (cudaStream_t); the current device
 * must be the one that owns dev_ptr.  Uses a cudaMalloc'ed scratch buffer that grows with n (kept per device). */
int tclb_scan_array(const void *dev_ptr, long long n, int real_size, void *stream, double *result3_dev);
const char *tclb_scan_last_error(void);

/* ---- device scan without any allocation (no counterpart in the reference) ----
 * Same result as tclb_scan_array, for callers that must not allocate: code running between the begin and end of a
 * region where the CUDA allocator is off limits (a memory-pool or allocator that is being replaced or audited,
 * a real-time loop with a fixed memory budget, a plugin that the host application forbids from calling cudaMalloc).
 * The scratch the kernel needs is a fixed array inside the library's device image (393,472 bytes per device), so the
 * call never asks CUDA, the C++ runtime or torch for memory; the caller supplies the input, the 3-double result buffer
 * and, optionally, a buffer for the error text.
 *   - n must be 1 .. TCLB_SCAN_STATIC_MAX_ELEMS; anything longer returns -3 with a message naming the limit.
 *   - The first call on each device looks up the device address of the array (that device must be current, as for
 *     every call).  Make that first call on each device before any stream capture starts.
 *   - All calls on a device share the one array.  Calls are safe back to back on one stream, or on streams ordered by
 *     events; calls that can run concurrently on the same device (unordered streams, several host threads) race on the
 *     array and give wrong results.  The array keeps the previous call's partial results; the only part that has to
 *     start at zero (an arrival counter) is cleared by every call on `stream` before the kernel.
 * Returns 0 on success, a negative code otherwise (-1 bad arguments, -2 bad real_size, -3 too long, -4 no device,
 * -5 symbol lookup failed, -6 counter clear failed, -7 launch failed); if err != NULL the reason is written there
 * (at most err_cap bytes, always terminated). */
#define TCLB_SCAN_STATIC_MAX_ELEMS 67108864LL
int tclb_scan_array_static(const void *dev_ptr, long long n, int real_size, void *stream, double *result3_dev,
                           char *err, int err_cap);

/* ---- discrete adjoint (csrc/core/adjoint.h; reference: Tapenade adjoint + acUSAdjoint/acSAdjoint handlers) ----
 * Needs the adjoint companion of the model library (libtclb_<model>_<precision>_ad.so, built by
 * tclb_b200.build.build_adjoint_lib).  One subdomain, single-stage Iteration action.
 *   record(n):  n primal iterations, inputs kept on a tape, Objective summed into tclb_adjoint_objective
 *   reverse():  backward sweep over the tape; afterwards get_field(f) = dObjective / d(field f at the start);
 *               for a parameter field (constant in time) that is the gradient with respect to the parameter
 *   steady(n):  n adjoint iterations around the current state
 *   set_field:  seed the adjoint (e.g. the derivative of a terminal cost) before reverse()/steady() */
typedef struct tclb_adjoint_h tclb_adjoint_h;
tclb_adjoint_h *tclb_adjoint_create(tclb_domain_h *);
void tclb_adjoint_destroy(tclb_adjoint_h *);
int tclb_adjoint_record(tclb_adjoint_h *, int n);
int tclb_adjoint_reverse(tclb_adjoint_h *);
int tclb_adjoint_steady(tclb_adjoint_h *, int n);
int tclb_adjoint_zero(tclb_adjoint_h *);
/* bound the number of stored states of the tape (0 = every iteration): checkpoints + recomputation in reverse() */
int tclb_adjoint_set_tape_frames(tclb_adjoint_h *, int n);
int tclb_adjoint_tape_frames(tclb_adjoint_h *);       /* states currently held */
double tclb_adjoint_objective(tclb_adjoint_h *);
/* gradients with respect to non-zonal settings (comma separated names; call before record()); the value returned is
 * the total derivative, following the derived-settings chain (e.g. nu -> omega) */
int tclb_adjoint_want_setting(tclb_adjoint_h *, tclb_domain_h *, const char *names_csv);
int tclb_adjoint_setting_gradient(tclb_adjoint_h *, tclb_domain_h *, const char *name, double *out);
int tclb_adjoint_get_field(tclb_adjoint_h *, tclb_domain_h *, const char *field, void *dst_real);
int tclb_adjoint_set_field(tclb_adjoint_h *, tclb_domain_h *, const char *field, const void *src_real);

/* ---- units (reference src/unit.{h,cpp}) ---- */
typedef struct tclb_units_h tclb_units_h;
tclb_units_h *tclb_units_create(void);
void tclb_units_destroy(tclb_units_h *);
int tclb_units_set(tclb_units_h *, const char *name, const char *value, const char *gauge);
int tclb_units_make_gauge(tclb_units_h *);
int tclb_units_alt(tclb_units_h *, const char *expr, double *out);                  /* SI text -> lattice units */
int tclb_units_si(tclb_units_h *, const char *expr, double *out, double *exponents9); /* SI value + unit exponents */

/* ---- case runner (reference src/main.cpp + Handlers) ---- */
tclb_solver_h *tclb_solver_create(const char *model, const char *precision, int backend, int nranks, const int *local_ranks,
                                  int n_local, const int *devices, int n_devices, int direct_halo);
void tclb_solver_destroy(tclb_solver_h *);
int tclb_solver_load_file(tclb_solver_h *, const char *case_xml, const char *const *xpath_args, int n_args);
int tclb_solver_load_string(tclb_solver_h *, const char *xml_text, const char *case_name, const char *const *xpath_args, int n_args);
int tclb_solver_run(tclb_solver_h *);
/* <RunPython>/<RunR> elements call this hook (element name, code text, iteration); non-zero return aborts the run */
typedef int (*tclb_script_fn)(void *user, const char *element, const char *code, int iteration);
int tclb_solver_set_script_hook(tclb_solver_h *, tclb_script_fn fn, void *user);
tclb_domain_h *tclb_solver_domain(tclb_solver_h *);     /* borrowed handle, valid while the solver lives */
int tclb_solver_iter(tclb_solver_h *);
int tclb_solver_zone_id(tclb_solver_h *, const char *zone_name);     /* setting-zone id of a named zone, -1 if unknown */
const char *tclb_solver_outpath(tclb_solver_h *);

#ifdef __cplusplus
}
#endif
#endif
