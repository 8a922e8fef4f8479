/* Binary interface between the runtime core (libtclb_core) and a generated model library
 * (libtclb_<model>_<precision>.so).  One model library = one model x one precision x one
 * option set, like one reference binary `CLB/<model>/main`; the metadata tables below carry
 * what the reference generates into `Model_m` (src/Lists.cpp.Rt:79-192) and the launch record
 * replaces its `__constant__ LatticeContainer` (src/LatticeContainer.h.Rt:38-146): it is passed
 * by value to every kernel as a __grid_constant__ parameter, so no symbol copy precedes a stage.
 */
#ifndef TCLB_MODEL_ABI_H
#define TCLB_MODEL_ABI_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TCLB_ABI_VERSION 11
#define TCLB_MAX_NEIGH 9            /* (dy,dz) in {-1,0,1}^2 ; x is never decomposed */
#define TCLB_MAX_RANKS 16           /* subdomains that exchange Globals through peer memory (one box) */
#define TCLB_GL_RING 8              /* depth of the Globals publication ring */
#define TCLB_GL_SLOTS 128           /* replicas of the local accumulators (spreads same-address atomics over L2 slices) */

enum { TCLB_BACKEND_CPU = 0, TCLB_BACKEND_CUDA = 1 };
enum { TCLB_GLOBALS_NONE = 0, TCLB_GLOBALS_INTEGRATE = 1, TCLB_GLOBALS_OBJECTIVE = 2 };
enum { TCLB_HALO_NONE = 0, TCLB_HALO_DIRECT = 1 };
enum { TCLB_OP_SUM = 0, TCLB_OP_MAX = 1 };

typedef struct { const char *name, *field, *group, *comment; int dx, dy, dz, field_index;
                 int parameter, average, adjoint, has_default; double default_value; } tclb_density_t;
typedef struct { const char *name, *group, *comment; int minx, maxx, miny, maxy, minz, maxz;
                 int parameter, average, adjoint; } tclb_field_t;
typedef struct { const char *name, *unit, *default_value, *comment; int zonal;
                 int derived_target; double (*derive)(double); int adjoint; } tclb_setting_t;
typedef struct { const char *name, *unit, *comment; int op, adjoint; double base; } tclb_global_t;
typedef struct { const char *name, *unit, *comment; int vector, adjoint; } tclb_quantity_t;
typedef struct { const char *name, *group; uint32_t value, mask; int shift, bits, save; } tclb_nodetype_t;
typedef struct { const char *name; uint32_t mask; int shift, bits, capacity, save; } tclb_nodegroup_t;
typedef struct { const char *name, *main; int particle;
                 const unsigned char *load_density, *save_field, *read_field; } tclb_stage_t;
typedef struct { const char *name; int n_stages; const int *stages; } tclb_action_t;
typedef struct { const char *name; int value; } tclb_option_t;
/* objective = expression of the Globals (reference AddObjective, src/conf.R:349-372; every Global is one by default):
 * value and gradient with respect to every Global at the given values (src/Lists.cpp.Rt:17-34 OF_<name>) */
typedef struct { const char *name; void (*fun)(const double *glob, double *obj, double *grad); } tclb_objective_t;

/* Synthetic-turbulence mode set (reference src/SyntheticTurbulence.h:20-110) */
typedef struct { double k[3]; double a[3]; double b[3]; double freq; } tclb_st_mode_t;

/* One stage / quantity launch.  Pointers are device pointers for the CUDA backend, host
 * pointers for the CPU backend. */
typedef struct {
    int backend;
    void *stream;                       /* cudaStream_t (CUDA backend) */
    int nx, ny, nz;                     /* local interior extent (nx already padded) */
    int gy, gz;                         /* ghost thickness in y and z */
    int off_x, off_y, off_z;            /* position of the local block in the global lattice */
    double pos_x, pos_y, pos_z;         /* coordinate of the centre of global node (0,0,0): origin + 0.5 (reference src/Lattice.cu.Rt:66-71) */
    int tot_nx, tot_ny, tot_nz;         /* global lattice extent */
    void *snap_in, *snap_out;           /* real_t[n_fields][nz+2gz][ny+2gy][nx] */
    const void *flags;                  /* flag_t[nz+2gz][ny+2gy][nx] */
    const void *cuts;                   /* uint16 cut distances [26][...] or NULL */
    double *globals;                    /* local accumulators [TCLB_GL_SLOTS][n_globals] (kept across the stages of an action) */
    /* fused Globals exchange: the last block of the last stage stores this rank's totals into slot
     * [gl_epoch % TCLB_GL_RING][gl_rank] of EVERY rank's ring (peer stores) followed by an epoch stamp,
     * then resets the accumulators.  Ring layout: double[TCLB_GL_RING][TCLB_MAX_RANKS][n_globals + 1]. */
    int gl_publish, gl_rank, gl_nranks;
    double *gl_ring[TCLB_MAX_RANKS];
    unsigned *gl_counter;
    void *gl_partial;                   /* real_t[n_globals][gl_partial_stride]: one partial per thread block of a stage launch */
    long long gl_partial_stride;        /* capacity in blocks (>= blocks of any launch) */
    unsigned gl_epoch;
    const double *settings;             /* host array [n_settings] */
    const void *zone_table;             /* real_t[zone_len][n_zonesettings][zone_stride] */
    int zone_len, zone_index, zone_stride;
    int iter, reset_iter;
    int globals_mode, halo_mode;
    int obj_enabled;                    /* 0: every <Global>InObj weight is zero in every zone, the kernels skip the Objective terms */
    void *halo_snap[TCLB_MAX_NEIGH];    /* snapshot that receives my border rows, index (dy+1)+3(dz+1) */
    int halo_ny[TCLB_MAX_NEIGH], halo_nz[TCLB_MAX_NEIGH];
    /* fused halo signalling (sequence numbers, see csrc/core/lattice.cpp) */
    unsigned *sig_count;                /* local arrival counters [TCLB_MAX_NEIGH] */
    unsigned *sig_remote[TCLB_MAX_NEIGH];/* flag word in the neighbour that consumes my rows */
    const unsigned *sig_wait[TCLB_MAX_NEIGH]; /* local flag words written by neighbours */
    unsigned sig_epoch;                 /* value to publish / await */
    int sig_enable;
    /* region (quantity / colour / scan kernels) */
    int rx, ry, rz, rnx, rny, rnz;
    void *out;                          /* quantity output */
    double scale;
    /* synthetic turbulence */
    const tclb_st_mode_t *st_modes; int st_n;
    /* block shape request (0 = model default) */
    int block_x, block_y;
    /* particles: rows of 16 real_t (r, pos, vel, angular vel, force, moment); optional uniform-grid container over them:
     * pgrid = int[pg_n[2]][pg_n[1]][pg_n[0]][pg_depth] particle indices (-1 = empty slot), cell of a position p is
     * floor((p - pg_origin) * pg_inv_delta) per axis; NULL = look at every particle.  pg_depth == -1: pgrid holds the pg_n[0]
     * nodes of a threaded box tree instead, 8 words each: float lo[3], hi[3]; int skip; int particle (-1 = inner node) */
    const void *particles; int n_particles;
    const int *pgrid; int pg_n[3]; int pg_depth; double pg_origin[3]; double pg_inv_delta;
    /* CUDA-graph launches: device int[2] = {iteration, epoch} at the start of the graph + this launch's constant offsets */
    const int *dyn; int dyn_iter_off; unsigned dyn_epoch_off;
} tclb_launch_t;

typedef struct {
    int abi_version;
    const char *name;                   /* e.g. "d2q9" or "d2q9_bc" */
    const char *base_name;
    int real_size, flag_size;
    int zone_shift, zone_max, sum_globals, in_obj_offset;
    int n_densities; const tclb_density_t *densities;
    int n_fields; const tclb_field_t *fields;
    int n_settings; const tclb_setting_t *settings;
    int n_zonesettings; const tclb_setting_t *zonesettings;
    int n_globals; const tclb_global_t *globals;
    int n_quantities; const tclb_quantity_t *quantities;
    int n_nodetypes; const tclb_nodetype_t *nodetypes;
    int n_nodegroups; const tclb_nodegroup_t *nodegroups;
    int n_stages; const tclb_stage_t *stages;
    int n_actions; const tclb_action_t *actions;
    int n_options; const tclb_option_t *options;
    int gy, gz;                          /* ghost thickness the kernels assume */
    int (*run_stage)(int stage, const tclb_launch_t *);
    int (*get_quantity)(int quantity, const tclb_launch_t *);
    int (*color)(const tclb_launch_t *);          /* out = uchar4[rny][rnx] of plane rz */
    int (*kernel_info)(int stage, int globals_mode, int *regs, int *smem, int *lmem, int *max_threads);
    const char *(*last_error)(void);
    /* storage of the lattice in memory (reference --with-storage): element size of a snapshot value, whether the
     * per-field shift is subtracted on store, the type name, and the shift of every field */
    int storage_size, storage_shifted;
    const char *storage_name;
    const double *field_shift;
    double part_margin;                  /* reach of a particle beyond its radius in the particle stages (model's PartMargin) */
    int n_objectives; const tclb_objective_t *objectives;
} tclb_model_t;

/* every model library exports exactly this symbol */
const tclb_model_t *tclb_get_model(void);

#ifdef __cplusplus
}
#endif
#endif
