// extern "C" surface of the runtime core (see include/tclb_b200.h)
#include "tclb_b200.h"
#include "common.h"
#include "lattice.h"
#include "adjoint.h"
#include <cmath>
#include <cstring>
#include <cuda_runtime.h>

using namespace tclb;

static thread_local std::string g_err;
const char *tclb_last_error(void) { return g_err.c_str(); }
void tclb_log_init(int rank, int level) { log_init(rank, level); }

#define GUARD_BEGIN try {
#define GUARD_END_INT } catch (const std::exception &e) { g_err = e.what(); return -1; } return 0;
#define GUARD_END_PTR } catch (const std::exception &e) { g_err = e.what(); return nullptr; }

struct tclb_model_h { Model *m; };
struct tclb_domain_h { Domain *d; tclb_exchange_fn ex = nullptr; tclb_reduce_fn red = nullptr; void *user = nullptr;
                       std::vector<void *> ipc_opened; };

tclb_model_h *tclb_model_load(const char *path) {
    GUARD_BEGIN
    auto *h = new tclb_model_h{Model::load(path)};
    return h;
    GUARD_END_PTR
}
const tclb_model_t *tclb_model_desc(tclb_model_h *h) { return h->m->d; }

tclb_domain_h *tclb_domain_create(tclb_model_h *m, int backend, int nx, int ny, int nz, int nranks, const int *lr, int nl,
                                  const int *dev, int nd) {
    GUARD_BEGIN
    std::vector<int> L(lr, lr + nl), D(dev, dev + (dev ? nd : 0));
    auto *h = new tclb_domain_h();
    h->d = new Domain(m->m, backend, nx, ny, nz, nranks, L, D);
    return h;
    GUARD_END_PTR
}
void tclb_domain_destroy(tclb_domain_h *h) {
    if (!h) return;
    for (void *p : h->ipc_opened) cudaIpcCloseMemHandle(p);
    delete h->d;
    delete h;
}
int tclb_domain_decomposition(tclb_domain_h *h, int *divy, int *divz, int *b6) {
    GUARD_BEGIN
    *divy = h->d->dec.divy; *divz = h->d->dec.divz;
    for (int r = 0; r < h->d->dec.nranks; r++) {
        const Box &b = h->d->dec.boxes[r];
        int v[6] = {b.x0, b.y0, b.z0, b.nx, b.ny, b.nz};
        memcpy(b6 + 6 * r, v, sizeof(v));
    }
    GUARD_END_INT
}
int tclb_domain_set_halo_mode(tclb_domain_h *h, int direct) {
    GUARD_BEGIN
    h->d->use_direct_halo = direct != 0;
    h->d->finalize_wiring();
    GUARD_END_INT
}
static Lattice *rank_lat(tclb_domain_h *h, int rank) {
    if (rank < 0 || rank >= (int)h->d->lat.size() || !h->d->lat[rank]) fail("rank %d is not owned by this process", rank);
    return h->d->lat[rank];
}
int tclb_domain_set_setting(tclb_domain_h *h, const char *name, double v) {
    GUARD_BEGIN
    int i = h->d->model->find_setting(name);
    if (i >= 0) { h->d->set_setting(i, v); return 0; }
    i = h->d->model->find_zonesetting(name);
    if (i < 0) fail("unknown setting %s", name);
    h->d->set_zone_setting(i, -1, v);
    GUARD_END_INT
}
double tclb_domain_get_setting(tclb_domain_h *h, const char *name) {
    int i = h->d->model->find_setting(name);
    if (i >= 0) return h->d->lat[h->d->local[0]]->get_setting(i);
    // a zonal setting: its value in the default zone at the current time level
    i = h->d->model->find_zonesetting(name);
    if (i < 0) return NAN;
    Lattice *l = h->d->lat[h->d->local[0]];
    return l->zs.get(i, 0, l->zs.len() > 1 ? l->iter % l->zs.len() : 0);
}
int tclb_domain_set_zone_setting(tclb_domain_h *h, const char *name, int zone, double v) {
    GUARD_BEGIN
    int i = h->d->model->find_zonesetting(name);
    if (i < 0) fail("unknown zonal setting %s", name);
    h->d->set_zone_setting(i, zone, v);
    GUARD_END_INT
}
int tclb_domain_set_zone_series(tclb_domain_h *h, const char *name, int zone, const double *v, int n) {
    GUARD_BEGIN
    int i = h->d->model->find_zonesetting(name);
    if (i < 0) fail("unknown zonal setting %s", name);
    h->d->set_zone_series(i, zone, std::vector<double>(v, v + n));
    GUARD_END_INT
}
int tclb_domain_set_flags(tclb_domain_h *h, int rank, const uint32_t *fl) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, rank);
    Box b; b.nx = l->nx; b.ny = l->ny; b.nz = l->nz;
    l->flag_overwrite(fl, b);
    // make sure every zone id in use has a column in the zone table
    uint32_t zmax = 0;
    for (long long i = 0; i < b.size(); i++) zmax = std::max(zmax, fl[i] >> h->d->model->d->zone_shift);
    h->d->set_zones((int)zmax + 1);
    GUARD_END_INT
}
int tclb_domain_get_flags(tclb_domain_h *h, int rank, uint32_t *fl) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, rank);
    Box b; b.nx = l->nx; b.ny = l->ny; b.nz = l->nz;
    l->get_flags(b, fl);
    GUARD_END_INT
}
int tclb_domain_set_cuts(tclb_domain_h *h, int rank, const uint16_t *c) {
    GUARD_BEGIN
    rank_lat(h, rank)->cuts_overwrite(c);
    GUARD_END_INT
}
int tclb_domain_init(tclb_domain_h *h) { GUARD_BEGIN h->d->init(); GUARD_END_INT }
int tclb_domain_iterate(tclb_domain_h *h, int n, int gl, int ge) { GUARD_BEGIN h->d->iterate(n, gl != 0, ge != 0); GUARD_END_INT }
int tclb_domain_run_action(tclb_domain_h *h, const char *a, int gm) {
    GUARD_BEGIN
    int i = h->d->model->find_action(a);
    if (i < 0) fail("unknown action %s", a);
    h->d->run_action(i, gm);
    GUARD_END_INT
}
int tclb_domain_sync(tclb_domain_h *h) { GUARD_BEGIN h->d->sync(); GUARD_END_INT }
int tclb_domain_iter(tclb_domain_h *h) { return h->d->iter; }
int tclb_domain_get_globals(tclb_domain_h *h, double *out) {
    GUARD_BEGIN
    memcpy(out, h->d->globals.data(), sizeof(double) * h->d->model->d->n_globals);
    GUARD_END_INT
}
int tclb_domain_get_field(tclb_domain_h *h, int rank, const char *f, void *dst) {
    GUARD_BEGIN
    int i = h->d->model->find_field(f);
    if (i < 0) fail("unknown field %s", f);
    rank_lat(h, rank)->copy_field_out(i, dst);
    GUARD_END_INT
}
int tclb_domain_set_field(tclb_domain_h *h, int rank, const char *f, const void *src) {
    GUARD_BEGIN
    int i = h->d->model->find_field(f);
    if (i < 0) fail("unknown field %s", f);
    rank_lat(h, rank)->copy_field_in(i, src);
    // keep the ghost layers consistent with the new interior: plain copies into the neighbours' snapshots (local objects or
    // mapped peer memory; the other processes must not run before a barrier of the caller), else the exchange hook
    const int cur = rank_lat(h, rank)->cur;
    h->d->copy_exchange(cur);
    bool unwired = false;
    for (int d = 0; d < TCLB_MAX_NEIGH; d++) if (d != 4 && !rank_lat(h, rank)->halo[d].snap[cur]) unwired = true;
    if (unwired && h->d->exchange_hook) h->d->exchange_hook(cur);
    GUARD_END_INT
}
int tclb_domain_get_quantity(tclb_domain_h *h, int rank, const char *q, const int *b6, void *dst, double scale) {
    GUARD_BEGIN
    int i = h->d->model->find_quantity(q);
    if (i < 0) fail("unknown quantity %s", q);
    Lattice *l = rank_lat(h, rank);
    Box b;
    if (b6) { b.x0 = b6[0]; b.y0 = b6[1]; b.z0 = b6[2]; b.nx = b6[3]; b.ny = b6[4]; b.nz = b6[5]; }
    else { b.nx = l->nx; b.ny = l->ny; b.nz = l->nz; }
    l->get_quantity(i, b, dst, scale);
    GUARD_END_INT
}
int tclb_domain_color(tclb_domain_h *h, int rank, int z, float scale, unsigned char *rgba) {
    GUARD_BEGIN
    rank_lat(h, rank)->color(z, scale, rgba);
    GUARD_END_INT
}
int tclb_domain_reset_average(tclb_domain_h *h) {
    GUARD_BEGIN
    for (int r : h->d->local) h->d->lat[r]->reset_iter = h->d->lat[r]->iter;
    GUARD_END_INT
}
int tclb_domain_use_graphs(tclb_domain_h *h, int on) {
    GUARD_BEGIN h->d->use_graphs = on != 0; if (!on) h->d->drop_graphs(); GUARD_END_INT
}
long long tclb_domain_graph_iterations(tclb_domain_h *h) { return h->d->graph_replays; }
int tclb_domain_set_block(tclb_domain_h *h, int bx, int by) {
    GUARD_BEGIN
    for (int r : h->d->local) { h->d->lat[r]->block_x = bx; h->d->lat[r]->block_y = by; }
    GUARD_END_INT
}
int tclb_domain_pri_blocks(tclb_domain_h *h, int rank, long long *out27) { GUARD_BEGIN rank_lat(h, rank)->pri_block_elems(out27); GUARD_END_INT }
int tclb_domain_save_pri(tclb_domain_h *h, int rank, const char *f) { GUARD_BEGIN rank_lat(h, rank)->save_pri(f); GUARD_END_INT }
int tclb_domain_load_pri(tclb_domain_h *h, int rank, const char *f) { GUARD_BEGIN rank_lat(h, rank)->load_pri(f); GUARD_END_INT }
void *tclb_domain_snapshot_ptr(tclb_domain_h *h, int rank, int snap, int64_t shape[4], int *gy, int *gz, int *current) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, rank);
    shape[0] = h->d->model->d->n_fields; shape[1] = l->NZ; shape[2] = l->NY; shape[3] = l->nx;
    if (gy) *gy = l->gy;
    if (gz) *gz = l->gz;
    if (current) *current = l->cur;
    return l->snap_ptr(snap < 0 ? l->cur : snap);
    GUARD_END_PTR
}
void *tclb_domain_flags_ptr(tclb_domain_h *h, int rank) { GUARD_BEGIN return rank_lat(h, rank)->flags_ptr(); GUARD_END_PTR }
void *tclb_domain_stream(tclb_domain_h *h, int rank) { GUARD_BEGIN return rank_lat(h, rank)->stream; GUARD_END_PTR }

int tclb_domain_export_ipc(tclb_domain_h *h, int rank, void *handle64, int64_t *offs4) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, rank);
    if (l->backend != TCLB_BACKEND_CUDA) fail("IPC export needs the CUDA backend");
    TCLB_CUDA(cudaSetDevice(l->device));
    // the whole subdomain is one allocation, so one handle + four offsets describe it
    cudaIpcMemHandle_t hd;
    char *base = (char *)l->arena_base();
    TCLB_CUDA(cudaIpcGetMemHandle(&hd, base));
    static_assert(sizeof(hd) == 64, "cudaIpcMemHandle_t size");
    memcpy(handle64, &hd, 64);
    offs4[0] = (char *)l->snap_ptr(0) - base;
    offs4[1] = (char *)l->snap_ptr(1) - base;
    offs4[2] = (char *)l->flags_in() - base;
    offs4[3] = (char *)l->gl_ring_ptr() - base;
    GUARD_END_INT
}
int tclb_domain_import_peer(tclb_domain_h *h, int my_rank, int peer_rank, const void *handle64, const int64_t *offs4, int pny, int pnz) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, my_rank);
    TCLB_CUDA(cudaSetDevice(l->device));
    cudaIpcMemHandle_t hd;
    memcpy(&hd, handle64, 64);
    void *base = nullptr;
    TCLB_CUDA(cudaIpcOpenMemHandle(&base, hd, cudaIpcMemLazyEnablePeerAccess));
    h->ipc_opened.push_back(base);
    PeerInfo p;
    p.valid = true;
    p.snap[0] = (char *)base + offs4[0];
    p.snap[1] = (char *)base + offs4[1];
    p.flags_in = (unsigned *)((char *)base + offs4[2]);
    p.gl_ring = (double *)((char *)base + offs4[3]);
    p.ny = pny; p.nz = pnz;
    h->d->set_peer(my_rank, peer_rank, p);
    GUARD_END_INT
}
int tclb_domain_io_bytes(tclb_domain_h *h, int rank, int64_t *h2d, int64_t *d2h) {
    GUARD_BEGIN
    Lattice *l = rank_lat(h, rank);
    *h2d = (int64_t)l->last_h2d_bytes; *d2h = (int64_t)l->last_d2h_bytes;
    GUARD_END_INT
}
int tclb_domain_finalize_wiring(tclb_domain_h *h) { GUARD_BEGIN h->d->finalize_wiring(); GUARD_END_INT }
int tclb_domain_set_hooks(tclb_domain_h *h, tclb_exchange_fn ex, tclb_reduce_fn red, void *user) {
    GUARD_BEGIN
    h->ex = ex; h->red = red; h->user = user;
    if (ex) h->d->exchange_hook = [h](int s) { h->ex(h->user, s); };
    else h->d->exchange_hook = nullptr;
    if (red) h->d->reduce_hook = [h](double *v, int n, int op) { h->red(h->user, v, n, op); };
    else h->d->reduce_hook = nullptr;
    GUARD_END_INT
}

// ---- units and case runner (host layer) -------------------------------------------------------
#include "../host/solver.h"
#include "../host/units.h"
struct tclb_units_h { UnitEnv env; };
tclb_units_h *tclb_units_create(void) { return new tclb_units_h(); }
void tclb_units_destroy(tclb_units_h *u) { delete u; }
int tclb_units_set(tclb_units_h *u, const char *name, const char *value, const char *gauge) {
    GUARD_BEGIN
    u->env.set_unit(name, u->env.read_text(value) / u->env.read_text(gauge));
    GUARD_END_INT
}
int tclb_units_make_gauge(tclb_units_h *u) { GUARD_BEGIN u->env.make_gauge(); GUARD_END_INT }
int tclb_units_alt(tclb_units_h *u, const char *expr, double *out) { GUARD_BEGIN *out = u->env.alt(std::string(expr)); GUARD_END_INT }
int tclb_units_si(tclb_units_h *u, const char *expr, double *out, double *exponents9) {
    GUARD_BEGIN
    UnitVal v = u->env.read_text(expr);
    *out = v.val;
    if (exponents9) for (int i = 0; i < N_BASE_UNITS; i++) exponents9[i] = v.uni[i];
    GUARD_END_INT
}

struct tclb_solver_h { Solver *s; tclb_domain_h dom; };
tclb_solver_h *tclb_solver_create(const char *model, const char *precision, int backend, int nranks, const int *local_ranks, int n_local,
                                  const int *devices, int n_devices, int direct_halo) {
    GUARD_BEGIN
    SolverOptions o;
    o.model = model;
    o.precision = precision;
    o.backend = backend;
    o.nranks = nranks;
    if (local_ranks) o.local_ranks.assign(local_ranks, local_ranks + n_local);
    if (devices) o.devices.assign(devices, devices + n_devices);
    o.direct_halo = direct_halo != 0;
    auto *h = new tclb_solver_h();
    h->s = new Solver(o);
    h->dom.d = nullptr;
    return h;
    GUARD_END_PTR
}
void tclb_solver_destroy(tclb_solver_h *h) {
    if (!h) return;
    delete h->s;
    delete h;
}
static std::vector<std::string> to_args(const char *const *argv, int argc) {
    std::vector<std::string> a;
    for (int i = 0; i < argc; i++) a.push_back(argv[i]);
    return a;
}
int tclb_solver_load_file(tclb_solver_h *h, const char *file, const char *const *xargs, int n) {
    GUARD_BEGIN h->s->load_case(file, to_args(xargs, n)); h->dom.d = h->s->domain.get(); GUARD_END_INT
}
int tclb_solver_load_string(tclb_solver_h *h, const char *xml, const char *name, const char *const *xargs, int n) {
    GUARD_BEGIN h->s->load_case_string(xml, name, to_args(xargs, n)); h->dom.d = h->s->domain.get(); GUARD_END_INT
}
int tclb_solver_run(tclb_solver_h *h) {
    GUARD_BEGIN
    if (h->s->run() != 0) fail("case run failed (see log)");
    GUARD_END_INT
}
int tclb_solver_set_script_hook(tclb_solver_h *h, tclb_script_fn fn, void *user) {
    GUARD_BEGIN h->s->script_hook = fn; h->s->script_user = user; GUARD_END_INT
}
int tclb_solver_zone_id(tclb_solver_h *h, const char *name) {
    auto it = h->s->setting_zones.find(name);
    return it == h->s->setting_zones.end() ? -1 : it->second;
}
tclb_domain_h *tclb_solver_domain(tclb_solver_h *h) { return h->dom.d ? &h->dom : nullptr; }
int tclb_solver_iter(tclb_solver_h *h) { return h->s->iter; }
const char *tclb_solver_outpath(tclb_solver_h *h) { return h->s->outpath.c_str(); }

// ---- adjoint (csrc/core/adjoint.h) ----
struct tclb_adjoint_h { Adjoint *a; };
tclb_adjoint_h *tclb_adjoint_create(tclb_domain_h *h) {
    GUARD_BEGIN
    return new tclb_adjoint_h{new Adjoint(h->d)};
    GUARD_END_PTR
}
void tclb_adjoint_destroy(tclb_adjoint_h *a) { if (a) { delete a->a; delete a; } }
int tclb_adjoint_record(tclb_adjoint_h *a, int n) { GUARD_BEGIN a->a->record(n); GUARD_END_INT }
int tclb_adjoint_reverse(tclb_adjoint_h *a) { GUARD_BEGIN a->a->reverse(); GUARD_END_INT }
int tclb_adjoint_steady(tclb_adjoint_h *a, int n) { GUARD_BEGIN a->a->steady(n); GUARD_END_INT }
int tclb_adjoint_set_tape_frames(tclb_adjoint_h *a, int n) { GUARD_BEGIN a->a->set_tape_frames(n); GUARD_END_INT }
int tclb_adjoint_tape_frames(tclb_adjoint_h *a) { return (int)a->a->tape_frames(); }
int tclb_adjoint_zero(tclb_adjoint_h *a) { GUARD_BEGIN a->a->zero(); a->a->objective = 0; GUARD_END_INT }
double tclb_adjoint_objective(tclb_adjoint_h *a) { return a->a->objective; }
int tclb_adjoint_want_setting(tclb_adjoint_h *a, tclb_domain_h *h, const char *names_csv) {
    GUARD_BEGIN
    std::vector<int> idx;
    std::string all = names_csv ? names_csv : "";
    size_t p0 = 0;
    while (p0 <= all.size()) {
        size_t c = all.find(',', p0);
        std::string nm = all.substr(p0, c == std::string::npos ? std::string::npos : c - p0);
        if (!nm.empty()) {
            int i = h->d->model->find_setting(nm);
            if (i < 0) fail("unknown (non-zonal) setting %s", nm.c_str());
            idx.push_back(i);
        }
        if (c == std::string::npos) break;
        p0 = c + 1;
    }
    a->a->want_setting_gradients(idx);
    GUARD_END_INT
}
int tclb_adjoint_setting_gradient(tclb_adjoint_h *a, tclb_domain_h *h, const char *name, double *out) {
    GUARD_BEGIN
    int i = h->d->model->find_setting(name);
    if (i < 0) fail("unknown (non-zonal) setting %s", name);
    *out = a->a->setting_gradient(i);
    GUARD_END_INT
}
int tclb_adjoint_get_field(tclb_adjoint_h *a, tclb_domain_h *h, const char *f, void *dst) {
    GUARD_BEGIN
    int i = h->d->model->find_field(f);
    if (i < 0) fail("unknown field %s", f);
    a->a->get_field(i, dst);
    GUARD_END_INT
}
int tclb_adjoint_set_field(tclb_adjoint_h *a, tclb_domain_h *h, const char *f, const void *src) {
    GUARD_BEGIN
    int i = h->d->model->find_field(f);
    if (i < 0) fail("unknown field %s", f);
    a->a->set_field(i, src);
    GUARD_END_INT
}
