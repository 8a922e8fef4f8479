#include "lattice.h"
#include "common.h"
#include "tclb_b200.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>

namespace tclb {

const DebugFlags &debug_flags() {
    static const DebugFlags f = [] {
        DebugFlags d;
        if (const char *e = getenv("TCLB_DEBUG")) {
            d.sync = strstr(e, "sync") != nullptr;
            d.poison = strstr(e, "poison") != nullptr;
        }
        if (const char *p = getenv("TCLB_PROFILE")) d.profile = p[0] && p[0] != '0';
        return d;
    }();
    return f;
}

Box Box::intersect(const Box &o) const {
    Box r;
    r.x0 = std::max(x0, o.x0); r.y0 = std::max(y0, o.y0); r.z0 = std::max(z0, o.z0);
    r.nx = std::min(x0 + nx, o.x0 + o.nx) - r.x0;
    r.ny = std::min(y0 + ny, o.y0 + o.ny) - r.y0;
    r.nz = std::min(z0 + nz, o.z0 + o.nz) - r.z0;
    if (r.nx < 0) r.nx = 0;
    if (r.ny < 0) r.ny = 0;
    if (r.nz < 0) r.nz = 0;
    return r;
}

// ------------------------------------------------------------------------------------------
Lattice::Lattice(Model *m, int backend_, int device_, int rank_, const Box &local, const int total[3], void *shared_stream)
    : model(m), backend(backend_), device(device_), rank(rank_), region(local), cur(0), iter(0), reset_iter(0), epoch(0),
      halo_mode(TCLB_HALO_DIRECT), sig_enable(false), stream(nullptr), arena_(backend_, device_), zone_dev_(nullptr),
      zone_dev_bytes_(0), zone_stride_(1), cuts_dev_(nullptr), st_dev_(nullptr), st_n_(0), scratch_(nullptr), scratch_bytes_(0) {
    tot[0] = total[0]; tot[1] = total[1]; tot[2] = total[2];
    nx = local.nx; ny = local.ny; nz = local.nz;
    gy = m->d->gy; gz = m->d->gz;
    NY = ny + 2 * gy; NZ = nz + 2 * gz;
    plane_elems = (size_t)nx * NY * NZ;
    // the kernels index one field with 32-bit integers (node index, row offsets, tile numbers)
    if (plane_elems >= (size_t)1 << 31)
        fail("subdomain of %d x %d x %d nodes (with ghost layers %zu values per field) exceeds the 2^31 values per field the kernels index; "
             "use more subdomains (--ranks / more processes)", nx, ny, nz, plane_elems);
    snap_bytes = plane_elems * m->d->n_fields * m->storage_size();
    int s0 = arena_.reserve(snap_bytes, "snapshot0");
    int s1 = arena_.reserve(snap_bytes, "snapshot1");
    int sf = arena_.reserve(plane_elems * m->flag_size(), "flags");
    int sg = arena_.reserve(sizeof(double) * TCLB_GL_SLOTS * std::max(1, m->d->n_globals), "globals");
    int ss = arena_.reserve(sizeof(unsigned) * 64, "halo flags");
    int sr = arena_.reserve(sizeof(double) * TCLB_GL_RING * TCLB_MAX_RANKS * (std::max(1, m->d->n_globals) + 1), "globals ring");
    // one Globals partial per thread block of a stage launch: sized for the block shape with the most blocks
    gl_partial_stride_ = 0;
    if (backend == TCLB_BACKEND_CUDA)
        for (int bx : {32, 64, 128}) {
            const int by = 128 / bx;
            gl_partial_stride_ = std::max(gl_partial_stride_, 4LL * ((nx + bx - 1) / bx) * ((ny + by - 1) / by) * nz);
        }
    int sp = arena_.reserve((size_t)gl_partial_stride_ * std::max(1, m->d->n_globals) * m->real_size(), "globals partials");
    int sd = arena_.reserve(64, "graph iteration/epoch");
    arena_.finalize();
    dyn_dev_ = (int *)arena_.ptr(sd);
    gl_partial_ = gl_partial_stride_ > 0 ? arena_.ptr(sp) : nullptr;
    gl_ring_ = (double *)arena_.ptr(sr);
    snap_[0] = arena_.ptr(s0); snap_[1] = arena_.ptr(s1);
    flags_ = arena_.ptr(sf);
    globals_dev_ = (double *)arena_.ptr(sg);
    sig_ = (unsigned *)arena_.ptr(ss);
    gl_counter_ = sig_ + 16;
    settings.assign(std::max(1, m->d->n_settings), 0.0);
    globals.assign(std::max(1, m->d->n_globals), 0.0);
    zs.configure(m->d->n_zonesettings, m->d->zone_max + 1);
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        // subdomains that share a device share ONE stream: their launches are then ordered, which is what
        // makes several ranks per GPU correct without halo flags (and flags between co-resident kernels on one
        // GPU could dead-lock: nothing guarantees that two launches run at the same time)
        if (shared_stream) { stream = shared_stream; owns_stream_ = false; }
        else {
            cudaStream_t s;
            TCLB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
            stream = (void *)s;
            owns_stream_ = true;
        }
        pinned_bytes_ = 1 << 17;
        TCLB_CUDA(cudaHostAlloc(&pinned_, pinned_bytes_, cudaHostAllocDefault));
    }
    for (int d = 0; d < TCLB_MAX_NEIGH; d++) halo[d] = HaloTarget();
    clear_globals();
}
PeerInfo Lattice::self_info() const {
    PeerInfo p;
    p.valid = true;
    p.snap[0] = snap_[0]; p.snap[1] = snap_[1];
    p.flags_in = sig_;
    p.gl_ring = gl_ring_;
    p.ny = ny; p.nz = nz;
    return p;
}

Lattice::~Lattice() {
    if (backend == TCLB_BACKEND_CUDA) {
        cudaSetDevice(device);
        if (stream && owns_stream_) cudaStreamDestroy((cudaStream_t)stream);
        if (zone_dev_) cudaFree(zone_dev_);
        if (cuts_dev_) cudaFree(cuts_dev_);
        if (particles_dev_) cudaFree(particles_dev_);
        if (pgrid_dev_) cudaFree(pgrid_dev_);
        if (ptree_dev_) cudaFree(ptree_dev_);
        if (pgrid_counts_) cudaFree(pgrid_counts_);
        if (pgrid_overflow_) cudaFree(pgrid_overflow_);
        if (st_dev_) cudaFree(st_dev_);
        if (scratch_) cudaFree(scratch_);
        if (pinned_) cudaFreeHost(pinned_);
    } else {
        free(zone_dev_); free(cuts_dev_); free(st_dev_); free(scratch_); free(particles_dev_); free(ptree_dev_); free(pgrid_dev_); free(pgrid_counts_); free(pgrid_overflow_);
    }
}

void Lattice::to_host(void *dst, const void *src, size_t bytes) const {
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
        TCLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    } else memcpy(dst, src, bytes);
}
void Lattice::to_dev(void *dst, const void *src, size_t bytes) const {
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
        TCLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    } else memcpy(dst, src, bytes);
}
void Lattice::sync() const {
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    }
}
void *Lattice::scratch(size_t bytes) {
    if (bytes <= scratch_bytes_) return scratch_;
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        if (scratch_) TCLB_CUDA(cudaFree(scratch_));
        TCLB_CUDA(cudaMalloc(&scratch_, bytes));
    } else {
        free(scratch_);
        scratch_ = malloc(bytes);
    }
    scratch_bytes_ = bytes;
    return scratch_;
}
// ---- particles ---------------------------------------------------------------------------
void Lattice::set_particles(const std::vector<double> &rows) {
    const int n = (int)(rows.size() / 16), rs = model->real_size();
    if (n > particles_cap_) {
        if (backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaSetDevice(device));
            sync();
            if (particles_dev_) TCLB_CUDA(cudaFree(particles_dev_));
            TCLB_CUDA(cudaMalloc(&particles_dev_, (size_t)n * 2 * 16 * rs));
        } else {
            free(particles_dev_);
            particles_dev_ = malloc((size_t)n * 2 * 16 * rs);
        }
        particles_cap_ = n * 2;
    }
    n_particles_ = n;
    pgrid_on_ = false;
    if (n == 0) return;
    if (rs == 4) { std::vector<float> f(rows.begin(), rows.end()); to_dev(particles_dev_, f.data(), f.size() * 4); }
    else to_dev(particles_dev_, rows.data(), rows.size() * 8);
    ptree_on_ = false;
    const char *sel = getenv("TCLB_SOLID_CONTAINER");
    if (sel && !strcmp(sel, "tree")) build_particle_tree(rows);
    else build_particle_grid(rows);
}
std::vector<int> build_particle_tree_words(const std::vector<double> &rows16, int n, double margin);
// TCLB_SOLID_CONTAINER=tree (the reference's --with-solid-container=tree): threaded box tree, built on the host, walked on the device
void Lattice::build_particle_tree(const std::vector<double> &rows) {
    if (model->d->part_margin > 0) part_margin = model->d->part_margin;
    std::vector<int> words = build_particle_tree_words(rows, n_particles_, part_margin);
    ptree_nodes_ = (int)(words.size() / 8);
    if (words.empty()) return;
    if (words.size() > ptree_cap_) {
        const size_t cap = words.size() * 2;
        if (backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaSetDevice(device));
            sync();
            if (ptree_dev_) cudaFree(ptree_dev_);
            TCLB_CUDA(cudaMalloc((void **)&ptree_dev_, cap * sizeof(int)));
        } else {
            free(ptree_dev_);
            ptree_dev_ = (int *)malloc(cap * sizeof(int));
        }
        ptree_cap_ = cap;
    }
    to_dev(ptree_dev_, words.data(), words.size() * sizeof(int));
    ptree_on_ = true;
}
extern "C" int tclb_particle_grid_fill(const void *rows, int n, int real_size, const double *origin, double inv_delta, const int *gn,
                                       int depth, int *grid, int *counts, int *overflow, int on_device, void *stream);
// Container over the particle table (TCLB_SOLID_CONTAINER=all switches it off, like the reference's
// --with-solid-container=all): cell edge = 2 * (largest radius + margin), so a node looks at 2 cells per axis.
void Lattice::build_particle_grid(const std::vector<double> &rows) {
    const char *sel = getenv("TCLB_SOLID_CONTAINER");
    if (sel && !strcmp(sel, "all")) return;
    const int n = n_particles_;
    if (model->d->part_margin > 0) part_margin = model->d->part_margin;
    double rmax = 0;
    for (int i = 0; i < n; i++) rmax = std::max(rmax, rows[(size_t)i * 16]);
    const double delta = 2.0 * (rmax + part_margin) * (1.0 + 1e-6) + 1e-9;
    const int nn[3] = {nx, ny, nz};
    const double lo[3] = {region.x0 + origin[0] + 0.5, region.y0 + origin[1] + 0.5, region.z0 + origin[2] + 0.5};
    size_t cells = 1;
    for (int k = 0; k < 3; k++) {
        pg_origin_[k] = lo[k] - delta;              // one cell of room around the nodes of this subdomain
        pg_n_[k] = (int)std::ceil((nn[k] - 1 + 2.0 * delta) / delta) + 1;
        cells *= (size_t)pg_n_[k];
    }
    pg_inv_delta_ = 1.0 / delta;
    const bool dev = backend == TCLB_BACKEND_CUDA;
    for (int attempt = 0; attempt < 12; attempt++) {
        if (cells * pg_depth_ > pgrid_cap_ || cells > pgrid_cells_cap_) {
            if (dev) {
                TCLB_CUDA(cudaSetDevice(device));
                sync();
                if (pgrid_dev_) cudaFree(pgrid_dev_);
                if (pgrid_counts_) cudaFree(pgrid_counts_);
                if (!pgrid_overflow_) TCLB_CUDA(cudaMalloc((void **)&pgrid_overflow_, sizeof(int)));
                TCLB_CUDA(cudaMalloc((void **)&pgrid_dev_, cells * pg_depth_ * sizeof(int)));
                TCLB_CUDA(cudaMalloc((void **)&pgrid_counts_, cells * sizeof(int)));
            } else {
                free(pgrid_dev_); free(pgrid_counts_);
                if (!pgrid_overflow_) pgrid_overflow_ = (int *)malloc(sizeof(int));
                pgrid_dev_ = (int *)malloc(cells * pg_depth_ * sizeof(int));
                pgrid_counts_ = (int *)malloc(cells * sizeof(int));
            }
            pgrid_cap_ = cells * pg_depth_;
            pgrid_cells_cap_ = cells;
        }
        if (tclb_particle_grid_fill(particles_dev_, n, model->real_size(), pg_origin_, pg_inv_delta_, pg_n_, pg_depth_, pgrid_dev_,
                                    pgrid_counts_, pgrid_overflow_, dev ? 1 : 0, stream) != 0)
            fail("particle grid build failed");
        int over = 0;
        to_host(&over, pgrid_overflow_, sizeof(int));
        if (!over) { pgrid_on_ = true; return; }
        pg_depth_ *= 2;                              // a cell holds more particles than slots: deeper cells, again
    }
    logf(L_WARNING, "particle grid: cells stay overfull, falling back to the scan over every particle");
}
void Lattice::get_particles(std::vector<double> &rows) {
    const int rs = model->real_size();
    rows.assign((size_t)n_particles_ * 16, 0.0);
    if (n_particles_ == 0) return;
    if (rs == 4) {
        std::vector<float> f(rows.size());
        to_host(f.data(), particles_dev_, f.size() * 4);
        for (size_t i = 0; i < f.size(); i++) rows[i] = f[i];
    } else to_host(rows.data(), particles_dev_, rows.size() * 8);
}
// debug aid: every ghost value of `snap` becomes NaN (all-ones bytes are a NaN in half, float and double)
void Lattice::poison_ghosts(int snap, int stage) {
    const size_t ss = model->storage_size(), row = (size_t)nx * ss, plane = (size_t)NY * row;
    const unsigned char *saved = model->d->stages[stage].save_field;
    for (int f = 0; f < model->d->n_fields; f++) {
        if (saved && !saved[f]) continue;
        char *base = (char *)field_ptr(snap, f);
        if (backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaSetDevice(device));
            cudaStream_t st = (cudaStream_t)stream;
            if (gz > 0) {
                TCLB_CUDA(cudaMemsetAsync(base, 0xFF, (size_t)gz * plane, st));
                TCLB_CUDA(cudaMemsetAsync(base + (size_t)(nz + gz) * plane, 0xFF, (size_t)gz * plane, st));
            }
            if (gy > 0) {
                TCLB_CUDA(cudaMemset2DAsync(base + (size_t)gz * plane, plane, 0xFF, (size_t)gy * row, nz, st));
                TCLB_CUDA(cudaMemset2DAsync(base + (size_t)gz * plane + (size_t)(ny + gy) * row, plane, 0xFF, (size_t)gy * row, nz, st));
            }
        } else {
            if (gz > 0) { memset(base, 0xFF, (size_t)gz * plane); memset(base + (size_t)(nz + gz) * plane, 0xFF, (size_t)gz * plane); }
            for (int z = 0; z < nz && gy > 0; z++) {
                memset(base + (size_t)(z + gz) * plane, 0xFF, (size_t)gy * row);
                memset(base + (size_t)(z + gz) * plane + (size_t)(ny + gy) * row, 0xFF, (size_t)gy * row);
            }
        }
    }
}
void *Lattice::field_ptr(int snap, int field) const {
    return (char *)snap_[snap] + (size_t)field * plane_elems * model->storage_size();
}

// ---- flags -------------------------------------------------------------------------------
// Flags of a box: only the z-planes the box touches travel (they are one contiguous run of the slab), not the whole array.
void Lattice::flag_overwrite(const uint32_t *fl, const Box &b) {
    if (b.nx <= 0 || b.ny <= 0 || b.nz <= 0) return;
    const int fs = model->flag_size();
    const size_t plane = (size_t)NY * nx, first = (size_t)(b.z0 + gz) * plane;
    std::vector<unsigned char> host((size_t)b.nz * plane * fs);
    const bool whole_rows = b.x0 == 0 && b.nx == nx && b.y0 == 0 && b.ny == ny && gy == 0;
    if (!whole_rows) to_host(host.data(), (const char *)flags_ + first * fs, host.size());     // keep what the box does not cover
    for (int z = 0; z < b.nz; z++)
        for (int y = 0; y < b.ny; y++)
            for (int x = 0; x < b.nx; x++) {
                size_t i = ((size_t)z * NY + (y + b.y0 + gy)) * nx + (x + b.x0);
                uint32_t v = fl[((size_t)z * b.ny + y) * b.nx + x];
                if (fs == 2) ((uint16_t *)host.data())[i] = (uint16_t)v;
                else ((uint32_t *)host.data())[i] = v;
            }
    to_dev((char *)flags_ + first * fs, host.data(), host.size());
}
void Lattice::get_flags(const Box &b, uint32_t *out) const {
    if (b.nx <= 0 || b.ny <= 0 || b.nz <= 0) return;
    const int fs = model->flag_size();
    const size_t plane = (size_t)NY * nx, first = (size_t)(b.z0 + gz) * plane;
    std::vector<unsigned char> host((size_t)b.nz * plane * fs);
    to_host(host.data(), (const char *)flags_ + first * fs, host.size());
    for (int z = 0; z < b.nz; z++)
        for (int y = 0; y < b.ny; y++)
            for (int x = 0; x < b.nx; x++) {
                size_t i = ((size_t)z * NY + (y + b.y0 + gy)) * nx + (x + b.x0);
                out[((size_t)z * b.ny + y) * b.nx + x] = fs == 2 ? ((uint16_t *)host.data())[i] : ((uint32_t *)host.data())[i];
            }
}
void Lattice::cuts_overwrite(const uint16_t *cuts) {
    size_t bytes = (size_t)26 * plane_elems * sizeof(uint16_t);
    if (!cuts) return;
    graph_version++;
    if (!cuts_dev_) {
        if (backend == TCLB_BACKEND_CUDA) { TCLB_CUDA(cudaSetDevice(device)); TCLB_CUDA(cudaMalloc(&cuts_dev_, bytes)); }
        else cuts_dev_ = malloc(bytes);
    }
    std::vector<uint16_t> host((size_t)26 * plane_elems, 65535);
    for (int d = 0; d < 26; d++)
        for (int z = 0; z < nz; z++)
            for (int y = 0; y < ny; y++)
                memcpy(&host[(size_t)d * plane_elems + ((size_t)(z + gz) * NY + (y + gy)) * nx],
                       &cuts[(((size_t)d * nz + z) * ny + y) * nx], sizeof(uint16_t) * nx);
    to_dev(cuts_dev_, host.data(), bytes);
}

// ---- fields ------------------------------------------------------------------------------
// IEEE binary16 <-> float on the host (the lattice may be stored narrower than the node code computes)
static float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16, exp = (h >> 10) & 31u, man = h & 1023u;
    uint32_t bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {                                   // subnormal: renormalise
            int e = -1; uint32_t m = man;
            do { e++; m <<= 1; } while (!(m & 1024u));
            bits = sign | (uint32_t)(127 - 15 - e) << 23 | (m & 1023u) << 13;
        }
    } else if (exp == 31) bits = sign | 0x7f800000u | man << 13;
    else bits = sign | (exp + 127 - 15) << 23 | man << 13;
    float f; memcpy(&f, &bits, 4); return f;
}
static uint16_t float_to_half(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    const int32_t exp = (int32_t)((x >> 23) & 255u) - 127 + 15;
    uint32_t man = x & 0x7fffffu;
    if (((x >> 23) & 255u) == 255u) return (uint16_t)(sign | 0x7c00u | (man ? 0x200u : 0));
    if (exp >= 31) return (uint16_t)(sign | 0x7c00u);
    if (exp <= 0) {
        if (exp < -10) return (uint16_t)sign;
        man |= 0x800000u;
        const int shift = 14 - exp;
        uint32_t h = man >> shift;
        const uint32_t rem = man & ((1u << shift) - 1), half = 1u << (shift - 1);
        if (rem > half || (rem == half && (h & 1))) h++;
        return (uint16_t)(sign | h);
    }
    uint32_t h = (uint32_t)exp << 10 | man >> 13;
    const uint32_t rem = man & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (h & 1))) h++;       // round to nearest even (may carry into the exponent)
    return (uint16_t)(sign | h);
}
// storage_t -> real_t (+ shift) and back, for the host-side field accessors
void Lattice::storage_to_real(int field, const void *src, void *dst, size_t n) const {
    const int rs = model->real_size(), ss = model->storage_size();
    const double sh = model->field_shift(field);
    for (size_t i = 0; i < n; i++) {
        const double v = (ss == 2 ? (double)half_to_float(((const uint16_t *)src)[i]) : ss == 4 ? (double)((const float *)src)[i] : ((const double *)src)[i]) + sh;
        if (rs == 4) ((float *)dst)[i] = (float)v; else ((double *)dst)[i] = v;
    }
}
void Lattice::real_to_storage(int field, const void *src, void *dst, size_t n) const {
    const int rs = model->real_size(), ss = model->storage_size();
    const double sh = model->field_shift(field);
    for (size_t i = 0; i < n; i++) {
        const double v = (rs == 4 ? (double)((const float *)src)[i] : ((const double *)src)[i]) - sh;
        if (ss == 2) ((uint16_t *)dst)[i] = float_to_half((float)v);
        else if (ss == 4) ((float *)dst)[i] = (float)v;
        else ((double *)dst)[i] = v;
    }
}
void Lattice::copy_field_out(int field, void *dst_real) const {
    const int rs = model->storage_size();
    const bool convert = rs != model->real_size() || model->field_shift(field) != 0.0;
    std::vector<char> staged(convert ? (size_t)nz * ny * nx * rs : 0);
    void *dst = convert ? (void *)staged.data() : dst_real;
    copy_field_out_raw(field, dst);
    if (convert) storage_to_real(field, staged.data(), dst_real, (size_t)nz * ny * nx);
}
void Lattice::copy_field_out_raw(int field, void *dst) const {
    const int rs = model->storage_size();
    const char *src = (const char *)field_ptr(cur, field);
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaMemcpy2DAsync(dst, (size_t)ny * nx * rs, src + ((size_t)gz * NY + gy) * nx * rs, (size_t)NY * nx * rs,
                                    (size_t)ny * nx * rs, (size_t)nz, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
        TCLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    } else {
        for (int z = 0; z < nz; z++)
            memcpy((char *)dst + (size_t)z * ny * nx * rs, src + ((size_t)(z + gz) * NY + gy) * nx * rs, (size_t)ny * nx * rs);
    }
}
void Lattice::copy_field_in(int field, const void *src_real) {
    const int ss = model->storage_size();
    if (ss != model->real_size() || model->field_shift(field) != 0.0) {
        std::vector<char> staged((size_t)nz * ny * nx * ss);
        real_to_storage(field, src_real, staged.data(), (size_t)nz * ny * nx);
        copy_field_in_raw(field, staged.data());
    } else copy_field_in_raw(field, src_real);
}
void Lattice::copy_field_in_raw(int field, const void *srcv) {
    const int rs = model->storage_size();
    char *dst = (char *)field_ptr(cur, field);
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        // the interior of a z-plane is one contiguous run (ny rows of nx): the whole field is ONE pitched copy
        TCLB_CUDA(cudaMemcpy2DAsync(dst + ((size_t)gz * NY + gy) * nx * rs, (size_t)NY * nx * rs, srcv, (size_t)ny * nx * rs,
                                    (size_t)ny * nx * rs, (size_t)nz, cudaMemcpyHostToDevice, (cudaStream_t)stream));
        TCLB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    } else {
        for (int z = 0; z < nz; z++)
            memcpy(dst + ((size_t)(z + gz) * NY + gy) * nx * rs, (const char *)srcv + (size_t)z * ny * nx * rs, (size_t)ny * nx * rs);
    }
}
void Lattice::fill_field(int field, double v) {
    const int rs = model->real_size();
    std::vector<char> buf((size_t)nz * ny * nx * rs);
    if (rs == 4) std::fill((float *)buf.data(), (float *)buf.data() + (size_t)nz * ny * nx, (float)v);
    else std::fill((double *)buf.data(), (double *)buf.data() + (size_t)nz * ny * nx, v);
    copy_field_in(field, buf.data());
}

// ---- settings ----------------------------------------------------------------------------
void Lattice::set_setting(int idx, double v) {
    // value, then every setting derived from it, transitively (reference Lattice::SetSetting
    // src/Lattice.cu.Rt:1350-1358 chains the generated DF_<name> functions the same way)
    int guard = 0;
    graph_version++;                       // settings travel by value in the captured kernel parameters
    while (idx >= 0 && guard++ < 64) {
        settings[idx] = v;
        const tclb_setting_t &s = model->d->settings[idx];
        if (s.derived_target < 0 || !s.derive) break;
        v = s.derive(v);
        idx = s.derived_target;
    }
}
void Lattice::set_synthetic_turbulence(const std::vector<tclb_st_mode_t> &modes) {
    size_t bytes = modes.size() * sizeof(tclb_st_mode_t);
    if (backend == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device));
        if (st_dev_) TCLB_CUDA(cudaFree(st_dev_));
        st_dev_ = nullptr;
        if (bytes) TCLB_CUDA(cudaMalloc(&st_dev_, bytes));
    } else {
        free(st_dev_);
        st_dev_ = bytes ? malloc(bytes) : nullptr;
    }
    if (bytes) to_dev(st_dev_, modes.data(), bytes);
    st_n_ = (int)modes.size();
    graph_version++;
}
void Lattice::upload_zone_table_if_dirty() {
    if (!zs.dirty() && zone_dev_) return;
    int stride = zs.zones();
    std::vector<double> tab;
    zs.build_table(tab, stride);
    {
        // table layout [len][n_zonesettings][stride]; the <Global>InObj weights are the settings from in_obj_offset on
        const int ns = zs.n_settings(), first = model->d->in_obj_offset, ng = model->d->sum_globals;
        obj_enabled_ = false;
        for (int t = 0; t < zs.len() && !obj_enabled_; t++)
            for (int s = std::max(0, first); s < std::min(ns, first + ng) && !obj_enabled_; s++)
                for (int z = 0; z < stride; z++)
                    if (tab[((size_t)t * ns + s) * stride + z] != 0.0) { obj_enabled_ = true; break; }
    }
    const int rs = model->real_size();
    size_t bytes = std::max<size_t>(tab.size(), 1) * rs;
    if (bytes > zone_dev_bytes_) {
        if (backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaSetDevice(device));
            sync();
            if (zone_dev_) TCLB_CUDA(cudaFree(zone_dev_));
            TCLB_CUDA(cudaMalloc(&zone_dev_, bytes));
        } else {
            free(zone_dev_);
            zone_dev_ = malloc(bytes);
        }
        zone_dev_bytes_ = bytes;
    }
    // staged through page-locked host memory so the upload is a true asynchronous DMA
    if (tab.empty()) tab.push_back(0.0);
    std::vector<float> f32;
    const void *src = tab.data();
    if (rs == 4) { f32.assign(tab.begin(), tab.end()); src = f32.data(); }
    if (pinned_ && bytes <= pinned_bytes_ / 2) { memcpy(pinned_, src, bytes); src = pinned_; }
    to_dev(zone_dev_, src, bytes);
    last_h2d_bytes = bytes;
    zone_stride_ = stride;
    graph_version++;
    zs.clear_dirty();
}

// ---- launches ----------------------------------------------------------------------------
void Lattice::fill_launch(tclb_launch_t &L) const {
    memset(&L, 0, sizeof(L));
    L.backend = backend;
    L.stream = stream;
    L.nx = nx; L.ny = ny; L.nz = nz; L.gy = gy; L.gz = gz;
    L.off_x = region.x0; L.off_y = region.y0; L.off_z = region.z0;
    L.pos_x = origin[0] + 0.5; L.pos_y = origin[1] + 0.5; L.pos_z = origin[2] + 0.5;
    L.tot_nx = tot[0]; L.tot_ny = tot[1]; L.tot_nz = tot[2];
    L.flags = flags_;
    L.cuts = cuts_dev_;
    L.globals = globals_dev_;
    L.settings = settings.data();
    L.zone_table = zone_dev_;
    L.zone_len = zs.len();
    L.zone_index = zs.len() > 1 ? iter % zs.len() : 0;
    L.zone_stride = zone_stride_;
    L.iter = iter; L.reset_iter = reset_iter;
    L.st_modes = (const tclb_st_mode_t *)st_dev_; L.st_n = st_n_;
    L.block_x = block_x; L.block_y = block_y;
    L.particles = particles_dev_; L.n_particles = n_particles_;
    L.pgrid = pgrid_on_ ? pgrid_dev_ : nullptr;
    for (int k = 0; k < 3; k++) { L.pg_n[k] = pg_n_[k]; L.pg_origin[k] = pg_origin_[k]; }
    L.pg_depth = pg_depth_; L.pg_inv_delta = pg_inv_delta_;
    if (ptree_on_) { L.pgrid = ptree_dev_; L.pg_depth = -1; L.pg_n[0] = ptree_nodes_; }      // depth -1: pgrid holds tree nodes
    L.scale = 1.0;
    L.obj_enabled = obj_enabled_ ? 1 : 0;
}
void Lattice::launch_stage(int stage, int in_snap, int out_snap, int gmode, bool publish) {
    upload_zone_table_if_dirty();
    tclb_launch_t L;
    fill_launch(L);
    L.snap_in = snap_[in_snap];
    L.snap_out = snap_[out_snap];
    L.globals_mode = gmode;
    L.halo_mode = halo_mode;
    epoch++;
    L.sig_epoch = epoch;
    L.sig_enable = (sig_enable && halo_mode == TCLB_HALO_DIRECT) ? 1 : 0;
    L.sig_count = sig_ + 9;
    if (capturing) { L.dyn = dyn_dev_; L.dyn_iter_off = cap_iter_off; L.dyn_epoch_off = epoch - cap_epoch_base; }
    L.gl_counter = gl_counter_;
    L.gl_partial = gl_partial_; L.gl_partial_stride = gl_partial_stride_;
    L.gl_epoch = epoch;
    L.gl_publish = (gmode != TCLB_GLOBALS_NONE && publish) ? 1 : 0;
    if (gl_complete) {
        L.gl_rank = rank; L.gl_nranks = (int)peers.size();
        for (size_t p = 0; p < peers.size(); p++) L.gl_ring[p] = peers[p].gl_ring;
    } else {
        L.gl_rank = 0; L.gl_nranks = 1;
        L.gl_ring[0] = gl_ring_;
    }
    if (L.gl_publish) last_glob_epoch = epoch;
    for (int d = 0; d < TCLB_MAX_NEIGH; d++) {
        if (d == 4) continue;
        L.halo_snap[d] = halo[d].snap[out_snap];
        L.halo_ny[d] = halo[d].ny; L.halo_nz[d] = halo[d].nz;
        // flags only towards neighbours whose kernels can overlap mine (another device or process); a neighbour on
        // this device shares my stream, never signals, and must not be waited for
        if (L.sig_enable && halo[d].concurrent) {
            L.sig_remote[d] = halo[d].flag;
            L.sig_wait[d] = sig_ + d;
        }
    }
    if (backend == TCLB_BACKEND_CUDA) TCLB_CUDA(cudaSetDevice(device));
    // ---- observability / debugging switches (TCLB_DEBUG=sync,poison  TCLB_PROFILE=1) ----
    const DebugFlags &dbg = debug_flags();
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    const bool timed = dbg.profile && backend == TCLB_BACKEND_CUDA && !capturing;
    if (timed) {
        TCLB_CUDA(cudaEventCreate(&ev0)); TCLB_CUDA(cudaEventCreate(&ev1));
        TCLB_CUDA(cudaEventRecord(ev0, (cudaStream_t)stream));
    }
    nvtxRangePushA(model->d->stages[stage].name);
    const int rc = model->d->run_stage(stage, &L);
    nvtxRangePop();
    if (rc != 0) fail("stage %s failed: %s", model->d->stages[stage].name, model->d->last_error());
    if (timed) {
        TCLB_CUDA(cudaEventRecord(ev1, (cudaStream_t)stream));
        TCLB_CUDA(cudaEventSynchronize(ev1));
        float ms = 0;
        TCLB_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
        cudaEventDestroy(ev0); cudaEventDestroy(ev1);
        StageStat &s = stage_stats[{stage, gmode}];
        s.launches++; s.ms += ms;
    }
    if (dbg.sync && backend == TCLB_BACKEND_CUDA && !capturing) {
        // like the reference's --enable-debug CROSS_SYNC: every launch is followed by a device sync and an error check
        cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) fail("stage %s: %s", model->d->stages[stage].name, cudaGetErrorString(e));
    }
}
void Lattice::clear_globals() {
    const int ng = model->d->n_globals;
    std::vector<double> z((size_t)TCLB_GL_SLOTS * std::max(1, ng), 0.0);
    for (int s = 0; s < TCLB_GL_SLOTS; s++)
        for (int i = 0; i < ng; i++) z[(size_t)s * ng + i] = model->d->globals[i].op == TCLB_OP_MAX ? -1e300 : 0.0;
    to_dev(globals_dev_, z.data(), z.size() * sizeof(double));
}
void Lattice::fetch_globals() {
    // totals of every publishing rank sit in ring slot [epoch % RING]; a stamp equal to the epoch says the
    // slot is complete.  Ranks run in lock step within a launch or two, so the loop below normally passes once.
    const int ng = model->d->n_globals, row = ng + 1;
    const int nr = gl_complete ? (int)peers.size() : 1;
    const unsigned e = last_glob_epoch;
    std::vector<double> slot_v;
    double *slot = pinned_ ? (double *)((char *)pinned_ + pinned_bytes_ / 2) : nullptr;
    if (!slot || sizeof(double) * (size_t)TCLB_MAX_RANKS * row > pinned_bytes_ / 2) { slot_v.resize((size_t)TCLB_MAX_RANKS * row); slot = slot_v.data(); }
    const double *src = gl_ring_ + (size_t)(e % TCLB_GL_RING) * TCLB_MAX_RANKS * row;
    double t0 = walltime();
    last_d2h_bytes = sizeof(double) * (size_t)nr * row;
    while (true) {
        to_host(slot, src, sizeof(double) * (size_t)nr * row);
        bool ok = true;
        for (int p = 0; p < nr; p++) if (slot[(size_t)p * row + ng] != (double)e) ok = false;
        if (ok) break;
        if (walltime() - t0 > 30.0) fail("rank %d: Globals of launch %u never arrived from every rank (a peer stopped?)", rank, e);
    }
    for (int i = 0; i < ng; i++) {
        const bool mx = model->d->globals[i].op == TCLB_OP_MAX;
        double v = mx ? model->d->globals[i].base : 0.0;
        for (int p = 0; p < nr; p++) v = mx ? std::max(v, slot[(size_t)p * row + i]) : v + slot[(size_t)p * row + i];
        globals[i] = v;
    }
}
void Lattice::get_quantity(int q, const Box &b, void *dst, double scale, bool dst_on_device) {
    upload_zone_table_if_dirty();
    const int nc = model->d->quantities[q].vector ? 3 : 1;
    size_t bytes = (size_t)b.size() * nc * model->real_size();
    tclb_launch_t L;
    fill_launch(L);
    L.snap_in = snap_[cur]; L.snap_out = snap_[cur];
    L.rx = b.x0; L.ry = b.y0; L.rz = b.z0; L.rnx = b.nx; L.rny = b.ny; L.rnz = b.nz;
    L.scale = scale;
    bool direct = (backend == TCLB_BACKEND_CPU) || dst_on_device;
    L.out = direct ? dst : scratch(bytes);
    if (backend == TCLB_BACKEND_CUDA) TCLB_CUDA(cudaSetDevice(device));
    if (model->d->get_quantity(q, &L) != 0) fail("quantity %s failed: %s", model->d->quantities[q].name, model->d->last_error());
    if (!direct) to_host(dst, L.out, bytes);
}
void Lattice::color(int zplane, float scale, unsigned char *rgba, bool dst_on_device) {
    upload_zone_table_if_dirty();
    size_t bytes = (size_t)nx * ny * 4;
    tclb_launch_t L;
    fill_launch(L);
    L.snap_in = snap_[cur]; L.snap_out = snap_[cur];
    L.rx = 0; L.ry = 0; L.rz = zplane; L.rnx = nx; L.rny = ny; L.rnz = 1;
    L.scale = scale;
    bool direct = (backend == TCLB_BACKEND_CPU) || dst_on_device;
    L.out = direct ? (void *)rgba : scratch(bytes);
    if (backend == TCLB_BACKEND_CUDA) TCLB_CUDA(cudaSetDevice(device));
    if (model->d->color(&L) != 0) fail("colour kernel failed: %s", model->d->last_error());
    if (!direct) to_host(rgba, L.out, bytes);
}
bool Lattice::has_nan(int q, const Box &b) {
    const int nc = model->d->quantities[q].vector ? 3 : 1;
    size_t n = (size_t)b.size() * nc;
    if (backend == TCLB_BACKEND_CUDA) {
        // quantity -> device staging -> device scan: only 24 bytes come back (reference copies the whole field)
        const size_t bytes = n * model->real_size(), off = (bytes + 255) & ~(size_t)255;
        char *buf = (char *)scratch(off + 64);
        get_quantity(q, b, buf, 1.0, true);
        if (tclb_scan_array(buf, (long long)n, model->real_size(), stream, (double *)(buf + off)) != 0)
            fail("device scan failed: %s", tclb_scan_last_error());
        double r[3];
        to_host(r, buf + off, sizeof(r));
        return r[0] > 0;
    }
    std::vector<char> buf(n * model->real_size());
    get_quantity(q, b, buf.data(), 1.0);
    if (model->real_size() == 4) { const float *p = (const float *)buf.data(); for (size_t i = 0; i < n; i++) if (!(p[i] == p[i])) return true; }
    else { const double *p = (const double *)buf.data(); for (size_t i = 0; i < n; i++) if (!(p[i] == p[i])) return true; }
    return false;
}
void Lattice::save_raw(const std::string &file) const {
    std::vector<char> buf(snap_bytes);
    to_host(buf.data(), snap_[cur], snap_bytes);
    FILE *f = fopen(file.c_str(), "wb");
    if (!f) fail("cannot write %s", file.c_str());
    fwrite(buf.data(), 1, buf.size(), f);
    fclose(f);
}
void Lattice::load_raw(const std::string &file) {
    std::vector<char> buf(snap_bytes);
    FILE *f = fopen(file.c_str(), "rb");
    if (!f) fail("cannot read %s", file.c_str());
    size_t n = fread(buf.data(), 1, buf.size(), f);
    fclose(f);
    if (n != buf.size()) fail("%s: expected %zu bytes, got %zu", file.c_str(), buf.size(), n);
    to_dev(snap_[cur], buf.data(), snap_bytes);
}

// ------------------------------------------------------------------------------------------
Decomposition Decomposition::make(int nx, int ny, int nz, int nranks) {
    Decomposition D;
    D.nranks = nranks;
    D.tot[0] = nx; D.tot[1] = ny; D.tot[2] = nz;
    D.boxes.assign(nranks, Box());
    float mincom = (1.0f + nranks) * (ny + nz);
    bool done = false;
    for (int divz = 1; divz <= nranks; divz++) {
        if (nranks % divz) continue;
        int divy = nranks / divz;
        float com = (float)divz * ny + (float)divy * nz;
        if (com < mincom) {
            mincom = com;
            if (nz >= divz && ny >= divy) {
                std::vector<int> yl(divy), zl(divz);
                int my = ny, mz = nz;
                for (int i = 0; i < divy; i++) { yl[i] = my / (divy - i); my -= yl[i]; }
                for (int i = 0; i < divz; i++) { zl[i] = mz / (divz - i); mz -= zl[i]; }
                int k = 0, dz = 0;
                for (int i = 0; i < divz; i++) {
                    int dy = 0;
                    for (int j = 0; j < divy; j++) {
                        Box &b = D.boxes[k++];
                        b.x0 = 0; b.nx = nx; b.y0 = dy; b.ny = yl[j]; b.z0 = dz; b.nz = zl[i];
                        dy += yl[j];
                    }
                    dz += zl[i];
                }
                D.divy = divy; D.divz = divz;
                done = true;
            }
        }
    }
    if (!done) fail("lattice %dx%dx%d is too small to be divided over %d ranks", nx, ny, nz, nranks);
    return D;
}
int Decomposition::neighbour(int rank, int dy, int dz) const {
    int iy = rank % divy, iz = rank / divy;
    iy = (iy + dy + divy) % divy;
    iz = (iz + dz + divz) % divz;
    return iy + divy * iz;
}

// ------------------------------------------------------------------------------------------
Domain::Domain(Model *m, int backend_, int nx, int ny, int nz, int nranks, const std::vector<int> &local_ranks,
               const std::vector<int> &devices)
    : model(m), backend(backend_), iter(0), use_direct_halo(true) {
    dec = Decomposition::make(nx, ny, nz, nranks);
    lat.assign(nranks, nullptr);
    local = local_ranks;
    for (size_t i = 0; i < local.size(); i++) {
        int r = local[i];
        int dev = devices.empty() ? 0 : devices[i % devices.size()];
        void *shared = nullptr;
        for (size_t k = 0; k < i; k++) if (lat[local[k]]->device == dev) shared = lat[local[k]]->stream;
        lat[r] = new Lattice(m, backend, dev, r, dec.boxes[r], dec.tot, shared);
    }
    globals.assign(std::max(1, m->d->n_globals), 0.0);
    // defaults of the settings are applied by the Solver (needs the unit system)
    connect_local();
    finalize_wiring();
}
Domain::~Domain() {
    try { sync(); } catch (...) {}
    drop_graphs();
    for (size_t i = lat.size(); i-- > 0;) delete lat[i];     // stream owners (first on a device) go last
}
static int opposite(int d) {
    int dy = d % 3 - 1, dz = d / 3 - 1;
    return (-dy + 1) + 3 * (-dz + 1);
}
void Domain::connect_local() {
    // peer access between the devices of local ranks
    if (backend == TCLB_BACKEND_CUDA) {
        for (int a : local)
            for (int b : local) {
                int da = lat[a]->device, db = lat[b]->device;
                if (da == db) continue;
                int can = 0;
                TCLB_CUDA(cudaDeviceCanAccessPeer(&can, da, db));
                if (can) {
                    TCLB_CUDA(cudaSetDevice(da));
                    cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) TCLB_CUDA(e);
                    cudaGetLastError();
                } else logf(L_WARNING, "device %d cannot access device %d directly; direct halo needs peer access", da, db);
            }
    }
    for (int r : local) {
        lat[r]->peers.assign(dec.nranks, PeerInfo());
        for (int q : local) lat[r]->peers[q] = lat[q]->self_info();
    }
}
void Domain::set_peer(int r, int peer, const PeerInfo &p) {
    if (!lat[r]) fail("set_peer: rank %d is not local", r);
    lat[r]->peers[peer] = p;
}
void Domain::finalize_wiring() {
    for (int r : local) {
        Lattice *L = lat[r];
        bool all_wired = true, concurrent = false, all_peers = true;
        for (int q = 0; q < dec.nranks; q++) if (!L->peers[q].valid) all_peers = false;
        for (int d = 0; d < TCLB_MAX_NEIGH; d++) {
            if (d == 4) continue;
            int dy = d % 3 - 1, dz = d / 3 - 1;
            int nb = dec.neighbour(r, dy, dz);
            HaloTarget &t = L->halo[d];
            t = HaloTarget();
            t.rank = nb;
            t.ny = dec.boxes[nb].ny; t.nz = dec.boxes[nb].nz;
            const PeerInfo &p = L->peers[nb];
            if (p.valid) {
                t.snap[0] = p.snap[0]; t.snap[1] = p.snap[1];
                t.flag = p.flags_in + opposite(d);
            } else all_wired = false;
            // sequence flags are only needed between kernels that may run concurrently (different devices or
            // processes); ranks sharing one stream-ordered device do not need them, and the CPU oracle runs
            // ranks one after another.
            t.concurrent = backend == TCLB_BACKEND_CUDA && nb != r && (!lat[nb] || lat[nb]->device != L->device);
            if (t.concurrent) concurrent = true;
        }
        L->sig_enable = concurrent && all_wired;
        L->halo_mode = (use_direct_halo && all_wired) ? TCLB_HALO_DIRECT : TCLB_HALO_NONE;
        L->gl_complete = all_peers && dec.nranks <= TCLB_MAX_RANKS;
    }
}
void Domain::set_setting(int idx, double v) { for (int r : local) lat[r]->set_setting(idx, v); }
void Domain::set_zone_setting(int s, int zone, double v) { for (int r : local) lat[r]->zs.set(s, zone, v); }
void Domain::set_zone_series(int s, int zone, const std::vector<double> &v) { for (int r : local) lat[r]->zs.set_series(s, zone, v); }
void Domain::set_zones(int n) { for (int r : local) lat[r]->zs.set_zones(n); }
void Domain::sync() { for (int r : local) lat[r]->sync(); }

void Domain::copy_exchange(int s) {
    // Baseline halo path: copy the border rows of every local rank into the ghost rows of its
    // neighbours with stream-ordered memcpys (plays the role of the reference's
    // D2H -> MPI_Isend/Irecv -> H2D, src/Lattice.cu.Rt:327-389, without the host bounce).
    const int rs = model->storage_size();
    const int nf = model->d->n_fields;
    if (backend == TCLB_BACKEND_CUDA) sync();
    for (int r : local) {
        Lattice *A = lat[r];
        for (int d = 0; d < TCLB_MAX_NEIGH; d++) {
            if (d == 4) continue;
            int dy = d % 3 - 1, dz = d / 3 - 1;
            if ((dy != 0 && A->gy == 0) || (dz != 0 && A->gz == 0)) continue;
            // the neighbour's snapshot: a local object, or (one process per GPU, direct wiring) its mapped peer memory;
            // a neighbour reachable only through the exchange hook is left to the hook
            const HaloTarget &T = A->halo[d];
            if (!T.snap[s]) continue;
            const int Bny = T.ny, Bnz = T.nz;
            const size_t BNY = (size_t)Bny + 2 * A->gy, BNZ = (size_t)Bnz + 2 * A->gz;
            // rows of A that B needs: y in [ny-gy, ny) for dy=+1, [0,gy) for dy=-1, all for 0
            int ys = dy > 0 ? A->ny - A->gy : 0, yn = dy != 0 ? A->gy : A->ny;
            int zs_ = dz > 0 ? A->nz - A->gz : 0, zn = dz != 0 ? A->gz : A->nz;
            for (int f = 0; f < nf; f++)
                for (int z = 0; z < zn; z++) {
                    int za = zs_ + z;
                    int zb = dz > 0 ? za - A->nz : (dz < 0 ? za + Bnz : za);
                    for (int y = 0; y < yn; y++) {
                        int ya = ys + y;
                        int yb = dy > 0 ? ya - A->ny : (dy < 0 ? ya + Bny : ya);
                        const char *src = (const char *)A->field_ptr(s, f) + (((size_t)(za + A->gz) * A->NY + (ya + A->gy)) * A->nx) * rs;
                        char *dst = (char *)T.snap[s] + ((size_t)f * (BNY * BNZ * A->nx) + ((size_t)(zb + A->gz) * BNY + (yb + A->gy)) * A->nx) * rs;
                        if (backend == TCLB_BACKEND_CUDA) {
                            if (dy == 0) {   // whole z-plane rows are contiguous in y: one copy per plane
                                if (y == 0) TCLB_CUDA(cudaMemcpyAsync(dst, src, (size_t)yn * A->nx * rs, cudaMemcpyDefault, (cudaStream_t)A->stream));
                            } else TCLB_CUDA(cudaMemcpyAsync(dst, src, (size_t)A->nx * rs, cudaMemcpyDefault, (cudaStream_t)A->stream));
                        } else memcpy(dst, src, (size_t)A->nx * rs);
                    }
                }
        }
    }
    if (backend == TCLB_BACKEND_CUDA) sync();
}

void Domain::run_action(int action, int gmode) {
    const tclb_action_t &a = model->d->actions[action];
    int tab0 = lat[local[0]]->cur, tab1 = tab0 ^ 1;
    int first_particle = -1, last_particle = -1;
    for (int s = 0; s < a.n_stages; s++)
        if (model->d->stages[a.stages[s]].particle) { if (first_particle < 0) first_particle = s; last_particle = s; }
    for (int s = 0; s < a.n_stages; s++) {
        int in = s == 0 ? tab0 : tab1;
        if (s == first_particle && particles_in_hook) particles_in_hook();
        // debug: NaN in the ghosts that this stage is about to fill (only safe when every writer runs after this point:
        // all subdomains in this process on one stream, and the stage does not work in place)
        if (debug_flags().poison && in != tab1 && (int)local.size() == dec.nranks && !lat[local[0]]->capturing) {
            bool one_stream = true;
            for (int r : local) one_stream = one_stream && (backend != TCLB_BACKEND_CUDA || lat[r]->stream == lat[local[0]]->stream);
            if (one_stream) for (int r : local) lat[r]->poison_ghosts(tab1, a.stages[s]);
        }
        for (int r : local) lat[r]->launch_stage(a.stages[s], in, tab1, gmode, s == a.n_stages - 1);
        if (s == last_particle && particles_out_hook) particles_out_hook();
        bool need_copy = false;
        for (int r : local) if (lat[r]->halo_mode == TCLB_HALO_NONE) need_copy = true;
        if (need_copy) {
            copy_exchange(tab1);
            if (exchange_hook) exchange_hook(tab1);
        }
    }
    for (int r : local) lat[r]->cur = tab1;
}
void Domain::init() {
    int a = model->find_action("Init");
    if (a < 0) fail("model has no Init action");
    for (int r : local) { lat[r]->iter = 0; lat[r]->reset_iter = 0; }
    iter = 0;
    run_action(a, TCLB_GLOBALS_NONE);
}
// ---- the iteration loop as a CUDA graph ------------------------------------------------------
extern "C" int tclb_dyn_set(int *dyn, int iter, unsigned epoch, void *stream);
extern "C" int tclb_dyn_advance(int *dyn, int diter, unsigned depoch, void *stream);

unsigned long Domain::graph_key() const {
    unsigned long k = 1469598103934665603ul;
    for (int r : local) {
        const Lattice *l = lat[r];
        for (unsigned long v : {l->graph_version, (unsigned long)l->reset_iter, (unsigned long)l->block_x, (unsigned long)l->block_y,
                                (unsigned long)l->halo_mode, (unsigned long)l->sig_enable, (unsigned long)(size_t)l->stream})
            k = (k ^ v) * 1099511628211ul;
    }
    return k;
}
bool Domain::graph_possible(int action) const {
    if (!use_graphs || backend != TCLB_BACKEND_CUDA || local.empty() || getenv("TCLB_NO_GRAPH")) return false;
    if (debug_flags().profile || debug_flags().sync || debug_flags().poison) return false;   // per-launch instrumentation
    const tclb_action_t &a = model->d->actions[action];
    for (int s = 0; s < a.n_stages; s++) if (model->d->stages[a.stages[s]].particle) return false;
    // launch latency only shows on small lattices; on large ones the graph variant of the kernels (which reads its
    // counters from memory) is a few percent slower than the plain one - measured - so the plain loop stays
    long long nodes = 0;
    for (int r : local) nodes += (long long)lat[r]->nx * lat[r]->ny * lat[r]->nz;
    if (nodes > (1ll << 22) && !getenv("TCLB_FORCE_GRAPH")) return false;
    for (int r : local) {
        const Lattice *l = lat[r];
        if (l->halo_mode != TCLB_HALO_DIRECT) return false;                    // the copy/NCCL exchange runs on the host
        if (l->device != lat[local[0]]->device || l->stream != lat[local[0]]->stream) return false;   // one stream to capture
    }
    return true;
}
void Domain::drop_graphs() {
    for (auto &g : graph_) {
        if (g.exec) cudaGraphExecDestroy((cudaGraphExec_t)g.exec);
        g = GraphSlot();
    }
}
void Domain::run_graph_pairs(int action, int pairs) {
    Lattice *l0 = lat[local[0]];
    cudaStream_t st = (cudaStream_t)l0->stream;
    TCLB_CUDA(cudaSetDevice(l0->device));
    for (int r : local) lat[r]->upload_zone_table_if_dirty();
    const int o = l0->cur & 1;
    const unsigned long key = graph_key();
    GraphSlot &g = graph_[o];
    if (!g.exec || g.key != key || g.action != action) {
        if (g.exec) { cudaGraphExecDestroy((cudaGraphExec_t)g.exec); g.exec = nullptr; }
        // capture two iterations (snapshot 0->1 and 1->0); iteration and epoch are read from device memory + offsets
        const int iter0 = iter;
        std::vector<unsigned> epoch0;
        for (int r : local) { epoch0.push_back(lat[r]->epoch); lat[r]->capturing = true; lat[r]->cap_epoch_base = lat[r]->epoch; }
        cudaGraph_t graph = nullptr;
        TCLB_CUDA(cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
        try {
            for (int k = 0; k < 2; k++) {
                for (int r : local) lat[r]->cap_iter_off = k;
                run_action(action, TCLB_GLOBALS_NONE);
            }
            g.launches = lat[local[0]]->epoch - epoch0[0];
            for (int r : local)
                if (tclb_dyn_advance(lat[r]->dyn_dev(), 2, lat[r]->epoch - lat[r]->cap_epoch_base, st) != 0) fail("graph capture: advance kernel failed");
        } catch (...) {
            cudaStreamEndCapture(st, &graph);
            if (graph) cudaGraphDestroy(graph);
            for (size_t i = 0; i < local.size(); i++) { lat[local[i]]->capturing = false; lat[local[i]]->epoch = epoch0[i]; }
            throw;
        }
        TCLB_CUDA(cudaStreamEndCapture(st, &graph));
        cudaGraphExec_t exec = nullptr;
        TCLB_CUDA(cudaGraphInstantiate(&exec, graph, 0));
        cudaGraphDestroy(graph);
        // nothing ran during the capture: put the host bookkeeping back
        for (size_t i = 0; i < local.size(); i++) { lat[local[i]]->capturing = false; lat[local[i]]->epoch = epoch0[i]; lat[local[i]]->iter = iter0; }
        iter = iter0;
        g.exec = exec; g.key = key; g.action = action;
    }
    for (int r : local)
        if (tclb_dyn_set(lat[r]->dyn_dev(), iter, lat[r]->epoch, st) != 0) fail("graph replay: cannot set the iteration counters");
    for (int p = 0; p < pairs; p++) TCLB_CUDA(cudaGraphLaunch((cudaGraphExec_t)g.exec, st));
    iter += 2 * pairs;
    for (int r : local) { lat[r]->iter = iter; lat[r]->epoch += g.launches * (unsigned)pairs; }
    graph_replays += 2LL * pairs;
}

void Domain::iterate(int n, bool globals_on_last, bool globals_every) {
    int a = model->find_action("Iteration");
    // iterations that need nothing from the host go through the captured graph, two at a time
    if (!globals_every && n >= 8 && graph_possible(a)) {
        const int plain = globals_on_last ? n - 1 : n;
        const int pairs = plain / 2;
        if (pairs > 0) { run_graph_pairs(a, pairs); n -= 2 * pairs; }
    }
    for (int i = 0; i < n; i++) {
        bool g = globals_every || (globals_on_last && i == n - 1);
        run_action(a, g ? TCLB_GLOBALS_INTEGRATE : TCLB_GLOBALS_NONE);
        iter++;
        for (int r : local) lat[r]->iter = iter;
    }
    if (globals_on_last || globals_every) reduce_globals();
}
void Domain::reduce_globals() {
    const int ng = model->d->n_globals;
    bool complete = true;
    for (int r : local) { lat[r]->fetch_globals(); if (!lat[r]->gl_complete) complete = false; }
    if (complete) {            // every rank already holds the totals of all ranks
        for (int i = 0; i < ng; i++) globals[i] = lat[local[0]]->globals[i];
        return;
    }
    for (int i = 0; i < ng; i++) globals[i] = model->d->globals[i].op == TCLB_OP_MAX ? model->d->globals[i].base : 0.0;
    for (int r : local)
        for (int i = 0; i < ng; i++) {
            if (model->d->globals[i].op == TCLB_OP_MAX) globals[i] = std::max(globals[i], lat[r]->globals[i]);
            else globals[i] += lat[r]->globals[i];
        }
    if (reduce_hook) {
        const int ns = model->d->sum_globals;
        reduce_hook(globals.data(), ns, TCLB_OP_SUM);
        if (ng > ns) reduce_hook(globals.data() + ns, ng - ns, TCLB_OP_MAX);
    }
}
}  // namespace tclb
