#include "adjoint.h"
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include "common.h"

extern "C" int tclb_adjoint_fold_ghosts(void *adj, int real_size, int on_device, int nfields, int nx, int ny, int nz, int gy, int gz,
                                        void *stream);

namespace tclb {

#define ADJ_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) fail("adjoint: %s", cudaGetErrorString(e__)); } while (0)

Adjoint::Adjoint(Domain *dom) : dom_(dom), lat_(nullptr) {
    if (dom->dec.nranks != 1 || dom->local.size() != 1)
        fail("adjoint: this version handles one subdomain (got %d ranks)", dom->dec.nranks);
    lat_ = dom->lat[dom->local[0]];
    Model *m = dom->model;
    if (m->storage_size() != m->real_size()) fail("adjoint: needs plain float/double lattice storage");
    action_ = m->find_action("Iteration");
    if (action_ < 0) fail("adjoint: model has no Iteration action");
    const tclb_action_t &a = m->d->actions[action_];
    if (a.n_stages != 1) fail("adjoint: action Iteration of %s has %d stages; only single-stage actions are differentiated", m->d->name, a.n_stages);
    stage_ = a.stages[0];
    for (int f = 0; f < m->d->n_fields; f++)
        if (!m->d->stages[stage_].save_field[f]) fail("adjoint: stage %s does not save field %s", m->d->stages[stage_].name, m->d->fields[f].name);
    // fields that the node code reads through a stencil accessor (phi(1,0) ...) have more cells in their declared reach
    // than densities pulling from them; those reads are constants for this adjoint: say so instead of being silently wrong
    for (int f = 0; f < m->d->n_fields; f++) {
        const tclb_field_t &F = m->d->fields[f];
        const long cells = (long)(F.maxx - F.minx + 1) * (F.maxy - F.miny + 1) * (F.maxz - F.minz + 1);
        long pulls = 0;
        for (int k = 0; k < m->d->n_densities; k++) if (m->d->densities[k].field_index == f) pulls++;
        if (cells > std::max(1L, pulls))
            logf(L_WARNING, "adjoint: field %s is read through a stencil accessor (%ld cells, %ld densities); those reads are treated "
                            "as constants, gradients through them are missing", F.name, cells, pulls);
    }
    // companion library: <primal>.so -> <primal>_ad.so
    std::string p = m->path;
    const size_t dot = p.rfind(".so");
    if (dot == std::string::npos) fail("adjoint: cannot derive the adjoint library name from %s", p.c_str());
    p = p.substr(0, dot) + "_ad.so";
    lib_ = dlopen(p.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!lib_) fail("adjoint: cannot load %s (%s); build it with tclb_b200.build.build_adjoint_lib", p.c_str(), dlerror());
    fn_ = (stage_fn)dlsym(lib_, "tclb_model_adjoint_stage");
    err_ = (err_fn)dlsym(lib_, "tclb_model_adjoint_error");
    typedef int (*info_fn)(int *, int *, int *);
    info_fn info = (info_fn)dlsym(lib_, "tclb_model_adjoint_info");
    if (!fn_ || !err_ || !info) fail("adjoint: %s does not export the adjoint entry points", p.c_str());
    int rs = 0, dirs = 0, nf = 0;
    if (info(&rs, &dirs, &nf) != TCLB_ABI_VERSION) fail("adjoint: %s was built for another ABI version (rebuild it)", p.c_str());
    if (rs != m->real_size() || nf != m->d->n_fields) fail("adjoint: %s does not match the primal library", p.c_str());
    rs_ = rs;
    dirs_ = dirs;
    sfn_ = (settings_fn)dlsym(lib_, "tclb_model_adjoint_settings");
    zfn_ = (zsettings_fn)dlsym(lib_, "tclb_model_adjoint_zsettings");
    bytes_ = (size_t)m->d->n_fields * lat_->plane_elems * rs_;
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    for (int i = 0; i < 2; i++) { adj_[i] = alloc(bytes_, dev); fill_zero(adj_[i]); }
    // tape budget on the device: TCLB_TAPE_DEVICE_BYTES, else 70 % of what is free now (a B200 has 180 GB); frames
    // beyond it go to page-locked host memory
    const char *cap = getenv("TCLB_TAPE_DEVICE_BYTES");
    if (cap) dev_tape_cap_ = (size_t)atoll(cap);
    else if (dev) { size_t fr = 0, tot = 0; ADJ_CUDA(cudaMemGetInfo(&fr, &tot)); dev_tape_cap_ = (size_t)(0.7 * (double)fr); }
    else dev_tape_cap_ = 0;
}
Adjoint::~Adjoint() {
    drop_tape();
    for (void *p : pool_dev_) release(p, true);
    for (void *p : pool_host_) release(p, false);
    const bool dev = lat_ && lat_->backend == TCLB_BACKEND_CUDA;
    for (int i = 0; i < 2; i++) if (adj_[i]) release(adj_[i], dev);
    if (sg_dev_) cudaFree(sg_dev_);
    if (zg_dev_) cudaFree(zg_dev_);
    if (lib_) dlclose(lib_);
}
void *Adjoint::alloc(size_t bytes, bool device) const {
    void *p = nullptr;
    const bool cuda = lat_->backend == TCLB_BACKEND_CUDA;
    if (device) { ADJ_CUDA(cudaSetDevice(lat_->device)); ADJ_CUDA(cudaMalloc(&p, bytes)); }
    else if (cuda) ADJ_CUDA(cudaMallocHost(&p, bytes));     // page-locked: the tape copies run at full PCIe/C2C rate
    else { p = malloc(bytes); if (!p) fail("adjoint: out of host memory (%zu bytes)", bytes); }
    return p;
}
void Adjoint::release(void *p, bool device) const {
    if (device) cudaFree(p);
    else if (lat_->backend == TCLB_BACKEND_CUDA) cudaFreeHost(p);
    else free(p);
}
// tape frames are recycled: a reverse sweep hands its frames to the next record() instead of freeing them
void *Adjoint::frame_get(bool device) {
    std::vector<void *> &pool = device ? pool_dev_ : pool_host_;
    if (!pool.empty()) { void *p = pool.back(); pool.pop_back(); return p; }
    return alloc(lat_->snap_bytes, device);
}
void Adjoint::frame_put(void *p, bool device) { (device ? pool_dev_ : pool_host_).push_back(p); }
void Adjoint::fill_zero(void *p) const {
    if (lat_->backend == TCLB_BACKEND_CUDA) ADJ_CUDA(cudaMemsetAsync(p, 0, bytes_, (cudaStream_t)lat_->stream));
    else memset(p, 0, bytes_);
}
void Adjoint::zero() { fill_zero(adj_[0]); fill_zero(adj_[1]); }
void Adjoint::drop_tape() {
    for (Frame &f : tape_) frame_put(f.data, f.on_device);
    tape_.clear();
    dev_tape_bytes_ = 0;
}
void Adjoint::record(int n) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    if (tape_.empty() && !zg_which_.empty()) clear_zone_gradients();
    const int obj = dom_->model->find_global("Objective");
    for (int i = 0; i < n; i++) {
        Frame f;
        f.iter = lat_->iter;
        f.on_device = dev && dev_tape_bytes_ + lat_->snap_bytes <= dev_tape_cap_;
        f.data = frame_get(f.on_device);
        const void *src = lat_->snap_ptr(lat_->cur);
        if (dev) {
            ADJ_CUDA(cudaMemcpyAsync(f.data, src, lat_->snap_bytes, f.on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                                     (cudaStream_t)lat_->stream));
            if (!f.on_device) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
            else dev_tape_bytes_ += lat_->snap_bytes;
        } else memcpy(f.data, src, lat_->snap_bytes);
        tape_.push_back(f);
        dom_->iterate(1, false, true);
        if (obj >= 0) objective += dom_->globals[obj];
    }
}
void Adjoint::step(const void *primal_in, int iter) {
    lat_->upload_zone_table_if_dirty();
    tclb_launch_t L;
    lat_->fill_launch(L);
    L.snap_in = const_cast<void *>(primal_in);
    L.snap_out = nullptr;
    L.iter = iter;
    L.zone_index = L.zone_len > 1 ? iter % L.zone_len : 0;
    L.globals_mode = TCLB_GLOBALS_NONE;
    L.halo_mode = TCLB_HALO_NONE;
    L.obj_enabled = 1;
    void *out = adj_[cur_], *in = adj_[cur_ ^ 1];
    fill_zero(in);
    if (lat_->backend == TCLB_BACKEND_CUDA) ADJ_CUDA(cudaSetDevice(lat_->device));
    if (fn_(stage_, &L, out, in) != 0) fail("adjoint stage failed: %s", err_());
    if (!sg_which_.empty()) settings_pass(L, out);
    if (!zg_which_.empty()) zsettings_pass(L, out, L.zone_index);
    if (tclb_adjoint_fold_ghosts(in, rs_, lat_->backend == TCLB_BACKEND_CUDA, dom_->model->d->n_fields, lat_->nx, lat_->ny, lat_->nz,
                                 lat_->gy, lat_->gz, lat_->stream) != 0)
        fail("adjoint: ghost fold failed");
    cur_ ^= 1;
}
void Adjoint::reverse() {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    void *stage_buf = nullptr;
    while (!tape_.empty()) {
        Frame f = tape_.back();
        tape_.pop_back();
        const void *src = f.data;
        if (dev && !f.on_device) {
            if (!stage_buf) stage_buf = alloc(lat_->snap_bytes, true);
            ADJ_CUDA(cudaMemcpyAsync(stage_buf, f.data, lat_->snap_bytes, cudaMemcpyHostToDevice, (cudaStream_t)lat_->stream));
            src = stage_buf;
        }
        step(src, f.iter);      // everything is ordered on the lattice's stream: no host synchronisation per step
        frame_put(f.data, f.on_device);
        if (f.on_device) dev_tape_bytes_ -= lat_->snap_bytes;
    }
    if (dev) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
    if (stage_buf) release(stage_buf, true);
}
void Adjoint::steady(int n) {
    for (int i = 0; i < n; i++) step(lat_->snap_ptr(lat_->cur), lat_->iter);
    if (lat_->backend == TCLB_BACKEND_CUDA) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
}
void Adjoint::want_setting_gradients(const std::vector<int> &settings) {
    if (!sfn_) fail("adjoint: the adjoint library has no settings entry point (rebuild it)");
    const tclb_model_t *d = dom_->model->d;
    sg_which_.clear();
    for (int s : settings) {
        int guard = 0;
        while (s >= 0 && s < d->n_settings && guard++ < 64) {
            if (std::find(sg_which_.begin(), sg_which_.end(), s) == sg_which_.end()) sg_which_.push_back(s);
            s = d->settings[s].derived_target;
        }
    }
    sg_sum_.assign(sg_which_.size(), 0.0);
    if (lat_->backend == TCLB_BACKEND_CUDA && !sg_which_.empty()) {
        if (sg_dev_) cudaFree(sg_dev_);
        ADJ_CUDA(cudaSetDevice(lat_->device));
        ADJ_CUDA(cudaMalloc((void **)&sg_dev_, sizeof(double) * (sg_which_.size() + dirs_)));
        ADJ_CUDA(cudaMemset(sg_dev_, 0, sizeof(double) * (sg_which_.size() + dirs_)));
    }
}
void Adjoint::settings_pass(const tclb_launch_t &L, const void *adj_out) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    for (size_t o = 0; o < sg_which_.size(); o += dirs_) {
        const int n = (int)std::min((size_t)dirs_, sg_which_.size() - o);
        double *dst = dev ? sg_dev_ + o : sg_sum_.data() + o;
        if (sfn_(stage_, &L, adj_out, sg_which_.data() + o, n, dst) != 0) fail("adjoint settings pass failed: %s", err_());
    }
}
void Adjoint::zg_resize() {
    const int len = std::max(1, lat_->zs.len());
    if (len == zg_len_ && zg_sum_.size() == (size_t)len * zg_which_.size() * 2) return;
    zg_len_ = len;
    zg_sum_.assign((size_t)len * zg_which_.size() * 2, 0.0);
    if (lat_->backend == TCLB_BACKEND_CUDA) {
        if (zg_dev_) cudaFree(zg_dev_);
        ADJ_CUDA(cudaSetDevice(lat_->device));
        ADJ_CUDA(cudaMalloc((void **)&zg_dev_, sizeof(double) * (zg_sum_.size() + dirs_)));
        ADJ_CUDA(cudaMemset(zg_dev_, 0, sizeof(double) * (zg_sum_.size() + dirs_)));
    }
}
void Adjoint::want_zone_gradient(int zonesetting, int zone) {
    if (!zfn_) fail("adjoint: the adjoint library has no zonal-settings entry point (rebuild it)");
    if (dirs_ < 2) fail("adjoint: zonal gradients need at least two tangent directions");
    for (auto &p : zg_which_) if (p.first == zonesetting && p.second == zone) return;
    zg_which_.push_back({zonesetting, zone});
    zg_len_ = 0;
    zg_resize();
}
void Adjoint::clear_zone_gradients() {
    zg_len_ = 0;
    if (!zg_which_.empty()) zg_resize();
}
void Adjoint::zsettings_pass(const tclb_launch_t &L, const void *adj_out, int time_level) {
    zg_resize();
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    const size_t per_level = zg_which_.size() * 2;
    const int pairs_per_pass = std::max(1, std::min(dirs_, 16) / 2);
    for (size_t o = 0; o < zg_which_.size(); o += pairs_per_pass) {
        const int np = (int)std::min((size_t)pairs_per_pass, zg_which_.size() - o);
        int zs[16], zone[16], dt[16];
        for (int k = 0; k < np; k++) {
            zs[2 * k] = zs[2 * k + 1] = zg_which_[o + k].first;
            zone[2 * k] = zone[2 * k + 1] = zg_which_[o + k].second;
            dt[2 * k] = 0; dt[2 * k + 1] = 1;
        }
        double *dst = (dev ? zg_dev_ : zg_sum_.data()) + (size_t)time_level * per_level + 2 * o;
        if (zfn_(stage_, &L, adj_out, zs, zone, dt, 2 * np, dst) != 0) fail("adjoint zonal-settings pass failed: %s", err_());
    }
}
void Adjoint::zone_gradient(int zonesetting, int zone, double *out, int len) {
    size_t k = 0;
    for (; k < zg_which_.size(); k++) if (zg_which_[k].first == zonesetting && zg_which_[k].second == zone) break;
    if (k == zg_which_.size()) fail("adjoint: the gradient of zonal setting %d in zone %d was not requested", zonesetting, zone);
    if (lat_->backend == TCLB_BACKEND_CUDA && zg_dev_) {
        ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
        ADJ_CUDA(cudaMemcpy(zg_sum_.data(), zg_dev_, sizeof(double) * zg_sum_.size(), cudaMemcpyDeviceToHost));
    }
    const size_t per_level = zg_which_.size() * 2;
    const int L = zg_len_;
    auto val = [&](int t) { return zg_sum_[(size_t)t * per_level + 2 * k]; };
    auto ddt = [&](int t) { return zg_sum_[(size_t)t * per_level + 2 * k + 1]; };
    if (len == 1) {               // a constant: every time level reads the same value, its central difference is zero
        double s = 0;
        for (int t = 0; t < L; t++) s += val(t);
        out[0] = s;
        return;
    }
    if (len != L) fail("adjoint: zone_gradient asked for %d time levels, the series has %d", len, L);
    for (int t = 0; t < L; t++) out[t] = val(t);
    // DT(t) = (v(t+1) - v(t-1)) / 2 with periodic wrap: transpose it
    for (int t = 0; t < L; t++) {
        const int ip = t + 1 < L ? t + 1 : 0, im = t > 0 ? t - 1 : L - 1;
        out[ip] += 0.5 * ddt(t);
        out[im] -= 0.5 * ddt(t);
    }
}
double Adjoint::setting_gradient(int setting) {
    const tclb_model_t *d = dom_->model->d;
    if (lat_->backend == TCLB_BACKEND_CUDA && sg_dev_) {
        ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
        ADJ_CUDA(cudaMemcpy(sg_sum_.data(), sg_dev_, sizeof(double) * sg_sum_.size(), cudaMemcpyDeviceToHost));
    }
    // total derivative along the chain of derived settings: dJ/dv_i = partial_i + f_i'(v_i) * dJ/dv_target
    double total = 0, factor = 1;
    int s = setting, guard = 0;
    while (s >= 0 && s < d->n_settings && guard++ < 64) {
        auto it = std::find(sg_which_.begin(), sg_which_.end(), s);
        if (it == sg_which_.end()) fail("adjoint: gradient of setting %s was not requested", d->settings[s].name);
        total += factor * sg_sum_[it - sg_which_.begin()];
        const tclb_setting_t &S = d->settings[s];
        if (S.derived_target < 0 || !S.derive) break;
        const double v = lat_->get_setting(s), h = std::max(1e-7, std::fabs(v) * 1e-6);
        factor *= (S.derive(v + h) - S.derive(v - h)) / (2 * h);
        s = S.derived_target;
    }
    return total;
}
// interior <-> dense copies (adjoint buffers have the snapshot layout [field][NZ][NY][nx] in real_t)
void Adjoint::get_field(int field, void *dst) const {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    const size_t row = (size_t)lat_->nx * rs_;
    const char *base = (const char *)adj_[cur_] + (size_t)field * lat_->plane_elems * rs_;
    if (dev) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
    for (int z = 0; z < lat_->nz; z++) {
        const char *src = base + ((size_t)(z + lat_->gz) * lat_->NY + lat_->gy) * row;
        char *d = (char *)dst + (size_t)z * lat_->ny * row;
        if (dev) ADJ_CUDA(cudaMemcpy2D(d, row, src, row, row, lat_->ny, cudaMemcpyDeviceToHost));
        else for (int y = 0; y < lat_->ny; y++) memcpy(d + y * row, src + y * row, row);
    }
}
void Adjoint::set_field(int field, const void *srcv) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    const size_t row = (size_t)lat_->nx * rs_;
    char *base = (char *)adj_[cur_] + (size_t)field * lat_->plane_elems * rs_;
    if (dev) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)lat_->stream));
    for (int z = 0; z < lat_->nz; z++) {
        char *d = base + ((size_t)(z + lat_->gz) * lat_->NY + lat_->gy) * row;
        const char *s = (const char *)srcv + (size_t)z * lat_->ny * row;
        if (dev) ADJ_CUDA(cudaMemcpy2D(d, row, s, row, row, lat_->ny, cudaMemcpyHostToDevice));
        else for (int y = 0; y < lat_->ny; y++) memcpy(d + y * row, s + y * row, row);
    }
}
double Adjoint::dot_field(int field, const void *dir) const {
    const size_t n = (size_t)lat_->nx * lat_->ny * lat_->nz;
    std::vector<char> tmp(n * rs_);
    get_field(field, tmp.data());
    double s = 0;
    if (rs_ == 8) { const double *a = (const double *)tmp.data(), *b = (const double *)dir; for (size_t i = 0; i < n; i++) s += a[i] * b[i]; }
    else { const float *a = (const float *)tmp.data(), *b = (const float *)dir; for (size_t i = 0; i < n; i++) s += (double)a[i] * b[i]; }
    return s;
}

}  // namespace tclb
