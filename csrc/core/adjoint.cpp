#include "adjoint.h"
#include <unistd.h>
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include "common.h"

extern "C" int tclb_adjoint_fold_ghosts(void *adj, int real_size, int on_device, int nfields, int nx, int ny, int nz, int gy, int gz,
                                        void *stream);
extern "C" int tclb_adjoint_fold_pair(void *dst, int dny, int dnz, void *src, int sny, int snz, int real_size, int on_device, int nfields,
                                      int nx, int gy, int gz, int dy, int dz, void *stream);
extern "C" int tclb_adjoint_add_into(void *dst, const void *src, long long n, int real_size, int on_device, void *stream);

namespace tclb {

#define ADJ_CUDA(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) fail("adjoint: %s", cudaGetErrorString(e__)); } while (0)

Adjoint::Adjoint(Domain *dom) : dom_(dom), lat_(nullptr) {
    if ((int)dom->local.size() != dom->dec.nranks)
        fail("adjoint: every subdomain has to live in this process (%zu of %d ranks are local)", dom->local.size(), dom->dec.nranks);
    lat_ = dom->lat[dom->local[0]];
    Model *m = dom->model;
    if (m->storage_size() != m->real_size()) fail("adjoint: needs plain float/double lattice storage");
    action_ = m->find_action("Iteration");
    if (action_ < 0) fail("adjoint: model has no Iteration action");
    const tclb_action_t &a = m->d->actions[action_];
    stages_.assign(a.stages, a.stages + a.n_stages);
    stage_ = stages_[0];
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
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    subs_.resize(dom->local.size());
    for (size_t k = 0; k < subs_.size(); k++) {
        Sub &s = subs_[k];
        s.lat = dom->lat[dom->local[k]];
        s.bytes = (size_t)m->d->n_fields * s.lat->plane_elems * rs_;
        for (int i = 0; i < 2; i++) { s.adj[i] = alloc(s, s.bytes, dev); fill_zero(s, s.adj[i]); }
        if (stages_.size() > 1) {
            s.tmp = alloc(s, s.bytes, dev);
            s.resid = alloc(s, s.bytes, dev);
            fill_zero(s, s.resid);
            for (size_t j = 1; j < stages_.size(); j++) s.inter.push_back(alloc(s, s.lat->snap_bytes, dev));
        }
    }
    bytes_ = subs_[0].bytes;
    // tape budget on the device: TCLB_TAPE_DEVICE_BYTES, else 70 % of what is free now (a B200 has 180 GB); frames
    // beyond it go to page-locked host memory.  TCLB_TAPE_FRAMES (or set_tape_frames) bounds the NUMBER of stored states
    // instead: the forward run keeps checkpoints and the reverse sweep recomputes between them.
    const char *cap = getenv("TCLB_TAPE_DEVICE_BYTES");
    if (cap) dev_tape_cap_ = (size_t)atoll(cap);
    else if (dev) { size_t fr = 0, tot = 0; ADJ_CUDA(cudaMemGetInfo(&fr, &tot)); dev_tape_cap_ = (size_t)(0.7 * (double)fr); }
    else dev_tape_cap_ = 0;
    if (const char *f = getenv("TCLB_TAPE_FRAMES")) max_frames_ = atoi(f);
    if (const char *h = getenv("TCLB_TAPE_HOST_BYTES")) host_tape_cap_ = (size_t)atoll(h);
    const char *td = getenv("TCLB_TAPE_DIR");
    tape_dir_ = td && *td ? td : (getenv("TMPDIR") ? getenv("TMPDIR") : "/tmp");
}
Adjoint::~Adjoint() {
    drop_tape();
    for (auto &pp : pool_) release(pp.first, pp.second);
    const bool dev = lat_ && lat_->backend == TCLB_BACKEND_CUDA;
    for (Sub &s : subs_) {
        for (int i = 0; i < 2; i++) if (s.adj[i]) release(s.adj[i], dev);
        if (s.tmp) release(s.tmp, dev);
        if (s.resid) release(s.resid, dev);
        for (void *q : s.inter) release(q, dev);
    }
    if (sg_dev_) cudaFree(sg_dev_);
    if (zg_dev_) cudaFree(zg_dev_);
    if (lib_) dlclose(lib_);
}
void *Adjoint::alloc(const Sub &s, size_t bytes, bool device) const {
    void *p = nullptr;
    const bool cuda = s.lat->backend == TCLB_BACKEND_CUDA;
    if (device) { ADJ_CUDA(cudaSetDevice(s.lat->device)); ADJ_CUDA(cudaMalloc(&p, bytes)); }
    else if (cuda) ADJ_CUDA(cudaMallocHost(&p, bytes));     // page-locked: the tape copies run at full PCIe/C2C rate
    else { p = malloc(bytes); if (!p) fail("adjoint: out of host memory (%zu bytes)", bytes); }
    return p;
}
void Adjoint::release(void *p, bool device) const {
    if (device) cudaFree(p);
    else if (lat_->backend == TCLB_BACKEND_CUDA) cudaFreeHost(p);
    else free(p);
}
void Adjoint::fill_zero(const Sub &s, void *p) const {
    if (s.lat->backend == TCLB_BACKEND_CUDA) { ADJ_CUDA(cudaSetDevice(s.lat->device)); ADJ_CUDA(cudaMemsetAsync(p, 0, s.bytes, (cudaStream_t)s.lat->stream)); }
    else memset(p, 0, s.bytes);
}
void Adjoint::zero() { for (Sub &s : subs_) { fill_zero(s, s.adj[0]); fill_zero(s, s.adj[1]); if (s.resid) fill_zero(s, s.resid); } }
void Adjoint::copy_snap(const Sub &s, void *dst, const void *src, bool dst_dev, bool src_dev) const {
    if (s.lat->backend == TCLB_BACKEND_CUDA) {
        ADJ_CUDA(cudaSetDevice(s.lat->device));
        ADJ_CUDA(cudaMemcpyAsync(dst, src, s.lat->snap_bytes, cudaMemcpyDefault, (cudaStream_t)s.lat->stream));
        if (!dst_dev || !src_dev) ADJ_CUDA(cudaStreamSynchronize((cudaStream_t)s.lat->stream));
    } else memcpy(dst, src, s.lat->snap_bytes);
}
// ---- tape frames: one stored state = the current snapshot of every subdomain (+ the other snapshot for multi-stage
// actions, whose later stages may read what the previous iteration left there) ----
Adjoint::Frame Adjoint::grab_frame(bool with_pre) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    Frame f;
    f.iter = lat_->iter;
    size_t need = 0;
    for (Sub &s : subs_) need += s.lat->snap_bytes * (with_pre ? 2 : 1);
    f.on_device = dev && dev_tape_bytes_ + need <= dev_tape_cap_;
    f.on_disk = !f.on_device && host_tape_bytes_ + need > host_tape_cap_;
    if (f.on_disk) {
        for (Sub &s : subs_) {
            std::string name;
            frame_to_disk(s, s.lat->snap_ptr(s.lat->cur), dev, name);
            f.data.push_back(nullptr); f.files.push_back(name);
            if (with_pre) {
                frame_to_disk(s, s.lat->snap_ptr(s.lat->cur ^ 1), dev, name);
                f.pre.push_back(nullptr); f.pre_files.push_back(name);
            }
        }
        return f;
    }
    if (!f.on_device) host_tape_bytes_ += need;
    for (Sub &s : subs_) {
        void *d = frame_get(s, f.on_device);
        copy_snap(s, d, s.lat->snap_ptr(s.lat->cur), f.on_device, dev);
        f.data.push_back(d);
        if (with_pre) {
            void *q = frame_get(s, f.on_device);
            copy_snap(s, q, s.lat->snap_ptr(s.lat->cur ^ 1), f.on_device, dev);
            f.pre.push_back(q);
        }
    }
    if (f.on_device) dev_tape_bytes_ += need;
    return f;
}
void Adjoint::frame_to_disk(const Sub &s, const void *snap, bool src_dev, std::string &name) {
    void *h = frame_get(s, false);
    copy_snap(s, h, snap, false, src_dev);
    name = tape_dir_ + "/tclb_tape_" + std::to_string((long)getpid()) + "_" + std::to_string((long)(size_t)this % 1000003) + "_" + std::to_string(tape_file_seq_++) + ".bin";
    FILE *fp = fopen(name.c_str(), "wb");
    if (!fp) fail("adjoint tape: cannot write %s (TCLB_TAPE_DIR)", name.c_str());
    const size_t n = fwrite(h, 1, s.lat->snap_bytes, fp);
    fclose(fp);
    frame_put(h, false, s.lat->snap_bytes);
    if (n != s.lat->snap_bytes) fail("adjoint tape: short write to %s", name.c_str());
}
void Adjoint::frame_load(const Frame &f, std::vector<void *> &data, std::vector<void *> &pre) {
    auto rd = [&](const Sub &s, const std::string &name) {
        void *h = frame_get(s, false);
        FILE *fp = fopen(name.c_str(), "rb");
        if (!fp) fail("adjoint tape: cannot read %s", name.c_str());
        const size_t n = fread(h, 1, s.lat->snap_bytes, fp);
        fclose(fp);
        if (n != s.lat->snap_bytes) fail("adjoint tape: short read from %s", name.c_str());
        return h;
    };
    for (size_t k = 0; k < subs_.size(); k++) {
        data.push_back(rd(subs_[k], f.files[k]));
        if (!f.pre_files.empty()) pre.push_back(rd(subs_[k], f.pre_files[k]));
    }
}
void Adjoint::frame_unload(std::vector<void *> &data, std::vector<void *> &pre) {
    for (size_t k = 0; k < data.size(); k++) frame_put(data[k], false, subs_[k].lat->snap_bytes);
    for (size_t k = 0; k < pre.size(); k++) frame_put(pre[k], false, subs_[k].lat->snap_bytes);
    data.clear(); pre.clear();
}
void Adjoint::free_frame(Frame &f) {
    if (f.on_disk) {
        for (auto &n : f.files) remove(n.c_str());
        for (auto &n : f.pre_files) remove(n.c_str());
        f.files.clear(); f.pre_files.clear(); f.data.clear(); f.pre.clear();
        return;
    }
    if (!f.on_device) {
        size_t need = 0;
        for (size_t k = 0; k < subs_.size(); k++) need += subs_[k].lat->snap_bytes * ((k < f.data.size() && f.data[k]) ? 1 : 0) + subs_[k].lat->snap_bytes * ((k < f.pre.size() && f.pre[k]) ? 1 : 0);
        host_tape_bytes_ -= std::min(host_tape_bytes_, need);
    }
    size_t k = 0;
    for (Sub &s : subs_) {
        if (k < f.data.size() && f.data[k]) { frame_put(f.data[k], f.on_device, s.lat->snap_bytes); if (f.on_device) dev_tape_bytes_ -= s.lat->snap_bytes; }
        if (k < f.pre.size() && f.pre[k]) { frame_put(f.pre[k], f.on_device, s.lat->snap_bytes); if (f.on_device) dev_tape_bytes_ -= s.lat->snap_bytes; }
        k++;
    }
    f.data.clear(); f.pre.clear();
}
void *Adjoint::frame_get(const Sub &s, bool device) {
    for (size_t i = 0; i < pool_.size(); i++)
        if (pool_[i].second == device && pool_bytes_[i] == s.lat->snap_bytes && pool_dev_id_[i] == s.lat->device) {
            void *p = pool_[i].first;
            pool_.erase(pool_.begin() + i); pool_bytes_.erase(pool_bytes_.begin() + i); pool_dev_id_.erase(pool_dev_id_.begin() + i);
            return p;
        }
    return alloc(s, s.lat->snap_bytes, device);
}
void Adjoint::frame_put(void *p, bool device, size_t bytes) {
    pool_.push_back({p, device}); pool_bytes_.push_back(bytes);
    int dv = 0;
    for (Sub &s : subs_) if (s.lat->snap_bytes == bytes) { dv = s.lat->device; break; }
    pool_dev_id_.push_back(dv);
}
void Adjoint::restore_frame(const Frame &f) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    std::vector<void *> ld, lp;
    if (f.on_disk) frame_load(f, ld, lp);
    const std::vector<void *> &data = f.on_disk ? ld : f.data, &pre = f.on_disk ? lp : f.pre;
    for (size_t k = 0; k < subs_.size(); k++) {
        Sub &s = subs_[k];
        copy_snap(s, s.lat->snap_ptr(s.lat->cur), data[k], dev, f.on_device);
        if (!pre.empty()) copy_snap(s, s.lat->snap_ptr(s.lat->cur ^ 1), pre[k], dev, f.on_device);
        s.lat->iter = f.iter;
    }
    dom_->iter = f.iter;
    if (f.on_disk) { lat_->sync(); frame_unload(ld, lp); }
}
void Adjoint::drop_tape() {
    for (Frame &f : tape_) free_frame(f);
    tape_.clear();
    tape_count_ = 0;
    stride_ = 1;
    dev_tape_bytes_ = 0;
}
void Adjoint::set_tape_frames(int n) { max_frames_ = n; }

// n primal iterations.  Without a frame budget the input of every iteration is kept.  With one (max_frames_ > 1) only
// every stride-th input is kept, at most max_frames_/2 of them: when that many exist the stride doubles and every
// second checkpoint is dropped (the reference keeps snapshots on binomial "levels" by the trailing zeros of the iteration
// number and spills high levels to disk, src/Lattice.cu.Rt:46-61,200-233; the effect - logarithmic memory, recomputation
// between snapshots - is the same).
void Adjoint::record(int n) {
    const int obj = dom_->model->find_global("Objective");
    const bool multi = stages_.size() > 1;
    if (tape_.empty()) { if (!zg_which_.empty()) clear_zone_gradients(); tape_count_ = 0; stride_ = 1; integrated.clear(); }
    bool first = integrated.empty();
    for (int i = 0; i < n; i++) {
        const long g = tape_count_++;
        if (max_frames_ <= 1 || g % stride_ == 0) {
            Frame f = grab_frame(multi);
            f.index = g;
            tape_.push_back(f);
            if (max_frames_ > 1 && (int)tape_.size() > std::max(1, max_frames_ / 2)) {
                stride_ *= 2;
                std::vector<Frame> keep;
                for (Frame &fr : tape_) { if (fr.index % stride_ == 0) keep.push_back(fr); else free_frame(fr); }
                tape_.swap(keep);
            }
        }
        dom_->iterate(1, false, true);
        if (obj >= 0) objective += dom_->globals[obj];
        // Globals of a recorded segment are integrated over its iterations (reference ITER_INTEG, src/Lattice.h.Rt:29,
        // calcGlobals src/Lattice.cu.Rt:1313): sums add up, maxima take the maximum
        const int ng = dom_->model->d->n_globals;
        if ((int)integrated.size() != ng) integrated.assign(ng, 0.0);
        for (int k = 0; k < ng; k++) {
            if (dom_->model->d->globals[k].op == TCLB_OP_MAX) integrated[k] = first ? dom_->globals[k] : std::max(integrated[k], dom_->globals[k]);
            else integrated[k] = (first ? 0.0 : integrated[k]) + dom_->globals[k];
        }
        first = false;
    }
    if (n > 0) for (size_t k = 0; k < integrated.size(); k++) dom_->globals[k] = integrated[k];
}

// ---- one reverse step: adj <- J^T adj + dObjective, for the iteration whose input state is in `in` (per subdomain) ----
void Adjoint::fold(int which_buf /*0 adj_in of stage 0, 1 tmp*/) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    const int nf = dom_->model->d->n_fields;
    if (subs_.size() == 1) {
        Sub &s = subs_[0];
        void *b = which_buf ? s.tmp : s.adj[cur_ ^ 1];
        if (tclb_adjoint_fold_ghosts(b, rs_, dev, nf, s.lat->nx, s.lat->ny, s.lat->nz, s.lat->gy, s.lat->gz, s.lat->stream) != 0)
            fail("adjoint: ghost fold failed");
        return;
    }
    // several subdomains: what was scattered into my ghost layers belongs to the neighbour that owns those cells (the
    // transpose of the primal halo: there the owner's rows are copied into my ghost, here my ghost is added to its rows)
    if (dev) dom_->sync();
    for (size_t k = 0; k < subs_.size(); k++) {
        Sub &s = subs_[k];
        void *src = which_buf ? s.tmp : s.adj[cur_ ^ 1];
        for (int d = 0; d < TCLB_MAX_NEIGH; d++) {
            if (d == 4) continue;
            const int dy = d % 3 - 1, dz = d / 3 - 1;
            if ((dy != 0 && s.lat->gy == 0) || (dz != 0 && s.lat->gz == 0)) continue;
            const int nb = s.lat->halo[d].rank;
            size_t kk = 0;
            for (; kk < subs_.size(); kk++) if (subs_[kk].lat->rank == nb) break;
            if (kk == subs_.size()) fail("adjoint: neighbour rank %d is not local", nb);
            Sub &t = subs_[kk];
            void *dst = which_buf ? t.tmp : t.adj[cur_ ^ 1];
            if (dev) ADJ_CUDA(cudaSetDevice(s.lat->device));
            if (tclb_adjoint_fold_pair(dst, t.lat->ny, t.lat->nz, src, s.lat->ny, s.lat->nz, rs_, dev, nf, s.lat->nx, s.lat->gy, s.lat->gz,
                                       dy, dz, s.lat->stream) != 0)
                fail("adjoint: ghost fold between subdomains failed");
        }
        if (dev) dom_->sync();
    }
}
void Adjoint::launch_adjoint(Sub &s, int stage, const void *primal_in, int iter, const void *adj_out, void *adj_in) {
    s.lat->upload_zone_table_if_dirty();
    tclb_launch_t L;
    s.lat->fill_launch(L);
    L.snap_in = const_cast<void *>(primal_in);
    L.snap_out = nullptr;
    L.iter = iter;
    L.zone_index = L.zone_len > 1 ? iter % L.zone_len : 0;
    L.globals_mode = TCLB_GLOBALS_NONE;
    L.halo_mode = TCLB_HALO_NONE;
    L.obj_enabled = 1;
    if (s.lat->backend == TCLB_BACKEND_CUDA) ADJ_CUDA(cudaSetDevice(s.lat->device));
    if (fn_(stage, &L, adj_out, adj_in) != 0) fail("adjoint stage failed: %s", err_());
    if (&s == &subs_[0] || true) {
        stage_ = stage;
        if (!sg_which_.empty()) settings_pass(L, adj_out);
        if (!zg_which_.empty()) zsettings_pass(L, adj_out, L.zone_index);
    }
}
void Adjoint::zero_fields(Sub &s, void *buf, const unsigned char *which) {
    const size_t fb = (size_t)s.lat->plane_elems * rs_;
    for (int f = 0; f < dom_->model->d->n_fields; f++) {
        if (!which[f]) continue;
        if (s.lat->backend == TCLB_BACKEND_CUDA) ADJ_CUDA(cudaMemsetAsync((char *)buf + f * fb, 0, fb, (cudaStream_t)s.lat->stream));
        else memset((char *)buf + f * fb, 0, fb);
    }
}
void Adjoint::step(const std::vector<const void *> &in, const std::vector<const void *> &pre, int iter) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    const tclb_model_t *d = dom_->model->d;
    const size_t ns = stages_.size();
    if (ns == 1) {
        for (size_t k = 0; k < subs_.size(); k++) {
            Sub &s = subs_[k];
            fill_zero(s, s.adj[cur_ ^ 1]);
            launch_adjoint(s, stages_[0], in[k], iter, s.adj[cur_], s.adj[cur_ ^ 1]);
        }
        fold(0);
        cur_ ^= 1;
        return;
    }
    // ---- several stages: replay the iteration on the lattice itself to get the state each later stage read
    // (stage j > 0 reads and writes the "output" snapshot in place), then go through the stages backwards ----
    for (size_t k = 0; k < subs_.size(); k++) {
        Sub &s = subs_[k];
        copy_snap(s, s.lat->snap_ptr(s.lat->cur), in[k], dev, dev);
        copy_snap(s, s.lat->snap_ptr(s.lat->cur ^ 1), pre[k], dev, dev);
        s.lat->iter = iter;
    }
    dom_->iter = iter;
    const int tab0 = lat_->cur, tab1 = tab0 ^ 1;
    for (size_t j = 0; j < ns; j++) {
        if (j > 0) for (Sub &s : subs_) copy_snap(s, s.inter[j - 1], s.lat->snap_ptr(tab1), dev, dev);
        if (j + 1 == ns) break;          // the last stage's output is not needed
        for (Sub &s : subs_) s.lat->launch_stage(stages_[j], j == 0 ? tab0 : tab1, tab1, TCLB_GLOBALS_NONE, false);
        bool need_copy = false;
        for (Sub &s : subs_) if (s.lat->halo_mode == TCLB_HALO_NONE) need_copy = true;
        if (need_copy) dom_->copy_exchange(tab1);
    }
    // adj[cur_] = adjoint of the state after the whole iteration (+ what the NEXT iteration's later stages read of it)
    for (size_t j = ns; j-- > 1;) {
        for (Sub &s : subs_) {
            fill_zero(s, s.tmp);
            launch_adjoint(s, stages_[j], s.inter[j - 1], iter, s.adj[cur_], s.tmp);
        }
        fold(1);
        for (Sub &s : subs_) {
            zero_fields(s, s.adj[cur_], d->stages[stages_[j]].save_field);      // those values were produced by this stage
            if (tclb_adjoint_add_into(s.adj[cur_], s.tmp, (long long)(s.bytes / rs_), rs_, dev, s.lat->stream) != 0) fail("adjoint: add failed");
        }
    }
    for (size_t k = 0; k < subs_.size(); k++) {
        Sub &s = subs_[k];
        fill_zero(s, s.adj[cur_ ^ 1]);
        launch_adjoint(s, stages_[0], in[k], iter, s.adj[cur_], s.adj[cur_ ^ 1]);
    }
    fold(0);
    for (Sub &s : subs_) {
        zero_fields(s, s.adj[cur_], d->stages[stages_[0]].save_field);
        // what is left in adj[cur_] belongs to values the iteration found in its output snapshot, i.e. the state one
        // iteration earlier: it joins the adjoint of that state when the sweep gets there
        if (tclb_adjoint_add_into(s.adj[cur_ ^ 1], s.resid, (long long)(s.bytes / rs_), rs_, dev, s.lat->stream) != 0) fail("adjoint: add failed");
        std::swap(s.resid, s.adj[cur_]);
    }
    cur_ ^= 1;
}
// reverse sweep over the iterations [g0, g0 + len) whose first input state is frame `f`; `slots` more states may be held
void Adjoint::reverse_segment(const Frame &f, long len, int slots) {
    const bool multi = stages_.size() > 1;
    if (len <= 0) return;
    if (len == 1) {
        std::vector<const void *> in, pre;
        std::vector<void *> staged;
        const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
        std::vector<void *> ld, lp;
        if (f.on_disk) frame_load(f, ld, lp);
        const std::vector<void *> &fdata = f.on_disk ? ld : f.data, &fpre = f.on_disk ? lp : f.pre;
        for (size_t k = 0; k < subs_.size(); k++) {
            const void *p = fdata[k], *q = multi ? fpre[k] : nullptr;
            if (dev && !f.on_device) {       // host frame: bring it to the device first
                void *b = frame_get(subs_[k], true); copy_snap(subs_[k], b, p, true, false); staged.push_back(b); p = b;
                if (multi) { void *c = frame_get(subs_[k], true); copy_snap(subs_[k], c, q, true, false); staged.push_back(c); q = c; }
            }
            in.push_back(p); pre.push_back(q);
        }
        step(in, pre, f.iter);
        for (size_t k = 0; k < staged.size(); k++) frame_put(staged[k], true, subs_[0].lat->snap_bytes);
        if (f.on_disk) { lat_->sync(); frame_unload(ld, lp); }
        return;
    }
    if (len - 1 <= slots) {
        // recompute the whole segment, keeping every input state, then step back through them
        restore_frame(f);
        std::vector<Frame> fine;
        for (long i = 1; i < len; i++) {
            dom_->iterate(1, false, false);
            fine.push_back(grab_frame(multi));
        }
        for (long i = len - 1; i >= 1; i--) { reverse_segment(fine[i - 1], 1, 0); free_frame(fine[i - 1]); }
        reverse_segment(f, 1, 0);
        return;
    }
    // not enough room: advance to the middle, keep that state, do the second half, then the first
    const long half = len / 2;
    restore_frame(f);
    dom_->iterate((int)half, false, false);
    Frame mid = grab_frame(multi);
    reverse_segment(mid, len - half, slots - 1);
    free_frame(mid);
    reverse_segment(f, half, slots);
}
void Adjoint::reverse() {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    if (tape_.empty()) return;
    const bool recompute = tape_count_ > (long)tape_.size();
    Frame final_state;
    if (recompute || stages_.size() > 1) final_state = grab_frame(true);       // the sweep uses the lattice as scratch
    long next = tape_count_;
    while (!tape_.empty()) {
        Frame f = tape_.back();
        tape_.pop_back();
        reverse_segment(f, next - f.index, std::max(1, max_frames_ - (int)tape_.size() - 2));
        next = f.index;
        free_frame(f);
    }
    if (!final_state.data.empty()) { restore_frame(final_state); free_frame(final_state); }
    tape_count_ = 0;
    stride_ = 1;
    if (dev) dom_->sync();
}
void Adjoint::steady(int n) {
    std::vector<const void *> in, pre;
    for (Sub &s : subs_) { in.push_back(s.lat->snap_ptr(s.lat->cur)); pre.push_back(s.lat->snap_ptr(s.lat->cur ^ 1)); }
    if (stages_.size() > 1) fail("adjoint: steady adjoint iterations need a single-stage Iteration action");
    for (int i = 0; i < n; i++) step(in, pre, lat_->iter);
    if (lat_->backend == TCLB_BACKEND_CUDA) dom_->sync();
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
// interior <-> dense copies: the caller's array is the interior of the WHOLE lattice [tot_nz][tot_ny][tot_nx] in real_t
// (adjoint buffers have the snapshot layout [field][NZ][NY][nx] per subdomain)
void Adjoint::get_field(int field, void *dst) const {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    if (dev) dom_->sync();
    for (const Sub &s : subs_) {
        const Lattice *l = s.lat;
        const size_t row = (size_t)l->nx * rs_, grow = (size_t)l->tot[0] * rs_;
        const char *base = (const char *)s.adj[cur_] + (size_t)field * l->plane_elems * rs_;
        for (int z = 0; z < l->nz; z++) {
            const char *src = base + ((size_t)(z + l->gz) * l->NY + l->gy) * row;
            char *d = (char *)dst + (((size_t)(z + l->region.z0) * l->tot[1] + l->region.y0) * l->tot[0] + l->region.x0) * rs_;
            if (dev) { ADJ_CUDA(cudaSetDevice(l->device)); ADJ_CUDA(cudaMemcpy2D(d, grow, src, row, row, l->ny, cudaMemcpyDeviceToHost)); }
            else for (int y = 0; y < l->ny; y++) memcpy(d + y * grow, src + y * row, row);
        }
    }
}
void Adjoint::set_field(int field, const void *srcv) {
    const bool dev = lat_->backend == TCLB_BACKEND_CUDA;
    if (dev) dom_->sync();
    for (Sub &s : subs_) {
        Lattice *l = s.lat;
        const size_t row = (size_t)l->nx * rs_, grow = (size_t)l->tot[0] * rs_;
        char *base = (char *)s.adj[cur_] + (size_t)field * l->plane_elems * rs_;
        for (int z = 0; z < l->nz; z++) {
            char *d = base + ((size_t)(z + l->gz) * l->NY + l->gy) * row;
            const char *src = (const char *)srcv + (((size_t)(z + l->region.z0) * l->tot[1] + l->region.y0) * l->tot[0] + l->region.x0) * rs_;
            if (dev) { ADJ_CUDA(cudaSetDevice(l->device)); ADJ_CUDA(cudaMemcpy2D(d, row, src, grow, row, l->ny, cudaMemcpyHostToDevice)); }
            else for (int y = 0; y < l->ny; y++) memcpy(d + y * row, src + y * grow, row);
        }
    }
}
double Adjoint::dot_field(int field, const void *dir) const {
    const size_t n = (size_t)lat_->tot[0] * lat_->tot[1] * lat_->tot[2];
    std::vector<char> tmp(n * rs_);
    get_field(field, tmp.data());
    double s = 0;
    if (rs_ == 8) { const double *a = (const double *)tmp.data(), *b = (const double *)dir; for (size_t i = 0; i < n; i++) s += a[i] * b[i]; }
    else { const float *a = (const float *)tmp.data(), *b = (const float *)dir; for (size_t i = 0; i < n; i++) s += (double)a[i] * b[i]; }
    return s;
}

}  // namespace tclb
