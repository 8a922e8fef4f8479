// Discrete adjoint driver: reverse sweep over a recorded primal run, or fixed-point iteration at a converged state.
//
// Role of the reference's adjoint machinery (Lattice::IterateTillAdjoint / snapshots "tape" levels,
// src/Lattice.cu.Rt:536-700; handlers acUSAdjoint / acSAdjoint), built differently: the stage adjoint comes from the
// dual-number build of the model library (libtclb_<model>_<prec>_ad.so, tclb_b200/gen/adjoint.cu.in), the tape is a
// list of whole input snapshots (device memory first, host memory beyond TCLB_TAPE_DEVICE_BYTES), and the halo
// transpose is a ghost fold.  This version covers actions made of ONE stage on ONE subdomain.
#pragma once
#include <string>
#include <vector>
#include "lattice.h"

namespace tclb {

class Adjoint {
  public:
    explicit Adjoint(Domain *dom);
    ~Adjoint();
    Adjoint(const Adjoint &) = delete;
    // ---- unsteady ----
    void record(int n);          // n primal iterations with the input of each kept on the tape; Objective accumulated
    void reverse();              // sweep the tape backwards; afterwards adjoint() is dObjective/d(state at the start)
    void drop_tape();
    size_t tape_length() const { return (size_t)tape_count_; }     // iterations recorded
    size_t tape_frames() const { return tape_.size(); }            // states actually held
    // bound the number of stored states (0 = keep every iteration's input): the forward run then keeps checkpoints at a
    // stride that doubles when the budget fills up, and reverse() recomputes the iterations in between (bisection when a
    // gap is longer than the free slots).  Also TCLB_TAPE_FRAMES=<n>.
    void set_tape_frames(int n);
    // ---- steady ----
    void steady(int n);          // n adjoint iterations around the current state (input of the next primal iteration)
    // ---- access ----
    double objective = 0.0;      // sum over the recorded iterations of the Objective global
    std::vector<double> integrated;   // every Global integrated over the recorded iterations (SUM: sum, MAX: maximum)
    void zero();
    void get_field(int field, void *dst_real) const;    // interior [nz][ny][nx] of the adjoint of `field`, in real_t
    void set_field(int field, const void *src_real);    // seed (e.g. d(cost)/d(final state))
    double dot_field(int field, const void *dir_real) const;   // sum adjoint * dir over the interior
    int real_size() const { return rs_; }
    // ---- gradients with respect to (non-zonal) settings, accumulated by reverse()/steady() ----
    void want_setting_gradients(const std::vector<int> &settings);   // call before record(); clears the sums
    double setting_gradient(int setting);        // total derivative: follows the derived-settings chain (nu -> omega ...)
    // ---- gradients with respect to zonal settings, per time level of their series (reference: the gradient series of
    // src/ZoneSettings.h:197-219, filled by the Tapenade adjoint; here by a tangent pass per reverse step) ----
    void want_zone_gradient(int zonesetting, int zone);     // call before record(); keeps earlier requests
    void clear_zone_gradients();                            // zero the sums (done by record() of a fresh tape)
    // d Objective / d value(t), t = 0..len-1 (len = series length, 1 for a constant); the dependence through <name>_DT
    // (central difference of the series) is folded back onto the neighbouring time levels
    void zone_gradient(int zonesetting, int zone, double *out, int len);
  private:
    typedef int (*zsettings_fn)(int, const tclb_launch_t *, const void *, const int *, const int *, const int *, int, double *);
    zsettings_fn zfn_ = nullptr;
    std::vector<std::pair<int, int>> zg_which_;      // (zonal setting, zone)
    std::vector<double> zg_sum_;                     // [time level][pair][value, dt]
    double *zg_dev_ = nullptr;
    int zg_len_ = 0;
    void zsettings_pass(const tclb_launch_t &L, const void *adj_out, int time_level);
    void zg_resize();
    typedef int (*settings_fn)(int, const tclb_launch_t *, const void *, const int *, int, double *);
    settings_fn sfn_ = nullptr;
    int dirs_ = 4;
    std::vector<int> sg_which_;          // requested settings + everything derived from them
    std::vector<double> sg_sum_;         // direct partial derivatives, same order
    double *sg_dev_ = nullptr;           // device accumulator (CUDA backend)
    void settings_pass(const tclb_launch_t &L, const void *adj_out);
    typedef int (*stage_fn)(int, const tclb_launch_t *, const void *, void *);
    typedef const char *(*err_fn)(void);
    // one stored state of the whole lattice: per subdomain the current snapshot (+ the other one for multi-stage actions)
    // a stored state lives in device memory, in (page-locked) host memory, or - beyond TCLB_TAPE_HOST_BYTES - in files under
    // TCLB_TAPE_DIR (the reference's snapshot levels end on disk the same way, src/Lattice.cu.Rt:200-233)
    struct Frame { std::vector<void *> data, pre; bool on_device = false, on_disk = false; int iter = 0; long index = 0;
                   std::vector<std::string> files, pre_files; };
    size_t host_tape_bytes_ = 0, host_tape_cap_ = (size_t)-1;
    std::string tape_dir_;
    long tape_file_seq_ = 0;
    // per subdomain: two adjoint buffers (snapshot layout in real_t), and for multi-stage actions a scratch adjoint, the
    // adjoint left over for the previous state, and the state before every later stage
    struct Sub { Lattice *lat = nullptr; void *adj[2] = {nullptr, nullptr}; void *tmp = nullptr, *resid = nullptr; std::vector<void *> inter; size_t bytes = 0; };
    void frame_to_disk(const Sub &s, const void *snap_dev_or_host, bool src_dev, std::string &name);
    void frame_load(const Frame &f, std::vector<void *> &data, std::vector<void *> &pre);     // host staging copies of an on-disk frame
    void frame_unload(std::vector<void *> &data, std::vector<void *> &pre);
    Domain *dom_;
    Lattice *lat_;
    std::vector<Sub> subs_;
    std::vector<int> stages_;            // stages of the Iteration action
    void *lib_ = nullptr;
    stage_fn fn_ = nullptr;
    err_fn err_ = nullptr;
    int rs_ = 8, action_ = -1, stage_ = -1;
    int cur_ = 0;
    size_t bytes_ = 0;
    std::vector<Frame> tape_;
    long tape_count_ = 0, stride_ = 1;   // iterations recorded since the tape was empty; spacing of the kept checkpoints
    int max_frames_ = 0;                 // 0: keep every input state
    size_t dev_tape_bytes_ = 0, dev_tape_cap_ = 0;
    std::vector<std::pair<void *, bool>> pool_;      // recycled frame buffers (pointer, on device)
    std::vector<size_t> pool_bytes_;
    std::vector<int> pool_dev_id_;
    Frame grab_frame(bool with_pre);
    void free_frame(Frame &f);
    void restore_frame(const Frame &f);
    void *frame_get(const Sub &s, bool device);
    void frame_put(void *p, bool device, size_t bytes);
    void copy_snap(const Sub &s, void *dst, const void *src, bool dst_dev, bool src_dev) const;
    void step(const std::vector<const void *> &in, const std::vector<const void *> &pre, int iter);
    void reverse_segment(const Frame &f, long len, int slots);
    void launch_adjoint(Sub &s, int stage, const void *primal_in, int iter, const void *adj_out, void *adj_in);
    void zero_fields(Sub &s, void *buf, const unsigned char *which);
    void fold(int which_buf);
    void *alloc(const Sub &s, size_t bytes, bool device) const;
    void release(void *p, bool device) const;
    void fill_zero(const Sub &s, void *p) const;
};

}  // namespace tclb
