// Lattice: owner of one subdomain's memory and the launcher of its stages.
//
// B200-first redesign of the reference's Lattice + LatticeContainer pair
// (src/Lattice.cu.Rt:87-151 ctor, :466-533 Iteration, :327-389 MPI halo streams, :1279-1315 globals;
//  src/LatticeContainer.inc.cpp.Rt:144-172 Alloc, :294-342 launches):
//   * one SoA slab per snapshot with explicit y/z ghost layers instead of 27 margin blocks; the
//     27-block layout only exists in the checkpoint packer (pri.cpp) for file compatibility;
//   * no border/interior kernel pair and no host-staged MPI: one fused kernel per stage writes its
//     outgoing rows directly into the neighbour subdomain's ghost layer (same GPU, a peer GPU over
//     NVLink, or host memory for the CPU oracle) and publishes a sequence flag;
//   * launch parameters travel as a __grid_constant__ struct, no constant-memory copy per stage.
#pragma once
#include <cstdint>
#include <functional>
#include <map>
#include <string>
#include <vector>
#include "arena.h"
#include "model.h"
#include "zonesettings.h"

namespace tclb {

struct Box {
    int x0 = 0, y0 = 0, z0 = 0, nx = 0, ny = 0, nz = 0;
    long long size() const { return (long long)nx * ny * nz; }
    Box intersect(const Box &o) const;
    bool empty() const { return nx <= 0 || ny <= 0 || nz <= 0; }
};

// run-time switches read once from the environment: TCLB_DEBUG=sync,poison  TCLB_PROFILE=1
struct DebugFlags { bool sync = false, poison = false, profile = false; };
const DebugFlags &debug_flags();

// what a rank knows about another rank's memory (own process: plain pointers; other process:
// pointers obtained from a CUDA IPC mapping of that rank's arena)
struct PeerInfo {
    bool valid = false;
    void *snap[2] = {nullptr, nullptr};
    unsigned *flags_in = nullptr;  // 9 halo sequence words of that rank
    double *gl_ring = nullptr;     // its Globals ring
    int ny = 0, nz = 0;
};

// where the rows that leave through face/edge d=(dy+1)+3(dz+1) have to be written
struct HaloTarget {
    int rank = -1;                 // owning rank (may be this one: periodic wrap)
    void *snap[2] = {nullptr, nullptr};
    int ny = 0, nz = 0;            // interior extent of the owner
    unsigned *flag = nullptr;      // owner's flag word that I publish into (its flags_in[opposite(d)])
    bool concurrent = false;       // owner runs on another device or in another process: its kernels are not ordered
                                   // with mine by a stream, so the halo needs the sequence flags
};

class Lattice {
  public:
    Lattice(Model *model, int backend, int device, int rank, const Box &local, const int total[3], void *shared_stream = nullptr);
    ~Lattice();
    Lattice(const Lattice &) = delete;

    Model *model;
    const int backend, device, rank;
    Box region;                    // position/extent inside the global lattice
    int tot[3];
    int nx, ny, nz, gy, gz, NY, NZ;
    size_t plane_elems;            // nx*NY*NZ  (elements of one field incl. ghosts)
    size_t snap_bytes;
    int cur;                       // snapshot holding the current state
    int iter, reset_iter;
    unsigned epoch;                // launches issued so far (halo sequence number)
    std::vector<double> settings;
    ZoneSettings zs;
    std::vector<double> globals;   // last reduced values (local part)
    int halo_mode;
    bool sig_enable;
    HaloTarget halo[TCLB_MAX_NEIGH];
    void *stream;                  // cudaStream_t
    int block_x = 0, block_y = 0;
    // CUDA-graph capture of the iteration loop (Domain::iterate): while `capturing`, launches read iteration/epoch from
    // dyn_dev() plus the offsets recorded here; graph_version changes whenever a captured launch parameter changes
    bool capturing = false;
    int cap_iter_off = 0;
    unsigned cap_epoch_base = 0;
    unsigned long graph_version = 0;
    int *dyn_dev() const { return dyn_dev_; }
    double origin[3] = {0.0, 0.0, 0.0};   // <Geometry px py pz>: node (i,j,k) sits at origin + (i,j,k) + 0.5

    // ---- memory access ----
    std::vector<PeerInfo> peers;   // indexed by rank (self included)
    bool gl_complete = false;      // every rank's ring is mapped: Globals are exchanged in-kernel
    unsigned last_glob_epoch = 0;
    size_t last_h2d_bytes = 0, last_d2h_bytes = 0;   // sizes of the last settings upload / Globals read-back
    PeerInfo self_info() const;
    double *gl_ring_ptr() const { return gl_ring_; }
    void *arena_base() const { return arena_.base(); }
    void *snap_ptr(int i) const { return snap_[i]; }
    void *flags_ptr() const { return flags_; }
    unsigned *flags_in() const { return sig_; }            // 9 words written by neighbours
    void *field_ptr(int snap, int field) const;
    void flag_overwrite(const uint32_t *flags_local, const Box &local_box);   // box in local coordinates
    void get_flags(const Box &local_box, uint32_t *out) const;
    void cuts_overwrite(const uint16_t *cuts_local26);     // [26][nz][ny][nx] or nullptr
    void copy_field_out(int field, void *dst_native) const;  // interior only, [nz][ny][nx] in real_t
    void copy_field_in(int field, const void *src_native);
    void copy_field_out_raw(int field, void *dst_storage) const;   // same, in storage_t without the shift
    void copy_field_in_raw(int field, const void *src_storage);
    void storage_to_real(int field, const void *src, void *dst, size_t n) const;
    void real_to_storage(int field, const void *src, void *dst, size_t n) const;
    void fill_field(int field, double v);
    void get_quantity(int q, const Box &local_box, void *dst_native, double scale, bool dst_on_device = false);
    void color(int zplane, float scale, unsigned char *rgba_host_or_dev, bool dst_on_device = false);
    // ---- settings ----
    void set_setting(int idx, double v);                   // applies derived-settings chain
    double get_setting(int idx) const { return settings[idx]; }
    void set_synthetic_turbulence(const std::vector<tclb_st_mode_t> &modes);
    // ---- observability: TCLB_PROFILE=1 times every stage launch with CUDA events (device time,
    // serialising), TCLB_DEBUG=sync checks for errors after every launch, TCLB_DEBUG=poison fills the ghost layers of
    // the output snapshot with NaN before each stage so a missing halo write shows up as NaN instead of stale data
    struct StageStat { long long launches = 0; double ms = 0; };
    std::map<std::pair<int, int>, StageStat> stage_stats;   // (stage, globals mode) -> device time
    void poison_ghosts(int snap, int stage);     // ghosts of the fields that `stage` saves
    // ---- particles (reference Lattice::CopyInParticles/CopyOutParticles, src/Lattice.cu.Rt:392-437) ----
    // rows of 16 doubles in the RFI layout (r, pos, vel, angular velocity, force, moment); forces start at zero
    void set_particles(const std::vector<double> &rows16);
    void get_particles(std::vector<double> &rows16);        // after the particle stages: force/moment filled in
    int n_particles() const { return n_particles_; }
    double part_margin = 0.5;         // reach of a particle beyond its radius (model's PartMargin), set by the solver
    int particle_grid_depth() const { return pgrid_on_ ? pg_depth_ : 0; }
    // ---- execution ----
    void launch_stage(int stage, int in_snap, int out_snap, int globals_mode, bool publish_globals = true);
    void clear_globals();
    void fetch_globals();                                  // ring slot of the last publishing launch -> this->globals
    void sync() const;
    void upload_zone_table_if_dirty();
    // raw state (all fields incl. ghosts, current snapshot) for tests / memory dumps
    void save_raw(const std::string &file) const;
    void load_raw(const std::string &file);
    // TCLB-compatible 27-block checkpoint image of the current snapshot (pri.cpp)
    void save_pri(const std::string &file) const;
    void load_pri(const std::string &file);
    size_t pri_bytes() const;
    void pri_block_elems(long long out[27]) const;      // elements of each of the 27 margin blocks of the image (0 = absent)
    bool has_nan(int q, const Box &local_box);

  private:
    friend class Adjoint;          // adjoint.h: launches the adjoint stage with this lattice's launch parameters
    Arena arena_;
    void *snap_[2];
    void *flags_;
    double *globals_dev_;
    double *gl_ring_;
    unsigned *gl_counter_;
    void *gl_partial_;             // real_t[n_globals][gl_partial_stride_]: per-block Globals partials of a stage launch
    long long gl_partial_stride_;
    unsigned *sig_;               // [0..8] flags_in, [9] arrival counter
    void *zone_dev_;
    size_t zone_dev_bytes_;
    int zone_stride_;
    bool obj_enabled_ = true;      // any <Global>InObj weight non-zero (refreshed with the zone table)
    void *cuts_dev_;
    void *st_dev_;
    int st_n_;
    int *dyn_dev_ = nullptr;          // device int[2]: iteration and epoch at the start of a replayed graph
    void *particles_dev_ = nullptr;   // real_t[n_particles_][16]
    int n_particles_ = 0, particles_cap_ = 0;
    // uniform-grid container over the particles (csrc/kernels/particle_grid.cu; reference SolidGrid), rebuilt by set_particles
    int *pgrid_dev_ = nullptr, *pgrid_counts_ = nullptr, *pgrid_overflow_ = nullptr;
    size_t pgrid_cap_ = 0, pgrid_cells_cap_ = 0;
    int pg_n_[3] = {0, 0, 0}, pg_depth_ = 4;
    double pg_origin_[3] = {0, 0, 0}, pg_inv_delta_ = 0;
    bool pgrid_on_ = false;
    void build_particle_grid(const std::vector<double> &rows16);
    // box tree over the particles (csrc/core/particle_tree.cpp; reference SolidTree), TCLB_SOLID_CONTAINER=tree
    int *ptree_dev_ = nullptr;
    size_t ptree_cap_ = 0;
    int ptree_nodes_ = 0;
    bool ptree_on_ = false;
    void build_particle_tree(const std::vector<double> &rows16);
    bool owns_stream_ = false;
    void *pinned_ = nullptr;       // page-locked staging: [0,half) host->device, [half,end) device->host
    size_t pinned_bytes_ = 0;
    void *scratch_;                // quantity staging
    size_t scratch_bytes_;
    void fill_launch(tclb_launch_t &L) const;
    void *scratch(size_t bytes);
    void to_host(void *dst, const void *src, size_t bytes) const;
    void to_dev(void *dst, const void *src, size_t bytes) const;
    void pri_stream(const std::string &file, bool save, void *dev_stage, char *host_stage, size_t stage_bytes);
};

// Decomposition of the global lattice over ranks: y x z split only, the reference's rule
// (src/Solver.cpp.Rt:288-370): minimise divz*ny + divy*nz, greedy remainder split, rank = iy + divy*iz.
struct Decomposition {
    int nranks = 1, divy = 1, divz = 1;
    int tot[3] = {0, 0, 0};
    std::vector<Box> boxes;
    static Decomposition make(int nx, int ny, int nz, int nranks);
    int neighbour(int rank, int dy, int dz) const;         // torus
};

// All subdomains this process owns + wiring between them.
class Domain {
  public:
    Domain(Model *model, int backend, int nx, int ny, int nz, int nranks,
           const std::vector<int> &local_ranks, const std::vector<int> &devices);
    ~Domain();
    Model *model;
    int backend;
    Decomposition dec;
    std::vector<Lattice *> lat;            // indexed by rank; nullptr when remote
    std::vector<int> local;                // ranks owned here
    int iter;
    bool use_direct_halo;                  // fused in-kernel halo (product path) vs copy exchange (baseline)
    // The iteration loop as a CUDA graph: two iterations (one per snapshot orientation) are captured once and replayed;
    // pays off when the lattice is small enough for launch latency to show.  Used by iterate() when nothing in the loop
    // needs the host (no Globals, no particle stages, fused halo, one device in this process).
    bool use_graphs = true;
    long long graph_replays = 0;           // statistics: iterations executed through graph replays
    bool graph_possible(int action) const;
    void run_graph_pairs(int action, int pairs);
    void drop_graphs();
    // hooks for multi-process operation
    std::function<void(int /*epoch*/)> exchange_hook;       // HALO_NONE: external (NCCL/gloo) exchange after every stage
    std::function<void(double *, int, int)> reduce_hook;    // (values, count, op) across processes
    // particle coupling: called around the particle stages of an action (before the first: load the particles of every
    // local subdomain; after the last: hand the forces back)
    std::function<void()> particles_in_hook, particles_out_hook;

    void connect_local();                  // publish local lattices to each other (enables peer access)
    void set_peer(int local_rank, int peer_rank, const PeerInfo &p);
    void finalize_wiring();                // halo targets from the peer tables; decide sig_enable / Globals mode
    // settings fan-out
    void set_setting(int idx, double v);
    void set_zone_setting(int s, int zone, double v);
    void set_zone_series(int s, int zone, const std::vector<double> &v);
    void set_zones(int n);
    // execution
    void init();
    void run_action(int action, int globals_mode);
    void iterate(int n, bool globals_on_last, bool globals_every = false);
    void copy_exchange(int snap_index);    // baseline: memcpy border rows -> neighbour ghosts (local ranks)
    void sync();
    std::vector<double> globals;           // reduced over all ranks
    void reduce_globals();
    long long nodes() const { return (long long)dec.tot[0] * dec.tot[1] * dec.tot[2]; }

  private:
    // one instantiated graph per snapshot orientation at entry (cur = 0 / 1)
    struct GraphSlot { void *exec = nullptr; unsigned long key = 0; int action = -1; unsigned launches = 0; };
    GraphSlot graph_[2];
    unsigned long graph_key() const;
};

}  // namespace tclb
