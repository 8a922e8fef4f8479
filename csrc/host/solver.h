// Solver: host orchestration of one case (reference src/Solver.{h,cpp}.Rt + src/main.cpp:173-425):
// units/gauge, lattice sizing + decomposition, geometry, settings defaults, output naming, CSV log,
// VTK/TXT/BIN writers, component save/load, throughput monitor, and the handler stack that the XML
// elements drive.  One Solver may own several subdomains ("ranks") in one process, each on its own
// GPU; rank-numbered outputs are written exactly as an MPI run of the reference would.
#pragma once
#include <functional>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <vector>
#include "../core/lattice.h"
#include "../core/model.h"
#include "geometry.h"
#include "particles.h"
#include "units.h"
#include "xml.h"

namespace tclb {

class Handler;
class Sampler;
class Adjoint;

// "what=" lists (reference src/utils.h name_set)
struct NameSet {
    bool all = false;
    std::set<std::string> names;
    void add_from_string(const std::string &s, char sep = ',');
    bool in(const std::string &n) const { return all || names.count(n) > 0; }
    bool explicitly_in(const std::string &n) const { return names.count(n) > 0; }
};

enum { ITERATION_STOP = 1 };
enum { ITER_NORM = 0x00, ITER_ADJOINT = 0x01, ITER_OPT = 0x02, ITER_TYPE = 0x03, ITER_GLOBS = 0x10, ITER_OBJ = 0x20,
       ITER_INIT = 0x30, ITER_STEADY = 0x40, ITER_INTEG = 0x100, ITER_LASTGLOB = 0x200, ITER_SKIPGRAD = 0x400 };

struct SolverOptions {
    std::string model;                 // variant name, e.g. d2q9_bc
    std::string precision = "float";
    int backend = TCLB_BACKEND_CUDA;
    int nranks = 1;                    // subdomains in the whole job
    std::vector<int> local_ranks;      // empty = all
    std::vector<int> devices;          // empty = rank % device_count
    bool direct_halo = true;
    std::string model_path;            // explicit library path (optional)
    int x_block = -1;                  // nx padding granule (-1: 32 on CUDA, 1 on CPU)
    std::string simplepart_config;     // <SimplePart> XML file for <RemoteForceInterface integrator="SIMPLEPART"> (bin/tclb --simplepart)
};

class Solver {
  public:
    explicit Solver(const SolverOptions &opt);
    ~Solver();
    SolverOptions opt;
    std::unique_ptr<Model> model;
    std::unique_ptr<Domain> domain;
    UnitEnv units;
    XmlDocument config;
    std::string conffile;              // case file path
    std::string outpath;               // output prefix = <output attr><case basename>
    int iter = 0, opt_iter = 0, steps = 0;
    int iter_type = ITER_NORM;
    Box total;
    std::vector<std::unique_ptr<Geometry>> geometry;    // per rank (nullptr if remote)
    std::map<std::string, int> setting_zones;
    std::vector<std::shared_ptr<Handler>> hands;        // active periodic callbacks / design handlers
    std::unique_ptr<Sampler> sampler;
    std::vector<double> log_scales;
    double px = 0, py = 0, pz = 0;
    bool stop_requested = false;
    std::unique_ptr<ParticleIntegrator> integrator;    // particle coupling (<RemoteForceInterface>)
    // in-process scripting hook (<RunPython>/<RunR>): set by the embedding program, called with the element name, the
    // code (text of the element or of its file= attribute) and the iteration; a non-zero return aborts the run
    int (*script_hook)(void *user, const char *element, const char *code, int iteration) = nullptr;
    void *script_user = nullptr;

    // ---- set-up ----
    void load_case(const std::string &file, const std::vector<std::string> &xpath_args);
    void load_case_string(const std::string &xml, const std::string &name, const std::vector<std::string> &xpath_args);
    int run();                                           // executes the <CLBConfig> handler tree
    void read_units();
    void set_size(int nx, int ny, int nz);
    void set_output(const std::string &out);
    // before the lattice exists (reading <Units>, naming the output) the first local rank is the one asked for on the command line
    int first_rank() const { return domain && !domain->local.empty() ? domain->local[0] : (opt.local_ranks.empty() ? 0 : opt.local_ranks[0]); }
    bool is_root() const { return first_rank() == 0; }
    // ---- file names ----
    std::string out_iter_file(const std::string &name, const std::string &suffix, int rank = -1) const;
    std::string out_global_file(const std::string &name, const std::string &suffix, int rank = -1) const;
    std::string out_iter_collective_file(const std::string &name, const std::string &suffix) const;
    // ---- outputs ----
    void init_log(const std::string &filename);
    void write_log(const std::string &filename);
    int write_vtk(const std::string &name, const NameSet &what, const Box &region);
    int write_txt(const std::string &name, const NameSet &what, bool gzip);
    int write_hdf5(const std::string &name, const NameSet &what, const Box &region, bool as_double, bool write_xdmf, bool point_data, int deflate = 0);
    int write_bin(const std::string &name);
    void save_solution(const std::string &prefix, std::vector<std::string> *files = nullptr);
    void load_solution(const std::string &prefix);
    void save_comp(const std::string &prefix, const std::string &comp);
    void load_comp(const std::string &prefix, const std::string &comp);
    // ---- parameters ----
    void set_param(const std::string &name, const std::string &value, const std::string &zone, bool permissive, const std::string &gauge);
    // ---- iteration ----
    void iterate(int n, int type);
    // ---- adjoint (csrc/core/adjoint.h): while `recording`, iterate() keeps every iteration's input on the tape ----
    std::unique_ptr<Adjoint> adjoint;
    bool recording = false;
    Adjoint *need_adjoint();           // created on first use (loads the _ad companion of the model library)
    std::vector<int> parameter_fields() const;     // fields declared with parameter=TRUE (design variables)
    void iterate_action(int action, int n, int type);
    void monitor(bool force = false);
    bool has_nan(const NameSet &what, const Box &region);
    std::vector<double> get_quantity_global(int q, const Box &region, double scale);   // gathered [nz][ny][nx][comp]
  private:
    void post_parse(const std::vector<std::string> &xpath_args);
    double mon_last_time_ = 0;
    int mon_last_iter_ = 0;
};

}  // namespace tclb
