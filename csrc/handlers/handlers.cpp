// All XML handlers of the case-file language (reference src/Handlers/*.cpp, one class per element;
// the table in SURVEY.md section 2.3 lists them).  Behaviour is re-stated per class below with the
// reference file it mirrors.
#include "handler.h"
#include "../host/rfi_bridge.h"
#include "../core/common.h"
#include "../host/sampler.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <queue>
#include <sstream>

namespace tclb {

// ------------------------------------------------------------------ infrastructure ----
std::map<std::string, HandlerFactory::Maker> &HandlerFactory::table() {
    static std::map<std::string, Maker> t;
    return t;
}
bool HandlerFactory::add(const std::string &n, Maker m) { table()[n] = m; return true; }
HandlerPtr HandlerFactory::produce(XmlNode *node, Solver *solver) {
    auto it = table().find(node->name);
    if (it == table().end()) { logf(L_FATAL, "Unknown Handler: %s", node->name.c_str()); return nullptr; }
    Handler *h = it->second();
    h->node = node;
    h->solver = solver;
    int ret;
    try { ret = h->init(); }
    catch (const Error &e) { logf(L_FATAL, "%s: %s", node->name.c_str(), e.what()); ret = -1; }
    if (ret) { delete h; return nullptr; }
    return HandlerPtr(h, [](Handler *p) { try { p->finish(); } catch (...) {} delete p; });
}
std::string Handler::context_attr(const std::string &name, bool *found) const {
    for (const XmlNode *n = node; n; n = n->parent)
        if (n->has_attr(name)) { if (found) *found = true; return n->attr(name); }
    if (found) *found = false;
    return "";
}
void Handler::read_iterations() {
    start_iter = solver->iter;
    if (node->has_attr("Iterations")) {
        every_iter = solver->units.alt(node->attr("Iterations"));
        logf(L_OUTPUT, "Setting %s at %lf iterations", node->name.c_str(), every_iter);
    } else every_iter = 0;
}
int Action::init() {
    read_iterations();
    if (node->has_attr("output")) solver->set_output(node->attr("output"));
    return 0;
}
int GenericAction::execute_internal() {
    // children: design handlers and periodic callbacks go on the solver's stack; callbacks without
    // Iterations fire once, immediately; actions/containers do their work in init()
    stack = 0;
    for (auto *par : node->children()) {
        HandlerPtr h = HandlerFactory::produce(par, solver);
        if (!h) { logf(L_FATAL, "Something wrong in %s", node->name.c_str()); return -1; }
        if (h->type() & HANDLER_DESIGN) { solver->hands.push_back(h); stack++; }
        else if (h->type() & HANDLER_CALLBACK) {
            if (h->every_iter != 0) { solver->hands.push_back(h); stack++; }
            else if (h->do_it() != 0) { logf(L_ERROR, "Handler call error: %s", par->name.c_str()); return -1; }
        }
    }
    return 0;
}
void GenericAction::unstack() {
    while (stack-- > 0) solver->hands.pop_back();
    stack = 0;
}
int GenericAction::finish() {
    if (stack > 0) { logf(L_WARNING_IMP, "Generic action still stacked at finish"); unstack(); }
    return 0;
}
int GenericAction::run_loop(std::function<void(int, int, bool)> step_fn) {
    // reference src/Handlers/acSolve.cpp:5-43
    bool stop = false;
    do {
        int my_next = next(solver->iter);
        int next_it = my_next;
        for (auto &h : solver->hands) {
            int it = h->next(solver->iter);
            if (it > 0 && it < next_it) next_it = it;
        }
        if (next_it <= 0) { logf(L_ERROR, "%s has no Iterations to run", node->name.c_str()); return -1; }
        solver->steps = next_it;
        solver->iter += next_it;
        int type = solver->iter_type;
        bool last = next_it == my_next;
        if (last) type |= ITER_LASTGLOB;
        step_fn(next_it, type, last);
        solver->domain->sync();
        for (size_t i = 0; i < solver->hands.size(); i++) {
            if (!solver->hands[i]->now(solver->iter)) continue;
            int ret = solver->hands[i]->do_it();
            if (ret == ITERATION_STOP) stop = true;
            else if (ret != 0) return -1;
        }
        if (stop) break;
    } while (!now(solver->iter));
    solver->domain->sync();
    return 0;
}

namespace {

// ------------------------------------------------------------------ containers ----
// <CLBConfig>: echo the (possibly modified) case as <out>_config_P00_00000000.xml with a <Run> stamp, run children
// (reference src/Handlers/MainContainer.cpp:5-27)
class MainContainer : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        if (node->has_attr("output")) solver->set_output(node->attr("output"));
        if (solver->is_root()) {
            XmlNode *run = node->append_element("Run");
            run->set_attr("model", solver->model->d->name);
            XmlNode *c = run->append_element("Code");
            c->set_attr("version", "tclb-b200");
            c->set_attr("precision", solver->model->real_size() == 8 ? "double" : "float");
            c->set_attr("cross", solver->opt.backend == TCLB_BACKEND_CUDA ? "GPU" : "CPU");
            solver->config.save_file(solver->out_iter_file("config", ".xml"));
            node->remove(run);
        }
        return execute_internal();
    }
    int finish() override { unstack(); return GenericAction::finish(); }
};
TCLB_REGISTER_HANDLER("CLBConfig", MainContainer)

class UnitsContainer : public GenericAction {      // <Units>: already consumed by Solver::read_units
  public:
    int init() override { return 0; }
};
TCLB_REGISTER_HANDLER("Units", UnitsContainer)

class AcContainer : public GenericAction {          // <Container> (reference acContainer.cpp)
  public:
    int init() override {
        GenericAction::init();
        if (execute_internal()) return -1;
        unstack();
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Container", AcContainer)

class AcRepeat : public GenericAction {             // <Repeat Times=> (reference acRepeat.cpp)
  public:
    int init() override {
        GenericAction::init();
        if (!node->has_attr("Times")) { logf(L_ERROR, "no Times parameter in %s", node->name.c_str()); return -1; }
        int times = atoi(node->attr("Times").c_str());
        for (int i = 0; i < times; i++) {
            if (execute_internal()) return -1;
            unstack();
        }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Repeat", AcRepeat)

class IfContainer : public GenericAction {          // <EvalIf opt_present= | opt_missing=> (reference IfContainer.cpp)
  public:
    int init() override {
        GenericAction::init();
        bool p = node->has_attr("opt_present"), m = node->has_attr("opt_missing");
        if (p == m) { logf(L_FATAL, "Use either opt_present or opt_missing (not both) in EvalIf"); return -1; }
        std::string option = p ? node->attr("opt_present") : node->attr("opt_missing");
        bool known = false;
        for (int i = 0; i < solver->model->d->n_options; i++) if (option == solver->model->d->options[i].name) known = true;
        if (!known) { logf(L_FATAL, "Unknown option in EvalIf: %s", option.c_str()); return -1; }
        if (solver->model->has_option(option) == p) return execute_internal();
        return 0;
    }
    int finish() override { unstack(); return GenericAction::finish(); }
};
TCLB_REGISTER_HANDLER("EvalIf", IfContainer)

class NullHandler : public Handler {                // <Run> marker written into config echoes
  public:
    int init() override { return 0; }
};
TCLB_REGISTER_HANDLER("Run", NullHandler)

// ------------------------------------------------------------------ actions ----
// <Geometry px py pz>: paint flags/cuts of every local subdomain, upload (reference acGeometry.cpp:5-24)
class AcGeometry : public Action {
  public:
    int init() override {
        Action::init();
        if (node->has_attr("px")) solver->px = solver->units.alt(node->attr("px"));
        if (node->has_attr("py")) solver->py = solver->units.alt(node->attr("py"));
        if (node->has_attr("pz")) solver->pz = solver->units.alt(node->attr("pz"));
        for (int r : solver->domain->local) {
            Geometry *g = solver->geometry[r].get();
            // every rank paints from a pristine copy of the subtree (load() prepends the built-in zones)
            std::unique_ptr<XmlNode> copy = node->clone();
            XmlNode *tmp = node->parent->insert_after(std::move(copy), node);
            try { g->load(tmp); }
            catch (...) { node->parent->remove(tmp); throw; }
            node->parent->remove(tmp);
            Lattice *l = solver->domain->lat[r];
            l->origin[0] = solver->px; l->origin[1] = solver->py; l->origin[2] = solver->pz;
            Box b; b.nx = l->nx; b.ny = l->ny; b.nz = l->nz;
            l->flag_overwrite(g->geom.data(), b);
            if (!g->Q.empty()) l->cuts_overwrite(g->Q.data());
            solver->setting_zones = g->zones;
            if (node->has_attr("save")) g->write_vti(node->attr("save"), r);
        }
        solver->domain->set_zones((int)solver->setting_zones.size());
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Geometry", AcGeometry)

class AcModel : public GenericContainer {            // <Model>: Params then Init action (reference acModel.cpp)
  public:
    int init() override {
        int ret = GenericContainer::init();
        if (ret) return ret;
        solver->domain->init();
        solver->iter = 0;
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Model", AcModel)

class AcInit : public Action {
  public:
    int init() override {
        Action::init();
        solver->domain->init();
        solver->iter = 0;
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Init", AcInit)

class AcParam : public Action {                      // <Param name value [zone] [gauge]> (reference acParam.cpp:5-60)
  public:
    int init() override {
        Action::init();
        bool found = false;
        std::string perm = context_attr("permissive", &found);
        bool permissive = found && !perm.empty() && (perm[0] == 't' || perm[0] == 'T' || perm[0] == '1' || perm[0] == 'y');
        solver->set_param(node->attr("name"), node->attr("value"), node->attr("zone"), permissive, node->attr("gauge"));
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Param", AcParam)

class AcSolve : public GenericAction {               // <Solve Iterations=> (reference acSolve.cpp:5-43)
  public:
    int init() override {
        GenericAction::init();
        if (execute_internal()) return -1;
        int ret = run_loop([this](int steps, int type, bool) { solver->iterate(steps, type); });
        unstack();
        return ret;
    }
};
TCLB_REGISTER_HANDLER("Solve", AcSolve)

class AcRunAction : public GenericAction {           // <RunAction name= Iterations=> (reference acRunAction.cpp)
  public:
    int init() override {
        GenericAction::init();
        std::string nm = node->attr("name");
        if (nm.empty()) { logf(L_FATAL, "Have to specify the name of the Action in RunAction"); return -1; }
        int action = solver->model->find_action(nm);
        if (action < 0) { logf(L_FATAL, "Unknown Action %s", nm.c_str()); return -1; }
        if (execute_internal()) return -1;
        int ret = run_loop([this, action](int steps, int type, bool) { solver->iterate_action(action, steps, type); });
        unstack();
        return ret;
    }
};
TCLB_REGISTER_HANDLER("RunAction", AcRunAction)

class AcLoadBinary : public Action {                 // <LoadBinary file|filename [comp]> (reference acLoadBinary.cpp)
  public:
    int init() override {
        Action::init();
        std::string f = node->has_attr("file") ? node->attr("file") : node->attr("filename");
        if (f.empty()) { logf(L_ERROR, "No file specified in %s", node->name.c_str()); return -1; }
        if (node->has_attr("comp")) solver->load_comp(f, node->attr("comp"));
        else solver->load_solution(f);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("LoadBinary", AcLoadBinary)
class AcLoadMemoryDump : public AcLoadBinary {};
TCLB_REGISTER_HANDLER("LoadMemoryDump", AcLoadMemoryDump)

// <SyntheticTurbulence Modes Spread Spectrum Main/Diffusion/Shortest/Longest/Time{WaveLength,WaveNumber,WaveFrequency}>
// (reference acSyntheticTurbulence.cpp:36-115, src/SyntheticTurbulence.cpp:49-131)
class AcSyntheticTurbulence : public Action {
    bool read_wn(const std::string &name, double *var) {
        int set = 0;
        double val = *var;
        if (node->has_attr(name + "WaveLength")) { set++; val = 1.0 / solver->units.alt(node->attr(name + "WaveLength")); }
        if (node->has_attr(name + "WaveNumber")) { set++; val = solver->units.alt(node->attr(name + "WaveNumber")); }
        if (node->has_attr(name + "WaveFrequency")) { set++; val = solver->units.alt(node->attr(name + "WaveFrequency")) * 8 * atan(1.0); }
        if (set > 1) logf(L_NOTICE_IMP, "Only one of WaveLength, WaveNumber or WaveFrequency can be set for \"%s\". Taking the last.", name.c_str());
        if (set) *var = val;
        return set > 0;
    }
    static void rnorm(int n, double *tab) {
        for (int i = 0; i < n; i++) {
            double a = (double)rand() / RAND_MAX * atan(1.0) * 8, b = sqrt(-log((double)rand() / RAND_MAX));
            tab[i] = cos(a) * b;
            if (++i >= n) break;
            tab[i] = sin(a) * b;
        }
    }
  public:
    int init() override {
        Action::init();
        int nmodes = node->has_attr("Modes") ? atoi(node->attr("Modes").c_str()) : 100;
        std::string spread = node->attr("Spread", "Even");
        if (spread != "Even" && spread != "Log" && spread != "Quantile") {
            logf(L_FATAL, "Unknown spread type \"%s\" in %s (available: Even, Log and Quantile)", spread.c_str(), node->name.c_str());
            return -1;
        }
        std::string spec = node->attr("Spectrum", "Von Karman");
        std::vector<double> amp, wl;
        if (spec == "Von Karman") {
            double mainWN = 0, diffWN = 0, maxWN, minWN;
            if (!read_wn("Main", &mainWN)) { logf(L_FATAL, "Must provide MainWaveNumber for synthetic turbulence Von Karman spectrum"); return -1; }
            if (!read_wn("Diffusion", &diffWN)) { logf(L_FATAL, "Must provide DiffusionWaveNumber for synthetic turbulence Von Karman spectrum"); return -1; }
            maxWN = 8 * atan(1.0) / 4;
            read_wn("Shortest", &maxWN);
            minWN = mainWN / 2;
            read_wn("Longest", &minWN);
            double dL = (maxWN - minWN) / nmodes, sum = 0;
            for (int i = 0; i < nmodes; i++) {
                double L = i * dL + dL / 2 + minWN;
                double E = 1.453 / mainWN * pow(L / mainWN, 4.0) / pow(1.0 + pow(L / mainWN, 2.0), 17. / 6.) * exp(-2.0 * pow(L / diffWN, 2.0));
                wl.push_back(L);
                amp.push_back(sqrt(E * dL));
                sum += E * dL;
            }
            logf(L_OUTPUT, "Total energy of synthetic turbulence: %2.0lf%% of spectrum", sum * 100);
        } else if (spec == "One Wave") {
            double WN = 0;
            if (!read_wn("", &WN)) { logf(L_FATAL, "Must provide WaveNumber for synthetic turbulence"); return -1; }
            amp.push_back(1);
            wl.push_back(WN);
        } else { logf(L_FATAL, "Unknown spectrum type \"%s\" in %s", spec.c_str(), node->name.c_str()); return -1; }
        double timeWN = 0;
        read_wn("Time", &timeWN);
        std::vector<tclb_st_mode_t> modes(amp.size());
        srand(node->has_attr("seed") ? atoi(node->attr("seed").c_str()) : 1);
        for (size_t j = 0; j < amp.size(); j++) {
            double t[6];
            rnorm(6, t);
            double l = sqrt(t[0] * t[0] + t[1] * t[1] + t[2] * t[2]);
            for (int i = 0; i < 3; i++) t[i] /= l;                       // unit wave direction n
            l = t[0] * t[3] + t[1] * t[4] + t[2] * t[5];
            for (int i = 0; i < 3; i++) t[i + 3] -= t[i] * l;            // amplitude vector a, orthogonal to n
            l = sqrt(t[3] * t[3] + t[4] * t[4] + t[5] * t[5]);
            for (int i = 0; i < 3; i++) t[i + 3] *= amp[j] / l;
            tclb_st_mode_t &m = modes[j];                                 // u += sin(w n.x) a + cos(w n.x) (n x a)
            for (int i = 0; i < 3; i++) { m.k[i] = t[i] * wl[j]; m.b[i] = t[i + 3]; }
            m.a[0] = t[1] * t[5] - t[2] * t[4];
            m.a[1] = t[2] * t[3] - t[0] * t[5];
            m.a[2] = t[0] * t[4] - t[1] * t[3];
            m.freq = 0;      // the reference stores TimeWN but its device evaluation does not use it
        }
        if (timeWN != 0) {
            // the time wave number travels as the frequency of one more mode of zero amplitude (it adds exactly nothing to the
            // velocity): node code that asks for `constContainer.ST.TimeWN` (models/heat/experimental/d3q27_cumulant_heat) reads it there
            tclb_st_mode_t m;
            for (int i = 0; i < 3; i++) { m.k[i] = 0; m.a[i] = 0; m.b[i] = 0; }
            m.freq = timeWN;
            modes.push_back(m);
        }
        for (int r : solver->domain->local) solver->domain->lat[r]->set_synthetic_turbulence(modes);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("SyntheticTurbulence", AcSyntheticTurbulence)

// <Control Iterations=N> <CSV file= Time=> <Param name zone value="Col*1m/s+..."/> </CSV> </Control>
// time-dependent zonal settings with linear interpolation (reference conControl.cpp:6-253)
class ConControl : public Action {
    typedef std::map<std::string, std::vector<double>> Context;
    Context context;
    int iters = 0;
    int get(Context &cont, const std::string &svar, double scale, std::vector<double> &fill) {
        size_t n = cont.empty() ? (size_t)iters : cont.begin()->second.size();
        fill.assign(n, 0.0);
        int sum_i = 0;
        for (auto &var : split(svar, '+')) {
            auto toks = split(var, '*');
            if (toks.size() > 2) { logf(L_ERROR, "Too many \"*\" in Control; syntax is [Variable]*[scale with unit]"); return -1; }
            auto it = cont.find(toks[0]);
            double nscale = 1;
            if (it == cont.end()) {
                if (toks.size() > 1 || (sum_i == 0 && cont.count(toks[0]) == 0 && !isdigit((unsigned char)toks[0][0]) && toks[0][0] != '-' && toks[0][0] != '.')) {
                    logf(L_ERROR, "Variable %s not found in context in Control in xml config", toks[0].c_str());
                    return -1;
                }
                nscale = solver->units.alt(toks[0]);
                for (auto &f : fill) f += nscale * scale;
            } else {
                if (toks.size() == 2) nscale = solver->units.alt(toks[1]);
                for (size_t i = 0; i < fill.size(); i++) fill[i] += it->second[i] * nscale * scale;
            }
            sum_i++;
        }
        return 0;
    }
    int param(XmlNode *n) {
        std::string par = n->attr("name"), zone = n->attr("zone");
        if (!n->has_attr("value")) { logf(L_FATAL, "Setting value not specified in Param element"); return -2; }
        int zn = -1;
        if (!zone.empty()) {
            auto it = solver->setting_zones.find(zone);
            if (it == solver->setting_zones.end()) { logf(L_FATAL, "Unknown zone %s (found while setting parameter %s)", zone.c_str(), par.c_str()); return -1; }
            zn = it->second;
        }
        int zs = solver->model->find_zonesetting(par);
        if (zs < 0) { logf(L_FATAL, "Unknown setting %s", par.c_str()); return -3; }
        std::vector<double> val;
        if (get(context, n->attr("value"), 1, val)) return -1;
        val.resize(iters, val.empty() ? 0.0 : val.back());
        solver->domain->set_zone_series(zs, zn, val);
        return 0;
    }
    int csv(XmlNode *n) {
        if (!n->has_attr("file")) { logf(L_ERROR, "No file attribute in CSV in xml config"); return -1; }
        std::string nm = n->attr("file");
        std::ifstream f(nm);
        if (!f.is_open()) { logf(L_ERROR, "Could not open file CSV %s", nm.c_str()); return -1; }
        std::string line;
        if (!std::getline(f, line)) { logf(L_ERROR, "Empty file CSV %s", nm.c_str()); return -1; }
        std::vector<std::string> names;
        for (auto t : split(trim(line), ',')) {
            t = trim(t);
            if (!t.empty() && t[0] == '"') t = t.substr(1, t.size() - 2);
            names.push_back(t);
        }
        Context data;
        int rows = 0;
        while (std::getline(f, line)) {
            if (trim(line).empty()) continue;
            auto toks = split(trim(line), ',');
            if (toks.size() != names.size()) { logf(L_ERROR, "Wrong number of columns in CSV file %s", nm.c_str()); return -1; }
            for (size_t i = 0; i < toks.size(); i++) data[names[i]].push_back(solver->units.alt(trim(toks[i])));
            data["_index"].push_back(rows++);
        }
        std::vector<double> time;
        if (n->has_attr("Time")) { if (get(data, n->attr("Time"), 1, time)) return -1; }
        else { if (get(data, "_index", (double)iters / rows, time)) return -1; }
        int k = 0, max_k = (int)time.size() - 2;
        if (max_k < 0) { logf(L_ERROR, "Not enough records in CSV file %s", nm.c_str()); return -1; }
        for (int i = 0; i < iters; i++) {
            while (i > time[k + 1] && k < max_k) k++;
            double alpha = i < time[k] ? 0 : (i > time[k + 1] ? 1 : (1.0 * i - time[k]) / (time[k + 1] - time[k]));
            for (auto &nmj : names) context[nmj].push_back(data[nmj][k] * (1 - alpha) + data[nmj][k + 1] * alpha);
        }
        for (auto *c : n->children()) {
            if (c->name != "Param") { logf(L_ERROR, "Only Param allowed in %s in Control", n->name.c_str()); return -1; }
            if (param(c)) return -1;
        }
        return 0;
    }
  public:
    int init() override {
        Action::init();
        iters = round_half_away(every_iter);
        if (iters <= 0) { logf(L_ERROR, "Zero (or less) iterations in Control element in config"); return -1; }
        for (auto *c : node->children()) {
            if (c->name == "Param") { if (param(c)) return -1; }
            else if (c->name == "CSV") { if (csv(c)) return -1; }
            else { logf(L_ERROR, "Element %s not allowed in Control element in config", c->name.c_str()); return -1; }
        }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Control", ConControl)

// ------------------------------------------------------------------ callbacks ----
static Box region_from_attrs(Handler *h, const Box &tot) {        // VTK/HDF5 sub-region (reference cbVTK.cpp:16-41)
    Box reg = tot;
    Solver *s = h->solver;
    XmlNode *n = h->node;
    if (n->has_attr("dx")) reg.x0 = (int)s->units.alt(n->attr("dx"));
    if (reg.x0 < 0) { reg.x0 = reg.nx + reg.x0; reg.nx = reg.nx - reg.x0; }
    if (n->has_attr("dy")) reg.y0 = (int)s->units.alt(n->attr("dy"));
    if (reg.y0 < 0) { reg.y0 = reg.ny + reg.y0; reg.ny = reg.ny - reg.y0; }
    if (n->has_attr("dz")) reg.z0 = (int)s->units.alt(n->attr("dz"));
    if (reg.z0 < 0) { reg.z0 = reg.nz + reg.z0; reg.nz = reg.nz - reg.z0; }
    if (n->has_attr("nx")) reg.nx = (int)s->units.alt(n->attr("nx"));
    if (reg.nx < 0) reg.nx = tot.nx - reg.x0 + reg.nx;
    if (n->has_attr("ny")) reg.ny = (int)s->units.alt(n->attr("ny"));
    if (reg.ny < 0) reg.ny = tot.ny - reg.y0 + reg.ny;
    if (n->has_attr("nz")) reg.nz = (int)s->units.alt(n->attr("nz"));
    if (reg.nz < 0) reg.nz = tot.nz - reg.z0 + reg.nz;
    return reg.intersect(tot);
}

class CbVTK : public Callback {
    std::string nm;
    NameSet what;
    Box reg;
  public:
    int init() override {
        Callback::init();
        nm = node->attr("name", "VTK");
        what.add_from_string(node->attr("what", "all"));
        reg = region_from_attrs(this, solver->total);
        if (reg.size() == 0) { logf(L_FATAL, "VTK \"%s\" output has size 0", nm.c_str()); return -1; }
        return 0;
    }
    int do_it() override { return solver->write_vtk(nm, what, reg); }
};
TCLB_REGISTER_HANDLER("VTK", CbVTK)

// <HDF5 name what compress write_xdmf point_data precision dx..nz> (reference cbHDF5.cpp): one .h5 for the whole output
// region + an .xmf index.  `compress` (true or a level 1..9) stores every dataset chunked (slabs along z) with the deflate filter; `chunk` is accepted and ignored.
class CbHDF5 : public Callback {
    std::string nm;
    NameSet what;
    Box reg;
    bool xdmf = true, point = false, dbl = false;
    int deflate = 0;
  public:
    int init() override {
        Callback::init();
        nm = node->attr("name", "HDF5");
        what.add_from_string(node->attr("what", "all"));
        reg = region_from_attrs(this, solver->total);
        if (reg.size() == 0) { logf(L_FATAL, "HDF5 \"%s\" output has size 0", nm.c_str()); return -1; }
        if (node->has_attr("write_xdmf")) xdmf = node->attr_bool("write_xdmf");
        point = node->attr_bool("point_data");
        std::string pr = node->attr("precision", "");
        if (pr == "double") dbl = true;
        else if (pr == "float" || pr.empty()) dbl = false;
        else { logf(L_ERROR, "HDF5: precision should be float or double"); return -1; }
        if (node->has_attr("compress")) {      // true/false, or a deflate level 1..9 (reference cbHDF5.cpp:102-127: deflate on chunked datasets)
            const std::string c = node->attr("compress");
            deflate = (c.size() == 1 && c[0] >= '0' && c[0] <= '9') ? c[0] - '0' : (node->attr_bool("compress") ? 6 : 0);
        }
        return 0;
    }
    int do_it() override { return solver->write_hdf5(nm, what, reg, dbl, xdmf, point, deflate); }
};
TCLB_REGISTER_HANDLER("HDF5", CbHDF5)

class CbTXT : public Callback {
    std::string nm;
    NameSet what;
    bool gzip = false;
  public:
    int init() override {
        Callback::init();
        nm = node->attr("name", "TXT");
        what.add_from_string(node->attr("what", "all"));
        gzip = node->attr_bool("gzip");
        return 0;
    }
    int do_it() override { return solver->write_txt(nm, what, gzip); }
};
TCLB_REGISTER_HANDLER("TXT", CbTXT)

class CbBIN : public Callback {
    std::string nm;
  public:
    int init() override { Callback::init(); nm = node->attr("name", "BIN"); return 0; }
    int do_it() override { return solver->write_bin(nm); }
};
TCLB_REGISTER_HANDLER("BIN", CbBIN)

// <Image Iterations=N [name="Image"] [plane=z] [scale=s]>: the picture the reference's interactive window shows
// (velocity magnitude through its colour map, walls grey: src/LatticeContainer.inc.cpp.Rt:355-423, drawn into an OpenGL
// pixel buffer by src/gpu_anim.h) written as a binary PPM per call; there is no window system here, the colour kernel is
// the same headless.  One file for the whole lattice when every subdomain is in this process, else one piece per rank.
class CbImage : public Callback {
    std::string nm;
    int plane = 0;
    float scale = 1.0f;
  public:
    int init() override {
        Callback::init();
        nm = node->attr("name", "Image");
        plane = node->has_attr("plane") ? atoi(node->attr("plane").c_str()) : solver->domain->dec.tot[2] / 2;
        scale = node->has_attr("scale") ? (float)atof(node->attr("scale").c_str()) : 1.0f;
        return 0;
    }
    static int write_ppm(const std::string &file, int w, int h, const std::vector<unsigned char> &rgba) {
        make_parent_dirs(file);
        FILE *f = fopen(file.c_str(), "wb");
        if (!f) { logf(L_ERROR, "cannot write %s", file.c_str()); return -1; }
        fprintf(f, "P6\n%d %d\n255\n", w, h);
        std::vector<unsigned char> row((size_t)w * 3);
        for (int y = h - 1; y >= 0; y--) {                 // image rows from the top: y grows upwards in the lattice
            for (int x = 0; x < w; x++) for (int c = 0; c < 3; c++) row[(size_t)x * 3 + c] = rgba[((size_t)y * w + x) * 4 + c];
            fwrite(row.data(), 1, row.size(), f);
        }
        fclose(f);
        return 0;
    }
    int do_it() override {
        Domain *D = solver->domain.get();
        bool all_local = true;
        for (int r = 0; r < D->dec.nranks; r++) if (!D->lat[r]) all_local = false;
        const int W = D->dec.tot[0], H = D->dec.tot[1];
        std::vector<unsigned char> whole(all_local ? (size_t)W * H * 4 : 0, 0);
        for (int r : D->local) {
            Lattice *l = D->lat[r];
            const Box &b = D->dec.boxes[r];
            if (plane < b.z0 || plane >= b.z0 + b.nz) continue;
            std::vector<unsigned char> piece((size_t)l->nx * l->ny * 4);
            l->color(plane - b.z0, scale, piece.data());
            if (all_local) {
                for (int y = 0; y < l->ny; y++)
                    memcpy(&whole[((size_t)(y + b.y0) * W + b.x0) * 4], &piece[(size_t)y * l->nx * 4], (size_t)l->nx * 4);
            } else if (write_ppm(solver->out_iter_file(nm, ".ppm", r), l->nx, l->ny, piece)) return -1;
        }
        if (all_local) return write_ppm(solver->out_iter_collective_file(nm, ".ppm"), W, H, whole);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Image", CbImage)

class CbLog : public Callback {                      // reference cbLog.cpp:5-29
    std::string filename;
    int old_type = 0;
  public:
    int init() override {
        Callback::init();
        filename = solver->out_iter_file(node->attr("name", "Log"), ".csv", 0);
        solver->init_log(filename);
        old_type = solver->iter_type;
        solver->iter_type |= ITER_LASTGLOB;
        return 0;
    }
    int do_it() override { solver->write_log(filename); return 0; }
    int finish() override { solver->iter_type = old_type; return 0; }
};
TCLB_REGISTER_HANDLER("Log", CbLog)

class CbStop : public Callback {                     // reference cbStop.cpp:12-99
    enum { CHANGE, PERCENT, ABOVE, BELOW };
    std::vector<int> what, kind;
    std::vector<double> limit, old;
    int times = 1, score = 0, old_type = 0;
  public:
    int init() override {
        Callback::init();
        const tclb_model_t *d = solver->model->d;
        for (int g = 0; g < d->n_globals; g++) {
            static const char *suf[] = {"Change", "PercentChange", "Above", "Below"};
            for (int k = 0; k < 4; k++) {
                std::string a = std::string(d->globals[g].name) + suf[k];
                if (node->has_attr(a)) { what.push_back(g); kind.push_back(k); limit.push_back(atof(node->attr(a).c_str())); old.push_back(-12341234); }
            }
        }
        if (what.empty()) { logf(L_ERROR, "No *Change attribute in %s", node->name.c_str()); return -1; }
        if (node->has_attr("Times")) {
            times = atoi(node->attr("Times").c_str());
            if (times < 1) { logf(L_ERROR, "Minimal number for Times attribute is 1"); return -1; }
        }
        old_type = solver->iter_type;
        solver->iter_type |= ITER_LASTGLOB;
        return 0;
    }
    int do_it() override {
        int any = 0;
        for (size_t i = 0; i < what.size(); i++) {
            double v = solver->domain->globals[what[i]];
            switch (kind[i]) {
            case CHANGE: if (fabs(old[i] - v) > limit[i]) any++; logf(L_OUTPUT, "    change:      %4lg < %4lg", fabs(old[i] - v), limit[i]); break;
            case PERCENT: if (fabs((old[i] - v) / v) > limit[i]) any++; logf(L_OUTPUT, "    change:    %4lg%% < %4lg%%", 100.0 * fabs((old[i] - v) / v), 100.0 * limit[i]); break;
            case ABOVE: if (v < limit[i]) any++; logf(L_OUTPUT, "    limit:       %4lg > %4lg", v, limit[i]); break;
            case BELOW: if (v > limit[i]) any++; logf(L_OUTPUT, "    limit:       %4lg < %4lg", v, limit[i]); break;
            }
            old[i] = v;
        }
        score = any ? 0 : score + 1;
        logf(L_OUTPUT, "Stop criterium score: %d", score);
        if (score >= times) {
            logf(L_NOTICE, "Stop.");
            for (auto &o : old) o = -12341234;
            score = 0;
            return ITERATION_STOP;
        }
        return 0;
    }
    int finish() override { solver->iter_type = old_type; return 0; }
};
TCLB_REGISTER_HANDLER("Stop", CbStop)

class CbFailcheck : public Callback {                // reference cbFailcheck.cpp:45-93
    Box reg;
    bool active = false;
  public:
    int init() override {
        Callback::init();
        reg = solver->total;
        if (node->has_attr("dx")) reg.x0 = (int)solver->units.alt(node->attr("dx"));
        if (node->has_attr("dy")) reg.y0 = (int)solver->units.alt(node->attr("dy"));
        if (node->has_attr("dz")) reg.z0 = (int)solver->units.alt(node->attr("dz"));
        if (node->has_attr("nx")) reg.nx = (int)solver->units.alt(node->attr("nx"));
        if (node->has_attr("ny")) reg.ny = (int)solver->units.alt(node->attr("ny"));
        if (node->has_attr("nz")) reg.nz = (int)solver->units.alt(node->attr("nz"));
        return 0;
    }
    int do_it() override {
        if (active) return 0;
        active = true;
        NameSet what;
        what.add_from_string(node->attr("what", "all"));
        int ret = 0;
        if (solver->has_nan(what, reg)) {
            logf(L_NOTICE, "NaN value discovered. Executing final actions from the Failcheck element before full stop...");
            for (auto *c : node->children()) {
                HandlerPtr h = HandlerFactory::produce(c, solver);
                if (h) h->do_it();
            }
            logf(L_NOTICE, "Stopping due to NaN value");
            ret = ITERATION_STOP;
        }
        active = false;
        return ret;
    }
};
TCLB_REGISTER_HANDLER("Failcheck", CbFailcheck)

class CbAveraging : public Callback {                // <Average>: reset averaging window (reference cbAveraging.cpp)
    void reset() {
        const tclb_model_t *d = solver->model->d;
        for (int r : solver->domain->local) {
            Lattice *l = solver->domain->lat[r];
            l->reset_iter = l->iter;
            for (int f = 0; f < d->n_fields; f++) if (d->fields[f].average) l->fill_field(f, 0.0);
        }
    }
  public:
    int init() override { Callback::init(); reset(); return 0; }
    int do_it() override { reset(); return 0; }
};
TCLB_REGISTER_HANDLER("Average", CbAveraging)

class CbSample : public Callback {                   // <Sample what Iterations><Point dx dy dz/> (reference cbSample.cpp)
  public:
    int init() override {
        Callback::init();
        if (every_iter == 0) { logf(L_ERROR, "Iteration value in sampler should not be zero"); return -1; }
        NameSet what;
        what.add_from_string(node->attr("what", "all"));
        solver->sampler = std::make_unique<Sampler>(solver);
        for (auto *p : node->children()) {
            if (p->name != "Point") { logf(L_ERROR, "Unknown element in Sampler"); return -1; }
            solver->sampler->add_point((int)solver->units.alt(p->attr("dx"), 0), (int)solver->units.alt(p->attr("dy"), 0), (int)solver->units.alt(p->attr("dz"), 0));
        }
        solver->sampler->allocate(what, start_iter, (int)every_iter);
        solver->sampler->init_csv(solver->out_iter_file("Sampler", ".csv"));
        return 0;
    }
    int do_it() override {
        solver->sampler->write_history(solver->iter);
        solver->sampler->start_iter = solver->iter;
        return 0;
    }
    int finish() override { if (solver->sampler) solver->sampler->finish(); solver->sampler.reset(); return 0; }
};
TCLB_REGISTER_HANDLER("Sample", CbSample)

class CbKeep : public Callback {                     // penalty weight on a Global (reference cbKeep.cpp)
    int what = -1, what_inobj = -1, my_type = 0, old_type = 0;
    double thr = 0, force = 1;
  public:
    int init() override {
        Callback::init();
        if (!node->has_attr("What")) { logf(L_ERROR, "No What attribute in %s", node->name.c_str()); return -1; }
        std::string nm = node->attr("What");
        what = solver->model->find_global(nm);
        what_inobj = solver->model->find_zonesetting(nm + "InObj");
        if (what < 0 || what_inobj < 0) { logf(L_ERROR, "Unknown Global %s in %s", nm.c_str(), node->name.c_str()); return -1; }
        if (node->has_attr("Above")) { thr = atof(node->attr("Above").c_str()); my_type = 1; }
        else if (node->has_attr("Below")) { thr = atof(node->attr("Below").c_str()); my_type = -1; }
        else if (node->has_attr("Equal")) { thr = atof(node->attr("Equal").c_str()); my_type = 0; }
        else { logf(L_ERROR, "%s should have Above, Below or Equal attribute", node->name.c_str()); return -1; }
        if (node->has_attr("Force")) force = atof(node->attr("Force").c_str());
        old_type = solver->iter_type;
        solver->iter_type |= ITER_LASTGLOB;
        return 0;
    }
    int do_it() override {
        double v = (thr - solver->domain->globals[what]) * force, s = 0;
        if (my_type == -1) { if (v < 0) s = v; } else if (my_type == 0) s = v; else if (v > 0) s = v;
        solver->domain->set_zone_setting(what_inobj, -1, s);
        return 0;
    }
    int finish() override { solver->iter_type = old_type; return 0; }
};
TCLB_REGISTER_HANDLER("Keep", CbKeep)

class CbPID : public Callback {                      // PID: Global (integral=) -> zonal setting (control=) (reference cbPID.cpp:5-125)
    int what = -1, setting = -1, zone_number = -1, old_type = 0;
    double target = 0, itime = 1, dtime = 1, scale = 1, integral = 0, old_err = 0, DT = 1, sval = 0;
  public:
    int init() override {
        Callback::init();
        if (!node->has_attr("target")) { logf(L_FATAL, "No target value provided in the PID element"); return -1; }
        target = solver->units.alt(node->attr("target"));
        what = solver->model->find_global(node->attr("integral"));
        if (what < 0) { logf(L_FATAL, "Unknown global name in the PID element"); return -1; }
        std::string par = node->attr("control"), zone = node->attr("zone");
        if (!zone.empty()) {
            auto it = solver->setting_zones.find(zone);
            if (it == solver->setting_zones.end()) { logf(L_FATAL, "Unknown zone %s", zone.c_str()); return -1; }
            zone_number = it->second;
        }
        setting = solver->model->find_zonesetting(par);
        if (setting < 0) { logf(L_ERROR, "%s is not a valid zonal setting supplied for control in %s", par.c_str(), node->name.c_str()); return -1; }
        itime = solver->units.alt(node->attr("IntegrationTime", "1s"));
        dtime = solver->units.alt(node->attr("DerivativeTime", "1s"));
        scale = solver->units.alt(node->attr("scale", "1s"));
        DT = every_iter;
        sval = solver->domain->lat[solver->first_rank()]->zs.get(setting, zone_number < 0 ? 0 : zone_number, 0);
        old_type = solver->iter_type;
        solver->iter_type |= ITER_LASTGLOB;
        return 0;
    }
    int do_it() override {
        double val = solver->domain->globals[what], err = target - val;
        integral += (old_err + err) * DT / 2.0;
        double derivative = (err - old_err) / DT;
        double control = err + integral / itime + derivative * dtime;
        old_err = err;
        double nval = sval + control * scale;
        logf(L_OUTPUT, "PID: %lg --> %lg (error %lg); setting %lg --> %lg", val, target, err, sval, nval);
        solver->domain->set_zone_setting(setting, zone_number, nval);
        return 0;
    }
    int finish() override { solver->iter_type = old_type; return 0; }
};
TCLB_REGISTER_HANDLER("PID", CbPID)

class CbDumpSettings : public Callback {
    std::string name;
  public:
    int init() override { Callback::init(); name = node->attr("name", "ZonalSettings"); return 0; }
    int do_it() override {
        std::vector<std::string> names;
        for (int i = 0; i < solver->model->d->n_zonesettings; i++) names.push_back(solver->model->d->zonesettings[i].name);
        solver->domain->lat[solver->first_rank()]->zs.dump(solver->out_iter_file(name, ".csv"), names);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("DumpSettings", CbDumpSettings)

// <SaveCheckpoint Iterations= keep=N|all>: .pri image per rank + restart XML (reference cbSaveCheckpoint.cpp:5-104)
class CbSaveCheckpoint : public Callback {
    int keep = 1;
    std::queue<std::vector<std::string>> files;
    std::queue<std::string> restarts;
  public:
    int init() override {
        Callback::init();
        if (node->has_attr("keep")) {
            if (node->attr("keep") == "all") keep = 0;
            else { keep = atoi(node->attr("keep").c_str()); if (keep < 0) { logf(L_ERROR, "Keeping a negative number of checkpoints is not allowed; using 1"); keep = 1; } }
        }
        return 0;
    }
    int do_it() override {
        logf(L_OUTPUT, "writing checkpoint");
        std::string prefix = solver->out_iter_collective_file("checkpoint", "");
        std::string restart = solver->out_iter_collective_file("restart", ".xml");
        std::vector<std::string> written;
        solver->save_solution(prefix, &written);
        if (solver->is_root()) {
            XmlDocument rf;
            for (auto &k : solver->config.kids) rf.append(k->clone());
            XmlNode *cfg = rf.child("CLBConfig");
            XmlNode *lb = cfg->child("LoadBinary");
            if (!lb) {
                auto n = std::make_unique<XmlNode>(XmlType::Element, "LoadBinary");
                n->set_attr("file", prefix);
                cfg->insert_before(std::move(n), cfg->child("Solve"));
            } else { lb->remove_attr("file"); lb->set_attr("file", prefix); }
            rf.save_file(restart);
        }
        if (keep != 0) {
            files.push(written);
            restarts.push(restart);
            if ((int)files.size() > keep) {
                for (auto &f : files.front()) if (remove(f.c_str()) != 0) logf(L_ERROR, "Checkpoint file was not deleted: %s", f.c_str());
                files.pop();
                if (solver->is_root() && remove(restarts.front().c_str()) != 0) logf(L_ERROR, "Restart file was not deleted: %s", restarts.front().c_str());
                restarts.pop();
            }
        }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("SaveCheckpoint", CbSaveCheckpoint)

class CbSaveBinary : public Callback {               // reference cbSaveBinary.cpp
    std::string fn;
  public:
    int init() override {
        Callback::init();
        if (node->has_attr("file")) fn = solver->outpath + "_" + node->attr("file");
        else if (node->has_attr("filename")) fn = node->attr("filename");
        else fn = solver->out_iter_file("Save", "");
        return 0;
    }
    int do_it() override {
        if (node->has_attr("comp")) solver->save_comp(fn, node->attr("comp"));
        else solver->save_solution(fn);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("SaveBinary", CbSaveBinary)
class CbSaveMemoryDump : public CbSaveBinary {};
TCLB_REGISTER_HANDLER("SaveMemoryDump", CbSaveMemoryDump)

// <RemoteForceInterface integrator="NAME" [omega torque] [spawn="program args" | connect="/path/to/socket"]>: particle coupling.
// The reference connects to an external program over MPI inter-communicators (acRemoteForceInterface.cpp:7-110, MPMD.hpp).
// Here: integrator="SIMPLEPART" without spawn/connect runs the integrator in this process (csrc/host/particles.h; its
// configuration is the <SimplePart> child element, or the file given with --simplepart); with spawn= or connect= the
// integrator is ANOTHER PROGRAM reached over a local socket with the same sizes/particles/forces sequence
// (csrc/host/rfi_bridge.h; bin/tclb-integrator provides `simplepart` and `empty`).
class AcRemoteForceInterface : public Action {
  public:
    int init() override {
        Action::init();
        std::string name = node->attr("integrator");
        if (name.empty()) { logf(L_ERROR, "You must specify RemoteForceInterface integrator name"); return -1; }
        // TCLB_RFI_CONNECT=/path/to/socket: the launcher started the integrator program next to this one (the reference's
        // `mpirun main case.xml : simplepart part.xml`), the case file itself names no socket
        const char *env_connect = getenv("TCLB_RFI_CONNECT");
        if (node->has_attr("spawn") || node->has_attr("connect") || (env_connect && *env_connect)) {
            std::unique_ptr<RemoteIntegrator> ri(new RemoteIntegrator());
            const int workers = (int)solver->domain->local.size();
            bool ok;
            if (node->has_attr("spawn")) {
                logf(L_OUTPUT, "Connecting RFI to %s (starting: %s)", name.c_str(), node->attr("spawn").c_str());
                ok = ri->spawn(node->attr("spawn"), workers, node->attr_bool("omega"), node->attr_bool("torque"));
            } else {
                const std::string sock = node->has_attr("connect") ? node->attr("connect") : std::string(env_connect);
                logf(L_OUTPUT, "Connecting RFI to %s (socket %s)", name.c_str(), sock.c_str());
                ok = ri->connect_to(sock, workers, node->attr_bool("omega"), node->attr_bool("torque"));
            }
            if (!ok) { logf(L_ERROR, "RemoteForceInterface: %s", ri->error.c_str()); return -1; }
            solver->integrator.reset(ri.release());
        } else {
            if (name != "SIMPLEPART" && name != "simplepart") {
                logf(L_ERROR, "RemoteForceInterface: integrator %s is not built in (built in: SIMPLEPART); give spawn=\"program ...\" or connect=\"socket\" for an external one", name.c_str());
                return -1;
            }
            std::unique_ptr<XmlDocument> doc;
            const XmlNode *cfg = node->child("SimplePart");
            if (!solver->opt.simplepart_config.empty()) {
                doc.reset(new XmlDocument());
                doc->load_file(solver->opt.simplepart_config);
                cfg = doc->child("SimplePart");
            }
            if (!cfg) { logf(L_ERROR, "RemoteForceInterface: no <SimplePart> configuration (child element or --simplepart file)"); return -1; }
            logf(L_OUTPUT, "Connecting RFI to %s (in-process)", name.c_str());
            solver->integrator.reset(new SimplePart(cfg, solver->outpath + "_SP_Log.csv", node->attr_bool("omega"), node->attr_bool("torque")));
        }
        Solver *s = solver;
        auto box_of = [s](int r) {
            const Box &b = s->domain->dec.boxes[r];
            WorkerBox w;
            w.lower[0] = s->px + b.x0; w.upper[0] = s->px + b.x0 + b.nx;
            w.lower[1] = s->py + b.y0; w.upper[1] = s->py + b.y0 + b.ny;
            w.lower[2] = s->pz + b.z0; w.upper[2] = s->pz + b.z0 + b.nz;
            return w;
        };
        solver->domain->particles_in_hook = [s, box_of]() {
            s->integrator->begin_exchange();
            std::vector<double> rows;
            for (int r : s->domain->local) { s->integrator->particles_for(box_of(r), rows); s->domain->lat[r]->set_particles(rows); }
        };
        solver->domain->particles_out_hook = [s, box_of]() {
            std::vector<double> rows;
            for (int r : s->domain->local) { s->domain->lat[r]->get_particles(rows); s->integrator->forces_from(box_of(r), rows); }
            s->integrator->step();
        };
        return 0;
    }
};
TCLB_REGISTER_HANDLER("RemoteForceInterface", AcRemoteForceInterface)

// <ESYSParticle>: the reference writes a driver script for ESyS-Particle and starts it with MPI_Comm_spawn
// (src/Handlers/acESYSParticle.cpp:6-260).  Neither ESyS-Particle nor MPI exists here; an ESyS driver that speaks the bridge
// protocol (csrc/host/rfi_bridge.h) is started like any external integrator, so the element says how instead of failing mutely.
class AcESYSParticle : public Action {
  public:
    int init() override {
        Action::init();
        logf(L_ERROR, "<ESYSParticle>: ESyS-Particle is not part of this installation. Couple an external integrator with "
                      "<RemoteForceInterface integrator=\"ESYSPARTICLE\" spawn=\"program args\"/> (or connect=\"socket\"); "
                      "the program serves tclb::rfi_serve() or speaks the protocol documented in csrc/host/rfi_bridge.h");
        return -1;
    }
};
TCLB_REGISTER_HANDLER("ESYSParticle", AcESYSParticle)

// <RunPython> / <RunR> (reference src/Handlers/cbRunR.cpp): the code runs inside the embedding Python process through
// the script hook (tclb_solver_set_script_hook; tclb_b200.runtime.Solver installs one that exposes the same object
// surface: Solver.Settings/Fields/Quantities/Globals/Geometry/Actions/Info).  Without a hook (plain bin/tclb) the
// element fails with a message that says how to run the case instead.
class CbRunScript : public Callback {
  public:
    std::string code;
    int init() override {
        Callback::init();
        if (node->has_attr("file")) {
            FILE *f = fopen(node->attr("file").c_str(), "rb");
            if (!f) { logf(L_ERROR, "<%s>: cannot open %s", node->name.c_str(), node->attr("file").c_str()); return -1; }
            char buf[4096]; size_t n;
            while ((n = fread(buf, 1, sizeof(buf), f)) > 0) code.append(buf, n);
            fclose(f);
        } else code = node->text();
        if (!solver->script_hook) {
            logf(L_ERROR, "<%s> needs the Python coupling: run the case with `python -m tclb_b200.run <model> <case.xml>`", node->name.c_str());
            return -1;
        }
        return 0;                                     // without Iterations= the container runs do_it() once, now
    }
    int do_it() override {
        for (int r : solver->domain->local) solver->domain->lat[r]->sync();
        int rc = solver->script_hook(solver->script_user, node->name.c_str(), code.c_str(), solver->iter);
        if (rc != 0) { logf(L_ERROR, "<%s> script failed", node->name.c_str()); return -1; }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("RunPython", CbRunScript)
class CbRunR : public CbRunScript {};
TCLB_REGISTER_HANDLER("RunR", CbRunR)

// elements whose back-ends are not available in this build: recognised, fail with a clear message
// (the reference does the same in its #else branches, e.g. src/Handlers/cbHDF5.cpp:129-132)
class Unavailable : public Handler {
  public:
    int init() override {
        logf(L_ERROR, "<%s> is recognised but its back-end is not part of this build", node->name.c_str());
        return -1;
    }
};
class CbCatalyst : public Unavailable {};
TCLB_REGISTER_HANDLER("Catalyst", CbCatalyst)

}  // namespace
}  // namespace tclb
