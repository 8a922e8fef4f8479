// Optimisation / adjoint elements of the case file.
//
// Reference: src/Handlers/acUSAdjoint.cpp, acSAdjoint.cpp (<Adjoint type=>), acOptSolve.cpp (<OptSolve>), acObjective.cpp
// (<Objective>), acThresholdNow.cpp / acThreshold.cpp, acFDTest.cpp.  The adjoint itself is csrc/core/adjoint.h (dual
// numbers instead of Tapenade), so these handlers drive a tape + reverse sweep instead of the reference's ITER_ADJOINT
// iteration type; the design variables are the model's parameter fields (AddDensity(parameter=TRUE)) restricted to
// nodes flagged DesignSpace when the model has that node type.
#include "handler.h"
#include "design.h"
#include "../core/adjoint.h"
#include "../core/common.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>

namespace tclb {
namespace {

Lattice *the_lattice(Solver *s) {
    if (s->domain->local.size() != 1 || s->domain->dec.nranks != 1) fail("optimisation elements handle one subdomain in this version");
    return s->domain->lat[s->domain->local[0]];
}
size_t n_nodes(Lattice *l) { return (size_t)l->nx * l->ny * l->nz; }
void get_field_d(Solver *s, int f, std::vector<double> &v) {
    Lattice *l = the_lattice(s);
    const size_t n = n_nodes(l);
    v.resize(n);
    if (s->model->real_size() == 8) l->copy_field_out(f, v.data());
    else { std::vector<float> t(n); l->copy_field_out(f, t.data()); for (size_t i = 0; i < n; i++) v[i] = t[i]; }
}
void set_field_d(Solver *s, int f, const std::vector<double> &v) {
    Lattice *l = the_lattice(s);
    if (s->model->real_size() == 8) l->copy_field_in(f, v.data());
    else { std::vector<float> t(v.begin(), v.end()); l->copy_field_in(f, t.data()); }
    s->domain->copy_exchange(l->cur);
}
void get_adj_d(Solver *s, int f, std::vector<double> &v) {
    Lattice *l = the_lattice(s);
    const size_t n = n_nodes(l);
    v.resize(n);
    if (s->model->real_size() == 8) s->need_adjoint()->get_field(f, v.data());
    else { std::vector<float> t(n); s->need_adjoint()->get_field(f, t.data()); for (size_t i = 0; i < n; i++) v[i] = t[i]; }
}
void set_adj_d(Solver *s, int f, const std::vector<double> &v) {
    if (s->model->real_size() == 8) s->need_adjoint()->set_field(f, v.data());
    else { std::vector<float> t(v.begin(), v.end()); s->need_adjoint()->set_field(f, t.data()); }
}
// 1 where a design variable may change: nodes flagged DesignSpace, or everywhere when the model has no such type
std::vector<char> design_mask(Solver *s) {
    Lattice *l = the_lattice(s);
    const size_t n = n_nodes(l);
    std::vector<char> m(n, 1);
    const tclb_nodetype_t *t = s->model->find_nodetype("DesignSpace");
    if (!t) return m;
    std::vector<uint32_t> fl(n);
    Box b; b.x0 = b.y0 = b.z0 = 0; b.nx = l->nx; b.ny = l->ny; b.nz = l->nz;
    l->get_flags(b, fl.data());
    for (size_t i = 0; i < n; i++) m[i] = (fl[i] & t->mask) == t->value;
    return m;
}
int objective_index(Solver *s) { return s->model->find_global("Objective"); }

// The design seen by an optimisation element: every design handler on the solver's stack (reference vHandler.cpp:30-66);
// when the case file declares none, all parameter fields on the DesignSpace nodes (pushed for the element's lifetime).
struct DesignScope {
    Solver *s;
    bool pushed = false;
    int n = 0;
    explicit DesignScope(Solver *s_) : s(s_) {
        if (design_count(s) == 0) {
            HandlerPtr d = make_default_design(s);
            if (d) { s->hands.push_back(d); pushed = true; }
        }
        n = design_number_of_parameters(s);
    }
    ~DesignScope() {
        if (!pushed) return;
        for (size_t i = s->hands.size(); i-- > 0;)
            if (s->hands[i]->type() & HANDLER_DESIGN) { s->hands.erase(s->hands.begin() + i); break; }
    }
    std::vector<double> get(int what) { std::vector<double> v(std::max(1, n), 0.0); design_parameters(s, what, v.data()); v.resize(n); return v; }
    void set(const std::vector<double> &x) { std::vector<double> v = x; v.resize(std::max(1, n)); design_parameters(s, PAR_SET, v.data()); }
};

// objective (and gradient) of "children run from the saved state with the design x": the evaluation every optimiser shares
struct Evaluator {
    GenericAction *self;
    Solver *solver;
    DesignScope &design;
    Lattice *l;
    Adjoint *A;
    std::string state;
    int iter0, diter0, evals = 0;
    Evaluator(GenericAction *self_, DesignScope &d, const char *tag) : self(self_), solver(self_->solver), design(d) {
        l = the_lattice(solver);
        A = solver->need_adjoint();
        state = solver->outpath + "_" + tag + "_state.raw";
        iter0 = solver->iter; diter0 = solver->domain->iter;
        l->save_raw(state);
    }
    void restore() {
        l->load_raw(state);
        solver->iter = iter0; solver->domain->iter = diter0; l->iter = diter0;
    }
    int run(const std::vector<double> &x, double &obj, std::vector<double> *grad) {
        restore();
        design.set(x);
        A->drop_tape(); A->zero(); A->objective = 0;
        solver->recording = true;
        int rc = self->execute_internal();
        solver->recording = false;
        self->unstack();
        if (rc) return rc;
        obj = A->objective;
        evals++;
        solver->opt_iter++;
        if (grad) { A->reverse(); *grad = design.get(PAR_GRAD); }
        else A->drop_tape();
        return 0;
    }
    void finish(const std::vector<double> *x) {
        restore();
        if (x) design.set(*x);
        remove(state.c_str());
    }
};

// <Adjoint type="unsteady"> children </Adjoint>: children run with every iteration on the tape, then the reverse sweep.
// <Adjoint type="steady" Iterations="N"> children </Adjoint>: children run, then N adjoint iterations around the state
// they left; parameter fields then hold the gradient contribution of ONE iteration (what a converged flow needs).
class AcAdjoint : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        std::string type = node->attr("type");
        if (type.empty()) type = node->has_attr("Iterations") ? "steady" : "unsteady";
        Adjoint *A = solver->need_adjoint();
        // tape="n": keep at most n states; the sweep recomputes the iterations in between (the reference's snapshot levels)
        if (node->has_attr("tape")) A->set_tape_frames(std::max(2, atoi(node->attr("tape").c_str())));
        if (type == "unsteady" && solver->recording) {
            // inside an evaluation of <Optimize>/<FDTest> (reference case files put the adjoint run there): the tape is
            // already being recorded and the optimiser sweeps it back itself
            const int rc = execute_internal();
            unstack();
            return rc;
        }
        if (type == "unsteady") {
            A->drop_tape(); A->zero(); A->objective = 0;
            solver->recording = true;
            int rc = execute_internal();
            solver->recording = false;
            if (rc) return -1;
            const size_t n = A->tape_length();
            if (n == 0) { logf(L_ERROR, "No iterations done inside of unsteady Adjoint: nothing to do"); unstack(); return -1; }
            A->reverse();
            const int oi = objective_index(solver);
            if (oi >= 0) solver->domain->globals[oi] = A->objective;
            logf(L_OUTPUT, "%8d it adjoint swept back over %zu iterations, Objective %.10lg", solver->iter, n, A->objective);
        } else if (type == "steady") {
            if (execute_internal()) return -1;
            const int n = std::max(1, (int)every_iter);
            A->drop_tape(); A->zero();
            std::vector<int> pf = solver->parameter_fields();
            std::vector<std::vector<double>> before(pf.size()), after(pf.size());
            if (n > 1) A->steady(n - 1);
            for (size_t k = 0; k < pf.size(); k++) get_adj_d(solver, pf[k], before[k]);
            A->steady(1);
            for (size_t k = 0; k < pf.size(); k++) {
                get_adj_d(solver, pf[k], after[k]);
                for (size_t i = 0; i < after[k].size(); i++) after[k][i] -= before[k][i];
                set_adj_d(solver, pf[k], after[k]);
            }
            logf(L_OUTPUT, "%8d it steady adjoint: %d iterations", solver->iter, n);
        } else { logf(L_ERROR, "Unknown type of adjoint in xml: %s", type.c_str()); return -1; }
        unstack();
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Adjoint", AcAdjoint)

// <Objective Drag="1" PressureLoss="-0.5"/>: weights of SUM Globals in the objective (sets <Global>InObj in every zone)
class AcObjective : public Action {
  public:
    // <Objective NAME="weight" .../>: NAME is a Global or an objective the model declared with AddObjective (an expression
    // of the Globals).  The weighted objectives are linearised at the CURRENT values of the Globals: the zonal settings
    // <Global>InObj become sum_k weight_k * d(objective_k)/d(Global), and the Objective global their weighted sum
    // (reference src/Handlers/acObjective.cpp:4-38, generated OF_<name> functions src/Lists.cpp.Rt:17-34).
    int init() override {
        Action::init();
        const tclb_model_t *d = solver->model->d;
        const int n = d->n_globals;
        std::vector<double> glob(n), grad(n), in_obj(n, 0.0);
        for (int g = 0; g < n; g++) glob[g] = solver->domain->globals[g];
        double obj = 0;
        int nset = 0;
        for (int k = 0; k < d->n_objectives; k++) {
            const std::string nm = d->objectives[k].name;
            if (nm == "Objective" || !node->has_attr(nm)) continue;
            const double w = atof(node->attr(nm).c_str());
            double val = 0;
            std::fill(grad.begin(), grad.end(), 0.0);
            d->objectives[k].fun(glob.data(), &val, grad.data());
            obj += w * val;
            for (int g = 0; g < n; g++) in_obj[g] += w * grad[g];
            if (w != 0.0) nset++;
        }
        for (int g = 0; g < d->sum_globals; g++) {
            const std::string nm = d->globals[g].name;
            if (nm == "Objective" || solver->model->find_zonesetting(nm + "InObj") < 0) continue;
            solver->set_param(nm + "InObj", strf("%.17g", in_obj[g]), "", true, "");
        }
        const int oi = objective_index(solver);
        if (oi >= 0) solver->domain->globals[oi] = obj;
        if (nset == 0) logf(L_WARNING, "<Objective> names no objective of this model: the objective is zero");
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Objective", AcObjective)

// <ThresholdNow Level="0.5"/>: design variables become 0/1
class AcThresholdNow : public Action {
  public:
    int init() override {
        Action::init();
        const double level = node->has_attr("Level") ? atof(node->attr("Level").c_str()) : 0.5;
        DesignScope design(solver);
        if (design.n == 0) { logf(L_ERROR, "No parameters defined!"); return -1; }
        std::vector<double> x = design.get(PAR_GET);
        for (double &v : x) v = v > level ? 1.0 : 0.0;
        design.set(x);
        logf(L_OUTPUT, "Parameters: %d", design.n);
        if (solver->model->find_setting("Threshold") >= 0) solver->set_param("Threshold", node->has_attr("Level") ? node->attr("Level") : "0.5", "", true, "");
        return 0;
    }
};
TCLB_REGISTER_HANDLER("ThresholdNow", AcThresholdNow)

// <Threshold Levels="N"> children </Threshold>: run the children once per threshold level i/N on a thresholded copy of
// the design, then put the design back
class AcThreshold : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        const int levels = node->has_attr("Levels") ? atoi(node->attr("Levels").c_str()) : 5;
        DesignScope design(solver);
        if (design.n == 0) { logf(L_ERROR, "No parameters defined!"); return -1; }
        const std::vector<double> orig = design.get(PAR_GET);
        for (int i = 0; i <= levels; i++) {
            const double level = (double)i / levels;
            std::vector<double> x = orig;
            for (double &v : x) v = v > level ? 1.0 : 0.0;
            design.set(x);
            logf(L_OUTPUT, "Threshold level %lg", level);
            if (execute_internal()) return -1;
            unstack();
        }
        design.set(orig);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Threshold", AcThreshold)

// <OptSolve Iterations="N" Descent="d" [direction="min"]>: one-shot optimisation, every iteration = one primal iteration (with the
// Objective), one adjoint iteration around the new state and a descent step of the design variables,
// clipped to [0,1] (reference ITER_OPT, src/cuda.cu.Rt:232-250: w += wb*Descent)
class AcOptSolve : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        if (execute_internal()) return -1;
        // step: Descent= attribute, else the model's Descent setting (<Param name="Descent">), else 1
        double descent = 1.0;
        const int ds = solver->model->find_setting("Descent");
        if (node->has_attr("Descent")) descent = atof(node->attr("Descent").c_str());
        else if (ds >= 0 && the_lattice(solver)->get_setting(ds) != 0.0) descent = the_lattice(solver)->get_setting(ds);
        // the reference's descent step is w += wb*Descent: the objective is MAXIMISED (DragInObj = -1 minimises the drag)
        const bool maximise = node->attr("direction") != "min";
        std::vector<int> pf = solver->parameter_fields();
        if (pf.empty()) { logf(L_ERROR, "No parameters defined!"); return -1; }
        std::vector<char> mask = design_mask(solver);
        Adjoint *A = solver->need_adjoint();
        A->drop_tape(); A->zero();
        const int oi = objective_index(solver);
        int ret = run_loop([&](int steps, int, bool) {
            std::vector<double> p, g, zero;
            for (int i = 0; i < steps; i++) {
                solver->domain->iterate(1, false, true);
                A->steady(1);
                for (int f : pf) {
                    get_field_d(solver, f, p);
                    get_adj_d(solver, f, g);
                    for (size_t j = 0; j < p.size(); j++)
                        if (mask[j]) p[j] = std::min(1.0, std::max(0.0, p[j] + (maximise ? descent : -descent) * g[j]));
                    set_field_d(solver, f, p);
                    zero.assign(p.size(), 0.0);
                    set_adj_d(solver, f, zero);
                }
            }
            if (oi >= 0) logf(L_OUTPUT, "%8d it OptSolve Objective %.10lg", solver->iter, solver->domain->globals[oi]);
        });
        unstack();
        return ret;
    }
};
TCLB_REGISTER_HANDLER("OptSolve", AcOptSolve)

// <FDTest [Samples="K"] [parameters="a:b"] [Epsilon="e"]> children </FDTest>: gradient of the objective accumulated by the
// children with respect to K components of the design vector, from the adjoint and from central differences (children
// re-run from the same state).  Reference: src/Handlers/acFDTest.cpp (there a sweep over step sizes).
class AcFDTest : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        DesignScope design(solver);
        if (design.n == 0) { logf(L_ERROR, "Error: No parameters defined!"); return -1; }
        int samples = node->has_attr("Samples") ? atoi(node->attr("Samples").c_str()) : 5;
        const double eps = node->has_attr("Epsilon") ? atof(node->attr("Epsilon").c_str()) : (solver->model->real_size() == 8 ? 1e-6 : 1e-3);
        int p0 = 0, pn = design.n;
        if (node->has_attr("parameters")) {       // "a", "a:", ":b", "a:b" (reference syntax)
            const std::string par = node->attr("parameters");
            const size_t i = par.find(':');
            if (i == std::string::npos) { p0 = atoi(par.c_str()); pn = 1; }
            else if (i == 0) pn = atoi(par.substr(1).c_str()) + 1;
            else if (i == par.size() - 1) { p0 = atoi(par.substr(0, i).c_str()); pn = design.n - p0; }
            else { p0 = atoi(par.substr(0, i).c_str()); pn = atoi(par.substr(i + 1).c_str()) - p0 + 1; }
            if (p0 < 0 || pn < 1 || p0 + pn > design.n) { logf(L_ERROR, "Parameter out of bounds in FDTest"); return -1; }
            if (!node->has_attr("Samples")) samples = pn;
        } else if (!node->has_attr("Samples")) samples = pn <= 16 ? pn : 5;     // every parameter of a small design, else a sample
        logf(L_NOTICE, "Parameters in test: %d:%d (%d) from %d", p0, p0 + pn - 1, pn, design.n);
        Evaluator ev(this, design, "fdtest");
        const std::vector<double> x0 = design.get(PAR_GET);
        double J0 = 0;
        std::vector<double> grad;
        if (ev.run(x0, J0, &grad)) return -1;
        std::mt19937 rng(12345);
        double worst = 0;
        logf(L_OUTPUT, "FDTest: objective %.12lg, %d samples, epsilon %lg", J0, samples, eps);
        for (int k = 0; k < samples; k++) {
            const int i = (samples >= pn) ? p0 + (k % pn) : p0 + (int)(rng() % pn);
            double Jp = 0, Jm = 0;
            std::vector<double> x = x0;
            x[i] = x0[i] + eps;
            if (ev.run(x, Jp, nullptr)) return -1;
            x[i] = x0[i] - eps;
            if (ev.run(x, Jm, nullptr)) return -1;
            const double fd = (Jp - Jm) / (2 * eps);
            const double err = std::fabs(fd - grad[i]) / std::max(1e-300, std::max(std::fabs(fd), std::fabs(grad[i])));
            worst = std::max(worst, err);
            logf(L_OUTPUT, "FDTest: parameter %d  adjoint %.10lg  finite difference %.10lg  relative difference %.3lg", i, grad[i], fd, err);
        }
        ev.finish(&x0);
        logf(L_OUTPUT, "FDTest: largest relative difference %.3lg", worst);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("FDTest", AcFDTest)

// <Optimize [method="MMA|LBFGS|NELDERMEAD|GD"] MaxEvaluations="N" [Material="more|less"] [XAbsTolerance RelTolerance
// AbsTolerance StopAtValue] [Step="s"]> children </Optimize>: MAXIMISE the objective accumulated by the children over
// the design vector within its bounds.  The reference hands this to NLopt (src/Handlers/acOptimize.cpp.Rt: LD_MMA,
// LD_LBFGS, LN_COBYLA, LN_NELDERMEAD, GN_DIRECT_L, GN_ESCH); NLopt is not available here, so the methods are built in:
//   MMA         method of moving asymptotes (Svanberg 1987), separable convex sub-problems solved in closed form, the
//               optional Material constraint (sum of the parameters not below / above its start value) by bisection
//               on its multiplier
//   LBFGS       limited-memory BFGS (two-loop recursion, 8 pairs) on the box, projected steps with Armijo backtracking
//   NELDERMEAD  derivative-free simplex (also answers to COBYLA), box by projection
//   GD          projected gradient ascent with step control (the method of round 1)
class AcOptimize : public GenericAction {
    std::vector<double> lo_, hi_;
    int max_eval_ = 20;
    double xtol_ = 0, ftol_rel_ = 0, ftol_abs_ = 0, stopval_ = HUGE_VAL;
    void project(std::vector<double> &x) const { for (size_t i = 0; i < x.size(); i++) x[i] = std::min(hi_[i], std::max(lo_[i], x[i])); }
    bool converged(double f_old, double f_new, const std::vector<double> &x_old, const std::vector<double> &x_new) const {
        if (f_new >= stopval_) return true;
        if (ftol_abs_ > 0 && std::fabs(f_new - f_old) < ftol_abs_) return true;
        if (ftol_rel_ > 0 && std::fabs(f_new - f_old) < ftol_rel_ * std::max(std::fabs(f_new), std::fabs(f_old))) return true;
        if (xtol_ > 0) {
            double d = 0;
            for (size_t i = 0; i < x_new.size(); i++) d = std::max(d, std::fabs(x_new[i] - x_old[i]));
            if (d < xtol_) return true;
        }
        return false;
    }
    int run_gd(Evaluator &ev, std::vector<double> &x, double &best, double step) {
        std::vector<double> g, trial, gt;
        if (ev.run(x, best, &g)) return -1;
        logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg", ev.evals, best);
        while (ev.evals < max_eval_) {
            double gmax = 0;
            for (size_t i = 0; i < g.size(); i++) if ((g[i] > 0 && x[i] < hi_[i]) || (g[i] < 0 && x[i] > lo_[i])) gmax = std::max(gmax, std::fabs(g[i]));
            if (gmax == 0) { logf(L_OUTPUT, "Optimize: zero projected gradient, stopping"); break; }
            bool improved = false;
            while (ev.evals < max_eval_ && !improved) {
                trial = x;
                for (size_t i = 0; i < x.size(); i++) trial[i] += step * g[i] / gmax * (hi_[i] - lo_[i] > 0 && std::isfinite(hi_[i] - lo_[i]) ? (hi_[i] - lo_[i]) : 1.0);
                project(trial);
                double obj = 0;
                if (ev.run(trial, obj, &gt)) return -1;
                const bool stop = obj > best && converged(best, obj, x, trial);
                if (obj > best) { best = obj; x = trial; g = gt; improved = true; step *= 1.5; }
                else step *= 0.5;
                logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg %s (step %lg)", ev.evals, obj, improved ? "accepted" : "rejected", step);
                if (stop) return 0;
                if (step < 1e-10) return 0;
            }
            if (!improved) break;
        }
        return 0;
    }
    int run_lbfgs(Evaluator &ev, std::vector<double> &x, double &best) {
        const size_t n = x.size(), M = 8;
        std::vector<double> g;
        if (ev.run(x, best, &g)) return -1;
        logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg", ev.evals, best);
        std::vector<std::vector<double>> S, Y;     // pairs for the MINIMISATION of -objective
        while (ev.evals < max_eval_) {
            // free variables: not pinned to a bound by the gradient
            std::vector<char> fr(n);
            size_t nfree = 0;
            for (size_t i = 0; i < n; i++) { fr[i] = !((x[i] <= lo_[i] && g[i] < 0) || (x[i] >= hi_[i] && g[i] > 0)); nfree += fr[i]; }
            if (nfree == 0) break;
            std::vector<double> q(n), alpha(S.size());
            for (size_t i = 0; i < n; i++) q[i] = fr[i] ? -g[i] : 0.0;       // gradient of -J
            double gn = 0;
            for (double v : q) gn += v * v;
            if (gn == 0) break;
            for (size_t k = S.size(); k-- > 0;) {
                double sy = 0, sq = 0;
                for (size_t i = 0; i < n; i++) if (fr[i]) { sy += S[k][i] * Y[k][i]; sq += S[k][i] * q[i]; }
                alpha[k] = sy > 0 ? sq / sy : 0.0;
                for (size_t i = 0; i < n; i++) if (fr[i]) q[i] -= alpha[k] * Y[k][i];
            }
            double gamma = 1.0;
            if (!S.empty()) {
                double sy = 0, yy = 0;
                for (size_t i = 0; i < n; i++) if (fr[i]) { sy += S.back()[i] * Y.back()[i]; yy += Y.back()[i] * Y.back()[i]; }
                if (sy > 0 && yy > 0) gamma = sy / yy;
            } else {
                // first step: the largest change of a variable is a tenth of its range
                double m = 0;
                for (size_t i = 0; i < n; i++) if (fr[i]) m = std::max(m, std::fabs(q[i]) / (std::isfinite(hi_[i] - lo_[i]) ? hi_[i] - lo_[i] : 1.0));
                gamma = m > 0 ? 0.1 / m : 1.0;
            }
            for (double &v : q) v *= gamma;
            for (size_t k = 0; k < S.size(); k++) {
                double sy = 0, yq = 0;
                for (size_t i = 0; i < n; i++) if (fr[i]) { sy += S[k][i] * Y[k][i]; yq += Y[k][i] * q[i]; }
                const double beta = sy > 0 ? yq / sy : 0.0;
                for (size_t i = 0; i < n; i++) if (fr[i]) q[i] += (alpha[k] - beta) * S[k][i];
            }
            // direction d = -H grad(-J) = -q ; ascent direction check
            double slope = 0;
            for (size_t i = 0; i < n; i++) slope += -q[i] * g[i];
            if (!(slope > 0)) { for (size_t i = 0; i < n; i++) q[i] = fr[i] ? -g[i] * gamma : 0.0; S.clear(); Y.clear(); slope = 0; for (size_t i = 0; i < n; i++) slope += -q[i] * g[i]; }
            double t = 1.0, obj = 0;
            std::vector<double> trial, gt;
            bool ok = false;
            while (ev.evals < max_eval_) {
                trial = x;
                for (size_t i = 0; i < n; i++) trial[i] -= t * q[i];
                project(trial);
                if (ev.run(trial, obj, &gt)) return -1;
                double lin = 0;
                for (size_t i = 0; i < n; i++) lin += g[i] * (trial[i] - x[i]);
                logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg (L-BFGS step %lg)", ev.evals, obj, t);
                if (obj >= best + 1e-4 * lin && obj > best) { ok = true; break; }
                t *= 0.5;
                if (t < 1e-10) break;
            }
            if (!ok) break;
            std::vector<double> sv(n), yv(n);
            for (size_t i = 0; i < n; i++) { sv[i] = trial[i] - x[i]; yv[i] = -(gt[i] - g[i]); }
            double sy = 0;
            for (size_t i = 0; i < n; i++) sy += sv[i] * yv[i];
            if (sy > 1e-300) { S.push_back(sv); Y.push_back(yv); if (S.size() > M) { S.erase(S.begin()); Y.erase(Y.begin()); } }
            const bool stop = converged(best, obj, x, trial);
            x = trial; g = gt; best = obj;
            if (stop) break;
        }
        return 0;
    }
    int run_mma(Evaluator &ev, std::vector<double> &x, double &best, int material, double mat0) {
        // maximise J  <=>  minimise f = -J.  Asymptotes L_i < x_i < U_i; sub-problem  min sum p_i/(U_i-x_i) + q_i/(x_i-L_i)
        // [+ lambda * (+-sum x_i)] over the move limits: x_i = (sqrt(p_i) L_i + sqrt(q_i) U_i) / (sqrt(p_i) + sqrt(q_i)), clipped
        const size_t n = x.size();
        std::vector<double> g, xold1 = x, xold2 = x, L(n), U(n), range(n);
        for (size_t i = 0; i < n; i++) range[i] = std::isfinite(hi_[i] - lo_[i]) && hi_[i] > lo_[i] ? hi_[i] - lo_[i] : 1.0;
        double f = 0;
        if (ev.run(x, f, &g)) return -1;
        best = f;
        std::vector<double> xbest = x;
        logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg", ev.evals, f);
        int it = 0;
        while (ev.evals < max_eval_) {
            for (size_t i = 0; i < n; i++) {
                if (it < 2) { L[i] = x[i] - 0.5 * range[i]; U[i] = x[i] + 0.5 * range[i]; }
                else {
                    const double osc = (x[i] - xold1[i]) * (xold1[i] - xold2[i]);
                    const double fac = osc < 0 ? 0.7 : (osc > 0 ? 1.2 : 1.0);
                    double l = x[i] - fac * (xold1[i] - L[i]), u = x[i] + fac * (U[i] - xold1[i]);
                    l = std::min(std::max(l, x[i] - 10 * range[i]), x[i] - 0.01 * range[i]);
                    u = std::max(std::min(u, x[i] + 10 * range[i]), x[i] + 0.01 * range[i]);
                    L[i] = l; U[i] = u;
                }
            }
            std::vector<double> p(n), q(n), a(n), b(n);
            for (size_t i = 0; i < n; i++) {
                const double df = -g[i];                                  // gradient of f = -J
                const double ux = U[i] - x[i], xl = x[i] - L[i];
                const double reg = 1e-3 * std::fabs(df) + 1e-6 / range[i];
                p[i] = ux * ux * (std::max(df, 0.0) + reg);
                q[i] = xl * xl * (std::max(-df, 0.0) + reg);
                a[i] = std::max(lo_[i], std::max(L[i] + 0.1 * xl, x[i] - 0.5 * range[i]));
                b[i] = std::min(hi_[i], std::min(U[i] - 0.1 * ux, x[i] + 0.5 * range[i]));
            }
            auto solve = [&](double lam, std::vector<double> &xn) {
                // material constraint c(x) = s * (mat0 - sum x) <= 0 with s = +1 ("more") or -1 ("less"): adds -s*lam to df_i
                xn.resize(n);
                for (size_t i = 0; i < n; i++) {
                    double pi = p[i], qi = q[i];
                    const double ux = U[i] - x[i], xl = x[i] - L[i];
                    if (material > 0) qi += lam * xl * xl; else if (material < 0) pi += lam * ux * ux;
                    const double sp = std::sqrt(pi), sq = std::sqrt(qi);
                    double v = (sp * L[i] + sq * U[i]) / (sp + sq);
                    xn[i] = std::min(b[i], std::max(a[i], v));
                }
            };
            std::vector<double> xn;
            solve(0.0, xn);
            if (material != 0) {
                auto viol = [&](const std::vector<double> &v) { double s = 0; for (double t : v) s += t; return material > 0 ? mat0 - s : s - mat0; };
                if (viol(xn) > 1e-3) {
                    double lo = 0, hi = 1;
                    std::vector<double> xt;
                    for (int k = 0; k < 60; k++) { solve(hi, xt); if (viol(xt) <= 1e-3) break; hi *= 4; }
                    for (int k = 0; k < 60; k++) { const double mid = 0.5 * (lo + hi); solve(mid, xt); if (viol(xt) > 1e-3) lo = mid; else hi = mid; }
                    solve(hi, xn);
                }
            }
            xold2 = xold1; xold1 = x;
            double fn = 0;
            std::vector<double> gn;
            if (ev.run(xn, fn, &gn)) return -1;
            logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg (MMA iteration %d)", ev.evals, fn, it + 1);
            const bool stop = converged(f, fn, x, xn);
            x = xn; g = gn; f = fn;
            if (f > best) { best = f; xbest = x; }
            it++;
            if (stop) break;
        }
        x = xbest;
        return 0;
    }
    int run_neldermead(Evaluator &ev, std::vector<double> &x, double &best) {
        const size_t n = x.size();
        std::vector<std::vector<double>> P(n + 1, x);
        std::vector<double> F(n + 1, 0.0);
        for (size_t k = 0; k < n; k++) {
            const double r = std::isfinite(hi_[k] - lo_[k]) && hi_[k] > lo_[k] ? hi_[k] - lo_[k] : 1.0;
            P[k + 1][k] += (P[k + 1][k] + 0.1 * r <= hi_[k]) ? 0.1 * r : -0.1 * r;
        }
        for (size_t k = 0; k <= n && ev.evals < max_eval_; k++) { project(P[k]); if (ev.run(P[k], F[k], nullptr)) return -1; }
        auto order = [&]() { std::vector<size_t> o(n + 1); std::iota(o.begin(), o.end(), (size_t)0); std::sort(o.begin(), o.end(), [&](size_t a, size_t b) { return F[a] > F[b]; }); return o; };
        while (ev.evals < max_eval_) {
            auto o = order();
            const size_t w = o[n], b0 = o[0], w2 = o[n - (n > 0 ? 1 : 0)];
            std::vector<double> c(n, 0.0);
            for (size_t k = 0; k < n; k++) for (size_t i = 0; i < n; i++) c[i] += P[o[k]][i] / n;
            auto point = [&](double t) { std::vector<double> v(n); for (size_t i = 0; i < n; i++) v[i] = c[i] + t * (c[i] - P[w][i]); project(v); return v; };
            std::vector<double> xr = point(1.0);
            double fr = 0;
            if (ev.run(xr, fr, nullptr)) return -1;
            if (fr > F[b0] && ev.evals < max_eval_) {
                std::vector<double> xe = point(2.0);
                double fe = 0;
                if (ev.run(xe, fe, nullptr)) return -1;
                if (fe > fr) { P[w] = xe; F[w] = fe; } else { P[w] = xr; F[w] = fr; }
            } else if (fr > F[w2]) { P[w] = xr; F[w] = fr; }
            else if (ev.evals < max_eval_) {
                std::vector<double> xc = point(fr > F[w] ? 0.5 : -0.5);
                double fc = 0;
                if (ev.run(xc, fc, nullptr)) return -1;
                if (fc > std::max(F[w], fr)) { P[w] = xc; F[w] = fc; }
                else for (size_t k = 1; k <= n && ev.evals < max_eval_; k++) {      // shrink towards the best
                    for (size_t i = 0; i < n; i++) P[o[k]][i] = P[b0][i] + 0.5 * (P[o[k]][i] - P[b0][i]);
                    if (ev.run(P[o[k]], F[o[k]], nullptr)) return -1;
                }
            }
            auto o2 = order();
            logf(L_OUTPUT, "Optimize: evaluation %d best objective %.10lg (simplex)", ev.evals, F[o2[0]]);
            if (converged(F[o2[n]], F[o2[0]], P[o2[n]], P[o2[0]]) && F[o2[0]] - F[o2[n]] >= 0 && (ftol_abs_ > 0 || ftol_rel_ > 0 || xtol_ > 0)) break;
        }
        auto o = order();
        x = P[o[0]]; best = F[o[0]];
        return 0;
    }
  public:
    int init() override {
        GenericAction::init();
        std::string method = node->has_attr("method") ? node->attr("method") : (node->has_attr("Method") ? node->attr("Method") : "MMA");
        if (method == "COBYLA") { logf(L_NOTICE, "Optimize: COBYLA is served by the built-in simplex method"); method = "NELDERMEAD"; }
        if (method != "GD" && method != "LBFGS" && method != "MMA" && method != "NELDERMEAD") {
            logf(L_ERROR, "Unknown Method in Optimize: %s (built in: MMA, LBFGS, NELDERMEAD/COBYLA, GD; NLopt's DIRECT_L and ESCH are not available)", method.c_str());
            return -1;
        }
        max_eval_ = node->has_attr("MaxEvaluations") ? atoi(node->attr("MaxEvaluations").c_str()) : 20;
        auto crit = [&](const char *name, double &dst, bool positive) -> int {
            if (!node->has_attr(name)) return 0;
            dst = atof(node->attr(name).c_str());
            if (positive && !(dst > 0)) { logf(L_ERROR, "%s in Optimize have to be above 0", name); return -1; }
            return 0;
        };
        if (crit("XAbsTolerance", xtol_, true) || crit("RelTolerance", ftol_rel_, true) || crit("AbsTolerance", ftol_abs_, true) || crit("StopAtValue", stopval_, false)) return -1;
        const double step = node->has_attr("Step") ? atof(node->attr("Step").c_str()) : 0.1;
        DesignScope design(solver);
        if (design.n == 0) { logf(L_ERROR, "Error: No parameters defined!"); return -1; }
        logf(L_NOTICE, "Parameters in optimization: %d", design.n);
        lo_ = design.get(PAR_LOWER);
        hi_ = design.get(PAR_UPPER);
        std::vector<double> x = design.get(PAR_GET);
        project(x);
        int material = 0;
        double mat0 = 0;
        if (node->has_attr("Material")) {
            const std::string m = node->attr("Material");
            if (m == "more") material = +1; else if (m == "less") material = -1;
            else { logf(L_ERROR, "Material attribute in Optimize should be \"more\" or \"less\""); return -1; }
            for (double v : x) mat0 += v;
            if (method != "MMA") logf(L_WARNING, "Optimize: the Material constraint is honoured by method MMA only");
        }
        Evaluator ev(this, design, "optimize");
        double best = 0;
        int rc = 0;
        logf(L_OUTPUT, "Setting the algorithm to %s", method.c_str());
        if (method == "GD") rc = run_gd(ev, x, best, step);
        else if (method == "LBFGS") rc = run_lbfgs(ev, x, best);
        else if (method == "MMA") rc = run_mma(ev, x, best, material, mat0);
        else rc = run_neldermead(ev, x, best);
        if (rc) return -1;
        ev.finish(&x);          // the solver is left at the start state with the best design
        logf(L_OUTPUT, "Optimize: best objective %.10lg after %d evaluations", best, ev.evals);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Optimize", AcOptimize)

// <Andersen Directions="m" [Times="k"] [Eps="e"]> children </Andersen>: Anderson mixing of the fixed-point iteration
// "state -> children(state)" (reference src/Handlers/acAndersen.cpp).  The last m residuals r_i = children(x_i) - x_i are
// kept orthonormal (Gram-Schmidt carries the states x_i and the sum-of-weights p_i along); the next state is the
// combination with the smallest residual norm whose original weights sum to one: x = sum_i x_i p_i / sum_i p_i^2.
class AcAndersen : public GenericAction {
    struct Entry { std::vector<double> x, r; double p; };
    void get_state(std::vector<double> &v) {
        v.clear();
        const int nf = solver->model->d->n_fields;
        for (int rk : solver->domain->local) {
            Lattice *l = solver->domain->lat[rk];
            const size_t n = n_nodes(l);
            for (int f = 0; f < nf; f++) {
                const size_t o = v.size();
                v.resize(o + n);
                if (solver->model->real_size() == 8) l->copy_field_out(f, v.data() + o);
                else { std::vector<float> t(n); l->copy_field_out(f, t.data()); for (size_t i = 0; i < n; i++) v[o + i] = t[i]; }
            }
        }
    }
    void set_state(const std::vector<double> &v) {
        const int nf = solver->model->d->n_fields;
        size_t o = 0;
        for (int rk : solver->domain->local) {
            Lattice *l = solver->domain->lat[rk];
            const size_t n = n_nodes(l);
            for (int f = 0; f < nf; f++, o += n) {
                if (solver->model->real_size() == 8) l->copy_field_in(f, v.data() + o);
                else { std::vector<float> t(v.begin() + o, v.begin() + o + n); l->copy_field_in(f, t.data()); }
            }
        }
        if ((int)solver->domain->local.size() == solver->domain->dec.nranks)
            solver->domain->copy_exchange(solver->domain->lat[solver->domain->local[0]]->cur);
    }
    double dot(const std::vector<double> &a, const std::vector<double> &b) {
        double s = 0;
        for (size_t i = 0; i < a.size(); i++) s += a[i] * b[i];
        if (solver->domain->reduce_hook) solver->domain->reduce_hook(&s, 1, TCLB_OP_SUM);
        return s;
    }
  public:
    int init() override {
        GenericAction::init();
        if (!node->has_attr("Directions")) { logf(L_ERROR, "no Directions parameter in %s", node->name.c_str()); return -1; }
        const int directions = std::max(1, atoi(node->attr("Directions").c_str()));
        const int times = node->has_attr("Times") ? atoi(node->attr("Times").c_str()) : directions;
        const double eps = node->has_attr("Eps") ? atof(node->attr("Eps").c_str()) : 0.0;
        std::vector<Entry> hist;
        for (int it = 0; it < times; it++) {
            Entry e;
            get_state(e.x);
            if (it == 0) logf(L_OUTPUT, "Size of vector in Andersen: %zu", e.x.size());
            if (execute_internal()) return -1;
            unstack();
            get_state(e.r);
            for (size_t i = 0; i < e.r.size(); i++) e.r[i] -= e.x[i];
            const double res = dot(e.r, e.r);
            logf(L_OUTPUT, "Residual in Andersen: %lg", res);
            if (!std::isfinite(res) || res < eps || res == 0.0) break;
            e.p = 1.0;
            for (const Entry &o : hist) {                 // the stored residuals are orthonormal
                const double a = dot(e.r, o.r);
                for (size_t i = 0; i < e.r.size(); i++) { e.r[i] -= a * o.r[i]; e.x[i] -= a * o.x[i]; }
                e.p -= a * o.p;
            }
            const double nrm = std::sqrt(dot(e.r, e.r));
            if (!(nrm > 0)) break;
            for (size_t i = 0; i < e.r.size(); i++) { e.r[i] /= nrm; e.x[i] /= nrm; }
            e.p /= nrm;
            hist.insert(hist.begin(), std::move(e));
            if ((int)hist.size() > directions) hist.pop_back();
            double psum = 0;
            for (const Entry &o : hist) psum += o.p * o.p;
            std::vector<double> nx(hist[0].x.size(), 0.0);
            for (const Entry &o : hist) {
                const double c = o.p / psum;
                for (size_t i = 0; i < nx.size(); i++) nx[i] += c * o.x[i];
            }
            set_state(nx);
            if (execute_internal()) return -1;
            unstack();
        }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("Andersen", AcAndersen)

}  // namespace
}  // namespace tclb
