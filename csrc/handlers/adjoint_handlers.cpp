// Optimisation / adjoint elements of the case file.
//
// Reference: src/Handlers/acUSAdjoint.cpp, acSAdjoint.cpp (<Adjoint type=>), acOptSolve.cpp (<OptSolve>), acObjective.cpp
// (<Objective>), acThresholdNow.cpp / acThreshold.cpp, acFDTest.cpp.  The adjoint itself is csrc/core/adjoint.h (dual
// numbers instead of Tapenade), so these handlers drive a tape + reverse sweep instead of the reference's ITER_ADJOINT
// iteration type; the design variables are the model's parameter fields (AddDensity(parameter=TRUE)) restricted to
// nodes flagged DesignSpace when the model has that node type.
#include "handler.h"
#include "../core/adjoint.h"
#include "../core/common.h"
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
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
    int init() override {
        Action::init();
        const tclb_model_t *d = solver->model->d;
        int nset = 0;
        for (int g = 0; g < d->sum_globals; g++) {
            std::string nm = d->globals[g].name;
            if (nm == "Objective") continue;
            const double w = node->has_attr(nm) ? atof(node->attr(nm).c_str()) : 0.0;
            if (solver->model->find_zonesetting(nm + "InObj") < 0) continue;
            solver->set_param(nm + "InObj", node->has_attr(nm) ? node->attr(nm) : "0", "", true, "");
            if (w != 0.0) nset++;
        }
        if (nset == 0) logf(L_WARNING, "<Objective> names no Global of this model: the objective is zero");
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
        std::vector<int> pf = solver->parameter_fields();
        if (pf.empty()) { logf(L_ERROR, "No parameters defined!"); return -1; }
        std::vector<char> mask = design_mask(solver);
        size_t npar = 0;
        for (int f : pf) {
            std::vector<double> v;
            get_field_d(solver, f, v);
            for (size_t i = 0; i < v.size(); i++) if (mask[i]) { v[i] = v[i] > level ? 1.0 : 0.0; npar++; }
            set_field_d(solver, f, v);
        }
        logf(L_OUTPUT, "Parameters: %zu", npar);
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
        std::vector<int> pf = solver->parameter_fields();
        if (pf.empty()) { logf(L_ERROR, "No parameters defined!"); return -1; }
        std::vector<char> mask = design_mask(solver);
        std::vector<std::vector<double>> orig(pf.size());
        for (size_t k = 0; k < pf.size(); k++) get_field_d(solver, pf[k], orig[k]);
        for (int i = 0; i <= levels; i++) {
            const double level = (double)i / levels;
            for (size_t k = 0; k < pf.size(); k++) {
                std::vector<double> v = orig[k];
                for (size_t j = 0; j < v.size(); j++) if (mask[j]) v[j] = v[j] > level ? 1.0 : 0.0;
                set_field_d(solver, pf[k], v);
            }
            logf(L_OUTPUT, "Threshold level %lg", level);
            if (execute_internal()) return -1;
            unstack();
        }
        for (size_t k = 0; k < pf.size(); k++) set_field_d(solver, pf[k], orig[k]);
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

// <FDTest Samples="K" Epsilon="e"> children </FDTest>: gradient of the objective accumulated by the children with respect to
// K design variables, from the adjoint and from central differences (children re-run from the same state)
class AcFDTest : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        const int samples = node->has_attr("Samples") ? atoi(node->attr("Samples").c_str()) : 5;
        const double eps = node->has_attr("Epsilon") ? atof(node->attr("Epsilon").c_str()) : (solver->model->real_size() == 8 ? 1e-6 : 1e-3);
        std::vector<int> pf = solver->parameter_fields();
        if (pf.empty()) { logf(L_ERROR, "No parameters defined!"); return -1; }
        Lattice *l = the_lattice(solver);
        Adjoint *A = solver->need_adjoint();
        const std::string state = solver->outpath + "_fdtest_state.raw";
        const int iter0 = solver->iter, diter0 = solver->domain->iter;
        l->save_raw(state);
        auto restore = [&]() {
            l->load_raw(state);
            solver->iter = iter0; solver->domain->iter = diter0; l->iter = diter0;
        };
        auto run_children = [&](double &obj) -> int {
            A->drop_tape(); A->zero(); A->objective = 0;
            solver->recording = true;
            int rc = execute_internal();
            solver->recording = false;
            unstack();
            obj = A->objective;
            return rc;
        };
        double J0 = 0;
        if (run_children(J0)) return -1;
        A->reverse();
        std::vector<char> mask = design_mask(solver);
        std::vector<size_t> cand;
        for (size_t i = 0; i < mask.size(); i++) if (mask[i]) cand.push_back(i);
        if (cand.empty()) { logf(L_ERROR, "FDTest: no node in the design space"); return -1; }
        std::mt19937 rng(12345);
        const int f = pf[0];
        std::vector<double> grad;
        get_adj_d(solver, f, grad);
        double worst = 0;
        logf(L_OUTPUT, "FDTest: objective %.12lg, %d samples, epsilon %lg", J0, samples, eps);
        for (int k = 0; k < samples; k++) {
            const size_t i = cand[rng() % cand.size()];
            double Jp = 0, Jm = 0;
            for (int sgn = +1; sgn >= -1; sgn -= 2) {
                restore();
                std::vector<double> p;
                get_field_d(solver, f, p);
                p[i] += sgn * eps;
                set_field_d(solver, f, p);
                double J = 0;
                if (run_children(J)) return -1;
                A->drop_tape();
                (sgn > 0 ? Jp : Jm) = J;
            }
            const double fd = (Jp - Jm) / (2 * eps);
            const double err = std::fabs(fd - grad[i]) / std::max(1e-300, std::max(std::fabs(fd), std::fabs(grad[i])));
            worst = std::max(worst, err);
            logf(L_OUTPUT, "FDTest: parameter %zu  adjoint %.10lg  finite difference %.10lg  relative difference %.3lg", i, grad[i], fd, err);
        }
        restore();
        remove(state.c_str());
        logf(L_OUTPUT, "FDTest: largest relative difference %.3lg", worst);
        return 0;
    }
};
TCLB_REGISTER_HANDLER("FDTest", AcFDTest)

// <Optimize MaxEvaluations="N" [method="GD"] [Step="s"]> children </Optimize>: maximise the objective accumulated by the
// children over the design variables, bounds [0,1].  The reference hands this to NLopt (MMA, L-BFGS, COBYLA ...,
// src/Handlers/acOptimize.cpp.Rt); NLopt is not available here, so the built-in method is projected gradient ascent
// with backtracking on the step: every evaluation = children from the same start state on the tape + reverse sweep.
class AcOptimize : public GenericAction {
  public:
    int init() override {
        GenericAction::init();
        const std::string method = node->has_attr("method") ? node->attr("method") : "GD";
        if (method != "GD") { logf(L_ERROR, "Unknown Method in Optimize: %s (this build has GD; NLopt methods are not available)", method.c_str()); return -1; }
        if (node->has_attr("Material")) logf(L_WARNING, "Optimize: the Material constraint is not implemented by the GD method");
        const int max_eval = node->has_attr("MaxEvaluations") ? atoi(node->attr("MaxEvaluations").c_str()) : 20;
        double step = node->has_attr("Step") ? atof(node->attr("Step").c_str()) : 1.0;
        std::vector<int> pf = solver->parameter_fields();
        if (pf.empty()) { logf(L_ERROR, "Error: No parameters defined!"); return -1; }
        Lattice *l = the_lattice(solver);
        Adjoint *A = solver->need_adjoint();
        std::vector<char> mask = design_mask(solver);
        const std::string state = solver->outpath + "_optimize_state.raw";
        const int iter0 = solver->iter, diter0 = solver->domain->iter;
        l->save_raw(state);
        size_t npar = 0;
        for (char c : mask) npar += c ? pf.size() : 0;
        logf(L_NOTICE, "Parameters in optimization: %zu", npar);
        std::vector<std::vector<double>> x(pf.size()), g(pf.size()), trial(pf.size());
        for (size_t k = 0; k < pf.size(); k++) get_field_d(solver, pf[k], x[k]);
        auto evaluate = [&](const std::vector<std::vector<double>> &p, double &obj, bool want_grad) -> int {
            l->load_raw(state);
            solver->iter = iter0; solver->domain->iter = diter0; l->iter = diter0;
            for (size_t k = 0; k < pf.size(); k++) set_field_d(solver, pf[k], p[k]);
            A->drop_tape(); A->zero(); A->objective = 0;
            solver->recording = true;
            int rc = execute_internal();
            solver->recording = false;
            unstack();
            if (rc) return rc;
            obj = A->objective;
            if (want_grad) { A->reverse(); for (size_t k = 0; k < pf.size(); k++) get_adj_d(solver, pf[k], g[k]); }
            else A->drop_tape();
            return 0;
        };
        double best = 0;
        int evals = 0;
        if (evaluate(x, best, true)) return -1;
        evals++;
        logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg", evals, best);
        while (evals < max_eval) {
            double gmax = 0;
            for (size_t k = 0; k < pf.size(); k++) for (size_t i = 0; i < g[k].size(); i++) if (mask[i]) gmax = std::max(gmax, std::fabs(g[k][i]));
            if (gmax == 0) { logf(L_OUTPUT, "Optimize: zero gradient, stopping"); break; }
            bool improved = false;
            while (evals < max_eval && !improved) {
                // step scaled so that the largest change of a design variable is `step`
                for (size_t k = 0; k < pf.size(); k++) {
                    trial[k] = x[k];
                    for (size_t i = 0; i < trial[k].size(); i++)
                        if (mask[i]) trial[k][i] = std::min(1.0, std::max(0.0, x[k][i] + step * g[k][i] / gmax));
                }
                double obj = 0;
                std::vector<std::vector<double>> gkeep = g;
                if (evaluate(trial, obj, true)) return -1;
                evals++;
                if (obj > best) { best = obj; x = trial; improved = true; step *= 1.5; }
                else { g = gkeep; step *= 0.5; }
                logf(L_OUTPUT, "Optimize: evaluation %d objective %.10lg %s (step %lg)", evals, obj, improved ? "accepted" : "rejected", step);
                if (step < 1e-8) break;
            }
            if (!improved) break;
        }
        // leave the solver at the start state with the best design
        l->load_raw(state);
        solver->iter = iter0; solver->domain->iter = diter0; l->iter = diter0;
        for (size_t k = 0; k < pf.size(); k++) set_field_d(solver, pf[k], x[k]);
        remove(state.c_str());
        logf(L_OUTPUT, "Optimize: best objective %.10lg after %d evaluations", best, evals);
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
