// Design elements of the case file: handlers that expose a vector of optimisation parameters.
//
// Reference: src/Handlers/Design.cpp (base), vHandler.h:13-50 + vHandler.cpp:30-66 (the protocol: NumberOfParameters,
// Parameters(PAR_GET | PAR_SET | PAR_GRAD | PAR_UPPER | PAR_LOWER | PAR_X..T) and the concatenation over every design
// handler on the solver's stack), OptimalControl.cpp, OptimalControlSecond.cpp, Fourier.cpp, BSpline.cpp,
// RepeatControl.cpp, InternalTopology.cpp, conFieldParameter.cpp, conExtrude.cpp, conControlParameter.cpp and the
// Solver side src/Solver.cpp.Rt:435-731 (getPars/getPar/getDPar/setPar).  Written from that behaviour, not from the
// code: one file, value semantics, gradients from this framework's adjoint (csrc/core/adjoint.h).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include "../core/adjoint.h"
#include "../core/common.h"
#include "design.h"
#include "handler.h"

namespace tclb {

// ---------------------------------------------------------------- protocol over the solver's stack
int design_count(Solver *s) {
    int n = 0;
    for (auto &h : s->hands) if (h->type() & HANDLER_DESIGN) n++;
    return n;
}
int design_number_of_parameters(Solver *s) {
    int n = 0;
    for (auto &h : s->hands) if (h->type() & HANDLER_DESIGN) n += h->number_of_parameters();
    return n;
}
int design_parameters(Solver *s, int what, double *tab) {
    int off = 0;
    for (auto &h : s->hands) if (h->type() & HANDLER_DESIGN) {
        const int n = h->number_of_parameters();
        if (int rc = h->parameters(what, tab + off)) return rc;
        off += n;
    }
    return 0;
}

namespace {

double attr_value(Handler *h, const char *name, double dflt, bool *given = nullptr) {
    if (given) *given = h->node->has_attr(name);
    if (!h->node->has_attr(name)) {
        logf(L_NOTICE, "%s bound not set in %s - setting to %lg", name, h->node->name.c_str(), dflt);
        return dflt;
    }
    return h->solver->units.alt(h->node->attr(name));
}

int zone_index_of(Solver *s, const std::string &name) {
    auto it = s->setting_zones.find(name);
    return it == s->setting_zones.end() ? -1 : it->second;
}

Lattice *only_lattice(Solver *s, const char *who) {
    if (s->domain->local.size() != 1 || s->domain->dec.nranks != 1) fail("%s: design elements handle one subdomain in this version", who);
    return s->domain->lat[s->domain->local[0]];
}

// ---- <OptimalControl what="Setting-Zone" lower= upper=/>: the time series of one zonal setting in one zone -----------
class OptimalControl : public Design {
  protected:
    int par_ = -1, zone_ = -1, len_ = 0;
    double lower_ = -1, upper_ = 1;
    std::string par_name_, zone_name_;
    FILE *f_ = nullptr;
    void trace(const char *tag, const double *tab, int n) {
        if (!f_) return;
        fprintf(f_, "%s", tag);
        for (int i = 0; i < n; i++) fprintf(f_, ",%lg", tab[i]);
        fprintf(f_, "\n");
        fflush(f_);
    }
    int select() {
        if (!node->has_attr("what")) { logf(L_ERROR, "Parameter \"what\" needed in %s", node->name.c_str()); return -1; }
        std::string what = node->attr("what");
        const size_t i = what.find('-');
        if (i == std::string::npos) { logf(L_ERROR, "Can only optimal control a parameters in a specific zone"); return -1; }
        par_name_ = what.substr(0, i);
        zone_name_ = what.substr(i + 1);
        zone_ = zone_index_of(solver, zone_name_);
        if (zone_ < 0) { logf(L_ERROR, "Unknown zone %s (found while setting parameter %s)", zone_name_.c_str(), par_name_.c_str()); return -1; }
        par_ = solver->model->find_zonesetting(par_name_);
        if (par_ < 0) { logf(L_ERROR, "Unknown param %s in %s", par_name_.c_str(), node->name.c_str()); return -1; }
        Lattice *l = only_lattice(solver, node->name.c_str());
        len_ = l->zs.is_series(par_, zone_) ? l->zs.len() : 1;
        logf(L_OUTPUT, "Selected %s (%d) in zone \"%s\" (%d) for optimal control, length of the control: %d", par_name_.c_str(), par_, zone_name_.c_str(), zone_, len_);
        solver->iter_type |= ITER_GLOBS;
        lower_ = attr_value(this, "lower", -1);
        upper_ = attr_value(this, "upper", 1);
        f_ = fopen((solver->outpath + "_OC_" + par_name_ + "_" + zone_name_ + ".dat").c_str(), "w");
        solver->need_adjoint()->want_zone_gradient(par_, zone_);
        return 0;
    }
    void get_series(std::vector<double> &v) {
        Lattice *l = only_lattice(solver, "OptimalControl");
        v.resize(len_);
        for (int t = 0; t < len_; t++) v[t] = l->zs.get(par_, zone_, t);
    }
    void set_series(const std::vector<double> &v) {
        for (int r : solver->domain->local) {
            if (len_ == 1) solver->domain->lat[r]->zs.set(par_, zone_, v[0]);
            else solver->domain->lat[r]->zs.set_series(par_, zone_, v);
        }
    }
    void get_grad(std::vector<double> &g) {
        g.assign(len_, 0.0);
        solver->need_adjoint()->zone_gradient(par_, zone_, g.data(), len_);
    }
  public:
    ~OptimalControl() override { if (f_) fclose(f_); }
    int init() override { if (select()) return -1; return Design::init(); }
    int number_of_parameters() override { return len_; }
    int parameters(int what, double *tab) override {
        std::vector<double> v;
        switch (what) {
        case PAR_GET: get_series(v); std::copy(v.begin(), v.end(), tab); trace("GET", tab, len_); return 0;
        case PAR_SET: v.assign(tab, tab + len_); set_series(v); trace("SET", tab, len_); return 0;
        case PAR_GRAD: get_grad(v); std::copy(v.begin(), v.end(), tab); trace("GRAD", tab, len_); return 0;
        case PAR_UPPER: std::fill(tab, tab + len_, upper_); return 0;
        case PAR_LOWER: std::fill(tab, tab + len_, lower_); return 0;
        case PAR_X: case PAR_Y: case PAR_Z: std::fill(tab, tab + len_, 0.0); return 0;
        case PAR_T: for (int i = 0; i < len_; i++) tab[i] = i; return 0;
        }
        logf(L_ERROR, "Unknown type %d in call to Parameters in %s", what, node->name.c_str());
        return -1;
    }
};
TCLB_REGISTER_HANDLER("OptimalControl", OptimalControl)

// ---- <OptimalControlSecond>: every second time level is a parameter, the levels in between are their means ----------
class OptimalControlSecond : public OptimalControl {
    int pars_ = 0;
  public:
    int init() override {
        if (select()) return -1;
        pars_ = (len_ + 1) / 2;
        return Design::init();
    }
    int number_of_parameters() override { return pars_; }
    int parameters(int what, double *tab) override {
        std::vector<double> v;
        switch (what) {
        case PAR_GET:
            get_series(v);
            for (int i = 0; i < len_; i += 2) tab[i / 2] = v[i];
            trace("GET", tab, pars_);
            return 0;
        case PAR_SET:
            trace("SET", tab, pars_);
            v.assign(len_, 0.0);
            for (int i = 0; i < pars_; i++) {
                if (2 * i < len_) v[2 * i] = tab[i];
                if (2 * i + 1 < len_) v[2 * i + 1] = i + 1 < pars_ ? 0.5 * (tab[i] + tab[i + 1]) : tab[i];
            }
            set_series(v);
            return 0;
        case PAR_GRAD:
            get_grad(v);
            std::fill(tab, tab + pars_, 0.0);
            for (int i = 0; i < pars_; i++) {
                if (2 * i < len_) tab[i] += v[2 * i];
                if (2 * i + 1 < len_) {
                    if (i + 1 < pars_) { tab[i] += 0.5 * v[2 * i + 1]; tab[i + 1] += 0.5 * v[2 * i + 1]; }
                    else tab[i] += v[2 * i + 1];
                }
            }
            trace("GRAD", tab, pars_);
            return 0;
        case PAR_UPPER: std::fill(tab, tab + pars_, upper_); return 0;
        case PAR_LOWER: std::fill(tab, tab + pars_, lower_); return 0;
        case PAR_X: case PAR_Y: case PAR_Z: std::fill(tab, tab + pars_, 0.0); return 0;
        case PAR_T: for (int i = 0; i < pars_; i++) tab[i] = 2 * i; return 0;
        }
        return -1;
    }
};
TCLB_REGISTER_HANDLER("OptimalControlSecond", OptimalControlSecond)

// ---- wrappers around ONE child design element ---------------------------------------------------------------------------
class WrapDesign : public Design {
  protected:
    HandlerPtr child_;
    int inner_ = 0;
    std::vector<double> tab2_;
    int make_child() {
        auto kids = node->children();
        if (kids.empty()) { logf(L_ERROR, "%s needs exacly one child! - none found", node->name.c_str()); return -1; }
        if (kids.size() > 1) { logf(L_ERROR, "%s needs exacly one child! - more found", node->name.c_str()); return -1; }
        child_ = HandlerFactory::produce(kids[0], solver);
        if (!child_) return -1;
        if (!(child_->type() & HANDLER_DESIGN)) { logf(L_ERROR, "%s needs child of design type!", node->name.c_str()); return -1; }
        inner_ = child_->number_of_parameters();
        tab2_.assign(std::max(1, inner_), 0.0);
        return 0;
    }
  public:
    int finish() override { if (child_) child_->finish(); child_.reset(); return 0; }
};

// <Fourier modes="m" lower= upper=> child </Fourier>: truncated Fourier series of the child's (time) parameters
class Fourier : public WrapDesign {
    int pars_ = 0;
    double lower_ = -1, upper_ = 1;
    double basis(int i, int j) const {
        const int i0 = (i + 1) >> 1;
        const double a = i0 * 2.0 * M_PI * j / inner_;
        return (i & 1) ? std::sin(a) : std::cos(a);
    }
  public:
    int init() override {
        if (make_child()) return -1;
        pars_ = node->has_attr("modes") ? atoi(node->attr("modes").c_str()) : 10;
        if (pars_ % 2 != 1) { pars_++; logf(L_NOTICE, "number of modes in %s not odd - setting to %d", node->name.c_str(), pars_); }
        lower_ = attr_value(this, "lower", -1);
        upper_ = attr_value(this, "upper", 1);
        logf(L_OUTPUT, "Lenght of time-resolved control: %d, of fourier control: %d", inner_, pars_);
        return Design::init();
    }
    int number_of_parameters() override { return pars_; }
    int parameters(int what, double *tab) override {
        const double s = 0.5 * (upper_ - lower_);
        switch (what) {
        case PAR_GET:
        case PAR_GRAD:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            for (int i = 0; i < pars_; i++) {
                double v = 0;
                for (int j = 0; j < inner_; j++) v += basis(i, j) * tab2_[j];
                if (what == PAR_GET) v = (i == 0 ? 1.0 : 2.0) * v / inner_;
                tab[i] = v;
            }
            return 0;
        case PAR_SET:
            for (int j = 0; j < inner_; j++) {
                double v = 0;
                for (int i = 0; i < pars_; i++) v += basis(i, j) * tab[i];
                tab2_[j] = v;
            }
            return child_->parameters(what, tab2_.data());
        case PAR_UPPER: for (int i = 1; i < pars_; i++) tab[i] = s / ((i + 1) >> 1); tab[0] = upper_; return 0;
        case PAR_LOWER: for (int i = 1; i < pars_; i++) tab[i] = -s / ((i + 1) >> 1); tab[0] = lower_; return 0;
        case PAR_X: case PAR_Y: case PAR_Z: case PAR_T: std::fill(tab, tab + pars_, 0.0); return 0;
        }
        return -1;
    }
};
TCLB_REGISTER_HANDLER("Fourier", Fourier)

// B-spline basis of the reference's src/spline.h (uniform knots, optional periodic wrap), value of basis function i of n
double bspline_basis(double x, int n, int i, int order, bool periodic) {
    // uniform B-spline of degree `order` on [0,1] with n control points
    if (periodic) {
        const double t = x * n - i;
        auto cardinal = [&](double u) {          // cardinal B-spline of degree `order` supported on [0, order+1]
            std::vector<double> b(order + 2, 0.0);
            for (int k = 0; k <= order; k++) b[k] = (u >= k && u < k + 1) ? 1.0 : 0.0;
            for (int d = 1; d <= order; d++)
                for (int k = 0; k + d <= order; k++)
                    b[k] = ((u - k) / d) * b[k] + ((k + d + 1 - u) / d) * b[k + 1];
            return b[0];
        };
        double v = 0;
        for (int w = -1; w <= 1; w++) v += cardinal(t + order + w * n);      // support [(i-k)/n, (i+1)/n], wrapped: the reference's placement
        return v;
    }
    // clamped (open uniform) knot vector
    const int m = n + order + 1;
    std::vector<double> knots(m);
    for (int k = 0; k < m; k++) {
        if (k <= order) knots[k] = 0;
        else if (k >= n) knots[k] = 1;
        else knots[k] = (double)(k - order) / (n - order);
    }
    std::vector<double> b(m - 1, 0.0);
    for (int k = 0; k < m - 1; k++) b[k] = (x >= knots[k] && x < knots[k + 1]) ? 1.0 : 0.0;
    if (x >= 1.0) { for (int k = m - 2; k >= 0; k--) if (knots[k] < knots[k + 1]) { b[k] = 1.0; break; } }
    for (int d = 1; d <= order; d++)
        for (int k = 0; k + d < m - 1 + 0; k++) {
            const double d1 = knots[k + d] - knots[k], d2 = knots[k + d + 1] - knots[k + 1];
            const double a = d1 > 0 ? (x - knots[k]) / d1 * b[k] : 0.0;
            const double c = d2 > 0 ? (knots[k + d + 1] - x) / d2 * b[k + 1] : 0.0;
            b[k] = a + c;
        }
    return b[i];
}
void gauss_solve(std::vector<double> A, std::vector<double> y, double *x, int n) {
    for (int c = 0; c < n; c++) {
        int p = c;
        for (int r = c + 1; r < n; r++) if (std::fabs(A[r * n + c]) > std::fabs(A[p * n + c])) p = r;
        if (std::fabs(A[p * n + c]) < 1e-300) continue;
        if (p != c) { for (int k = 0; k < n; k++) std::swap(A[c * n + k], A[p * n + k]); std::swap(y[c], y[p]); }
        for (int r = c + 1; r < n; r++) {
            const double f = A[r * n + c] / A[c * n + c];
            for (int k = c; k < n; k++) A[r * n + k] -= f * A[c * n + k];
            y[r] -= f * y[c];
        }
    }
    for (int r = n - 1; r >= 0; r--) {
        double s = y[r];
        for (int k = r + 1; k < n; k++) s -= A[r * n + k] * x[k];
        x[r] = std::fabs(A[r * n + r]) > 1e-300 ? s / A[r * n + r] : 0.0;
    }
}

// <BSpline nodes="n" order="k" periodic="true|false" lower= upper=> child </BSpline>
class BSpline : public WrapDesign {
    int pars_ = 0, order_ = 3;
    bool per_ = false;
    double lower_ = -1, upper_ = 1;
    double pos(int j) const { return per_ ? (double)j / inner_ : j / (inner_ - 1.0); }
  public:
    int init() override {
        if (make_child()) return -1;
        pars_ = node->has_attr("nodes") ? atoi(node->attr("nodes").c_str()) : 10;
        order_ = node->has_attr("order") ? atoi(node->attr("order").c_str()) : 3;
        per_ = node->attr("periodic") == "true" || node->attr("periodic") == "yes" || node->attr("periodic") == "1";
        if (pars_ <= order_) { logf(L_ERROR, "%s needs more nodes (%d) than its order (%d)", node->name.c_str(), pars_, order_); return -1; }
        lower_ = attr_value(this, "lower", -1);
        upper_ = attr_value(this, "upper", 1);
        logf(L_OUTPUT, "Lenght of time-resolved control: %d, of b-spline control: %d", inner_, pars_);
        return Design::init();
    }
    int number_of_parameters() override { return pars_; }
    int parameters(int what, double *tab) override {
        switch (what) {
        case PAR_GET: {      // least-squares fit of the child's values
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            std::vector<double> mat((size_t)pars_ * pars_, 0.0), y(pars_, 0.0), w(pars_);
            for (int j = 0; j < inner_; j++) {
                const double x = pos(j);
                for (int i = 0; i < pars_; i++) w[i] = bspline_basis(x, pars_, i, order_, per_);
                for (int i = 0; i < pars_; i++) {
                    y[i] += w[i] * tab2_[j];
                    for (int k = 0; k < pars_; k++) mat[(size_t)i * pars_ + k] += w[i] * w[k];
                }
            }
            gauss_solve(mat, y, tab, pars_);
            return 0;
        }
        case PAR_SET:
            for (int j = 0; j < inner_; j++) {
                double v = 0;
                for (int i = 0; i < pars_; i++) v += bspline_basis(pos(j), pars_, i, order_, per_) * tab[i];
                tab2_[j] = v;
            }
            return child_->parameters(what, tab2_.data());
        case PAR_GRAD:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            for (int i = 0; i < pars_; i++) {
                double v = 0;
                for (int j = 0; j < inner_; j++) v += bspline_basis(pos(j), pars_, i, order_, per_) * tab2_[j];
                tab[i] = v;
            }
            return 0;
        case PAR_UPPER: std::fill(tab, tab + pars_, upper_); return 0;
        case PAR_LOWER: std::fill(tab, tab + pars_, lower_); return 0;
        case PAR_X: case PAR_Y: case PAR_Z: case PAR_T: std::fill(tab, tab + pars_, 0.0); return 0;
        }
        return -1;
    }
};
TCLB_REGISTER_HANDLER("BSpline", BSpline)

// <RepeatControl length="n" [flip="level"] lower= upper=> child </RepeatControl>: a control of period n repeated over the child
class RepeatControl : public WrapDesign {
    int pars_ = 0;
    bool flip_ = false;
    double flip_level_ = 0, lower_ = -1, upper_ = 1;
    double flip(double v, double l, int j) const {
        if (!flip_) return v;
        return ((j / pars_) % 2 == 0) ? v : l - v;
    }
  public:
    int init() override {
        if (make_child()) return -1;
        pars_ = node->has_attr("length") ? (int)solver->units.alt(node->attr("length")) : 10;
        if (pars_ < 1) pars_ = 1;
        if (pars_ > inner_) pars_ = std::max(1, inner_);
        lower_ = attr_value(this, "lower", -1);
        upper_ = attr_value(this, "upper", 1);
        flip_ = node->has_attr("flip");
        if (flip_) flip_level_ = solver->units.alt(node->attr("flip"));
        logf(L_OUTPUT, "Lenght of time-resolved control: %d, of repeated control: %d", inner_, pars_);
        return Design::init();
    }
    int number_of_parameters() override { return pars_; }
    int parameters(int what, double *tab) override {
        switch (what) {
        case PAR_GET:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            std::fill(tab, tab + pars_, 0.0);
            for (int j = 0; j < inner_; j++) tab[j % pars_] += flip(tab2_[j], flip_level_, j);
            for (int i = 0; i < pars_; i++) tab[i] /= std::floor((inner_ - i - 1.0) / pars_ + 1);
            return 0;
        case PAR_SET:
            for (int j = 0; j < inner_; j++) tab2_[j] = flip(tab[j % pars_], flip_level_, j);
            return child_->parameters(what, tab2_.data());
        case PAR_GRAD:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            std::fill(tab, tab + pars_, 0.0);
            for (int j = 0; j < inner_; j++) tab[j % pars_] += flip(tab2_[j], 0.0, j);
            return 0;
        case PAR_UPPER: std::fill(tab, tab + pars_, upper_); return 0;
        case PAR_LOWER: std::fill(tab, tab + pars_, lower_); return 0;
        case PAR_X: case PAR_Y: case PAR_Z: case PAR_T: std::fill(tab, tab + pars_, 0.0); return 0;
        }
        return -1;
    }
};
TCLB_REGISTER_HANDLER("RepeatControl", RepeatControl)

// ---- parameter fields on a set of nodes ---------------------------------------------------------------------------------
// base of <InternalTopology/> (every parameter field on the DesignSpace nodes: src/Solver.cpp.Rt:435-731) and
// <FieldParameter field= [where=NodeType] [zone=Zone]/> (one parameter field on the nodes of a type and/or zone)
class FieldDesign : public Design {
  protected:
    std::vector<int> fields_;
    uint32_t mask_ = 0, value_ = 0;
    std::vector<size_t> nodes_;          // linear indices of the design nodes
    Lattice *lat_ = nullptr;
    void find_nodes() {
        lat_ = only_lattice(solver, node ? node->name.c_str() : "design");
        const size_t n = (size_t)lat_->nx * lat_->ny * lat_->nz;
        std::vector<uint32_t> fl(n);
        Box b; b.x0 = b.y0 = b.z0 = 0; b.nx = lat_->nx; b.ny = lat_->ny; b.nz = lat_->nz;
        lat_->get_flags(b, fl.data());
        nodes_.clear();
        for (size_t i = 0; i < n; i++) if ((fl[i] & mask_) == value_) nodes_.push_back(i);
    }
    void get_field(int f, std::vector<double> &v, bool adjoint) {
        const size_t n = (size_t)lat_->nx * lat_->ny * lat_->nz;
        v.resize(n);
        if (solver->model->real_size() == 8) { if (adjoint) solver->need_adjoint()->get_field(f, v.data()); else lat_->copy_field_out(f, v.data()); }
        else {
            std::vector<float> t(n);
            if (adjoint) solver->need_adjoint()->get_field(f, t.data()); else lat_->copy_field_out(f, t.data());
            for (size_t i = 0; i < n; i++) v[i] = t[i];
        }
    }
    void set_field(int f, const std::vector<double> &v) {
        if (solver->model->real_size() == 8) lat_->copy_field_in(f, v.data());
        else { std::vector<float> t(v.begin(), v.end()); lat_->copy_field_in(f, t.data()); }
        solver->domain->copy_exchange(lat_->cur);
    }
  public:
    int number_of_parameters() override { return (int)(nodes_.size() * fields_.size()); }
    int parameters(int what, double *tab) override {
        const size_t np = nodes_.size();
        std::vector<double> v;
        switch (what) {
        case PAR_GET:
        case PAR_GRAD: {
            double sum = 0;
            for (size_t k = 0; k < fields_.size(); k++) {
                get_field(fields_[k], v, what == PAR_GRAD);
                for (size_t j = 0; j < np; j++) { tab[k * np + j] = v[nodes_[j]]; sum += v[nodes_[j]] * v[nodes_[j]]; }
            }
            if (what == PAR_GRAD) logf(L_OUTPUT, "L2 norm of gradient: %lg", std::sqrt(sum));
            return 0;
        }
        case PAR_SET: {
            double sum = 0;
            for (size_t k = 0; k < fields_.size(); k++) {
                get_field(fields_[k], v, false);
                for (size_t j = 0; j < np; j++) { const double d = v[nodes_[j]] - tab[k * np + j]; sum += d * d; v[nodes_[j]] = tab[k * np + j]; }
                set_field(fields_[k], v);
            }
            logf(L_OUTPUT, "L2 norm of parameter change: %lg", std::sqrt(sum));
            return 0;
        }
        case PAR_UPPER: std::fill(tab, tab + number_of_parameters(), 1.0); return 0;
        case PAR_LOWER: std::fill(tab, tab + number_of_parameters(), 0.0); return 0;
        case PAR_X: case PAR_Y: case PAR_Z: case PAR_T:
            for (size_t k = 0; k < fields_.size(); k++)
                for (size_t j = 0; j < np; j++) {
                    const size_t i = nodes_[j];
                    const int x = (int)(i % lat_->nx), y = (int)((i / lat_->nx) % lat_->ny), z = (int)(i / ((size_t)lat_->nx * lat_->ny));
                    tab[k * np + j] = what == PAR_X ? x : what == PAR_Y ? y : what == PAR_Z ? z : 0;
                }
            return 0;
        }
        logf(L_ERROR, "Unknown type %d in call to Parameters in %s", what, node ? node->name.c_str() : "design");
        return -1;
    }
};

class InternalTopology : public FieldDesign {
  public:
    int init() override {
        fields_ = solver->parameter_fields();
        if (fields_.empty()) { logf(L_ERROR, "%s: the model has no parameter fields", node ? node->name.c_str() : "InternalTopology"); return -1; }
        const tclb_nodetype_t *t = solver->model->find_nodetype("DesignSpace");
        if (t) { mask_ = t->mask; value_ = t->value; } else { mask_ = 0; value_ = 0; }
        find_nodes();
        logf(L_OUTPUT, "%s with %d parameters", node ? node->name.c_str() : "InternalTopology", number_of_parameters());
        return node ? Design::init() : 0;
    }
};
TCLB_REGISTER_HANDLER("InternalTopology", InternalTopology)

class FieldParameter : public FieldDesign {
  public:
    int init() override {
        if (!node->has_attr("field")) { logf(L_ERROR, "No \"field\" attribute in %s", node->name.c_str()); return -1; }
        const std::string field = node->attr("field");
        const int f = solver->model->find_field(field);
        if (f < 0) { logf(L_ERROR, "\"%s\" is not a field", field.c_str()); return -1; }
        if (!solver->model->d->fields[f].parameter) { logf(L_ERROR, "\"%s\" is not a valid parameter field", field.c_str()); return -1; }
        fields_.assign(1, f);
        mask_ = value_ = 0;
        if (node->has_attr("where")) {
            const tclb_nodetype_t *t = solver->model->find_nodetype(node->attr("where"));
            if (!t) { logf(L_ERROR, "\"%s\" is not a valid node type", node->attr("where").c_str()); return -1; }
            mask_ = t->mask; value_ = t->value;
        }
        if (node->has_attr("zone")) {
            const int z = zone_index_of(solver, node->attr("zone"));
            if (z < 0) { logf(L_ERROR, "Unknown zone %s in %s", node->attr("zone").c_str(), node->name.c_str()); return -1; }
            const tclb_model_t *d = solver->model->d;
            mask_ |= (uint32_t)d->zone_max << d->zone_shift;
            value_ |= (uint32_t)z << d->zone_shift;
        }
        find_nodes();
        logf(L_OUTPUT, "FieldParameter with %d parameters", number_of_parameters());
        return Design::init();
    }
};
TCLB_REGISTER_HANDLER("FieldParameter", FieldParameter)

// <ControlParameter field=/>: in the reference an action that only checks its attribute (conControlParameter.cpp)
class ControlParameter : public Action {
  public:
    int init() override {
        Action::init();
        if (!node->has_attr("field")) { logf(L_ERROR, "No \"field\" attribute in %s", node->name.c_str()); return -1; }
        return 0;
    }
};
TCLB_REGISTER_HANDLER("ControlParameter", ControlParameter)

// <Extrude direction="x|y|z|t" theta= margin=> child </Extrude>: every column of child parameters along `direction` is
// replaced by ONE parameter v, the position of a smooth front: child(j) = s((coord_j - v)/theta), s = logistic
class Extrude : public WrapDesign {
    int dir_ = -1, pars_ = 0;
    double theta_ = 1, margin_ = 1;
    std::vector<double> coords_[4], par_;
    std::vector<size_t> idx_;
    bool next(size_t k) const {
        if (!(k + 1 < (size_t)inner_)) return true;
        const size_t a = idx_[k], b = idx_[k + 1];
        for (int i = 0; i < 4; i++) if (i != dir_ && coords_[i][a] != coords_[i][b]) return true;
        return false;
    }
    double fun(double x, double v) const { const double e = std::exp((x - v) / theta_); return e / (e + 1); }
    double fund(double x, double v) const { const double e = std::exp((x - v) / theta_); return -e / (e + 1) / (e + 1) / theta_; }
  public:
    int init() override {
        if (make_child()) return -1;
        const std::string d = node->attr("direction");
        dir_ = d == "x" ? 0 : d == "y" ? 1 : d == "z" ? 2 : d == "t" ? 3 : -1;
        if (dir_ < 0) { logf(L_ERROR, "%s needs proper direction - \"%s\" given!", node->name.c_str(), d.c_str()); return -1; }
        if (node->has_attr("theta")) theta_ = solver->units.alt(node->attr("theta"));
        if (node->has_attr("margin")) margin_ = solver->units.alt(node->attr("margin"));
        for (int i = 0; i < 4; i++) { coords_[i].assign(inner_, 0.0); child_->parameters(PAR_X + i, coords_[i].data()); }
        idx_.resize(inner_);
        std::iota(idx_.begin(), idx_.end(), (size_t)0);
        const bool positive = theta_ > 0;
        std::sort(idx_.begin(), idx_.end(), [&](size_t a, size_t b) {
            for (int i = 0; i < 4; i++) if (i != dir_) {
                if (coords_[i][a] < coords_[i][b]) return true;
                if (coords_[i][a] > coords_[i][b]) return false;
            }
            return positive ? coords_[dir_][a] < coords_[dir_][b] : coords_[dir_][a] > coords_[dir_][b];
        });
        pars_ = 0;
        for (size_t i = 0; i < (size_t)inner_; i++) if (next(i)) pars_++;
        par_.assign(pars_, 0.0);
        logf(L_OUTPUT, "%s with %d parameters", node->name.c_str(), pars_);
        return Design::init();
    }
    int number_of_parameters() override { return pars_; }
    int parameters(int what, double *tab) override {
        size_t k = 0;
        switch (what) {
        case PAR_SET:
            std::copy(tab, tab + pars_, par_.begin());        // the map is non-linear: the gradient needs the point
            for (size_t i = 0; i < (size_t)inner_; i++) { const size_t j = idx_[i]; tab2_[j] = fun(coords_[dir_][j], tab[k]); if (next(i)) k++; }
            return child_->parameters(what, tab2_.data());
        case PAR_GRAD:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            std::fill(tab, tab + pars_, 0.0);
            for (size_t i = 0; i < (size_t)inner_; i++) { const size_t j = idx_[i]; tab[k] += fund(coords_[dir_][j], par_[k]) * tab2_[j]; if (next(i)) k++; }
            return 0;
        case PAR_GET:
            if (int rc = child_->parameters(what, tab2_.data())) return rc;
            // fall through
        case PAR_UPPER:
        case PAR_LOWER: {
            bool st = true;
            for (size_t i = 0; i < (size_t)inner_; i++) {
                const size_t j = idx_[i];
                if (what == PAR_UPPER && tab[k] < coords_[dir_][j]) st = true;
                if (what == PAR_LOWER && tab[k] > coords_[dir_][j]) st = true;
                if (what == PAR_GET && tab2_[j] < 0.5) st = true;
                if (st) { tab[k] = coords_[dir_][j]; st = false; }
                if (next(i)) { k++; st = true; }
            }
            const double off = std::fabs(margin_ * theta_);
            for (int i = 0; i < pars_; i++) tab[i] += what == PAR_UPPER ? off : -off;
            if (what == PAR_GET) std::copy(tab, tab + pars_, par_.begin());
            return 0;
        }
        case PAR_X: case PAR_Y: case PAR_Z: case PAR_T: std::fill(tab, tab + pars_, 0.0); return 0;
        }
        return -1;
    }
};
TCLB_REGISTER_HANDLER("Extrude", Extrude)

}  // namespace

// the design every optimisation element falls back to when the case file declares none: all parameter fields on the
// DesignSpace nodes (what the reference's optimisers see through Solver::getPars when <InternalTopology/> is given)
HandlerPtr make_default_design(Solver *s) {
    auto *d = new InternalTopology();
    d->node = nullptr;
    d->solver = s;
    HandlerPtr h(d);
    if (d->init() != 0) return nullptr;
    return h;
}

}  // namespace tclb
