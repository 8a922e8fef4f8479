#include "units.h"
#include "../core/common.h"
#include <cctype>
#include <cmath>
#include <cstdlib>
#include <sstream>

namespace tclb {

const char *const BASE_UNIT_NAMES[N_BASE_UNITS] = {"m", "s", "kg", "K", "x", "y", "z", "A", "t"};

static std::string num_str(double v) {
    std::ostringstream s;
    s << v;
    return s.str();
}
UnitVal UnitVal::base(int k) {
    if (k < 0 || k >= N_BASE_UNITS) fail("Wrong number of unit at initializer");
    UnitVal u(1.0);
    u.uni[k] = 1;
    return u;
}
UnitVal UnitVal::operator*(const UnitVal &o) const {
    UnitVal r(val * o.val);
    for (int i = 0; i < N_BASE_UNITS; i++) r.uni[i] = uni[i] + o.uni[i];
    return r;
}
UnitVal UnitVal::operator/(const UnitVal &o) const {
    UnitVal r(val / o.val);
    for (int i = 0; i < N_BASE_UNITS; i++) r.uni[i] = uni[i] - o.uni[i];
    return r;
}
UnitVal UnitVal::operator+(const UnitVal &o) const {
    if (!same_unit(o)) fail("Different units in addition");
    UnitVal r = *this;
    r.val += o.val;
    return r;
}
UnitVal UnitVal::pow(double p) const {
    UnitVal r(std::pow(val, p));
    for (int i = 0; i < N_BASE_UNITS; i++) r.uni[i] = uni[i] * p;
    return r;
}
bool UnitVal::same_unit(const UnitVal &o) const {
    for (int i = 0; i < N_BASE_UNITS; i++) if (uni[i] != o.uni[i]) return false;
    return true;
}
std::string UnitVal::to_string() const {
    std::string s = num_str(val), div = "/";
    for (int i = 0; i < N_BASE_UNITS; i++) if (uni[i] > 0) s += BASE_UNIT_NAMES[i] + num_str(uni[i]);
    for (int i = 0; i < N_BASE_UNITS; i++) if (uni[i] < 0) { s += div + BASE_UNIT_NAMES[i] + num_str(-uni[i]); div = ""; }
    return s;
}
std::string UnitVal::describe() const {
    std::string s = strf("%lg [ ", val);
    for (int i = 0; i < N_BASE_UNITS; i++) s += strf("%s^%lg ", BASE_UNIT_NAMES[i], uni[i]);
    return s + "]";
}

UnitEnv::UnitEnv() {
    for (int i = 0; i < N_BASE_UNITS; i++) { scale_[i] = 1.0; units_[BASE_UNIT_NAMES[i]] = UnitVal::base(i); }
    static const char *derived[][2] = {
        {"N", "1kgm/s2"}, {"Pa", "1N/m2"}, {"J", "1Nm"}, {"W", "1J/s"}, {"V", "1kgm2/t3/A"}, {"C", "1tA"},
        {"nm", "1e-9m"}, {"um", "1e-6m"}, {"mm", "1e-3m"}, {"cm", "1e-2m"}, {"km", "1e+3m"},
        {"h", "3600s"}, {"ns", "1e-9s"}, {"us", "1e-6s"}, {"ms", "1e-3s"}, {"g", "1e-3kg"}, {"mg", "1e-6kg"}};
    for (auto &d : derived) units_[d[0]] = read_text(d[1]);
    units_["d"] = UnitVal(std::atan(1.0) * 4.0 / 180.0);
    units_["%"] = UnitVal(1.0 / 100.0);
    units_["An"] = UnitVal(6.022e23);
}
UnitVal UnitEnv::read_one(const std::string &s) const {
    auto it = units_.find(s);
    return it == units_.end() ? UnitVal(0.0) : it->second;
}
// A run of letters is a product of 1- or 2-letter unit names; the exponent applies to the last one.
UnitVal UnitEnv::read_alpha(const std::string &s, double p) const {
    UnitVal one = read_one(s.substr(0, 1));
    if (s.size() < 2) return one.pow(p);
    UnitVal a = one * read_alpha(s.substr(1), p);
    UnitVal two = read_one(s.substr(0, 2));
    UnitVal b = s.size() > 2 ? two * read_alpha(s.substr(2), p) : two.pow(p);
    if (a.val == 0) return b.val == 0 ? UnitVal(0.0) : b;
    if (b.val == 0) return a;
    if (s[0] == 'm') {
        logf(L_WARNING, "Ambiguous unit: \"%s\". Interpreting \"m\" as \"milli\"", s.c_str());
        return b;
    }
    fail("Ambiguous unit: \"%s\"", s.c_str());
}
UnitVal UnitEnv::read_unit(const std::string &s) const {
    UnitVal ret(1.0);
    size_t i = 0;
    bool divide = false;
    while (i < s.size()) {
        size_t j = i;
        while (i < s.size() && (isalpha((unsigned char)s[i]) || s[i] == '%')) i++;
        size_t k = i;
        while (i < s.size() && (isdigit((unsigned char)s[i]) || s[i] == '.')) i++;
        double p = i > k ? atof(s.substr(k, i - k).c_str()) : 1.0;
        UnitVal last = k > j ? read_alpha(s.substr(j, k - j), p) : UnitVal(1.0);
        if (k > j && last.val == 0) fail("Unknown unit \"%s\"", s.substr(j, k - j).c_str());
        ret = divide ? ret / last : ret * last;
        j = i;
        while (i < s.size() && !isalnum((unsigned char)s[i]) && s[i] != '%') i++;
        if (i - j > 1) fail("Too many non-alpha-numeric characters in units: \"%s\"", s.substr(j, i - j).c_str());
        if (i - j == 1) {
            if (s[j] == '/') divide = true;
            else fail("Only \"/\" allowed in units: \"%c\"", s[j]);
        }
    }
    return ret;
}
UnitVal UnitEnv::read_text(const std::string &s) const {
    size_t i = 0;
    while (i < s.size() && (isdigit((unsigned char)s[i]) || s[i] == '-' || s[i] == '+' || s[i] == '.' || s[i] == 'e' || s[i] == 'E')) i++;
    UnitVal ret = read_unit(s.substr(i));
    if (i > 0) ret = ret * UnitVal(atof(s.substr(0, i).c_str()));
    return ret;
}
double UnitEnv::alt(const UnitVal &v) const {
    double r = v.val;
    for (int i = 0; i < N_BASE_UNITS; i++) if (v.uni[i] != 0) r *= std::pow(scale_[i], v.uni[i]);
    return r;
}
double UnitEnv::alt(const std::string &s) const {
    // sum of terms separated by +/- (a sign right after e/E belongs to the exponent)
    double ret = 0;
    size_t i = 0, j = 0;
    while (i < s.size()) {
        char c = j < s.size() ? s[j] : '\0';
        if (c == '-' || c == '+' || c == '\0') {
            if (j > i) ret += alt(read_text(s.substr(i, j - i)));
            i = j;
            if (c == '\0') break;
        } else if ((c == 'e' || c == 'E') && j + 1 < s.size() && (s[j + 1] == '-' || s[j + 1] == '+')) {
            // only an exponent if what precedes is numeric
            if (j > i && (isdigit((unsigned char)s[j - 1]) || s[j - 1] == '.')) j++;
        }
        j++;
    }
    return ret;
}
void UnitEnv::set_unit(const std::string &name, const UnitVal &v, double lattice) { gauge_[name] = v / UnitVal(lattice); }
void UnitEnv::set_unit(const std::string &name, const UnitVal &ratio) { gauge_[name] = ratio; }

void UnitEnv::make_gauge() {
    const int n = N_BASE_UNITS;
    std::vector<double> A(n * n, 0.0), b(n, 0.0), x(n, 0.0);
    int rows = 0;
    for (auto &g : gauge_) {
        if (rows >= n) fail("Gauge variables over-constructed");
        for (int j = 0; j < n; j++) A[rows * n + j] = g.second.uni[j];
        b[rows] = std::log(g.second.val);
        rows++;
    }
    for (int j = 0; j < n; j++) {          // base units no gauge touches keep scale 1
        bool used = false;
        for (int r = 0; r < rows; r++) if (A[r * n + j] != 0) used = true;
        if (!used) {
            if (rows >= n) fail("Gauge variables over-constructed");
            A[rows * n + j] = 1;
            b[rows] = 0;
            rows++;
        }
    }
    if (rows < n) fail("Gauge variables under-constructed");
    // Gaussian elimination with partial pivoting
    for (int k = 0; k < n; k++) {
        int piv = k;
        for (int i = k + 1; i < n; i++) if (std::fabs(A[i * n + k]) > std::fabs(A[piv * n + k])) piv = i;
        if (std::fabs(A[piv * n + k]) < 1e-12) fail("Gauge variables are not independent");
        if (piv != k) { for (int j = 0; j < n; j++) std::swap(A[k * n + j], A[piv * n + j]); std::swap(b[k], b[piv]); }
        for (int i = k + 1; i < n; i++) {
            double m = A[i * n + k] / A[k * n + k];
            if (m == 0) continue;
            for (int j = k; j < n; j++) A[i * n + j] -= m * A[k * n + j];
            b[i] -= m * b[k];
        }
    }
    for (int i = n - 1; i >= 0; i--) {
        double s = b[i];
        for (int j = i + 1; j < n; j++) s -= A[i * n + j] * x[j];
        x[i] = s / A[i * n + i];
    }
    for (int j = 0; j < n; j++) scale_[j] = std::exp(-x[j]);
}
void UnitEnv::print_gauge() const {
    logf(L_OUTPUT, "/---------------[ GAUGE ]-----------------");
    for (auto &g : gauge_) logf(L_OUTPUT, "|  %s", g.second.describe().c_str());
    logf(L_OUTPUT, "------------------------------------------");
    for (int j = 0; j < N_BASE_UNITS; j++) logf(L_OUTPUT, "| 1 %s = %lf units", BASE_UNIT_NAMES[j], scale_[j]);
    logf(L_OUTPUT, "\\-----------------------------------------");
}
}  // namespace tclb
