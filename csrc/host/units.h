// Physical units: values carrying exponents of 9 base units, a text parser ("10cm", "0.001sPa",
// "1kgm2/t3/A", "120d", "5%", sums "10cm+2") and a gauge that fixes the SI -> lattice scaling of every
// base unit from a few (value, lattice value) pairs by solving a log-linear system.
// Same behaviour as the reference's UnitVal/UnitEnv (src/unit.h:17-192, src/unit.cpp:58-278),
// written independently (std::array exponents, recursive-descent tokenizer, LU with pivoting).
#pragma once
#include <array>
#include <map>
#include <string>
#include <vector>

namespace tclb {

constexpr int N_BASE_UNITS = 9;
extern const char *const BASE_UNIT_NAMES[N_BASE_UNITS];   // m s kg K x y z A t

struct UnitVal {
    double val = 0.0;
    std::array<double, N_BASE_UNITS> uni{};
    UnitVal() {}
    UnitVal(double v) : val(v) {}
    static UnitVal base(int k);
    UnitVal operator*(const UnitVal &o) const;
    UnitVal operator/(const UnitVal &o) const;
    UnitVal operator+(const UnitVal &o) const;      // throws on unit mismatch
    UnitVal pow(double p) const;
    bool same_unit(const UnitVal &o) const;
    std::string to_string() const;                   // "0.01m1/s1" style
    std::string describe() const;                    // "0.01 [ m^1 s^-1 ... ]"
};

class UnitEnv {
  public:
    UnitEnv();
    UnitVal read_text(const std::string &s) const;   // "<number><unit expr>"
    UnitVal read_unit(const std::string &s) const;   // unit expression only
    double si(const std::string &s) const { return read_text(s).val; }
    double si(const std::string &s, double def) const { return s.empty() ? def : si(s); }
    double alt(const UnitVal &v) const;              // SI -> lattice units
    double alt(const std::string &s) const;          // sum of signed terms
    double alt(const std::string &s, double def) const { return s.empty() ? def : alt(s); }
    void set_unit(const std::string &name, const UnitVal &si_value, double lattice_value);
    void set_unit(const std::string &name, const UnitVal &ratio);
    void make_gauge();
    void print_gauge() const;
    double scale(int k) const { return scale_[k]; }
    size_t n_gauge() const { return gauge_.size(); }
  private:
    UnitVal read_one(const std::string &s) const;
    UnitVal read_alpha(const std::string &s, double p) const;
    std::map<std::string, UnitVal> units_;
    std::map<std::string, UnitVal> gauge_;
    double scale_[N_BASE_UNITS];
};

}  // namespace tclb
