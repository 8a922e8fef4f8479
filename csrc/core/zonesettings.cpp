#include "zonesettings.h"
#include "common.h"
#include <cstdio>

namespace tclb {

void ZoneSettings::configure(int ns, int cap) {
    ns_ = ns;
    cap_ = cap < 1 ? 1 : cap;
    consts_.assign((size_t)ns_ * cap_, 0.0);
    nzones_ = 1;
    len_ = 1;
    dirty_ = true;
}
void ZoneSettings::set_zones(int n) {
    if (n > cap_) fail("too many setting zones (%d > %d)", n, cap_);
    if (n > nzones_) { nzones_ = n; dirty_ = true; }
}
void ZoneSettings::set(int s, int zone, double v) {
    if (s < 0 || s >= ns_) fail("bad zonal setting index %d", s);
    if (zone < 0) {
        for (int z = 0; z < cap_; z++) { consts_[(size_t)s * cap_ + z] = v; series_.erase({s, z}); }
    } else {
        if (zone >= cap_) fail("zone %d out of range", zone);
        consts_[(size_t)s * cap_ + zone] = v;
        series_.erase({s, zone});
        if (zone + 1 > nzones_) nzones_ = zone + 1;
    }
    dirty_ = true;
}
void ZoneSettings::set_series(int s, int zone, const std::vector<double> &v) {
    if (v.empty()) fail("empty time series for zonal setting");
    if (len_ != 1 && (int)v.size() != len_) fail("time series length %zu differs from the one already set (%d)", v.size(), len_);
    len_ = (int)v.size();
    if (zone < 0) { for (int z = 0; z < nzones_; z++) series_[{s, z}] = v; }
    else { series_[{s, zone}] = v; if (zone + 1 > nzones_) nzones_ = zone + 1; }
    dirty_ = true;
}
bool ZoneSettings::is_series(int s, int zone) const { return series_.count({s, zone}) > 0; }
double ZoneSettings::get(int s, int zone, int t) const {
    auto it = series_.find({s, zone});
    if (it != series_.end()) return it->second[(size_t)(t % len_)];
    return consts_[(size_t)s * cap_ + zone];
}
void ZoneSettings::build_table(std::vector<double> &out, int stride) const {
    out.assign((size_t)len_ * ns_ * stride, 0.0);
    for (int t = 0; t < len_; t++)
        for (int s = 0; s < ns_; s++)
            for (int z = 0; z < stride && z < cap_; z++)
                out[((size_t)t * ns_ + s) * stride + z] = get(s, z, t);
}
void ZoneSettings::dump(const std::string &filename, const std::vector<std::string> &names) const {
    FILE *f = fopen(filename.c_str(), "wt");
    if (!f) fail("cannot write %s", filename.c_str());
    fprintf(f, "\"Iteration\"");
    for (int s = 0; s < ns_; s++)
        for (int z = 0; z < nzones_; z++) fprintf(f, ",\"%s-%d\"", names[s].c_str(), z);
    fprintf(f, "\n");
    for (int t = 0; t < len_; t++) {
        fprintf(f, "%d", t);
        for (int s = 0; s < ns_; s++)
            for (int z = 0; z < nzones_; z++) fprintf(f, ", %.13le", get(s, z, t));
        fprintf(f, "\n");
    }
    fclose(f);
}
}  // namespace tclb
