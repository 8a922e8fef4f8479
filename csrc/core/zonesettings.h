// Per-zone, optionally time-dependent settings (reference src/ZoneSettings.h:9-276):
// a value (or a time series) per (zonal setting, zone); zone id = NodeType >> ZONE_SHIFT.
// The device sees one dense table real_t[len][n_zonesettings][stride]; kernels index it with
// the current time index, and form the _DT central difference from its two time neighbours.
#pragma once
#include <map>
#include <string>
#include <vector>

namespace tclb {

class ZoneSettings {
  public:
    ZoneSettings() : ns_(0), nzones_(1), len_(1), dirty_(true) {}
    void configure(int n_zonesettings, int zone_capacity);
    int n_settings() const { return ns_; }
    int zones() const { return nzones_; }          // zones in use (max zone index + 1)
    int len() const { return len_; }
    void set_zones(int n);                         // grow the number of zones in use
    // zone < 0 : all zones
    void set(int s, int zone, double v);
    void set_series(int s, int zone, const std::vector<double> &v);
    double get(int s, int zone, int t) const;
    bool is_series(int s, int zone) const;
    bool dirty() const { return dirty_; }
    void clear_dirty() { dirty_ = false; }
    // dense table [len][ns][stride] as doubles
    void build_table(std::vector<double> &out, int stride) const;
    void dump(const std::string &filename, const std::vector<std::string> &names) const;
  private:
    int ns_, nzones_, len_;
    bool dirty_;
    int cap_;
    std::vector<double> consts_;                              // [ns][cap]
    std::map<std::pair<int, int>, std::vector<double>> series_;
};

}  // namespace tclb
