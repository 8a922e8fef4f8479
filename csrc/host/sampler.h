// Point probes (reference src/Sampler.{h,cpp}, Lattice::updateAllSamples src/Lattice.cu.Rt:1376-1389):
// after every iteration the selected quantities at each probe are evaluated ON the device into a
// device-resident history buffer; the history is flushed to a CSV file when the callback fires.
#pragma once
#include <map>
#include <string>
#include <vector>
#include "solver.h"

namespace tclb {

class Sampler {
  public:
    explicit Sampler(Solver *s) : solver_(s) {}
    ~Sampler();
    void add_point(int gx, int gy, int gz);
    void allocate(const NameSet &what, int start_iter, int total_iter);
    void init_csv(const std::string &filename);
    void update(Domain *d);                      // one sample row per point (called after an iteration)
    void write_history(int current_iter);
    void finish();
    bool active() const { return size_ > 0 && !points_.empty(); }
    int start_iter = 0;
  private:
    struct Point { int x, y, z, rank; void *buf; };
    Solver *solver_;
    std::vector<Point> points_;
    std::vector<int> quantities_, offsets_;
    int size_ = 0, total_iter_ = 0;
    std::string filename_;
};

}  // namespace tclb
