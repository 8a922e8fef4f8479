// Particle coupling without MPI: the force calculator (this solver) and the integrator live in one process.
//
// The reference couples external integrators over MPI inter-communicators (RemoteForceInterface + MPMD,
// src/RemoteForceInterface.hpp, src/MPMD.hpp) and ships `simplepart` (src/simplepart.cpp) as the reference
// integrator.  Here an integrator is an object behind a small interface; SimplePart is the same integrator
// (prescribed oscillating acceleration, optional mass, periodic images, CSV log) configured by the same
// <SimplePart> XML.  One exchange per particle action: particles in -> particle stages -> forces out -> step.
#pragma once
#include <array>
#include <cstdio>
#include <string>
#include <vector>
#include "xml.h"

namespace tclb {

struct WorkerBox { double lower[3], upper[3]; };          // subdomain of one rank, padded by the particle margin

class ParticleIntegrator {
  public:
    virtual ~ParticleIntegrator() {}
    // rows of 16 doubles (RFI layout) for every particle image that touches `box`; forces/moments zero
    virtual void particles_for(const WorkerBox &box, std::vector<double> &rows16) = 0;
    // the same rows back with force/moment filled in by the particle stages (call once per box, same order)
    virtual void forces_from(const WorkerBox &box, const std::vector<double> &rows16) = 0;
    virtual void begin_exchange() = 0;                      // forces of all particles := 0
    virtual void step() = 0;                                // log, integrate, advance the iteration counter
    virtual void finish() = 0;                              // what the integrator does when the calculator closes
};

class SimplePart : public ParticleIntegrator {
  public:
    struct P {
        double x[3] = {0, 0, 0}, r = 0, m = 0, v[3] = {0, 0, 0}, v0[3] = {0, 0, 0}, f[3] = {0, 0, 0}, favg[3] = {0, 0, 0};
        double omega[3] = {0, 0, 0}, omega0[3] = {0, 0, 0}, torque[3] = {0, 0, 0}, ease_in = 0;
        size_t n = 0;
        bool logging = false;
    };
    explicit SimplePart(const XmlNode *config, const std::string &default_log, bool rot, bool torque);
    ~SimplePart() override;
    void particles_for(const WorkerBox &box, std::vector<double> &rows16) override;
    void forces_from(const WorkerBox &box, const std::vector<double> &rows16) override;
    void begin_exchange() override;
    void step() override;
    void finish() override;
    std::vector<P> particles;
    int iter = 0;
  private:
    double dt_ = 1, acc_[3] = {0, 0, 0}, acc_freq_ = 0, period_[3] = {0, 0, 0};
    bool periodic_[3] = {false, false, false}, rot_, torque_;
    bool logging_ = false, avg_ = false, log_omega_ = false;
    int log_iter_ = 1;
    std::string log_name_;
    FILE *log_ = nullptr;
    template <class F> void images(const WorkerBox &box, const P &p, F &&fn) const;
};

}  // namespace tclb
