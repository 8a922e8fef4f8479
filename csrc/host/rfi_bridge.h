// External particle integrators without MPI: the force calculator (this solver) and an integrator PROGRAM exchange the
// message sequence of the reference's RemoteForceInterface (src/RemoteForceInterface.hpp:596-735: sizes -> particles ->
// forces, once per particle action; src/MPMD.hpp connects the two programs through an MPI inter-communicator) over a
// local stream socket instead.  Either side can be started first:
//   <RemoteForceInterface integrator="simplepart" spawn="bin/tclb-integrator simplepart SP.xml"/>   the solver starts the program
//   <RemoteForceInterface integrator="simplepart" connect="/tmp/rfi.sock"/>                        the program listens already
// bin/tclb-integrator is the counterpart: `simplepart <config.xml>` (the reference's src/simplepart.cpp) and `empty`
// (src/empty.cpp: takes part in every exchange with no particles).  Any other integrator can speak the same protocol:
//
//   hello      calculator -> integrator   "TCLBRFI1", int32 workers, int32 rot, int32 torque, int32 row_doubles (16)
//              integrator -> calculator   "TCLBRFI1", int32 accepted (1)
//   exchange   calculator -> integrator   int32 MSG_PARTICLES, per worker 6 doubles (lower, upper of its padded box)
//              integrator -> calculator   per worker: int64 n, n*16 doubles (RFI rows, forces zero)
//              calculator -> integrator   int32 MSG_FORCES, per worker: int64 n, n*16 doubles (forces and moments filled in)
//              calculator -> integrator   int32 MSG_STEP
//   close      calculator -> integrator   int32 MSG_CLOSE
#pragma once
#include <string>
#include <vector>
#include "particles.h"

namespace tclb {

enum { RFI_MSG_PARTICLES = 1, RFI_MSG_FORCES = 2, RFI_MSG_STEP = 3, RFI_MSG_CLOSE = 4 };
static const char RFI_MAGIC[8] = {'T', 'C', 'L', 'B', 'R', 'F', 'I', '1'};

int rfi_read_all(int fd, void *buf, size_t n);          // 0 = ok, -1 = peer closed / error
int rfi_write_all(int fd, const void *buf, size_t n);

// Calculator side: a ParticleIntegrator whose particles live in another process.
class RemoteIntegrator : public ParticleIntegrator {
  public:
    RemoteIntegrator() {}
    ~RemoteIntegrator() override;
    // one of the two; returns false with `error` set
    bool spawn(const std::string &command_line, int workers, bool rot, bool torque);
    bool connect_to(const std::string &socket_path, int workers, bool rot, bool torque, double timeout_s = 30);
    void begin_exchange() override;
    void particles_for(const WorkerBox &box, std::vector<double> &rows16) override;
    void forces_from(const WorkerBox &box, const std::vector<double> &rows16) override;
    void step() override;
    void finish() override;
    std::string error;
  private:
    bool hello(int workers, bool rot, bool torque);
    void fail(const char *what);
    int fd_ = -1, child_ = -1, workers_ = 1;
    bool dead_ = false;
    std::vector<WorkerBox> boxes_;                      // boxes asked for in this exchange, in call order
    std::vector<std::vector<double>> in_, out_;         // rows per worker: received / to send back
    size_t asked_ = 0, answered_ = 0;
    bool fetched_ = false;
};

// Integrator side: serve one calculator on `fd` with `integ` until it closes.  Returns the number of exchanges served.
long rfi_serve(int fd, ParticleIntegrator *integ);
// listen on a unix socket path and return the first connection (-1 on error)
int rfi_listen_accept(const std::string &socket_path);

// the `empty` integrator (src/empty.cpp): no particles, every exchange answered
class EmptyIntegrator : public ParticleIntegrator {
  public:
    void particles_for(const WorkerBox &, std::vector<double> &rows16) override { rows16.clear(); }
    void forces_from(const WorkerBox &, const std::vector<double> &) override {}
    void begin_exchange() override {}
    void step() override { steps++; }
    void finish() override {}
    long steps = 0;
};

}  // namespace tclb
