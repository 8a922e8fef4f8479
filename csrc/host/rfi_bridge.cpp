// see rfi_bridge.h
#include "rfi_bridge.h"
#include <cerrno>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <sstream>
#include <sys/socket.h>
#include <sys/types.h>
#include <sys/un.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>
#include "../core/common.h"

namespace tclb {

int rfi_read_all(int fd, void *buf, size_t n) {
    char *p = (char *)buf;
    while (n > 0) {
        ssize_t k = ::read(fd, p, n);
        if (k == 0) return -1;
        if (k < 0) { if (errno == EINTR) continue; return -1; }
        p += k; n -= (size_t)k;
    }
    return 0;
}
int rfi_write_all(int fd, const void *buf, size_t n) {
    const char *p = (const char *)buf;
    while (n > 0) {
        ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
        if (k < 0) { if (errno == EINTR) continue; return -1; }
        p += k; n -= (size_t)k;
    }
    return 0;
}
static int put_i32(int fd, int32_t v) { return rfi_write_all(fd, &v, 4); }
static int get_i32(int fd, int32_t *v) { return rfi_read_all(fd, v, 4); }

RemoteIntegrator::~RemoteIntegrator() {
    if (fd_ >= 0) { if (!dead_) put_i32(fd_, RFI_MSG_CLOSE); ::close(fd_); }
    if (child_ > 0) { int st = 0; waitpid(child_, &st, 0); }
}
void RemoteIntegrator::fail(const char *what) {
    if (!dead_) logf(L_ERROR, "RemoteForceInterface: connection to the integrator lost (%s)", what);
    dead_ = true;
}
bool RemoteIntegrator::hello(int workers, bool rot, bool torque) {
    workers_ = workers;
    int32_t head[4] = {workers, rot ? 1 : 0, torque ? 1 : 0, 16};
    char magic[8];
    int32_t ok = 0;
    if (rfi_write_all(fd_, RFI_MAGIC, 8) || rfi_write_all(fd_, head, sizeof(head)) || rfi_read_all(fd_, magic, 8) || get_i32(fd_, &ok) ||
        memcmp(magic, RFI_MAGIC, 8) != 0 || ok != 1) {
        error = "the integrator did not answer the handshake";
        return false;
    }
    return true;
}
bool RemoteIntegrator::spawn(const std::string &command_line, int workers, bool rot, bool torque) {
    int sv[2];
    if (socketpair(AF_UNIX, SOCK_STREAM, 0, sv) != 0) { error = std::string("socketpair: ") + strerror(errno); return false; }
    std::vector<std::string> words;
    { std::istringstream is(command_line); std::string w; while (is >> w) words.push_back(w); }
    if (words.empty()) { error = "empty spawn command"; return false; }
    pid_t pid = fork();
    if (pid < 0) { error = std::string("fork: ") + strerror(errno); return false; }
    if (pid == 0) {
        ::close(sv[0]);
        std::string fdarg = std::to_string(sv[1]);
        std::vector<char *> argv;
        for (auto &w : words) argv.push_back(const_cast<char *>(w.c_str()));
        argv.push_back(const_cast<char *>("--fd"));
        argv.push_back(const_cast<char *>(fdarg.c_str()));
        argv.push_back(nullptr);
        execvp(argv[0], argv.data());
        fprintf(stderr, "RemoteForceInterface: cannot start %s: %s\n", argv[0], strerror(errno));
        _exit(127);
    }
    ::close(sv[1]);
    fd_ = sv[0];
    child_ = pid;
    return hello(workers, rot, torque);
}
bool RemoteIntegrator::connect_to(const std::string &path, int workers, bool rot, bool torque, double timeout_s) {
    sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    if (path.size() >= sizeof(addr.sun_path)) { error = "socket path too long"; return false; }
    strcpy(addr.sun_path, path.c_str());
    timespec t0; clock_gettime(CLOCK_MONOTONIC, &t0);
    for (;;) {
        fd_ = socket(AF_UNIX, SOCK_STREAM, 0);
        if (fd_ < 0) { error = std::string("socket: ") + strerror(errno); return false; }
        if (::connect(fd_, (sockaddr *)&addr, sizeof(addr)) == 0) break;
        ::close(fd_); fd_ = -1;
        timespec t; clock_gettime(CLOCK_MONOTONIC, &t);
        if ((t.tv_sec - t0.tv_sec) + 1e-9 * (t.tv_nsec - t0.tv_nsec) > timeout_s) { error = "no integrator listening on " + path; return false; }
        usleep(50000);
    }
    return hello(workers, rot, torque);
}
void RemoteIntegrator::begin_exchange() {
    boxes_.clear(); in_.clear(); out_.clear();
    asked_ = answered_ = 0;
    fetched_ = false;
}
// The calculator asks worker by worker, the wire carries all workers of this process in one message: the first
// particles_for() of an exchange cannot know how many will follow, so the request is sent per worker (workers = 1 on the
// wire per request) and the integrator keeps answering until MSG_FORCES arrives.
void RemoteIntegrator::particles_for(const WorkerBox &box, std::vector<double> &rows16) {
    rows16.clear();
    boxes_.push_back(box);
    in_.emplace_back();
    if (dead_) return;
    int64_t n = 0;
    if (put_i32(fd_, RFI_MSG_PARTICLES) || rfi_write_all(fd_, box.lower, 3 * sizeof(double)) || rfi_write_all(fd_, box.upper, 3 * sizeof(double)) ||
        rfi_read_all(fd_, &n, 8) || n < 0) { fail("particles"); return; }
    in_.back().resize((size_t)n * 16);
    if (n > 0 && rfi_read_all(fd_, in_.back().data(), (size_t)n * 16 * sizeof(double))) { fail("particle rows"); in_.back().clear(); return; }
    rows16 = in_.back();
}
void RemoteIntegrator::forces_from(const WorkerBox &box, const std::vector<double> &rows16) {
    (void)box;
    out_.push_back(rows16);
    if (out_.size() < boxes_.size() || dead_) return;
    if (put_i32(fd_, RFI_MSG_FORCES) || put_i32(fd_, (int32_t)out_.size())) { fail("forces"); return; }
    for (size_t w = 0; w < out_.size(); w++) {
        int64_t n = (int64_t)(out_[w].size() / 16);
        if (rfi_write_all(fd_, boxes_[w].lower, 3 * sizeof(double)) || rfi_write_all(fd_, boxes_[w].upper, 3 * sizeof(double)) ||
            rfi_write_all(fd_, &n, 8) || (n > 0 && rfi_write_all(fd_, out_[w].data(), out_[w].size() * sizeof(double)))) { fail("force rows"); return; }
    }
}
void RemoteIntegrator::step() {
    if (!dead_ && put_i32(fd_, RFI_MSG_STEP)) fail("step");
}
void RemoteIntegrator::finish() {
    if (fd_ >= 0 && !dead_) put_i32(fd_, RFI_MSG_CLOSE);
    if (fd_ >= 0) { ::close(fd_); fd_ = -1; }
    if (child_ > 0) { int st = 0; waitpid(child_, &st, 0); child_ = -1; }
}

long rfi_serve(int fd, ParticleIntegrator *integ) {
    char magic[8];
    int32_t head[4];
    if (rfi_read_all(fd, magic, 8) || memcmp(magic, RFI_MAGIC, 8) != 0 || rfi_read_all(fd, head, sizeof(head)) || head[3] != 16) return -1;
    if (rfi_write_all(fd, RFI_MAGIC, 8) || put_i32(fd, 1)) return -1;
    long exchanges = 0;
    bool open_exchange = false;
    std::vector<double> rows;
    for (;;) {
        int32_t msg = 0;
        if (get_i32(fd, &msg)) { integ->finish(); return exchanges; }       // calculator went away
        if (msg == RFI_MSG_PARTICLES) {
            WorkerBox box;
            if (rfi_read_all(fd, box.lower, 3 * sizeof(double)) || rfi_read_all(fd, box.upper, 3 * sizeof(double))) break;
            if (!open_exchange) { integ->begin_exchange(); open_exchange = true; }
            integ->particles_for(box, rows);
            int64_t n = (int64_t)(rows.size() / 16);
            if (rfi_write_all(fd, &n, 8) || (n > 0 && rfi_write_all(fd, rows.data(), rows.size() * sizeof(double)))) break;
        } else if (msg == RFI_MSG_FORCES) {
            int32_t nw = 0;
            if (get_i32(fd, &nw)) break;
            bool bad = false;
            for (int w = 0; w < nw && !bad; w++) {
                WorkerBox box;
                int64_t n = 0;
                if (rfi_read_all(fd, box.lower, 3 * sizeof(double)) || rfi_read_all(fd, box.upper, 3 * sizeof(double)) || rfi_read_all(fd, &n, 8) || n < 0) { bad = true; break; }
                rows.resize((size_t)n * 16);
                if (n > 0 && rfi_read_all(fd, rows.data(), rows.size() * sizeof(double))) { bad = true; break; }
                integ->forces_from(box, rows);
            }
            if (bad) break;
        } else if (msg == RFI_MSG_STEP) {
            integ->step();
            open_exchange = false;
            exchanges++;
        } else if (msg == RFI_MSG_CLOSE) {
            integ->finish();
            return exchanges;
        } else break;
    }
    integ->finish();
    return exchanges;
}

int rfi_listen_accept(const std::string &path) {
    sockaddr_un addr;
    memset(&addr, 0, sizeof(addr));
    addr.sun_family = AF_UNIX;
    if (path.size() >= sizeof(addr.sun_path)) return -1;
    strcpy(addr.sun_path, path.c_str());
    ::unlink(path.c_str());
    int ls = socket(AF_UNIX, SOCK_STREAM, 0);
    if (ls < 0) return -1;
    if (bind(ls, (sockaddr *)&addr, sizeof(addr)) != 0 || listen(ls, 1) != 0) { ::close(ls); return -1; }
    int fd = accept(ls, nullptr, nullptr);
    ::close(ls);
    ::unlink(path.c_str());
    return fd;
}

}  // namespace tclb
