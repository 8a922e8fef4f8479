#include "particles.h"
#include "../core/common.h"
#include <cmath>
#include <cstdlib>

namespace tclb {

static int axis_of(const std::string &name, std::string &stem) {
    stem = name;
    if (name.empty()) return -1;
    const char c = name.back();
    const int d = c == 'x' ? 0 : c == 'y' ? 1 : c == 'z' ? 2 : -1;
    if (d >= 0) stem = name.substr(0, name.size() - 1);
    return d;
}

SimplePart::SimplePart(const XmlNode *cfg, const std::string &default_log, bool rot, bool torque) : rot_(rot), torque_(torque) {
    if (!cfg || cfg->name != "SimplePart") fail("SimplePart: the configuration needs a <SimplePart> element");
    for (auto &a : cfg->attrs) {
        if (a.first == "dt") dt_ = atof(a.second.c_str());
        else if (a.first == "ax") acc_[0] = atof(a.second.c_str());
        else if (a.first == "ay") acc_[1] = atof(a.second.c_str());
        else if (a.first == "az") acc_[2] = atof(a.second.c_str());
        else if (a.first == "afreq") acc_freq_ = atof(a.second.c_str());
        else fail("SimplePart: unknown attribute '%s'", a.first.c_str());
    }
    for (auto *n : cfg->children()) {
        if (n->name == "Particle") {
            P p;
            for (auto &a : n->attrs) {
                std::string stem;
                const int d = axis_of(a.first, stem);
                const double v = atof(a.second.c_str());
                if (d >= 0 && stem == "") p.x[d] = v;
                else if (d >= 0 && stem == "v") p.v0[d] = v;
                else if (d >= 0 && stem == "omega") p.omega0[d] = v;
                else if (a.first == "r") p.r = v;
                else if (a.first == "m") p.m = v;
                else if (a.first == "log") p.logging = n->attr_bool("log");
                else if (a.first == "ease-in") p.ease_in = v;
                else fail("SimplePart: unknown attribute '%s' in Particle", a.first.c_str());
            }
            if (p.r <= 0) fail("SimplePart: specify the radius with the 'r' attribute");
            for (int i = 0; i < 3; i++) { p.v[i] = p.ease_in > 0 ? 0 : p.v0[i]; p.omega[i] = p.ease_in > 0 ? 0 : p.omega0[i]; }
            p.n = particles.size();
            particles.push_back(p);
        } else if (n->name == "Periodic") {
            for (auto &a : n->attrs) {
                std::string stem;
                const int d = axis_of(a.first, stem);
                if (d >= 0 && stem == "") { periodic_[d] = true; period_[d] = atof(a.second.c_str()); }
                else if (d >= 0 && stem == "p") { /* periodic origin: accepted, unused (as in the reference) */ }
                else fail("SimplePart: unknown attribute '%s' in Periodic", a.first.c_str());
            }
        } else if (n->name == "Log") {
            if (logging_) fail("SimplePart: there can be only one Log element");
            logging_ = true;
            for (auto &a : n->attrs) {
                if (a.first == "name") log_name_ = a.second;
                else if (a.first == "Iterations") { log_iter_ = atoi(a.second.c_str()); if (log_iter_ < 1) fail("SimplePart: Log Iterations must be at least 1"); }
                else if (a.first == "average") avg_ = n->attr_bool("average");
                else if (a.first == "rotation") log_omega_ = n->attr_bool("rotation");
                else fail("SimplePart: unknown attribute '%s' in Log", a.first.c_str());
            }
            if (log_name_.empty()) log_name_ = default_log;
        } else fail("SimplePart: unknown element '%s'", n->name.c_str());
    }
    if (logging_) {
        make_parent_dirs(log_name_);
        log_ = fopen(log_name_.c_str(), "w");
        if (!log_) fail("SimplePart: cannot write %s", log_name_.c_str());
        fprintf(log_, "Iteration,Time");
        for (auto &p : particles) if (p.logging) {
            const long n = (long)p.n;
            fprintf(log_, ",p%ld_x,p%ld_y,p%ld_z,p%ld_vx,p%ld_vy,p%ld_vz,p%ld_fx,p%ld_fy,p%ld_fz", n, n, n, n, n, n, n, n, n);
            if (log_omega_) fprintf(log_, ",p%ld_ox,p%ld_oy,p%ld_oz,p%ld_tx,p%ld_ty,p%ld_tz", n, n, n, n, n, n);
        }
        fprintf(log_, "\n");
    }
    // leap-frog start of the prescribed oscillation
    for (auto &p : particles) for (int i = 0; i < 3; i++) p.v[i] -= acc_[i] / 2.0;
}
SimplePart::~SimplePart() { if (log_) fclose(log_); }

template <class F>
void SimplePart::images(const WorkerBox &box, const P &p, F &&fn) const {
    int lo[3], hi[3];
    for (int j = 0; j < 3; j++) {
        if (periodic_[j]) {
            hi[j] = (int)std::floor((box.upper[j] - p.x[j] + p.r) / period_[j]);
            lo[j] = (int)std::ceil((box.lower[j] - p.x[j] - p.r) / period_[j]);
        } else if (p.x[j] + p.r >= box.lower[j] && p.x[j] - p.r <= box.upper[j]) { lo[j] = 0; hi[j] = 0; }
        else { lo[j] = 0; hi[j] = -1; }
    }
    for (int a = lo[0]; a <= hi[0]; a++)
        for (int b = lo[1]; b <= hi[1]; b++)
            for (int c = lo[2]; c <= hi[2]; c++) {
                const double px[3] = {p.x[0] + a * period_[0], p.x[1] + b * period_[1], p.x[2] + c * period_[2]};
                fn(px);
            }
}
void SimplePart::begin_exchange() {
    for (auto &p : particles) for (int i = 0; i < 3; i++) { p.f[i] = 0; p.torque[i] = 0; }
}
void SimplePart::particles_for(const WorkerBox &box, std::vector<double> &rows) {
    rows.clear();
    for (auto &p : particles)
        images(box, p, [&](const double px[3]) {
            double row[16] = {0};
            row[0] = p.r;
            for (int i = 0; i < 3; i++) { row[1 + i] = px[i]; row[4 + i] = p.v[i]; row[7 + i] = rot_ ? p.omega[i] : 0.0; }
            rows.insert(rows.end(), row, row + 16);
        });
}
void SimplePart::forces_from(const WorkerBox &box, const std::vector<double> &rows) {
    size_t k = 0;
    for (auto &p : particles)
        images(box, p, [&](const double *) {
            if ((k + 1) * 16 > rows.size()) fail("SimplePart: fewer force rows than particle images");
            const double *row = rows.data() + k * 16;
            for (int i = 0; i < 3; i++) {
                const double f = row[10 + i], t = row[13 + i];
                p.f[i] += std::isfinite(f) ? f : 0.0;
                if (torque_) p.torque[i] += std::isfinite(t) ? t : 0.0;
            }
            k++;
        });
}
void SimplePart::step() {
    static const double pi = 4 * atan(1.0), twopi = 8 * atan(1.0);
    if (logging_ && iter % log_iter_ == 0) {
        fprintf(log_, "%d,%.15lg", iter, dt_ * iter);
        for (auto &p : particles) if (p.logging) {
            for (int i = 0; i < 3; i++) fprintf(log_, ",%.15lg", p.x[i]);
            for (int i = 0; i < 3; i++) fprintf(log_, ",%.15lg", p.v[i]);
            for (int i = 0; i < 3; i++) fprintf(log_, ",%.15lg", avg_ ? p.favg[i] / log_iter_ : p.f[i]);
            if (avg_) for (int i = 0; i < 3; i++) p.favg[i] = 0;
            if (log_omega_) { for (int i = 0; i < 3; i++) fprintf(log_, ",%.15lg", p.omega[i]); for (int i = 0; i < 3; i++) fprintf(log_, ",%.15lg", p.torque[i]); }
        }
        fprintf(log_, "\n");
        fflush(log_);
    }
    for (auto &p : particles) {
        const double t = dt_ * iter;
        for (int i = 0; i < 3; i++) p.favg[i] += p.f[i];
        if (p.ease_in > 0) {
            if (p.ease_in > t) {
                const double fac = (1 - cos(pi * t / p.ease_in)) * 0.5;
                for (int i = 0; i < 3; i++) { p.v[i] = p.v0[i] * fac; p.omega[i] = p.omega0[i] * fac; }
            } else {
                for (int i = 0; i < 3; i++) { p.v[i] = p.v0[i]; p.omega[i] = p.omega0[i]; }
                p.ease_in = 0;
            }
        } else {
            if (p.m > 0) for (int i = 0; i < 3; i++) p.v[i] += p.f[i] / p.m * dt_;
            for (int i = 0; i < 3; i++) p.v[i] += acc_[i] * cos(twopi * t * acc_freq_);
        }
        for (int i = 0; i < 3; i++) p.x[i] += p.v[i] * dt_;
    }
    iter++;
}
void SimplePart::finish() {
    // the stand-alone integrator notices the closed connection only after one more pass of its loop, in which the
    // force buffers still hold the last values: one more log row and one more step (kept for identical logs)
    step();
}

}  // namespace tclb
