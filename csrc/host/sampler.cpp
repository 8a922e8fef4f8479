#include "sampler.h"
#include "../core/common.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

namespace tclb {

Sampler::~Sampler() { finish(); }
void Sampler::add_point(int gx, int gy, int gz) {
    Domain *d = solver_->domain.get();
    for (int r : d->local) {
        const Box &b = d->dec.boxes[r];
        if (gx >= b.x0 && gx < b.x0 + b.nx && gy >= b.y0 && gy < b.y0 + b.ny && gz >= b.z0 && gz < b.z0 + b.nz)
            points_.push_back({gx, gy, gz, r, nullptr});
    }
}
void Sampler::allocate(const NameSet &what, int start, int total) {
    const tclb_model_t *m = solver_->model->d;
    start_iter = start;
    total_iter_ = total;
    quantities_.clear(); offsets_.clear();
    int i = 0;
    for (int q = 0; q < m->n_quantities; q++)
        if (what.in(m->quantities[q].name)) { quantities_.push_back(q); offsets_.push_back(i); i += m->quantities[q].vector ? 3 : 1; }
    size_ = i;
    size_t bytes = (size_t)size_ * total_iter_ * solver_->model->real_size();
    for (auto &p : points_) {
        Lattice *l = solver_->domain->lat[p.rank];
        if (l->backend == TCLB_BACKEND_CUDA) { TCLB_CUDA(cudaSetDevice(l->device)); TCLB_CUDA(cudaMalloc(&p.buf, bytes)); TCLB_CUDA(cudaMemset(p.buf, 0, bytes)); }
        else p.buf = calloc(1, bytes);
    }
}
void Sampler::init_csv(const std::string &filename) {
    filename_ = filename;
    const tclb_model_t *m = solver_->model->d;
    FILE *f = fopen(filename.c_str(), "wt");
    if (!f) fail("cannot write %s", filename.c_str());
    fprintf(f, "Iteration,X,Y,Z");
    for (int q : quantities_) {
        const char *n = m->quantities[q].name;
        if (m->quantities[q].vector) fprintf(f, ",%s.x,%s.y,%s.z", n, n, n);
        else fprintf(f, ",%s", n);
    }
    fprintf(f, "\n");
    fclose(f);
}
void Sampler::update(Domain *d) {
    const int rs = solver_->model->real_size();
    for (auto &p : points_) {
        Lattice *l = d->lat[p.rank];
        int row = l->iter - 1 - start_iter;           // sample of the iteration just finished
        if (row < 0 || row >= total_iter_) continue;
        Box b; b.x0 = p.x - l->region.x0; b.y0 = p.y - l->region.y0; b.z0 = p.z - l->region.z0; b.nx = b.ny = b.nz = 1;
        for (size_t k = 0; k < quantities_.size(); k++) {
            int q = quantities_[k];
            double v = solver_->units.alt(solver_->model->d->quantities[q].unit);
            char *dst = (char *)p.buf + ((size_t)row * size_ + offsets_[k]) * rs;
            l->get_quantity(q, b, dst, 1 / v, true);
        }
    }
}
void Sampler::write_history(int cur) {
    const int rs = solver_->model->real_size();
    FILE *f = fopen(filename_.c_str(), "at");
    if (!f) fail("cannot append to %s", filename_.c_str());
    int rows = cur - start_iter;
    if (rows > total_iter_) rows = total_iter_;
    std::vector<char> host((size_t)size_ * total_iter_ * rs);
    for (auto &p : points_) {
        Lattice *l = solver_->domain->lat[p.rank];
        if (l->backend == TCLB_BACKEND_CUDA) {
            TCLB_CUDA(cudaSetDevice(l->device));
            l->sync();
            TCLB_CUDA(cudaMemcpy(host.data(), p.buf, host.size(), cudaMemcpyDeviceToHost));
        } else memcpy(host.data(), p.buf, host.size());
        for (int i = 0; i < rows; i++) {
            fprintf(f, "%d,%d,%d,%d", start_iter + i, p.x, p.y, p.z);
            for (int c = 0; c < size_; c++) {
                double v = rs == 4 ? ((float *)host.data())[(size_t)i * size_ + c] : ((double *)host.data())[(size_t)i * size_ + c];
                fprintf(f, ",%.13le", v);
            }
            fprintf(f, "\n");
        }
    }
    fclose(f);
}
void Sampler::finish() {
    for (auto &p : points_) {
        if (!p.buf) continue;
        Lattice *l = solver_->domain ? solver_->domain->lat[p.rank] : nullptr;
        if (l && l->backend == TCLB_BACKEND_CUDA) { cudaSetDevice(l->device); cudaFree(p.buf); }
        else free(p.buf);
        p.buf = nullptr;
    }
    points_.clear();
    size_ = 0;
}
}  // namespace tclb
