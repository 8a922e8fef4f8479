// Checkpoint images in the reference's on-disk layout (".pri"): the raw concatenation of the 27
// margin blocks (index (dx+1)+3(dy+1)+9(dz+1)+1, interior = block 14) of the current snapshot,
// each block holding, field after field, as many ghost layers as that field's stencil reaches in
// that direction (layout algebra: reference src/conf.R:897-999; writer src/Lattice.cu.Rt:642-661,
// 708-737).  Internally this framework keeps ONE ghost-padded SoA slab, so the 27-block image only
// exists here, produced by a gather (save) / scatter (load) over the slab.
#include "lattice.h"
#include "common.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

namespace tclb {

namespace {
struct BlockField { int c[3]; int th[3]; int ext[3]; size_t elems; };

// enumerate (margin, field) pieces in file order; returns false if the field is absent there
bool piece(const Lattice &L, int m, int k, BlockField &p) {
    const tclb_field_t &f = L.model->d->fields[k];
    const int mins[3] = {f.minx < 0 ? f.minx : 0, f.miny < 0 ? f.miny : 0, f.minz < 0 ? f.minz : 0};
    const int maxs[3] = {f.maxx > 0 ? f.maxx : 0, f.maxy > 0 ? f.maxy : 0, f.maxz > 0 ? f.maxz : 0};
    const int n[3] = {L.nx, L.ny, L.nz};
    p.c[0] = m % 3 - 1; p.c[1] = (m / 3) % 3 - 1; p.c[2] = m / 9 - 1;
    p.elems = 1;
    for (int d = 0; d < 3; d++) {
        p.th[d] = p.c[d] < 0 ? -mins[d] : (p.c[d] > 0 ? maxs[d] : 1);
        p.ext[d] = p.c[d] == 0 ? n[d] : 1;
        p.elems *= (size_t)p.th[d] * p.ext[d];
    }
    return p.elems > 0;
}
}  // namespace

size_t Lattice::pri_bytes() const {
    size_t tot = 0;
    BlockField p;
    for (int m = 0; m < 27; m++)
        for (int k = 0; k < model->d->n_fields; k++)
            if (piece(*this, m, k, p)) tot += p.elems;
    return tot * model->storage_size();       // the image holds storage_t values, like the reference's checkpoints
}

void Lattice::pri_block_elems(long long out[27]) const {
    BlockField p;
    for (int m = 0; m < 27; m++) {
        out[m] = 0;
        for (int k = 0; k < model->d->n_fields; k++)
            if (piece(*this, m, k, p)) out[m] += (long long)p.elems;
    }
}

template <class T, bool SAVE>
static void pri_transfer(const Lattice &L, T *slab, T *image) {
    size_t pos = 0;
    BlockField p;
    const int n[3] = {L.nx, L.ny, L.nz};
    for (int m = 0; m < 27; m++)
        for (int k = 0; k < L.model->d->n_fields; k++) {
            if (!piece(L, m, k, p)) continue;
            const int X = p.th[0] * p.ext[0], Y = p.th[1] * p.ext[1], Z = p.th[2] * p.ext[2];
            T *fld = slab + (size_t)k * L.plane_elems;
            for (int zz = 0; zz < Z; zz++) {
                int z = p.c[2] == 0 ? zz : (p.c[2] < 0 ? zz - p.th[2] : n[2] + zz);
                for (int yy = 0; yy < Y; yy++) {
                    int y = p.c[1] == 0 ? yy : (p.c[1] < 0 ? yy - p.th[1] : n[1] + yy);
                    T *row = fld + ((size_t)(z + L.gz) * L.NY + (y + L.gy)) * L.nx;
                    T *img = image + pos + ((size_t)zz * Y + yy) * X;
                    if (p.c[0] == 0) {
                        if (SAVE) memcpy(img, row, sizeof(T) * X);
                        else memcpy(row, img, sizeof(T) * X);
                    } else if (SAVE) {     // x margins are periodic images of the own columns
                        for (int xx = 0; xx < X; xx++) {
                            int x = p.c[0] < 0 ? xx - p.th[0] : n[0] + xx;
                            x = ((x % n[0]) + n[0]) % n[0];
                            img[xx] = row[x];
                        }
                    }
                }
            }
            pos += p.elems;
        }
}

extern "C" int tclb_pri_piece(void *field_dev, void *image_dev, int elem_size, int save, const int *c, const int *th, int X, int Y,
                              int z0, int nzz, const int *geom, void *stream);

// CUDA backend: the image streams through a fixed device staging buffer and a page-locked host buffer, piece by piece
// (a 1024x512x512 d3q27 checkpoint is 29 GB; nothing here is larger than the staging buffers).  TCLB_PRI_CHUNK_MB sets
// their size (default 64).
void Lattice::pri_stream(const std::string &file, bool save, void *dev_stage, char *host_stage, size_t stage_bytes) {
    Lattice &L = *this;
    const int rs = L.model->storage_size();
    FILE *f = fopen(file.c_str(), save ? "wb" : "rb");
    if (!f) fail(save ? "cannot write %s" : "cannot read %s", file.c_str());
    const int geom[6] = {L.nx, L.ny, L.nz, L.gy, L.gz, L.NY};
    BlockField p;
    size_t done = 0;
    for (int m = 0; m < 27; m++)
        for (int k = 0; k < L.model->d->n_fields; k++) {
            if (!piece(L, m, k, p)) continue;
            const int X = p.th[0] * p.ext[0], Y = p.th[1] * p.ext[1], Z = p.th[2] * p.ext[2];
            const size_t slice = (size_t)X * Y * rs;
            if (slice > stage_bytes) { fclose(f); fail("checkpoint staging buffer (%zu bytes) is smaller than one z-slice of the lattice (%zu)", stage_bytes, slice); }
            const int per = (int)std::max<size_t>(1, stage_bytes / slice);
            char *fld = (char *)L.field_ptr(L.cur, k);
            for (int z0 = 0; z0 < Z; z0 += per) {
                const int nzz = std::min(per, Z - z0);
                const size_t bytes = slice * nzz;
                if (save) {
                    if (tclb_pri_piece(fld, dev_stage, rs, 1, p.c, p.th, X, Y, z0, nzz, geom, L.stream) != 0) { fclose(f); fail("checkpoint pack kernel failed"); }
                    L.to_host(host_stage, dev_stage, bytes);
                    if (fwrite(host_stage, 1, bytes, f) != bytes) { fclose(f); fail("short write to %s", file.c_str()); }
                } else {
                    if (fread(host_stage, 1, bytes, f) != bytes) { fclose(f); fail("%s: size mismatch (expected %zu bytes): the file was written for another model, precision, lattice size or decomposition", file.c_str(), L.pri_bytes()); }
                    if (p.c[0] == 0) {
                        L.to_dev(dev_stage, host_stage, bytes);
                        if (tclb_pri_piece(fld, dev_stage, rs, 0, p.c, p.th, X, Y, z0, nzz, geom, L.stream) != 0) { fclose(f); fail("checkpoint unpack kernel failed"); }
                        L.sync();
                    }
                }
                done += bytes;
            }
        }
    if (!save) {
        const int extra = fgetc(f);
        if (extra != EOF) { fclose(f); fail("%s: size mismatch (expected %zu bytes): the file was written for another model, precision, lattice size or decomposition", file.c_str(), L.pri_bytes()); }
    }
    fclose(f);
    (void)done;
}

void Lattice::save_pri(const std::string &file) const {
    if (backend == TCLB_BACKEND_CUDA) {
        Lattice &self = const_cast<Lattice &>(*this);
        make_parent_dirs(file);
        const char *mb = getenv("TCLB_PRI_CHUNK_MB");
        const size_t stage = (size_t)std::max(1, mb ? atoi(mb) : 64) << 20;
        void *dev = self.scratch(stage);
        void *host = nullptr;
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaHostAlloc(&host, stage, cudaHostAllocDefault));
        try { self.pri_stream(file, true, dev, (char *)host, stage); } catch (...) { cudaFreeHost(host); throw; }
        cudaFreeHost(host);
        return;
    }
    std::vector<char> slab(snap_bytes), image(pri_bytes());
    to_host(slab.data(), snap_[cur], snap_bytes);
    if (model->storage_size() == 2) pri_transfer<uint16_t, true>(*this, (uint16_t *)slab.data(), (uint16_t *)image.data());
    else if (model->storage_size() == 4) pri_transfer<float, true>(*this, (float *)slab.data(), (float *)image.data());
    else pri_transfer<double, true>(*this, (double *)slab.data(), (double *)image.data());
    make_parent_dirs(file);
    FILE *f = fopen(file.c_str(), "wb");
    if (!f) fail("cannot write %s", file.c_str());
    size_t n = fwrite(image.data(), 1, image.size(), f);
    fclose(f);
    if (n != image.size()) fail("short write to %s", file.c_str());
}

void Lattice::load_pri(const std::string &file) {
    if (backend == TCLB_BACKEND_CUDA) {
        const char *mb = getenv("TCLB_PRI_CHUNK_MB");
        const size_t stage = (size_t)std::max(1, mb ? atoi(mb) : 64) << 20;
        void *dev = scratch(stage);
        void *host = nullptr;
        TCLB_CUDA(cudaSetDevice(device));
        TCLB_CUDA(cudaHostAlloc(&host, stage, cudaHostAllocDefault));
        try { pri_stream(file, false, dev, (char *)host, stage); } catch (...) { cudaFreeHost(host); throw; }
        cudaFreeHost(host);
        return;
    }
    std::vector<char> slab(snap_bytes), image(pri_bytes());
    FILE *f = fopen(file.c_str(), "rb");
    if (!f) fail("cannot read %s", file.c_str());
    size_t n = fread(image.data(), 1, image.size(), f);
    int extra = fgetc(f);
    fclose(f);
    if (n != image.size() || extra != EOF)
        fail("%s: size mismatch (expected %zu bytes): the file was written for another model, precision, lattice size or decomposition",
             file.c_str(), image.size());
    to_host(slab.data(), snap_[cur], snap_bytes);
    if (model->storage_size() == 2) pri_transfer<uint16_t, false>(*this, (uint16_t *)slab.data(), (uint16_t *)image.data());
    else if (model->storage_size() == 4) pri_transfer<float, false>(*this, (float *)slab.data(), (float *)image.data());
    else pri_transfer<double, false>(*this, (double *)slab.data(), (double *)image.data());
    to_dev(snap_[cur], slab.data(), snap_bytes);
}
}  // namespace tclb
