// Checkpoint images in the reference's on-disk layout (".pri"): the raw concatenation of the 27
// margin blocks (index (dx+1)+3(dy+1)+9(dz+1)+1, interior = block 14) of the current snapshot,
// each block holding, field after field, as many ghost layers as that field's stencil reaches in
// that direction (layout algebra: reference src/conf.R:897-999; writer src/Lattice.cu.Rt:642-661,
// 708-737).  Internally this framework keeps ONE ghost-padded SoA slab, so the 27-block image only
// exists here, produced by a gather (save) / scatter (load) over the slab.
#include "lattice.h"
#include "common.h"
#include <cstdio>
#include <cstring>

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

void Lattice::save_pri(const std::string &file) const {
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
