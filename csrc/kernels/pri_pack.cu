// Pack / unpack of one piece of a checkpoint image (csrc/core/pri.cpp) on the device: a piece is the box of one field
// inside one of the 27 margin blocks; the kernel gathers it from the ghost-padded slab into a dense run of the image
// (save) or scatters it back (load).  With it a checkpoint streams through a fixed staging buffer piece by piece
// instead of needing the whole snapshot twice in host memory (reference writer: src/Lattice.cu.Rt:642-661,708-737,
// which copies each margin block to the host in one piece).
#include <cuda_runtime.h>
#include <cstdint>

namespace {
struct Piece {
    int c[3], th[3];          // block direction and thickness per axis
    int X, Y;                 // extent of the piece in x and y (elements)
    int z0, nzz;              // slice [z0, z0 + nzz) of the piece handled by this launch
    int nx, ny, nz, gy, gz, NY;
};
template <class T, bool SAVE>
__global__ void k_pri_piece(T *__restrict__ fld, T *__restrict__ img, Piece p) {
    const long long total = (long long)p.X * p.Y * p.nzz;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int xx = (int)(i % p.X);
        const long long r = i / p.X;
        const int yy = (int)(r % p.Y), zz = p.z0 + (int)(r / p.Y);
        const int z = p.c[2] == 0 ? zz : (p.c[2] < 0 ? zz - p.th[2] : p.nz + zz);
        const int y = p.c[1] == 0 ? yy : (p.c[1] < 0 ? yy - p.th[1] : p.ny + yy);
        int x = xx;
        if (p.c[0] != 0) { x = p.c[0] < 0 ? xx - p.th[0] : p.nx + xx; x = ((x % p.nx) + p.nx) % p.nx; }   // periodic images of own columns
        T *cell = fld + ((size_t)(z + p.gz) * p.NY + (y + p.gy)) * p.nx + x;
        if (SAVE) img[i] = *cell;
        else if (p.c[0] == 0) *cell = img[i];
    }
}
template <class T>
int launch(void *fld, void *img, const Piece &p, int save, cudaStream_t st) {
    const long long total = (long long)p.X * p.Y * p.nzz;
    if (total <= 0) return 0;
    const int tb = 256;
    const int nb = (int)((total + tb - 1) / tb < 148 * 16 ? (total + tb - 1) / tb : 148 * 16);
    if (save) k_pri_piece<T, true><<<nb, tb, 0, st>>>((T *)fld, (T *)img, p);
    else k_pri_piece<T, false><<<nb, tb, 0, st>>>((T *)fld, (T *)img, p);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
}  // namespace

// geom = {nx, ny, nz, gy, gz, NY}; c/th = block direction / thickness per axis
extern "C" int tclb_pri_piece(void *field_dev, void *image_dev, int elem_size, int save, const int *c, const int *th, int X, int Y,
                              int z0, int nzz, const int *geom, void *stream) {
    Piece p;
    for (int k = 0; k < 3; k++) { p.c[k] = c[k]; p.th[k] = th[k]; }
    p.X = X; p.Y = Y; p.z0 = z0; p.nzz = nzz;
    p.nx = geom[0]; p.ny = geom[1]; p.nz = geom[2]; p.gy = geom[3]; p.gz = geom[4]; p.NY = geom[5];
    cudaStream_t st = (cudaStream_t)stream;
    if (elem_size == 2) return launch<uint16_t>(field_dev, image_dev, p, save, st);
    if (elem_size == 4) return launch<uint32_t>(field_dev, image_dev, p, save, st);
    if (elem_size == 8) return launch<unsigned long long>(field_dev, image_dev, p, save, st);
    return -1;
}
