// Uniform-grid container over the particle table: which particles can touch a node is answered by looking at the (at most)
// eight cells around it instead of at every particle.
// Role of the reference's SolidGrid (src/SolidGrid.h, src/SolidGrid.hpp:11-60: cell edge = 2 * largest radius, a fixed number
// of slots ("depth") per cell, rebuilt with a larger depth when a cell overflows); there it is built on the host every
// iteration and copied, here the table is filled on the device right after the particle rows arrive (one thread per
// particle, one atomic per particle), the host only picks the cell size and checks the overflow word.
#include <cuda_runtime.h>
#include <cmath>
#include <cstring>

namespace {
template <class T>
__host__ __device__ inline void put_particle(const T *rows, int i, const double *origin, double inv_delta, const int *gn, int depth,
                                             int *grid, int *counts, int *overflow) {
    const T *row = rows + (size_t)i * 16;
    int c[3];
    for (int k = 0; k < 3; k++) {
        c[k] = (int)floor(((double)row[1 + k] - origin[k]) * inv_delta);
        if (c[k] < 0 || c[k] >= gn[k]) return;        // too far from this subdomain to touch any of its nodes
    }
    const size_t cell = ((size_t)c[2] * gn[1] + c[1]) * gn[0] + c[0];
#ifdef __CUDA_ARCH__
    const int slot = atomicAdd(counts + cell, 1);
#else
    const int slot = counts[cell]++;
#endif
    if (slot < depth) grid[cell * depth + slot] = i;
    else *overflow = 1;
}
struct GridArgs { double origin[3]; double inv_delta; int gn[3]; int depth; };
template <class T>
__global__ void k_particle_grid(const T *rows, int n, GridArgs g, int *grid, int *counts, int *overflow) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) put_particle(rows, i, g.origin, g.inv_delta, g.gn, g.depth, grid, counts, overflow);
}
}  // namespace

// grid: int[cells][depth] (-1 = empty), counts: int[cells], overflow: int[1]; all three are cleared here.
extern "C" int tclb_particle_grid_fill(const void *rows, int n, int real_size, const double *origin, double inv_delta, const int *gn,
                                       int depth, int *grid, int *counts, int *overflow, int on_device, void *stream) {
    const size_t cells = (size_t)gn[0] * gn[1] * gn[2];
    if (!on_device) {
        memset(grid, 0xFF, cells * depth * sizeof(int));
        memset(counts, 0, cells * sizeof(int));
        *overflow = 0;
        for (int i = 0; i < n; i++) {
            if (real_size == 8) put_particle((const double *)rows, i, origin, inv_delta, gn, depth, grid, counts, overflow);
            else put_particle((const float *)rows, i, origin, inv_delta, gn, depth, grid, counts, overflow);
        }
        return 0;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (cudaMemsetAsync(grid, 0xFF, cells * depth * sizeof(int), st) != cudaSuccess) return -1;
    if (cudaMemsetAsync(counts, 0, cells * sizeof(int), st) != cudaSuccess) return -1;
    if (cudaMemsetAsync(overflow, 0, sizeof(int), st) != cudaSuccess) return -1;
    if (n == 0) return 0;
    GridArgs g;
    for (int k = 0; k < 3; k++) { g.origin[k] = origin[k]; g.gn[k] = gn[k]; }
    g.inv_delta = inv_delta; g.depth = depth;
    const int tb = 128, nb = (n + tb - 1) / tb;
    if (real_size == 8) k_particle_grid<double><<<nb, tb, 0, st>>>((const double *)rows, n, g, grid, counts, overflow);
    else k_particle_grid<float><<<nb, tb, 0, st>>>((const float *)rows, n, g, grid, counts, overflow);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
