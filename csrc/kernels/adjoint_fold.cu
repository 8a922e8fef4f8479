// Helpers of the adjoint sweep that do not depend on the model (float and double lattices).
//
// tclb_adjoint_fold_ghosts: the adjoint stage scatters into the cells its inputs were pulled from, which includes the
// y/z ghost layers; a ghost cell is an image of an interior cell of the periodic neighbour (here: the same subdomain,
// wrapped), so its value is added to that cell and cleared.  The primal halo is a copy owner -> ghost; this is its
// transpose.  (Reference: the adjoint border kernels + reversed MPI exchange, src/Lattice.cu.Rt:589-640.)
#include <cuda_runtime.h>
#include <cstdio>

namespace {
// one thread per ghost cell: blockIdx.y enumerates the ghost rows (the 2*gz whole z-planes first, then the 2*gy rows
// of every interior plane), blockIdx.z the field; interior cells are never visited
template <class T>
__global__ void k_fold_ghosts(T *adj, int nx, int ny, int nz, int gy, int gz) {
    const int NY = ny + 2 * gy, NZ = nz + 2 * gz;
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= nx) return;
    int r = blockIdx.y, Y, Z;
    const int nzg = 2 * gz * NY;
    if (r < nzg) { const int zi = r / NY; Y = r - zi * NY; Z = zi < gz ? zi : zi + nz; }
    else { r -= nzg; const int zi = r / (2 * gy), k = r - zi * 2 * gy; Z = gz + zi; Y = k < gy ? k : k + ny; }
    T *base = adj + (size_t)blockIdx.z * ((size_t)nx * NY * NZ);
    T *src = base + ((size_t)Z * NY + Y) * nx + x;
    const T v = *src;
    if (v == (T)0) return;
    const int y = Y - gy, z = Z - gz;
    const int yy = ((y % ny) + ny) % ny, zz = ((z % nz) + nz) % nz;
    atomicAdd(base + ((size_t)(zz + gz) * NY + (yy + gy)) * nx + x, v);
    *src = (T)0;
}
template <class T>
void fold_host(T *adj, int nfields, int nx, int ny, int nz, int gy, int gz) {
    const int NY = ny + 2 * gy, NZ = nz + 2 * gz;
    const long long plane = (long long)nx * NY * NZ;
    for (int f = 0; f < nfields; f++)
        for (int Z = 0; Z < NZ; Z++)
            for (int Y = 0; Y < NY; Y++) {
                const int y = Y - gy, z = Z - gz;
                if (y >= 0 && y < ny && z >= 0 && z < nz) continue;
                const int yy = ((y % ny) + ny) % ny, zz = ((z % nz) + nz) % nz;
                T *src = adj + f * plane + ((long long)Z * NY + Y) * nx;
                T *dst = adj + f * plane + ((long long)(zz + gz) * NY + (yy + gy)) * nx;
                for (int x = 0; x < nx; x++) { dst[x] += src[x]; src[x] = (T)0; }
            }
}
}  // namespace

extern "C" int tclb_adjoint_fold_ghosts(void *adj, int real_size, int on_device, int nfields, int nx, int ny, int nz, int gy, int gz,
                                        void *stream) {
    if (gy == 0 && gz == 0) return 0;
    if (!on_device) {
        if (real_size == 8) fold_host((double *)adj, nfields, nx, ny, nz, gy, gz);
        else fold_host((float *)adj, nfields, nx, ny, nz, gy, gz);
        return 0;
    }
    const int NY = ny + 2 * gy;
    const int rows = 2 * gz * NY + nz * 2 * gy;
    const int tb = nx >= 128 ? 128 : 32;
    dim3 grid((nx + tb - 1) / tb, rows, nfields);
    if (rows > 65535 || nfields > 65535) return -1;
    if (real_size == 8) k_fold_ghosts<double><<<grid, tb, 0, (cudaStream_t)stream>>>((double *)adj, nx, ny, nz, gy, gz);
    else k_fold_ghosts<float><<<grid, tb, 0, (cudaStream_t)stream>>>((float *)adj, nx, ny, nz, gy, gz);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// dst += src over n values (adjoint buffers of the intermediate stages of a multi-stage action are merged with this)
namespace {
template <class T>
__global__ void k_add_into(T *dst, const T *src, long long n) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const T v = src[i]; if (v != (T)0) dst[i] += v; }
}
}  // namespace
extern "C" int tclb_adjoint_add_into(void *dst, const void *src, long long n, int real_size, int on_device, void *stream) {
    if (!on_device) {
        if (real_size == 8) { double *d = (double *)dst; const double *s = (const double *)src; for (long long i = 0; i < n; i++) d[i] += s[i]; }
        else { float *d = (float *)dst; const float *s = (const float *)src; for (long long i = 0; i < n; i++) d[i] += s[i]; }
        return 0;
    }
    const int tb = 256;
    const long long nb = (n + tb - 1) / tb;
    if (real_size == 8) k_add_into<double><<<(unsigned)nb, tb, 0, (cudaStream_t)stream>>>((double *)dst, (const double *)src, n);
    else k_add_into<float><<<(unsigned)nb, tb, 0, (cudaStream_t)stream>>>((float *)dst, (const float *)src, n);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// Ghost fold between two subdomains: the ghost cells of `src` on its (dy,dz) side are images of interior cells of the
// neighbour `dst` on that side; their adjoint values are added there and cleared.  (dy,dz) = (0,0) is not a ghost side.
namespace {
template <class T>
__host__ __device__ inline void fold_pair_cell(T *dst, int dny, int dnz, T *src, int sny, int snz, int nx, int gy, int gz, int dy, int dz,
                                               int f, int iy, int iz, int x) {
    // (iy, iz) enumerate the ghost region of src: iy in [0, dy ? gy : sny), iz in [0, dz ? gz : snz)
    const int sy = dy > 0 ? sny + iy : (dy < 0 ? -gy + iy : iy), sz = dz > 0 ? snz + iz : (dz < 0 ? -gz + iz : iz);
    const int ty = dy > 0 ? iy : (dy < 0 ? dny - gy + iy : iy), tz = dz > 0 ? iz : (dz < 0 ? dnz - gz + iz : iz);
    const int sNY = sny + 2 * gy, sNZ = snz + 2 * gz, dNY = dny + 2 * gy, dNZ = dnz + 2 * gz;
    T *ps = src + (size_t)f * ((size_t)nx * sNY * sNZ) + ((size_t)(sz + gz) * sNY + (sy + gy)) * nx + x;
    const T v = *ps;
    if (v == (T)0) return;
    T *pd = dst + (size_t)f * ((size_t)nx * dNY * dNZ) + ((size_t)(tz + gz) * dNY + (ty + gy)) * nx + x;
#ifdef __CUDA_ARCH__
    atomicAdd(pd, v);
#else
    *pd += v;
#endif
    *ps = (T)0;
}
template <class T>
__global__ void k_fold_pair(T *dst, int dny, int dnz, T *src, int sny, int snz, int nx, int gy, int gz, int dy, int dz, int ry) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= nx) return;
    const int iy = blockIdx.y % ry, iz = blockIdx.y / ry;
    fold_pair_cell(dst, dny, dnz, src, sny, snz, nx, gy, gz, dy, dz, (int)blockIdx.z, iy, iz, x);
}
}  // namespace
extern "C" int tclb_adjoint_fold_pair(void *dst, int dny, int dnz, void *src, int sny, int snz, int real_size, int on_device, int nfields,
                                      int nx, int gy, int gz, int dy, int dz, void *stream) {
    const int ry = dy ? gy : sny, rz = dz ? gz : snz;
    if (ry <= 0 || rz <= 0 || (dy == 0 && dz == 0)) return 0;
    if (!on_device) {
        for (int f = 0; f < nfields; f++)
            for (int iz = 0; iz < rz; iz++)
                for (int iy = 0; iy < ry; iy++)
                    for (int x = 0; x < nx; x++) {
                        if (real_size == 8) fold_pair_cell((double *)dst, dny, dnz, (double *)src, sny, snz, nx, gy, gz, dy, dz, f, iy, iz, x);
                        else fold_pair_cell((float *)dst, dny, dnz, (float *)src, sny, snz, nx, gy, gz, dy, dz, f, iy, iz, x);
                    }
        return 0;
    }
    const int tb = nx >= 128 ? 128 : 32;
    if ((long long)ry * rz > 65535 || nfields > 65535) return -1;
    dim3 grid((nx + tb - 1) / tb, ry * rz, nfields);
    if (real_size == 8) k_fold_pair<double><<<grid, tb, 0, (cudaStream_t)stream>>>((double *)dst, dny, dnz, (double *)src, sny, snz, nx, gy, gz, dy, dz, ry);
    else k_fold_pair<float><<<grid, tb, 0, (cudaStream_t)stream>>>((float *)dst, dny, dnz, (float *)src, sny, snz, nx, gy, gz, dy, dz, ry);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
