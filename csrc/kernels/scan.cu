// Device-side scan of a real array: number of NaNs, minimum and maximum of the finite values.
// Used by <Failcheck> (reference src/Handlers/cbFailcheck.cpp:45-93 copies every quantity to the host and loops
// there; SURVEY section 8 W8 asks for the scan on the device) and exposed through the C ABI for callers that
// hold device arrays (torch tensors, lattice fields).
//
// One block reduces one fixed chunk of SCAN_CHUNK elements into a partial (count, min, max) in global scratch
// memory; the block that finishes last (arrival counter) folds the partials into the 3-double result.
#include <cfloat>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include "tclb_b200.h"

namespace {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_CHUNK = 256 * 16;     // elements per block

struct Partial { double nan_count, vmin, vmax; };

__device__ __forceinline__ Partial combine(const Partial &a, const Partial &b) {
    Partial r;
    r.nan_count = a.nan_count + b.nan_count;
    r.vmin = fmin(a.vmin, b.vmin);
    r.vmax = fmax(a.vmax, b.vmax);
    return r;
}
__device__ __forceinline__ Partial warp_fold(Partial p) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        Partial q;
        q.nan_count = __shfl_xor_sync(0xffffffffu, p.nan_count, o);
        q.vmin = __shfl_xor_sync(0xffffffffu, p.vmin, o);
        q.vmax = __shfl_xor_sync(0xffffffffu, p.vmax, o);
        p = combine(p, q);
    }
    return p;
}

// scratch layout: Partial[n_blocks] followed (at `counter`) by one unsigned arrival counter, which must be 0 on entry
// and is left at 0 on exit
template <class T>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan(const T *__restrict__ data, long long n, Partial *partials, unsigned *counter,
                                                        double *result3) {
    __shared__ Partial sh[SCAN_THREADS / 32];
    __shared__ bool is_last;
    const long long base = (long long)blockIdx.x * SCAN_CHUNK;
    Partial p{0.0, DBL_MAX, -DBL_MAX};
#pragma unroll 4
    for (int k = threadIdx.x; k < SCAN_CHUNK; k += SCAN_THREADS) {
        const long long i = base + k;
        if (i < n) {
            const double v = (double)data[i];
            if (v != v) p.nan_count += 1.0;
            else { p.vmin = fmin(p.vmin, v); p.vmax = fmax(p.vmax, v); }
        }
    }
    p = warp_fold(p);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = p;
    __syncthreads();
    if (threadIdx.x == 0) {
        Partial t = sh[0];
        for (int w = 1; w < SCAN_THREADS / 32; w++) t = combine(t, sh[w]);
        partials[blockIdx.x] = t;
        __threadfence();
        is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    Partial t{0.0, DBL_MAX, -DBL_MAX};
    for (unsigned b = threadIdx.x; b < gridDim.x; b += SCAN_THREADS) {
        const volatile Partial *q = partials + b;
        Partial r{q->nan_count, q->vmin, q->vmax};
        t = combine(t, r);
    }
    t = warp_fold(t);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        Partial r = sh[0];
        for (int w = 1; w < SCAN_THREADS / 32; w++) r = combine(r, sh[w]);
        result3[0] = r.nan_count; result3[1] = r.vmin; result3[2] = r.vmax;
        *counter = 0u;      // re-armed for the next call
    }
}

thread_local std::string g_scan_err;

int launch_scan(const void *dev, long long n, int real_size, cudaStream_t st, void *scratch, size_t n_blocks, double *result_dev) {
    Partial *partials = (Partial *)scratch;
    unsigned *counter = (unsigned *)((char *)scratch + n_blocks * sizeof(Partial));
    if (real_size == 4) k_scan<float><<<(unsigned)n_blocks, SCAN_THREADS, 0, st>>>((const float *)dev, n, partials, counter, result_dev);
    else if (real_size == 8) k_scan<double><<<(unsigned)n_blocks, SCAN_THREADS, 0, st>>>((const double *)dev, n, partials, counter, result_dev);
    else { g_scan_err = "tclb_scan_array: real_size must be 4 or 8"; return -1; }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { g_scan_err = std::string("tclb_scan_array: launch failed: ") + cudaGetErrorString(e); return -1; }
    return 0;
}

// scratch owned by the allocating entry point: one growing buffer per device
struct DevScratch { void *buf = nullptr; size_t bytes = 0; };
DevScratch g_scratch[64];
std::mutex g_mu;

}  // namespace

extern "C" const char *tclb_scan_last_error(void) { return g_scan_err.c_str(); }

extern "C" int tclb_scan_array(const void *dev_ptr, long long n, int real_size, void *stream, double *result3_dev) {
    if (n <= 0 || !dev_ptr || !result3_dev) { g_scan_err = "tclb_scan_array: empty input or missing output"; return -1; }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) { g_scan_err = "tclb_scan_array: no current CUDA device"; return -1; }
    const size_t n_blocks = (size_t)((n + SCAN_CHUNK - 1) / SCAN_CHUNK);
    const size_t need = n_blocks * sizeof(Partial) + 256;
    void *scratch;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        DevScratch &s = g_scratch[dev];
        if (need > s.bytes) {
            if (s.buf) { cudaDeviceSynchronize(); cudaFree(s.buf); s.buf = nullptr; s.bytes = 0; }
            size_t bytes = need * 2;
            if ((e = cudaMalloc(&s.buf, bytes)) != cudaSuccess) { g_scan_err = std::string("tclb_scan_array: cudaMalloc: ") + cudaGetErrorString(e); return -1; }
            s.bytes = bytes;
        }
        // the counter sits right behind the partials of THIS call, so it is cleared per call
        scratch = s.buf;
        if ((e = cudaMemsetAsync((char *)scratch + n_blocks * sizeof(Partial), 0, sizeof(unsigned), (cudaStream_t)stream)) != cudaSuccess) {
            g_scan_err = std::string("tclb_scan_array: memset: ") + cudaGetErrorString(e); return -1;
        }
    }
    return launch_scan(dev_ptr, n, real_size, (cudaStream_t)stream, scratch, n_blocks, result3_dev);
}

// ---------------------------------------------------------------------------------------------------------------
// Scan entry point that allocates nothing (declared in include/tclb_b200.h; no counterpart in the reference).
//
// Same k_scan kernel, but the partials and the arrival counter live in a fixed array that is part of this library's
// device image, so nothing is obtained from the CUDA allocator or any other allocating routine, on the first call or
// later.  The array is sized for TCLB_SCAN_STATIC_MAX_ELEMS elements; longer inputs are refused.
//
// Layout of the array: bytes [0,256) hold the arrival counter (one unsigned), the partials start at byte 256.
//   * the counter must be 0 when the kernel starts: every call clears it with a memset ordered on the caller's
//     stream ahead of the launch, so a call that was cut short earlier cannot poison the next one;
//   * the partials keep whatever the previous call left there; each block writes its own slot before the last block
//     reads slots [0, n_blocks), so stale slots are never read.
// There is ONE array per device and every call on that device uses it: two calls whose work can overlap on the GPU
// (different streams without an ordering between them, or two host threads) race on it and return garbage.
namespace {

constexpr long long SCAN_STATIC_MAX_BLOCKS = 16384;
constexpr long long SCAN_STATIC_MAX_ELEMS = SCAN_STATIC_MAX_BLOCKS * SCAN_CHUNK;       // 67,108,864 elements
constexpr size_t SCAN_STATIC_HEADER = 256;
constexpr size_t SCAN_STATIC_BYTES = SCAN_STATIC_HEADER + (size_t)SCAN_STATIC_MAX_BLOCKS * sizeof(Partial);   // 393,472 B

static __device__ __align__(256) unsigned char g_scan_static[SCAN_STATIC_BYTES];

// address of g_scan_static on each device, looked up once per device (first call there) with that device current
void *g_scan_static_ptr[64] = {};
std::mutex g_scan_static_mu;

int scan_static_fail(char *err, int err_cap, int code, const char *fmt, long long a = 0, long long b = 0) {
    if (err && err_cap > 0) snprintf(err, (size_t)err_cap, fmt, a, b);
    return code;
}

}  // namespace

static_assert(SCAN_STATIC_MAX_ELEMS == TCLB_SCAN_STATIC_MAX_ELEMS, "header and kernel disagree on the static scan limit");

extern "C" int tclb_scan_array_static(const void *dev_ptr, long long n, int real_size, void *stream, double *result3_dev,
                                      char *err, int err_cap) {
    if (err && err_cap > 0) err[0] = 0;
    if (!dev_ptr || !result3_dev || n <= 0)
        return scan_static_fail(err, err_cap, -1, "tclb_scan_array_static: empty input or missing output (n=%lld)", n);
    if (real_size != 4 && real_size != 8)
        return scan_static_fail(err, err_cap, -2, "tclb_scan_array_static: real_size must be 4 or 8, got %lld", real_size);
    if (n > SCAN_STATIC_MAX_ELEMS)
        return scan_static_fail(err, err_cap, -3, "tclb_scan_array_static: n=%lld exceeds the fixed scratch limit of %lld elements; "
                                                  "use tclb_scan_array for longer arrays", n, SCAN_STATIC_MAX_ELEMS);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64)
        return scan_static_fail(err, err_cap, -4, "tclb_scan_array_static: no usable current CUDA device (device %lld, CUDA error %lld)", dev, (long long)e);
    void *scratch;
    {
        std::lock_guard<std::mutex> lk(g_scan_static_mu);
        if (!g_scan_static_ptr[dev]) {
            // set-up, once per device, with that device current
            void *p = nullptr;
            e = cudaGetSymbolAddress((void **)&p, g_scan_static);
            if (e != cudaSuccess || !p)
                return scan_static_fail(err, err_cap, -5, "tclb_scan_array_static: cudaGetSymbolAddress failed on device %lld (CUDA error %lld)", dev, (long long)e);
            g_scan_static_ptr[dev] = p;
        }
        scratch = g_scan_static_ptr[dev];
    }
    cudaStream_t st = (cudaStream_t)stream;
    unsigned *counter = (unsigned *)scratch;
    Partial *partials = (Partial *)((char *)scratch + SCAN_STATIC_HEADER);
    const unsigned n_blocks = (unsigned)((n + SCAN_CHUNK - 1) / SCAN_CHUNK);
    if ((e = cudaMemsetAsync(counter, 0, sizeof(unsigned), st)) != cudaSuccess)
        return scan_static_fail(err, err_cap, -6, "tclb_scan_array_static: clearing the arrival counter failed (CUDA error %lld)", (long long)e);
    if (real_size == 4) k_scan<float><<<n_blocks, SCAN_THREADS, 0, st>>>((const float *)dev_ptr, n, partials, counter, result3_dev);
    else k_scan<double><<<n_blocks, SCAN_THREADS, 0, st>>>((const double *)dev_ptr, n, partials, counter, result3_dev);
    if ((e = cudaGetLastError()) != cudaSuccess)
        return scan_static_fail(err, err_cap, -7, "tclb_scan_array_static: kernel launch failed (CUDA error %lld)", (long long)e);
    return 0;
}
