// Shared device/host helpers for generated model libraries.
// Counterpart (new implementation) of the reference's cross-platform shim and reductions
// (src/cross.h, src/cross.hpp:45-233) reduced to the two targets this framework has:
// sm_100a device code and a plain C++ host restatement of the same node code (the oracle).
#pragma once
#include <cstdint>
#include <cmath>
#include <cstring>
#include <cstdio>
#include "tclb_model_abi.h"

#if defined(__CUDACC__)
#define TCLB_HD __host__ __device__ __forceinline__
#define TCLB_D __device__ __forceinline__
#else
#define TCLB_HD inline
#define TCLB_D inline
#endif
#define CudaDeviceFunction TCLB_HD

#ifndef __CUDACC__
struct float2 { float x, y; };
struct uchar4 { unsigned char x, y, z, w; };
#endif

template <class T> struct tclb_vec3 { T x, y, z; };

// value part of a scalar (the adjoint build adds an overload for dual numbers, tclb_dual.h)
TCLB_HD int tclb_value(int x) { return x; }
TCLB_HD float tclb_value(float x) { return x; }
TCLB_HD double tclb_value(double x) { return x; }

namespace tclbrt {

template <class T> TCLB_HD T tmax(T a, T b) { return a > b ? a : b; }
template <class T> TCLB_HD T tmin(T a, T b) { return a < b ? a : b; }

// wrapped x coordinate for an arbitrary (small) offset
TCLB_HD int wrapx(int x, int dx, int nx) {
    int v = x + dx;
    if (v < 0) v += nx;
    if (v >= nx) v -= nx;
    return v;
}

#if defined(__CUDACC__)
// ---- reductions: warp shuffle -> shared -> one atomic per block (reference K6) -------------
TCLB_D double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
TCLB_D double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
TCLB_D float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
TCLB_D float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
TCLB_D void atomic_max_double(double *addr, double v) {
    unsigned long long *p = reinterpret_cast<unsigned long long *>(addr);
    unsigned long long old = *p, assumed;
    do {
        assumed = old;
        if (__longlong_as_double((long long)assumed) >= v) break;
        old = atomicCAS(p, assumed, (unsigned long long)__double_as_longlong(v));
    } while (assumed != old);
}
// acquire / release accesses at system scope for the cross-GPU halo flags
TCLB_D unsigned ld_acquire_sys(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
TCLB_D void st_release_sys(unsigned *p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
TCLB_D unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#ifndef TCLB_HALO_TIMEOUT_NS
#define TCLB_HALO_TIMEOUT_NS 20000000000ull      // 20 s without a neighbour's halo flag = the peer is gone
#endif

// ---- bulk asynchronous copies into shared memory (TMA engine, SASS UBLKCP) + mbarrier completion ----
// Used by the persistent stage kernel: one elected thread streams the rows of the next tile into a
// shared-memory ring while the block computes the current one.  All sources are this GPU's own snapshot.
TCLB_D unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
TCLB_D void mbar_init(void *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
TCLB_D void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
TCLB_D void mbar_arrive_expect_tx(void *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
TCLB_D void mbar_wait(void *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "TCLB_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra TCLB_MBAR_DONE;\n"
        "bra TCLB_MBAR_WAIT;\n"
        "TCLB_MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// true in exactly one lane of a fully converged warp (elect.sync): unlike `lane == 0` the compiler then knows a single thread
// issues what follows, so a bulk copy is one instruction instead of a loop over the active lanes
TCLB_D bool elect_one() {
    unsigned pred = 0;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(pred));
    return pred != 0;
}
// global -> shared bulk copy (size and both addresses multiples of 16 bytes); completion is counted on `bar`
TCLB_D void bulk_g2s(void *dst, const void *src, unsigned bytes, void *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// one element (4 or 8 bytes) global -> shared without a register in between (LDGSTS)
template <int BYTES> TCLB_D void cp_async_elem(void *dst, const void *src) {
    if constexpr (BYTES == 4) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
    else if constexpr (BYTES == 8) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
// this thread's earlier cp.async copies arrive on `bar` when they have landed (the barrier's count includes them)
TCLB_D void cp_async_arrive_noinc(void *bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
#endif

// piecewise colour map used by the headless raster kernel (values in [0,1])
TCLB_HD void colormap(float v, float mask, unsigned char *rgba) {
    if (!(v == v)) { rgba[0] = 255; rgba[1] = 0; rgba[2] = 255; rgba[3] = 255; return; }
    v = v < 0.f ? 0.f : (v > 1.f ? 1.f : v);
    // blue -> cyan -> green -> yellow -> red
    float r, g, b;
    float t = v * 4.f;
    if (t < 1.f) { r = 0.f; g = t; b = 1.f; }
    else if (t < 2.f) { r = 0.f; g = 1.f; b = 2.f - t; }
    else if (t < 3.f) { r = t - 2.f; g = 1.f; b = 0.f; }
    else { r = 1.f; g = 4.f - t; b = 0.f; }
    rgba[0] = (unsigned char)(255.f * r * mask);
    rgba[1] = (unsigned char)(255.f * g * mask);
    rgba[2] = (unsigned char)(255.f * b * mask);
    rgba[3] = 255;
}

}  // namespace tclbrt
