// Tiny kernels behind the CUDA-graph iteration loop (Domain::iterate_graph): the per-launch iteration number and halo
// epoch live in device memory so that an instantiated graph can be replayed without patching its nodes.
#include <cuda_runtime.h>

namespace {
__global__ void k_dyn_set(int *dyn, int iter, int epoch) { dyn[0] = iter; dyn[1] = epoch; }
__global__ void k_dyn_advance(int *dyn, int diter, int depoch) { dyn[0] += diter; dyn[1] += depoch; }
}  // namespace

extern "C" int tclb_dyn_set(int *dyn, int iter, unsigned epoch, void *stream) {
    k_dyn_set<<<1, 1, 0, (cudaStream_t)stream>>>(dyn, iter, (int)epoch);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
extern "C" int tclb_dyn_advance(int *dyn, int diter, unsigned depoch, void *stream) {
    k_dyn_advance<<<1, 1, 0, (cudaStream_t)stream>>>(dyn, diter, (int)depoch);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
