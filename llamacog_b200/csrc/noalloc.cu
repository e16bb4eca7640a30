// noalloc.cu -- b200_mul_mat_q_noalloc: the tensor-core prefill mat-mul (mmq.cu) for callers that cannot allocate
// (code under CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// b200_mul_mat_q wants a caller-provided scratch for the bf16 activation rows [M][K].  Here that scratch is a file-scope
// __device__ array: every GPU that loads the library holds B200_NOALLOC_SCRATCH_BYTES of it until the process exits.
// In host code the array's NAME is only a host-side stand-in, not a device address, so its device address is fetched once
// per device with cudaGetSymbolAddress (that device current) and the pointer is what the kernels get.
//
// The entry point calls no cudaMalloc / cudaFree / allocating library routine: cudaGetDevice, cudaGetSymbolAddress (first
// call per device), cuTensorMapEncodeTiled (fills a host struct) and two kernel launches.
// Every call shares the one array per device, and it keeps the last call's contents.  Nothing in it has to start at zero:
// the conversion kernel writes all M*K values of a call before the mat-mul reads them (same stream), and token rows
// beyond M are never read from memory (the tensor map's extent is M; TMA fills out-of-range rows with zeros).  Two calls
// that overlap on the GPU (different streams, or graphs replayed concurrently) race on the array: serialise them.
#include "common.cuh"
#include "../../include/b200_ops.h"

#include <cstdio>
#include <mutex>

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int MAX_DEV = 64;
std::mutex g_mu;
void * g_ptr[MAX_DEV] = {nullptr};        // device address of g_noalloc_scratch on device i (each device has its own instance)

// set-up, once per device, with that device current
int scratch_of_current_device(void ** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= MAX_DEV) return -1;
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_ptr[dev]) {
        void * p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) {
            fprintf(stderr, "b200_mul_mat_q_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
            return (int)e;
        }
        if (!p || ((uintptr_t)p & 127)) return -1;
        g_ptr[dev] = p;
    }
    *out = g_ptr[dev];
    return 0;
}
}  // namespace

extern "C" int b200_mul_mat_q_noalloc(int type, const void * W, int64_t nb1, const float * X, int64_t xs, float * dst, int64_t ds,
                                      int N, int K, int M, void * stream) {
    if (!b200_mul_mat_q_supported(type, W, nb1, K) || N <= 0 || M <= 0) return -1;
    const int64_t need = b200_acts_bf16_bytes(K, M);
    if (need > (int64_t)B200_NOALLOC_SCRATCH_BYTES) {
        fprintf(stderr, "b200_mul_mat_q_noalloc: K = %d x M = %d needs %lld bytes of activation scratch, the static array holds %lld "
                        "(B200_NOALLOC_SCRATCH_BYTES): split the batch or use b200_mul_mat_q with your own scratch\n",
                K, M, (long long)need, (long long)B200_NOALLOC_SCRATCH_BYTES);
        return B200_ERR_NOALLOC_TOO_LARGE;
    }
    void * xb = nullptr;
    const int rc0 = scratch_of_current_device(&xb);
    if (rc0) return rc0;
    const int rc = b200_convert_acts_bf16(X, xs, xb, K, M, stream);
    if (rc) return rc;
    return b200_mul_mat_q_bf16(type, W, nb1, xb, dst, ds, N, K, M, stream);
}
