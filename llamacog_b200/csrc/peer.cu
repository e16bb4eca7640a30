// peer.cu -- device-side synchronisation of the row-split matmul over NVLink peer memory.
//
// The reference forks and joins its per-device streams with cudaEventRecord / cudaStreamWaitEvent around every
// split MUL_MAT and moves the activations and the results with peer memcpys (ggml-cuda/ggml-cuda.cu:1605-1608,
// :1628-1630, :1701-1745).  A cross-device event wait costs ~10 us each way on this machine, which is several
// times the duration of a decode GEMV.  Here the devices never meet on the host's timeline: their streams run
// free and hand over through flags in each other's HBM.
//
//   main:  b200_peer_bcast   copies the prepared activations into every peer's buffer with peer stores, then
//                            (last block, after a system-scope fence) raises the flag in each peer's memory
//   peer:  b200_peer_wait    one thread spins on its local flag (ld.acquire.sys), lowers it again
//          ... the shard's GEMV / MMQ kernels; they store their slice of dst straight into main's buffer
//          b200_peer_signal  fence + st.release.sys of the peer's "done" flag in main's memory
//   main:  ... its own shard, then b200_peer_wait on the done flags of all peers
//
// The flags return to 0 after every hand-over, so a captured CUDA graph replays the same protocol.  Every spin is
// bounded by a wall-clock budget (B200_SPLIT_TIMEOUT_MS, default 30 s: tools that serialise kernels, time slicing or a long
// backlog on the other device must not trip it).  On expiry b200_peer_wait_ex sets the caller's error word and carries on --
// the plugin then fails the backend cleanly; the plain b200_peer_wait has nowhere to report to and traps.
//
// For decode the wait and the signal no longer are launches of their own: the mat-vec kernels take the flags as arguments
// (b200_mmvq_args.hs_*, common.cuh hs_params) -- wait in the prologue of every CTA, signal / join from the CTA that finishes last.
#include "common.cuh"
#include "../../include/b200_ops.h"

#include <cstdlib>

namespace b200 {

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t * p) {
    uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t * p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t;
}

struct bcast_args {
    const uint4 * src;
    uint4 * dst[B200_PEER_MAX];
    uint32_t * flag[B200_PEER_MAX];
    int n_peers;
    unsigned long long n16;             // 16-byte words
    unsigned * counter;                 // in main's memory; wraps back to 0 by itself
};

// grid (blocks, n_peers): block (x, p) copies its slice of src to peer p
__global__ void __launch_bounds__(512) peer_bcast_kernel(const bcast_args a) {
    const int p = blockIdx.y;
    uint4 * dst = a.dst[p];
    const unsigned long long step = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * step < a.n16; i += 4 * step) {                  // four loads in flight per thread
        const uint4 v0 = a.src[i], v1 = a.src[i + step], v2 = a.src[i + 2 * step], v3 = a.src[i + 3 * step];
        dst[i] = v0; dst[i + step] = v1; dst[i + 2 * step] = v2; dst[i + 3 * step] = v3;
    }
    for (; i < a.n16; i += step) dst[i] = a.src[i];
    __syncthreads();
    if (threadIdx.x == 0) {
        // gpu-scope fence per block, ONE system-scope fence by the block that draws the last ticket (it has observed all the
        // others; the fence is cumulative): 512 threads x 128 blocks fencing at system scope serialise for tens of microseconds
        __threadfence();
        const unsigned total = gridDim.x * gridDim.y;
        if (atomicInc(a.counter, total - 1) == total - 1) {        // every block's stores are ordered before its increment
            __threadfence_system();
            for (int q = 0; q < a.n_peers; ++q) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(a.flag[q]), "r"(1u) : "memory");
        }
    }
}

// thread t (if its bit is set in mask) waits for flags[t * stride] to be raised and lowers it
__global__ void peer_wait_kernel(uint32_t * flags, int stride, uint32_t mask, unsigned long long timeout_ns, uint32_t * error) {
    if (!((mask >> threadIdx.x) & 1u)) return;
    uint32_t * f = flags + (size_t)threadIdx.x * stride;
    const unsigned long long t0 = globaltimer_ns();
    while (ld_acquire_sys(f) == 0u) {
        __nanosleep(40);
        if (timeout_ns && globaltimer_ns() - t0 > timeout_ns) {
            if (!error) __trap();
            *(volatile uint32_t *)error = 1u;
            break;
        }
    }
    *f = 0u;
    __threadfence_system();
}

__global__ void peer_signal_kernel(uint32_t * remote_flag) {
    __threadfence_system();
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(remote_flag), "r"(1u) : "memory");
}

}  // namespace b200

using namespace b200;

extern "C" int b200_peer_bcast(const void * src, void * const * dsts, uint32_t * const * flags, int n_peers, int64_t bytes,
                               unsigned * counter, void * stream) {
    if (n_peers < 1 || n_peers > B200_PEER_MAX || bytes < 0 || (bytes & 15) || ((uintptr_t)src & 15)) return -1;      // bytes == 0: flags only
    bcast_args a = {};
    a.src = (const uint4 *)src; a.n_peers = n_peers; a.n16 = (unsigned long long)(bytes / 16); a.counter = counter;
    for (int p = 0; p < n_peers; ++p) {
        if ((uintptr_t)dsts[p] & 15) return -1;
        a.dst[p] = (uint4 *)dsts[p]; a.flag[p] = flags[p];
    }
    unsigned blocks = (unsigned)((a.n16 + 511) / 512);
    if (blocks > 128) blocks = 128;        // 1 MiB per pass over the grid; NVLink egress is the limit, not SMs
    if (blocks < 1) blocks = 1;
    peer_bcast_kernel<<<dim3(blocks, (unsigned)n_peers), 512, 0, (cudaStream_t)stream>>>(a);
    B200_CUDA_CHECK(cudaGetLastError());
    ++g_launches;
    return 0;
}

extern "C" unsigned b200_peer_timeout_ms(void) {
    static const unsigned v = [] { const char * e = getenv("B200_SPLIT_TIMEOUT_MS"); return e ? (unsigned)atol(e) : 30000u; }();
    return v;
}
extern "C" int b200_peer_wait_ex(uint32_t * flags, int stride, uint32_t mask, unsigned timeout_ms, uint32_t * error, void * stream) {
    if (mask == 0) return 0;
    peer_wait_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(flags, stride, mask, (unsigned long long)timeout_ms * 1000000ull, error);
    B200_CUDA_CHECK(cudaGetLastError());
    ++g_launches;
    return 0;
}
extern "C" int b200_peer_wait(uint32_t * flags, int stride, uint32_t mask, void * stream) {
    return b200_peer_wait_ex(flags, stride, mask, b200_peer_timeout_ms(), nullptr, stream);
}

extern "C" int b200_peer_signal(uint32_t * remote_flag, void * stream) {
    peer_signal_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(remote_flag);
    B200_CUDA_CHECK(cudaGetLastError());
    ++g_launches;
    return 0;
}
