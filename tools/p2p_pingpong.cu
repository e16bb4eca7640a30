// p2p_pingpong.cu -- latency of the flag primitives the row-split hand-over is built from, between two B200s over NVLink.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/p2p_pingpong tools/p2p_pingpong.cu ; run on a 2-GPU box.
// Measures (a) a one-thread ping-pong round trip with different store / poll flavours, (b) the cost of a system-scope fence
// after remote stores, (c) flag propagation when N CTAs poll the same local word, (d) the full folded pattern: remote stores from
// many CTAs -> per-CTA fence -> ticket -> flag.
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ uint32_t ld_acq_sys(const uint32_t * p) { uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ uint32_t ld_rlx_sys(const uint32_t * p) { uint32_t v; asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ uint32_t ld_vol(const uint32_t * p) { return *(volatile const uint32_t *)p; }
__device__ __forceinline__ void st_rel_sys(uint32_t * p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_rlx_sys(uint32_t * p, uint32_t v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }

// mode: poll flavour 0 acquire.sys, 1 relaxed.sys, 2 volatile ; store flavour (mode >> 4): 0 release.sys, 1 relaxed.sys, 2 fence.sys + volatile
// each side polls its LOCAL word and writes the REMOTE one
__global__ void pingpong(uint32_t * mine, uint32_t * theirs, int iters, int first, int mode, unsigned long long * out) {
    const int pm = mode & 15, sm = mode >> 4;
    unsigned long long t0 = gt();
    for (int i = 1; i <= iters; ++i) {
        if (first) {
            if (sm == 0) st_rel_sys(theirs, i); else if (sm == 1) st_rlx_sys(theirs, i); else { __threadfence_system(); *(volatile uint32_t *)theirs = i; }
        }
        for (;;) {
            const uint32_t v = pm == 0 ? ld_acq_sys(mine) : pm == 1 ? ld_rlx_sys(mine) : ld_vol(mine);
            if (v >= (uint32_t)i) break;
        }
        if (!first) {
            if (sm == 0) st_rel_sys(theirs, i); else if (sm == 1) st_rlx_sys(theirs, i); else { __threadfence_system(); *(volatile uint32_t *)theirs = i; }
        }
    }
    if (out) *out = gt() - t0;
}

// cost of fences after n remote 4-byte stores by one thread
__global__ void fence_cost(float * remote, int n, int iters, int scope, unsigned long long * out) {
    unsigned long long t0 = gt();
    for (int i = 0; i < iters; ++i) {
        for (int j = 0; j < n; ++j) remote[j * 32] = (float)i;
        if (scope == 0) __threadfence(); else if (scope == 1) __threadfence_system();
    }
    *out = gt() - t0;
}

// folded pattern, peer side: `ctas` CTAs spin on the local flag (value >= round), then each stores `rows` floats to remote memory, fences
// (scope per `fmode`: 0 = every lane fence.sys, 1 = lane 0 fence.sys, 2 = lane 0 fence.gpu and only the last CTA fence.sys), takes a ticket;
// the last CTA raises the remote done flag.  main side (single thread kernel): raises the flag, polls done; times the round trip.
__global__ void folded_peer(const uint32_t * flag, float * remote_dst, int rows, unsigned * ticket, uint32_t * remote_done, int rounds, int fmode, int pollmode) {
    const int lane = threadIdx.x;
    for (int r = 1; r <= rounds; ++r) {
        if (lane == 0) {
            int n = 0;
            for (;;) {
                const uint32_t v = pollmode == 0 ? ld_acq_sys(flag) : ld_rlx_sys(flag);
                if (v >= (uint32_t)r) break;
                if (++n > 64) __nanosleep(32);
            }
            if (pollmode != 0) __threadfence_system();
        }
        __syncwarp();
        for (int o = lane; o < rows; o += 32) remote_dst[(size_t)blockIdx.x * rows + o] = (float)r;
        if (fmode == 0) __threadfence_system();
        __syncwarp();
        if (lane == 0) {
            if (fmode <= 1) __threadfence_system(); else __threadfence();
            if (atomicInc(ticket, gridDim.x - 1) == gridDim.x - 1) {
                __threadfence_system();
                st_rel_sys(remote_done, (uint32_t)r);
            }
        }
        __syncwarp();
    }
}
__global__ void folded_main(uint32_t * remote_flag, const uint32_t * done, const float * dst, int n_dst, int rounds, unsigned long long * out, int * bad) {
    unsigned long long t0 = gt();
    int nbad = 0;
    for (int r = 1; r <= rounds; ++r) {
        st_rel_sys(remote_flag, (uint32_t)r);
        while (ld_acq_sys(done) < (uint32_t)r) {}
        // the peer's stores must be visible now: sample a few
        for (int j = 0; j < n_dst; j += n_dst / 8 + 1) if (*(volatile const float *)(dst + j) != (float)r) ++nbad;
    }
    *out = gt() - t0; *bad = nbad;
}

int main() {
    int n = 0; CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0));
    CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0));
    uint32_t * f0, * f1; unsigned long long * o0; float * d0; unsigned * tk1; int * bad0;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&f0, 256)); CK(cudaMemset(f0, 0, 256)); CK(cudaMalloc(&o0, 64)); CK(cudaMalloc(&d0, 4 << 20)); CK(cudaMalloc(&bad0, 64));
    CK(cudaSetDevice(1)); CK(cudaMalloc(&f1, 256)); CK(cudaMemset(f1, 0, 256)); CK(cudaMalloc(&tk1, 64)); CK(cudaMemset(tk1, 0, 64));
    cudaStream_t s0, s1;
    CK(cudaSetDevice(0)); CK(cudaStreamCreate(&s0));
    CK(cudaSetDevice(1)); CK(cudaStreamCreate(&s1));
    const int iters = 2000;
    const char * pn[3] = {"ld.acquire.sys", "ld.relaxed.sys", "ld.volatile"}, * sn[3] = {"st.release.sys", "st.relaxed.sys", "fence.sys+st.volatile"};
    for (int sm = 0; sm < 3; ++sm) for (int pm = 0; pm < 3; ++pm) {
        CK(cudaSetDevice(0)); CK(cudaMemset(f0, 0, 256)); CK(cudaSetDevice(1)); CK(cudaMemset(f1, 0, 256)); CK(cudaDeviceSynchronize()); CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize());
        CK(cudaSetDevice(1)); pingpong<<<1, 1, 0, s1>>>(f1, f0, iters, 0, pm | (sm << 4), nullptr);
        CK(cudaSetDevice(0)); pingpong<<<1, 1, 0, s0>>>(f0, f1, iters, 1, pm | (sm << 4), o0);
        CK(cudaDeviceSynchronize()); CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize());
        unsigned long long ns; CK(cudaMemcpy(&ns, o0, 8, cudaMemcpyDeviceToHost));
        printf("pingpong  %-22s / %-15s : %.2f us per round trip\n", sn[sm], pn[pm], ns / 1e3 / iters);
    }
    for (int scope = 0; scope < 3; ++scope) for (int nst : {1, 16}) {
        CK(cudaSetDevice(1));
        fence_cost<<<1, 1, 0, s1>>>((float *)d0, nst, 500, scope, (unsigned long long *)(f1 + 32));
        CK(cudaDeviceSynchronize());
        unsigned long long ns; CK(cudaMemcpy(&ns, f1 + 32, 8, cudaMemcpyDeviceToHost));
        printf("fence     %d remote store(s) + %-18s : %.2f us per iteration\n", nst, scope == 0 ? "__threadfence" : scope == 1 ? "__threadfence_system" : "no fence", ns / 1e3 / 500);
    }
    const int rounds = 300;
    for (int ctas : {1, 32, 148, 296}) for (int fmode = 0; fmode < 3; ++fmode) for (int pollmode = 0; pollmode < 2; ++pollmode) {
        CK(cudaSetDevice(0)); CK(cudaMemset(f0, 0, 256)); CK(cudaSetDevice(1)); CK(cudaMemset(f1, 0, 256)); CK(cudaMemset(tk1, 0, 64)); CK(cudaDeviceSynchronize());
        CK(cudaSetDevice(0)); CK(cudaDeviceSynchronize());
        const int rows = 16;
        CK(cudaSetDevice(1)); folded_peer<<<ctas, 32, 0, s1>>>(f1, d0, rows, tk1, f0, rounds, fmode, pollmode);
        CK(cudaSetDevice(0)); folded_main<<<1, 1, 0, s0>>>(f1, f0, d0, ctas * rows, rounds, o0, bad0);
        CK(cudaDeviceSynchronize()); CK(cudaSetDevice(1)); CK(cudaDeviceSynchronize());
        unsigned long long ns; int bad; CK(cudaMemcpy(&ns, o0, 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&bad, bad0, 4, cudaMemcpyDeviceToHost));
        printf("folded    %3d CTAs, fence mode %d (%s), poll %s : %.2f us per hand-over, stale reads %d\n", ctas, fmode,
               fmode == 0 ? "all lanes sys" : fmode == 1 ? "lane0 sys" : "lane0 gpu + last sys", pollmode ? "relaxed+fence" : "acquire", ns / 1e3 / rounds, bad);
    }
    return 0;
}
