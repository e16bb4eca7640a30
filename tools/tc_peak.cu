// tc_peak.cu -- empirical tensor-pipe peaks of tcgen05.mma on this B200 for kind::f16 (bf16), kind::i8 (s8 x s8 -> s32) and
// kind::f8f6f4 (e4m3): what the prefill MMQ's achieved TFLOP/s is a fraction of ("measure an empirical
// tcgen05 kind::i8 peak on the box first"; MEASURED_PEAKS.json only holds the cuBLAS bf16 figure).
// One CTA per SM; operands (random bytes) sit in shared memory in the canonical K-major 128-byte-swizzle layout, one elected thread
// issues M=128 x N=256 MMAs back to back into two TMEM accumulators; no loads in the loop: this is the issue-rate ceiling of the
// pipe, the number an SS-mode kernel could reach with perfect feeding.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -o tools/tc_peak.bin tools/tc_peak.cu
#include "../llamacog_b200/csrc/tc.cuh"

#include <cstdio>
#include <cstdlib>
#include <vector>

using namespace b200;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

template <int KIND>      // 0 f16 (bf16 operands), 1 i8, 2 f8f6f4 (e4m3)
__device__ __forceinline__ void mma(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    if constexpr (KIND == 0)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" :: "r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else if constexpr (KIND == 1)
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" :: "r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
    else
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" :: "r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

template <int KIND>
__global__ void __launch_bounds__(128, 1) peak_kernel(const uint4 * __restrict__ rnd, int iters, int n) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint32_t tmem_base;
    __shared__ __align__(8) uint64_t bar;
    uint8_t * sa = smem;                 // 128 rows x 128 bytes (one swizzle atom along K: 4 k-steps of 32 bytes)
    uint8_t * sb = smem + 128 * 128;     // 256 rows x 128 bytes
    for (int i = threadIdx.x; i < (128 + 256) * 128 / 16; i += blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = rnd[(i + 17 * blockIdx.x) & 4095];
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x < 32) { tmem_alloc(smem_u32(&tmem_base), 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t d0 = tmem_base;
    // instruction descriptor: D format (1 f32, 2 s32), A / B formats, N >> 3 at bit 17, M >> 4 at bit 24; K-major operands
    const uint32_t cfmt = KIND == 1 ? 2u : 1u, abfmt = KIND == 0 ? 1u : KIND == 1 ? 1u : 0u;
    const uint32_t idesc = (cfmt << 4) | (abfmt << 7) | (abfmt << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    if (threadIdx.x < 32 && elect_one()) {
        const uint64_t ad = umma_desc_sw128(smem_u32(sa)), bd = umma_desc_sw128(smem_u32(sb));
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) mma<KIND>(d0 + ((it & 1) ? 256u : 0u), ad + 2 * ks, bd + 2 * ks, idesc, (it > 1 || ks > 0) ? 1u : 0u);
        }
        tc_commit(smem_u32(&bar));
    }
    __syncwarp();
    mbar_wait(smem_u32(&bar), 0);
    tc_fence_after();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem_base, 512);
}

template <int KIND>
double run(const char * name, int kelems, const uint4 * rnd, int sms, int n) {
    const int iters = 8192;
    const size_t smem = (128 + 256) * 128 + 1024;
    CK(cudaFuncSetAttribute(peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
        CK(cudaEventRecord(e0));
        peak_kernel<KIND><<<sms, 128, smem>>>(rnd, iters, n);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        const double ops = 2.0 * 128 * n * kelems * 4.0 * iters * sms;
        const double t = ops / (ms * 1e-3) / 1e12;
        if (rep > 0 && t > best) best = t;
        if (rep == 4) printf("%-28s M=128 N=%d K=%d per instruction: %8.1f T%s/s (best of 4 timed launches, %.3f ms each, %d CTAs)\n", name, n, kelems, best, KIND == 1 ? "OP" : "FLOP", ms, sms);
    }
    return best;
}

int main() {
    int sms = 0; CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    std::vector<uint32_t> h(4096 * 4);
    uint32_t x = 12345u;
    for (auto & v : h) { x = x * 1664525u + 1013904223u; v = (x >> 3) & 0x3F3F3F3Fu; }      // small magnitudes: finite in every format
    uint4 * rnd; CK(cudaMalloc(&rnd, h.size() * 4)); CK(cudaMemcpy(rnd, h.data(), h.size() * 4, cudaMemcpyHostToDevice));
    const double f16 = run<0>("kind::f16 (bf16 x bf16 -> f32)", 16, rnd, sms, 256);
    const double i8 = run<1>("kind::i8 (s8 x s8 -> s32)", 32, rnd, sms, 256);
    const double f8 = run<2>("kind::f8f6f4 (e4m3 -> f32)", 32, rnd, sms, 256);
    printf("{\"tc_peak\": {\"bf16_tflops\": %.1f, \"int8_tops\": %.1f, \"fp8_tflops\": %.1f, \"sms\": %d}}\n", f16, i8, f8, sms);
    return 0;
}
