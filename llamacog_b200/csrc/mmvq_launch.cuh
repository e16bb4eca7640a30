// mmvq_launch.cuh -- per-format instantiation table of mmvq_kernel (one translation unit per format
// so the formats compile in parallel).
#pragma once
#include "mmvq.cuh"

namespace b200 {

struct mmvq_launch_cfg {
    int M, A16;
    dim3 grid, block;
    size_t smem;
    int pdl;
};

template <int TYPE> int mmvq_launch_type(const mmvq_params & p, const mmvq_launch_cfg & c, cudaStream_t stream);

template <int TYPE, int M, bool A16, bool DBG>
static int mmvq_launch_one(const mmvq_params & p, const mmvq_launch_cfg & c, cudaStream_t stream) {
    auto kern = mmvq_kernel<TYPE, M, A16, DBG>;
    static int smem_set[16] = {0};   // per device
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 16 && !smem_set[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        smem_set[dev] = 1;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = c.grid; lc.blockDim = c.block; lc.dynamicSmemBytes = c.smem; lc.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = c.pdl ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, kern, p));
    return 0;
}

#define B200_MMVQ_CASE(MM) \
    if (c.M == MM) return p.dbg_sumi ? mmvq_launch_one<TYPE, MM, false, true>(p, c, stream) : \
        c.A16 ? mmvq_launch_one<TYPE, MM, A16_OK, false>(p, c, stream) : mmvq_launch_one<TYPE, MM, false, false>(p, c, stream);

// A16 (16-byte aligned items, LDS.128) exists for every format whose item size is a multiple of 16
#define B200_MMVQ_DEFINE(TYPE_) \
    template <> int mmvq_launch_type<TYPE_>(const mmvq_params & p, const mmvq_launch_cfg & c, cudaStream_t stream) { \
        constexpr int TYPE = TYPE_; \
        constexpr bool A16_OK = (TYPE_ != T_Q6_K && TYPE_ != T_Q2_K && TYPE_ != T_Q3_K && TYPE_ != T_IQ4_XS); \
        B200_MMVQ_CASE(1) B200_MMVQ_CASE(2) B200_MMVQ_CASE(3) B200_MMVQ_CASE(4) \
        B200_MMVQ_CASE(5) B200_MMVQ_CASE(6) B200_MMVQ_CASE(7) B200_MMVQ_CASE(8) \
        return -1; \
    }

}  // namespace b200
