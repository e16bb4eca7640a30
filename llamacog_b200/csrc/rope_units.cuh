// rope_units.cuh -- the rope arithmetic, shared by the stand-alone kernels (glue.cu) and the q/k/v GEMV's fused epilogue
// (mmvq_r.cuh) so that both give the same bits.  Reference: ggml-cpu/ops.cpp:5080-5118, 5178-5360 (modes normal and neox).
#pragma once
#include "common.cuh"
#include <cuda_bf16.h>
#include <cmath>

namespace b200 {

struct rope_dev { int n_dims, neox; float freq_scale, ext_factor, attn_factor, theta_scale, corr0, corr1; };
// host: the device form of the op parameters (YaRN correction dims as ggml_rope_yarn_corr_dims, reference ggml.c)
template <typename PRM> static inline rope_dev rope_dev_from(const PRM & prm) {
    rope_dev p;
    p.n_dims = prm.n_dims; p.neox = prm.mode == 2; p.freq_scale = prm.freq_scale; p.ext_factor = prm.ext_factor;
    p.attn_factor = prm.attn_factor; p.theta_scale = powf(prm.freq_base, -2.0f / prm.n_dims);
    auto corr_dim = [&](float n_rot) { return prm.n_dims * logf(prm.n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(prm.freq_base)); };
    p.corr0 = fmaxf(0.0f, floorf(corr_dim(prm.beta_fast)));
    p.corr1 = fminf((float)prm.n_dims - 1, ceilf(corr_dim(prm.beta_slow)));
    return p;
}

// rotation of pair ip of a row at position p: returns cos/sin (already scaled)
__device__ __forceinline__ void rope_cs(const rope_dev & p, int64_t ip, int32_t posv, const float * __restrict__ ff, float & c, float & sn) {
    const float theta_extrap = (float)posv * powf(p.theta_scale, (float)ip) / (ff ? ff[ip] : 1.0f);
    float theta = p.freq_scale * theta_extrap, mscale = p.attn_factor;
    if (p.ext_factor != 0.0f) {
        const float y = ((float)ip - p.corr0) / fmaxf(0.001f, p.corr1 - p.corr0);
        const float ramp = (1.0f - fminf(1.0f, fmaxf(0.0f, y))) * p.ext_factor;
        theta = theta * (1.0f - ramp) + theta_extrap * ramp;
        mscale *= 1.0f + 0.1f * logf(1.0f / p.freq_scale);
    }
    c = cosf(theta) * mscale; sn = sinf(theta) * mscale;
}
// the rotation itself; explicit roundings so that no context can contract it differently
__device__ __forceinline__ void rope_pair(float x0, float x1, float c, float sn, float & ya, float & yb) {
    ya = __fsub_rn(__fmul_rn(x0, c), __fmul_rn(x1, sn));
    yb = __fadd_rn(__fmul_rn(x0, sn), __fmul_rn(x1, c));
}
// f32 -> element of a KV-cache cell (type id: T_F16, T_BF16, else f32)
__device__ __forceinline__ void store_as(char * p, int type, float v) {
    if (type == T_F16) *(__half *)p = __float2half_rn(v);
    else if (type == T_BF16) *(__nv_bfloat16 *)p = __float2bfloat16_rn(v);
    else *(float *)p = v;
}

}  // namespace b200
