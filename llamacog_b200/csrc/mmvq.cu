// mmvq.cu -- host side of the decode GEMV: launch geometry + C-ABI entry points.
// Replaces reference ggml-cuda/mmvq.cu:491-561 (ggml_cuda_mul_mat_vec_q) and the launch plumbing
// of :226-489; see mmvq.cuh for the kernel.
#include "mmvq_launch.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cstdlib>

namespace b200 {

template <> int mmvq_launch_type<T_Q4_0>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q8_0>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q4_K>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q5_K>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q6_K>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);

static int num_sms() {
    static int n[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 16) dev = 0;
    if (!n[dev]) cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    return n[dev] > 0 ? n[dev] : 148;
}

static const size_t SMEM_MAX = 232448 - 1024;   // 227 KB opt-in limit minus static/alignment slack

// Largest number of activation columns one launch can take for this K (register budget:
// CPT*M <= 8 unit-columns per thread, team <= 15 warps).
static int max_cols_per_launch(int K) {
    const int U = K / 32;
    for (int m = 8; m >= 1; --m) {
        for (int cpt = 1; cpt * m <= 8; cpt <<= 1) {
            if ((U + cpt - 1) / cpt <= 32 * 15) return m;
        }
    }
    return 0;
}

// Choose geometry.  Returns 0 on success.
static int configure(int type, int N, int K, int M, const b200_mmvq_args & a, mmvq_params & p, mmvq_launch_cfg & c) {
    const int U = K / 32;
    const int rb = K / block_elems(type) * block_bytes(type);
    // units per thread: smallest CPT that brings a team to <= 7 warps, within the register budget
    int cpt = 1;
    while (cpt * 2 * M <= 8 && (U + cpt - 1) / cpt > 224) cpt <<= 1;
    const int team_threads = (U + cpt - 1) / cpt;
    const int wpt = (team_threads + 31) / 32;
    if (wpt > 15) return -2;
    int nt = std::max(1, std::min(4, 16 / wpt));
    const int sms = num_sms();
    // CTAs: one per SM at most, and at least ~one row per team
    int gx = std::min(sms, std::max(1, (N + nt - 1) / nt));
    const int rows_per_cta = (N + gx - 1) / gx;
    nt = std::max(1, std::min(nt, rows_per_cta));

    const bool w_aligned = (((uintptr_t)a.W | (uintptr_t)a.nb1 | (uintptr_t)a.w_ch_stride) & 15) == 0;
    const bool contig = w_aligned && a.nb1 == rb && (rb % 16) == 0;
    const int pitch = contig ? rb : ((rb + 16 + 15) & ~15);
    // rows per stage: aim at ~32 KB per bulk transaction group, a multiple of NT, no more than the CTA has
    int x = std::max(1, (32 * 1024) / (nt * pitch));
    x = std::min(x, std::max(1, (rows_per_cta + nt - 1) / nt));
    int R = nt * x, stage_bytes = 0, nstages = 0;
    for (;;) {
        stage_bytes = ((R * pitch + 15) & ~15) + 16;
        const size_t per_stage = (size_t)stage_bytes + (size_t)wpt * nt * (R / nt) * M * 4 + 16;
        nstages = (int)std::min<size_t>(8, SMEM_MAX / per_stage);
        if (nstages >= 2 || R == 1) break;
        if (x > 1) { x = (x + 1) / 2; R = nt * x; }
        else if (nt > 1) { nt = nt / 2; R = nt; }
        else break;
    }
    if (nstages < 1) return -3;    // a single row does not fit in shared memory
    const int groups_per_cta = (rows_per_cta + R - 1) / R;
    nstages = std::max(1, std::min(nstages, groups_per_cta));

    p.R = R; p.nstages = nstages; p.pitch = pitch; p.stage_bytes = stage_bytes; p.contig = contig ? 1 : 0;
    p.WPT = wpt; p.NT = nt;
    c.M = M; c.CPT = cpt; c.A16 = (contig || w_aligned) ? 1 : 0;
    c.grid = dim3(gx, std::max(1, a.nchannels), 1);
    c.block = dim3(32 * (1 + wpt * nt), 1, 1);
    c.smem = (size_t)nstages * stage_bytes + (((size_t)nstages * wpt * nt * (R / nt) * M + 1) & ~(size_t)1) * 4 + (size_t)nstages * 16;
    return 0;
}

static int mmvq_dispatch(int type, const mmvq_params & p, const mmvq_launch_cfg & c, cudaStream_t s) {
    switch (type) {
        case T_Q4_0: return mmvq_launch_type<T_Q4_0>(p, c, s);
        case T_Q8_0: return mmvq_launch_type<T_Q8_0>(p, c, s);
        case T_Q4_K: return mmvq_launch_type<T_Q4_K>(p, c, s);
        case T_Q5_K: return mmvq_launch_type<T_Q5_K>(p, c, s);
        case T_Q6_K: return mmvq_launch_type<T_Q6_K>(p, c, s);
    }
    return -1;
}

int g_launches = 0;

}  // namespace b200

using namespace b200;

extern "C" int b200_mmvq_supported_type(int type) {
    return type == T_Q4_0 || type == T_Q8_0 || type == T_Q4_K || type == T_Q5_K || type == T_Q6_K;
}

extern "C" int b200_kernel_launch_count(void) { return g_launches; }

extern "C" int b200_mul_mat_vec_q_ex(const b200_mmvq_args * a, void * stream) {
    if (!a || !b200_mmvq_supported_type(a->type)) return -1;
    if (a->K <= 0 || a->K % block_elems(a->type) || a->N <= 0 || a->M <= 0) return -1;
    const int mmax = max_cols_per_launch(a->K);
    if (mmax <= 0) return -2;
    static const int pdl_env = [] { const char * e = getenv("B200_PDL"); return e ? atoi(e) : 0; }();
    for (int m0 = 0; m0 < a->M; m0 += mmax) {
        const int mc = std::min(mmax, a->M - m0);
        mmvq_params p = {};
        p.W = (const uint8_t *)a->W; p.nb1 = a->nb1;
        p.X = a->X + (int64_t)m0 * a->xs; p.xs = a->xs;
        p.dst = a->dst + (int64_t)m0 * a->ds; p.ds = a->ds;
        p.N = a->N; p.K = a->K;
        p.w_ch_stride = a->w_ch_stride; p.x_ch_stride = a->x_ch_stride; p.d_ch_stride = a->d_ch_stride;
        p.ids = a->ids; p.x_ch_mod = a->x_ch_mod > 0 ? a->x_ch_mod : 1 << 30; p.w_ch_div = a->w_ch_div > 0 ? a->w_ch_div : 1;
        p.vec_ok = ((((uintptr_t)p.X) | ((uintptr_t)a->xs * 4) | ((uintptr_t)a->x_ch_stride * 4)) & 15) == 0;
        p.dbg_sumi = a->dbg_sumi ? a->dbg_sumi + (int64_t)m0 * a->N * (a->K / 32) : nullptr;
        mmvq_launch_cfg c = {};
        int rc = configure(a->type, a->N, a->K, mc, *a, p, c);
        if (rc) return rc;
        c.pdl = a->pdl >= 0 ? a->pdl : pdl_env;
        rc = mmvq_dispatch(a->type, p, c, (cudaStream_t)stream);
        if (rc) return rc;
        ++g_launches;
    }
    return 0;
}

extern "C" int b200_mul_mat_vec_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs,
                                  float * dst, int64_t ds, int N, int K, int M, void * stream) {
    b200_mmvq_args a = {};
    a.type = type; a.W = W; a.nb1 = nb1; a.X = X; a.xs = xs; a.dst = dst; a.ds = ds; a.N = N; a.K = K; a.M = M;
    a.nchannels = 1; a.pdl = -1;
    return b200_mul_mat_vec_q_ex(&a, stream);
}
