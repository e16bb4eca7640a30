// mmvq_r.cu -- host side of the register-resident-activation decode GEMV (mmvq_r.cuh): eligibility, launch geometry.
// Called first by b200_mul_mat_vec_q_ex / b200_mul_mat_vec_q_group (mmvq.cu); anything it declines goes to the general
// kernel of mmvq.cuh.
#include "mmvq_r.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cstdlib>

namespace b200 {

B200_MMVQR_DEFINE(T_Q4_K)
B200_MMVQR_DEFINE(T_Q6_K)

static int env_i(const char * name, int dflt) { const char * e = getenv(name); return e ? atoi(e) : dflt; }

static int sm_count() {
    static int n[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 16) dev = 0;
    if (!n[dev]) cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    return n[dev] > 0 ? n[dev] : 148;
}

// returns MMVQR_SKIP when the call is not eligible (the caller falls through to the general kernel), 0 when launched, else an error
int mmvqr_try(const b200_mmvq_args * a, int ng, const void * const * W, float * const * dst, const int * N, cudaStream_t stream) {
    static const int on = env_i("B200_MMVQR", 1), cps = std::max(1, std::min(2, env_i("B200_MMVQR_CPS", 2))),
                     smem_kb = env_i("B200_MMVQR_SMEM_KB", 110), stage_kb = env_i("B200_MMVQR_STAGE_KB", 36),
                     max_stages = std::max(1, std::min(8, env_i("B200_MMVQR_NS", 8))), eU = env_i("B200_MMVQR_U", 0);
    if (!on || a->M != 1 || !a->Xq || a->ids || std::max(1, a->ny) != 1 || std::max(1, a->nz) != 1) return MMVQR_SKIP;
    if (a->type != T_Q4_K && a->type != T_Q6_K) return MMVQR_SKIP;
    if (a->dbg_sumi && ng != 1) return MMVQR_SKIP;
    const int K = a->K;
    if (K <= 0 || K % 256) return MMVQR_SKIP;
    const int bpr = K / 256, hpr = 2 * bpr;
    if (hpr > 256) return MMVQR_SKIP;
    int tpr = 1, lg = 0;
    while (tpr < hpr) { tpr <<= 1; ++lg; }
    if (hpr * 10 < tpr * 6) return MMVQR_SKIP;                       // more than 40 % of the lanes would idle
    const int rb = bpr * block_bytes(a->type);
    if (a->nb1 != rb || rb % 16 || ((uintptr_t)a->Xq & 15)) return MMVQR_SKIP;
    int ntot = 0;
    for (int g = 0; g < ng; ++g) {
        if (((uintptr_t)W[g] & 15) || N[g] <= 0) return MMVQR_SKIP;
        ntot += N[g];
    }
    const int rpp = 256 / tpr, nw = tpr > 32 ? tpr / 32 : 1;
    const int gx_max = std::max(ng, std::min(sm_count() * cps, (ntot + rpp - 1) / rpp));

    mmvqr_params p = {};
    p.ngroup = ng;
    int used = 0, rows_per_cta = 1;
    for (int g = 0; g < ng; ++g) {
        int share = g + 1 == ng ? std::max(1, gx_max - used) : std::max(1, (int)((int64_t)gx_max * N[g] / ntot));
        share = std::max(1, std::min(share, (N[g] + rpp - 1) / rpp));
        p.cta0[g] = used; used += share;
        p.Wg[g] = (const uint8_t *)W[g]; p.Dg[g] = dst[g];
        p.rows_qg[g] = N[g] / share; p.rows_remg[g] = N[g] % share;
        rows_per_cta = std::max(rows_per_cta, (N[g] + share - 1) / share);
    }
    p.cta0[ng] = used;

    // rows per stage: about stage_kb of weights, at least four stages per CTA when it has the rows for them
    auto round_up = [](int x, int m) { return (x + m - 1) / m * m; };
    // U passes in flight per thread.  Measured (profiles/r01_mmvqr_sweep.log): one pass with 16-row stages beats two passes
    // (register pressure, spills for Q6_K); B200_MMVQR_U=2 keeps the variant reachable for tuning runs.
    const int U = eU > 0 ? std::min(2, eU) : 1;
    const int rq = rpp * U;                                  // rows per stage come in multiples of a full U-pass
    int R = std::max(rq, (stage_kb * 1024 / rb) / rq * rq);
    R = std::min(R, std::max(rq, round_up((rows_per_cta + 3) / 4, rq)));
    R = std::min(R, round_up(rows_per_cta, rpp));
    const int stage_bytes = round_up(R * rb + 16, 128);     // slack: the 2-byte aligned loads of Q6_K over-read 4 bytes
    const size_t budget = (size_t)smem_kb * 1024;
    const size_t per_stage = (size_t)stage_bytes + (size_t)R * nw * 4 + 16;
    const size_t act_bytes = (size_t)bpr * ACT_ITEM;
    if (budget < act_bytes + 128 + per_stage) return MMVQR_SKIP;
    int nstages = (int)std::min<size_t>(max_stages, (budget - act_bytes - 128) / per_stage);
    nstages = std::min(nstages, (rows_per_cta + R - 1) / R);
    if (nstages < 1) return MMVQR_SKIP;

    p.Xq = (const uint8_t *)a->Xq; p.K = K; p.BPR = bpr; p.TPR = tpr; p.lgTPR = lg; p.rb = rb;
    p.R = R; p.nstages = nstages; p.stage_bytes = stage_bytes;
    p.dbg_sumi = a->dbg_sumi;
    mmvqr_launch_cfg c;
    c.grid = dim3(used, 1, 1);
    c.smem = (size_t)nstages * stage_bytes + act_bytes + (((size_t)nstages * R * nw + 1) & ~(size_t)1) * 4 + (size_t)nstages * 16 + 16;
    c.U = U;
    c.pdl = a->pdl >= 0 ? a->pdl : pdl_default();
    const int rc = a->type == T_Q4_K ? mmvqr_launch_type<T_Q4_K>(p, c, stream) : mmvqr_launch_type<T_Q6_K>(p, c, stream);
    if (rc) return rc;
    ++g_launches;
    return 0;
}

}  // namespace b200
