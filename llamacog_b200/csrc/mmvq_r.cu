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
B200_MMVQR_DEFINE(T_Q8_0)
B200_MMVQR_DEFINE(T_Q4_0)
B200_MMVQR_DEFINE(T_Q5_K)
B200_MMVQR_DEFINE(T_Q4_1)
B200_MMVQR_DEFINE(T_Q5_0)
B200_MMVQR_DEFINE(T_Q5_1)
B200_MMVQR_DEFINE(T_IQ4_NL)
B200_MMVQR_DEFINE(T_Q2_K)
B200_MMVQR_DEFINE(T_Q3_K)
B200_MMVQR_DEFINE(T_IQ4_XS)

static int env_i(const char * name, int dflt) { const char * e = getenv(name); return e ? atoi(e) : dflt; }

static int sm_count() {
    static int n[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 16) dev = 0;
    if (!n[dev]) cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    return n[dev] > 0 ? n[dev] : 148;
}

static bool al16(const void * p) { return ((uintptr_t)p & 15) == 0; }
// the fused prologue / epilogue of this launch, if any, are ones the kernel can do
static bool fuse_ok(const b200_mmvq_args * a, int ng) {
    if (a->fuse_resid || a->fuse_sum) {
        if (!a->fuse_resid || !a->fuse_sum || ng != 1) return false;
    }
    if (a->rope) {
        const b200_gemv_rope * r = a->rope;
        if (a->fuse_resid || a->dbg_sumi || !r->pos || r->prm.mode != 0 || r->prm.n_dims % 2 || r->prm.n_dims <= 0) return false;
        for (int g = 0; g < ng; ++g) {
            const int m = r->mode[g];
            if (m < 0 || m > 3) return false;
            if (m == 0) continue;
            if ((m == 1 || m == 2) && (!r->out[g] || r->head_dim[g] <= 0 || r->head_dim[g] % 2 || r->prm.n_dims > r->head_dim[g])) return false;
            if ((m == 2 || m == 3) && (!(r->cell[g] || r->cell_slot[g]) || (r->cell_type[g] != T_F16 && r->cell_type[g] != T_F32 && r->cell_type[g] != T_BF16))) return false;
        }
    }
    if (a->norm_x) {
        if (!a->norm_w || !a->norm_y || !a->norm_out || a->K > 4096 || a->dbg_sumi) return false;
        if (!al16(a->norm_x) || !al16(a->norm_w) || !al16(a->norm_y) || !al16(a->norm_out)) return false;
    }
    if (a->sm_act) {      // silu(gate) * up in the epilogue of a gate / up launch
        if (a->sm_op != B200_UNARY_SILU || !a->sm_out || a->fuse_resid || a->rope || a->dbg_sumi || a->hs_wait || a->hs_counter) return false;
        if (a->ids ? (ng != 1 || !a->sm_up_W || !a->sm_up_dst || ((uintptr_t)a->sm_up_W & 15) || a->norm_x) : ng != 2) return false;
        if (!al16(a->sm_act) || !al16(a->sm_out) || !al16(a->sm_xq)) return false;
    }
    return true;
}

// returns MMVQR_SKIP when the call is not eligible (the caller falls through to the general kernel), 0 when launched, else an error.
// stream == (cudaStream_t)-1: eligibility check only (returns 0 when the call would be launched)
int mmvqr_try(const b200_mmvq_args * a, int ng, const void * const * W, float * const * dst, const int * N, cudaStream_t stream) {
    const bool probe = stream == (cudaStream_t)-1;
    // MUL_MAT_ID for one token: slot y of the router's choice becomes matrix y of a group launch; the kernel adds ids[y] * w_sy to
    // the base of the expert stack after its PDL wait (mmvq_r.cuh)
    const void * Wid[3]; float * Did[3]; int Nid[3];
    const bool by_id = a->ids != nullptr;
    if (by_id && a->sm_act && !fuse_ok(a, ng)) return MMVQR_SKIP;        // (asked with ng = 1 here: the slots become groups below)
    if (by_id) {
        const int ny = std::max(1, a->ny);
        if (ng != 1 || ny > 3 || std::max(1, a->nz) != 1 || !a->Xq || a->fuse_resid || a->fuse_sum || a->norm_x || a->rope || a->dbg_sumi ||
            a->hs_wait || a->hs_counter || std::max(1, a->w_div_y) != 1 || a->xq_sy || (a->w_sy & 15)) return MMVQR_SKIP;
        for (int y = 0; y < ny; ++y) { Wid[y] = W[0]; Did[y] = probe ? nullptr : dst[0] + (int64_t)y * a->d_sy; Nid[y] = N[0]; }
        W = Wid; dst = Did; N = Nid; ng = ny;
    }
    static const int on = env_i("B200_MMVQR", 1), cps = std::max(1, std::min(2, env_i("B200_MMVQR_CPS", 2))),
                     smem_kb = env_i("B200_MMVQR_SMEM_KB", 112), stage_kb = env_i("B200_MMVQR_STAGE_KB", 36),
                     max_stages = std::max(1, std::min(8, env_i("B200_MMVQR_NS", 8))), eU = env_i("B200_MMVQR_U", 0),
                     ramp = env_i("B200_MMVQR_RAMP", 0),          // short first stages: measured slower (per-stage hand-over cost), kept for tuning runs
                     small_mb = env_i("B200_MMVQR_SMALL_MB", 0), stagger = env_i("B200_MMVQR_STAGGER", 0);
    if (!(by_id && a->sm_act) && !fuse_ok(a, ng)) return MMVQR_SKIP;
    if (!on || a->M != 1 || (!a->Xq && !a->norm_x) || (!by_id && std::max(1, a->ny) != 1) || std::max(1, a->nz) != 1) return MMVQR_SKIP;
    static const int more_on = env_i("B200_MMVQR_MORE", 1);      // Q8_0 / Q4_0 / Q5_K (round 2); 0: only Q4_K and Q6_K take this kernel
    static const int wide_on = env_i("B200_MMVQR_WIDE", 1);      // the formats of the round-2 widening
    const bool more_t = a->type == T_Q8_0 || a->type == T_Q4_0 || a->type == T_Q5_K;
    const bool wide_t = a->type == T_Q4_1 || a->type == T_Q5_0 || a->type == T_Q5_1 || a->type == T_IQ4_NL || a->type == T_Q2_K || a->type == T_Q3_K || a->type == T_IQ4_XS;
    if (a->type != T_Q4_K && a->type != T_Q6_K && !(more_t && more_on) && !(wide_t && wide_on)) return MMVQR_SKIP;
    if (a->dbg_sumi && ng != 1) return MMVQR_SKIP;
    const int K = a->K;
    if (K <= 0 || K % 256) return MMVQR_SKIP;
    const int bpr = K / 256, hpr = 2 * bpr;
    if (hpr > 256) return MMVQR_SKIP;
    int tpr = 1, lg = 0;
    while (tpr < hpr) { tpr <<= 1; ++lg; }
    if (hpr * 10 < tpr * 6) return MMVQR_SKIP;                       // more than 40 % of the lanes would idle
    const int rb = K / block_elems(a->type) * block_bytes(a->type);     // (Q8_0: eight 34-byte blocks per 256-weight item)
    if (a->nb1 != rb || rb % 16 || ((uintptr_t)a->Xq & 15)) return MMVQR_SKIP;
    int ntot = 0;
    const bool pair = a->sm_act != nullptr;                           // gate / up pair mode: every CTA owns whole 32-row blocks of both matrices
    const int unit = pair ? 32 : a->rope ? 2 : 1;                     // rope epilogue: rows are dealt to the CTAs in pairs
    if (a->sm_act && ((!by_id && N[0] != N[1]) || N[0] % 32 || (by_id && a->sm_xq && N[0] % 256))) return MMVQR_SKIP;
    for (int g = 0; g < ng; ++g) {
        if (((uintptr_t)W[g] & 15) || N[g] <= 0 || N[g] % unit) return MMVQR_SKIP;
        if (a->rope && (a->rope->mode[g] == 1 || a->rope->mode[g] == 2) && N[g] % a->rope->head_dim[g]) return MMVQR_SKIP;
        ntot += N[g];
    }
    const int rpp = 256 / tpr, nw = tpr > 32 ? tpr / 32 : 1;
    // Tuning knob (B200_MMVQR_SMALL_MB, default 0 = off): a matrix below the threshold takes one CTA per SM only, leaving the
    // other register/thread slot of the SM to the mat-vec that follows under PDL.  No effect on tg128 (471-474 tok/s either way).
    const int cps_eff = (int64_t)ntot * rb <= (int64_t)small_mb << 20 ? 1 : cps;
    const int gx_max = std::max(ng, std::min(sm_count() * cps_eff, (ntot + rpp - 1) / rpp));

    mmvqr_params p = {};
    p.ngroup = pair && !by_id ? 1 : ng;
    int used = 0, rows_per_cta = 1;
    if (pair) {
        // the same number of blocks for every CTA where possible: CTAs share HBM evenly, so the launch ends with its largest CTA
        // (448 blocks of a 14336-row pair: 224 CTAs of two blocks, not 152 of two and 144 of one).  MoE: one such group per slot.
        const int ngr = by_id ? ng : 1;
        // Two geometries, measured A/B on tg128 (profiles/r02_ab_pair_geometry.log).  0: every CTA the same number of blocks (448 blocks of
        // an 8B gate/up pair: 224 CTAs of two) -- CTAs share HBM evenly, so a launch ends with its largest CTA; better for the dense pair
        // (8B: 1.685 vs 1.717 ms per token).  1: all the CTAs the SMs hold, the first ones a block more than the rest (CTA i and CTA i + #SMs
        // land on one SM: the SMs come out even) -- better for the MoE form, whose groups would otherwise leave a quarter of the CTA slots
        // empty (Mixtral: 2.77 vs 2.81 ms per token).  B200_PAIR_GEOM=0/1 forces one.
        static const int geom_env = env_i("B200_PAIR_GEOM", -1);
        const int geom = geom_env >= 0 ? geom_env : (by_id ? 1 : 0);
        const int units = N[0] / 32, cap = std::max(1, sm_count() * cps_eff / ngr), per = (units + cap - 1) / cap,
                  share = geom == 1 ? std::min(cap, units) : (units + per - 1) / per;
        for (int g = 0; g < ngr; ++g) {
            p.cta0[g] = used; used += share;
            p.rows_qg[g] = units / share; p.rows_remg[g] = units % share;
        }
        if (by_id) {
            for (int g = 0; g < ngr; ++g) { p.Wg[g] = (const uint8_t *)W[g]; p.Dg[g] = probe ? nullptr : dst[g]; p.Du[g] = probe ? nullptr : a->sm_up_dst + (int64_t)g * a->d_sy; }
            p.Wup = (const uint8_t *)a->sm_up_W; p.sm_slot = a->d_sy; p.sm_xq_col = (int64_t)(N[0] / 256) * ACT_ITEM;
        } else {
            for (int g = 0; g < 2; ++g) { p.Wg[g] = (const uint8_t *)W[g]; p.Dg[g] = probe ? nullptr : dst[g]; }
        }
        rows_per_cta = 32 * ((units + share - 1) / share);
    } else
    for (int g = 0; g < ng; ++g) {
        int share = g + 1 == ng ? std::max(1, gx_max - used) : std::max(1, (int)((int64_t)gx_max * N[g] / ntot));
        share = std::max(1, std::min(share, (N[g] / unit + rpp - 1) / rpp));
        p.cta0[g] = used; used += share;
        p.Wg[g] = (const uint8_t *)W[g]; p.Dg[g] = probe ? nullptr : dst[g];
        p.rows_qg[g] = N[g] / unit / share; p.rows_remg[g] = N[g] / unit % share;      // in units of `unit` rows
        rows_per_cta = std::max(rows_per_cta, unit * ((N[g] / unit + share - 1) / share));
    }
    p.cta0[p.ngroup] = used;

    // rows per stage: about stage_kb of weights, at least four stages per CTA when it has the rows for them
    auto round_up = [](int x, int m) { return (x + m - 1) / m * m; };
    // U passes in flight per thread.  Measured (profiles/r01_mmvqr_sweep.log): one pass with 16-row stages beats two passes
    // (register pressure, spills for Q6_K); B200_MMVQR_U=2 keeps the variant reachable for tuning runs.
    const int U = eU > 0 ? std::min(2, eU) : 1;
    const int rq = rpp * U;                                  // rows per stage come in multiples of a full U-pass
    int R = std::max(rq, (stage_kb * 1024 / rb) / rq * rq);
    R = std::min(R, std::max(rq, round_up((rows_per_cta + 3) / 4, rq)));
    R = std::min(R, round_up(rows_per_cta, rpp));
    if (a->rope && (R & 1)) return MMVQR_SKIP;               // pairs must not straddle stages
    const int stage_bytes = round_up(R * rb + 16, 128);     // slack: the 2-byte aligned loads of Q6_K over-read 4 bytes
    const size_t budget = (size_t)smem_kb * 1024;
    const size_t per_stage = (size_t)stage_bytes + (size_t)R * nw * 4 + 16;
    const size_t act_bytes = (size_t)bpr * ACT_ITEM + (a->norm_x ? (size_t)K * 8 : 0);      // + the fp32 row of the fused norm and the norm weights
    const size_t tail_bytes = 16 + 128 + ((a->fuse_resid || a->rope) ? (size_t)(rows_per_cta + 2) * 4 : 0) + (pair ? (size_t)rows_per_cta * 8 : 0) + 16;   // act barrier, norm partials, residual rows / rope cos-sin
    if (budget < act_bytes + tail_bytes + 128 + per_stage) return MMVQR_SKIP;
    int nstages = (int)std::min<size_t>(max_stages, (budget - act_bytes - tail_bytes - 128) / per_stage);
    const bool rmp = ramp && !a->rope;
    const int q0 = rmp ? std::max(1, R >> 3) : R, q1 = rmp ? std::max(1, R >> 2) : R, q2 = rmp ? std::max(1, R >> 1) : R;
    const int need = rows_per_cta <= q0 ? 1 : rows_per_cta <= q0 + q1 ? 2 : rows_per_cta <= q0 + q1 + q2 ? 3 : 3 + (rows_per_cta - q0 - q1 - q2 + R - 1) / R;
    nstages = std::min(nstages, pair ? 2 * need : need);
    if (nstages < 1) return MMVQR_SKIP;
    if (probe) return 0;

    p.nx = a->norm_x; p.nw = a->norm_w; p.nnorm = a->norm_y; p.nout = a->norm_out; p.nxq = a->norm_x ? (uint8_t *)a->Xq : nullptr; p.neps = a->norm_eps;
    p.resid = a->fuse_resid; p.dsum = a->fuse_sum; p.resid_rows = (a->fuse_resid || pair) ? rows_per_cta : 0;      // pair mode: the CTA's gate sums live there
    if (a->rope) {
        const b200_gemv_rope * r = a->rope;
        p.rope_on = 1; p.rope_rows = rows_per_cta + 2;
        for (int g = 0; g < ng; ++g) {
            p.rmode[g] = r->mode[g]; p.rout[g] = r->out[g]; p.rcell[g] = (char *)r->cell[g]; p.rslot[g] = (char * const *)r->cell_slot[g];
            p.rctype[g] = r->cell_type[g]; p.rcesz[g] = r->cell_type[g] == T_F32 ? 4 : 2; p.rhd[g] = std::max(2, r->head_dim[g]);
        }
        p.rpos = r->pos; p.rff = r->ffac; p.rp = rope_dev_from(r->prm);
    }
    p.Xq = (const uint8_t *)a->Xq; p.K = K; p.BPR = bpr; p.TPR = tpr; p.lgTPR = lg; p.rb = rb;
    p.R = R; p.nstages = nstages; p.stage_bytes = stage_bytes;
    p.r0 = q0; p.r1 = q1; p.r2 = q2; p.stagger = stagger;
    p.dbg_sumi = a->dbg_sumi;
    if (a->sm_act) {
        if (!al16(dst[0]) || (by_id ? (!al16(a->sm_up_dst) || (a->d_sy & 3)) : !al16(dst[1]))) return MMVQR_SKIP;
        p.sm_act = a->sm_act; p.sm_out = a->sm_out; p.sm_xq = (uint8_t *)a->sm_xq;
        p.rope_rows = rows_per_cta;                       // ... and its products behind them (the rope slot of the tail: never both)
    }
    if (by_id) {
        p.ids = a->ids; p.w_expert = a->w_sy;
        for (int g = 0; g < ng; ++g) p.xq_col[g] = a->x_mod_y > 0 ? g % a->x_mod_y : g;
    }
    B200_HS_FROM_ARGS(p.hs, a);
    mmvqr_launch_cfg c;
    c.grid = dim3(used, 1, 1);
    c.smem = (size_t)nstages * stage_bytes + act_bytes + (((size_t)nstages * R * nw + 1) & ~(size_t)1) * 4 + (size_t)nstages * 16 + tail_bytes;
    c.U = U;
    c.pdl = a->pdl >= 0 ? a->pdl : pdl_default();
    int rc = -1;
    switch (a->type) {
        case T_Q4_K: rc = mmvqr_launch_type<T_Q4_K>(p, c, stream); break;
        case T_Q6_K: rc = mmvqr_launch_type<T_Q6_K>(p, c, stream); break;
        case T_Q8_0: rc = mmvqr_launch_type<T_Q8_0>(p, c, stream); break;
        case T_Q4_0: rc = mmvqr_launch_type<T_Q4_0>(p, c, stream); break;
        case T_Q5_K: rc = mmvqr_launch_type<T_Q5_K>(p, c, stream); break;
        case T_Q4_1: rc = mmvqr_launch_type<T_Q4_1>(p, c, stream); break;
        case T_Q5_0: rc = mmvqr_launch_type<T_Q5_0>(p, c, stream); break;
        case T_Q5_1: rc = mmvqr_launch_type<T_Q5_1>(p, c, stream); break;
        case T_IQ4_NL: rc = mmvqr_launch_type<T_IQ4_NL>(p, c, stream); break;
        case T_Q2_K: rc = mmvqr_launch_type<T_Q2_K>(p, c, stream); break;
        case T_Q3_K: rc = mmvqr_launch_type<T_Q3_K>(p, c, stream); break;
        case T_IQ4_XS: rc = mmvqr_launch_type<T_IQ4_XS>(p, c, stream); break;
    }
    if (rc) return rc;
    ++g_launches;
    return 0;
}

}  // namespace b200

extern "C" int b200_mmvq_fusable(const b200_mmvq_args * a, int ng, const void * const * W, const int * N) {
    if (!a) return 0;
    const void * W1[1] = { a->W }; const int N1[1] = { a->N };
    if (!W || !N) { W = W1; N = N1; ng = 1; }
    if (ng < 1 || ng > 3) return 0;
    b200_mmvq_args t = *a;
    static const char dummy[16] __attribute__((aligned(16))) = {0};
    if (!t.Xq && !t.norm_x) t.Xq = dummy;                  // asked before the activations exist: only the shape matters
    return b200::mmvqr_try(&t, ng, W, nullptr, N, (cudaStream_t)-1) == 0 ? 1 : 0;
}
