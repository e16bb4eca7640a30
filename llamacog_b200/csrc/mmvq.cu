// mmvq.cu -- host side of the decode GEMV: launch geometry + C-ABI entry points.
// Replaces reference ggml-cuda/mmvq.cu:491-561 (ggml_cuda_mul_mat_vec_q) and the launch plumbing
// of :226-489; see mmvq.cuh for the kernel.
#include "mmvq_launch.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cstdlib>
#include <cstdint>

namespace b200 {

template <> int mmvq_launch_type<T_Q4_0>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q8_0>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q4_1>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q2_K>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_IQ4_NL>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_IQ4_XS>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q3_K>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q5_0>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
template <> int mmvq_launch_type<T_Q5_1>(const mmvq_params &, const mmvq_launch_cfg &, cudaStream_t);
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

static const size_t SMEM_1CTA = 232448 - 1024;   // 227 KB opt-in limit minus static/alignment slack
static const size_t SMEM_2CTA = 115712 - 1024;   // half an SM: two CTAs (or two kernels under PDL) co-reside

static int env_int(const char * name, int dflt) { const char * e = getenv(name); return e ? atoi(e) : dflt; }

// Largest number of activation columns one launch can take for this K: the quantised activations
// of all columns live in shared memory (400 B per 256 elements per column), at most ~96 KB.
static int max_cols_per_launch(int K) {
    const int bpr = (K + 255) / 256;
    const int m = (int)((96 * 1024) / ((size_t)bpr * ACT_ITEM));
    return std::max(0, std::min(8, m));
}

static int item_bytes(int type) {
    switch (type) { case T_Q4_0: return 144; case T_Q2_K: return 84; case T_Q3_K: return 110; case T_IQ4_NL: return 144; case T_IQ4_XS: return 136; case T_Q4_1: return 160; case T_Q5_0: return 176; case T_Q5_1: return 192; case T_Q8_0: return 272; case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210; }
    return 0;
}

// Choose geometry.  Returns 0 on success.
static int configure(int type, int N, int K, int M, const b200_mmvq_args & a, mmvq_params & p, mmvq_launch_cfg & c) {
    const int rb = K / block_elems(type) * block_bytes(type);
    const int bpr = (K + 255) / 256;
    int G = 1;
    while (G < 32 && bpr % (2 * G) == 0) G <<= 1;
    const int nseg = bpr / G;
    const int sms = num_sms();

    const bool w_aligned = (((uintptr_t)a.W | (uintptr_t)a.nb1 | (uintptr_t)a.w_sy | (uintptr_t)a.w_sz) & 15) == 0;
    const bool contig = w_aligned && a.nb1 == rb && (rb % 16) == 0;
    const int pitch = contig ? rb : ((rb + 16 + 15) & ~15);
    const size_t act_bytes = (size_t)M * bpr * ACT_ITEM;

    // static knobs for tuning runs
    static const int eR = env_int("B200_MMVQ_R", 0), eCW = env_int("B200_MMVQ_CW", 0), eNG = env_int("B200_MMVQ_NG", 0),
                     eNS = env_int("B200_MMVQ_NS", 0), eCPS = env_int("B200_MMVQ_CPS", 0), eSB = env_int("B200_MMVQ_STAGE_KB", 32);

    int cps = eCPS > 0 ? eCPS : (M <= 4 ? 2 : 1);                  // CTAs per SM the kernel is sized for
    if (act_bytes + 2 * (size_t)pitch + 4096 > SMEM_2CTA) cps = 1;
    const int nct_max = sms * cps;
    const int rows_per_cta = (N + nct_max - 1) / nct_max;

    // search rows-per-stage R and consumer geometry (CW warps per group x NG groups)
    int bR = 1, bCW = 1, bNG = 1; double best = -1.0;
    const int stage_cap = eSB * 1024;
    for (int R = 1; R <= std::min(rows_per_cta, 256); ++R) {
        if (R > 1 && (size_t)R * pitch > (size_t)stage_cap) break;
        const int items = R * bpr;
        const int stages_per_cta = (rows_per_cta + R - 1) / R;
        for (int CW = 1; CW <= 8; ++CW) {
            const int passes = (items + 32 * CW - 1) / (32 * CW);
            const double util = (double)items / (passes * 32.0 * CW);
            const int NG = std::max(1, std::min(8 / CW, stages_per_cta));
            const int warps = CW * NG;
            // tail efficiency: the last stage of a CTA is partially filled
            const double fill = (double)rows_per_cta / (stages_per_cta * R);
            const double bytes = std::min(1.0, (double)R * pitch / 16384.0);
            const double score = util * fill * (0.5 + 0.5 * std::min(1.0, warps / 8.0)) * (0.75 + 0.25 * bytes);
            if (score > best + 1e-9) { best = score; bR = R; bCW = CW; bNG = NG; }
        }
    }
    if (eR > 0) bR = eR;
    if (eCW > 0) bCW = std::min(8, eCW);
    if (eNG > 0) bNG = eNG;
    if (bCW * bNG > 8) bNG = std::max(1, 8 / bCW);

    const int R = bR;
    const int stage_bytes = (((size_t)R * pitch + 15) & ~(size_t)15) + 288;   // slack: a partial tail item over-reads
    const int groups_per_cta = (rows_per_cta + R - 1) / R;
    const size_t fixed = act_bytes + (size_t)4 * R * nseg * M * 4 + 8 + 8 * 16 + 256;
    auto stages_for = [&](size_t budget) -> int {
        if (budget <= fixed) return 0;
        return (int)std::min<size_t>(8, (budget - fixed) / ((size_t)stage_bytes + (size_t)R * nseg * M * 4));
    };
    int want = std::min(eNS > 0 ? eNS : 4, std::max(1, groups_per_cta));
    want = std::max(want, std::min(bNG, groups_per_cta));
    int nstages = std::min(want, stages_for(cps == 2 ? SMEM_2CTA : SMEM_1CTA));
    if (nstages < std::min(2, want) && cps == 2) { cps = 1; nstages = std::min(want, stages_for(SMEM_1CTA)); }
    if (nstages < 1) return -3;    // a single stage does not fit in shared memory
    if (bNG > nstages) bNG = nstages;

    static const int eGPS = env_int("B200_MMVQ_GRID_PER_SM", 0);
    int gx = std::min(sms * (eGPS > 0 ? eGPS : cps), (N + R - 1) / R);
    gx = std::max(gx, 1);

    p.R = R; p.nstages = nstages; p.pitch = pitch; p.stage_bytes = stage_bytes; p.contig = contig ? 1 : 0;
    p.CW = bCW; p.NG = bNG; p.BPR = bpr; p.G = G; p.act_bytes = (int)act_bytes;
    p.rows_q = N / gx; p.rows_rem = N % gx;
    c.M = M; c.A16 = (w_aligned && item_bytes(type) % 16 == 0) ? 1 : 0;
    c.grid = dim3(gx, std::max(1, a.ny), std::max(1, a.nz));
    c.block = dim3(288, 1, 1);   // producer warp + 8 warps (all quantise; CW*NG of them consume)
    c.smem = (size_t)nstages * stage_bytes + act_bytes + (((size_t)nstages * R * nseg * M + 1) & ~(size_t)1) * 4 + (size_t)nstages * 16 + 8;
    return 0;
}

static int mmvq_dispatch(int type, const mmvq_params & p, const mmvq_launch_cfg & c, cudaStream_t s) {
    switch (type) {
        case T_Q4_0: return mmvq_launch_type<T_Q4_0>(p, c, s);
        case T_Q8_0: return mmvq_launch_type<T_Q8_0>(p, c, s);
        case T_Q4_1: return mmvq_launch_type<T_Q4_1>(p, c, s);
        case T_Q2_K: return mmvq_launch_type<T_Q2_K>(p, c, s);
        case T_IQ4_NL: return mmvq_launch_type<T_IQ4_NL>(p, c, s);
        case T_IQ4_XS: return mmvq_launch_type<T_IQ4_XS>(p, c, s);
        case T_Q3_K: return mmvq_launch_type<T_Q3_K>(p, c, s);
        case T_Q5_0: return mmvq_launch_type<T_Q5_0>(p, c, s);
        case T_Q5_1: return mmvq_launch_type<T_Q5_1>(p, c, s);
        case T_Q4_K: return mmvq_launch_type<T_Q4_K>(p, c, s);
        case T_Q5_K: return mmvq_launch_type<T_Q5_K>(p, c, s);
        case T_Q6_K: return mmvq_launch_type<T_Q6_K>(p, c, s);
    }
    return -1;
}

int g_launches = 0;
int mmvqr_try(const b200_mmvq_args * a, int ng, const void * const * W, float * const * dst, const int * N, cudaStream_t stream);   // mmvq_r.cu
constexpr int MMVQR_SKIP = -777;
int pdl_default() { static const int v = [] { const char * e = getenv("B200_PDL"); return e ? atoi(e) : 1; }(); return v; }

}  // namespace b200

using namespace b200;

extern "C" int b200_mmvq_supported_type(int type) {
    return type == T_Q4_0 || type == T_Q8_0 || type == T_Q4_K || type == T_Q5_K || type == T_Q6_K || type == T_Q4_1 || type == T_Q5_0 || type == T_Q5_1 || type == T_Q2_K || type == T_Q3_K || type == T_IQ4_NL || type == T_IQ4_XS;
}

extern "C" int b200_kernel_launch_count(void) { return g_launches; }
extern "C" int b200_kernel_launch_add(int n) { g_launches += n; return g_launches; }

static void set_prefetch(mmvq_params & p, const b200_mmvq_args * a);
extern "C" int b200_mul_mat_vec_q_ex(const b200_mmvq_args * a, void * stream) {
    if (!a || !b200_mmvq_supported_type(a->type)) return -1;
    if (a->K <= 0 || a->K % block_elems(a->type) || a->N <= 0 || a->M <= 0) return -1;
    const int mmax = max_cols_per_launch(a->K);
    if (mmax <= 0) return -2;
    if ((a->hs_wait || a->hs_counter) && (!a->Xq || a->ids || a->M > mmax)) return -1;      // the hand-over hooks guard Xq, one launch
    {   // batch-1 decode on pre-quantised activations: register-resident kernel (mmvq_r.cuh)
        const void * W1[1] = { a->W }; float * D1[1] = { a->dst }; const int N1[1] = { a->N };
        const int rc = mmvqr_try(a, 1, W1, D1, N1, (cudaStream_t)stream);
        if (rc != MMVQR_SKIP) return rc;
        if (a->fuse_resid || a->fuse_sum || a->norm_x || a->rope || a->sm_act) return -4;     // only that kernel has the fused prologue / epilogue
    }
    const int pdl_env = pdl_default();
    for (int m0 = 0; m0 < a->M; m0 += mmax) {
        const int mc = std::min(mmax, a->M - m0);
        mmvq_params p = {};
        p.W = (const uint8_t *)a->W; p.nb1 = a->nb1;
        p.X = a->X ? a->X + (int64_t)m0 * a->xs : nullptr; p.xs = a->xs;
        p.Xq = a->Xq ? (const uint8_t *)a->Xq + (size_t)m0 * ((a->K + 255) / 256) * ACT_ITEM : nullptr;
        p.dst = a->dst + (int64_t)m0 * a->ds; p.ds = a->ds;
        p.N = a->N; p.K = a->K;
        p.w_sy = a->w_sy; p.w_sz = a->w_sz; p.w_div_y = std::max(1, a->w_div_y); p.w_div_z = std::max(1, a->w_div_z);
        p.x_sy = a->x_sy; p.x_sz = a->x_sz; p.x_mod_y = a->x_mod_y > 0 ? a->x_mod_y : 1 << 30;
        p.d_sy = a->d_sy; p.d_sz = a->d_sz; p.ids = a->ids; p.ids_sz = a->ids_sz;
        p.xq_sy = a->xq_sy; p.xq_sz = a->xq_sz;
        p.vec_ok = ((((uintptr_t)p.X) | ((uintptr_t)a->xs * 4) | ((uintptr_t)a->x_sy * 4) | ((uintptr_t)a->x_sz * 4)) & 15) == 0;
        p.dbg_sumi = a->dbg_sumi ? a->dbg_sumi + (int64_t)m0 * a->N * (a->K / 32) : nullptr;
        if (m0 + mmax >= a->M) set_prefetch(p, a);
        B200_HS_FROM_ARGS(p.hs, a);
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

static void set_prefetch(mmvq_params & p, const b200_mmvq_args * a) {
    p.pf_n = 0;
    for (int i = 0; i < a->pf_n && i < 3; ++i) {
        const uintptr_t q = ((uintptr_t)a->pf_ptr[i] + 15) & ~(uintptr_t)15;
        const int64_t bytes = a->pf_bytes[i] - (int64_t)(q - (uintptr_t)a->pf_ptr[i]);
        if (!a->pf_ptr[i] || bytes < 16) continue;
        p.pf_ptr[p.pf_n] = (const uint8_t *)q; p.pf_bytes[p.pf_n] = bytes; ++p.pf_n;
    }
}

// Several weight matrices (same type, K and row stride) times the same activations in one launch: q/k/v, gate/up.
extern "C" int b200_mul_mat_vec_q_group(const b200_mmvq_args * a, int ng, const void * const * W, float * const * dst, const int * N,
                                        const int64_t * ds, void * stream) {
    if (!a || ng < 1 || ng > 3 || !b200_mmvq_supported_type(a->type)) return -1;
    if (a->K <= 0 || a->K % block_elems(a->type) || a->M <= 0 || a->M > max_cols_per_launch(a->K)) return -1;
    if (std::max(1, a->ny) != 1 || std::max(1, a->nz) != 1 || a->ids || a->dbg_sumi) return -1;
    if ((a->hs_wait || a->hs_counter) && !a->Xq) return -1;
    int ntot = 0;
    for (int g = 0; g < ng; ++g) { if (N[g] <= 0) return -1; ntot += N[g]; }
    {
        const int rc = mmvqr_try(a, ng, W, dst, N, (cudaStream_t)stream);
        if (rc != MMVQR_SKIP) return rc;
        if (a->fuse_resid || a->fuse_sum || a->norm_x || a->rope || a->sm_act) return -4;
    }
    mmvq_params p = {};
    p.W = (const uint8_t *)W[0]; p.nb1 = a->nb1;
    p.X = a->X; p.xs = a->xs; p.Xq = (const uint8_t *)a->Xq;
    p.dst = dst[0]; p.ds = ds[0];
    p.N = ntot; p.K = a->K;
    p.w_div_y = 1; p.w_div_z = 1; p.x_mod_y = 1 << 30;
    p.vec_ok = ((((uintptr_t)p.X) | ((uintptr_t)a->xs * 4)) & 15) == 0;
    set_prefetch(p, a);
    B200_HS_FROM_ARGS(p.hs, a);
    b200_mmvq_args cfg = *a;
    uintptr_t align = 0;
    for (int g = 0; g < ng; ++g) align |= (uintptr_t)W[g];
    cfg.W = (const void *)align;                     // configure() only looks at the alignment bits
    cfg.ny = 1; cfg.nz = 1;
    mmvq_launch_cfg c = {};
    int rc = configure(a->type, ntot, a->K, a->M, cfg, p, c);
    if (rc) return rc;
    // split the CTAs over the matrices in proportion to their rows (at least one each)
    const int gx = (int)c.grid.x;
    int used = 0;
    p.ngroup = ng;
    for (int g = 0; g < ng; ++g) {
        int share = g + 1 == ng ? std::max(1, gx - used) : std::max(1, (int)((int64_t)gx * N[g] / ntot));
        share = std::min(share, (N[g] + p.R - 1) / p.R);
        share = std::max(share, 1);
        p.cta0[g] = used; used += share;
        p.Wg[g] = (const uint8_t *)W[g]; p.Dg[g] = dst[g]; p.dsg[g] = ds[g];
        p.rows_qg[g] = N[g] / share; p.rows_remg[g] = N[g] % share;
    }
    p.cta0[ng] = used;
    c.grid = dim3(used, 1, 1);
    c.pdl = a->pdl >= 0 ? a->pdl : pdl_default();
    rc = mmvq_dispatch(a->type, p, c, (cudaStream_t)stream);
    if (rc) return rc;
    ++g_launches;
    return 0;
}

extern "C" int b200_mul_mat_vec_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs,
                                  float * dst, int64_t ds, int N, int K, int M, void * stream) {
    b200_mmvq_args a = {};
    a.type = type; a.W = W; a.nb1 = nb1; a.X = X; a.xs = xs; a.dst = dst; a.ds = ds; a.N = N; a.K = K; a.M = M;
    a.ny = 1; a.nz = 1; a.pdl = -1;
    return b200_mul_mat_vec_q_ex(&a, stream);
}
