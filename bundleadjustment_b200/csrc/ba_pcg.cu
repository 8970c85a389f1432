// ba_pcg.cu -- Schur-Jacobi preconditioned conjugate gradients on the IMPLICIT reduced camera system
// (SURVEY.md 8a C6'; Ceres ITERATIVE_SCHUR + SCHUR_JACOBI).  Not what the reference configures
// (ba_project/src/ba/Optimizer.cpp:85 selects DENSE_SCHUR) but what BASELINE.json's K5 prescribes: S is never
// formed, S x = (U + D_c^2) x - sum_p W_p V_p'^-1 (W_p^T x) is applied from the stored W / V'^-1 blocks.
// The iteration follows ConjugateGradientsSolver::Solve (Ceres 2.0/2.1): q-tolerance stop rule, residual reset
// every 10 iterations, optional relative-residual stop.
#include "ba_launch.h"

#include <cooperative_groups.h>

#include <algorithm>
#include <cstdlib>

namespace ba {

__device__ __forceinline__ double lm_diag_pcg(const Ctx &ctx, double jtj, double radius) {
    return fmin(fmax(jtj, ctx.opt.min_lm_diagonal), ctx.opt.max_lm_diagonal) / radius;
}

#define CG_GUARD(st) if ((st)->done || (st)->cg_done) return;

// landmark contributions to block-diag(S) and rhs, CAMERA-major and deterministic: one warp per chunk of <= CHUNK
// observations of one camera gathers W_o (144 B), V'^-1 (48 B) and V'^-1 b_p (24 B) of its observations, accumulates the
// 21 unique entries of W V'^-1 W^T plus the 6 rhs entries per lane, reduces them with a fixed shuffle tree; the chunk
// partials are then summed per camera in chunk order (pcg_setup_reduce_kernel).  No atomics.
__global__ void __launch_bounds__(128) pcg_setup_cams_kernel(Ctx ctx) {
    const LMState *st = ctx.state;
    if (st->done) return;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= ctx.n_chunks) return;
    const int beg = ctx.chunk_beg[w], end = ctx.chunk_end[w];
    double acc[27];
#pragma unroll
    for (int k = 0; k < 27; ++k) acc[k] = 0.0;
    int e_n = 0, p_n = 0;                                  // indices one iteration ahead: they head a dependent chain
    if (beg + lane < end) { e_n = ctx.cm_obs[beg + lane]; p_n = ctx.cm_pt[beg + lane]; }
    for (int d = beg + lane; d < end; d += 32) {
        const int e = e_n, p = p_n;
        if (d + 32 < end) { e_n = ctx.cm_obs[d + 32]; p_n = ctx.cm_pt[d + 32]; }
        double wv[18], vi[6];                              // gathers as aligned 256-bit windows (7 requests instead of 12)
        load_row18(ctx.W, (size_t)e, wv);
        load_row6(ctx.Vinv, (size_t)p, vi);
        const double *w3 = ctx.wv + 3 * (size_t)p;
        const double t0 = w3[0], t1 = w3[1], t2 = w3[2];
        double Y[18];
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            Y[a * 3 + 0] = wv[a * 3] * vi[0] + wv[a * 3 + 1] * vi[1] + wv[a * 3 + 2] * vi[2];
            Y[a * 3 + 1] = wv[a * 3] * vi[1] + wv[a * 3 + 1] * vi[3] + wv[a * 3 + 2] * vi[4];
            Y[a * 3 + 2] = wv[a * 3] * vi[2] + wv[a * 3 + 1] * vi[4] + wv[a * 3 + 2] * vi[5];
            acc[21 + a] += wv[a * 3] * t0 + wv[a * 3 + 1] * t1 + wv[a * 3 + 2] * t2;
        }
        int k = 0;
#pragma unroll
        for (int a = 0; a < 6; ++a)
#pragma unroll
            for (int b = a; b < 6; ++b) { acc[k] += Y[a * 3] * wv[b * 3] + Y[a * 3 + 1] * wv[b * 3 + 1] + Y[a * 3 + 2] * wv[b * 3 + 2]; ++k; }
    }
#pragma unroll
    for (int k = 0; k < 27; ++k) acc[k] = warp_sum(acc[k]);
    if (lane == 0) {
        double *dst = ctx.cam_part + 27 * (size_t)w;
#pragma unroll
        for (int k = 0; k < 27; ++k) dst[k] = acc[k];
    }
}

// per camera: Minv <- (U + D^2) - sum of chunk partials (symmetric), rhs <- b_c - sum; replicated terms from rank 0 only
__global__ void pcg_setup_reduce_kernel(Ctx ctx) {
    const LMState *st = ctx.state;
    if (st->done) return;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ctx.Nc) return;
    double *M = ctx.Minv + 36 * (size_t)c;
    const bool lead = ctx.rank == 0;
    if (ctx.cam_kind[c] != CAM_FREE) {
#pragma unroll
        for (int k = 0; k < 36; ++k) M[k] = (lead && (k % 7 == 0)) ? 1.0 : 0.0;
#pragma unroll
        for (int a = 0; a < 6; ++a) ctx.rhs[6 * (size_t)c + a] = 0.0;
        return;
    }
    double acc[27];
#pragma unroll
    for (int k = 0; k < 27; ++k) acc[k] = 0.0;
    for (int w = ctx.cam_chunk_ptr[c]; w < ctx.cam_chunk_ptr[c + 1]; ++w) {
        const double *src = ctx.cam_part + 27 * (size_t)w;
#pragma unroll
        for (int k = 0; k < 27; ++k) acc[k] += src[k];
    }
    const double *U = ctx.U + 36 * (size_t)c;
    int k = 0;
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int b = a; b < 6; ++b) {
            double v = lead ? U[a * 6 + b] + (a == b ? lm_diag_pcg(ctx, U[a * 6 + a], st->radius) : 0.0) : 0.0;
            v -= acc[k++];
            M[a * 6 + b] = v; M[b * 6 + a] = v;
        }
#pragma unroll
    for (int a = 0; a < 6; ++a) ctx.rhs[6 * (size_t)c + a] = (lead ? ctx.bc[6 * (size_t)c + a] : 0.0) - acc[21 + a];
}

// The camera-side vector kernels of one CG iteration are COOPERATIVE launches (one CTA per SM at most): a dot product is
// reduced per CTA, published, and after grid.sync() EVERY CTA sums the published partials in the same fixed order, so all
// of them hold the identical scalar and go on to the dependent vector update -- one launch instead of a single-CTA kernel
// (which left 147 SMs idle for ~120 us per call, the dominant replicated cost on 8 GPUs).
constexpr int CG_THREADS = 256;
constexpr int CG_MAX_GRID = 1024;              // partial slots (grid is capped by co-residency, far below this)

// identical in every CTA: fixed-order sum of the per-CTA partials
__device__ __forceinline__ double grid_total(const double *part, int nblocks, double *red) {
    double v = 0.0;
    for (int i = threadIdx.x; i < nblocks; i += blockDim.x) v += part[i];
    v = block_sum(v, red);
    __shared__ double s_total;
    __syncthreads();
    if (threadIdx.x == 0) s_total = v;
    __syncthreads();
    return s_total;
}

// invert the 6x6 diagonal blocks in place (one thread per camera)
__global__ void __launch_bounds__(128) pcg_invert_blocks_kernel(Ctx ctx) {
    LMState *st = ctx.state;
    if (st->done) return;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ctx.Nc) return;
    double *M = ctx.Minv + 36 * (size_t)c;
    double A[36], inv[36];
#pragma unroll
    for (int k = 0; k < 36; ++k) A[k] = M[k];
    if (!inv_spd6(A, inv)) {
        atomicExch(&st->cg_bad_block, 1);
#pragma unroll
        for (int k = 0; k < 36; ++k) inv[k] = 0.0;
    }
#pragma unroll
    for (int k = 0; k < 36; ++k) M[k] = inv[k];
}

// start CG: x = 0, r = b, |b|.   Multi-CTA; the last CTA to arrive holds the total and sets up the CG scalars.
__global__ void __launch_bounds__(RED_THREADS) pcg_start_kernel(Ctx ctx) {
    LMState *st = ctx.state;
    if (st->done) return;
    __shared__ double red[128];
    double nb = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        const double b = ctx.rhs[i];
        ctx.cg_x[i] = 0.0;
        ctx.cg_r[i] = b;
        nb += b * b;
    }
    double u0 = 0.0, u1 = 0.0, u2 = 0.0;
    if (!grid_sum4(nb, u0, u1, u2, red_slot(ctx, RED_CG_START), ctx.red_count + RED_CG_START, red)) return;
    if (threadIdx.x == 0) {
        st->cg_norm_b = sqrt(nb);
        st->cg_rho = 1.0; st->cg_last_rho = 1.0; st->cg_Q0 = 0.0; st->cg_iter = 0; st->cg_done = 0;
        if (st->cg_bad_block || !isfinite(nb)) { st->chol_fail = 1; st->cg_done = 1; }
        st->cg_bad_block = 0;
        // ConjugateGradientsSolver::Solve: |b| = 0 -> x = 0, success, no iterations (all cameras fixed, or a converged point)
        if (nb == 0.0) st->cg_done = 1;
        // |r| <= r_tolerance * |b| at x = 0 (only when the relative-residual rule is enabled)
        if (ctx.opt.cg_r_tolerance > 0 && sqrt(nb) <= ctx.opt.cg_r_tolerance * sqrt(nb)) st->cg_done = 1;
        ctx.host_flags[1] = st->cg_done;
    }
}

// z = M^-1 r ; rho = r.z ; p = z + beta p ; q = 0.   Cooperative, grid-stride over cameras.
__global__ void __launch_bounds__(CG_THREADS) pcg_precond_kernel(Ctx ctx, double *part) {
    LMState *st = ctx.state;
    CG_GUARD(st);
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    __shared__ double red[32];
    const int it = st->cg_iter + 1;                 // read before anyone advances it (block 0 writes after the grid barrier)
    const double last_rho = st->cg_rho;
    double rho = 0.0;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ctx.Nc; c += gridDim.x * blockDim.x) {
        const double *M = ctx.Minv + 36 * (size_t)c;
        double r[6];
#pragma unroll
        for (int a = 0; a < 6; ++a) r[a] = ctx.cg_r[6 * (size_t)c + a];
#pragma unroll
        for (int a = 0; a < 6; ++a) {
            double z = 0.0;
#pragma unroll
            for (int b = 0; b < 6; ++b) z += M[a * 6 + b] * r[b];
            ctx.cg_z[6 * (size_t)c + a] = z;
            rho += r[a] * z;
        }
    }
    rho = block_sum(rho, red);
    if (threadIdx.x == 0) part[blockIdx.x] = rho;
    grid.sync();
    rho = grid_total(part, gridDim.x, red);
    bool stop = false;
    double beta = 0.0;
    if (rho == 0.0 || !isfinite(rho)) stop = true;
    else if (it > 1) {
        beta = rho / last_rho;
        if (beta == 0.0 || !isfinite(beta)) stop = true;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        st->cg_iter = it;
        st->cg_last_rho = last_rho;
        st->cg_rho = rho;
        if (stop) { st->chol_fail = 1; st->cg_done = 1; __threadfence_system(); ctx.host_flags[1] = 1; }
    }
    if (stop) return;
    const bool first = it == 1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        const double z = ctx.cg_z[i];
        ctx.cg_p[i] = first ? z : z + beta * ctx.cg_p[i];
        ctx.cg_q[i] = 0.0;
    }
}

// q -= sum_p W_p V'^-1 W_p^T v   (v = p or x), ATOMIC variant (pcg mode 2; the default when the ordered kernel's private copies
// of q would not fit in L2).  Tile kernel: the block's W records are read from HBM ONCE (coalesced,
// into shared memory) and serve both halves of the product: t_o = W_o^T v_c per observation, a_p = sum over the track
// (in track order), s_p = V'^-1 a_p per landmark, then q_c -= W_o s_p per observation (fp64 reductions into the
// L2-resident 6 Nc vector).
__global__ void __launch_bounds__(TILE_THREADS, 5) pcg_matvec_atomic_kernel(Ctx ctx, const double *__restrict__ v, double *__restrict__ qp) {
    const LMState *st = ctx.state;
    CG_GUARD(st);
    // the reductions go to one of CG_QCOPIES private copies of the vector (by CTA): when landmarks that are neighbours in
    // index share cameras, concurrent CTAs would otherwise serialise on the same few addresses
    double *__restrict__ q = qp + (size_t)(blockIdx.x % (unsigned)ctx.mv_grid) * ctx.n;
    extern __shared__ __align__(16) double tsm[];
    double *sW = tsm;                              // [cap][18]
    double *sT = tsm + 18 * (size_t)ctx.tile_cap;  // [cap][3]
    const int4 bi = ctx.blk_info[blockIdx.x];
    const int o0 = bi.z, n = bi.w;
    if (n == 0) return;
    // in flight across the barrier below: the thread's camera word, track word and landmark (V'^-1 on its way into L2)
    unsigned ck_first = 0u, trk_first = 0u;
    int p_first = 0;
    if (threadIdx.x < n) {
        ck_first = ctx.pm_camk[o0 + threadIdx.x]; trk_first = ctx.pm_trk[o0 + threadIdx.x]; p_first = ctx.pm_pt[o0 + threadIdx.x];
        prefetch_l2(ctx.Vinv + 6 * (size_t)p_first);
    }
    // the block's W records: one TMA bulk copy (thread 0) that streams in while the threads gather their camera entries of v
    __shared__ unsigned long long wbar;
    if (threadIdx.x == 0) mbar_init(&wbar, 1);
    __syncthreads();
    if (threadIdx.x == 0) bulk_load(sW, ctx.W + 18 * (size_t)o0, 144u * (unsigned)n, &wbar);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int k = i;
        const unsigned ck = i == (int)threadIdx.x ? ck_first : ctx.pm_camk[o0 + i];
        const int c = camk_cam(ck);
        double a3[3] = {0, 0, 0};
        if (camk_kind(ck) == CAM_FREE) {
            double x[6];
            load_row6(v, (size_t)c, x);
            mbar_wait(&wbar, 0);
            double Wo[18];
            load_record18(sW, k, Wo);
#pragma unroll
            for (int a = 0; a < 6; ++a) { a3[0] += Wo[a * 3] * x[a]; a3[1] += Wo[a * 3 + 1] * x[a]; a3[2] += Wo[a * 3 + 2] * x[a]; }
        }
        sT[3 * (size_t)k] = a3[0]; sT[3 * (size_t)k + 1] = a3[1]; sT[3 * (size_t)k + 2] = a3[2];
    }
    // The thread that started the bulk copy always sees it complete before the CTA can leave the tile: a block whose
    // observations all belong to fixed cameras has no other waiter, and a CTA must not exit with a copy in flight.
    if (threadIdx.x == 0) mbar_wait(&wbar, 0);
    __syncthreads();
    // Second half, still one thread per OBSERVATION and no further barrier: every thread sums its landmark's track in track
    // order (the position/length word says where it starts: t shared-memory reads, the same for all observations of the
    // landmark), applies V'^-1 and scatters W_o s_p.  Redundant per observation, but all eight warps work; the former
    // thread-per-landmark pass kept two warps busy while six waited at a second barrier (33 % of the stall samples).
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned ck = i == (int)threadIdx.x ? ck_first : ctx.pm_camk[o0 + i];
        if (camk_kind(ck) != CAM_FREE) continue;
        const int c = camk_cam(ck);
        const unsigned trk = i == (int)threadIdx.x ? trk_first : ctx.pm_trk[o0 + i];
        const int p = i == (int)threadIdx.x ? p_first : ctx.pm_pt[o0 + i];
        const int lo = i - (int)(trk & 0xffffu), hi = lo + (int)(trk >> 16);
        double vi[6];
        load_row6(ctx.Vinv, (size_t)p, vi);
        double a3[3] = {0, 0, 0};
        for (int k = lo; k < hi; ++k) { a3[0] += sT[3 * (size_t)k]; a3[1] += sT[3 * (size_t)k + 1]; a3[2] += sT[3 * (size_t)k + 2]; }
        const double t0 = vi[0] * a3[0] + vi[1] * a3[1] + vi[2] * a3[2];
        const double t1 = vi[1] * a3[0] + vi[3] * a3[1] + vi[4] * a3[2];
        const double t2 = vi[2] * a3[0] + vi[4] * a3[1] + vi[5] * a3[2];
        mbar_wait(&wbar, 0);                       // (a thread whose camera was fixed in the first pass never waited)
        double Wo[18];
        load_record18(sW, i, Wo);
#pragma unroll
        for (int a = 0; a < 6; ++a)
            atomicAdd(q + 6 * (size_t)c + a, -(Wo[a * 3] * t0 + Wo[a * 3 + 1] * t1 + Wo[a * 3 + 2] * t2));
    }
}

// q -= sum_p W_p V'^-1 W_p^T v   (v = p or x): the landmark part of the implicit Schur product, one pass over W.
//
// PERSISTENT and DETERMINISTIC.  CTA g walks a contiguous range of landmark blocks with a 2-stage TMA ring: while block b is
// computed, the W tile of block b + 1 (one cp.async.bulk) and its index words are already on their way, so a tile buffer is
// being filled for the whole life of the CTA.  Per block:
//   1. thread / observation:  t_o = W_o^T v_c                      (W from shared memory, v_c gathered)
//   2. thread / observation:  a_p = sum of its landmark's t (track order), s_p = V'^-1 a_p, contribution c_o = -W_o s_p -> smem
//   3. thread / observation in CAMERA order (blk_ord, sorted once per problem): the first observation of each camera sums its
//      camera's contributions in that fixed order and adds the 6-vector to the CTA's PRIVATE copy of q -- a plain
//      read-modify-write: within a block every camera has exactly one writer, between blocks the CTA barrier orders them, and
//      no other CTA ever touches this copy.  No floating-point atomics: the product is bit-reproducible run to run
//      (north_star (1): "segmented reductions keyed by camera").
// pcg_fold_kernel then adds the copies in CTA order.
constexpr int MV_CTAS_PER_SM = 2;
__global__ void __launch_bounds__(TILE_THREADS, MV_CTAS_PER_SM) pcg_matvec_ordered_kernel(Ctx ctx, const double *__restrict__ v, double *__restrict__ qp,
                                                                                          int blocks_per_cta) {
    const LMState *st = ctx.state;
    CG_GUARD(st);
    const int b0 = blockIdx.x * blocks_per_cta, b1 = min(ctx.n_blocks, b0 + blocks_per_cta);
    if (b0 >= b1) return;
    double *q = qp + (size_t)blockIdx.x * ctx.n;
    extern __shared__ __align__(16) double tsm[];
    const size_t cap = ((size_t)ctx.tile_cap + 1) & ~(size_t)1;      // even: every region below stays 16-byte aligned
    double *sT = tsm + 36 * cap;                   // [cap][3]
    double *sC = sT + 4 * cap;                     // [cap][6] (double2 accesses)
    __shared__ unsigned long long bars[2];
    if (threadIdx.x == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
    __syncthreads();
    int4 bi = ctx.blk_info[b0];
    if (threadIdx.x == 0 && bi.w > 0) bulk_load(tsm, ctx.W + 18 * (size_t)bi.z, 144u * (unsigned)bi.w, &bars[0]);
    unsigned ck_n = 0u, trk_n = 0u, ord_n = 0u;
    int p_n = 0;
    if ((int)threadIdx.x < bi.w) {
        const int o = bi.z + threadIdx.x;
        ck_n = ctx.pm_camk[o]; trk_n = ctx.pm_trk[o]; p_n = ctx.pm_pt[o]; ord_n = ctx.blk_ord[o];
        prefetch_l2(ctx.Vinv + 6 * (size_t)p_n);
    }
    unsigned parity0 = 0u, parity1 = 0u;
    for (int b = b0; b < b1; ++b) {
        const int stage = (b - b0) & 1;
        const int o0 = bi.z, n = bi.w;
        const unsigned ck_first = ck_n, trk_first = trk_n, ord_first = ord_n;
        const int p_first = p_n;
        // the CTA's next block: W into the other stage (free since the barrier that ended the previous iteration), index words
        // into registers
        int4 bin = make_int4(0, 0, 0, 0);
        if (b + 1 < b1) {
            bin = ctx.blk_info[b + 1];
            if (threadIdx.x == 0 && bin.w > 0)
                bulk_load(tsm + 18 * cap * (size_t)(stage ^ 1), ctx.W + 18 * (size_t)bin.z, 144u * (unsigned)bin.w, &bars[stage ^ 1]);
            if ((int)threadIdx.x < bin.w) {
                const int o = bin.z + threadIdx.x;
                ck_n = ctx.pm_camk[o]; trk_n = ctx.pm_trk[o]; p_n = ctx.pm_pt[o]; ord_n = ctx.blk_ord[o];
                prefetch_l2(ctx.Vinv + 6 * (size_t)p_n);
            }
        }
        if (n > 0) {
            const double *sW = tsm + 18 * cap * (size_t)stage;
            unsigned long long *bar = &bars[stage];
            const unsigned par = stage ? parity1 : parity0;
            // ---- 1: t_o = W_o^T v_c ----
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const unsigned ck = i == (int)threadIdx.x ? ck_first : ctx.pm_camk[o0 + i];
                double a3[3] = {0, 0, 0};
                if (camk_kind(ck) == CAM_FREE) {
                    double x[6];
                    load_row6(v, (size_t)camk_cam(ck), x);
                    mbar_wait(bar, par);
                    double Wo[18];
                    load_record18(sW, i, Wo);
#pragma unroll
                    for (int a = 0; a < 6; ++a) { a3[0] += Wo[a * 3] * x[a]; a3[1] += Wo[a * 3 + 1] * x[a]; a3[2] += Wo[a * 3 + 2] * x[a]; }
                }
                sT[3 * (size_t)i] = a3[0]; sT[3 * (size_t)i + 1] = a3[1]; sT[3 * (size_t)i + 2] = a3[2];
            }
            // the issuing thread always sees the copy complete in THIS phase of the barrier: the stage may only be re-armed (and its
            // parity flipped) once nobody can still be waiting on it
            if (threadIdx.x == 0) mbar_wait(bar, par);
            __syncthreads();
            // ---- 2: per observation, its landmark's s_p (redundantly per observation: all eight warps work) and c_o = -W_o s_p ----
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const unsigned ck = i == (int)threadIdx.x ? ck_first : ctx.pm_camk[o0 + i];
                double c6[6] = {0, 0, 0, 0, 0, 0};
                if (camk_kind(ck) == CAM_FREE) {
                    const unsigned trk = i == (int)threadIdx.x ? trk_first : ctx.pm_trk[o0 + i];
                    const int p = i == (int)threadIdx.x ? p_first : ctx.pm_pt[o0 + i];
                    const int lo = i - (int)(trk & 0xffffu), hi = lo + (int)(trk >> 16);
                    double vi[6];
                    load_row6(ctx.Vinv, (size_t)p, vi);
                    double a3[3] = {0, 0, 0};
                    for (int k = lo; k < hi; ++k) { a3[0] += sT[3 * (size_t)k]; a3[1] += sT[3 * (size_t)k + 1]; a3[2] += sT[3 * (size_t)k + 2]; }
                    const double t0 = vi[0] * a3[0] + vi[1] * a3[1] + vi[2] * a3[2];
                    const double t1 = vi[1] * a3[0] + vi[3] * a3[1] + vi[4] * a3[2];
                    const double t2 = vi[2] * a3[0] + vi[4] * a3[1] + vi[5] * a3[2];
                    double Wo[18];
                    load_record18(sW, i, Wo);
#pragma unroll
                    for (int a = 0; a < 6; ++a) c6[a] = -(Wo[a * 3] * t0 + Wo[a * 3 + 1] * t1 + Wo[a * 3 + 2] * t2);
                }
                double2 *dst = reinterpret_cast<double2 *>(sC + 6 * (size_t)i);
                dst[0] = make_double2(c6[0], c6[1]); dst[1] = make_double2(c6[2], c6[3]); dst[2] = make_double2(c6[4], c6[5]);
            }
            __syncthreads();
            // ---- 3: camera-keyed ordered fold into the CTA's private copy of q ----
            for (int j = threadIdx.x; j < n; j += blockDim.x) {
                const unsigned w = j == (int)threadIdx.x ? ord_first : ctx.blk_ord[o0 + j];
                if (!(w & 2048u)) continue;                                  // not the first observation of its camera in this block
                const unsigned cam = w >> 12;
                double acc[6];
                {
                    const double2 *src = reinterpret_cast<const double2 *>(sC + 6 * (size_t)(w & 2047u));
                    const double2 u0 = src[0], u1 = src[1], u2 = src[2];
                    acc[0] = u0.x; acc[1] = u0.y; acc[2] = u1.x; acc[3] = u1.y; acc[4] = u2.x; acc[5] = u2.y;
                }
                for (int k = j + 1; k < n; ++k) {
                    const unsigned wk = ctx.blk_ord[o0 + k];
                    if (wk & 2048u) break;                                   // next camera's segment
                    const double2 *src = reinterpret_cast<const double2 *>(sC + 6 * (size_t)(wk & 2047u));
                    const double2 u0 = src[0], u1 = src[1], u2 = src[2];
                    acc[0] += u0.x; acc[1] += u0.y; acc[2] += u1.x; acc[3] += u1.y; acc[4] += u2.x; acc[5] += u2.y;
                }
                double2 *dst = reinterpret_cast<double2 *>(q + 6 * (size_t)cam);
                double2 r0 = dst[0], r1 = dst[1], r2 = dst[2];
                r0.x += acc[0]; r0.y += acc[1]; r1.x += acc[2]; r1.y += acc[3]; r2.x += acc[4]; r2.y += acc[5];
                dst[0] = r0; dst[1] = r1; dst[2] = r2;
            }
            if (stage) parity1 ^= 1u; else parity0 ^= 1u;
        }
        __syncthreads();                           // sT, sC and this stage's W tile are free again; q updates ordered before the next block's
        bi = bin;
    }
}

// out = sum of the private copies (fixed order), copies zeroed for the next product
__global__ void pcg_fold_kernel(Ctx ctx, double *__restrict__ qp, double *__restrict__ out, int copies) {
    const LMState *st = ctx.state;
    CG_GUARD(st);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ctx.n) return;
    double t = 0.0;
#pragma unroll 8
    for (int k = 0; k < copies; ++k) { t += qp[(size_t)k * ctx.n + i]; qp[(size_t)k * ctx.n + i] = 0.0; }
    out[i] = t;
}

// out[i] = in[i] + ((U + D^2) v)[i]  helper (per camera row)
__device__ __forceinline__ double cam_block_row(const Ctx &ctx, const LMState *st, int c, int a, const double *v) {
    if (ctx.cam_kind[c] != CAM_FREE) return v[6 * (size_t)c + a];
    const double *U = ctx.U + 36 * (size_t)c;
    double s = 0.0;
#pragma unroll
    for (int b = 0; b < 6; ++b) s += U[a * 6 + b] * v[6 * (size_t)c + b];
    s += lm_diag_pcg(ctx, U[a * 6 + a], st->radius) * v[6 * (size_t)c + a];
    return s;
}

// q += (U + D^2) p ; alpha = rho / p.q ; x += alpha p ; r -= alpha q ; stop rules.   Cooperative.
__global__ void __launch_bounds__(CG_THREADS) pcg_update_kernel(Ctx ctx, double *part) {
    LMState *st = ctx.state;
    CG_GUARD(st);
    cooperative_groups::grid_group grid = cooperative_groups::this_grid();
    __shared__ double red[32];
    const double rho = st->cg_rho;
    double pq = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        const double q = ctx.cg_q[i] + cam_block_row(ctx, st, i / 6, i % 6, ctx.cg_p);
        ctx.cg_q[i] = q;
        pq += ctx.cg_p[i] * q;
    }
    pq = block_sum(pq, red);
    if (threadIdx.x == 0) part[blockIdx.x] = pq;
    grid.sync();
    pq = grid_total(part, gridDim.x, red);
    bool stop = false, fail = false;
    double alpha = 0.0;
    if (pq <= 0.0 || isinf(pq) || isnan(pq)) {
        stop = true;                                  // "Matrix is indefinite, no more progress can be made": keep the current x
    } else {
        alpha = rho / pq;
        if (isinf(alpha)) { stop = true; fail = true; }
    }
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            if (fail) st->chol_fail = 1;
            st->cg_done = 1; __threadfence_system(); ctx.host_flags[1] = 1;
        }
        return;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        ctx.cg_x[i] += alpha * ctx.cg_p[i];
        ctx.cg_r[i] -= alpha * ctx.cg_q[i];
    }
}

// residual reset: r = b - (q + (U + D^2) x), with q holding the landmark part of S x.
__global__ void __launch_bounds__(256) pcg_residual_reset_kernel(Ctx ctx) {
    LMState *st = ctx.state;
    CG_GUARD(st);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        const double ax = ctx.cg_z[i] + cam_block_row(ctx, st, i / 6, i % 6, ctx.cg_x);
        ctx.cg_r[i] = ctx.rhs[i] - ax;
    }
}

// Q1 = -x.(b + r); zeta = it (Q1 - Q0) / Q1 < eta -> stop; |r| <= r_tol |b| -> stop; it >= max -> stop
// ... and, in the same launch, the verdict as of the end of this CG iteration for the host's enqueue loop (see HF_CG_RING)
__global__ void __launch_bounds__(RED_THREADS) pcg_check_kernel(Ctx ctx, int slot) {
    LMState *st = ctx.state;
    if (st->done || st->cg_done) {                 // (what CG_GUARD tests) already over: only the verdict is due
        if (blockIdx.x == 0 && threadIdx.x == 0) ctx.host_flags[HF_CG_RING + slot] = 1;
        return;
    }
    __shared__ double red[128];
    double s = 0.0, rr = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n; i += gridDim.x * blockDim.x) {
        const double r = ctx.cg_r[i];
        s += ctx.cg_x[i] * (ctx.rhs[i] + r);
        rr += r * r;
    }
    double u0 = 0.0, u1 = 0.0;
    if (!grid_sum4(s, rr, u0, u1, red_slot(ctx, RED_CG_CHECK), ctx.red_count + RED_CG_CHECK, red)) return;
    if (threadIdx.x == 0) {
        const int it = st->cg_iter;
        const double Q1 = -1.0 * s;
        const double zeta = it * (Q1 - st->cg_Q0) / Q1;
        bool stop = false;
        if (zeta < ctx.opt.eta) stop = true;
        st->cg_Q0 = Q1;
        if (!stop && ctx.opt.cg_r_tolerance > 0 && sqrt(rr) <= ctx.opt.cg_r_tolerance * st->cg_norm_b) stop = true;
        if (!stop && it >= ctx.opt.max_linear_solver_iterations) stop = true;
        if (stop) { st->cg_done = 1; __threadfence_system(); ctx.host_flags[1] = 1; }
        ctx.host_flags[HF_CG_RING + slot] = stop ? 1 : 0;
    }
}

__global__ void pcg_finish_kernel(Ctx ctx) {
    LMState *st = ctx.state;
    if (st->done) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < ctx.n) ctx.yc[i] = (ctx.cam_kind[i / 6] == CAM_FREE) ? ctx.cg_x[i] : 0.0;
    if (i == 0) st->lin_iters = st->cg_iter;
}

static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// local part of the setup; the caller all-reduces Minv (36 Nc) and rhs (6 Nc) when sharded, then calls
// launch_pcg_iteration repeatedly (first call starts CG).
void launch_pcg_setup(const Ctx &ctx, cudaStream_t s) {
    if (ctx.n_chunks > 0) pcg_setup_cams_kernel<<<cdiv(ctx.n_chunks * 32, 128), 128, 0, s>>>(ctx);
    pcg_setup_reduce_kernel<<<cdiv(ctx.Nc, 128), 128, 0, s>>>(ctx);
}
static inline size_t pcg_tile_smem(const Ctx &ctx) { return sizeof(double) * 46 * (((size_t)ctx.tile_cap + 1) & ~(size_t)1); }   // 2 x 18 (W ring) + 4 (t, padded) + 6 (c)
static inline size_t pcg_atomic_smem(const Ctx &ctx) { return sizeof(double) * 21 * (size_t)ctx.tile_cap; }
// Private copies of q.  Ordered (bit-reproducible) kernel: one per persistent CTA, MV_CTAS_PER_SM per SM when shared memory allows,
// never more than blocks.  Atomic kernel: 32 copies, picked by block index, so that concurrent CTAs rarely hit the same address.
int pcg_matvec_grid(const Ctx &ctx) {
    if (!ctx.pcg_ordered) return 32;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int per_sm = pcg_tile_smem(ctx) * MV_CTAS_PER_SM + 4096 <= 227 * 1024 ? MV_CTAS_PER_SM : 1;
    return std::max(1, std::min(ctx.n_blocks, std::max(sms, 1) * per_sm));
}
// ordered kernel's copies for a problem of this shape, in bytes (the automatic mode picks it while they stay L2-resident)
size_t pcg_ordered_copy_bytes(const Ctx &ctx) {
    Ctx c = ctx;
    c.pcg_ordered = 1;
    return sizeof(double) * (size_t)pcg_matvec_grid(c) * (size_t)ctx.n;
}
int pcg_tiles_init(const Ctx &ctx) {
    cudaError_t e = cudaFuncSetAttribute(pcg_matvec_atomic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pcg_atomic_smem(ctx));
    if (e != cudaSuccess) return (int)e;
    if (pcg_tile_smem(ctx) > 227 * 1024) return 0;           // (blocks too long for the ordered kernel's ring: atomic kernel only)
    return (int)cudaFuncSetAttribute(pcg_matvec_ordered_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pcg_tile_smem(ctx));
}
// local landmark part of  qp -= W V'^-1 W^T v
static void launch_matvec(const Ctx &ctx, const double *v, cudaStream_t s) {
    if (ctx.n_blocks <= 0 || ctx.mv_grid <= 0) return;
    if (ctx.pcg_ordered) {
        const int per = (ctx.n_blocks + ctx.mv_grid - 1) / ctx.mv_grid;
        pcg_matvec_ordered_kernel<<<ctx.mv_grid, TILE_THREADS, pcg_tile_smem(ctx), s>>>(ctx, v, ctx.cg_qp, per);
    } else {
        pcg_matvec_atomic_kernel<<<ctx.n_blocks, TILE_THREADS, pcg_atomic_smem(ctx), s>>>(ctx, v, ctx.cg_qp);
    }
}
void launch_pcg_start(const Ctx &ctx, cudaStream_t s) {
    pcg_invert_blocks_kernel<<<cdiv(ctx.Nc, 128), 128, 0, s>>>(ctx);
    pcg_start_kernel<<<std::max(1, std::min(cdiv(ctx.n, 2 * RED_THREADS), 256)), RED_THREADS, 0, s>>>(ctx);
}

// grid of the cooperative camera-vector kernels: enough CTAs for the vector, never more than are co-resident
static int cg_grid(const Ctx &ctx, const void *kernel) {
    static int per_dev[64];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!per_dev[dev]) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, CG_THREADS, 0);
        per_dev[dev] = std::max(1, std::min(sms * std::max(per_sm, 1), sms));      // one CTA per SM is plenty
    }
    return std::max(1, std::min(std::min(per_dev[dev], CG_MAX_GRID), cdiv(ctx.n, CG_THREADS)));
}

void launch_pcg_iter_a(const Ctx &ctx, cudaStream_t s, cudaEvent_t ev0, cudaEvent_t ev1) {      // precondition + direction + local matvec
    Ctx c = ctx;
    double *part = ctx.cg_part;
    void *args[] = {&c, &part};
    cudaLaunchCooperativeKernel((const void *)pcg_precond_kernel, dim3(cg_grid(ctx, (const void *)pcg_precond_kernel)), dim3(CG_THREADS), args, 0, s);
    if (ev0) cudaEventRecord(ev0, s);
    launch_matvec(ctx, ctx.cg_p, s);
    if (ev1) cudaEventRecord(ev1, s);
    pcg_fold_kernel<<<cdiv(ctx.n, 256), 256, 0, s>>>(ctx, ctx.cg_qp, ctx.cg_q, ctx.mv_grid);
}
void launch_pcg_iter_b(const Ctx &ctx, cudaStream_t s) {
    Ctx c = ctx;
    double *part = ctx.cg_part;
    void *args[] = {&c, &part};
    cudaLaunchCooperativeKernel((const void *)pcg_update_kernel, dim3(cg_grid(ctx, (const void *)pcg_update_kernel)), dim3(CG_THREADS), args, 0, s);
}
void launch_pcg_reset_a(const Ctx &ctx, cudaStream_t s) {     // landmark part of S x into cg_z
    launch_matvec(ctx, ctx.cg_x, s);
    pcg_fold_kernel<<<cdiv(ctx.n, 256), 256, 0, s>>>(ctx, ctx.cg_qp, ctx.cg_z, ctx.mv_grid);
}
void launch_pcg_reset_b(const Ctx &ctx, cudaStream_t s) { pcg_residual_reset_kernel<<<cdiv(ctx.n, 256), 256, 0, s>>>(ctx); }
void launch_pcg_check(const Ctx &ctx, int iteration, cudaStream_t s) {
    pcg_check_kernel<<<std::max(1, std::min(cdiv(ctx.n, 2 * RED_THREADS), 256)), RED_THREADS, 0, s>>>(ctx, iteration % HF_CG_SLOTS);
}
void launch_pcg_finish(const Ctx &ctx, cudaStream_t s) { pcg_finish_kernel<<<cdiv(ctx.n, 256), 256, 0, s>>>(ctx); }

} // namespace ba
