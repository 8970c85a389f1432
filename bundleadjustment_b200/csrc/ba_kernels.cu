// ba_kernels.cu -- residual/Jacobian evaluation, normal-equation blocks, Schur assembly, back-substitution and
// the device-resident trust-region controller.
//
// Reference lines replaced: the per-residual AutoDiffCostFunction::Evaluate of AngleReprojectionError and
// PointOnlyReprojectionError (ba_project/src/ba/Optimizer.h:54-76, 96-107) + HuberLoss (Optimizer.cpp:312) and
// everything ceres::Solve (Optimizer.cpp:242) does with them per LM iteration.
#include "ba_launch.h"

#include <algorithm>
#include <cstdlib>

namespace ba {

#define DEV_STATE_GUARD(ctx)                 \
    const LMState *st__ = (ctx).state;       \
    if (st__->done) return;

// -------------------------------------------------------------------------------------------------------------
// C1 prologue: per-camera R, dR/dw (no sincos in the per-observation kernels)
// -------------------------------------------------------------------------------------------------------------
__global__ void cam_prep_kernel(Ctx ctx, int which /*0: x (with derivs, needs need_linearize), 1: candidate*/, int force) {
    const LMState *st = ctx.state;
    if (!force) {
        if (st->done) return;
        if (which == 0 && !st->need_linearize) return;
    }
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ctx.Nc) return;
    const int buf = which == 0 ? st->cur : (st->cur ^ 1);
    const int kind = ctx.cam_kind[c];
    const float *K = ctx.cam_K + 9 * (size_t)c;
    if (which == 1) {
        // candidate table: the 3 x 4 projection only.  Fixed cameras: K E[0:3, :] (written once, at set_problem time).
        double P[12];
        if (kind == CAM_FIXED) {
            const float *E = ctx.cam_extr + 16 * (size_t)c;
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    P[4 * i + j] = (double)K[3 * i] * (double)E[j] + (double)K[3 * i + 1] * (double)E[4 + j] + (double)K[3 * i + 2] * (double)E[8 + j];
        } else {
            const double *x = ctx.cam[buf] + 6 * (size_t)c;
            const double w[3] = {x[0], x[1], x[2]}, t[3] = {x[3], x[4], x[5]};
            double R[9];
            rodrigues(w, R, nullptr);
            cam_projection(K, R, t, P);
        }
        double *dst = ctx.tab_c[c].P;
#pragma unroll
        for (int i = 0; i < 12; ++i) dst[i] = P[i];
        return;
    }
    CamEntry *ce = ctx.tab_x + c;
#pragma unroll
    for (int i = 0; i < 9; ++i) ce->K[i] = (double)K[i];
    if (kind == CAM_FIXED) {
        const float *E = ctx.cam_extr + 16 * (size_t)c;
        double *P = ce->Jr;                       // 16 contiguous doubles: Jr[9] | sc[6] | pad1
#pragma unroll
        for (int i = 0; i < 16; ++i) P[i] = (double)E[i];
        return;
    }
    const double *x = ctx.cam[buf] + 6 * (size_t)c;
    double w[3] = {x[0], x[1], x[2]};
    double R[9], dR[27];
    rodrigues(w, R, dR);
#pragma unroll
    for (int i = 0; i < 9; ++i) ce->R[i] = R[i];
    // right Jacobian from the exact derivative: R^T dR_j is the skew matrix [Jr_j]x
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        const double *D = dR + 9 * j;
        double S[9];                          // S[a][b] = sum_i R[i][a] D[i][b]   (R, D column-major)
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
            for (int b = 0; b < 3; ++b) S[a * 3 + b] = R[0 + 3 * a] * D[0 + 3 * b] + R[1 + 3 * a] * D[1 + 3 * b] + R[2 + 3 * a] * D[2 + 3 * b];
        ce->Jr[3 * j + 0] = 0.5 * (S[2 * 3 + 1] - S[1 * 3 + 2]);
        ce->Jr[3 * j + 1] = 0.5 * (S[0 * 3 + 2] - S[2 * 3 + 0]);
        ce->Jr[3 * j + 2] = 0.5 * (S[1 * 3 + 0] - S[0 * 3 + 1]);
    }
#pragma unroll
    for (int a = 0; a < 6; ++a) ce->sc[a] = ctx.scale_cam[6 * (size_t)c + a];
    ce->t[0] = x[3]; ce->t[1] = x[4]; ce->t[2] = x[5];
}

// -------------------------------------------------------------------------------------------------------------
// Tile kernels (landmark side).  A CTA owns one landmark block = whole landmarks, a CONTIGUOUS run of <= tile_cap
// landmark-major observations (ba_prep.cu).  Work is split one THREAD PER OBSERVATION (no divergence over track
// lengths), per-observation results meet in shared memory, and per-landmark sums are taken there in track order
// (deterministic).  Everything that touches W moves through shared memory so that global traffic is 16-byte coalesced.
// -------------------------------------------------------------------------------------------------------------
struct Tile { int p0, p1, o0, n; };
__device__ __forceinline__ Tile tile_of(const Ctx &ctx, int b) {
    const int4 v = ctx.blk_info[b];
    Tile t;
    t.p0 = v.x; t.p1 = v.y; t.o0 = v.z; t.n = v.w;
    return t;
}
// n records of 18 doubles between global W (at observation o0) and shared memory: one TMA bulk copy each way.
// tile_store_W_begin: call after a CTA barrier that follows the last write of sW; tile_store_W_end: before the CTA exits.
// Only thread 0 does anything; the wait sits at the very end so that the per-landmark sums and the block reduction overlap
// with the drain of the tile (thread 0 waiting right after the issue held the whole CTA at its last barrier: 27 % of the
// kernel's stall samples).
__device__ __forceinline__ void tile_store_W_begin(double *__restrict__ W, int o0, int n, const double *__restrict__ sW) {
    if (threadIdx.x == 0 && n > 0) bulk_store_begin(W + 18 * (size_t)o0, sW, 144u * (unsigned)n);
}
__device__ __forceinline__ void tile_store_W_end() {
    if (threadIdx.x == 0) bulk_store_end();
}
// tile_load_W: starts the copy (thread 0) and returns; every thread calls tile_wait_W before touching sW.  bar is a shared
// 8-byte word used for this one transfer.
__device__ __forceinline__ void tile_load_W(const double *__restrict__ W, int o0, int n, double *__restrict__ sW, unsigned long long *bar) {
    if (threadIdx.x == 0) mbar_init(bar, 1);
    __syncthreads();
    if (threadIdx.x == 0 && n > 0) bulk_load(sW, W + 18 * (size_t)o0, 144u * (unsigned)n, bar);
}
__device__ __forceinline__ void tile_wait_W(unsigned long long *bar) { mbar_wait(bar, 0); }

// C1+C2+C3 (landmark side): residual, analytic Jacobian, Huber corrector, Jacobi scaling per observation;
// W = Jc^T Jp per observation; V_p, b_p per landmark; cost and gradient-norm partials per block.
template <int MIN_CTAS>
__global__ void __launch_bounds__(TILE_THREADS, MIN_CTAS) linearize_points_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && (st->done || !st->need_linearize)) return;
    extern __shared__ __align__(16) double tsm[];
    double *sW = tsm;                              // [cap][18]
    double *sV = tsm + 18 * (size_t)ctx.tile_cap;  // [cap][9]: Jp^T Jp (6) | Jp^T r (3)
    __shared__ double red[128];
    const Tile t = tile_of(ctx, blockIdx.x);
    const double *ptx = ctx.pt[st->cur];
    double cost = 0.0, gsq = 0.0, gmax = 0.0, xsq = 0.0;
    // second phase (thread per landmark): its track bounds are fetched now, off the critical path
    const int p_mine = t.p0 + threadIdx.x;
    int lo_mine = 0, hi_mine = 0;
    if (p_mine < t.p1) { lo_mine = ctx.pt_ptr[p_mine] - t.o0; hi_mine = ctx.pt_ptr[p_mine + 1] - t.o0; }
    for (int i = threadIdx.x; i < t.n; i += blockDim.x) {
        const int o = t.o0 + i, k = i;
        const unsigned ck = ctx.pm_camk[o];
        const int c = camk_cam(ck), kind = camk_kind(ck), p = ctx.pm_pt[o];
        const float2 uv = ctx.pm_uv[o];
        const double X[3] = {ptx[3 * (size_t)p], ptx[3 * (size_t)p + 1], ptx[3 * (size_t)p + 2]};
        const double sp[3] = {ctx.scale_pt[3 * (size_t)p], ctx.scale_pt[3 * (size_t)p + 1], ctx.scale_pt[3 * (size_t)p + 2]};
        double r[2], Jc[12], Jp[6], sc[6];
        double sr = 0.0;
        if (kind != OBS_MASKED) {
            obs_linearize<true, true>(ctx.tab_x + c, kind, X, uv, r, Jc, Jp, sc);
            double rho0;
            const double rho1 = huber_rho(r[0] * r[0] + r[1] * r[1], ctx.opt.huber_delta, ctx.opt.use_huber, rho0);
            sr = sqrt(rho1);
            cost += 0.5 * rho0;
        } else {                                   // dropped by the outer loop's prune: no residual block
            r[0] = r[1] = 0.0;
#pragma unroll
            for (int j = 0; j < 6; ++j) Jp[j] = 0.0;
        }
        r[0] *= sr; r[1] *= sr;
#pragma unroll
        for (int j = 0; j < 3; ++j) { const double f = sr * sp[j]; Jp[j] *= f; Jp[3 + j] *= f; }
        double *v = sV + 9 * (size_t)k;
        v[0] = Jp[0] * Jp[0] + Jp[3] * Jp[3];
        v[1] = Jp[0] * Jp[1] + Jp[3] * Jp[4];
        v[2] = Jp[0] * Jp[2] + Jp[3] * Jp[5];
        v[3] = Jp[1] * Jp[1] + Jp[4] * Jp[4];
        v[4] = Jp[1] * Jp[2] + Jp[4] * Jp[5];
        v[5] = Jp[2] * Jp[2] + Jp[5] * Jp[5];
#pragma unroll
        for (int j = 0; j < 3; ++j) v[6 + j] = Jp[j] * r[0] + Jp[3 + j] * r[1];
        // the record leaves as nine 16-byte shared stores: at the 144-byte record stride those are bank-conflict free,
        // eighteen 8-byte stores are 4-way conflicted
        double wr[18];
        if (kind == CAM_FREE) {
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double f = sr * sc[a];
                const double j0 = Jc[a] * f, j1 = Jc[6 + a] * f;
#pragma unroll
                for (int q = 0; q < 3; ++q) wr[a * 3 + q] = j0 * Jp[q] + j1 * Jp[3 + q];
            }
        } else {
#pragma unroll
            for (int q = 0; q < 18; ++q) wr[q] = 0.0;
        }
        double2 *Wo2 = reinterpret_cast<double2 *>(sW + 18 * (size_t)k);
#pragma unroll
        for (int q = 0; q < 9; ++q) Wo2[q] = make_double2(wr[2 * q], wr[2 * q + 1]);
    }
    __syncthreads();
    tile_store_W_begin(ctx.W, t.o0, t.n, sW);      // TMA store issued by thread 0; everybody goes on with the per-landmark sums
    for (int p = p_mine; p < t.p1; p += blockDim.x) {
        const int lo = p == p_mine ? lo_mine : ctx.pt_ptr[p] - t.o0, hi = p == p_mine ? hi_mine : ctx.pt_ptr[p + 1] - t.o0;
        double V[6] = {0, 0, 0, 0, 0, 0}, b[3] = {0, 0, 0};
        for (int k = lo; k < hi; ++k) {                     // track order: the same sum order as a serial walk
            const double *v = sV + 9 * (size_t)k;
#pragma unroll
            for (int j = 0; j < 6; ++j) V[j] += v[j];
#pragma unroll
            for (int j = 0; j < 3; ++j) b[j] += v[6 + j];
        }
        // part of the problem = has a residual block; with masked observations (outer loop) that is "some Jp is nonzero"
        if (hi > lo && (!ctx.masked || V[0] + V[3] + V[5] > 0.0)) {
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                const double x = ptx[3 * (size_t)p + j];
                const double g = b[j] / ctx.scale_pt[3 * (size_t)p + j];
                const double d = x - (x + (-g));
                gsq += d * d;
                gmax = fmax(gmax, fabs(d));
                xsq += x * x;
            }
        }
        double *Vp = ctx.V + 6 * (size_t)p;
#pragma unroll
        for (int j = 0; j < 6; ++j) Vp[j] = V[j];
        double *bp = ctx.bp + 3 * (size_t)p;
        bp[0] = b[0]; bp[1] = b[1]; bp[2] = b[2];
    }
    block_sum3_max(cost, gsq, xsq, gmax, red);
    if (threadIdx.x == 0) {
        double4 *pp = reinterpret_cast<double4 *>(ctx.part + 4 * (size_t)blockIdx.x);
        *pp = make_double4(cost, gsq, xsq, gmax);
    }
    tile_store_W_end();                            // shared memory must stay intact until the copy engine has read the tile
}

// -------------------------------------------------------------------------------------------------------------
// C3 (camera side): camera-major pass, one warp per chunk of <= CHUNK observations of ONE camera: U_c, b_c partials.
// U_c = sum_o Jc_o^T Jc_o and b_c = sum_o Jc_o^T r_o are ONE small product  A^T A  with A = the stacked rows [Jc | r] of
// the chunk (2 rows per observation, 7 columns): per batch of 32 observations every lane evaluates its observation, parks its
// two rows k-major in shared memory and the warp issues 16 DMMAs (m8n8k4; both operands are the same fragment).  Two
// accumulator registers per lane instead of 27 doubles and no 27-fold shuffle reduction; deterministic (fixed order).
// -------------------------------------------------------------------------------------------------------------
constexpr int CAMS_WARPS = 4;
constexpr int CAMS_LD = 64 + 4;              // 2 rows x 32 observations per batch; 68 mod 16 == 4 -> conflict-free fragment loads
__device__ __forceinline__ void dmma884_acc(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
template <int MIN_CTAS>
__global__ void __launch_bounds__(CAMS_WARPS * 32, MIN_CTAS) linearize_cams_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && (st->done || !st->need_linearize)) return;
    __shared__ __align__(16) double sJ[CAMS_WARPS][7 * CAMS_LD];       // [column a of [Jc | r]][k = 2 lane + row]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tig = lane & 3;
    const int w = blockIdx.x * CAMS_WARPS + warp;
    if (w >= ctx.n_chunks) return;
    const int c = ctx.chunk_cam[w];
    const int beg = ctx.chunk_beg[w], end = ctx.chunk_end[w];
    const CamEntry *ce = ctx.tab_x + c;
    const double *ptx = ctx.pt[st->cur];
    double *Js = sJ[warp];
    double acc[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};                // C[g][2 tig], C[g][2 tig + 1]: four independent DMMA chains
    // the next batch's pixel and landmark are fetched while this batch is evaluated (index -> landmark is a dependent chain)
    float2 uv_n = make_float2(0.f, 0.f);
    double Xn[3] = {0.0, 0.0, 0.0};
    if (beg + lane < end) { uv_n = ctx.cm_uv[beg + lane]; load_pt(ptx, (size_t)ctx.cm_pt[beg + lane], Xn); }
    for (int base = beg; base < end; base += 32) {
        const int e = base + lane;
        double r[2] = {0.0, 0.0}, Jc[12];
#pragma unroll
        for (int a = 0; a < 12; ++a) Jc[a] = 0.0;
        const float2 uv = uv_n;
        const double X[3] = {Xn[0], Xn[1], Xn[2]};
        if (e + 32 < end) { uv_n = ctx.cm_uv[e + 32]; load_pt(ptx, (size_t)ctx.cm_pt[e + 32], Xn); }
        if (e < end && !(ctx.masked && camk_kind(ctx.pm_camk[ctx.cm_obs[e]]) == OBS_MASKED)) {
            double Jp[6], sc[6];
            obs_linearize<true, false>(ce, CAM_FREE, X, uv, r, Jc, Jp, sc);
            double rho0;
            const double rho1 = huber_rho(r[0] * r[0] + r[1] * r[1], ctx.opt.huber_delta, ctx.opt.use_huber, rho0);
            const double sr = sqrt(rho1);
            r[0] *= sr; r[1] *= sr;
#pragma unroll
            for (int a = 0; a < 6; ++a) { const double f = sr * sc[a]; Jc[a] *= f; Jc[6 + a] *= f; }
        }
        __syncwarp();                                                   // the previous batch's DMMAs have consumed the buffer
#pragma unroll
        for (int a = 0; a < 6; ++a) *reinterpret_cast<double2 *>(Js + a * CAMS_LD + 2 * lane) = make_double2(Jc[a], Jc[6 + a]);
        *reinterpret_cast<double2 *>(Js + 6 * CAMS_LD + 2 * lane) = make_double2(r[0], r[1]);
        __syncwarp();
        const int left = end - base;
        const int ksteps = left >= 32 ? 16 : (2 * left + 3) / 4;        // dead lanes of the last batch wrote zeros
        const double *frag = Js + (g < 7 ? g : 0) * CAMS_LD + tig;          // A[g][k] and B[k][g] are the same number
        const double keep = g < 7 ? 1.0 : 0.0;
        int s4 = 0;
        for (; s4 + 4 <= ksteps; s4 += 4) {
#pragma unroll
            for (int q = 0; q < 4; ++q) { const double v = frag[4 * (s4 + q)] * keep; dmma884_acc(acc[q][0], acc[q][1], v, v); }
        }
        for (; s4 < ksteps; ++s4) { const double v = frag[4 * s4] * keep; dmma884_acc(acc[0][0], acc[0][1], v, v); }
    }
    const double c0 = (acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]), c1 = (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]);
    // upper triangle of U (row-major, 21 entries) | b_c (6): the layout cam_reduce_kernel sums
    if (g < 6) {
        double *dst = ctx.cam_part + 27 * (size_t)w;
        const int base_k = g * 6 - (g * (g - 1)) / 2 - g;               // + column = index of (g, column), column >= g
        const int col = 2 * tig;
        if (col >= g && col < 6) dst[base_k + col] = c0;
        if (col + 1 >= g && col + 1 < 6) dst[base_k + col + 1] = c1;
        if (col == 6) dst[21 + g] = c0;
    }
}

// per camera: ordered sum of its chunk partials -> U (full 6x6 row-major), b_c.  One warp per camera, lane k owns entry k
// of the 27 (21 upper-triangle entries of U + 6 of b_c): coalesced reads, the same chunk order as a serial walk.
__global__ void cam_reduce_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && (st->done || !st->need_linearize)) return;
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, k = threadIdx.x & 31;
    if (c >= ctx.Nc || k >= 27) return;
    double acc = 0.0;
    for (int w = ctx.cam_chunk_ptr[c]; w < ctx.cam_chunk_ptr[c + 1]; ++w) acc += ctx.cam_part[27 * (size_t)w + k];
    if (k < 21) {
        int a = 0, rem = k;                                   // k-th entry of the row-major upper triangle
        while (rem >= 6 - a) { rem -= 6 - a; ++a; }
        const int b = a + rem;
        double *U = ctx.U_loc + 36 * (size_t)c;
        U[a * 6 + b] = acc; U[b * 6 + a] = acc;
    } else ctx.bc_loc[6 * (size_t)c + (k - 21)] = acc;
}

// landmark-side partials -> sums / maxs (multi-CTA, fixed order)
__global__ void __launch_bounds__(RED_THREADS) reduce_lin_partials_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && (st->done || !st->need_linearize)) return;
    __shared__ double red[128];
    double cost = 0, gsq = 0, xsq = 0, gmax = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n_part; i += gridDim.x * blockDim.x) {
        const double4 pp = reinterpret_cast<const double4 *>(ctx.part)[i];
        cost += pp.x; gsq += pp.y; xsq += pp.z; gmax = fmax(gmax, pp.w);
    }
    if (!grid_sum3_max(cost, gsq, xsq, gmax, red_slot(ctx, RED_LIN), ctx.red_count + RED_LIN, red)) return;
    if (threadIdx.x == 0) {
        ctx.sums_loc[SUM_LIN_COST] = cost; ctx.sums_loc[SUM_LIN_GSQ] = gsq; ctx.sums_loc[SUM_LIN_XSQ] = xsq;
        ctx.maxs_loc[0] = gmax;
    }
}

__device__ void lm_begin_iteration(const Ctx &ctx);      // controller, part 1 (below)

// camera-side gradient / parameter norms from the (all-reduced) b_c; multi-CTA.  with_begin (LM loop): the last CTA to arrive
// -- every CTA has read the controller state by then -- also runs controller part 1, which used to be a launch of its own;
// when there was no re-linearisation the reduction runs on zeros just for that hand-over.
__global__ void __launch_bounds__(RED_THREADS) cam_norms_kernel(Ctx ctx, int force, int with_begin) {
    const LMState *st = ctx.state;
    const bool active = force || (!st->done && st->need_linearize);
    if (!active && !with_begin) return;
    __shared__ double red[128];
    const double *x = ctx.cam[st->cur];
    double gsq = 0, xsq = 0, gmax = 0;
    if (active) {
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 6 * ctx.Nc; i += gridDim.x * blockDim.x) {
            if (ctx.cam_kind[i / 6] != CAM_FREE) continue;
            const double g = ctx.bc[i] / ctx.scale_cam[i];
            const double d = x[i] - (x[i] + (-g));
            gsq += d * d; gmax = fmax(gmax, fabs(d)); xsq += x[i] * x[i];
        }
    }
    double unused = 0.0;
    if (!grid_sum3_max(gsq, xsq, unused, gmax, red_slot(ctx, RED_CAM_NORMS), ctx.red_count + RED_CAM_NORMS, red)) return;
    if (threadIdx.x == 0) {
        if (active) { ctx.camsums[CS_GSQ] = gsq; ctx.camsums[CS_GMAX] = gmax; ctx.camsums[CS_XSQ] = xsq; }
        if (with_begin) lm_begin_iteration(ctx);
    }
}

// Jacobi scaling from the iteration-0 (unscaled) blocks: scale = 1 / (1 + sqrt(diag(J^T J)))
__global__ void compute_scale_kernel(Ctx ctx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < 6 * ctx.Nc) {
        const int c = i / 6, a = i % 6;
        ctx.scale_cam[i] = (ctx.cam_kind[c] == CAM_FREE) ? 1.0 / (1.0 + sqrt(ctx.U[36 * (size_t)c + a * 6 + a])) : 1.0;
    }
    if (i < 3 * ctx.Np) {
        const int p = i / 3, a = i % 3;
        const int idx = (a == 0) ? 0 : (a == 1 ? 3 : 5);
        const bool active = ctx.pt_ptr[p + 1] > ctx.pt_ptr[p];
        ctx.scale_pt[i] = active ? 1.0 / (1.0 + sqrt(ctx.V[6 * (size_t)p + idx])) : 1.0;
    }
}

__global__ void fill_kernel(double *p, size_t n, double v) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// -------------------------------------------------------------------------------------------------------------
// C4 + C5: LM diagonal, V'^-1, Schur complement S = (U + D_c^2) - sum_p W V'^-1 W^T, rhs = b_c - sum_p W V'^-1 b_p
// -------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double lm_diag(const Ctx &ctx, double jtj, double radius) {
    return fmin(fmax(jtj, ctx.opt.min_lm_diagonal), ctx.opt.max_lm_diagonal) / radius;
}

__device__ __forceinline__ void point_invert_block(const Ctx &ctx, double radius_override, int bid) {
    const LMState *st = ctx.state;
    if (radius_override <= 0 && st->done) return;
    const double radius = radius_override > 0 ? radius_override : st->radius;
    const int p = bid * blockDim.x + threadIdx.x;
    if (p >= ctx.Np) return;
    const double *V = ctx.V + 6 * (size_t)p;
    double m[6] = {V[0], V[1], V[2], V[3], V[4], V[5]};
    m[0] += lm_diag(ctx, V[0], radius);
    m[3] += lm_diag(ctx, V[3], radius);
    m[5] += lm_diag(ctx, V[5], radius);
    double inv[6];
    if (!inv_sym3(m, inv)) {
#pragma unroll
        for (int k = 0; k < 6; ++k) inv[k] = nan("");
    }
    double *Vi = ctx.Vinv + 6 * (size_t)p;
#pragma unroll
    for (int k = 0; k < 6; ++k) Vi[k] = inv[k];
    const double *b = ctx.bp + 3 * (size_t)p;
    double *w = ctx.wv + 3 * (size_t)p;
    w[0] = inv[0] * b[0] + inv[1] * b[1] + inv[2] * b[2];
    w[1] = inv[1] * b[0] + inv[3] * b[1] + inv[4] * b[2];
    w[2] = inv[2] * b[0] + inv[4] * b[1] + inv[5] * b[2];
}

__global__ void __launch_bounds__(128) point_invert_kernel(Ctx ctx, double radius_override) { point_invert_block(ctx, radius_override, blockIdx.x); }

// diagonal blocks + rhs initialisation (S has been zeroed).  Only rank 0 contributes the replicated terms.
__device__ __forceinline__ void schur_init_block(const Ctx &ctx, double radius_override, int bid) {
    const LMState *st = ctx.state;
    if (radius_override <= 0 && st->done) return;
    const double radius = radius_override > 0 ? radius_override : st->radius;
    const int i = bid * blockDim.x + threadIdx.x;             // one thread per row of S
    if (i >= ctx.ldS) return;
    const bool lead = ctx.rank == 0;
    const int src = ctx.row_src[i];
    if (src < 0) { *s_elem(ctx, i, i) = lead ? 1.0 : 0.0; return; }     // identity padding
    const int c = src / 6, a = src % 6;
    double *row = s_elem(ctx, i, ctx.cam_row[c]);        // a camera's 6 x 6 block never straddles a tile
    if (ctx.cam_kind[c] == CAM_FREE) {
        const double *U = ctx.U + 36 * (size_t)c;
#pragma unroll
        for (int b = 0; b < 6; ++b) row[b] = lead ? U[a * 6 + b] + (a == b ? lm_diag(ctx, U[a * 6 + a], radius) : 0.0) : 0.0;
        ctx.rhs[src] = lead ? ctx.bc[src] : 0.0;
    } else {
#pragma unroll
        for (int b = 0; b < 6; ++b) row[b] = (lead && a == b) ? 1.0 : 0.0;
        ctx.rhs[src] = 0.0;
    }
}

__global__ void schur_init_kernel(Ctx ctx, double radius_override) { schur_init_block(ctx, radius_override, blockIdx.x); }
// LM loop, dense Schur: (V + D^2)^-1 per landmark and the initialisation of S in ONE launch (independent work: the first
// n_point_blocks CTAs take the landmarks, the rest the rows of S)
__global__ void __launch_bounds__(128) point_invert_schur_init_kernel(Ctx ctx, int n_point_blocks) {
    if ((int)blockIdx.x < n_point_blocks) point_invert_block(ctx, -1.0, blockIdx.x);
    else schur_init_block(ctx, -1.0, blockIdx.x - n_point_blocks);
}

// (The permutations rhs (camera order) -> rhs row of S (elimination order) and y (elimination order) -> yc (camera order)
// happen inside the Cholesky kernel: CholArgs::rhs_cam / y_cam / row_src.)

// C5 proper: S -= sum_p W_i V'^-1 W_j^T and rhs -= sum_p W_i V'^-1 b_p as DETERMINISTIC segmented reductions over the
// camera-pair-major index (ba_prep.cu).  One warp per unit (<= PAIR_UNIT entries of ONE S block): each lane accumulates
// the 6x6 block (and, on diagonal blocks, the 6 rhs entries) of its entries in registers, the 32 lane partials are
// transposed through shared memory and summed in a fixed order.  No atomics anywhere: every S block has exactly one
// writer (schur_segments_kernel), so the reduced system is bit-reproducible run to run.
// Traffic per entry: 2 x 144 B of W + 48 B V'^-1 (+ 24 B on diagonal blocks); consecutive segments share cameras, so all
// but the first touch of a W block come from L2.
constexpr int SCHUR_WARPS = 4;
constexpr int SCHUR_K = 96;                  // stacked reduction index of one batch: 32 entries x 3 landmark coordinates
constexpr int SCHUR_LD = SCHUR_K + 4;        // row stride 100 doubles: 100 mod 16 == 4 -> conflict-free DMMA fragment loads
// The block of one camera pair is  sum_e Y_e W_b,e^T  with Y_e = W_a,e V'^-1 (6 x 3): stacked over the entries this is ONE
// small GEMM  [Y_1 .. Y_E] (6 x 3E) * [W_b,1 .. W_b,E]^T (3E x 6), so the reduction over entries runs on the FP64 tensor
// pipe (DMMA m8n8k4, 6 of 8 rows/columns used; column 6 carries the stacked b_p, which makes rhs_c -= W V'^-1 b_p fall out
// of the same product).  Per batch of 32 entries each lane gathers its entry's rows, forms Y in registers and parks Y and
// W_b k-major in shared memory; the warp then issues <= 24 DMMAs.  Two accumulator registers per lane instead of 42.
// (Tried in round 2 and dropped after measuring: streaming each entry's rows into a per-warp shared-memory stage with cp.async one
// batch ahead -- 32 KB of shared memory per warp leaves 6 warps per SM, 807 us on K4 against 381 us for this register-gather
// form at 13 warps per SM: profiles/r02_K4_schur_units_cpasync_ncu.txt.)
__global__ void __launch_bounds__(SCHUR_WARPS * 32) schur_units_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && st->done) return;
    __shared__ __align__(16) double sY[SCHUR_WARPS][6 * SCHUR_LD];      // [a][k]: Y_e[a][q] at k = 3 e + q
    __shared__ __align__(16) double sB[SCHUR_WARPS][7 * SCHUR_LD];      // [b][k]: W_b,e[b][q]; row 6: b_p of self entries
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 2, tig = lane & 3;
    const int u = blockIdx.x * SCHUR_WARPS + warp;
    const PairIndex &pi = ctx.pairs;
    if (u >= pi.n_units) return;
    const int info = pi.unit_seg[u];
    const long long beg = pi.unit_beg[u];
    const long long end = beg + (info >> 1);
    const bool diag = info & 1;
    double *Ys = sY[warp], *Bs = sB[warp];
    double acc[4][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}};                // C[g][2 tig], C[g][2 tig + 1]: four independent DMMA chains
    long long e = beg + lane;
    int2 ob_n = make_int2(0, 0);
    int p_n = 0;
    if (e < end) { ob_n = pi.pair_obs[e]; p_n = ctx.pm_pt[ob_n.x]; }
    for (long long base = beg; base < end; base += 32, e += 32) {
        const bool live = e < end;
        const int2 ob = live ? ob_n : make_int2(0, 0);                  // dead lanes read row 0 (finite data) and contribute Y = 0
        const int p = live ? p_n : 0;
        if (e + 32 < end) { ob_n = pi.pair_obs[e + 32]; p_n = ctx.pm_pt[ob_n.x]; }
        // gathers as aligned 256-bit windows (12 sector requests per entry instead of 21); W_b is W_a on self entries
        double vi[6], wa[18], wb[18];
        load_row6(ctx.Vinv, (size_t)p, vi);
        load_row18(ctx.W, (size_t)ob.x, wa);
        if (ob.x != ob.y) load_row18(ctx.W, (size_t)ob.y, wb);
        else {
#pragma unroll
            for (int k = 0; k < 18; ++k) wb[k] = wa[k];
        }
        double b3[3] = {0.0, 0.0, 0.0};
        if (diag && live && ob.x == ob.y) {                             // the observation itself: its share of the rhs
            const double *bp = ctx.bp + 3 * (size_t)p;
            b3[0] = bp[0]; b3[1] = bp[1]; b3[2] = bp[2];
        }
        __syncwarp();                                                   // previous batch's DMMAs have consumed the buffers
        const double lv = live ? 1.0 : 0.0;
#pragma unroll
        for (int h = 0; h < 3; ++h) {                                   // rows 2h, 2h+1 of W_a
            const double r0[3] = {wa[6 * h] * lv, wa[6 * h + 1] * lv, wa[6 * h + 2] * lv};
            const double r1[3] = {wa[6 * h + 3] * lv, wa[6 * h + 4] * lv, wa[6 * h + 5] * lv};
            double *y0 = Ys + (2 * h) * SCHUR_LD + 3 * lane, *y1 = y0 + SCHUR_LD;
            y0[0] = r0[0] * vi[0] + r0[1] * vi[1] + r0[2] * vi[2];
            y0[1] = r0[0] * vi[1] + r0[1] * vi[3] + r0[2] * vi[4];
            y0[2] = r0[0] * vi[2] + r0[1] * vi[4] + r0[2] * vi[5];
            y1[0] = r1[0] * vi[0] + r1[1] * vi[1] + r1[2] * vi[2];
            y1[1] = r1[0] * vi[1] + r1[1] * vi[3] + r1[2] * vi[4];
            y1[2] = r1[0] * vi[2] + r1[1] * vi[4] + r1[2] * vi[5];
            double *w0 = Bs + (2 * h) * SCHUR_LD + 3 * lane, *w1 = w0 + SCHUR_LD;
            w0[0] = wb[6 * h]; w0[1] = wb[6 * h + 1]; w0[2] = wb[6 * h + 2];
            w1[0] = wb[6 * h + 3]; w1[1] = wb[6 * h + 4]; w1[2] = wb[6 * h + 5];
        }
        if (diag) {
            double *w6 = Bs + 6 * SCHUR_LD + 3 * lane;
            w6[0] = b3[0]; w6[1] = b3[1]; w6[2] = b3[2];
        }
        __syncwarp();
        const long long left = end - base;
        const int ksteps = left >= 32 ? SCHUR_K / 4 : (int)((3 * left + 3) / 4);
        const bool arow = g < 6, brow = g < 6 || (diag && g == 6);
        int s4 = 0;
        for (; s4 + 4 <= ksteps; s4 += 4) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const double av = arow ? Ys[g * SCHUR_LD + 4 * (s4 + q) + tig] : 0.0;
                const double bv = brow ? Bs[g * SCHUR_LD + 4 * (s4 + q) + tig] : 0.0;
                dmma884_acc(acc[q][0], acc[q][1], av, bv);
            }
        }
        for (; s4 < ksteps; ++s4) {
            const double av = arow ? Ys[g * SCHUR_LD + 4 * s4 + tig] : 0.0;
            const double bv = brow ? Bs[g * SCHUR_LD + 4 * s4 + tig] : 0.0;
            dmma884_acc(acc[0][0], acc[0][1], av, bv);
        }
    }
    const double c0 = (acc[0][0] + acc[1][0]) + (acc[2][0] + acc[3][0]), c1 = (acc[0][1] + acc[1][1]) + (acc[2][1] + acc[3][1]);
    double *dst = pi.unit_part + 42 * (size_t)u;
    if (g < 6) {
        const int c = 2 * tig;
        if (c < 6) { dst[g * 6 + c] = c0; dst[g * 6 + c + 1] = c1; }
        else dst[36 + g] = c0;                                          // column 6: the rhs share (zero on off-diagonal blocks)
    }
}

// one thread per (segment, entry): ordered sum of the segment's unit partials, single writer per S element
__global__ void schur_segments_kernel(Ctx ctx, int force) {
    const LMState *st = ctx.state;
    if (!force && st->done) return;
    const PairIndex &pi = ctx.pairs;
    const long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int seg = (int)(idx / 42), k = (int)(idx % 42);
    if (seg >= pi.n_seg) return;
    const int2 cams = pi.seg_cams[seg];
    if (k >= 36 && cams.x != cams.y) return;
    double t = 0.0;
    for (int u = pi.unit_ptr[seg]; u < pi.unit_ptr[seg + 1]; ++u) t += pi.unit_part[42 * (size_t)u + k];
    if (k < 36) {
        // block (camera x, camera y) of the lower triangle IN ELIMINATION ORDER: transposed if y comes later than x
        const int r = ctx.cam_row[cams.x] + k / 6, q = ctx.cam_row[cams.y] + k % 6;
        if (ctx.cam_row[cams.x] >= ctx.cam_row[cams.y]) *s_elem(ctx, r, q) -= t;
        else *s_elem(ctx, q, r) -= t;
    }
    else ctx.rhs[6 * (size_t)cams.x + (k - 36)] -= t;
}

// -------------------------------------------------------------------------------------------------------------
// C7: back-substitution y_p = V'^-1 (b_p - sum W^T y_c), candidate = x + (-y) * scale, model-cost pieces
// -------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(TILE_THREADS) backsub_kernel(Ctx ctx) {
    const LMState *st = ctx.state;
    if (st->done) return;
    extern __shared__ __align__(16) double tsm[];
    double *sW = tsm;                              // [cap][18]
    double *sT = tsm + 18 * (size_t)ctx.tile_cap;  // [cap][3]: W_o^T y_c
    __shared__ double red[128];
    const Tile t = tile_of(ctx, blockIdx.x);
    const double *ptx = ctx.pt[st->cur];
    double *ptc = ctx.pt[st->cur ^ 1];
    __shared__ unsigned long long wbar;
    const unsigned ck_first = threadIdx.x < t.n ? ctx.pm_camk[t.o0 + threadIdx.x] : 0u;   // in flight across the barrier below
    // second phase (thread per landmark): track bounds now, the landmark's rows on their way into L2
    const int p_mine = t.p0 + threadIdx.x;
    int lo_mine = 0, hi_mine = 0;
    if (p_mine < t.p1) {
        lo_mine = ctx.pt_ptr[p_mine] - t.o0; hi_mine = ctx.pt_ptr[p_mine + 1] - t.o0;
        prefetch_l2(ctx.Vinv + 6 * (size_t)p_mine); prefetch_l2(ctx.V + 6 * (size_t)p_mine);
        prefetch_l2(ctx.bp + 3 * (size_t)p_mine); prefetch_l2(ptx + 3 * (size_t)p_mine); prefetch_l2(ctx.scale_pt + 3 * (size_t)p_mine);
    }
    tile_load_W(ctx.W, t.o0, t.n, sW, &wbar);      // TMA: the tile streams in while the threads gather their camera steps
    for (int i = threadIdx.x; i < t.n; i += blockDim.x) {
        const int k = i;
        const unsigned ck = i == (int)threadIdx.x ? ck_first : ctx.pm_camk[t.o0 + i];
        const int c = camk_cam(ck);
        double a3[3] = {0, 0, 0};
        if (camk_kind(ck) == CAM_FREE) {
            double y[6];
            load_row6(ctx.yc, (size_t)c, y);
            tile_wait_W(&wbar);
            double Wo[18];
            load_record18(sW, k, Wo);
#pragma unroll
            for (int a = 0; a < 6; ++a) { a3[0] += Wo[a * 3] * y[a]; a3[1] += Wo[a * 3 + 1] * y[a]; a3[2] += Wo[a * 3 + 2] * y[a]; }
        }
        sT[3 * (size_t)k] = a3[0]; sT[3 * (size_t)k + 1] = a3[1]; sT[3 * (size_t)k + 2] = a3[2];
    }
    // the issuing thread always sees the bulk copy complete (a block seen only by fixed cameras has no other waiter;
    // the CTA must not exit with the copy still writing into its shared memory)
    if (threadIdx.x == 0 && t.n > 0) tile_wait_W(&wbar);
    __syncthreads();
    double dxsq = 0, yb = 0, yhy = 0;
    for (int p = p_mine; p < t.p1; p += blockDim.x) {
        const int lo = p == p_mine ? lo_mine : ctx.pt_ptr[p] - t.o0, hi = p == p_mine ? hi_mine : ctx.pt_ptr[p + 1] - t.o0;
        const double X[3] = {ptx[3 * (size_t)p], ptx[3 * (size_t)p + 1], ptx[3 * (size_t)p + 2]};
        if (hi > lo) {
            double a3[3] = {0, 0, 0};
            for (int k = lo; k < hi; ++k) { a3[0] += sT[3 * (size_t)k]; a3[1] += sT[3 * (size_t)k + 1]; a3[2] += sT[3 * (size_t)k + 2]; }
            const double *Vi = ctx.Vinv + 6 * (size_t)p;
            const double *b = ctx.bp + 3 * (size_t)p;
            const double e0 = b[0] - a3[0], e1 = b[1] - a3[1], e2 = b[2] - a3[2];
            const double y0 = Vi[0] * e0 + Vi[1] * e1 + Vi[2] * e2;
            const double y1 = Vi[1] * e0 + Vi[3] * e1 + Vi[4] * e2;
            const double y2 = Vi[2] * e0 + Vi[4] * e1 + Vi[5] * e2;
            const double *V = ctx.V + 6 * (size_t)p;
            yb += y0 * b[0] + y1 * b[1] + y2 * b[2];
            yhy += y0 * (V[0] * y0 + V[1] * y1 + V[2] * y2) + y1 * (V[1] * y0 + V[3] * y1 + V[4] * y2) +
                   y2 * (V[2] * y0 + V[4] * y1 + V[5] * y2) + 2.0 * (a3[0] * y0 + a3[1] * y1 + a3[2] * y2);
            const double *sp = ctx.scale_pt + 3 * (size_t)p;
            const double c0 = X[0] + (-y0) * sp[0], c1 = X[1] + (-y1) * sp[1], c2 = X[2] + (-y2) * sp[2];
            ptc[3 * (size_t)p] = c0; ptc[3 * (size_t)p + 1] = c1; ptc[3 * (size_t)p + 2] = c2;
            const double d0 = X[0] - c0, d1 = X[1] - c1, d2 = X[2] - c2;
            dxsq += d0 * d0 + d1 * d1 + d2 * d2;
        } else {
            ptc[3 * (size_t)p] = X[0]; ptc[3 * (size_t)p + 1] = X[1]; ptc[3 * (size_t)p + 2] = X[2];
        }
    }
    double unused = 0.0;
    block_sum3_max(dxsq, yb, yhy, unused, red);
    if (threadIdx.x == 0) {
        double4 *pp = reinterpret_cast<double4 *>(ctx.part + 4 * (size_t)blockIdx.x);
        *pp = make_double4(dxsq, yb, yhy, 0.0);
    }
}

// cameras: candidate = x + (-y_c) * scale, camera-side model-cost pieces, candidate R table.  Multi-CTA, thread per camera.
__global__ void __launch_bounds__(128) cam_update_kernel(Ctx ctx) {
    const LMState *st = ctx.state;
    if (st->done) return;
    __shared__ double red[128];
    const double *x = ctx.cam[st->cur];
    double *xc = ctx.cam[st->cur ^ 1];
    double dxsq = 0, yb = 0, yhy = 0;
    for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ctx.Nc; c += gridDim.x * blockDim.x) {
        const int kind = ctx.cam_kind[c];
        CamProj *cp = ctx.tab_c + c;
        double cand[6];
        if (kind == CAM_FREE) {
            const double *y = ctx.yc + 6 * (size_t)c;
            const double *U = ctx.U + 36 * (size_t)c;
            const double *sc = ctx.scale_cam + 6 * (size_t)c;
#pragma unroll
            for (int a = 0; a < 6; ++a) {
                const double xa = x[6 * (size_t)c + a];
                cand[a] = xa + (-y[a]) * sc[a];
                const double d = xa - cand[a];
                dxsq += d * d;
                yb += y[a] * ctx.bc[6 * (size_t)c + a];
                double uy = 0;
#pragma unroll
                for (int b = 0; b < 6; ++b) uy += U[a * 6 + b] * y[b];
                yhy += y[a] * uy;
            }
        } else {
#pragma unroll
            for (int a = 0; a < 6; ++a) cand[a] = x[6 * (size_t)c + a];
        }
#pragma unroll
        for (int a = 0; a < 6; ++a) xc[6 * (size_t)c + a] = cand[a];
        if (kind == CAM_FREE) {
            double R[9], P[12];
            rodrigues(cand, R, nullptr);
            const double tc[3] = {cand[3], cand[4], cand[5]};
            cam_projection(ctx.cam_K + 9 * (size_t)c, R, tc, P);
#pragma unroll
            for (int i = 0; i < 12; ++i) cp->P[i] = P[i];
        }
        // the projections of fixed cameras were written once, at set_problem time
    }
    double unused = 0.0;
    if (!grid_sum3_max(dxsq, yb, yhy, unused, red_slot(ctx, RED_CAM_UPDATE), ctx.red_count + RED_CAM_UPDATE, red)) return;
    if (threadIdx.x == 0) { ctx.camsums[CS_DXSQ] = dxsq; ctx.camsums[CS_YB] = yb; ctx.camsums[CS_YHY] = yhy; }
}

// -------------------------------------------------------------------------------------------------------------
// C8 (ii): cost-only evaluation at the candidate, one thread per observation (landmark-major order)
// -------------------------------------------------------------------------------------------------------------
// Persistent (grid-stride) with the next observation's index triplet prefetched: the chain  index -> gathers -> arithmetic
// of one observation overlaps with the index load of the next instead of being paid per CTA.
__global__ void __launch_bounds__(256) eval_candidate_kernel(Ctx ctx, double *part_out) {
    const LMState *st = ctx.state;
    if (st->done) return;
    __shared__ double red[32];
    const double *ptc = ctx.pt[st->cur ^ 1];
    double cost = 0.0;
    const int stride = gridDim.x * blockDim.x;
    int o = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned ck_n = 0u;
    int p_n = 0;
    float2 uv_n = make_float2(0.f, 0.f);
    if (o < ctx.No) { ck_n = ctx.pm_camk[o]; p_n = ctx.pm_pt[o]; uv_n = ctx.pm_uv[o]; }
    for (; o < ctx.No; o += stride) {
        const int c = camk_cam(ck_n), p = p_n;
        const bool live = camk_kind(ck_n) != OBS_MASKED;
        const float2 uv = uv_n;
        const int on = o + stride;
        if (on < ctx.No) { ck_n = ctx.pm_camk[on]; p_n = ctx.pm_pt[on]; uv_n = ctx.pm_uv[on]; }
        if (!live) continue;
        double X[3];
        load_pt(ptc, (size_t)p, X);
        double r[2];
        obs_residual(ctx.tab_c + c, X, uv, r);
        double rho0;
        huber_rho(r[0] * r[0] + r[1] * r[1], ctx.opt.huber_delta, ctx.opt.use_huber, rho0);
        cost += 0.5 * rho0;
    }
    cost = block_sum(cost, red);
    if (threadIdx.x == 0) part_out[blockIdx.x] = cost;
}

// step partials -> sums (multi-CTA, fixed order)
__global__ void __launch_bounds__(RED_THREADS) reduce_step_partials_kernel(Ctx ctx, const double *cost_part, int n_cost_part) {
    const LMState *st = ctx.state;
    if (st->done) return;
    __shared__ double red[128];
    double dxsq = 0, yb = 0, yhy = 0, cc = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < ctx.n_part; i += gridDim.x * blockDim.x) {
        const double4 pp = reinterpret_cast<const double4 *>(ctx.part)[i];
        dxsq += pp.x; yb += pp.y; yhy += pp.z;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_cost_part; i += gridDim.x * blockDim.x) cc += cost_part[i];
    if (!grid_sum4(dxsq, yb, yhy, cc, red_slot(ctx, RED_STEP), ctx.red_count + RED_STEP, red)) return;
    if (threadIdx.x == 0) {
        ctx.sums_loc[SUM_STEP_DXSQ] = dxsq; ctx.sums_loc[SUM_STEP_YB] = yb; ctx.sums_loc[SUM_STEP_YHY] = yhy;
        ctx.sums_loc[SUM_CAND_COST] = cc;
    }
}

// -------------------------------------------------------------------------------------------------------------
// C8/C9 controller, part 1 (single thread): runs after a (re)linearisation.  Fills gradient norms / cost of the
// iteration that triggered it, records the iteration (FinalizeIterationAndCheckIfMinimizerCanContinue) and opens
// the next one.
// -------------------------------------------------------------------------------------------------------------
__device__ void lm_begin_iteration(const Ctx &ctx) {
    LMState *st = ctx.state;
    if (st->done) return;
    const ba_options &o = ctx.opt;
    ba_iteration &it = st->it;
    if (st->need_linearize) {
        st->x_cost = ctx.sums[SUM_LIN_COST];
        it.cost = st->x_cost;
        double gm = ctx.camsums[CS_GMAX];
        for (int r = 0; r < max(ctx.n_ranks, 1); ++r) gm = fmax(gm, ctx.maxs[r]);      // one slot per rank (summed message)
        it.gradient_max_norm = gm;
        it.gradient_norm = sqrt(ctx.sums[SUM_LIN_GSQ] + ctx.camsums[CS_GSQ]);
        st->x_norm = sqrt(ctx.sums[SUM_LIN_XSQ] + ctx.camsums[CS_XSQ]);
        if (it.iteration == 0) {
            st->initial_cost = st->x_cost;
            it.step_is_valid = 1;
            it.step_is_successful = 1;
        }
        st->need_linearize = 0;
        if (!isfinite(st->x_cost)) {
            st->termination = BA_FAILURE; st->msg = MSG_EVAL_FAILED; st->done = 1; ctx.host_flags[0] = 1;
            return;
        }
    }
    // ---- FinalizeIterationAndCheckIfMinimizerCanContinue ----
    if (it.step_is_successful) ++st->num_successful; else ++st->num_unsuccessful;
    it.trust_region_radius = st->radius;
    ctx.iters[st->n_recorded++] = it;
    bool stop = false;
    if (it.iteration >= o.max_num_iterations) { st->termination = BA_NO_CONVERGENCE; st->msg = MSG_MAX_ITER; stop = true; }
    else if (it.step_is_successful && it.gradient_max_norm <= o.gradient_tolerance) { st->termination = BA_CONVERGENCE; st->msg = MSG_GRADIENT_TOL; stop = true; }
    else if (it.trust_region_radius <= o.min_trust_region_radius) { st->termination = BA_CONVERGENCE; st->msg = MSG_MIN_RADIUS; stop = true; }
    if (stop) { st->done = 1; __threadfence_system(); ctx.host_flags[0] = 1; return; }
    // ---- open the next iteration ----
    const double pg = it.gradient_max_norm, pn = it.gradient_norm;
    const int next = it.iteration + 1;
    it.iteration = next;
    it.step_is_valid = 0; it.step_is_successful = 0; it.linear_solver_iterations = 0;
    it.cost = 0; it.cost_change = 0; it.step_norm = 0; it.relative_decrease = 0; it.trust_region_radius = 0;
    it.gradient_max_norm = pg; it.gradient_norm = pn;    // carried over unless the step is accepted
    it.eta = o.eta;
    st->chol_fail = 0;
    st->lin_iters = 0;
}

// controller, part 2 (single thread): step validity, tolerances, rho test, accept/reject, radius update
__device__ void lm_decide(const Ctx &ctx) {
    LMState *st = ctx.state;
    if (st->done) return;
    const ba_options &o = ctx.opt;
    ba_iteration &it = st->it;
    it.linear_solver_iterations = st->lin_iters;
    const double yb = ctx.sums[SUM_STEP_YB] + ctx.camsums[CS_YB];
    const double yhy = ctx.sums[SUM_STEP_YHY] + ctx.camsums[CS_YHY];
    // model_cost_change = -(J s).(r + J s/2) with s = -y  ==  y.b - y^T (J^T J) y / 2
    const double mcc = yb - 0.5 * yhy;
    const double dxsq = ctx.sums[SUM_STEP_DXSQ] + ctx.camsums[CS_DXSQ];
    const bool solver_ok = !st->chol_fail && isfinite(mcc) && isfinite(dxsq);
    it.step_is_valid = (solver_ok && mcc > 0.0) ? 1 : 0;
    if (!it.step_is_valid) {
        // ---- HandleInvalidStep ----
        if (++st->num_invalid >= o.max_num_consecutive_invalid_steps) {
            st->termination = BA_FAILURE; st->msg = MSG_INVALID_STEPS; st->done = 1;
            __threadfence_system(); ctx.host_flags[0] = 1;
            return;
        }
        st->radius = st->radius / st->decrease_factor;      // StepIsInvalid() == StepRejected(0)
        st->decrease_factor *= 2.0;
        it.cost = st->x_cost; it.cost_change = 0; it.step_norm = 0; it.relative_decrease = 0;
        return;
    }
    st->num_invalid = 0;
    double cand_cost = ctx.sums[SUM_CAND_COST];
    if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
    // ---- ParameterToleranceReached ----
    it.step_norm = sqrt(dxsq);
    if (it.step_norm <= o.parameter_tolerance * (st->x_norm + o.parameter_tolerance)) {
        st->termination = BA_CONVERGENCE; st->msg = MSG_PARAM_TOL; st->done = 1;
        __threadfence_system(); ctx.host_flags[0] = 1;
        return;                                              // step NOT taken, iteration NOT recorded
    }
    // ---- FunctionToleranceReached ----
    it.cost_change = st->x_cost - cand_cost;
    if (fabs(it.cost_change) <= o.function_tolerance * st->x_cost) {
        st->termination = BA_CONVERGENCE; st->msg = MSG_FUNCTION_TOL; st->done = 1;
        __threadfence_system(); ctx.host_flags[0] = 1;
        return;
    }
    // ---- IsStepSuccessful ----
    it.relative_decrease = (st->x_cost - cand_cost) / mcc;
    if (it.relative_decrease > o.min_relative_decrease) {
        // HandleSuccessfulStep: x <- candidate; the re-linearisation runs at the top of the next body
        st->cur ^= 1;
        st->need_linearize = 1;
        it.step_is_successful = 1;
        const double t = 2.0 * it.relative_decrease - 1.0;
        st->radius = st->radius / fmax(1.0 / 3.0, 1.0 - t * t * t);
        st->radius = fmin(o.max_trust_region_radius, st->radius);
        st->decrease_factor = 2.0;
    } else {
        it.step_is_successful = 0;
        it.cost = cand_cost;
        st->radius = st->radius / st->decrease_factor;
        st->decrease_factor *= 2.0;
    }
}

// ... and, in the same launch, the verdict as of the end of this LM body for the host's enqueue loop (see HF_LM_RING)
__global__ void lm_decide_kernel(Ctx ctx, int slot) {
    lm_decide(ctx);
    ctx.host_flags[HF_LM_RING + slot] = ctx.state->done;
}

__global__ void lm_init_state_kernel(Ctx ctx) {
    LMState *st = ctx.state;
    memset(st, 0, sizeof(LMState));
    st->need_linearize = 1;
    st->radius = ctx.opt.initial_trust_region_radius;
    st->decrease_factor = 2.0;
    st->it.eta = ctx.opt.eta;
    ctx.host_flags[0] = 0;
    ctx.host_flags[1] = 0;
}

// -------------------------------------------------------------------------------------------------------------
// Materialised evaluation (ba_evaluate): residual + both Jacobian blocks per observation, robustified, unscaled.
// 176 B / observation of output; this is the "residual+Jacobian obs/sec" kernel.
//
// Persistent, WARP-granular: a warp owns runs of 32 consecutive observations.  Every lane evaluates one observation and
// parks its record in the warp's shared-memory stage (16-byte stores); the three output runs of the warp (res 512 B,
// J_cam 3072 B, J_pt 1536 B, each contiguous in global memory) then leave as TMA bulk stores issued by lane 0.  The stage
// is double buffered (wait_group.read 1 before a buffer is rewritten), so the stores of run i drain while the warp gathers
// and computes run i + 1, whose indices were prefetched one run ahead.  No CTA barrier anywhere: warps drift freely.
// -------------------------------------------------------------------------------------------------------------
constexpr int EV_WARPS = 4;
struct __align__(128) EvStage { double jc[32 * 12]; double jp[32 * 6]; double res[32 * 2]; };   // 5120 B
static_assert(sizeof(EvStage) == 5120, "EvStage layout");

__global__ void __launch_bounds__(EV_WARPS * 32, 5) evaluate_materialize_kernel(Ctx ctx, double *res, double *jc, double *jp, double *cost_part) {
    __shared__ EvStage stage[EV_WARPS][2];
    const LMState *st = ctx.state;
    const double *ptx = ctx.pt[st->cur];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int gw = blockIdx.x * EV_WARPS + warp, nw = gridDim.x * EV_WARPS;
    const int n_runs = (ctx.No + 31) >> 5;
    double cost = 0.0;
    unsigned c_n = 0u;
    int p_n = 0;
    float2 uv_n = make_float2(0.f, 0.f);
    if (gw < n_runs && gw * 32 + lane < ctx.No) { const int o = gw * 32 + lane; c_n = ctx.pm_camk[o]; p_n = ctx.pm_pt[o]; uv_n = ctx.pm_uv[o]; }
    int it = 0;
    for (int run = gw; run < n_runs; run += nw, ++it) {
        const int o0 = run * 32, o = o0 + lane;
        const int n = min(32, ctx.No - o0);
        const int c = camk_cam(c_n), kind = camk_kind(c_n), p = p_n;
        const float2 uv = uv_n;
        {   // indices of the warp's next run: in flight during this run's gathers and arithmetic
            const int on = o + nw * 32;
            if (run + nw < n_runs && on < ctx.No) { c_n = ctx.pm_camk[on]; p_n = ctx.pm_pt[on]; uv_n = ctx.pm_uv[on]; }
        }
        double r[2] = {0, 0}, Jc[12], Jp[6], sr = 0.0;
        if (lane < n && kind == OBS_MASKED) {
#pragma unroll
            for (int k = 0; k < 12; ++k) Jc[k] = 0.0;
#pragma unroll
            for (int k = 0; k < 6; ++k) Jp[k] = 0.0;
        } else if (lane < n) {
            const double X[3] = {ptx[3 * (size_t)p], ptx[3 * (size_t)p + 1], ptx[3 * (size_t)p + 2]};
            obs_linearize<true, true>(ctx.tab_x + c, kind, X, uv, r, Jc, Jp);
            double rho0;
            const double rho1 = huber_rho(r[0] * r[0] + r[1] * r[1], ctx.opt.huber_delta, ctx.opt.use_huber, rho0);
            sr = sqrt(rho1);
            cost += 0.5 * rho0;
        }
        EvStage &sg = stage[warp][it & 1];
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory");    // the stores of run it - 2 have read sg
        __syncwarp();
        if (lane < n) {
            double2 *c2 = reinterpret_cast<double2 *>(sg.jc + 12 * lane);
#pragma unroll
            for (int k = 0; k < 6; ++k) c2[k] = make_double2(Jc[2 * k] * sr, Jc[2 * k + 1] * sr);
            double2 *p2 = reinterpret_cast<double2 *>(sg.jp + 6 * lane);
#pragma unroll
            for (int k = 0; k < 3; ++k) p2[k] = make_double2(Jp[2 * k] * sr, Jp[2 * k + 1] * sr);
            reinterpret_cast<double2 *>(sg.res)[lane] = make_double2(r[0] * sr, r[1] * sr);
        }
        asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");                     // writers: generic -> async proxy
        __syncwarp();
        if (lane == 0) {
            bulk_store_issue(jc + 12 * (size_t)o0, sg.jc, 96u * (unsigned)n);
            bulk_store_issue(jp + 6 * (size_t)o0, sg.jp, 48u * (unsigned)n);
            bulk_store_issue(res + 2 * (size_t)o0, sg.res, 16u * (unsigned)n);
            asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");        // shared memory stays valid until read
    cost = warp_sum(cost);
    if (lane == 0) cost_part[gw] = cost;
}

// =============================================================================================================
// host-side launchers
// =============================================================================================================
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }
// grid of a multi-CTA reduction over n items: enough CTAs to spread the reads, never more than the partial slots
static inline int red_grid(int n, int per_cta) { return std::max(1, std::min(cdiv(n, per_cta), 256)); }
static inline size_t tile_smem(const Ctx &ctx, int doubles_per_obs) { return sizeof(double) * (size_t)doubles_per_obs * (size_t)ctx.tile_cap; }

// dynamic shared memory of the tile kernels grows with the longest landmark block; returns a cudaError
int tiles_init(const Ctx &ctx) {
    const size_t need = tile_smem(ctx, 27);
    if (need > 227 * 1024) return (int)cudaErrorInvalidConfiguration;
    cudaError_t e = cudaFuncSetAttribute(linearize_points_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)need);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(backsub_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem(ctx, 21));
    if (e != cudaSuccess) return (int)e;
    return pcg_tiles_init(ctx);
}

void launch_init_state(const Ctx &ctx, cudaStream_t s) { lm_init_state_kernel<<<1, 1, 0, s>>>(ctx); }

void launch_fill(double *p, size_t n, double v, cudaStream_t s) {
    if (n == 0) return;
    fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p, n, v);
}

void launch_cam_prep(const Ctx &ctx, int which, int force, cudaStream_t s) {
    cam_prep_kernel<<<cdiv(ctx.Nc, 128), 128, 0, s>>>(ctx, which, force);
}

// landmark + camera side of one linearisation (local part; the caller all-reduces U/bc/sums when sharded)
void launch_linearize_local(const Ctx &ctx, int force, cudaStream_t s) {
    launch_cam_prep(ctx, 0, force, s);
    if (ctx.n_blocks > 0) {
        // 2 CTAs/SM at 128 registers (no spills) beat 3 CTAs/SM at 80 registers + 128 B of spills by 4-5 % (K4 and K5)
        linearize_points_kernel<2><<<ctx.n_blocks, TILE_THREADS, tile_smem(ctx, 27), s>>>(ctx, force);
    }
    if (ctx.n_chunks > 0) linearize_cams_kernel<4><<<cdiv(ctx.n_chunks, CAMS_WARPS), CAMS_WARPS * 32, 0, s>>>(ctx, force);
    cam_reduce_kernel<<<cdiv(ctx.Nc * 32, 128), 128, 0, s>>>(ctx, force);
    reduce_lin_partials_kernel<<<red_grid(ctx.n_part, 4 * RED_THREADS), RED_THREADS, 0, s>>>(ctx, force);
}
void launch_cam_norms(const Ctx &ctx, int force, int with_begin, cudaStream_t s) {
    cam_norms_kernel<<<red_grid(6 * ctx.Nc, 2 * RED_THREADS), RED_THREADS, 0, s>>>(ctx, force, with_begin);
}

void launch_compute_scale(const Ctx &ctx, cudaStream_t s) {
    const int n = max(6 * ctx.Nc, 3 * ctx.Np);
    compute_scale_kernel<<<cdiv(n, 256), 256, 0, s>>>(ctx);
}

void launch_decide(const Ctx &ctx, int body, cudaStream_t s) { lm_decide_kernel<<<1, 1, 0, s>>>(ctx, body % HF_LM_SLOTS); }

// LM loop, dense Schur (S already cleared): point_invert + schur_init + schur_units + schur_segments as three launches
void launch_point_invert_schur_build(const Ctx &ctx, cudaStream_t s) {
    const int nA = cdiv(ctx.Np, 128);
    point_invert_schur_init_kernel<<<nA + cdiv(ctx.ldS, 128), 128, 0, s>>>(ctx, nA);
    if (ctx.pairs.n_units > 0) {
        schur_units_kernel<<<cdiv(ctx.pairs.n_units, SCHUR_WARPS), SCHUR_WARPS * 32, 0, s>>>(ctx, 0);
        schur_segments_kernel<<<(unsigned)(((long long)ctx.pairs.n_seg * 42 + 255) / 256), 256, 0, s>>>(ctx, 0);
    }
}
void launch_point_invert(const Ctx &ctx, double radius_override, cudaStream_t s) {
    point_invert_kernel<<<cdiv(ctx.Np, 128), 128, 0, s>>>(ctx, radius_override);
}

void launch_schur_build(const Ctx &ctx, double radius_override, cudaStream_t s) {
    const int force = radius_override > 0;
    schur_init_kernel<<<cdiv(ctx.ldS, 128), 128, 0, s>>>(ctx, radius_override);
    if (ctx.pairs.n_units > 0) {
        schur_units_kernel<<<cdiv(ctx.pairs.n_units, SCHUR_WARPS), SCHUR_WARPS * 32, 0, s>>>(ctx, force);
        schur_segments_kernel<<<(unsigned)(((long long)ctx.pairs.n_seg * 42 + 255) / 256), 256, 0, s>>>(ctx, force);
    }
}


void launch_backsub(const Ctx &ctx, cudaStream_t s) {
    if (ctx.n_blocks > 0) backsub_kernel<<<ctx.n_blocks, TILE_THREADS, tile_smem(ctx, 21), s>>>(ctx);
    cam_update_kernel<<<red_grid(ctx.Nc, 128), 128, 0, s>>>(ctx);
}

void launch_eval_candidate(const Ctx &ctx, double *cost_part, cudaStream_t s) {
    static int resident = 0;                      // CTAs of the persistent grid: every SM full
    if (!resident) {
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, eval_candidate_kernel, 256, 0);
        resident = std::max(1, sms * std::max(per_sm, 1));
    }
    const int nb = std::min(cdiv(ctx.No, 256), resident);
    if (nb > 0) eval_candidate_kernel<<<nb, 256, 0, s>>>(ctx, cost_part);
    reduce_step_partials_kernel<<<red_grid(std::max(ctx.n_part, nb), 4 * RED_THREADS), RED_THREADS, 0, s>>>(ctx, cost_part, nb);
}

// returns the number of cost partials written (one per warp)
int launch_materialize(const Ctx &ctx, double *res, double *jc, double *jp, double *cost_part, cudaStream_t s) {
    if (ctx.No <= 0) return 0;
    const int n_runs = cdiv(ctx.No, 32);
    static int ctas_per_sm = 0, sms = 0;
    if (!ctas_per_sm) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, evaluate_materialize_kernel, EV_WARPS * 32, 0);
        if (ctas_per_sm < 1) ctas_per_sm = 1;
    }
    const int grid = std::min(cdiv(n_runs, EV_WARPS), sms * ctas_per_sm);
    evaluate_materialize_kernel<<<grid, EV_WARPS * 32, 0, s>>>(ctx, res, jc, jp, cost_part);
    return grid * EV_WARPS;
}

} // namespace ba
