// ba_cholesky.cu -- dense Cholesky solve of the reduced camera system on the FP64 tensor pipe.
//
// Replaces Ceres' DenseSchurComplementSolver::SolveReducedLinearSystem (Eigen LLT) that the reference selects with
// options.linear_solver_type = ceres::DENSE_SCHUR (ba_project/src/ba/Optimizer.cpp:85).
//
// ONE persistent cooperative kernel (one CTA per SM) runs the whole factorisation + both substitutions as a tile
// dataflow: 128 x 128 tiles, LEFT-looking, tasks handed out column-major round-robin, tile readiness published through
// per-row progress counters (release/acquire), so there is no launch gap, no global barrier and the serial chain
// potrf(j) -> trsm(j+1, j) -> last k-slice of (j+1, j+1) -> potrf(j+1) overlaps with the bulk updates of later columns.
//
//   task (i, j), i >= j:   C = A_ij - sum_{k<j} L_ik L_jk^T          one long-K DMMA GEMM (K = 128 j), accumulators in
//                                                                     registers, A/B k-slices through a 4-stage cp.async ring
//        i == j:           C = L_jj L_jj^T in shared memory (warp-register 32 x 32 base case + DMMA panel/trailing updates),
//                          L_jj^-1 falls out of the same elimination ([A; I] trick) and is stored for the TRSMs
//        i >  j:           L_ij = C L_jj^-T as a second DMMA GEMM (C staged in shared memory)
//   rhs: the right-hand side is row n_pad of A (tile row T), so the forward substitution z = L^-1 rhs IS the tasks (T, j).
//   back substitution:     block m's owner applies z_m -= L_km^T y_k as the y_k appear, then y_m = L_mm^-T z_m.
//
// Layout contract: the matrix is stored TILE-PACKED -- only the 128 x 128 tiles of the plan exist, each contiguous and
// row-major (row stride 128), at offset plan->tile_off[i * T + j] (doubles) of one buffer of plan->total_doubles.  Logical
// matrix: n_pad = ceil(n / 128) * 128 rows/columns, lower triangle of the leading n x n block = S, rows/cols [n, n_pad)
// identity padding; tile row T (i = T) is the right-hand side: row 0 of tile (T, j) holds rhs[128 j ..], its other rows are
// zero.  The tiles S itself can touch come FIRST (plan->front_doubles), followed by plan->extra_doubles caller-owned doubles
// (the reduced right-hand side in camera order), then the fill-in tiles and the rhs tiles: a sharded solve all-reduces the
// prefix [S tiles | rhs] in ONE call straight out of the buffer the Schur kernels wrote (no staging copy).  FP64 has no tcgen05 kind on sm_100a: mma.sync.m8n8k4.f64 (SASS DMMA) is the tensor-pipe
// instruction (measured 37.1 TFLOP/s register-only peak on B200, tools/dmma_bench.cu).
#include "ba_launch.h"

#include <cooperative_groups.h>

#include <algorithm>
#include <cstring>
#include <vector>

namespace ba {

constexpr int BK = 32;                 // k-slice of the cp.async ring
constexpr int SPAD = BK + 4;           // 36 doubles: row stride == 4 (mod 16) -> conflict-free DMMA fragment loads
constexpr int NSTAGE = 3;
constexpr int PB = 32;                 // potrf panel width (warp-register base case)
constexpr int LDP = PB + 4;            // 36
constexpr int LDQ = PB + 2;            // 34: stride of the potrf hand-off buffers -- rows stay 16-byte aligned, column access is 2-way at worst
// warp 0 -> warp 1 hand-off of a finished pivot column inside potrf: 2 = one mbarrier per pivot (arrive has CTA-scope release
// semantics; try_wait suspends in hardware), 0 = st.release / ld.acquire on a progress word (the fence costs warp 0 ~50 clocks
// per pivot), 1 = volatile progress word without a fence (measurement only: not ordered by the memory model)
#ifndef POTRF_HANDOFF
#define POTRF_HANDOFF 2
#endif

// Tile edge NB is a template parameter: 128 for (near-)dense reduced systems (highest arithmetic intensity: 67 % of the FP64
// GEMM rate on a dense matrix), 64 for tile-sparse ones.  A sparse factorisation is a CHAIN of potrf -> trsm -> update tasks
// along the nested-dissection separators; per 128 rows of that chain a 64-row tiling pays two links of 1/4 .. 1/8 the work each
// instead of one (measured on K4: tools/chol_chain), so the chain -- the whole kernel there -- gets markedly shorter.
template <int NB_>
struct ChCfg {
    static constexpr int NB = NB_;
    static constexpr int KSL = NB / BK;                    // k-slices per tile
    static constexpr int LDT = NB + 4;                     // stride of the staged C tile (conflict-free like SPAD)
    static constexpr int MI = NB / 16, NJ = NB / 32;       // m8 x n8 DMMA tiles per warp: 8 warps as 2 x 4, warp tile (NB/2) x (NB/4)
    static constexpr int STAGE_DOUBLES = 2 * NB * SPAD;    // A slice + B slice
    static constexpr int TILE_DOUBLES = NB * LDT;
    static constexpr int RING_DOUBLES = NSTAGE * STAGE_DOUBLES;
    // Shared memory: [C tile | scratch].  The main-loop ring overlays both; after it the scratch holds either the TRSM's
    // 2-stage ring of L_jj^-1 slices or the potrf work arrays (Pv | invd | dinv).
    static constexpr int SCRATCH_DOUBLES = 2 * NB * SPAD;
    static constexpr int SMEM_DOUBLES = (TILE_DOUBLES + SCRATCH_DOUBLES > RING_DOUBLES) ? TILE_DOUBLES + SCRATCH_DOUBLES : RING_DOUBLES;
    static constexpr int SMEM = SMEM_DOUBLES * (int)sizeof(double);
    static_assert(NB * LDP + PB * LDP + NB + 2 * PB * LDQ + 2 * PB <= SMEM_DOUBLES - TILE_DOUBLES, "potrf scratch (Pv | invd | dinv | Lc | XuT | rdv | sqv) must fit behind the tile");
    static_assert(SMEM <= 227 * 1024, "shared memory budget");
    static_assert(NB == 64 || NB == 128, "tile edge");
};

int chol_ld(int n, int nb) { return ((n + nb - 1) / nb) * nb; }
// work area: L_jj^-1 blocks, y blocks, the chain-fusion hand-off tiles (one per tile row), then int flags
// [tiledone ((T + 1) x T) | ydone (T) | c2done (T) | abort (1) | next task (1) | CTAs started per SM (256)]
static inline int chol_flag_count(int T) { return (T + 1) * T + 2 * T + 2 + 256; }
size_t chol_work_doubles(int n, int nb) {
    const int T = chol_ld(n, nb) / nb;
    return 2 * (size_t)T * nb * nb + (size_t)T * nb + (size_t)(chol_flag_count(T) + 1) / 2 + 1;
}

// ---- small PTX helpers -------------------------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}
__device__ __forceinline__ int ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_relaxed(const int *p) {
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

// CTA-scope release / acquire on a shared-memory word (warp-to-warp hand-off inside potrf)
__device__ __forceinline__ void st_release_cta(int *p, int v) {
    asm volatile("st.release.cta.shared.s32 [%0], %1;\n" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}
__device__ __forceinline__ int ld_acquire_cta(const int *p) {
    int v;
    asm volatile("ld.acquire.cta.shared.s32 %0, [%1];\n" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
    return v;
}

// one arrival on an mbarrier (CTA-scope release: the arriving thread's earlier shared-memory writes, and those it has
// synchronised with, are visible to whoever observes the phase completion)
__device__ __forceinline__ void mbar_arrive(unsigned long long *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

// 1 / d to ~1 ulp: MUFU seed (20+ bits) + two Newton steps; d must be a normal positive number
__device__ __forceinline__ double fast_rcp(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;\n" : "=d"(r) : "d"(d));
    double e = fma(-d, r, 1.0);
    r = fma(r, e, r);
    e = fma(-d, r, 1.0);
    return fma(r, e, r);
}

// the same with ONE third-order step, r (1 + e + e^2) with e = 1 - d r: three dependent FP64 operations after the seed instead of
// four (relative error e^3 < 2^-60 for the >= 20-bit seed) -- for the serial pivot chain of potrf
__device__ __forceinline__ double fast_rcp3(double d) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;\n" : "=d"(r) : "d"(d));
    const double e = fma(-d, r, 1.0);
    const double t = fma(e, e, e);
    return fma(r, t, r);
}

struct CholArgs {
    double *A;            // tile-packed matrix (see the layout contract)
    const long long *toff;   // [(T + 1) * T] offset of tile (i, j) in doubles, -1 = not part of the plan
    int T;                // column tiles = n_pad / 128; row tiles = T + 1 (the last one is the rhs row)
    int n;                // logical order (entries of y written)
    double *invL;         // [T][NB*NB]
    double *ypad;         // [T][NB] solution blocks as they are produced (cross-CTA exchange)
    int *tiledone;        // [T + 1][T] tile (i, j) of L (row T: of z) is final
    int *ydone;           // [T]
    double *c2buf;        // [T][NB*NB] chain fusion: row i's last off-diagonal tile BEFORE its TRSM (C2 = A_ic - sum_k L_ik L_ck^T)
    int *c2done;          // [T] c2buf[i] is complete
    int *next_task;       // [1] tickets handed out beyond the first gridDim.x tasks
    int *sm_slot;         // [256] CTAs that have started on each SM
    // optional (LM loop): right-hand side read as rhs_cam[row_src[i]] (0 where row_src[i] < 0) instead of the rhs tiles, solution also
    // written to y_cam[row_src[i]] -- the camera-order <-> elimination-order permutations without launches of their own
    const int *row_src;
    const double *rhs_cam;
    double *y_cam;
    const unsigned char *partner;   // [ntasks] != 0: off-diagonal task (i, c) whose column c is the LAST k-block of diagonal task (i, i)
    int *abort_flag;      // [1] set when a wait timed out (never expected; keeps a bug from hanging the GPU)
    double *y;            // [n]
    LMState *state;       // may be nullptr
    int *fail_flag;       // may be nullptr
    long long *task_prof; // may be nullptr (tools only): [ntasks][4] ns: picked, operands consumed, factor/diag ready, published; then [T][2] phase 2
    long long *prof;      // may be nullptr; [grid][8] ns: start, phase-1 end, end, waited (ph 1), waited (ph 2), potrf, tasks, trsm
    // tile-sparsity plan (chol_plan_create): only filled tiles are tasks, only filled k-blocks are multiplied
    const int4 *task;     // [ntasks + 1] (i, j, dependency level, first klist entry), sorted by level
    const int *klist;     // k-blocks of each task's update, in the order their operands become ready
    const unsigned char *fill;   // [(T + 1)][T] filled tiles of L (row T = rhs)
    int ntasks;
};

constexpr long long SPIN_TIMEOUT_NS = 4000000000LL;    // 4 s

__device__ __forceinline__ long long global_ns() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;\n" : "=l"(t));
    return t;
}

// spin until *flag >= need; returns false if aborted.  Called by ONE thread per CTA; the verdict reaches the others through
// a CTA barrier, and everything they then read was written through L2 (cp.async.cg / __ldcg), never a stale L1 line.
__device__ __forceinline__ bool wait_ge(const int *flag, int need, int *abort_flag, long long &waited) {
    if (ld_acquire(flag) >= need) return true;
    const long long t0 = global_ns();
    unsigned spins = 0;
    while (true) {
        __nanosleep(32);
        if (ld_acquire(flag) >= need) { waited += global_ns() - t0; return true; }
        if ((++spins & 255u) == 0) {
            if (ld_acquire(abort_flag)) return false;
            if (global_ns() - t0 > SPIN_TIMEOUT_NS) { atomicExch(abort_flag, 1); return false; }
        }
    }
}

// -------------------------------------------------------------------------------------------------------------
// Diagonal tile: Cholesky + inverse of the factor, in shared memory, one CTA (8 warps).
// In-place augmented elimination of [A; I]: after the sweep the tile holds L in the lower triangle (diag included)
// and E = L^-T in the STRICT upper triangle (E[r][c], r < c), with diag(E) = 1 / diag(L) in dinv.  Per 32-column panel:
//   (A) warp 0 factors the 32 x 32 diagonal block in REGISTERS (lane = row, shuffles broadcast the pivot column) and
//       forward-substitutes its inverse (lane = column of L_d^-1);
//   (B) every other row x (L rows below, E rows above) gets  row <- row * L_d^-T  as DMMA tiles;
//   (C) masked DMMA trailing update of the columns to the right:  T[x][c] -= Pv[x] . Pv[c]  for x >= c (L) or x < c0 + 32 (E).
// Non-positive / non-finite pivots are replaced by 1 and reported (no hang, Ceres' LINEAR_SOLVER_FAILURE semantics
// are applied by the LM controller).
// -------------------------------------------------------------------------------------------------------------
__device__ long long g_potrf_phase[12];  // tools/chol_prof.cu: accumulated ns in (A), (B), (C), calls; [4..8]: clocks inside (A): load, factor sweep, scale + store L, inverse sweep, stores

template <int NB>
__device__ __noinline__ bool potrf_inverse_tile(double *__restrict__ Ts, double *__restrict__ Pv, double *__restrict__ invd,
                                                double *__restrict__ dinv, unsigned long long *piv, unsigned phase, int w0) {
    constexpr int LDT = ChCfg<NB>::LDT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, tig = lane & 3;
    double *Lc = dinv + NB;                                  // [PB][LDQ] row j = column j of the matrix at pivot j (unscaled)
    double *XuT = Lc + PB * LDQ;                             // [PB][LDQ] row c = column c of the unit factor's inverse
    double *rdv = XuT + PB * LDQ, *sqv = rdv + PB;           // [PB] 1 / d_j, sqrt(d_j)
    __shared__ int s_bad, s_prog;
    int *prog = &s_prog;                                     // (hand-off variants 0 / 1) pivots published, counted over the tile
    (void)prog; (void)piv; (void)phase;
    if (tid == 0) { s_bad = 0; s_prog = 0; }
    __syncthreads();
#ifdef CHOL_PROFILE_PHASES
    long long tA = 0, tB = 0, tC = 0;
#endif
    for (int c0 = 0; c0 < NB; c0 += PB) {
#ifdef CHOL_PROFILE_PHASES
        const long long t0 = global_ns();
#endif
        // ---- (A) ----
        // Two warps, pipelined.  Warp 0 runs the square-root-free factor sweep (L D L^T: only a reciprocal sits on the serial
        // pivot chain) and publishes every pivot column as soon as it is final; warp 1 runs the forward substitution for the
        // inverse of the UNIT factor one pivot behind it (lane = column of the inverse), so the second serial sweep overlaps
        // the first instead of following it.  The scaling by sqrt(D) and every store into the tile happen afterwards in an
        // epilogue spread over all eight warps.  (Before: one warp ran factor sweep, scaling, inverse sweep and stores back to
        // back, 10.3 k clocks per panel -- profiles/r02_chol_potrf_variants.txt.)
        if (warp == w0) {                                    // (w0 = 0 or 2: two co-resident CTAs keep their sweeps on different SM sub-partitions)
#ifdef CHOL_PROFILE_PHASES
            const long long ck0 = clock64();
#endif
            double a[PB];
            const double *row = Ts + (size_t)(c0 + lane) * LDT + c0;
#pragma unroll
            for (int k = 0; k < PB; ++k) a[k] = row[k];
            bool bad = false;
#ifdef CHOL_PROFILE_PHASES
            const long long ck1 = clock64();
#endif
            // Column j of the current matrix is broadcast through shared memory (one store per lane, 16-byte broadcast loads)
            // instead of 2 x (31 - j) shuffles; column j of the unit factor stays UNSCALED in Lc[j][.] (= l_ij d_j).
            double myd = 1.0;                                // d_lane
            // Software-pipelined by one pivot: pivot j's update is applied EARLY only to column j + 1 (the next pivot column);
            // its update of columns j + 2 .. is issued in the next iteration, in the shadow of that pivot's serial chain
            // (store -> broadcast load -> reciprocal).  Every element still receives its updates in pivot order.
            double ajs_prev = 0.0;
#pragma unroll
            for (int j = 0; j < PB; ++j) {
                Lc[j * LDQ + lane] = a[j];
                __syncwarp();
                double d = Lc[j * LDQ + j];
                if (!(d > 1e-300) || isinf(d)) { d = 1.0; bad = true; }
                const double rd = fast_rcp3(d);
                if (lane == j) myd = d;
                if (j >= 1) {                                // deferred: pivot j - 1 on columns j + 1 .. PB - 1
                    if ((j + 1) & 1) { if (j + 1 < PB) a[j + 1] = fma(-ajs_prev, Lc[(j - 1) * LDQ + j + 1], a[j + 1]); }
#pragma unroll
                    for (int c = (j + 2) & ~1; c < PB; c += 2) {      // 16-byte broadcast loads of column pairs
                        const double2 ac = *reinterpret_cast<const double2 *>(Lc + (j - 1) * LDQ + c);
                        a[c] = fma(-ajs_prev, ac.x, a[c]);
                        a[c + 1] = fma(-ajs_prev, ac.y, a[c + 1]);
                    }
                }
                const double ajs = a[j] * rd;
                if (j + 1 < PB) a[j + 1] = fma(-ajs, Lc[j * LDQ + j + 1], a[j + 1]);      // early: the next pivot column
                ajs_prev = ajs;
                if (lane == 0) {                                 // column j and 1 / d_j are final: hand them to warp 1
                    rdv[j] = rd;
#if POTRF_HANDOFF == 2
                    mbar_arrive(&piv[c0 + j]);
#elif POTRF_HANDOFF == 1
                    *reinterpret_cast<volatile int *>(prog) = c0 + j + 1;
#else
                    st_release_cta(prog, c0 + j + 1);
#endif
                }
            }
            // (Round 2 measured two alternatives of this sweep on the 64-row banded system and dropped both, profiles/
            // r02_chol_potrf_variants.txt: augmented elimination [A | I] with the pivot and row j of E travelling by shuffle
            // -- 496 double shuffles per panel, 12.5 k clocks; and rsqrt + shuffled 1/sqrt(d_j) in a scaling loop: 29 k clocks.)
            const double mys = sqrt(myd);
            sqv[lane] = mys;                                 // l_jj
            dinv[c0 + lane] = 1.0 / mys;                     // 1 / l_jj
            if (bad && lane == 0) s_bad = 1;
#ifdef CHOL_PROFILE_PHASES
            if (lane == 0) {
                const long long ck2 = clock64();
                atomicAdd((unsigned long long *)&g_potrf_phase[4], (unsigned long long)(ck1 - ck0));
                atomicAdd((unsigned long long *)&g_potrf_phase[5], (unsigned long long)(ck2 - ck1));
            }
#endif
        } else if (warp == w0 + 1) {
            // Xu = (unit factor)^-1, right-looking: x[i] -= l_ik x[k] with l_ik = Lc[k][i] / d_k; x[i] = 0 for i < lane falls
            // out of the recurrence.  Step k only needs column k, which warp 0 has just published.
            double x[PB];
#pragma unroll
            for (int i = 0; i < PB; ++i) x[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
            for (int k = 0; k < PB - 1; ++k) {
#if POTRF_HANDOFF == 2
                mbar_wait(&piv[c0 + k], phase);
#elif POTRF_HANDOFF == 1
                while (*reinterpret_cast<volatile int *>(prog) < c0 + k + 1) { }
#else
                while (ld_acquire_cta(prog) < c0 + k + 1) { }
#endif
                const double t = rdv[k] * x[k];
                if (((k + 1) & 1)) x[k + 1] = fma(-Lc[k * LDQ + k + 1], t, x[k + 1]);
#pragma unroll
                for (int i = (k + 2) & ~1; i < PB; i += 2) {
                    const double2 l2 = *reinterpret_cast<const double2 *>(Lc + k * LDQ + i);
                    x[i] = fma(-l2.x, t, x[i]);
                    x[i + 1] = fma(-l2.y, t, x[i + 1]);
                }
            }
#if POTRF_HANDOFF == 2
            mbar_wait(&piv[c0 + PB - 1], phase);             // every barrier completes exactly one phase per call
#endif
#pragma unroll
            for (int k = 0; k < PB; k += 2)                  // column `lane` of Xu, contiguous (16-byte stores, conflict-free at LDQ)
                *reinterpret_cast<double2 *>(XuT + lane * LDQ + k) = make_double2(x[k], x[k + 1]);
        }
        __syncthreads();
#ifdef CHOL_PROFILE_PHASES
        const long long ckE = clock64();
#endif
        // Epilogue, all warps.  X = L_d^-1 = D^-1/2 Xu (row k scaled by 1 / l_kk), L_d = (unit factor) sqrt(D).
        // Row pass (lane = k): row c0 + c of the tile = [L_d row c | E row c] (E[r][c] = X[c][r], strict upper) and row c0 + c
        // of Pv = column c of X (the E row incl. its diagonal 1 / l_cc and the zeros left of it) -- contiguous stores.
#pragma unroll
        for (int q = 0; q < PB / 8; ++q) {
            const int c = warp + 8 * q, k = lane;
            const double isk = dinv[c0 + k];
            const double v = XuT[c * LDQ + k] * isk;
            Pv[(size_t)(c0 + c) * LDP + k] = v;
            Ts[(size_t)(c0 + c) * LDT + c0 + k] = (k > c) ? v : (k == c) ? sqv[k] : Lc[k * LDQ + c] * isk;
        }
        // Column pass (lane = c): X row-major into invd, the B operand of (B)
#pragma unroll
        for (int q = 0; q < PB / 8; ++q) {
            const int k = warp + 8 * q;
            invd[(size_t)k * LDP + lane] = XuT[lane * LDQ + k] * dinv[c0 + k];
        }
#ifdef CHOL_PROFILE_PHASES
        if (tid == 0) atomicAdd((unsigned long long *)&g_potrf_phase[7], (unsigned long long)(clock64() - ckE));
#endif
        __syncthreads();
#ifdef CHOL_PROFILE_PHASES
        const long long t1 = global_ns();
#endif
        // ---- (B) rows outside the diagonal block: 96 rows = 12 m-tiles of 8 ----
        for (int mt = warp; mt < (NB - PB) / 8; mt += 8) {
            const int r = mt * 8 + g;
            const int xrow = r < c0 ? r : r + PB;
            double af[PB / 4];
#pragma unroll
            for (int ks = 0; ks < PB / 4; ++ks) af[ks] = Ts[(size_t)xrow * LDT + c0 + ks * 4 + tig];
            double o[PB / 8][2];
#pragma unroll
            for (int nt = 0; nt < PB / 8; ++nt) o[nt][0] = o[nt][1] = 0.0;
#pragma unroll
            for (int ks = 0; ks < PB / 4; ++ks)
#pragma unroll
                for (int nt = 0; nt < PB / 8; ++nt)
                    dmma884(o[nt][0], o[nt][1], af[ks], invd[(size_t)(nt * 8 + g) * LDP + ks * 4 + tig]);
            __syncwarp();
#pragma unroll
            for (int nt = 0; nt < PB / 8; ++nt) {
                const int col = nt * 8 + 2 * tig;
                Ts[(size_t)xrow * LDT + c0 + col] = o[nt][0];
                Ts[(size_t)xrow * LDT + c0 + col + 1] = o[nt][1];
                Pv[(size_t)xrow * LDP + col] = o[nt][0];
                Pv[(size_t)xrow * LDP + col + 1] = o[nt][1];
            }
        }
        __syncthreads();
#ifdef CHOL_PROFILE_PHASES
        const long long t2 = global_ns();
#endif
        // ---- (C) trailing columns c >= c0 + 32: one m-tile x up to 4 n-tiles per pass (independent DMMA chains) ----
        const int nb0 = (c0 + PB) / 8, nbn = NB / 8 - nb0;
        const int ngrp = (nbn + 3) / 4;
        for (int idx = warp; idx < (NB / 8) * ngrp; idx += 8) {
            const int am = idx / ngrp, bn0 = nb0 + (idx % ngrp) * 4;
            const bool erow = 8 * am < c0 + PB;
            if (!erow && am < bn0) continue;                 // every n-tile of the group lies right of the diagonal
            double af[PB / 4];
#pragma unroll
            for (int ks = 0; ks < PB / 4; ++ks) af[ks] = Pv[(size_t)(am * 8 + g) * LDP + ks * 4 + tig];
            double o[4][2];
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q][0] = o[q][1] = 0.0;
#pragma unroll
            for (int ks = 0; ks < PB / 4; ++ks)
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int bn = min(bn0 + q, NB / 8 - 1);
                    dmma884(o[q][0], o[q][1], af[ks], Pv[(size_t)(bn * 8 + g) * LDP + ks * 4 + tig]);
                }
            const int x = am * 8 + g;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int bn = bn0 + q;
                if (bn >= NB / 8) break;
                const int c = bn * 8 + 2 * tig;
                if (erow || x >= c) Ts[(size_t)x * LDT + c] -= o[q][0];
                if (erow || x >= c + 1) Ts[(size_t)x * LDT + c + 1] -= o[q][1];
            }
        }
        __syncthreads();
#ifdef CHOL_PROFILE_PHASES
        const long long t3 = global_ns();
        tA += t1 - t0; tB += t2 - t1; tC += t3 - t2;
#endif
    }
#ifdef CHOL_PROFILE_PHASES
    if (tid == 0) {
        atomicAdd((unsigned long long *)&g_potrf_phase[0], (unsigned long long)tA);
        atomicAdd((unsigned long long *)&g_potrf_phase[1], (unsigned long long)tB);
        atomicAdd((unsigned long long *)&g_potrf_phase[2], (unsigned long long)tC);
        atomicAdd((unsigned long long *)&g_potrf_phase[3], 1ull);
    }
#endif
    return s_bad != 0;
}

// -------------------------------------------------------------------------------------------------------------
// The dataflow kernel
// -------------------------------------------------------------------------------------------------------------
// one k-slice (BK columns) of an NB-row operand tile into shared memory: NB x BK/2 16-byte segments
template <int NB>
__device__ __forceinline__ void load_slice(double *__restrict__ dst, const double *__restrict__ src, int ld, int tid) {
    constexpr int SEGS = BK / 2;
#pragma unroll
    for (int r = 0; r < NB * SEGS / 256; ++r) {
        const int idx = tid + 256 * r;
        const int row = idx / SEGS, seg = idx % SEGS;
        cp_async16(dst + (size_t)row * SPAD + seg * 2, src + (size_t)row * ld + seg * 2);
    }
}

// acc += A_s (NB x BK, stride lda_s) * B_s^T (NB x BK, stride SPAD): warp (wm, wn) owns an (NB/2) x (NB/4) sub-tile
template <int MI, int NJ>
__device__ __forceinline__ void mma_slice(double (&acc)[MI][NJ][2], const double *__restrict__ a_s, int lda_s,
                                          const double *__restrict__ b_s) {
#pragma unroll
    for (int ks = 0; ks < BK / 4; ++ks) {
        double a[MI], b[NJ];
#pragma unroll
        for (int i = 0; i < MI; ++i) a[i] = a_s[(size_t)i * 8 * lda_s + ks * 4];
#pragma unroll
        for (int j = 0; j < NJ; ++j) b[j] = b_s[(size_t)j * 8 * SPAD + ks * 4];
#pragma unroll
        for (int i = 0; i < MI; ++i)
#pragma unroll
            for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
}

// 64-row tiles: TWO CTAs per SM (110 KB of shared memory and <= 128 registers each).  A tile-sparse factorisation has far more
// tasks in flight than arithmetic (each holds its CTA while it waits for operands), so the leaf levels were limited by the number
// of CTAs, not by the SMs.  128-row tiles (dense systems, flop-bound) keep one CTA per SM with the full register file.
template <int NB>
__global__ void __launch_bounds__(256, NB == 64 ? 2 : 1) chol_dataflow_kernel(CholArgs p) {
    using Cfg = ChCfg<NB>;
    constexpr int KSL = Cfg::KSL, LDT = Cfg::LDT, MI = Cfg::MI, NJ = Cfg::NJ, STAGE_DOUBLES = Cfg::STAGE_DOUBLES, TILE_DOUBLES = Cfg::TILE_DOUBLES;
    constexpr int WR = NB / 2, WC = NB / 4;                // warp tile
    if (p.state && p.state->done) return;
    extern __shared__ __align__(16) double smem[];
    double *tileC = smem;                                  // [NB][LDT]   (main loop: part of the ring)
    double *scratch = smem + TILE_DOUBLES;                 // epilogue ring (2 stages of B) / potrf scratch
    double *Pv = scratch, *invd = scratch + NB * LDP, *dinv = invd + PB * LDP;
    // chain fusion: a diagonal task's own tile A_ii [NB][LDT], staged behind the TRSM's scratch while the ring is idle
    // (consumed before potrf, whose work arrays reach into this area)
    double *own = scratch + Cfg::SCRATCH_DOUBLES;
    static_assert(NB != 64 || TILE_DOUBLES + Cfg::SCRATCH_DOUBLES + TILE_DOUBLES <= Cfg::SMEM_DOUBLES, "own-tile stage must fit the idle ring");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ unsigned long long s_piv[NB];               // potrf: one mbarrier per pivot column (warp 0 -> warp 1 hand-off)
    if (tid < NB) mbar_init(&s_piv[tid], 1);
    unsigned potrf_calls = 0;                              // every call completes one phase of each of them
    // which of the SM's co-resident CTAs this is (64-row tiles run two per SM): the second one runs potrf's two serial sweeps on
    // warps 2 / 3, i.e. on the SM sub-partitions the first one leaves alone
    __shared__ int s_slot;
    if (tid == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;\n" : "=r"(smid));
        s_slot = NB == 64 ? (atomicAdd(p.sm_slot + (smid & 255u), 1) & 1) : 0;
    }
    __syncthreads();
    const int sweep_warp = 2 * s_slot;
    // 2 x 4 warps -> (NB/2) x (NB/4) per warp.  Warps w and w + 4 share an SM sub-partition (and its DMMA pipe): the column blocks are
    // mirrored in the lower warp row so that triangular work (TRSM against the lower-triangular L_jj^-1) is balanced.
    const int wm = warp >> 2, wn = wm ? 3 - (warp & 3) : (warp & 3);
    const int g = lane >> 2, tig = lane & 3;
    const int T = p.T;
    constexpr int ld = NB;                                 // row stride inside a packed tile
    auto tile_ptr = [&](int ti, int tj) -> double * { return p.A + __ldg(p.toff + (size_t)ti * T + tj); };
    __shared__ int s_have;
    bool aborted = false;
    long long waited = 0, t_potrf = 0, t_trsm = 0, n_tasks = 0;
    const long long t_start = p.prof ? global_ns() : 0;

    // ================= phase 1: factorisation + forward substitution =================
    // Tasks are handed out DYNAMICALLY in plan order (a topological order, so whoever holds the lowest unfinished task can always
    // finish it: no deadlock): the first gridDim.x by CTA index, every later one through an atomic ticket.  With a static
    // round-robin a CTA that sat waiting for the operands of one task also held back the tasks queued behind it, which showed as
    // late pick-ups of critical-path tasks in the leaf levels of a tile-sparse system (tools/chol_chain).
    __shared__ int s_next;
    auto next_task = [&]() -> int {
        __syncthreads();
        if (tid == 0) s_next = atomicAdd(p.next_task, 1) + (int)gridDim.x;
        __syncthreads();
        return s_next;
    };
    for (int t = blockIdx.x; t < p.ntasks && !aborted; t = next_task()) {
        const int4 tk = __ldg(p.task + t);
        const int i = tk.x, j = tk.y, kbeg = tk.w;       // klist[kbeg ..): k-blocks with L_ik and L_jk both filled
        const int nkb = __ldg(&p.task[t + 1].w) - tk.w;
        const bool diag = (i == j);
        if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t] = global_ns();
        double acc[MI][NJ][2];
#pragma unroll
        for (int a = 0; a < MI; ++a)
#pragma unroll
            for (int b = 0; b < NJ; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
        // ---- chain fusion (64-row tiles) ----
        // On a tile-sparse system the kernel IS the chain  potrf(c) -> trsm(i, c) -> last k-block of (i, i) -> potrf(i)  along the
        // separators, and between two potrf's the data crossed the chip twice: the trsm task waited for L_cc^-1, published L_ic
        // (global stores, fence, flag), and only then did the diagonal task find the flag and fetch the tile.  Instead, the
        // off-diagonal task whose column c is the diagonal task's LAST k-block (plan: partner) hands over its tile BEFORE the
        // TRSM (C2 = A_ic - sum_k L_ik L_ck^T, ready one link earlier); the diagonal task runs the TRSM itself as soon as
        // L_cc^-1 is published, applies L_ic to its own tile straight from shared memory and goes into potrf(i).  Both tasks
        // run the same TRSM code on the same C2, so the tile the trsm task publishes for everybody else is bit-identical.
        constexpr bool FUSE = (NB == 64);                  // (128-row tiles: no registers for the second accumulator set)
        const bool fused = FUSE && diag && nkb > 0;
        const bool hands_over = FUSE && !diag && i < T && p.partner[t] != 0;
        const int fc = fused ? __ldg(p.klist + kbeg + nkb - 1) : -1;
        const int nkb_main = fused ? nkb - 1 : nkb;        // the fused k-block is applied from shared memory after the loop

        const int nk = nkb_main * KSL;                     // k-slices of the bulk update
        int have = 0;                                      // leading klist entries known to be final in both row panels
        // L_ik and L_jk final for klist entry q.  Called at uniform points by ALL threads; only warp 0 touches the flags:
        // lane l probes the flag of entry q + l / 2 (row panel i or j) with a relaxed load, one ballot finds how many
        // leading entries are complete, one fence turns the relaxed loads into acquires.  Only if entry q itself is not
        // there yet does lane 0 spin.  The verdict travels through shared memory + a barrier; the operand loads that follow
        // are cp.async.cg (L2), so the CTA barrier is enough to order them after the acquire.
        auto ensure = [&](int q) -> bool {
            if (q < have) return true;
            if (warp == 0) {
                const int e = q + (lane >> 1);
                int ok = 1;
                if (e < nkb_main && !(diag && (lane & 1))) {
                    const int kb = __ldg(p.klist + kbeg + e);
                    ok = ld_relaxed(p.tiledone + (size_t)((lane & 1) ? j : i) * T + kb);
                }
                const unsigned m = __ballot_sync(0xffffffffu, ok == 0);
                int h = m ? (__ffs(m) - 1) >> 1 : 16;      // leading complete entries
                h = min(q + h, nkb_main);
                if (h == q) {                              // entry q itself is pending: wait for it
                    if (lane == 0) {
                        const int kb = __ldg(p.klist + kbeg + q);
                        const bool fine = wait_ge(p.tiledone + (size_t)i * T + kb, 1, p.abort_flag, waited) &&
                                          (diag || wait_ge(p.tiledone + (size_t)j * T + kb, 1, p.abort_flag, waited));
                        s_have = fine ? q + 1 : -1;
                    }
                } else {
                    __threadfence();                       // relaxed loads above + fence = acquire
                    if (lane == 0) s_have = h;
                }
            }
            __syncthreads();
            const int h = s_have;
            if (h < 0) return false;
            have = h;
            return true;
        };
        auto issue = [&](int kt) {
            double *st = smem + (size_t)(kt % NSTAGE) * STAGE_DOUBLES;
            const int kb = __ldg(p.klist + kbeg + kt / KSL);
            const int col = (kt % KSL) * BK;
            load_slice<NB>(st, tile_ptr(i, kb) + col, ld, tid);
            if (!diag) load_slice<NB>(st + NB * SPAD, tile_ptr(j, kb) + col, ld, tid);
        };
        // prologue
        for (int s = 0; s < NSTAGE - 1; ++s) {
            if (s < nk) {
                if ((s % KSL) == 0 && !ensure(s / KSL)) { aborted = true; }
                if (!aborted) issue(s);
            }
            cp_async_commit();
        }
        for (int kt = 0; kt < nk; ++kt) {
            cp_async_wait<NSTAGE - 2>();
            __syncthreads();
            const int nx = kt + NSTAGE - 1;
            if (nx < nk && !aborted) {
                if ((nx % KSL) == 0 && !ensure(nx / KSL)) aborted = true;
                if (!aborted) issue(nx);
            }
            cp_async_commit();
            const double *st = smem + (size_t)(kt % NSTAGE) * STAGE_DOUBLES;
            const double *a_s = st + (size_t)(wm * WR + g) * SPAD + tig;
            const double *b_s = (diag ? st : st + NB * SPAD) + (size_t)(wn * WC + g) * SPAD + tig;
            mma_slice<MI, NJ>(acc, a_s, SPAD, b_s);
        }
        cp_async_wait<0>();
        aborted = __syncthreads_or(aborted);
        if (aborted) break;
        if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t + 1] = global_ns();

        // r = (tile staged in tileC) * L_cc^-T : A operand = the staged tile, B operand = invL of column jc through a 2-stage ring
        // in the scratch.  Ends with a CTA barrier (tileC and the scratch are free again).
        auto trsm_mul = [&](double (&r)[MI][NJ][2], int jc) {
            const double *Ig = p.invL + (size_t)jc * NB * NB;
#pragma unroll
            for (int a = 0; a < MI; ++a)
#pragma unroll
                for (int b = 0; b < NJ; ++b) r[a][b][0] = r[a][b][1] = 0.0;
            if constexpr (NB / BK <= 2) {                  // both k-slices fit the scratch: one round trip to L2 instead of two
#pragma unroll
                for (int kt = 0; kt < NB / BK; ++kt) load_slice<NB>(scratch + (size_t)kt * NB * SPAD, Ig + (size_t)kt * BK, NB, tid);
                cp_async_commit();
                cp_async_wait<0>();
                __syncthreads();
#pragma unroll
                for (int kt = 0; kt < NB / BK; ++kt) {
                    const double *a_s = tileC + (size_t)(wm * WR + g) * LDT + kt * BK + tig;
                    const double *b_s = scratch + (size_t)kt * NB * SPAD + (size_t)(wn * WC + g) * SPAD + tig;
                    // L_cc^-1 is lower triangular: k-slices entirely right of this warp's columns are zero
                    if (kt * BK <= wn * WC + WC - 1) mma_slice<MI, NJ>(r, a_s, LDT, b_s);
                }
                __syncthreads();
            } else {
                load_slice<NB>(scratch, Ig, NB, tid);
                cp_async_commit();
                for (int kt = 0; kt < NB / BK; ++kt) {
                    if (kt + 1 < NB / BK) load_slice<NB>(scratch + (size_t)((kt + 1) & 1) * NB * SPAD, Ig + (size_t)(kt + 1) * BK, NB, tid);
                    cp_async_commit();
                    cp_async_wait<1>();
                    __syncthreads();
                    const double *a_s = tileC + (size_t)(wm * WR + g) * LDT + kt * BK + tig;
                    const double *b_s = scratch + (size_t)(kt & 1) * NB * SPAD + (size_t)(wn * WC + g) * SPAD + tig;
                    if (kt * BK <= wn * WC + WC - 1) mma_slice<MI, NJ>(r, a_s, LDT, b_s);
                    __syncthreads();
                }
            }
        };
        if constexpr (FUSE) {
            if (fused) {
                // the own tile is needed right after the TRSM: it travels into shared memory (behind the TRSM's scratch; the ring is
                // idle now) while the task waits for the hand-over and for potrf(c)
                {
                    const double *Aii = tile_ptr(i, i);
                    for (int idx = tid; idx < NB * NB / 2; idx += 256) {
                        const int r = idx / (NB / 2), c2 = (idx % (NB / 2)) * 2;
                        cp_async16(own + (size_t)r * LDT + c2, Aii + (size_t)r * ld + c2);
                    }
                    cp_async_commit();
                }
                if (tid == 0 && !wait_ge(p.c2done + i, 1, p.abort_flag, waited)) aborted = true;
                aborted = __syncthreads_or(aborted);
                if (aborted) break;
                const double *C2 = p.c2buf + (size_t)i * NB * NB;
                for (int idx = tid; idx < NB * NB / 2; idx += 256) {
                    const int r = idx / (NB / 2), c2 = (idx % (NB / 2)) * 2;
                    *reinterpret_cast<double2 *>(tileC + (size_t)r * LDT + c2) = __ldcg(reinterpret_cast<const double2 *>(C2 + (size_t)r * NB + c2));
                }
                if (tid == 0 && !wait_ge(p.tiledone + (size_t)fc * T + fc, 1, p.abort_flag, waited)) aborted = true;
                aborted = __syncthreads_or(aborted);       // (also: tileC is complete)
                if (aborted) break;
                if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t + 1] = global_ns();
                double r[MI][NJ][2];
                trsm_mul(r, fc);                           // r = L_ic
                // L_ic as KSL k-slices in the scratch, then the diagonal tile's last k-block straight from shared memory
#pragma unroll
                for (int a = 0; a < MI; ++a) {
                    const int row = wm * WR + a * 8 + g;
#pragma unroll
                    for (int b = 0; b < NJ; ++b) {
                        const int col = wn * WC + b * 8 + 2 * tig;
                        *reinterpret_cast<double2 *>(scratch + (size_t)(col / BK) * NB * SPAD + (size_t)row * SPAD + (col % BK)) =
                            make_double2(r[a][b][0], r[a][b][1]);
                    }
                }
                __syncthreads();
#pragma unroll
                for (int ks = 0; ks < KSL; ++ks) {
                    const double *sl = scratch + (size_t)ks * NB * SPAD;
                    mma_slice<MI, NJ>(acc, sl + (size_t)(wm * WR + g) * SPAD + tig, SPAD, sl + (size_t)(wn * WC + g) * SPAD + tig);
                }
                __syncthreads();                           // the scratch is potrf's work area next
            }
        }

        // C = A_ij - acc  -> shared tile
        {
            const double *Cg = tile_ptr(i, j);
#pragma unroll
            for (int a = 0; a < MI; ++a) {
                const int row = wm * WR + a * 8 + g;
#pragma unroll
                for (int b = 0; b < NJ; ++b) {
                    const int col = wn * WC + b * 8 + 2 * tig;
                    double2 v;
                    if (fused) v = *reinterpret_cast<const double2 *>(own + (size_t)row * LDT + col);
                    else if (i == T && p.rhs_cam) {        // rhs tile: row 0 = the camera-ordered right-hand side, permuted on the fly
                        v = make_double2(0.0, 0.0);
                        if (row == 0) {
                            const int s0 = __ldg(p.row_src + j * NB + col), s1 = __ldg(p.row_src + j * NB + col + 1);
                            v.x = s0 >= 0 ? __ldcg(p.rhs_cam + s0) : 0.0;
                            v.y = s1 >= 0 ? __ldcg(p.rhs_cam + s1) : 0.0;
                        }
                    } else v = *reinterpret_cast<const double2 *>(Cg + (size_t)row * ld + col);
                    v.x -= acc[a][b][0]; v.y -= acc[a][b][1];
                    if (diag) {                            // strict upper starts as the zero part of [A; I]'s identity
                        if (col > row) v.x = 0.0;
                        if (col + 1 > row) v.y = 0.0;
                    }
                    *reinterpret_cast<double2 *>(tileC + (size_t)row * LDT + col) = v;
                    if (hands_over) __stcg(reinterpret_cast<double2 *>(p.c2buf + (size_t)i * NB * NB + (size_t)row * NB + col), v);
                }
            }
        }
        if (hands_over) {                                  // chain fusion: the row's diagonal task takes the tile from here
            __threadfence();
            __syncthreads();
            if (tid == 0) st_release(p.c2done + i, 1);
        }
        __syncthreads();
        if (diag) {
            const long long tp0 = p.prof ? global_ns() : 0;
            const bool bad = potrf_inverse_tile<NB>(tileC, Pv, invd, dinv, s_piv, potrf_calls++ & 1u, sweep_warp);
            if (p.prof) t_potrf += global_ns() - tp0;
            if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t + 2] = global_ns();
            // L_jj^-1 first: it is all the TRSMs (and the back substitution) wait for; L_jj itself is only stored for the
            // caller and leaves after the flag is up
            double *Ig = p.invL + (size_t)j * NB * NB;
            // (only the lower triangle is written: the work area is zeroed when it is allocated and nothing else ever writes
            // the strict upper part, so half of the 128 KB leaves the SM before the flag can go up)
            for (int idx = tid; idx < NB * NB; idx += 256) {
                const int r = idx / NB, c = idx % NB;
                if (c <= r) Ig[idx] = (c < r) ? tileC[(size_t)c * LDT + r] : dinv[r];
            }
            if (bad && tid == 0) {
                if (p.state) p.state->chol_fail = 1;
                if (p.fail_flag) *p.fail_flag = 1;
            }
            __threadfence();
            __syncthreads();
            if (tid == 0) st_release(p.tiledone + (size_t)j * T + j, 1);
            double *Lg = tile_ptr(j, j);
            for (int idx = tid; idx < NB * NB; idx += 256) {
                const int r = idx / NB, c = idx % NB;
                if (c <= r) Lg[(size_t)r * ld + c] = tileC[(size_t)r * LDT + c];
            }
            __syncthreads();                               // tileC is reused by the next task
        } else {
            // L_ij = C * L_jj^-T : A operand = the staged tile, B operand = invL_j through a 2-stage ring
            if (tid == 0 && !wait_ge(p.tiledone + (size_t)j * T + j, 1, p.abort_flag, waited)) aborted = true;
            aborted = __syncthreads_or(aborted);
            if (aborted) break;
            const long long tt0 = p.prof ? global_ns() : 0;
            if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t + 2] = global_ns();
            trsm_mul(acc, j);
            double *Cg = tile_ptr(i, j);
#pragma unroll
            for (int a = 0; a < MI; ++a) {
                const int row = wm * WR + a * 8 + g;
#pragma unroll
                for (int b = 0; b < NJ; ++b) {
                    const int col = wn * WC + b * 8 + 2 * tig;
                    *reinterpret_cast<double2 *>(Cg + (size_t)row * ld + col) = make_double2(acc[a][b][0], acc[a][b][1]);
                }
            }
            __threadfence();
            __syncthreads();
            if (tid == 0) st_release(p.tiledone + (size_t)i * T + j, 1);
            if (p.prof) t_trsm += global_ns() - tt0;
        }
        if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)t + 3] = global_ns();
        ++n_tasks;
    }
    const long long t_phase1 = p.prof ? global_ns() : 0, waited1 = waited;
    if (aborted) {
        if (tid == 0) {
            if (p.state) p.state->chol_fail = 1;
            if (p.fail_flag) *p.fail_flag = 1;
        }
        return;
    }

    // ================= phase 2: back substitution L^T y = z (z = row 0 of the rhs tiles) =================
    // 256 threads = PARTS groups of NB threads; group q sums rows [q RP, (q + 1) RP) of the tile for column c
    constexpr int PARTS = 256 / NB, RP = NB / PARTS, RC = RP < 32 ? RP : 32;
    double *zs = smem;                 // [NB] running z_m
    double *ys = smem + NB;            // [NB] y_k being applied
    double *part = smem + 2 * NB;      // [PARTS][NB]
    for (int m = T - 1 - (int)blockIdx.x; m >= 0; m -= gridDim.x) {
        // z_m is final once tile (T, m) is
        if (!__syncthreads_and(tid != 0 || (wait_ge(p.tiledone + (size_t)T * T + m, 1, p.abort_flag, waited) && wait_ge(p.tiledone + (size_t)m * T + m, 1, p.abort_flag, waited)))) break;
        if (tid < NB) zs[tid] = __ldcg(tile_ptr(T, m) + tid);      // row 0 of the rhs tile
        const int c = tid % NB, q = tid / NB;
        bool ok = true;
        for (int k = T - 1; k > m; --k) {
            if (!__ldg(p.fill + (size_t)k * T + m)) continue;          // L_km is structurally zero
            ok = tid != 0 || (wait_ge(p.ydone + k, 1, p.abort_flag, waited) && wait_ge(p.tiledone + (size_t)k * T + m, 1, p.abort_flag, waited));
            if (!__syncthreads_and(ok)) { ok = false; break; }
            if (tid < NB) ys[tid] = __ldcg(p.ypad + (size_t)k * NB + tid);
            __syncthreads();
            const double *Lkm = tile_ptr(k, m) + (size_t)(q * RP) * ld + c;
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
            for (int rb = 0; rb < RP; rb += RC) {
                double lv[RC];
#pragma unroll
                for (int r = 0; r < RC; ++r) lv[r] = __ldcg(Lkm + (size_t)(rb + r) * ld);    // RC loads in flight per thread
#pragma unroll
                for (int r = 0; r < RC; r += 4) {
                    const double *yy = ys + q * RP + rb + r;
                    a0 = fma(lv[r], yy[0], a0); a1 = fma(lv[r + 1], yy[1], a1);
                    a2 = fma(lv[r + 2], yy[2], a2); a3 = fma(lv[r + 3], yy[3], a3);
                }
            }
            part[q * NB + c] = (a0 + a1) + (a2 + a3);
            __syncthreads();
            if (tid < NB) {
                double t = 0.0;
#pragma unroll
                for (int qq = 0; qq < PARTS; ++qq) t += part[qq * NB + tid];
                zs[tid] -= t;
            }
        }
        if (!ok) break;
        __syncthreads();
        // y_m[c] = sum_{r >= c} invL_m[r][c] z_m[r]
        const double *Ig = p.invL + (size_t)m * NB * NB;
        double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
#pragma unroll
        for (int rb = 0; rb < RP; rb += RC) {
            double iv[RC];
#pragma unroll
            for (int r = 0; r < RC; ++r) iv[r] = (q * RP + rb + r >= c) ? __ldcg(Ig + (size_t)(q * RP + rb + r) * NB + c) : 0.0;
#pragma unroll
            for (int r = 0; r < RC; r += 4) {
                const double *zz = zs + q * RP + rb + r;
                a0 = fma(iv[r], zz[0], a0); a1 = fma(iv[r + 1], zz[1], a1);
                a2 = fma(iv[r + 2], zz[2], a2); a3 = fma(iv[r + 3], zz[3], a3);
            }
        }
        part[q * NB + c] = (a0 + a1) + (a2 + a3);
        __syncthreads();
        if (tid < NB) {
            double v = 0.0;
#pragma unroll
            for (int qq = 0; qq < PARTS; ++qq) v += part[qq * NB + tid];
            p.ypad[(size_t)m * NB + tid] = v;
            if (m * NB + tid < p.n) {
                p.y[m * NB + tid] = v;
                if (p.y_cam) { const int src = __ldg(p.row_src + m * NB + tid); if (src >= 0) p.y_cam[src] = v; }
            }
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) st_release(p.ydone + m, 1);
        if (p.task_prof && tid == 0) p.task_prof[4 * (size_t)p.ntasks + m] = global_ns();
    }
    if (p.prof && tid == 0) {
        long long *o = p.prof + 8 * (size_t)blockIdx.x;
        o[0] = t_start; o[1] = t_phase1; o[2] = global_ns(); o[3] = waited1; o[4] = waited - waited1; o[5] = t_potrf;
        o[6] = n_tasks; o[7] = t_trsm;
    }
}

// -------------------------------------------------------------------------------------------------------------
// host side
// -------------------------------------------------------------------------------------------------------------
static int g_chol_grid[64];     // per device: CTAs of the persistent grid of the 128-row kernel (0 = not initialised)
static int g_chol_grid64[64];   // ... and of the 64-row kernel (co-resident CTAs per SM from the occupancy query)
long long *g_chol_prof = nullptr;
long long *g_chol_task_prof = nullptr;   // tools/chol_prof.cu only: [ntasks][4] + [T]   // tools/chol_prof.cu only: device buffer [grid][8] of per-CTA timings

int chol_init() {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    if (g_chol_grid[dev]) return 0;
    e = cudaFuncSetAttribute(chol_dataflow_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, ChCfg<128>::SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(chol_dataflow_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, ChCfg<64>::SMEM);
    if (e != cudaSuccess) return (int)e;
    int coop = 0, sms = 0, per_sm = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (!coop) return (int)cudaErrorNotSupported;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_dataflow_kernel<128>, 256, ChCfg<128>::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (per_sm < 1) return (int)cudaErrorLaunchOutOfResources;
    int per_sm64 = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm64, chol_dataflow_kernel<64>, 256, ChCfg<64>::SMEM);
    if (e != cudaSuccess) return (int)e;
    if (per_sm64 < 1) return (int)cudaErrorLaunchOutOfResources;
    g_chol_grid64[dev] = sms * std::min(per_sm64, 2);
    g_chol_grid[dev] = sms;       // one CTA per SM (shared memory allows exactly one)
    return 0;
}

// ---- elimination order of the tiles -----------------------------------------------------------------------------
// Nested dissection on the tile graph (George's automatic scheme: the middle level of a BFS level structure rooted at a
// pseudo-peripheral node is a separator).  Parts on the two sides of a separator share no tile, so their columns
// factorise concurrently and the dependency chain shrinks from T tiles to roughly leaf + sum of the separators above
// it.  Graphs without a usable level structure (dense co-visibility) come back in BFS order, which is harmless.
namespace {
struct TileGraph {
    int G;
    const unsigned char *adj;
    bool edge(int a, int b) const { return a != b && adj[(size_t)std::max(a, b) * G + std::min(a, b)] != 0; }
};

// BFS over the nodes with in[v] == tag; fills lvl (levels) and out (visit order); returns the number of levels
int bfs(const TileGraph &g, const std::vector<int> &in, int tag, int root, std::vector<int> &lvl, std::vector<int> &out) {
    out.clear();
    out.push_back(root);
    lvl[root] = 0;
    int nlev = 1;
    for (size_t h = 0; h < out.size(); ++h) {
        const int v = out[h];
        for (int w = 0; w < g.G; ++w)
            if (in[w] == tag && lvl[w] < 0 && g.edge(v, w)) { lvl[w] = lvl[v] + 1; nlev = std::max(nlev, lvl[w] + 1); out.push_back(w); }
    }
    return nlev;
}

void dissect(const TileGraph &g, std::vector<int> nodes, std::vector<int> &in, int &next_tag, std::vector<int> &order) {
    const int tag = next_tag++;
    for (int v : nodes) in[v] = tag;
    std::vector<int> lvl(g.G, -1), comp, tmp;
    for (int seed : nodes) {
        if (lvl[seed] >= 0 || in[seed] != tag) continue;
        // connected component of seed, then a pseudo-peripheral root (two sweeps towards a far, low-degree node)
        int nlev = bfs(g, in, tag, seed, lvl, comp);
        for (int sweep = 0; sweep < 3; ++sweep) {
            int best = -1, bestdeg = 1 << 30;
            for (int v : comp) {
                if (lvl[v] != nlev - 1) continue;
                int deg = 0;
                for (int w : comp) deg += g.edge(v, w);
                if (deg < bestdeg) { bestdeg = deg; best = v; }
            }
            for (int v : comp) lvl[v] = -1;
            const int nl2 = bfs(g, in, tag, best, lvl, tmp);
            const bool deeper = nl2 > nlev;
            nlev = nl2; comp = tmp;
            if (!deeper) break;
        }
        // separator = the level that balances the two sides best
        std::vector<int> cnt(nlev, 0);
        for (int v : comp) cnt[lvl[v]]++;
        int m = -1, best_cost = 1 << 30, before = 0;
        const int total = (int)comp.size();
        for (int l = 0; l < nlev; ++l) {
            if (l >= 1 && l <= nlev - 2) {
                const int after = total - before - cnt[l];
                const int cost = std::max(before, after) * 4 + cnt[l];
                if (cost < best_cost) { best_cost = cost; m = l; }
            }
            before += cnt[l];
        }
        const bool leaf = m < 0 || total <= cnt[m] + 1 || 2 * cnt[m] > total;
        if (leaf) {
            for (int v : comp) { order.push_back(v); in[v] = -1; }     // BFS order keeps the leaf banded
            continue;
        }
        std::vector<int> rest, sep;
        for (int v : comp) (lvl[v] == m ? sep : rest).push_back(v);
        for (int v : comp) in[v] = -1;                                  // leave this level's tag (lvl stays set: visited)
        dissect(g, rest, in, next_tag, order);
        for (int v : sep) order.push_back(v);
    }
}
} // namespace

void chol_tile_order(int G, const unsigned char *adj, std::vector<int> &order) {
    order.clear();
    order.reserve(G);
    TileGraph g{G, adj};
    std::vector<int> nodes(G), in(G, -1);
    for (int v = 0; v < G; ++v) nodes[v] = v;
    int next_tag = 0;
    dissect(g, nodes, in, next_tag, order);
}

// ---- tile-sparsity plan ------------------------------------------------------------------------------------------
// The reduced camera matrix of a SLAM / SfM map is block-banded when cameras only share landmarks with their
// neighbours.  A tile of L can only be nonzero if it is nonzero in S or filled by the elimination of an earlier column
// (tile-level symbolic factorisation); everything else is skipped.  The arithmetic on the filled tiles is exactly the
// dense one (the skipped terms are exact zeros), so results do not depend on the plan.
namespace {
struct HostPlan {
    std::vector<unsigned char> blob;
    std::vector<long long> toff;          // [(T + 1) * T]
    long long front_doubles = 0, extra_doubles = 0, total_doubles = 0;
    size_t o_task = 0, o_kl = 0, o_fill = 0, o_toff = 0, o_partner = 0;
    int T = 0, ntasks = 0, depth = 0;
    long long filled = 0;
    double flops = 0.0;
};

void plan_analyse(int n, int NB, const unsigned char *nz, HostPlan &hp, size_t extra_doubles = 0) {
    const int T = chol_ld(n, NB) / NB;
    std::vector<unsigned char> f((size_t)(T + 1) * T, 0);
    for (int i = 0; i < T; ++i)
        for (int j = 0; j <= i; ++j) f[(size_t)i * T + j] = (i == j) || !nz || nz[(size_t)i * T + j] || nz[(size_t)j * T + i];
    // storage: [tiles S can touch | extra (caller's rhs) | fill-in tiles | rhs tiles]
    hp.toff.assign((size_t)(T + 1) * T, -1);
    long long off = 0;
    for (int i = 0; i < T; ++i)
        for (int j = 0; j <= i; ++j) if (f[(size_t)i * T + j]) { hp.toff[(size_t)i * T + j] = off; off += (long long)NB * NB; }
    hp.front_doubles = off;
    hp.extra_doubles = (long long)((extra_doubles + 1) & ~(size_t)1);     // keeps every tile 16-byte aligned
    off += hp.extra_doubles;
    std::vector<int> rows;
    for (int j = 0; j < T; ++j) {                       // eliminating column j couples all its rows pairwise
        rows.clear();
        for (int i = j + 1; i < T; ++i) if (f[(size_t)i * T + j]) rows.push_back(i);
        for (size_t a = 0; a < rows.size(); ++a)
            for (size_t b = a + 1; b < rows.size(); ++b) f[(size_t)rows[b] * T + rows[a]] = 1;
    }
    for (int j = 0; j < T; ++j) f[(size_t)T * T + j] = 1;   // the rhs row
    for (int i = 0; i <= T; ++i)
        for (int j = 0; j < T && j <= i; ++j)
            if (f[(size_t)i * T + j] && hp.toff[(size_t)i * T + j] < 0) { hp.toff[(size_t)i * T + j] = off; off += (long long)NB * NB; }
    hp.total_doubles = off;
    // tasks = filled tiles; dependency level = longest chain of tasks below it.  Tasks are handed out in level order
    // (a topological order, so the round-robin dataflow cannot deadlock): independent sub-chains of a sparse pattern
    // run side by side.  For a dense matrix this is exactly the column-major order.
    struct HostTask { int i, j, level, kbeg, kend; };
    std::vector<HostTask> ht;
    std::vector<int> kl_tmp, level((size_t)(T + 1) * T, 0);
    double flops = 0.0;
    long long filled = 0;
    const double t3 = (double)NB * NB * NB;
    for (int j = 0; j < T; ++j)
        for (int i = j; i <= T; ++i) {
            if (!f[(size_t)i * T + j]) continue;
            const int kbeg = (int)kl_tmp.size();
            int lv = 0;
            for (int k = 0; k < j; ++k)
                if (f[(size_t)i * T + k] && f[(size_t)j * T + k]) {
                    kl_tmp.push_back(k);
                    lv = std::max(lv, 1 + std::max(level[(size_t)i * T + k], level[(size_t)j * T + k]));
                }
            if (i != j) lv = std::max(lv, 1 + level[(size_t)j * T + j]);
            level[(size_t)i * T + j] = lv;
            // consume the k-blocks in the order their operands become available (dependency level, then k): a separator
            // tile with many contributions then only waits for the last ones.  Fixed per plan, so sums stay reproducible.
            std::stable_sort(kl_tmp.begin() + kbeg, kl_tmp.end(), [&](int a, int b) {
                return std::max(level[(size_t)i * T + a], level[(size_t)j * T + a]) < std::max(level[(size_t)i * T + b], level[(size_t)j * T + b]);
            });
            const double nkb = (double)((int)kl_tmp.size() - kbeg);
            if (i == T) flops += 2.0 * NB * NB * nkb + (double)NB * NB;             // one rhs row block
            else if (i == j) { flops += t3 * nkb + t3 / 3.0; ++filled; }
            else { flops += 2.0 * t3 * nkb + t3; ++filled; }
            ht.push_back({i, j, lv, kbeg, (int)kl_tmp.size()});
        }
    std::stable_sort(ht.begin(), ht.end(), [](const HostTask &a, const HostTask &b) { return a.level < b.level; });
    std::vector<int4> task;
    std::vector<int> klist;
    klist.reserve(kl_tmp.size());
    int depth = 0;
    for (const HostTask &h : ht) {
        task.push_back(make_int4(h.i, h.j, h.level, (int)klist.size()));
        klist.insert(klist.end(), kl_tmp.begin() + h.kbeg, kl_tmp.begin() + h.kend);
        depth = std::max(depth, h.level + 1);
    }
    task.push_back(make_int4(0, 0, 0, (int)klist.size()));
    flops += 2.0 * (double)filled * NB * NB;                                        // back substitution
    const size_t o_task = 0, o_kl = o_task + sizeof(int4) * task.size();
    const size_t o_fill = o_kl + sizeof(int) * std::max<size_t>(klist.size(), 1);
    const size_t o_toff = (o_fill + f.size() + 15) & ~(size_t)15;
    const size_t o_partner = o_toff + sizeof(long long) * hp.toff.size();
    const size_t bytes = o_partner + task.size();
    std::vector<unsigned char> &blob = hp.blob;
    blob.assign(bytes, 0);
    // chain fusion: the off-diagonal task (i, c) whose column is the LAST k-block of diagonal task (i, i) hands its
    // pre-TRSM tile to that task
    {
        std::vector<int> at((size_t)(T + 1) * T, -1);
        for (size_t q = 0; q + 1 < task.size(); ++q) at[(size_t)task[q].x * T + task[q].y] = (int)q;
        for (size_t q = 0; q + 1 < task.size(); ++q) {
            const int nkb = task[q + 1].w - task[q].w;
            if (task[q].x != task[q].y || nkb <= 0) continue;
            const int c = klist[(size_t)task[q].w + nkb - 1];
            blob[o_partner + (size_t)at[(size_t)task[q].x * T + c]] = 1;
        }
    }
    hp.o_partner = o_partner;
    std::memcpy(blob.data() + o_toff, hp.toff.data(), sizeof(long long) * hp.toff.size());
    hp.o_toff = o_toff;
    std::memcpy(blob.data() + o_task, task.data(), sizeof(int4) * task.size());
    if (!klist.empty()) std::memcpy(blob.data() + o_kl, klist.data(), sizeof(int) * klist.size());
    std::memcpy(blob.data() + o_fill, f.data(), f.size());
    hp.o_task = o_task; hp.o_kl = o_kl; hp.o_fill = o_fill;
    hp.T = T; hp.ntasks = (int)task.size() - 1; hp.depth = depth; hp.filled = filled; hp.flops = flops;
}
} // namespace

// host-only: dependency depth (tasks) and filled tiles of the plan for a tile pattern
void chol_plan_stats(int n, int nb, const unsigned char *nz, int *depth, long long *filled, double *flops) {
    HostPlan hp;
    plan_analyse(n, nb, nz, hp);
    if (depth) *depth = hp.depth;
    if (filled) *filled = hp.filled;
    if (flops) *flops = hp.flops;
}

int chol_plan_create(int n, int nb, const unsigned char *nz, CholPlan *pl, cudaStream_t s, size_t extra_doubles) {
    if (nb != 64 && nb != 128) return (int)cudaErrorInvalidValue;
    HostPlan hp;
    plan_analyse(n, nb, nz, hp, extra_doubles);
    const size_t bytes = hp.blob.size();
    if (pl->cap < bytes) {
        if (pl->blob) cudaFree(pl->blob);
        pl->blob = nullptr; pl->cap = 0;
        cudaError_t e = cudaMalloc(&pl->blob, bytes);
        if (e != cudaSuccess) return (int)e;
        pl->cap = bytes;
    }
    cudaError_t e = cudaMemcpyAsync(pl->blob, hp.blob.data(), bytes, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return (int)e;
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return (int)e;
    unsigned char *b = static_cast<unsigned char *>(pl->blob);
    pl->n = n; pl->nb = nb; pl->T = hp.T; pl->ntasks = hp.ntasks;
    pl->task = reinterpret_cast<const int4 *>(b + hp.o_task);
    pl->klist = reinterpret_cast<const int *>(b + hp.o_kl);
    pl->fill = b + hp.o_fill;
    pl->tile_off = reinterpret_cast<const long long *>(b + hp.o_toff);
    pl->partner = b + hp.o_partner;
    pl->h_tile_off = hp.toff;
    pl->front_doubles = hp.front_doubles; pl->extra_doubles = hp.extra_doubles; pl->total_doubles = hp.total_doubles;
    pl->filled_tiles = hp.filled;
    pl->depth = hp.depth;
    pl->flops = hp.flops;
    return 0;
}

// zero the whole packed matrix (every tile of the plan + the caller's extra region) before S is assembled
int chol_clear(double *A, const CholPlan *plan, cudaStream_t s) {
    return (int)cudaMemsetAsync(A, 0, sizeof(double) * (size_t)plan->total_doubles, s);
}

void chol_plan_destroy(CholPlan *pl) {
    if (pl->blob) cudaFree(pl->blob);
    *pl = CholPlan();
}

// host matrix (n x n row-major, lower triangle read) + rhs -> the plan's tile-packed storage (host vector of total_doubles)
void chol_pack_host(const CholPlan *pl, const double *A, const double *b, std::vector<double> &out) {
    const int n = pl->n, T = pl->T, NB = pl->nb;
    out.assign((size_t)pl->total_doubles, 0.0);
    for (int ti = 0; ti < T; ++ti)
        for (int tj = 0; tj <= ti; ++tj) {
            const long long o = pl->h_tile_off[(size_t)ti * T + tj];
            if (o < 0) continue;
            double *t = out.data() + o;
            for (int r = 0; r < NB; ++r) {
                const int gi = ti * NB + r;
                for (int c = 0; c < NB; ++c) {
                    const int gj = tj * NB + c;
                    if (gj > gi) continue;
                    t[(size_t)r * NB + c] = (gi < n && gj < n) ? A[(size_t)gi * n + gj] : (gi == gj ? 1.0 : 0.0);
                }
            }
        }
    for (int tj = 0; tj < T; ++tj) {
        double *t = out.data() + pl->h_tile_off[(size_t)T * T + tj];
        for (int c = 0; c < NB; ++c) { const int gj = tj * NB + c; t[c] = gj < n ? b[gj] : 0.0; }
    }
}

int chol_factor_solve(double *A, int n, double *work, double *y, LMState *state, int *fail_flag, const CholPlan *plan, cudaStream_t s,
                      const CholPerm *perm) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    if (!plan || plan->n != n) return (int)cudaErrorInvalidValue;       // the plan defines the storage (and the tile edge)
    const int NB = plan->nb, T = plan->T;
    CholArgs p;
    p.A = A; p.toff = plan->tile_off; p.T = T; p.n = n;
    p.invL = work;
    p.ypad = work + (size_t)T * NB * NB;
    p.c2buf = p.ypad + (size_t)T * NB;
    int *flags = reinterpret_cast<int *>(p.c2buf + (size_t)T * NB * NB);
    p.tiledone = flags; p.ydone = flags + (size_t)(T + 1) * T; p.c2done = p.ydone + T; p.abort_flag = p.c2done + T; p.next_task = p.abort_flag + 1; p.sm_slot = p.next_task + 1;
    p.partner = plan->partner;
    p.row_src = perm ? perm->row_src : nullptr; p.rhs_cam = perm ? perm->rhs_cam : nullptr; p.y_cam = perm ? perm->y_cam : nullptr;
    p.y = y; p.state = state; p.fail_flag = fail_flag; p.prof = g_chol_prof; p.task_prof = g_chol_task_prof;
    p.task = plan->task; p.klist = plan->klist; p.fill = plan->fill; p.ntasks = plan->ntasks;
    cudaError_t e = cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)chol_flag_count(T), s);
    if (e != cudaSuccess) return (int)e;
    const int cap = NB == 64 ? g_chol_grid64[dev] : g_chol_grid[dev];
    int grid = cap < plan->ntasks ? cap : plan->ntasks;
    if (grid < 1) return (int)cudaErrorInitializationError;
    void *args[] = {&p};
    if (NB == 128) e = cudaLaunchCooperativeKernel((const void *)chol_dataflow_kernel<128>, dim3(grid), dim3(256), args, ChCfg<128>::SMEM, s);
    else e = cudaLaunchCooperativeKernel((const void *)chol_dataflow_kernel<64>, dim3(grid), dim3(256), args, ChCfg<64>::SMEM, s);
    return (int)e;
}

} // namespace ba
