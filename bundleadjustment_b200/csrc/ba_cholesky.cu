// ba_cholesky.cu -- dense blocked Cholesky of the reduced camera system on the FP64 tensor pipe.
//
// Replaces Ceres' DenseSchurComplementSolver::SolveReducedLinearSystem (Eigen LLT) that the reference selects with
// options.linear_solver_type = ceres::DENSE_SCHUR (ba_project/src/ba/Optimizer.cpp:85).
//
// Layout: A is (n+1) x n row-major (ld = n).  Lower triangle holds S; the extra row n holds rhs^T, so the forward
// substitution z = L^-1 rhs falls out of the factorisation itself (row n is just one more row of every panel).
// Right-looking, block size NB = 128:
//   potrf_diag : one CTA factors the NB x NB diagonal block in shared memory and inverts it (L_kk^-1)
//   gemm_nt    : FP64 tensor-core (DMMA m8n8k4) C = A * B^T tiles, 128x128x16, cp.async double-buffered
//                 - MODE_SET : panel TRSM as a GEMM with L_kk^-T     A_ik <- A_ik * L_kk^-T
//                 - MODE_SUB : trailing SYRK update (lower tiles)     A_ij <- A_ij - A_ik * A_jk^T
//   back_solve : one launch per block row for L^T y = z
// FP64 on sm_100a has no tcgen05 kind; mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4) is the tensor-pipe instruction.
#include "ba_launch.h"

namespace ba {

constexpr int NB = CHOL_NB;        // 128
constexpr int BM = 128, BN = 128, BK = 16;
constexpr int SPAD = BK + 4;       // smem row stride in doubles: 160 B == 32 mod 128 -> conflict-free fragments

size_t chol_work_doubles(int n) { return (size_t)((n + NB - 1) / NB) * NB * NB; }

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, int src_bytes) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

__device__ __forceinline__ void dmma884(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

enum { MODE_SET = 0, MODE_SUB = 1 };

// C[m][n] (op)= sum_k A[m][k] * B[n][k];  A: M x K (lda), B: N x K (ldb), C: M x N (ldc); all row-major.
// TRI: blockIdx.x enumerates lower-triangular tile pairs (ti >= tj) of a square tile grid whose tile rows start at
// row_tile0 * 0 ... ; A and B are then the same panel (B tile = rows of the panel at tj).
template <int MODE, bool TRI>
__global__ void __launch_bounds__(256, 1)
gemm_nt_kernel(const double *A, int lda, const double *B, int ldb, double *C,
               int ldc, int M, int N, int K, int tiles_n, const LMState *state) {
    if (state && state->done) return;
    extern __shared__ __align__(16) double smem[];
    double *sA = smem;                       // [2][BM][SPAD]
    double *sB = smem + 2 * BM * SPAD;       // [2][BN][SPAD]
    int tm, tn;
    if (TRI) {
        // linear index -> (tm >= tn)
        const int t = blockIdx.x;
        int r = (int)((sqrt(8.0 * (double)t + 1.0) - 1.0) * 0.5);
        while ((r + 1) * (r + 2) / 2 <= t) ++r;
        while (r * (r + 1) / 2 > t) --r;
        tm = r; tn = t - r * (r + 1) / 2;
    } else {
        tm = blockIdx.x / tiles_n; tn = blockIdx.x % tiles_n;
    }
    const int m0 = tm * BM, n0 = tn * BN;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wm = warp >> 2, wn = warp & 3;          // 2 x 4 warps -> 64 x 32 per warp
    const int g = lane >> 2, tig = lane & 3;

    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    auto load_stage = [&](int stage, int kk) {
        // each operand tile: 128 rows x 16 doubles = 128 x 8 segments of 16 B; 256 threads x 4 each
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int idx = tid + 256 * r;          // 0..1023
            const int row = idx >> 3, seg = idx & 7;
            const int k = kk + seg * 2;
            int kb = (K - k) * 8;
            kb = kb < 0 ? 0 : (kb > 16 ? 16 : kb);
            {
                const int gr = m0 + row;
                const int bytes = (gr < M) ? kb : 0;
                const double *src = A + (size_t)(gr < M ? gr : 0) * lda + (bytes ? k : 0);
                cp_async16(sA + ((size_t)stage * BM + row) * SPAD + seg * 2, src, bytes);
            }
            {
                const int gr = n0 + row;
                const int bytes = (gr < N) ? kb : 0;
                const double *src = B + (size_t)(gr < N ? gr : 0) * ldb + (bytes ? k : 0);
                cp_async16(sB + ((size_t)stage * BN + row) * SPAD + seg * 2, src, bytes);
            }
        }
    };

    const int nk = (K + BK - 1) / BK;
    load_stage(0, 0);
    cp_async_commit();
    for (int kt = 0; kt < nk; ++kt) {
        if (kt + 1 < nk) {
            load_stage((kt + 1) & 1, (kt + 1) * BK);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        const double *a_s = sA + (size_t)(kt & 1) * BM * SPAD + (size_t)(wm * 64 + g) * SPAD + tig;
        const double *b_s = sB + (size_t)(kt & 1) * BN * SPAD + (size_t)(wn * 32 + g) * SPAD + tig;
#pragma unroll
        for (int ks = 0; ks < BK / 4; ++ks) {
            double a[8], b[4];
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = a_s[(size_t)i * 8 * SPAD + ks * 4];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = b_s[(size_t)j * 8 * SPAD + ks * 4];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
        }
        __syncthreads();
    }
    // epilogue
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = m0 + wm * 64 + i * 8 + g;
        if (row >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = n0 + wn * 32 + j * 8 + 2 * tig;
            if (col >= N) continue;
            double *dst = C + (size_t)row * ldc + col;
            if (col + 1 < N) {
                double2 v;
                if (MODE == MODE_SUB) {
                    v = *reinterpret_cast<double2 *>(dst);
                    v.x -= acc[i][j][0]; v.y -= acc[i][j][1];
                } else {
                    v.x = acc[i][j][0]; v.y = acc[i][j][1];
                }
                *reinterpret_cast<double2 *>(dst) = v;
            } else {
                if (MODE == MODE_SUB) dst[0] -= acc[i][j][0]; else dst[0] = acc[i][j][0];
            }
        }
    }
}

// -------------------------------------------------------------------------------------------------------------
// Diagonal block: Cholesky in shared memory + inverse of the factor, one CTA of 256 threads.
// The inverse costs no extra sequential depth: the block is factored as the augmented panel [A_kk ; I], whose
// bottom half turns into L^-T under the very same right-looking updates.  Both halves share one NB x (NB+1) tile:
// L in the lower triangle, E = L^-T in the strict upper triangle (E[r][c], r < c), diag(E) = 1/diag(L) in dinv.
// Blocked by SB = 16 columns: (A) one warp factors + inverts the 16x16 diagonal sub-block, (B) one thread per row
// applies L_d^-T to the sub-panel, (C) all threads apply the rank-16 trailing update with 4x4 register tiles.
// -------------------------------------------------------------------------------------------------------------
constexpr int TP = NB + 1;
constexpr int SB = 16;
constexpr int PVP = SB + 1;

__global__ void __launch_bounds__(256, 1)
potrf_diag_kernel(double *__restrict__ A, int ld, int k0, int nb, double *__restrict__ invL, LMState *state, int *fail_flag) {
    if (state && state->done) return;
    extern __shared__ __align__(16) double T[];       // [NB][TP] + dinv[NB] + Pv[NB][PVP]
    double *dinv = T + NB * TP;
    double *Pv = dinv + NB;
    __shared__ int s_bad;
    const int tid = threadIdx.x;
    if (tid == 0) s_bad = 0;
    for (int idx = tid; idx < NB * NB; idx += 256) {
        const int i = idx / NB, j = idx % NB;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < nb && j < nb && j <= i) v = A[(size_t)(k0 + i) * ld + k0 + j];
        T[i * TP + j] = v;
    }
    __syncthreads();
    for (int j0 = 0; j0 < NB; j0 += SB) {
        // ---- (A) 16x16 diagonal sub-block: potrf + inverse, warp 0 ----
        if (tid < 32) {
            const int i = tid;
            for (int j = 0; j < SB; ++j) {
                double d = T[(j0 + j) * TP + j0 + j];
                const bool bad = !(d > 0.0) || !isfinite(d);
                if (bad) d = 1.0;
                const double sd = sqrt(d), inv = 1.0 / sd;
                __syncwarp();
                if (bad && i == 0) s_bad = 1;
                if (i < SB && i > j) T[(j0 + i) * TP + j0 + j] *= inv;
                if (i == j) { T[(j0 + j) * TP + j0 + j] = sd; dinv[j0 + j] = inv; }
                __syncwarp();
                if (i < SB && i > j) {
                    const double lij = T[(j0 + i) * TP + j0 + j];
                    for (int c = j + 1; c <= i; ++c) T[(j0 + i) * TP + j0 + c] -= lij * T[(j0 + c) * TP + j0 + j];
                }
                __syncwarp();
            }
            if (i < SB) {
                // lane c: column c of X = L_d^-1, stored transposed in the strict upper triangle
                const int c = i;
                const double xc = dinv[j0 + c];
                for (int ii = 1; ii < SB; ++ii) {
                    double acc = 0.0;
                    for (int k = 0; k < ii; ++k) {
                        const double lik = T[(j0 + ii) * TP + j0 + k];
                        if (k >= c) acc += lik * ((k == c) ? xc : T[(j0 + c) * TP + j0 + k]);
                    }
                    if (ii > c) T[(j0 + c) * TP + j0 + ii] = -acc * dinv[j0 + ii];
                }
            }
        }
        __syncthreads();
        // ---- (B) sub-panel rows: x <- x * L_d^-T ; x < j0: rows of E, x >= j0 + SB: rows of L ----
        if (tid < NB) {
            const int x = tid;
            double w[SB];
            if (x < j0 || x >= j0 + SB) {
                double v[SB];
#pragma unroll
                for (int k = 0; k < SB; ++k) v[k] = T[x * TP + j0 + k];
#pragma unroll
                for (int k = 0; k < SB; ++k) {
                    double acc = v[k] * dinv[j0 + k];
#pragma unroll
                    for (int m = 0; m < k; ++m) acc += v[m] * T[(j0 + m) * TP + j0 + k];
                    w[k] = acc;
                }
#pragma unroll
                for (int k = 0; k < SB; ++k) T[x * TP + j0 + k] = w[k];
            } else {
#pragma unroll
                for (int k = 0; k < SB; ++k) w[k] = (x < j0 + k) ? T[x * TP + j0 + k] : ((x == j0 + k) ? dinv[x] : 0.0);
            }
#pragma unroll
            for (int k = 0; k < SB; ++k) Pv[x * PVP + k] = w[k];
        }
        __syncthreads();
        // ---- (C) trailing update of columns c >= j0 + SB: rows x < j0 + SB (E) and x >= c (L) ----
        const int c_begin = j0 + SB;
        const int ntask = 32 * ((NB - c_begin) / 4);
        for (int t = tid; t < ntask; t += 256) {
            const int tx = t & 31, c0 = c_begin + (t >> 5) * 4;
            double acc[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[r][q] = 0.0;
#pragma unroll 4
            for (int k = 0; k < SB; ++k) {
                double a[4], b[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) a[r] = Pv[(tx + 32 * r) * PVP + k];
#pragma unroll
                for (int q = 0; q < 4; ++q) b[q] = Pv[(c0 + q) * PVP + k];
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[r][q] += a[r] * b[q];
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int x = tx + 32 * r;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c = c0 + q;
                    if (x < c_begin || x >= c) T[x * TP + c] -= acc[r][q];
                }
            }
        }
        __syncthreads();
    }
    // write L back (lower triangle of the block)
    for (int idx = tid; idx < nb * nb; idx += 256) {
        const int i = idx / nb, j = idx % nb;
        if (j <= i) A[(size_t)(k0 + i) * ld + k0 + j] = T[i * TP + j];
    }
    // invL: full NB x NB row-major, zeros above the diagonal; X[i][c] = E[c][i]
    for (int idx = tid; idx < NB * NB; idx += 256) {
        const int i = idx / NB, c = idx % NB;
        double v = 0.0;
        if (i == c) v = dinv[c];
        else if (i > c) v = T[c * TP + i];
        invL[idx] = v;
    }
    if (tid == 0 && s_bad) {
        if (state) state->chol_fail = 1;
        if (fail_flag) *fail_flag = 1;
    }
}

// Backward substitution L^T y = z, one launch per block row k (descending).  z lives in row n of A.
// Every CTA recomputes y_k = L_kk^-T z_k (128 x 128 GEMV from the stored inverse), then owns 64 columns of the
// update z[col] -= sum_r L[k0 + r][col] y_k[r], rows split over 4 thread groups.
__global__ void __launch_bounds__(256)
back_solve_kernel(double *__restrict__ A, int ld, int n, int k0, int nb, const double *__restrict__ invL,
                  double *__restrict__ y, const LMState *state) {
    if (state && state->done) return;
    __shared__ double ys[NB];
    __shared__ double part[4][64];
    __shared__ double yp[2][NB];
    double *z = A + (size_t)n * ld;
    const int tid = threadIdx.x;
    {
        // y_k[c] = sum_{r >= c} invL[r][c] * z[k0 + r]; two threads per column (even / odd r)
        const int c = tid & (NB - 1), half = tid >> 7;
        double acc = 0.0;
        if (c < nb)
            for (int r = c + half; r < nb; r += 2) acc += invL[r * NB + c] * z[k0 + r];
        yp[half][c] = acc;
    }
    __syncthreads();
    if (tid < NB) ys[tid] = yp[0][tid] + yp[1][tid];
    __syncthreads();
    if (blockIdx.x == 0 && tid < nb) y[k0 + tid] = ys[tid];
    const int cl = tid & 63, rg = tid >> 6;
    const int col = blockIdx.x * 64 + cl;
    double acc = 0.0;
    if (col < k0) {
        const int r_lo = rg * 32, r_hi = min(nb, r_lo + 32);
        const double *src = A + (size_t)(k0 + r_lo) * ld + col;
#pragma unroll 8
        for (int r = r_lo; r < r_hi; ++r, src += ld) acc += (*src) * ys[r];
    }
    part[rg][cl] = acc;
    __syncthreads();
    if (tid < 64 && col < k0) z[col] -= part[0][tid] + part[1][tid] + part[2][tid] + part[3][tid];
}

static const int GEMM_SMEM = 2 * (BM + BN) * SPAD * (int)sizeof(double);     // 81920
static const int POTRF_SMEM = (NB * TP + NB + NB * PVP) * (int)sizeof(double);   // 150528

int chol_init() {
    cudaError_t e;
    e = cudaFuncSetAttribute(gemm_nt_kernel<MODE_SET, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(gemm_nt_kernel<MODE_SUB, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(potrf_diag_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, POTRF_SMEM);
    return (int)e;
}

void chol_factor_solve(double *A, int n, double *invL, double *y, LMState *state, int *fail_flag, cudaStream_t s) {
    const int M = n + 1;                 // rows including the rhs row
    const int nblk = (n + NB - 1) / NB;
    for (int kb = 0; kb < nblk; ++kb) {
        const int k0 = kb * NB;
        const int nb = (n - k0 < NB) ? (n - k0) : NB;
        double *inv = invL + (size_t)kb * NB * NB;
        potrf_diag_kernel<<<1, 256, POTRF_SMEM, s>>>(A, n, k0, nb, inv, state, fail_flag);
        const int r0 = k0 + nb;          // first row below the diagonal block
        const int rows = M - r0;         // >= 1 (the rhs row)
        // panel: A[r0.., k0..k0+nb) <- A * invL^T
        {
            const int tiles_m = (rows + BM - 1) / BM;
            gemm_nt_kernel<MODE_SET, false><<<tiles_m, 256, GEMM_SMEM, s>>>(
                A + (size_t)r0 * n + k0, n, inv, NB, A + (size_t)r0 * n + k0, n, rows, nb, nb, 1, state);
        }
        // trailing update: lower tiles of A[r0.., r0..n) -= P P^T with P = A[r0.., k0..k0+nb)
        const int cols = n - r0;
        if (cols > 0) {
            const int t = (rows + BM - 1) / BM;            // tile rows (rows >= cols, same tile origin)
            const int tiles = t * (t + 1) / 2;
            gemm_nt_kernel<MODE_SUB, true><<<tiles, 256, GEMM_SMEM, s>>>(
                A + (size_t)r0 * n + k0, n, A + (size_t)r0 * n + k0, n, A + (size_t)r0 * n + r0, n, rows, cols, nb, 0, state);
        }
    }
    for (int kb = nblk - 1; kb >= 0; --kb) {
        const int k0 = kb * NB;
        const int nb = (n - k0 < NB) ? (n - k0) : NB;
        const int blocks = (k0 + 63) / 64 > 0 ? (k0 + 63) / 64 : 1;
        back_solve_kernel<<<blocks, 256, 0, s>>>(A, n, n, k0, nb, invL + (size_t)kb * NB * NB, y, state);
    }
}

} // namespace ba
