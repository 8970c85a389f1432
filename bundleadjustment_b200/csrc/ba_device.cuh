// ba_device.cuh -- device-side building blocks shared by the BA kernels (sm_100a).
//
// Math follows the reference functors (ba_project/src/ba/Optimizer.h:54-76 R1, 96-107 R2, 163-182 R3) with
// float32 inputs promoted to double exactly where the functors cast to T, and Ceres' HuberLoss + Corrector
// semantics.
#pragma once

#include <cstddef>

#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include <math.h>

namespace ba {

// ---------------------------------------------------------------------------------------------------------
// Per-camera table entry, rebuilt once per linearisation point by cam_prep_kernel so that the per-observation
// kernels never evaluate sincos.  40 doubles = 320 B in three 32-byte-aligned groups, so that a thread gathers its camera
// with 256-bit loads (LDG.E.ENL2.256; the landmark-major kernels gather a different camera in every lane, and a gather
// costs L1 data-pipe passes per load instruction and lane, not per byte).
//   free camera : R (col-major 3x3) | t,   K (row-major, promoted from f32),   Jr | sc
//                 Jr = right Jacobian of the rotation, stored as its three columns: d(R X)/dw_j = R (Jr_j x X).  It is
//                 derived in cam_prep from the exact dR/dw_j (Jr_j = vee(R^T dR_j)), 9 doubles instead of 27.
//                 sc = Jacobi scale of the six camera parameters (travels with the entry instead of a second gather).
//   fixed camera: P = 4x4 row-major world->camera extrinsics (promoted from f32) in Jr[0..8] | sc[0..5] | pad, K
// ---------------------------------------------------------------------------------------------------------
struct __align__(32) CamEntry {
    double R[9];
    double t[3];
    double K[9];
    double pad0[3];
    double Jr[9];
    double sc[6];
    double pad1;
};
static_assert(sizeof(CamEntry) == 40 * 8, "CamEntry layout");
static_assert(offsetof(CamEntry, K) % 32 == 0 && offsetof(CamEntry, Jr) % 32 == 0, "CamEntry groups must be 32-byte aligned");

// Candidate-point table (cost-only evaluation): the 3 x 4 projection P = K [R | t] (free camera) or K E[0:3, :] (fixed camera;
// the homogeneous divide of the reference's (E X).hnormalized() cancels in the pixel ratio), row-major.  Three 256-bit loads per
// observation instead of six plus a kind test; the pixel differs from the two-stage product by rounding only (~1e-16 relative).
struct __align__(32) CamProj { double P[12]; };
static_assert(sizeof(CamProj) == 96, "CamProj layout");

struct __align__(32) Double4 { double x, y, z, w; };
// N4 x 256-bit loads of 32-byte-aligned doubles into registers
template <int N4>
__device__ __forceinline__ void load_vec4(const double *__restrict__ src, double *dst) {
#pragma unroll
    for (int i = 0; i < N4; ++i) {
        const Double4 v = *reinterpret_cast<const Double4 *>(src + 4 * i);
        dst[4 * i] = v.x; dst[4 * i + 1] = v.y; dst[4 * i + 2] = v.z; dst[4 * i + 3] = v.w;
    }
}

// landmark p of a tight [.][3] parameter array
__device__ __forceinline__ void load_pt(const double *__restrict__ base, size_t p, double (&X)[3]) {
    X[0] = base[3 * p]; X[1] = base[3 * p + 1]; X[2] = base[3 * p + 2];
}

// Row `row` of a [.][18]-double array (W: 144-byte rows, so odd rows start 16 bytes off a 32-byte boundary) as five 256-bit
// loads of the enclosing 32-byte-aligned window + a parity select.  A per-lane gather costs one L1 sector request per load
// instruction and lane, whatever the width: 5 requests instead of 9 (128-bit) or 18 (64-bit).  The array needs 16 bytes of
// slack behind its last row.  load_row6: the same for [.][6] arrays (48-byte rows, two loads).
__device__ __forceinline__ void load_row18(const double *__restrict__ base, size_t row, double (&w)[18]) {
    const bool odd = row & 1;
    double raw[20];
    load_vec4<5>(base + 18 * row - (odd ? 2 : 0), raw);
#pragma unroll
    for (int k = 0; k < 18; ++k) w[k] = odd ? raw[k + 2] : raw[k];
}
__device__ __forceinline__ void load_row6(const double *__restrict__ base, size_t row, double (&v)[6]) {
    const bool odd = row & 1;
    double raw[8];
    load_vec4<2>(base + 6 * row - (odd ? 2 : 0), raw);
#pragma unroll
    for (int k = 0; k < 6; ++k) v[k] = odd ? raw[k + 2] : raw[k];
}

// ---- TMA 1-D bulk copies (cp.async.bulk, SASS UBLKCP): a landmark block's W records are ONE contiguous, 16-byte
// aligned run of 144 n bytes, so a single thread moves the whole tile between HBM and shared memory; the copy engine
// does not occupy the LSU / L1 data pipe that the per-observation gathers and reductions need.
__device__ __forceinline__ unsigned smem_addr(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(smem_addr(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
}
// one thread: arm the barrier with the byte count and start global -> shared
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gsrc, unsigned bytes, unsigned long long *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
                 ::"r"(smem_addr(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity) {
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}
// one thread, after a CTA barrier that ordered the generic-proxy writes of the tile: shared -> global, then wait until the
// source has been read (the CTA may exit / reuse the buffer afterwards)
__device__ __forceinline__ void bulk_store(void *gdst, const void *smem_src, unsigned bytes) {
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(smem_addr(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
    asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory");
}

// split form: bulk_store_begin after the CTA barrier that ordered the writes of the tile, bulk_store_end (same thread) before
// the CTA exits or reuses the buffer -- whatever lies between overlaps with the drain of the tile
__device__ __forceinline__ void bulk_store_begin(void *gdst, const void *smem_src, unsigned bytes) {
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(smem_addr(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_store_end() { asm volatile("cp.async.bulk.wait_group.read 0;\n" ::: "memory"); }

// the copy alone (no fence / commit / wait): for callers that batch several copies into one group and double buffer
__device__ __forceinline__ void bulk_store_issue(void *gdst, const void *smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;\n" ::"l"(gdst), "r"(smem_addr(smem_src)), "r"(bytes) : "memory");
}

// record k of a shared-memory tile of 18-double records as nine 16-byte loads (conflict free at the 144-byte stride)
__device__ __forceinline__ void load_record18(const double *__restrict__ sW, int k, double (&w)[18]) {
    const double2 *src = reinterpret_cast<const double2 *>(sW + 18 * (size_t)k);
#pragma unroll
    for (int q = 0; q < 9; ++q) { const double2 v = src[q]; w[2 * q] = v.x; w[2 * q + 1] = v.y; }
}

enum : int { CAM_FREE = 0, CAM_FIXED = 1, CAM_INACTIVE = 2, OBS_MASKED = 3 };
// pm_camk word: camera index | kind << 30.  kind = the camera's CAM_* kind, or OBS_MASKED for an observation that the device-
// resident outer loop (ba_solve_outer) dropped after pruneCorrespondences flagged its keypoint: it contributes nothing, exactly
// like a residual block the reference does not add when it rebuilds the problem (Optimizer.cpp:308: isOutlier -> continue).
__device__ __forceinline__ int camk_cam(unsigned v) { return (int)(v & 0x3fffffffu); }
__device__ __forceinline__ int camk_kind(unsigned v) { return (int)(v >> 30); }

// ceres::AngleAxisToRotationMatrix + closed-form derivative of the branch taken (autodiff-equivalent)
__device__ __forceinline__ void rodrigues(const double w[3], double R[9], double *dR /*27 or nullptr*/) {
    const double theta2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    if (theta2 > DBL_EPSILON) {
        const double theta = sqrt(theta2);
        const double k0 = w[0] / theta, k1 = w[1] / theta, k2 = w[2] / theta;
        double s, c;
        sincos(theta, &s, &c);
        const double omc = 1.0 - c;
        R[0] = c + k0 * k0 * omc;       R[1] = k2 * s + k0 * k1 * omc;  R[2] = -k1 * s + k0 * k2 * omc;
        R[3] = k0 * k1 * omc - k2 * s;  R[4] = c + k1 * k1 * omc;       R[5] = k0 * s + k1 * k2 * omc;
        R[6] = k1 * s + k0 * k2 * omc;  R[7] = -k0 * s + k1 * k2 * omc; R[8] = c + k2 * k2 * omc;
        if (dR) {
            const double k[3] = {k0, k1, k2};
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                double dk[3];
#pragma unroll
                for (int i = 0; i < 3; ++i) dk[i] = ((i == j ? 1.0 : 0.0) - k[i] * k[j]) / theta;
                const double dth = k[j];
                double *D = dR + 9 * j;
#pragma unroll
                for (int col = 0; col < 3; ++col)
#pragma unroll
                    for (int row = 0; row < 3; ++row) {
                        double v = s * dth * k[row] * k[col] + omc * (dk[row] * k[col] + k[row] * dk[col]);
                        if (row == col) v += -s * dth;
                        D[row + 3 * col] = v;
                    }
                const double a0 = c * dth * k[0] + s * dk[0];
                const double a1 = c * dth * k[1] + s * dk[1];
                const double a2 = c * dth * k[2] + s * dk[2];
                D[0 + 3 * 1] += -a2; D[0 + 3 * 2] += a1;
                D[1 + 3 * 0] += a2;  D[1 + 3 * 2] += -a0;
                D[2 + 3 * 0] += -a1; D[2 + 3 * 1] += a0;
            }
        }
    } else {
        R[0] = 1.0;   R[1] = w[2];  R[2] = -w[1];
        R[3] = -w[2]; R[4] = 1.0;   R[5] = w[0];
        R[6] = w[1];  R[7] = -w[0]; R[8] = 1.0;
        if (dR) {
#pragma unroll
            for (int i = 0; i < 27; ++i) dR[i] = 0.0;
            // d/dw0: [e0]x ; d/dw1: [e1]x ; d/dw2: [e2]x  (col-major)
            dR[0 * 9 + 2 + 3 * 1] = 1.0;  dR[0 * 9 + 1 + 3 * 2] = -1.0;
            dR[1 * 9 + 0 + 3 * 2] = 1.0;  dR[1 * 9 + 2 + 3 * 0] = -1.0;
            dR[2 * 9 + 1 + 3 * 0] = 1.0;  dR[2 * 9 + 0 + 3 * 1] = -1.0;
        }
    }
}

// ceres::HuberLoss::Evaluate: returns rho'(s) and sets rho0 = rho(s).  a = delta, b = a*a.
__device__ __forceinline__ double huber_rho(double s, double a, int use_huber, double &rho0) {
    if (!use_huber) { rho0 = s; return 1.0; }
    const double b = a * a;
    if (s > b) {
        const double r = sqrt(s);
        rho0 = 2.0 * a * r - b;
        return fmax(DBL_MIN, a / r);
    }
    rho0 = s;
    return 1.0;
}

// pinhole tail: q = K p, pix = q.hnormalized(), G = d pix / d p (2x3 row-major)
__device__ __forceinline__ void project(const double *__restrict__ K, const double p[3], double pix[2], double G[6], bool want_G) {
    const double q0 = K[0] * p[0] + K[1] * p[1] + K[2] * p[2];
    const double q1 = K[3] * p[0] + K[4] * p[1] + K[5] * p[2];
    const double q2 = K[6] * p[0] + K[7] * p[1] + K[8] * p[2];
    pix[0] = q0 / q2;
    pix[1] = q1 / q2;
    if (want_G) {
        const double iq = 1.0 / q2;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            G[j] = (K[j] - pix[0] * K[6 + j]) * iq;
            G[3 + j] = (K[3 + j] - pix[1] * K[6 + j]) * iq;
        }
    }
}

// residual only (cost evaluation at a candidate point) from the camera's 3 x 4 projection
__device__ __forceinline__ void obs_residual(const CamProj *__restrict__ cp, const double X[3], float2 uv, double r[2]) {
    double P[12];
    load_vec4<3>(cp->P, P);
    const double q0 = P[0] * X[0] + P[1] * X[1] + P[2] * X[2] + P[3];
    const double q1 = P[4] * X[0] + P[5] * X[1] + P[6] * X[2] + P[7];
    const double q2 = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11];
    r[0] = q0 / q2 - (double)uv.x;
    r[1] = q1 / q2 - (double)uv.y;
}
// P = K [R | t] with R column-major; K row-major promoted from f32
__device__ __forceinline__ void cam_projection(const float *__restrict__ K, const double R[9], const double t[3], double P[12]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const double k0 = (double)K[3 * i], k1 = (double)K[3 * i + 1], k2 = (double)K[3 * i + 2];
#pragma unroll
        for (int j = 0; j < 3; ++j) P[4 * i + j] = k0 * R[3 * j] + k1 * R[3 * j + 1] + k2 * R[3 * j + 2];
        P[4 * i + 3] = k0 * t[0] + k1 * t[1] + k2 * t[2];
    }
}

// residual + Jacobians (unrobustified, unscaled).  Jc 2x6 row-major (zero for fixed), Jp 2x3 row-major.
// sc (may be nullptr): receives the camera's six Jacobi scales from the table entry (1 for fixed cameras).
template <bool WANT_JC, bool WANT_JP>
__device__ __forceinline__ void obs_linearize(const CamEntry *__restrict__ ce, int kind, const double X[3], float2 uv,
                                              double r[2], double Jc[12], double Jp[6], double *sc = nullptr) {
    double p[3], pix[2], G[6], K[12];
    load_vec4<3>(ce->K, K);
    if (kind == CAM_FREE) {
        double Rt[12];
        load_vec4<3>(ce->R, Rt);
#pragma unroll
        for (int i = 0; i < 3; ++i) p[i] = Rt[i] * X[0] + Rt[i + 3] * X[1] + Rt[i + 6] * X[2] + Rt[9 + i];
        project(K, p, pix, G, true);
        double GR[6];                                     // d pix / d X = G R (2 x 3)
#pragma unroll
        for (int rr = 0; rr < 2; ++rr)
#pragma unroll
            for (int j = 0; j < 3; ++j)
                GR[rr * 3 + j] = G[rr * 3 + 0] * Rt[0 + 3 * j] + G[rr * 3 + 1] * Rt[1 + 3 * j] + G[rr * 3 + 2] * Rt[2 + 3 * j];
        if (WANT_JP) {
#pragma unroll
            for (int k = 0; k < 6; ++k) Jp[k] = GR[k];
        }
        if (WANT_JC) {
            double Js[16];
            load_vec4<4>(ce->Jr, Js);
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                // d(R X)/dw_j = R (Jr_j x X)  ->  d pix / dw_j = (G R) (Jr_j x X)
                const double a0 = Js[3 * j], a1 = Js[3 * j + 1], a2 = Js[3 * j + 2];
                const double c0 = a1 * X[2] - a2 * X[1], c1 = a2 * X[0] - a0 * X[2], c2 = a0 * X[1] - a1 * X[0];
                Jc[j] = GR[0] * c0 + GR[1] * c1 + GR[2] * c2;
                Jc[6 + j] = GR[3] * c0 + GR[4] * c1 + GR[5] * c2;
            }
#pragma unroll
            for (int j = 0; j < 3; ++j) { Jc[3 + j] = G[j]; Jc[9 + j] = G[3 + j]; }
            if (sc) {
#pragma unroll
                for (int a = 0; a < 6; ++a) sc[a] = Js[9 + a];
            }
        }
    } else {
        double P[16];
        load_vec4<4>(ce->Jr, P);
        double ph[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) ph[i] = P[4 * i] * X[0] + P[4 * i + 1] * X[1] + P[4 * i + 2] * X[2] + P[4 * i + 3];
        const double ih = 1.0 / ph[3];
        p[0] = ph[0] / ph[3]; p[1] = ph[1] / ph[3]; p[2] = ph[2] / ph[3];
        project(K, p, pix, G, true);
        if (WANT_JP) {
            double dp[9];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 3; ++j) dp[i * 3 + j] = (P[4 * i + j] - p[i] * P[12 + j]) * ih;
#pragma unroll
            for (int rr = 0; rr < 2; ++rr)
#pragma unroll
                for (int j = 0; j < 3; ++j)
                    Jp[rr * 3 + j] = G[rr * 3 + 0] * dp[j] + G[rr * 3 + 1] * dp[3 + j] + G[rr * 3 + 2] * dp[6 + j];
        }
        if (WANT_JC) {
#pragma unroll
            for (int j = 0; j < 12; ++j) Jc[j] = 0.0;
            if (sc) {
#pragma unroll
                for (int a = 0; a < 6; ++a) sc[a] = 1.0;
            }
        }
    }
    r[0] = pix[0] - (double)uv.x;
    r[1] = pix[1] - (double)uv.y;
}

// L2 prefetch of a line the thread will read a few microseconds later (per-landmark rows of the tile kernels' second phase):
// costs no registers, turns a DRAM round trip on the critical path of the CTA into an L2 hit
__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];\n" ::"l"(p)); }

// ---- reductions --------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// deterministic block sum (fixed tree); result valid in thread 0.  smem: >= 32 doubles.
__device__ __forceinline__ double block_sum(double v, double *smem) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? smem[threadIdx.x] : 0.0;
    if (wid == 0) v = warp_sum(v);
    return v;
}
__device__ __forceinline__ double block_max(double v, double *smem) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) smem[wid] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? smem[threadIdx.x] : 0.0;
    if (wid == 0) v = warp_max(v);
    return v;
}

// three sums and one max in ONE pass (fixed trees; results valid in thread 0).  smem: >= 4 * 32 doubles, not in use by
// anyone else; one CTA barrier instead of eight.
// MAX4: the fourth value is a maximum of non-negative numbers (else a fourth sum).
template <bool MAX4>
__device__ __forceinline__ void block_red4(double &a, double &b, double &c, double &m, double *smem) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c); m = MAX4 ? warp_max(m) : warp_sum(m);
    if (lane == 0) { smem[wid] = a; smem[32 + wid] = b; smem[64 + wid] = c; smem[96 + wid] = m; }
    __syncthreads();
    if (wid == 0) {
        const int nw = (blockDim.x + 31) >> 5;
        a = lane < nw ? smem[lane] : 0.0; b = lane < nw ? smem[32 + lane] : 0.0;
        c = lane < nw ? smem[64 + lane] : 0.0; m = lane < nw ? smem[96 + lane] : 0.0;
        a = warp_sum(a); b = warp_sum(b); c = warp_sum(c); m = MAX4 ? warp_max(m) : warp_sum(m);
    }
}
__device__ __forceinline__ void block_sum3_max(double &a, double &b, double &c, double &m, double *smem) { block_red4<true>(a, b, c, m, smem); }

// ---- multi-CTA reductions with a "last CTA sums" tail (no second launch, no cooperative launch) ---------------------------
// Every CTA publishes its four block-reduced values, the last one to arrive (atomic ticket) adds the published partials in
// CTA order with the same fixed tree and returns true with the totals valid in thread 0.  Deterministic for a given grid.
// part: [gridDim.x] double4 slots owned by this kernel; count: zero before the launch, reset by the last CTA.
constexpr int RED_MAX_GRID = 1024;
template <bool MAX4>
__device__ __forceinline__ bool grid_red4(double &a, double &b, double &c, double &m, double *part, unsigned *count, double *smem) {
    block_red4<MAX4>(a, b, c, m, smem);
    if (gridDim.x == 1) return true;
    __shared__ int s_last;
    if (threadIdx.x == 0) {
        reinterpret_cast<double4 *>(part)[blockIdx.x] = make_double4(a, b, c, m);
        __threadfence();
        const unsigned t = atomicAdd(count, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return false;
    __threadfence();
    a = b = c = m = 0.0;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) {
        const double2 v0 = __ldcg(reinterpret_cast<const double2 *>(part) + 2 * i), v1 = __ldcg(reinterpret_cast<const double2 *>(part) + 2 * i + 1);
        a += v0.x; b += v0.y; c += v1.x; m = MAX4 ? fmax(m, v1.y) : m + v1.y;
    }
    block_red4<MAX4>(a, b, c, m, smem);
    if (threadIdx.x == 0) *count = 0;
    return true;
}
__device__ __forceinline__ bool grid_sum3_max(double &a, double &b, double &c, double &m, double *part, unsigned *count, double *smem) {
    return grid_red4<true>(a, b, c, m, part, count, smem);
}
__device__ __forceinline__ bool grid_sum4(double &a, double &b, double &c, double &d, double *part, unsigned *count, double *smem) {
    return grid_red4<false>(a, b, c, d, part, count, smem);
}

// symmetric 3x3 SPD inverse via Cholesky (InvertPSDMatrix).  m = {00,01,02,11,12,22}; returns false if not PD.
__device__ __forceinline__ bool inv_sym3(const double m[6], double inv[6]) {
    const double l00sq = m[0];
    if (!(l00sq > 0.0)) return false;
    const double l00 = sqrt(l00sq);
    const double l10 = m[1] / l00, l20 = m[2] / l00;
    const double d1 = m[3] - l10 * l10;
    if (!(d1 > 0.0)) return false;
    const double l11 = sqrt(d1);
    const double l21 = (m[4] - l20 * l10) / l11;
    const double d2 = m[5] - l20 * l20 - l21 * l21;
    if (!(d2 > 0.0)) return false;
    const double l22 = sqrt(d2);
    // inverse of L (lower)
    const double i00 = 1.0 / l00, i11 = 1.0 / l11, i22 = 1.0 / l22;
    const double i10 = -l10 * i00 * i11;
    const double i21 = -l21 * i11 * i22;
    const double i20 = -(l20 * i00 + l21 * i10) * i22;
    // A^-1 = L^-T L^-1
    inv[0] = i00 * i00 + i10 * i10 + i20 * i20;
    inv[1] = i10 * i11 + i20 * i21;
    inv[2] = i20 * i22;
    inv[3] = i11 * i11 + i21 * i21;
    inv[4] = i21 * i22;
    inv[5] = i22 * i22;
    return true;
}

// dense 6x6 SPD inverse (row-major full in/out) via Cholesky; returns false if not PD.
__device__ __forceinline__ bool inv_spd6(const double *A, double *inv) {
    double L[36];
#pragma unroll
    for (int i = 0; i < 36; ++i) L[i] = 0.0;
    for (int j = 0; j < 6; ++j) {
        double d = A[j * 6 + j];
        for (int k = 0; k < j; ++k) d -= L[j * 6 + k] * L[j * 6 + k];
        if (!(d > 0.0)) return false;
        d = sqrt(d);
        L[j * 6 + j] = d;
        for (int i = j + 1; i < 6; ++i) {
            double v = A[i * 6 + j];
            for (int k = 0; k < j; ++k) v -= L[i * 6 + k] * L[j * 6 + k];
            L[i * 6 + j] = v / d;
        }
    }
    for (int c = 0; c < 6; ++c) {
        double y[6];
        for (int i = 0; i < 6; ++i) {
            double v = (i == c) ? 1.0 : 0.0;
            for (int k = 0; k < i; ++k) v -= L[i * 6 + k] * y[k];
            y[i] = v / L[i * 6 + i];
        }
        for (int i = 5; i >= 0; --i) {
            double v = y[i];
            for (int k = i + 1; k < 6; ++k) v -= L[k * 6 + i] * inv[k * 6 + c];
            inv[i * 6 + c] = v / L[i * 6 + i];
        }
    }
    return true;
}

} // namespace ba
