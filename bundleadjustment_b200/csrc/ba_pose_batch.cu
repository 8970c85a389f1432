// ba_pose_batch.cu -- batched frame-to-frame "BA" pose estimation and correspondence pruning.
//
// Replaces, per problem, one ceres::Solve of MotionOnlyBAOptimizerAngles::optimizeCameraPose
// (ba_project/src/ba/Optimizer.cpp:417-498: PoseOnlyAngleReprojectionError<2,6> blocks with HuberLoss on ONE 6-dof
// parameter block) and BAOptimizer::pruneCorrespondences (Optimizer.cpp:6-79).  With a single parameter block
// Ceres' Schur ordering eliminates it as the only e-block, i.e. each LM step is the damped 6x6 normal-equation
// solve ("degenerate case for R3/K3").  One CTA per problem runs the WHOLE trust-region loop.
#include "ba_launch.h"
#include "ba_pose_math.cuh"

namespace ba {

constexpr int PB_THREADS = 128;

// camera record of one pose problem (shared memory, read uniformly by the CTA): R and dR/dw_k column-major, t, K
struct PoseCam {
    double R[9];
    double dR[27];
    double t[3];
    double K[9];
};

struct PoseShared {
    PoseCam ce;               // R, dR, t, K at the point being evaluated
    double x[6], cand[6], scale[6];
    double H[36];             // J^T J (unscaled, robustified)
    double g[6];              // J^T r (unscaled)
    double red[4][28];
    double cost, cand_cost, x_cost, x_norm, radius, decrease_factor, mcc, dxsq;
    int need_linearize, done, num_invalid, n_recorded, termination, msg, num_successful, num_unsuccessful, solver_fail;
    double initial_cost;
    ba_iteration it;
};

// residual (+ 2x6 Jacobian) of PoseOnlyAngleReprojectionError: p = [R t] * Xh
template <bool WANT_J>
__device__ __forceinline__ void pose_obs(const PoseCam &ce, float4 Xh, float2 uv, double r[2], double Jc[12]) {
    const double X[3] = {(double)Xh.x, (double)Xh.y, (double)Xh.z};
    const double w = (double)Xh.w;
    double p[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) p[i] = ce.R[i] * X[0] + ce.R[i + 3] * X[1] + ce.R[i + 6] * X[2] + ce.t[i] * w;
    double pix[2], G[6];
    project(ce.K, p, pix, G, WANT_J);
    r[0] = pix[0] - (double)uv.x;
    r[1] = pix[1] - (double)uv.y;
    if (WANT_J) {
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const double *D = ce.dR + 9 * j;
            double d[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) d[i] = D[i] * X[0] + D[i + 3] * X[1] + D[i + 6] * X[2];
            Jc[j] = G[0] * d[0] + G[1] * d[1] + G[2] * d[2];
            Jc[6 + j] = G[3] * d[0] + G[4] * d[1] + G[5] * d[2];
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) { Jc[3 + j] = G[j] * w; Jc[9 + j] = G[3 + j] * w; }
    }
}

__device__ __forceinline__ void pose_set_cam(PoseShared &S, const double *x, bool derivs) {
    if (threadIdx.x == 0) {
        double R[9], dR[27];
        rodrigues(x, R, derivs ? dR : nullptr);
        for (int i = 0; i < 9; ++i) S.ce.R[i] = R[i];
        if (derivs) for (int i = 0; i < 27; ++i) S.ce.dR[i] = dR[i];
        S.ce.t[0] = x[3]; S.ce.t[1] = x[4]; S.ce.t[2] = x[5];
    }
    __syncthreads();
}

// block reduction of NV values per thread into S.red -> result in S.red[0][*] visible to all after return
template <int NV>
__device__ __forceinline__ void pose_reduce(PoseShared &S, double *v) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) v[k] = warp_sum(v[k]);
    __syncthreads();
    if (lane == 0)
        for (int k = 0; k < NV; ++k) S.red[wid][k] = v[k];
    __syncthreads();
    if (threadIdx.x < NV) S.red[0][threadIdx.x] = S.red[0][threadIdx.x] + S.red[1][threadIdx.x] + S.red[2][threadIdx.x] + S.red[3][threadIdx.x];
    __syncthreads();
}

__global__ void __launch_bounds__(PB_THREADS)
pose_batch_kernel(ba_options opt, int B, double *__restrict__ cam, const float *__restrict__ K, const int *__restrict__ prob_ptr,
                  const float4 *__restrict__ Xh, const float2 *__restrict__ uv, const uint8_t *__restrict__ mask,
                  ba_summary *__restrict__ summaries, ba_iteration *__restrict__ iters, int iters_cap) {
    __shared__ PoseShared S;
    const int b = blockIdx.x;
    if (b >= B) return;
    const int o0 = prob_ptr[b], o1 = prob_ptr[b + 1];
    const int tid = threadIdx.x;
    if (tid == 0) {
        for (int k = 0; k < 6; ++k) { S.x[k] = cam[6 * (size_t)b + k]; S.scale[k] = 1.0; }
        for (int k = 0; k < 9; ++k) S.ce.K[k] = (double)K[9 * (size_t)b + k];
        S.radius = opt.initial_trust_region_radius; S.decrease_factor = 2.0;
        S.need_linearize = 1; S.done = 0; S.num_invalid = 0; S.n_recorded = 0; S.termination = BA_NO_CONVERGENCE; S.msg = MSG_NONE;
        S.num_successful = 0; S.num_unsuccessful = 0; S.initial_cost = 0;
        memset(&S.it, 0, sizeof(S.it));
        S.it.eta = opt.eta;
    }
    __syncthreads();
    ba_iteration *my_iters = iters ? iters + (size_t)b * iters_cap : nullptr;

    for (;;) {
        // ---------------- (re)linearise at x ----------------
        if (S.need_linearize) {
            pose_set_cam(S, S.x, true);
            double v[28];
#pragma unroll
            for (int k = 0; k < 28; ++k) v[k] = 0.0;
            for (int o = o0 + tid; o < o1; o += PB_THREADS) {
                if (mask && !mask[o]) continue;
                double r[2], Jc[12];
                pose_obs<true>(S.ce, Xh[o], uv[o], r, Jc);
                double rho0;
                const double rho1 = huber_rho(r[0] * r[0] + r[1] * r[1], opt.huber_delta, opt.use_huber, rho0);
                const double sr = sqrt(rho1);
                r[0] *= sr; r[1] *= sr;
#pragma unroll
                for (int a = 0; a < 12; ++a) Jc[a] *= sr;
                int k = 0;
#pragma unroll
                for (int a = 0; a < 6; ++a)
#pragma unroll
                    for (int c = a; c < 6; ++c) v[k++] += Jc[a] * Jc[c] + Jc[6 + a] * Jc[6 + c];
#pragma unroll
                for (int a = 0; a < 6; ++a) v[21 + a] += Jc[a] * r[0] + Jc[6 + a] * r[1];
                v[27] += 0.5 * rho0;
            }
            pose_reduce<28>(S, v);
            if (tid == 0) {
                int k = 0;
                for (int a = 0; a < 6; ++a)
                    for (int c = a; c < 6; ++c) { S.H[a * 6 + c] = S.red[0][k]; S.H[c * 6 + a] = S.red[0][k]; ++k; }
                for (int a = 0; a < 6; ++a) S.g[a] = S.red[0][21 + a];
                S.x_cost = S.red[0][27];
            }
            __syncthreads();
        }
        // ---------------- controller part 1 + linear solve (thread 0) ----------------
        if (tid == 0) {
            ba_iteration &it = S.it;
            if (S.need_linearize) {
                it.cost = S.x_cost;
                if (it.iteration == 0) {
                    S.initial_cost = S.x_cost;
                    it.step_is_valid = 1; it.step_is_successful = 1;
                    if (opt.jacobi_scaling)
                        for (int a = 0; a < 6; ++a) S.scale[a] = 1.0 / (1.0 + sqrt(S.H[a * 6 + a]));
                }
                double gm = 0, gs = 0, xs = 0;
                for (int a = 0; a < 6; ++a) {
                    const double d = S.x[a] - (S.x[a] + (-S.g[a]));
                    gm = fmax(gm, fabs(d)); gs += d * d; xs += S.x[a] * S.x[a];
                }
                it.gradient_max_norm = gm; it.gradient_norm = sqrt(gs); S.x_norm = sqrt(xs);
                S.need_linearize = 0;
                if (!isfinite(S.x_cost)) { S.termination = BA_FAILURE; S.msg = MSG_EVAL_FAILED; S.done = 1; }
            }
            if (!S.done) {
                if (it.step_is_successful) ++S.num_successful; else ++S.num_unsuccessful;
                it.trust_region_radius = S.radius;
                if (my_iters && S.n_recorded < iters_cap) my_iters[S.n_recorded] = it;
                ++S.n_recorded;
                if (it.iteration >= opt.max_num_iterations) { S.termination = BA_NO_CONVERGENCE; S.msg = MSG_MAX_ITER; S.done = 1; }
                else if (it.step_is_successful && it.gradient_max_norm <= opt.gradient_tolerance) { S.termination = BA_CONVERGENCE; S.msg = MSG_GRADIENT_TOL; S.done = 1; }
                else if (it.trust_region_radius <= opt.min_trust_region_radius) { S.termination = BA_CONVERGENCE; S.msg = MSG_MIN_RADIUS; S.done = 1; }
            }
            if (!S.done) {
                const double pg = it.gradient_max_norm, pn = it.gradient_norm;
                const int next = it.iteration + 1;
                memset(&it, 0, sizeof(it));
                it.iteration = next; it.gradient_max_norm = pg; it.gradient_norm = pn; it.eta = opt.eta;
                it.linear_solver_iterations = 1;
                // (J_s^T J_s + D^2) y = J_s^T r in the Jacobi-scaled space
                double A[36], bs[6], inv[36], y[6];
                for (int a = 0; a < 6; ++a) {
                    for (int c = 0; c < 6; ++c) A[a * 6 + c] = S.H[a * 6 + c] * S.scale[a] * S.scale[c];
                    bs[a] = S.g[a] * S.scale[a];
                }
                double Hs_y = 0;
                for (int a = 0; a < 6; ++a) {
                    const double d = fmin(fmax(A[a * 6 + a], opt.min_lm_diagonal), opt.max_lm_diagonal) / S.radius;
                    inv[a] = d;   // stash
                }
                double Ad[36];
                for (int k = 0; k < 36; ++k) Ad[k] = A[k];
                for (int a = 0; a < 6; ++a) Ad[a * 6 + a] += inv[a];
                S.solver_fail = inv_spd6(Ad, inv) ? 0 : 1;
                double yb = 0;
                if (!S.solver_fail) {
                    for (int a = 0; a < 6; ++a) { double s = 0; for (int c = 0; c < 6; ++c) s += inv[a * 6 + c] * bs[c]; y[a] = s; if (!isfinite(s)) S.solver_fail = 1; }
                }
                if (!S.solver_fail) {
                    for (int a = 0; a < 6; ++a) { double s = 0; for (int c = 0; c < 6; ++c) s += A[a * 6 + c] * y[c]; Hs_y += y[a] * s; yb += y[a] * bs[a]; }
                    S.mcc = yb - 0.5 * Hs_y;
                    double dx = 0;
                    for (int a = 0; a < 6; ++a) { S.cand[a] = S.x[a] + (-y[a]) * S.scale[a]; const double d = S.x[a] - S.cand[a]; dx += d * d; }
                    S.dxsq = dx;
                } else {
                    S.mcc = -1.0;
                    for (int a = 0; a < 6; ++a) S.cand[a] = S.x[a];
                }
            }
        }
        __syncthreads();
        if (S.done) break;
        // ---------------- candidate cost ----------------
        pose_set_cam(S, S.cand, false);
        {
            double c1[1] = {0.0};
            for (int o = o0 + tid; o < o1; o += PB_THREADS) {
                if (mask && !mask[o]) continue;
                double r[2];
                pose_obs<false>(S.ce, Xh[o], uv[o], r, nullptr);
                double rho0;
                huber_rho(r[0] * r[0] + r[1] * r[1], opt.huber_delta, opt.use_huber, rho0);
                c1[0] += 0.5 * rho0;
            }
            pose_reduce<1>(S, c1);
        }
        // ---------------- controller part 2 (thread 0) ----------------
        if (tid == 0) {
            ba_iteration &it = S.it;
            double cand_cost = S.red[0][0];
            it.step_is_valid = (!S.solver_fail && isfinite(S.mcc) && S.mcc > 0.0) ? 1 : 0;
            if (!it.step_is_valid) {
                if (++S.num_invalid >= opt.max_num_consecutive_invalid_steps) { S.termination = BA_FAILURE; S.msg = MSG_INVALID_STEPS; S.done = 1; }
                else {
                    S.radius /= S.decrease_factor; S.decrease_factor *= 2.0;
                    it.cost = S.x_cost; it.cost_change = 0; it.step_norm = 0; it.relative_decrease = 0;
                }
            } else {
                S.num_invalid = 0;
                if (!isfinite(cand_cost)) cand_cost = DBL_MAX;
                it.step_norm = sqrt(S.dxsq);
                if (it.step_norm <= opt.parameter_tolerance * (S.x_norm + opt.parameter_tolerance)) { S.termination = BA_CONVERGENCE; S.msg = MSG_PARAM_TOL; S.done = 1; }
                else {
                    it.cost_change = S.x_cost - cand_cost;
                    if (fabs(it.cost_change) <= opt.function_tolerance * S.x_cost) { S.termination = BA_CONVERGENCE; S.msg = MSG_FUNCTION_TOL; S.done = 1; }
                    else {
                        it.relative_decrease = (S.x_cost - cand_cost) / S.mcc;
                        if (it.relative_decrease > opt.min_relative_decrease) {
                            for (int a = 0; a < 6; ++a) S.x[a] = S.cand[a];
                            S.need_linearize = 1;
                            it.step_is_successful = 1;
                            const double t = 2.0 * it.relative_decrease - 1.0;
                            S.radius = fmin(opt.max_trust_region_radius, S.radius / fmax(1.0 / 3.0, 1.0 - t * t * t));
                            S.decrease_factor = 2.0;
                        } else {
                            it.step_is_successful = 0;
                            it.cost = cand_cost;
                            S.radius /= S.decrease_factor; S.decrease_factor *= 2.0;
                        }
                    }
                }
            }
        }
        __syncthreads();
        if (S.done) break;
    }
    if (tid == 0) {
        for (int k = 0; k < 6; ++k) cam[6 * (size_t)b + k] = S.x[k];
        if (summaries) {
            ba_summary sm;
            memset(&sm, 0, sizeof(sm));
            sm.termination_type = S.termination;
            sm.num_iterations = S.n_recorded;
            sm.num_successful_steps = S.num_successful;
            sm.num_unsuccessful_steps = S.num_unsuccessful;
            sm.initial_cost = S.initial_cost;
            sm.final_cost = S.x_cost;       // refined on the host from the iteration list when it is kept
            sm.message[0] = (char)S.msg;    // message code; expanded on the host
            summaries[b] = sm;
        }
    }
}

// pruneCorrespondences, one thread per keypoint, float32 arithmetic as in the reference
__global__ void prune_kernel(int n_frames, const float *__restrict__ pose, const float *__restrict__ K, const int *__restrict__ frame_ptr,
                             const float2 *__restrict__ kp_xy, const int *__restrict__ kp_octave, const float *__restrict__ pt_xyz,
                             const float *__restrict__ pt_min, const float *__restrict__ pt_max, const uint8_t *__restrict__ has_point,
                             uint8_t *__restrict__ outlier, int consider_outlier, int n_kp, const float *__restrict__ extr) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_kp) return;
    if (!(consider_outlier || !outlier[i])) return;
    if (!has_point[i]) return;
    int lo = 0, hi = n_frames;                 // frame f with frame_ptr[f] <= i < frame_ptr[f+1]
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (frame_ptr[mid] <= i) lo = mid; else hi = mid; }
    const int f = lo;
    const float *E = extr + 16 * (size_t)f;
    const float *P = pose + 16 * (size_t)f;
    const float *Kf = K + 9 * (size_t)f;
    const float chiThresh = 5.991f;
    const float invSigmaP = (float)(1.0 / pow(1.2, (double)kp_octave[i]));
    const float X0 = pt_xyz[3 * (size_t)i], X1 = pt_xyz[3 * (size_t)i + 1], X2 = pt_xyz[3 * (size_t)i + 2];
    float ph[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) ph[r] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(E[r * 4 + 0], X0), __fmul_rn(E[r * 4 + 1], X1)), __fmul_rn(E[r * 4 + 2], X2)), __fmul_rn(E[r * 4 + 3], 1.0f));
    const float c0 = __fdiv_rn(ph[0], ph[3]), c1 = __fdiv_rn(ph[1], ph[3]), c2 = __fdiv_rn(ph[2], ph[3]);
    if (c2 <= 0) { outlier[i] = 1; return; }
    const float dx = __fsub_rn(X0, P[3]), dy = __fsub_rn(X1, P[7]), dz = __fsub_rn(X2, P[11]);
    const float worldDist = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)));
    if (worldDist > pt_max[i] || worldDist < pt_min[i]) { outlier[i] = 1; return; }
    float q[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) q[r] = __fadd_rn(__fadd_rn(__fmul_rn(Kf[r * 3 + 0], c0), __fmul_rn(Kf[r * 3 + 1], c1)), __fmul_rn(Kf[r * 3 + 2], c2));
    const float px = __fdiv_rn(q[0], q[2]), py = __fdiv_rn(q[1], q[2]);
    const float ex = __fsub_rn(px, kp_xy[i].x), ey = __fsub_rn(py, kp_xy[i].y);
    const float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));
    const float chi = __fmul_rn(dist, invSigmaP);
    if (chi > chiThresh) { outlier[i] = 1; return; }
    outlier[i] = 0;
}

// ---- device-resident outer loop of SfMHelper::estimatePoseUsingBA (SfMHelper.cpp:106-117 -> Optimizer.cpp:417-457) ----
// per frame: pose (f32, camera->world) -> extrinsics (f32 inverse) -> angle-axis + translation (f64)   [Optimizer.cpp:462-465]
__global__ void pose_to_cam_kernel(int F, const float *__restrict__ pose, double *__restrict__ cam) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    float extr[16];
    double c[6];
    dev_pose_to_cam(pose + 16 * (size_t)f, c, extr);
#pragma unroll
    for (int k = 0; k < 6; ++k) cam[6 * (size_t)f + k] = c[k];
}
// per frame: write-back (Optimizer.cpp:448-453) pose = inverse(f32(R(cam), t)), and the extrinsics pruneCorrespondences
// derives from that pose (Optimizer.cpp:10: extr = pose.inverse())
__global__ void cam_to_pose_kernel(int F, const double *__restrict__ cam, float *__restrict__ pose, float *__restrict__ extr) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    float P[16], E[16];
    double c[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) c[k] = cam[6 * (size_t)f + k];
    dev_cam_to_pose(c, P);
    dev_inverse4x4_f32(P, E);
#pragma unroll
    for (int k = 0; k < 16; ++k) { pose[16 * (size_t)f + k] = P[k]; extr[16 * (size_t)f + k] = E[k]; }
}
// extr = pose.inverse() alone (ba_prune_frames)
__global__ void pose_inverse_kernel(int F, const float *__restrict__ pose, float *__restrict__ extr) {
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f >= F) return;
    float E[16];
    dev_inverse4x4_f32(pose + 16 * (size_t)f, E);
#pragma unroll
    for (int k = 0; k < 16; ++k) extr[16 * (size_t)f + k] = E[k];
}
// per keypoint: the observation list of the next solve = keypoints with a map point that are not outliers (Optimizer.cpp:472);
// Xh = MapPoint::getPosition().homogeneous() (Optimizer.cpp:482), written once
__global__ void pose_mask_kernel(int n_kp, const uint8_t *__restrict__ has_point, const uint8_t *__restrict__ outlier, uint8_t *__restrict__ mask,
                                 const float *__restrict__ xyz, float4 *__restrict__ Xh) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_kp) return;
    mask[i] = (has_point[i] && !outlier[i]) ? 1 : 0;
    if (Xh) Xh[i] = make_float4(xyz[3 * (size_t)i], xyz[3 * (size_t)i + 1], xyz[3 * (size_t)i + 2], 1.0f);
}

// ---- device-resident outer loop of Global / LocalBAOptimizerAngles (Optimizer.cpp:222-271, 550-603) ------------------------
// Write-back of one ceres::Solve and the start of the next: per camera with wb[c] != 0 (the frames the reference writes back:
// its optimised list, frame ID 0 included)  pose = inverse(f32(R(cam), t))  (Optimizer.cpp:262-267), and the next problem build
// reads that pose again:  extr = pose.inverse() (f32), cam = (angle-axis(extr), t)  (Optimizer.cpp:296-299) -- extr is what a
// fixed-by-ID camera's point-only residuals bake in.  Cameras with wb == 0 (Local BA's fixed observers) keep pose and extr.
__global__ void outer_cam_writeback_kernel(int Nc, const unsigned char *__restrict__ wb, const double *__restrict__ cam_cur,
                                           float *__restrict__ pose, double *__restrict__ cam_init, float *__restrict__ cam_extr) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc || !wb[c]) return;
    double x[6], y[6];
    float P[16], E[16];
#pragma unroll
    for (int k = 0; k < 6; ++k) x[k] = cam_cur[6 * (size_t)c + k];
    dev_cam_to_pose(x, P);
    dev_pose_to_cam(P, y, E);
#pragma unroll
    for (int k = 0; k < 16; ++k) { pose[16 * (size_t)c + k] = P[k]; cam_extr[16 * (size_t)c + k] = E[k]; }
#pragma unroll
    for (int k = 0; k < 6; ++k) cam_init[6 * (size_t)c + k] = y[k];
}
// landmarks: MapPoint::setPosition((float) x) (Optimizer.cpp:252-258), read back as double by the next build (:282-289)
__global__ void outer_pt_writeback_kernel(size_t n3, const double *__restrict__ pt_cur, float *__restrict__ pt32, double *__restrict__ pt_init) {
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const float v = (float)pt_cur[i];
    pt32[i] = v;
    pt_init[i] = (double)v;
}
// per keypoint: the position pruneCorrespondences sees = the keypoint's map point after the write-back (a problem landmark) or
// its constant position (a map point outside the problem)
__global__ void outer_kp_xyz_kernel(int n_kp, const int *__restrict__ kp_pt, const float *__restrict__ pt32, float *__restrict__ kp_xyz) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_kp) return;
    const int p = kp_pt[i];
    if (p >= 0) { kp_xyz[3 * (size_t)i] = pt32[3 * (size_t)p]; kp_xyz[3 * (size_t)i + 1] = pt32[3 * (size_t)p + 1]; kp_xyz[3 * (size_t)i + 2] = pt32[3 * (size_t)p + 2]; }
}
// observations whose keypoint pruneCorrespondences flagged leave the problem (Optimizer.cpp:308), cameras left without any
// observation leave it too (ceres: a parameter block without residual blocks is not part of the problem)
__global__ void outer_mask_obs_kernel(int n_kp, const int *__restrict__ kp_obs, const unsigned char *__restrict__ outlier, const int *__restrict__ perm,
                                      const int *__restrict__ pm_cam, unsigned char *__restrict__ obs_masked, int *__restrict__ cam_active) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_kp) return;
    const int o = kp_obs[i];
    if (o < 0) return;
    const int e = perm[o];
    const unsigned char m = outlier[i] ? 1 : 0;
    obs_masked[e] = m;
    if (!m) atomicAdd(cam_active + pm_cam[e], 1);
}
__global__ void outer_cam_kind_kernel(int Nc, const unsigned char *__restrict__ fixed, const int *__restrict__ cam_active, int *__restrict__ cam_kind) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc) return;
    cam_kind[c] = cam_active[c] > 0 ? (fixed[c] ? CAM_FIXED : CAM_FREE) : CAM_INACTIVE;
}
__global__ void outer_camk_kernel(int No, const int *__restrict__ pm_cam, const int *__restrict__ cam_kind, const unsigned char *__restrict__ obs_masked,
                                  unsigned *__restrict__ pm_camk) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= No) return;
    const int c = pm_cam[e];
    pm_camk[e] = (unsigned)c | ((unsigned)(obs_masked[e] ? OBS_MASKED : cam_kind[c]) << 30);
}

void launch_outer_writeback(int Nc, const unsigned char *wb, const double *cam_cur, float *pose, double *cam_init, float *cam_extr,
                            size_t n3, const double *pt_cur, float *pt32, double *pt_init, cudaStream_t s) {
    if (Nc > 0) outer_cam_writeback_kernel<<<(Nc + 127) / 128, 128, 0, s>>>(Nc, wb, cam_cur, pose, cam_init, cam_extr);
    if (n3 > 0) outer_pt_writeback_kernel<<<(unsigned)((n3 + 255) / 256), 256, 0, s>>>(n3, pt_cur, pt32, pt_init);
}
void launch_outer_kp_xyz(int n_kp, const int *kp_pt, const float *pt32, float *kp_xyz, cudaStream_t s) {
    if (n_kp > 0) outer_kp_xyz_kernel<<<(n_kp + 255) / 256, 256, 0, s>>>(n_kp, kp_pt, pt32, kp_xyz);
}
void launch_outer_mask(int n_kp, const int *kp_obs, const unsigned char *outlier, const int *perm, int Nc, const unsigned char *fixed, int No,
                       const int *pm_cam, unsigned char *obs_masked, int *cam_active, int *cam_kind, unsigned *pm_camk, cudaStream_t s) {
    cudaMemsetAsync(cam_active, 0, 4 * (size_t)Nc, s);
    if (n_kp > 0) outer_mask_obs_kernel<<<(n_kp + 255) / 256, 256, 0, s>>>(n_kp, kp_obs, outlier, perm, pm_cam, obs_masked, cam_active);
    outer_cam_kind_kernel<<<(Nc + 127) / 128, 128, 0, s>>>(Nc, fixed, cam_active, cam_kind);
    if (No > 0) outer_camk_kernel<<<(No + 255) / 256, 256, 0, s>>>(No, pm_cam, cam_kind, obs_masked, pm_camk);
}

void launch_pose_to_cam(int F, const float *pose, double *cam, cudaStream_t s) {
    if (F > 0) pose_to_cam_kernel<<<(F + 127) / 128, 128, 0, s>>>(F, pose, cam);
}
void launch_cam_to_pose(int F, const double *cam, float *pose, float *extr, cudaStream_t s) {
    if (F > 0) cam_to_pose_kernel<<<(F + 127) / 128, 128, 0, s>>>(F, cam, pose, extr);
}
void launch_pose_inverse(int F, const float *pose, float *extr, cudaStream_t s) {
    if (F > 0) pose_inverse_kernel<<<(F + 127) / 128, 128, 0, s>>>(F, pose, extr);
}
void launch_pose_mask(int n_kp, const uint8_t *has_point, const uint8_t *outlier, uint8_t *mask, const float *xyz, float4 *Xh, cudaStream_t s) {
    if (n_kp > 0) pose_mask_kernel<<<(n_kp + 255) / 256, 256, 0, s>>>(n_kp, has_point, outlier, mask, xyz, Xh);
}

void launch_pose_batch(const ba_options &opt, int B, double *cam, const float *K, const int *prob_ptr,
                       const float4 *Xh, const float2 *uv, const uint8_t *mask, ba_summary *summaries,
                       ba_iteration *iters, int iters_cap, cudaStream_t s) {
    if (B <= 0) return;
    pose_batch_kernel<<<B, PB_THREADS, 0, s>>>(opt, B, cam, K, prob_ptr, Xh, uv, mask, summaries, iters, iters_cap);
}

void launch_prune(int n_frames, const float *pose, const float *extr, const float *K, const int *frame_ptr, const float2 *kp_xy,
                  const int *kp_octave, const float *pt_xyz, const float *pt_min, const float *pt_max,
                  const uint8_t *has_point, uint8_t *outlier, int consider_outlier, int n_kp, cudaStream_t s) {
    if (n_kp <= 0) return;
    prune_kernel<<<(n_kp + 255) / 256, 256, 0, s>>>(n_frames, pose, K, frame_ptr, kp_xy, kp_octave, pt_xyz, pt_min, pt_max,
                                                  has_point, outlier, consider_outlier, n_kp, extr);
}

} // namespace ba

// =============================================================================================================
// ba_prune_frames_noalloc: the prune path with its scratch in static device memory (include/ba_b200.h; no counterpart in
// the reference -- added for callers that must not allocate)
// =============================================================================================================
#include <cstdio>
#include <mutex>

namespace {

// extr = pose.inverse() of every frame of one call (what ba_prune_frames keeps in Workspace::extr).  One array per GPU
// that loads this module, shared by every call on that GPU.  In HOST code this name is only a stand-in for the symbol:
// the kernels get the address cudaGetSymbolAddress returns for the current device.
static __device__ __align__(256) float g_prune_extr[16 * BA_PRUNE_NOALLOC_MAX_FRAMES];

constexpr int NOALLOC_MAX_DEVICES = 64;
std::mutex g_prune_mu;
float *g_prune_extr_ptr[NOALLOC_MAX_DEVICES];      // device address of g_prune_extr on device d, nullptr = not looked up yet

int noalloc_fail(int code, const char *what, cudaError_t e) {
    char buf[256];
    std::snprintf(buf, sizeof(buf), "ba_prune_frames_noalloc: %s: %s", what, cudaGetErrorString(e));
    return ba::report_error(code, buf);
}

// once per device, with that device current
int prune_scratch(int device, float **out) {
    std::lock_guard<std::mutex> lock(g_prune_mu);
    if (!g_prune_extr_ptr[device]) {
        int major = 0;
        cudaError_t e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device);
        if (e != cudaSuccess) return noalloc_fail(BA_ERR_CUDA, "cudaDeviceGetAttribute", e);
        if (major < 10) return ba::report_error(BA_ERR_CUDA, "ba_prune_frames_noalloc: this build targets sm_100a only");
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_prune_extr);
        if (e != cudaSuccess || !p) return noalloc_fail(BA_ERR_CUDA, "cudaGetSymbolAddress(g_prune_extr)", e);
        g_prune_extr_ptr[device] = static_cast<float *>(p);
    }
    *out = g_prune_extr_ptr[device];
    return BA_OK;
}

} // namespace

extern "C" int ba_prune_frames_noalloc(int device, void *stream, int32_t n_frames, int32_t n_kp, const float *d_pose,
                                       const float *d_K, const int32_t *d_frame_ptr, const float *d_kp_xy,
                                       const int32_t *d_kp_octave, const float *d_pt_xyz, const float *d_pt_min_dist,
                                       const float *d_pt_max_dist, const uint8_t *d_has_point, uint8_t *d_outlier,
                                       int32_t consider_outlier) {
    char buf[256];
    if (n_frames < 0 || n_kp < 0) return ba::report_error(BA_ERR_INVALID, "ba_prune_frames_noalloc: negative size");
    if (n_frames > BA_PRUNE_NOALLOC_MAX_FRAMES) {
        std::snprintf(buf, sizeof(buf), "ba_prune_frames_noalloc: n_frames %d exceeds BA_PRUNE_NOALLOC_MAX_FRAMES (%d): the scratch "
                      "is a fixed static device array; use ba_prune_frames or split the batch", (int)n_frames, BA_PRUNE_NOALLOC_MAX_FRAMES);
        return ba::report_error(BA_ERR_INVALID, buf);
    }
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        cudaGetLastError();
        return ba::report_error(BA_ERR_CUDA, "ba_prune_frames_noalloc: no CUDA device available; this library has no CPU fallback");
    }
    if (device < 0 || device >= n_dev || device >= NOALLOC_MAX_DEVICES) return ba::report_error(BA_ERR_INVALID, "ba_prune_frames_noalloc: device out of range");
    if (n_frames == 0 || n_kp == 0) return BA_OK;
    if (!d_pose || !d_K || !d_frame_ptr || !d_kp_xy || !d_kp_octave || !d_pt_xyz || !d_pt_min_dist || !d_pt_max_dist || !d_has_point || !d_outlier)
        return ba::report_error(BA_ERR_INVALID, "ba_prune_frames_noalloc: NULL device array");
    if ((e = cudaSetDevice(device)) != cudaSuccess) return noalloc_fail(BA_ERR_CUDA, "cudaSetDevice", e);
    float *extr = nullptr;
    int rc = prune_scratch(device, &extr);
    if (rc) return rc;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // the first n_frames entries of the shared scratch are rewritten here before prune_kernel reads them: nothing of the
    // previous call's contents is ever read, so no clear is needed
    ba::launch_pose_inverse(n_frames, d_pose, extr, s);                                   // extr = pose.inverse(), f32 (Optimizer.cpp:10)
    ba::launch_prune(n_frames, d_pose, extr, d_K, d_frame_ptr, reinterpret_cast<const float2 *>(d_kp_xy), d_kp_octave, d_pt_xyz,
                     d_pt_min_dist, d_pt_max_dist, d_has_point, d_outlier, consider_outlier, n_kp, s);
    if ((e = cudaPeekAtLastError()) != cudaSuccess) return noalloc_fail(BA_ERR_CUDA, "kernel launch", e);
    return BA_OK;
}
