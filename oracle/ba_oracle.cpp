// ba_oracle.cpp -- CPU restatement of the reference's BA optimisation path (see ba_oracle.h).
//
// TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (no Ceres in tree / image; no reference tests for the path).
//
// What follows which reference lines:
//   functors .................. ba_project/src/ba/Optimizer.h:49-88 (R1), 91-120 (R2), 156-194 (R3)
//   loss ...................... ceres::HuberLoss(sqrt(5.991)) at Optimizer.cpp:312, 489, 645, 686
//   solver options ............ Optimizer.cpp:80-90 (+ Ceres defaults)
//   problem build / solve ..... Optimizer.cpp:216-333 (global), 417-498 (motion only), 500-698 (local)
//   prune ..................... Optimizer.cpp:6-79
// and, from Ceres 2.x (recalled; [CERES-EXT] in SURVEY.md 8a): rotation.h, loss_function.cc,
// corrector.cc, trust_region_minimizer.cc, levenberg_marquardt_strategy.cc,
// trust_region_step_evaluator.cc, schur_eliminator_impl.h, schur_complement_solver.cc,
// implicit_schur_complement.cc, conjugate_gradients_solver.cc.

#include "ba_oracle.h"

#include <algorithm>
#include <chrono>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <limits>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

double now_s() {
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// ------------------------------------------------------------------------------------------------
// Dual numbers ("Jets"): value + N partials.  Only what the three functors need.
// ------------------------------------------------------------------------------------------------
template <int N>
struct Jet {
    double a;
    double v[N];
    Jet() : a(0) { for (int i = 0; i < N; ++i) v[i] = 0; }
    explicit Jet(double x) : a(x) { for (int i = 0; i < N; ++i) v[i] = 0; }
    Jet(double x, int k) : a(x) { for (int i = 0; i < N; ++i) v[i] = 0; v[k] = 1.0; }
};
template <int N> Jet<N> operator+(const Jet<N> &x, const Jet<N> &y) { Jet<N> r; r.a = x.a + y.a; for (int i = 0; i < N; ++i) r.v[i] = x.v[i] + y.v[i]; return r; }
template <int N> Jet<N> operator-(const Jet<N> &x, const Jet<N> &y) { Jet<N> r; r.a = x.a - y.a; for (int i = 0; i < N; ++i) r.v[i] = x.v[i] - y.v[i]; return r; }
template <int N> Jet<N> operator-(const Jet<N> &x) { Jet<N> r; r.a = -x.a; for (int i = 0; i < N; ++i) r.v[i] = -x.v[i]; return r; }
template <int N> Jet<N> operator*(const Jet<N> &x, const Jet<N> &y) { Jet<N> r; r.a = x.a * y.a; for (int i = 0; i < N; ++i) r.v[i] = x.a * y.v[i] + x.v[i] * y.a; return r; }
template <int N> Jet<N> operator/(const Jet<N> &x, const Jet<N> &y) {
    Jet<N> r; const double inv = 1.0 / y.a; r.a = x.a * inv;
    for (int i = 0; i < N; ++i) r.v[i] = (x.v[i] - r.a * y.v[i]) * inv;
    return r;
}
template <int N> Jet<N> jsqrt(const Jet<N> &x) { Jet<N> r; r.a = std::sqrt(x.a); const double d = 0.5 / r.a; for (int i = 0; i < N; ++i) r.v[i] = x.v[i] * d; return r; }
template <int N> Jet<N> jcos(const Jet<N> &x) { Jet<N> r; r.a = std::cos(x.a); const double d = -std::sin(x.a); for (int i = 0; i < N; ++i) r.v[i] = x.v[i] * d; return r; }
template <int N> Jet<N> jsin(const Jet<N> &x) { Jet<N> r; r.a = std::sin(x.a); const double d = std::cos(x.a); for (int i = 0; i < N; ++i) r.v[i] = x.v[i] * d; return r; }
inline double jsqrt(double x) { return std::sqrt(x); }
inline double jcos(double x) { return std::cos(x); }
inline double jsin(double x) { return std::sin(x); }
inline double val(double x) { return x; }
template <int N> double val(const Jet<N> &x) { return x.a; }

// ceres::AngleAxisToRotationMatrix (rotation.h): R(i,j) column-major -> R[i + 3*j]
template <typename T>
void angle_axis_to_rotation_T(const T *aa, T *R) {
    const T one(1.0);
    const T theta2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
    if (val(theta2) > std::numeric_limits<double>::epsilon()) {
        const T theta = jsqrt(theta2);
        const T wx = aa[0] / theta, wy = aa[1] / theta, wz = aa[2] / theta;
        const T c = jcos(theta), s = jsin(theta);
        R[0 + 3 * 0] = c + wx * wx * (one - c);
        R[1 + 3 * 0] = wz * s + wx * wy * (one - c);
        R[2 + 3 * 0] = -wy * s + wx * wz * (one - c);
        R[0 + 3 * 1] = wx * wy * (one - c) - wz * s;
        R[1 + 3 * 1] = c + wy * wy * (one - c);
        R[2 + 3 * 1] = wx * s + wy * wz * (one - c);
        R[0 + 3 * 2] = wy * s + wx * wz * (one - c);
        R[1 + 3 * 2] = -wx * s + wy * wz * (one - c);
        R[2 + 3 * 2] = c + wz * wz * (one - c);
    } else {
        // first-order Taylor branch; its derivative is the derivative of THIS expression
        R[0 + 3 * 0] = one;    R[1 + 3 * 0] = aa[2];  R[2 + 3 * 0] = -aa[1];
        R[0 + 3 * 1] = -aa[2]; R[1 + 3 * 1] = one;    R[2 + 3 * 1] = aa[0];
        R[0 + 3 * 2] = aa[1];  R[1 + 3 * 2] = -aa[0]; R[2 + 3 * 2] = one;
    }
}

// Closed-form dR/dw_j (three column-major 3x3) of the expression above, branch included.
void rotation_and_derivs(const double w[3], double R[9], double dR[3][9]) {
    angle_axis_to_rotation_T<double>(w, R);
    const double theta2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    if (theta2 > std::numeric_limits<double>::epsilon()) {
        const double theta = std::sqrt(theta2);
        const double k[3] = {w[0] / theta, w[1] / theta, w[2] / theta};
        const double c = std::cos(theta), s = std::sin(theta), omc = 1.0 - c;
        for (int j = 0; j < 3; ++j) {
            double dk[3];
            for (int i = 0; i < 3; ++i) dk[i] = ((i == j ? 1.0 : 0.0) - k[i] * k[j]) / theta;
            const double dth = k[j];
            // R = c I + omc k k^T + s [k]x
            double *D = dR[j];
            for (int col = 0; col < 3; ++col)
                for (int row = 0; row < 3; ++row) {
                    double v = s * dth * k[row] * k[col] + omc * (dk[row] * k[col] + k[row] * dk[col]);
                    if (row == col) v += -s * dth;
                    D[row + 3 * col] = v;
                }
            // + d(s [k]x) = c dth [k]x + s [dk]x ; [a]x = [0 -a2 a1; a2 0 -a0; -a1 a0 0]
            double a[3];
            for (int i = 0; i < 3; ++i) a[i] = c * dth * k[i] + s * dk[i];
            D[0 + 3 * 1] += -a[2]; D[0 + 3 * 2] += a[1];
            D[1 + 3 * 0] += a[2];  D[1 + 3 * 2] += -a[0];
            D[2 + 3 * 0] += -a[1]; D[2 + 3 * 1] += a[0];
        }
    } else {
        for (int j = 0; j < 3; ++j) {
            double *D = dR[j];
            for (int i = 0; i < 9; ++i) D[i] = 0.0;
            double a[3] = {0, 0, 0}; a[j] = 1.0;
            D[0 + 3 * 1] = -a[2]; D[0 + 3 * 2] = a[1];
            D[1 + 3 * 0] = a[2];  D[1 + 3 * 2] = -a[0];
            D[2 + 3 * 0] = -a[1]; D[2 + 3 * 1] = a[0];
        }
    }
}

// pinhole tail shared by all functors: q = K p, pix = q.hnormalized(); G = d pix / d p (2x3 row-major)
inline void project(const float K[9], const double p[3], double pix[2], double G[6]) {
    double Kd[9];
    for (int i = 0; i < 9; ++i) Kd[i] = (double)K[i];
    const double q0 = Kd[0] * p[0] + Kd[1] * p[1] + Kd[2] * p[2];
    const double q1 = Kd[3] * p[0] + Kd[4] * p[1] + Kd[5] * p[2];
    const double q2 = Kd[6] * p[0] + Kd[7] * p[1] + Kd[8] * p[2];
    pix[0] = q0 / q2;
    pix[1] = q1 / q2;
    if (G) {
        const double iq = 1.0 / q2;
        for (int j = 0; j < 3; ++j) {
            G[0 * 3 + j] = (Kd[0 + j] - pix[0] * Kd[6 + j]) * iq;
            G[1 * 3 + j] = (Kd[3 + j] - pix[1] * Kd[6 + j]) * iq;
        }
    }
}

// Per-camera cache: R, dR (free cameras) -- what the autodiff cost function recomputes per residual.
struct CamCache {
    double R[9];
    double dR[3][9];
    double t[3];
};

inline void cam_cache(const double cam[6], CamCache &cc) {
    rotation_and_derivs(cam, cc.R, cc.dR);
    cc.t[0] = cam[3]; cc.t[1] = cam[4]; cc.t[2] = cam[5];
}

// R1: AngleReprojectionError, Optimizer.h:54-76
inline void eval_cam_point_cached(const CamCache &cc, const double X[3], const float uv[2], const float K[9],
                                  double res[2], double *Jc, double *Jp) {
    double p[3];
    for (int i = 0; i < 3; ++i) p[i] = cc.R[i] * X[0] + cc.R[i + 3] * X[1] + cc.R[i + 6] * X[2] + cc.t[i];
    double pix[2], G[6];
    project(K, p, pix, (Jc || Jp) ? G : nullptr);
    res[0] = pix[0] - (double)uv[0];
    res[1] = pix[1] - (double)uv[1];
    if (Jp) {
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j)
                Jp[r * 3 + j] = G[r * 3 + 0] * cc.R[0 + 3 * j] + G[r * 3 + 1] * cc.R[1 + 3 * j] + G[r * 3 + 2] * cc.R[2 + 3 * j];
    }
    if (Jc) {
        for (int j = 0; j < 3; ++j) {
            const double *D = cc.dR[j];
            double d[3];
            for (int i = 0; i < 3; ++i) d[i] = D[i] * X[0] + D[i + 3] * X[1] + D[i + 6] * X[2];
            for (int r = 0; r < 2; ++r) Jc[r * 6 + j] = G[r * 3 + 0] * d[0] + G[r * 3 + 1] * d[1] + G[r * 3 + 2] * d[2];
        }
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j) Jc[r * 6 + 3 + j] = G[r * 3 + j];
    }
}

// R2: PointOnlyReprojectionError, Optimizer.h:96-107.  E row-major f32 4x4, promoted to double.
inline void eval_point_only_d(const double X[3], const float uv[2], const float E[16], const float K[9],
                              double res[2], double *Jp) {
    double Ed[16];
    for (int i = 0; i < 16; ++i) Ed[i] = (double)E[i];
    double ph[4];
    for (int r = 0; r < 4; ++r) ph[r] = Ed[r * 4 + 0] * X[0] + Ed[r * 4 + 1] * X[1] + Ed[r * 4 + 2] * X[2] + Ed[r * 4 + 3];
    double p[3] = {ph[0] / ph[3], ph[1] / ph[3], ph[2] / ph[3]};
    double pix[2], G[6];
    project(K, p, pix, Jp ? G : nullptr);
    res[0] = pix[0] - (double)uv[0];
    res[1] = pix[1] - (double)uv[1];
    if (Jp) {
        double dp[9]; // 3x3 row-major d p / d X
        const double ih = 1.0 / ph[3];
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) dp[i * 3 + j] = (Ed[i * 4 + j] - p[i] * Ed[3 * 4 + j]) * ih;
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j)
                Jp[r * 3 + j] = G[r * 3 + 0] * dp[0 * 3 + j] + G[r * 3 + 1] * dp[1 * 3 + j] + G[r * 3 + 2] * dp[2 * 3 + j];
    }
}

// R3: PoseOnlyAngleReprojectionError, Optimizer.h:163-182.  p = [R t] * Xh
inline void eval_pose_only_cached(const CamCache &cc, const float Xh[4], const float uv[2], const float K[9],
                                  double res[2], double *Jc) {
    const double X[3] = {(double)Xh[0], (double)Xh[1], (double)Xh[2]};
    const double w = (double)Xh[3];
    double p[3];
    for (int i = 0; i < 3; ++i) p[i] = cc.R[i] * X[0] + cc.R[i + 3] * X[1] + cc.R[i + 6] * X[2] + cc.t[i] * w;
    double pix[2], G[6];
    project(K, p, pix, Jc ? G : nullptr);
    res[0] = pix[0] - (double)uv[0];
    res[1] = pix[1] - (double)uv[1];
    if (Jc) {
        for (int j = 0; j < 3; ++j) {
            const double *D = cc.dR[j];
            double d[3];
            for (int i = 0; i < 3; ++i) d[i] = D[i] * X[0] + D[i + 3] * X[1] + D[i + 6] * X[2];
            for (int r = 0; r < 2; ++r) Jc[r * 6 + j] = G[r * 3 + 0] * d[0] + G[r * 3 + 1] * d[1] + G[r * 3 + 2] * d[2];
        }
        for (int r = 0; r < 2; ++r)
            for (int j = 0; j < 3; ++j) Jc[r * 6 + 3 + j] = G[r * 3 + j] * w;
    }
}

// ---- dual-number versions: literal transcription of the templated functors ----
template <typename T>
void functor_cam_point(const T *camera, const T *point, const float uv[2], const float K[9], T *res) {
    T R[9];
    angle_axis_to_rotation_T<T>(camera, R);
    T p[3];
    for (int i = 0; i < 3; ++i) p[i] = R[i] * point[0] + R[i + 3] * point[1] + R[i + 6] * point[2] + camera[3 + i] * T(1.0);
    T q[3];
    for (int i = 0; i < 3; ++i) q[i] = T((double)K[i * 3 + 0]) * p[0] + T((double)K[i * 3 + 1]) * p[1] + T((double)K[i * 3 + 2]) * p[2];
    res[0] = q[0] / q[2] - T((double)uv[0]);
    res[1] = q[1] / q[2] - T((double)uv[1]);
}
template <typename T>
void functor_point_only(const T *point, const float uv[2], const float E[16], const float K[9], T *res) {
    T ph[4];
    for (int r = 0; r < 4; ++r)
        ph[r] = T((double)E[r * 4 + 0]) * point[0] + T((double)E[r * 4 + 1]) * point[1] + T((double)E[r * 4 + 2]) * point[2] + T((double)E[r * 4 + 3]) * T(1.0);
    T p[3] = {ph[0] / ph[3], ph[1] / ph[3], ph[2] / ph[3]};
    T q[3];
    for (int i = 0; i < 3; ++i) q[i] = T((double)K[i * 3 + 0]) * p[0] + T((double)K[i * 3 + 1]) * p[1] + T((double)K[i * 3 + 2]) * p[2];
    res[0] = q[0] / q[2] - T((double)uv[0]);
    res[1] = q[1] / q[2] - T((double)uv[1]);
}
template <typename T>
void functor_pose_only(const T *camera, const float Xh[4], const float uv[2], const float K[9], T *res) {
    T R[9];
    angle_axis_to_rotation_T<T>(camera, R);
    T p[3];
    for (int i = 0; i < 3; ++i)
        p[i] = R[i] * T((double)Xh[0]) + R[i + 3] * T((double)Xh[1]) + R[i + 6] * T((double)Xh[2]) + camera[3 + i] * T((double)Xh[3]);
    T q[3];
    for (int i = 0; i < 3; ++i) q[i] = T((double)K[i * 3 + 0]) * p[0] + T((double)K[i * 3 + 1]) * p[1] + T((double)K[i * 3 + 2]) * p[2];
    res[0] = q[0] / q[2] - T((double)uv[0]);
    res[1] = q[1] / q[2] - T((double)uv[1]);
}

// ceres::HuberLoss::Evaluate (loss_function.cc); a = delta, b = a*a
inline void huber(double s, double a, double rho[3]) {
    const double b = a * a;
    if (s > b) {
        const double r = std::sqrt(s);
        rho[0] = 2.0 * a * r - b;
        rho[1] = std::max(std::numeric_limits<double>::min(), a / r);
        rho[2] = -rho[1] / (2.0 * s);
    } else {
        rho[0] = s; rho[1] = 1.0; rho[2] = 0.0;
    }
}

// ResidualBlock::Evaluate + Corrector (corrector.cc): for Huber rho''<=0 always, so alpha = 0:
// residual *= sqrt(rho'), jacobian *= sqrt(rho'); cost = 0.5 * rho(s).
inline double robustify(const orc_options &o, double res[2], double *J0, int n0, double *J1, int n1) {
    const double s = res[0] * res[0] + res[1] * res[1];
    if (!o.use_huber) return 0.5 * s;
    double rho[3];
    huber(s, o.huber_delta, rho);
    const double sr = std::sqrt(rho[1]);
    if (J0) for (int i = 0; i < n0; ++i) J0[i] *= sr;
    if (J1) for (int i = 0; i < n1; ++i) J1[i] *= sr;
    res[0] *= sr; res[1] *= sr;
    return 0.5 * rho[0];
}

// ------------------------------------------------------------------------------------------------
// Dense SPD machinery (Eigen::LLT stand-in).  A is n x n, row-major, UPPER triangle valid.
// Row-major upper == column-major lower, so we factor the column-major lower view L L^T in place.
// ------------------------------------------------------------------------------------------------
#define CM(i, j) A[(size_t)(j) * n + (i)] /* column-major lower view: i >= j */

int potrf_unblocked(int n0, int nb, double *A, int n) {
    // factor the nb x nb diagonal block starting at n0
    for (int j = n0; j < n0 + nb; ++j) {
        double d = CM(j, j);
        for (int k = n0; k < j; ++k) d -= CM(j, k) * CM(j, k);
        if (!(d > 0.0) || !std::isfinite(d)) return 1;
        d = std::sqrt(d);
        CM(j, j) = d;
        const double inv = 1.0 / d;
        for (int i = j + 1; i < n0 + nb; ++i) {
            double v = CM(i, j);
            for (int k = n0; k < j; ++k) v -= CM(i, k) * CM(j, k);
            CM(i, j) = v * inv;
        }
    }
    return 0;
}

// last (may be nullptr = dense): last[j] = index of the last row of column j of L that can be nonzero (monotone
// non-decreasing, >= j: the column envelope after symbolic elimination).  Rows below the panel's envelope are exact zeros
// before and after the panel's TRSM / SYRK, so skipping them changes no bit of the result.  orc_options.exploit_sparsity
// selects it: the CPU baseline is then given the same structural-zero skipping the CUDA path's tile plan has.
int cholesky_inplace(int n, double *A, int num_threads, const int *last = nullptr) {
    const int NB = 64;
    (void)num_threads;
    for (int k0 = 0; k0 < n; k0 += NB) {
        const int nb = std::min(NB, n - k0);
        if (potrf_unblocked(k0, nb, A, n)) return 1;
        const int r0 = k0 + nb;
        if (r0 >= n) break;
        const int nr = last ? std::min(n, last[k0 + nb - 1] + 1) : n;    // rows [r0, nr) of the panel can be nonzero
        if (r0 >= nr) continue;
        // TRSM: rows r0..nr of columns k0..k0+nb:  L21 = A21 * L11^{-T}
#pragma omp parallel for schedule(static) num_threads(num_threads)
        for (int ib = r0; ib < nr; ib += 256) {
            const int ie = std::min(nr, ib + 256);
            for (int j = k0; j < k0 + nb; ++j) {
                const double inv = 1.0 / CM(j, j);
                for (int k = k0; k < j; ++k) {
                    const double l = CM(j, k);
                    double *dst = &CM(ib, j);
                    const double *src = &CM(ib, k);
                    for (int i = 0; i < ie - ib; ++i) dst[i] -= src[i] * l;
                }
                double *dst = &CM(ib, j);
                for (int i = 0; i < ie - ib; ++i) dst[i] *= inv;
            }
        }
        // SYRK: A22 -= L21 L21^T (lower part), column tiles in parallel
        const int TJ = 32;
#pragma omp parallel for schedule(dynamic, 1) num_threads(num_threads)
        for (int jb = r0; jb < nr; jb += TJ) {
            const int je = std::min(nr, jb + TJ);
            for (int j = jb; j < je; ++j) {
                double *dst = &CM(j, j); // rows j..nr of column j
                const int len = nr - j;
                int k = k0;
                for (; k + 4 <= k0 + nb; k += 4) {
                    const double l0 = CM(j, k), l1 = CM(j, k + 1), l2 = CM(j, k + 2), l3 = CM(j, k + 3);
                    const double *s0 = &CM(j, k), *s1 = &CM(j, k + 1), *s2 = &CM(j, k + 2), *s3 = &CM(j, k + 3);
                    for (int i = 0; i < len; ++i) dst[i] -= s0[i] * l0 + s1[i] * l1 + s2[i] * l2 + s3[i] * l3;
                }
                for (; k < k0 + nb; ++k) {
                    const double l0 = CM(j, k);
                    const double *s0 = &CM(j, k);
                    for (int i = 0; i < len; ++i) dst[i] -= s0[i] * l0;
                }
            }
        }
    }
    return 0;
}

void cholesky_solve_inplace(int n, const double *A, double *b, const int *last = nullptr) {
    // L y = b
    for (int j = 0; j < n; ++j) {
        const double y = b[j] / CM(j, j);
        b[j] = y;
        const double *col = &CM(j, j);
        const int len = (last ? last[j] + 1 : n) - j;
        for (int i = 1; i < len; ++i) b[j + i] -= col[i] * y;
    }
    // L^T x = y
    for (int j = n - 1; j >= 0; --j) {
        const double *col = &CM(j, j);
        double v = b[j];
        const int len = (last ? last[j] + 1 : n) - j;
        for (int i = 1; i < len; ++i) v -= col[i] * b[j + i];
        b[j] = v / col[0];
    }
}
#undef CM

// Small fixed-size SPD inverse via Cholesky (InvertPSDMatrix with assume_full_rank_ete = true:
// m.selfadjointView<Upper>().llt().solve(Identity)).  Row-major n x n, n <= 6.
template <int N>
bool invert_spd(const double *M, double *inv) {
    double L[N * N];
    for (int i = 0; i < N * N; ++i) L[i] = 0.0;
    for (int j = 0; j < N; ++j) {
        double d = M[j * N + j];
        for (int k = 0; k < j; ++k) d -= L[j * N + k] * L[j * N + k];
        if (!(d > 0.0)) return false;
        d = std::sqrt(d);
        L[j * N + j] = d;
        for (int i = j + 1; i < N; ++i) {
            double v = M[j * N + i]; // upper triangle of M: M(j,i), i>j
            for (int k = 0; k < j; ++k) v -= L[i * N + k] * L[j * N + k];
            L[i * N + j] = v / d;
        }
    }
    for (int c = 0; c < N; ++c) {
        double y[N];
        for (int i = 0; i < N; ++i) {
            double v = (i == c) ? 1.0 : 0.0;
            for (int k = 0; k < i; ++k) v -= L[i * N + k] * y[k];
            y[i] = v / L[i * N + i];
        }
        for (int i = N - 1; i >= 0; --i) {
            double v = y[i];
            for (int k = i + 1; k < N; ++k) v -= L[k * N + i] * inv[k * N + c];
            inv[i * N + c] = v / L[i * N + i];
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------
// Generic trust-region minimiser (trust_region_minimizer.cc + levenberg_marquardt_strategy.cc)
// ------------------------------------------------------------------------------------------------
struct Model {
    int num_params = 0;
    int num_residuals = 0;
    double t_eval = 0, t_lin = 0;
    virtual ~Model() {}
    // cost at x.  with_jac: also (re)build residuals, corrected Jacobian and gradient = J^T r (unscaled J).
    virtual bool evaluate(const double *x, double *cost, bool with_jac) = 0;
    virtual const double *gradient() const = 0;
    virtual void squared_column_norms(double *out) const = 0; // of the stored Jacobian
    virtual void scale_columns(const double *scale) = 0;
    // solve (J^T J + D^2) y = J^T r.  returns 0 ok, 1 failure.
    virtual int solve(const double *D, double *y, int *lin_iters) = 0;
    virtual double model_cost_change(const double *step) const = 0; // -(J s).(r + J s / 2)
};

struct Minimizer {
    const orc_options &o;
    Model &m;
    std::vector<orc_iteration> iterations;
    orc_summary summary;
    std::vector<double> best_x;

    Minimizer(const orc_options &opt, Model &model) : o(opt), m(model) { std::memset(&summary, 0, sizeof(summary)); }

    void set_message(const char *msg) { std::snprintf(summary.message, sizeof(summary.message), "%s", msg); }

    void run(std::vector<double> &parameters) {
        const double t0 = now_s();
        const int n = m.num_params;
        std::vector<double> x = parameters, candidate_x(n), delta(n), step(n), scale(n, 1.0), diagonal(n), lm_diag(n);
        best_x = parameters;
        double x_norm = 0;
        for (int i = 0; i < n; ++i) x_norm += x[i] * x[i];
        x_norm = std::sqrt(x_norm);
        double x_cost = 0, candidate_cost = 0, minimum_cost = std::numeric_limits<double>::max();
        double model_cost_change = 0;
        // LM strategy state
        double radius = o.initial_trust_region_radius, decrease_factor = 2.0;
        bool reuse_diagonal = false;
        int num_consecutive_invalid = 0;

        orc_iteration it;
        std::memset(&it, 0, sizeof(it));
        it.eta = o.eta;
        it.trust_region_radius = radius;

        auto eval_grad_jac = [&](orc_iteration &its) -> bool {
            const double te = now_s();
            if (!m.evaluate(x.data(), &x_cost, true)) return false;
            m.t_eval += now_s() - te;
            its.cost = x_cost;
            if (o.jacobi_scaling) {
                if (its.iteration == 0) {
                    m.squared_column_norms(scale.data());
                    for (int i = 0; i < n; ++i) scale[i] = 1.0 / (1.0 + std::sqrt(scale[i]));
                }
                m.scale_columns(scale.data());
            }
            // |Plus(x, -g) - x| (trust_region_minimizer.cc EvaluateGradientAndJacobian)
            const double *g = m.gradient();
            double mx = 0, nn = 0;
            for (int i = 0; i < n; ++i) {
                const double d = x[i] - (x[i] + (-g[i]));
                mx = std::max(mx, std::fabs(d));
                nn += d * d;
            }
            its.gradient_max_norm = mx;
            its.gradient_norm = std::sqrt(nn);
            return true;
        };

        // ---- IterationZero ----
        it.iteration = 0;
        if (!eval_grad_jac(it)) {
            summary.termination_type = ORC_FAILURE;
            set_message("Initial residual and Jacobian evaluation failed.");
            finish(parameters, t0);
            return;
        }
        summary.initial_cost = x_cost;
        it.step_is_valid = 1;
        it.step_is_successful = 1;

        // ---- main loop: FinalizeIterationAndCheckIfMinimizerCanContinue at the top ----
        for (;;) {
            if (it.step_is_successful) {
                ++summary.num_successful_steps;
                if (x_cost < minimum_cost) {
                    minimum_cost = x_cost;
                    best_x = x;
                }
            } else {
                ++summary.num_unsuccessful_steps;
            }
            it.trust_region_radius = radius;
            iterations.push_back(it);

            if (it.iteration >= o.max_num_iterations) {
                summary.termination_type = ORC_NO_CONVERGENCE;
                set_message("Maximum number of iterations reached.");
                break;
            }
            if (it.step_is_successful && it.gradient_max_norm <= o.gradient_tolerance) {
                summary.termination_type = ORC_CONVERGENCE;
                set_message("Gradient tolerance reached.");
                break;
            }
            if (it.trust_region_radius <= o.min_trust_region_radius) {
                summary.termination_type = ORC_CONVERGENCE;
                set_message("Minimum trust region radius reached.");
                break;
            }

            const double prev_gmax = it.gradient_max_norm, prev_gnorm = it.gradient_norm;
            const int next = it.iteration + 1;
            std::memset(&it, 0, sizeof(it));
            it.iteration = next;
            it.eta = o.eta;

            // ---- ComputeTrustRegionStep (LevenbergMarquardtStrategy::ComputeStep) ----
            if (!reuse_diagonal) {
                m.squared_column_norms(diagonal.data());
                for (int i = 0; i < n; ++i) diagonal[i] = std::min(std::max(diagonal[i], o.min_lm_diagonal), o.max_lm_diagonal);
            }
            for (int i = 0; i < n; ++i) lm_diag[i] = std::sqrt(diagonal[i] / radius);
            int lin_iters = 0;
            const double tl = now_s();
            int lin_fail = m.solve(lm_diag.data(), step.data(), &lin_iters);
            m.t_lin += now_s() - tl;
            if (!lin_fail) {
                for (int i = 0; i < n; ++i)
                    if (!std::isfinite(step[i])) { lin_fail = 1; break; }
            }
            if (!lin_fail) for (int i = 0; i < n; ++i) step[i] = -step[i];
            reuse_diagonal = true;
            it.linear_solver_iterations = lin_iters;
            it.step_is_valid = 0;
            if (!lin_fail) {
                model_cost_change = m.model_cost_change(step.data());
                it.step_is_valid = (model_cost_change > 0.0) ? 1 : 0;
                if (it.step_is_valid) {
                    for (int i = 0; i < n; ++i) delta[i] = step[i] * scale[i];
                    num_consecutive_invalid = 0;
                }
            }

            if (!it.step_is_valid) {
                // ---- HandleInvalidStep ----
                ++num_consecutive_invalid;
                if (num_consecutive_invalid >= o.max_num_consecutive_invalid_steps) {
                    summary.termination_type = ORC_FAILURE;
                    set_message("Number of consecutive invalid steps more than Solver::Options::max_num_consecutive_invalid_steps.");
                    break;
                }
                // StepIsInvalid() == StepRejected(0)
                radius = radius / decrease_factor;
                decrease_factor *= 2.0;
                reuse_diagonal = true;
                it.cost = x_cost;
                it.cost_change = 0.0;
                it.gradient_max_norm = prev_gmax;
                it.gradient_norm = prev_gnorm;
                it.step_norm = 0.0;
                it.relative_decrease = 0.0;
                continue;
            }

            // ---- ComputeCandidatePointAndEvaluateCost ----
            for (int i = 0; i < n; ++i) candidate_x[i] = x[i] + delta[i];
            {
                const double te = now_s();
                if (!m.evaluate(candidate_x.data(), &candidate_cost, false) || !std::isfinite(candidate_cost))
                    candidate_cost = std::numeric_limits<double>::max();
                m.t_eval += now_s() - te;
            }
            // ---- ParameterToleranceReached ----
            {
                double s = 0;
                for (int i = 0; i < n; ++i) { const double d = x[i] - candidate_x[i]; s += d * d; }
                it.step_norm = std::sqrt(s);
                const double tol = o.parameter_tolerance * (x_norm + o.parameter_tolerance);
                if (it.step_norm <= tol) {
                    summary.termination_type = ORC_CONVERGENCE;
                    set_message("Parameter tolerance reached.");
                    break; // returns WITHOUT taking the step or recording the iteration
                }
            }
            // ---- FunctionToleranceReached ----
            it.cost_change = x_cost - candidate_cost;
            if (std::fabs(it.cost_change) <= o.function_tolerance * x_cost) {
                summary.termination_type = ORC_CONVERGENCE;
                set_message("Function tolerance reached.");
                break;
            }
            // ---- IsStepSuccessful (monotonic TrustRegionStepEvaluator) ----
            it.relative_decrease = (x_cost - candidate_cost) / model_cost_change;
            if (it.relative_decrease > o.min_relative_decrease) {
                // ---- HandleSuccessfulStep ----
                x = candidate_x;
                x_norm = 0;
                for (int i = 0; i < n; ++i) x_norm += x[i] * x[i];
                x_norm = std::sqrt(x_norm);
                if (!eval_grad_jac(it)) {
                    summary.termination_type = ORC_FAILURE;
                    set_message("Residual and Jacobian evaluation failed.");
                    break;
                }
                it.step_is_successful = 1;
                // StepAccepted
                radius = radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * it.relative_decrease - 1.0, 3));
                radius = std::min(o.max_trust_region_radius, radius);
                decrease_factor = 2.0;
                reuse_diagonal = false;
            } else {
                // ---- HandleUnsuccessfulStep ----
                it.step_is_successful = 0;
                it.cost = candidate_cost;
                it.gradient_max_norm = prev_gmax;
                it.gradient_norm = prev_gnorm;
                radius = radius / decrease_factor;
                decrease_factor *= 2.0;
                reuse_diagonal = true;
            }
        }
        finish(parameters, t0);
    }

    void finish(std::vector<double> &parameters, double t0) {
        parameters = best_x;
        // SetSummaryFinalCost (solver.cc): min over recorded iteration costs
        summary.final_cost = summary.initial_cost;
        for (const auto &i : iterations) summary.final_cost = std::min(summary.final_cost, i.cost);
        summary.num_iterations = (int)iterations.size();
        summary.seconds_total = now_s() - t0;
        summary.seconds_linear_solver = m.t_lin;
        summary.seconds_evaluate = m.t_eval;
    }
};

// ------------------------------------------------------------------------------------------------
// BA model: points are the e-blocks (Ceres' automatic Schur ordering), cameras the f-blocks.
// ------------------------------------------------------------------------------------------------
struct BAModel : Model {
    const orc_options &o;
    int Nc, Np, No;
    const uint8_t *cam_fixed;
    const float *fixed_extr, *K;
    const int32_t *obs_cam, *obs_pt;
    const float *obs_uv;
    // active parameter blocks (those that appear in at least one residual block)
    std::vector<int> pt_off, cam_off; // offset into x, or -1
    int n_pt_active = 0, n_cam_active = 0;
    std::vector<int> cam_slot;        // dense index among active free cameras, or -1
    // CSR by point and by camera over observations
    std::vector<int> pt_ptr, pt_obs, cam_ptr, cam_obs;
    // stored linearisation
    std::vector<double> res, Jc, Jp, grad;
    std::vector<double> x_lin;        // point at which the Jacobian was taken
    int nthreads;

    BAModel(const orc_options &opt, int Nc_, const uint8_t *cf, const float *fe, const float *K_, int Np_, int No_,
            const int32_t *oc, const int32_t *op, const float *uv)
        : o(opt), Nc(Nc_), Np(Np_), No(No_), cam_fixed(cf), fixed_extr(fe), K(K_), obs_cam(oc), obs_pt(op), obs_uv(uv) {
        nthreads = std::max(1, opt.num_threads);
        pt_off.assign(Np, -1);
        cam_off.assign(Nc, -1);
        cam_slot.assign(Nc, -1);
        std::vector<int> pc(Np, 0), ccnt(Nc, 0);
        for (int i = 0; i < No; ++i) { pc[obs_pt[i]]++; ccnt[obs_cam[i]]++; }
        int off = 0;
        for (int p = 0; p < Np; ++p) if (pc[p] > 0) { pt_off[p] = off; off += 3; n_pt_active++; }
        for (int c = 0; c < Nc; ++c) if (ccnt[c] > 0 && !cam_fixed[c]) { cam_off[c] = off; off += 6; cam_slot[c] = n_cam_active++; }
        num_params = off;
        num_residuals = 2 * No;
        pt_ptr.assign(Np + 1, 0);
        cam_ptr.assign(Nc + 1, 0);
        for (int p = 0; p < Np; ++p) pt_ptr[p + 1] = pt_ptr[p] + pc[p];
        for (int c = 0; c < Nc; ++c) cam_ptr[c + 1] = cam_ptr[c] + ccnt[c];
        pt_obs.resize(No); cam_obs.resize(No);
        std::vector<int> pf(pt_ptr.begin(), pt_ptr.end() - 1), cf2(cam_ptr.begin(), cam_ptr.end() - 1);
        for (int i = 0; i < No; ++i) { pt_obs[pf[obs_pt[i]]++] = i; cam_obs[cf2[obs_cam[i]]++] = i; }
        res.resize(2 * (size_t)No); Jc.resize(12 * (size_t)No); Jp.resize(6 * (size_t)No);
        grad.resize(num_params);
    }

    void pack(const double *cam, const double *pt, std::vector<double> &x) const {
        x.assign(num_params, 0.0);
        for (int p = 0; p < Np; ++p) if (pt_off[p] >= 0) for (int k = 0; k < 3; ++k) x[pt_off[p] + k] = pt[3 * p + k];
        for (int c = 0; c < Nc; ++c) if (cam_off[c] >= 0) for (int k = 0; k < 6; ++k) x[cam_off[c] + k] = cam[6 * c + k];
    }
    void unpack(const std::vector<double> &x, double *cam, double *pt) const {
        for (int p = 0; p < Np; ++p) if (pt_off[p] >= 0) for (int k = 0; k < 3; ++k) pt[3 * p + k] = x[pt_off[p] + k];
        for (int c = 0; c < Nc; ++c) if (cam_off[c] >= 0) for (int k = 0; k < 6; ++k) cam[6 * c + k] = x[cam_off[c] + k];
    }

    bool evaluate(const double *x, double *cost, bool with_jac) override {
        std::vector<CamCache> cc(Nc);
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int c = 0; c < Nc; ++c) if (cam_off[c] >= 0) cam_cache(x + cam_off[c], cc[c]);
        double total = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : total) num_threads(nthreads)
        for (int i = 0; i < No; ++i) {
            const int c = obs_cam[i], p = obs_pt[i];
            const double *X = x + pt_off[p];
            double r[2], jc[12], jp[6];
            if (cam_fixed[c]) {
                eval_point_only_d(X, obs_uv + 2 * i, fixed_extr + 16 * (size_t)c, K + 9 * (size_t)c, r, with_jac ? jp : nullptr);
                if (with_jac) for (int k = 0; k < 12; ++k) jc[k] = 0.0;
                total += robustify(o, r, nullptr, 0, with_jac ? jp : nullptr, 6);
            } else {
                eval_cam_point_cached(cc[c], X, obs_uv + 2 * i, K + 9 * (size_t)c, r, with_jac ? jc : nullptr, with_jac ? jp : nullptr);
                total += robustify(o, r, with_jac ? jc : nullptr, 12, with_jac ? jp : nullptr, 6);
            }
            if (with_jac) {
                res[2 * (size_t)i] = r[0]; res[2 * (size_t)i + 1] = r[1];
                std::memcpy(&Jc[12 * (size_t)i], jc, sizeof(jc));
                std::memcpy(&Jp[6 * (size_t)i], jp, sizeof(jp));
            }
        }
        *cost = total;
        if (with_jac) {
            // gradient = J^T r, per parameter block in observation order (deterministic)
#pragma omp parallel for schedule(static) num_threads(nthreads)
            for (int p = 0; p < Np; ++p) {
                if (pt_off[p] < 0) continue;
                double g[3] = {0, 0, 0};
                for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                    const size_t i = pt_obs[e];
                    for (int k = 0; k < 3; ++k) g[k] += Jp[6 * i + k] * res[2 * i] + Jp[6 * i + 3 + k] * res[2 * i + 1];
                }
                for (int k = 0; k < 3; ++k) grad[pt_off[p] + k] = g[k];
            }
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
            for (int c = 0; c < Nc; ++c) {
                if (cam_off[c] < 0) continue;
                double g[6] = {0, 0, 0, 0, 0, 0};
                for (int e = cam_ptr[c]; e < cam_ptr[c + 1]; ++e) {
                    const size_t i = cam_obs[e];
                    for (int k = 0; k < 6; ++k) g[k] += Jc[12 * i + k] * res[2 * i] + Jc[12 * i + 6 + k] * res[2 * i + 1];
                }
                for (int k = 0; k < 6; ++k) grad[cam_off[c] + k] = g[k];
            }
        }
        return std::isfinite(total);
    }
    const double *gradient() const override { return grad.data(); }

    void squared_column_norms(double *out) const override {
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int p = 0; p < Np; ++p) {
            if (pt_off[p] < 0) continue;
            double s[3] = {0, 0, 0};
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const size_t i = pt_obs[e];
                for (int k = 0; k < 3; ++k) s[k] += Jp[6 * i + k] * Jp[6 * i + k] + Jp[6 * i + 3 + k] * Jp[6 * i + 3 + k];
            }
            for (int k = 0; k < 3; ++k) out[pt_off[p] + k] = s[k];
        }
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
        for (int c = 0; c < Nc; ++c) {
            if (cam_off[c] < 0) continue;
            double s[6] = {0, 0, 0, 0, 0, 0};
            for (int e = cam_ptr[c]; e < cam_ptr[c + 1]; ++e) {
                const size_t i = cam_obs[e];
                for (int k = 0; k < 6; ++k) s[k] += Jc[12 * i + k] * Jc[12 * i + k] + Jc[12 * i + 6 + k] * Jc[12 * i + 6 + k];
            }
            for (int k = 0; k < 6; ++k) out[cam_off[c] + k] = s[k];
        }
    }

    void scale_columns(const double *scale) override {
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int i = 0; i < No; ++i) {
            const int c = obs_cam[i], p = obs_pt[i];
            const double *sp = scale + pt_off[p];
            for (int k = 0; k < 3; ++k) { Jp[6 * (size_t)i + k] *= sp[k]; Jp[6 * (size_t)i + 3 + k] *= sp[k]; }
            if (cam_off[c] >= 0) {
                const double *sc = scale + cam_off[c];
                for (int k = 0; k < 6; ++k) { Jc[12 * (size_t)i + k] *= sc[k]; Jc[12 * (size_t)i + 6 + k] *= sc[k]; }
            }
        }
    }

    // --- per-point elimination pieces (SchurEliminator) ---
    // inverse_ete (3x3) and g_p for point p given D
    inline void point_block(int p, const double *D, double inv[9], double g[3]) const {
        double ete[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        g[0] = g[1] = g[2] = 0;
        const double *Dp = D + pt_off[p];
        for (int k = 0; k < 3; ++k) ete[k * 3 + k] = Dp[k] * Dp[k];
        for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
            const size_t i = pt_obs[e];
            const double *J = &Jp[6 * i];
            for (int a = 0; a < 3; ++a) {
                for (int b = 0; b < 3; ++b) ete[a * 3 + b] += J[a] * J[b] + J[3 + a] * J[3 + b];
                g[a] += J[a] * res[2 * i] + J[3 + a] * res[2 * i + 1];
            }
        }
        if (!invert_spd<3>(ete, inv)) {
            for (int k = 0; k < 9; ++k) inv[k] = std::numeric_limits<double>::quiet_NaN();
        }
    }

    // Build reduced system over points [pb, pe).  S: n x n row-major with n = 6*n_cam_active (slot order), upper
    // blocks only.  add_diag: add U-side LM diagonal D_c^2 (once per system).  Full U_c (from all observations of
    // the shard's points) is accumulated.
    void schur_build(const double *D, int pb, int pe, bool add_diag, double *S, double *rhs) const {
        const int n = 6 * n_cam_active;
        // per-observation: Y_i = W_i * inv_p  (6x3), s_i = r_i - Jp_i * (inv_p g_p)
        std::vector<double> Y(18 * (size_t)No), sres(2 * (size_t)No);
        std::vector<double> invs(9 * (size_t)Np);
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int p = pb; p < pe; ++p) {
            if (pt_off[p] < 0) continue;
            double inv[9], g[3];
            point_block(p, D, inv, g);
            std::memcpy(&invs[9 * (size_t)p], inv, sizeof(inv));
            double ig[3];
            for (int a = 0; a < 3; ++a) ig[a] = inv[a * 3 + 0] * g[0] + inv[a * 3 + 1] * g[1] + inv[a * 3 + 2] * g[2];
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const size_t i = pt_obs[e];
                const double *J = &Jp[6 * i];
                for (int r = 0; r < 2; ++r) sres[2 * i + r] = res[2 * i + r] - (J[r * 3 + 0] * ig[0] + J[r * 3 + 1] * ig[1] + J[r * 3 + 2] * ig[2]);
                if (cam_off[obs_cam[i]] < 0) continue;
                const double *C = &Jc[12 * i];
                double W[18]; // 6x3 = Jc^T Jp
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 3; ++b) W[a * 3 + b] = C[a] * J[b] + C[6 + a] * J[3 + b];
                double *y = &Y[18 * i];
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 3; ++b) y[a * 3 + b] = W[a * 3 + 0] * inv[0 * 3 + b] + W[a * 3 + 1] * inv[1 * 3 + b] + W[a * 3 + 2] * inv[2 * 3 + b];
            }
        }
        // per camera block-row (deterministic, no atomics)
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
        for (int c = 0; c < Nc; ++c) {
            if (cam_off[c] < 0) continue;
            const int sc = cam_slot[c];
            double *Srow = S + (size_t)(6 * sc) * n;
            double r6[6] = {0, 0, 0, 0, 0, 0};
            if (add_diag) {
                const double *Dc = D + cam_off[c];
                for (int k = 0; k < 6; ++k) Srow[(size_t)k * n + 6 * sc + k] += Dc[k] * Dc[k];
            }
            for (int e = cam_ptr[c]; e < cam_ptr[c + 1]; ++e) {
                const size_t i = cam_obs[e];
                const int p = obs_pt[i];
                if (p < pb || p >= pe) continue;
                const double *C = &Jc[12 * i];
                // U_c += Jc^T Jc ; rhs += Jc^T s_i
                for (int a = 0; a < 6; ++a) {
                    for (int b = 0; b < 6; ++b) Srow[(size_t)a * n + 6 * sc + b] += C[a] * C[b] + C[6 + a] * C[6 + b];
                    r6[a] += C[a] * sres[2 * i] + C[6 + a] * sres[2 * i + 1];
                }
                const double *y = &Y[18 * i];
                for (int e2 = pt_ptr[p]; e2 < pt_ptr[p + 1]; ++e2) {
                    const size_t i2 = pt_obs[e2];
                    const int c2 = obs_cam[i2];
                    if (cam_off[c2] < 0) continue;
                    const int sc2 = cam_slot[c2];
                    if (sc2 < sc) continue;
                    const double *C2 = &Jc[12 * i2];
                    const double *J2 = &Jp[6 * i2];
                    // S[c][c2] -= Y_i * W_i2^T ; W_i2^T (3x6) = Jp2^T Jc2
                    for (int a = 0; a < 6; ++a) {
                        double *dst = Srow + (size_t)a * n + 6 * sc2;
                        for (int b = 0; b < 6; ++b) {
                            double w0 = C2[b] * J2[0] + C2[6 + b] * J2[3];
                            double w1 = C2[b] * J2[1] + C2[6 + b] * J2[4];
                            double w2 = C2[b] * J2[2] + C2[6 + b] * J2[5];
                            dst[b] -= y[a * 3 + 0] * w0 + y[a * 3 + 1] * w1 + y[a * 3 + 2] * w2;
                        }
                    }
                }
            }
            for (int k = 0; k < 6; ++k) rhs[6 * sc + k] += r6[k];
        }
        // symmetrise the diagonal blocks' lower part is NOT needed: only upper is read.
    }

    void back_substitute(const double *D, const double *yc /*slot order*/, double *y) const {
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int p = 0; p < Np; ++p) {
            if (pt_off[p] < 0) continue;
            double inv[9], g[3];
            point_block(p, D, inv, g); // g = E^T b; we need E^T (b - F z)
            double acc[3] = {0, 0, 0};
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const size_t i = pt_obs[e];
                const double *J = &Jp[6 * i];
                double s[2] = {res[2 * i], res[2 * i + 1]};
                const int c = obs_cam[i];
                if (cam_off[c] >= 0) {
                    const double *C = &Jc[12 * i];
                    const double *z = yc + 6 * cam_slot[c];
                    for (int r = 0; r < 2; ++r)
                        for (int k = 0; k < 6; ++k) s[r] -= C[r * 6 + k] * z[k];
                }
                for (int a = 0; a < 3; ++a) acc[a] += J[a] * s[0] + J[3 + a] * s[1];
            }
            for (int a = 0; a < 3; ++a) y[pt_off[p] + a] = inv[a * 3 + 0] * acc[0] + inv[a * 3 + 1] * acc[1] + inv[a * 3 + 2] * acc[2];
        }
        for (int c = 0; c < Nc; ++c)
            if (cam_off[c] >= 0) for (int k = 0; k < 6; ++k) y[cam_off[c] + k] = yc[6 * cam_slot[c] + k];
    }

    // ---- implicit Schur complement product (ImplicitSchurComplement::RightMultiply) ----
    struct Implicit {
        std::vector<double> invs;   // 9 per point
        std::vector<double> Ud;     // 36 per active cam: F^T F + D_f^2 block
        std::vector<double> Minv;   // 36 per active cam: inverse of block-diag(S)  (SCHUR_JACOBI)
        std::vector<double> rhs;
    };

    void implicit_setup(const double *D, Implicit &im) const {
        const int n = 6 * n_cam_active;
        im.invs.assign(9 * (size_t)Np, 0.0);
        im.Ud.assign(36 * (size_t)n_cam_active, 0.0);
        im.Minv.assign(36 * (size_t)n_cam_active, 0.0);
        im.rhs.assign(n, 0.0);
        std::vector<double> sres(2 * (size_t)No);
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int p = 0; p < Np; ++p) {
            if (pt_off[p] < 0) continue;
            double inv[9], g[3];
            point_block(p, D, inv, g);
            std::memcpy(&im.invs[9 * (size_t)p], inv, sizeof(inv));
            double ig[3];
            for (int a = 0; a < 3; ++a) ig[a] = inv[a * 3 + 0] * g[0] + inv[a * 3 + 1] * g[1] + inv[a * 3 + 2] * g[2];
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const size_t i = pt_obs[e];
                const double *J = &Jp[6 * i];
                for (int r = 0; r < 2; ++r) sres[2 * i + r] = res[2 * i + r] - (J[r * 3 + 0] * ig[0] + J[r * 3 + 1] * ig[1] + J[r * 3 + 2] * ig[2]);
            }
        }
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
        for (int c = 0; c < Nc; ++c) {
            if (cam_off[c] < 0) continue;
            const int sc = cam_slot[c];
            double U[36], Sd[36], r6[6] = {0, 0, 0, 0, 0, 0};
            for (int k = 0; k < 36; ++k) U[k] = 0.0;
            const double *Dc = D + cam_off[c];
            for (int k = 0; k < 6; ++k) U[k * 6 + k] = Dc[k] * Dc[k];
            for (int k = 0; k < 36; ++k) Sd[k] = 0.0;
            for (int e = cam_ptr[c]; e < cam_ptr[c + 1]; ++e) {
                const size_t i = cam_obs[e];
                const double *C = &Jc[12 * i];
                const double *J = &Jp[6 * i];
                const double *inv = &im.invs[9 * (size_t)obs_pt[i]];
                double W[18], Yv[18];
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 3; ++b) W[a * 3 + b] = C[a] * J[b] + C[6 + a] * J[3 + b];
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 3; ++b) Yv[a * 3 + b] = W[a * 3 + 0] * inv[0 * 3 + b] + W[a * 3 + 1] * inv[1 * 3 + b] + W[a * 3 + 2] * inv[2 * 3 + b];
                for (int a = 0; a < 6; ++a) {
                    for (int b = 0; b < 6; ++b) {
                        U[a * 6 + b] += C[a] * C[b] + C[6 + a] * C[6 + b];
                        Sd[a * 6 + b] += Yv[a * 3 + 0] * W[b * 3 + 0] + Yv[a * 3 + 1] * W[b * 3 + 1] + Yv[a * 3 + 2] * W[b * 3 + 2];
                    }
                    r6[a] += C[a] * sres[2 * i] + C[6 + a] * sres[2 * i + 1];
                }
            }
            std::memcpy(&im.Ud[36 * (size_t)sc], U, sizeof(U));
            double M[36];
            for (int k = 0; k < 36; ++k) M[k] = U[k] - Sd[k];
            if (!invert_spd<6>(M, &im.Minv[36 * (size_t)sc]))
                for (int k = 0; k < 36; ++k) im.Minv[36 * (size_t)sc + k] = std::numeric_limits<double>::quiet_NaN();
            for (int k = 0; k < 6; ++k) im.rhs[6 * sc + k] = r6[k];
        }
    }

    // out = S x = (F'F + D_f^2) x - F'E (E'E + D_e^2)^-1 E'F x
    void implicit_multiply(const Implicit &im, const double *x, double *out) const {
        std::vector<double> tp(3 * (size_t)Np, 0.0);
#pragma omp parallel for schedule(static) num_threads(nthreads)
        for (int p = 0; p < Np; ++p) {
            if (pt_off[p] < 0) continue;
            double acc[3] = {0, 0, 0};
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const size_t i = pt_obs[e];
                const int c = obs_cam[i];
                if (cam_off[c] < 0) continue;
                const double *C = &Jc[12 * i];
                const double *J = &Jp[6 * i];
                const double *z = x + 6 * cam_slot[c];
                double f[2] = {0, 0};
                for (int r = 0; r < 2; ++r) for (int k = 0; k < 6; ++k) f[r] += C[r * 6 + k] * z[k];
                for (int a = 0; a < 3; ++a) acc[a] += J[a] * f[0] + J[3 + a] * f[1];
            }
            const double *inv = &im.invs[9 * (size_t)p];
            for (int a = 0; a < 3; ++a) tp[3 * (size_t)p + a] = inv[a * 3 + 0] * acc[0] + inv[a * 3 + 1] * acc[1] + inv[a * 3 + 2] * acc[2];
        }
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
        for (int c = 0; c < Nc; ++c) {
            if (cam_off[c] < 0) continue;
            const int sc = cam_slot[c];
            const double *U = &im.Ud[36 * (size_t)sc];
            const double *z = x + 6 * sc;
            double r6[6];
            for (int a = 0; a < 6; ++a) { double v = 0; for (int b = 0; b < 6; ++b) v += U[a * 6 + b] * z[b]; r6[a] = v; }
            for (int e = cam_ptr[c]; e < cam_ptr[c + 1]; ++e) {
                const size_t i = cam_obs[e];
                const double *C = &Jc[12 * i];
                const double *J = &Jp[6 * i];
                const double *t = &tp[3 * (size_t)obs_pt[i]];
                double f[2];
                for (int r = 0; r < 2; ++r) f[r] = J[r * 3 + 0] * t[0] + J[r * 3 + 1] * t[1] + J[r * 3 + 2] * t[2];
                for (int a = 0; a < 6; ++a) r6[a] -= C[a] * f[0] + C[6 + a] * f[1];
            }
            for (int a = 0; a < 6; ++a) out[6 * sc + a] = r6[a];
        }
    }

    // ConjugateGradientsSolver::Solve (Ceres 2.0/2.1) with SCHUR_JACOBI preconditioner
    int pcg(const Implicit &im, double *xs, int *iters) const {
        const int n = 6 * n_cam_active;
        std::vector<double> r(n), p(n), z(n), tmp(n);
        const std::vector<double> &b = im.rhs;
        auto dot = [&](const double *a, const double *c) { double s = 0; for (int i = 0; i < n; ++i) s += a[i] * c[i]; return s; };
        auto precond = [&](const double *in, double *outv) {
            for (int c = 0; c < n_cam_active; ++c) {
                const double *M = &im.Minv[36 * (size_t)c];
                for (int a = 0; a < 6; ++a) { double v = 0; for (int k = 0; k < 6; ++k) v += M[a * 6 + k] * in[6 * c + k]; outv[6 * c + a] = v; }
            }
        };
        for (int i = 0; i < n; ++i) xs[i] = 0.0;
        implicit_multiply(im, xs, tmp.data());
        for (int i = 0; i < n; ++i) r[i] = b[i] - tmp[i];
        double norm_r = std::sqrt(dot(r.data(), r.data()));
        const double norm_b = std::sqrt(dot(b.data(), b.data()));
        const double tol_r = o.cg_r_tolerance * norm_b;
        *iters = 0;
        if (norm_r <= tol_r) return 0;
        double rho = 1.0;
        double Q0 = 0.0;
        { double s = 0; for (int i = 0; i < n; ++i) s += xs[i] * (b[i] + r[i]); Q0 = -1.0 * s; }
        int it = 1;
        for (;; ++it) {
            precond(r.data(), z.data());
            const double last_rho = rho;
            rho = dot(r.data(), z.data());
            if (rho == 0.0 || !std::isfinite(rho)) { *iters = it; return 1; }
            if (it == 1) p = z;
            else {
                const double beta = rho / last_rho;
                if (beta == 0.0 || !std::isfinite(beta)) { *iters = it; return 1; }
                for (int i = 0; i < n; ++i) p[i] = z[i] + beta * p[i];
            }
            std::vector<double> &q = z;
            implicit_multiply(im, p.data(), q.data());
            const double pq = dot(p.data(), q.data());
            if (pq <= 0 || std::isinf(pq)) break; // NO_CONVERGENCE: indefinite; keep current x
            const double alpha = rho / pq;
            if (std::isinf(alpha)) { *iters = it; return 1; }
            for (int i = 0; i < n; ++i) xs[i] += alpha * p[i];
            if (it % 10 == 0) { // residual_reset_period
                implicit_multiply(im, xs, tmp.data());
                for (int i = 0; i < n; ++i) r[i] = b[i] - tmp[i];
            } else {
                for (int i = 0; i < n; ++i) r[i] -= alpha * q[i];
            }
            double s = 0; for (int i = 0; i < n; ++i) s += xs[i] * (b[i] + r[i]);
            const double Q1 = -1.0 * s;
            const double zeta = it * (Q1 - Q0) / Q1;
            if (zeta < o.eta) break;
            Q0 = Q1;
            norm_r = std::sqrt(dot(r.data(), r.data()));
            if (norm_r <= tol_r) break;
            if (it >= o.max_linear_solver_iterations) break;
        }
        *iters = it;
        return 0;
    }

    // column envelope of the reduced camera matrix in slot order (row-major upper == column-major lower): row 6 sc + a
    // reaches column 6 * (largest slot sharing a landmark with sc) + 5, then the monotone closure that covers all fill
    std::vector<int> env_last;
    std::vector<double> S_env;
    void envelope_init(int n) {
        if ((int)env_last.size() == n) return;
        std::vector<int> far(n_cam_active);
        for (int s = 0; s < n_cam_active; ++s) far[s] = s;
        for (int p = 0; p < Np; ++p) {
            int lo = 1 << 30, hi = -1;
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const int sc = cam_slot[obs_cam[pt_obs[e]]];
                if (sc >= 0) { lo = std::min(lo, sc); hi = std::max(hi, sc); }
            }
            if (hi >= 0) far[lo] = std::max(far[lo], hi);       // every camera of the track reaches hi; lo's row is the longest
            for (int e = pt_ptr[p]; e < pt_ptr[p + 1]; ++e) {
                const int sc = cam_slot[obs_cam[pt_obs[e]]];
                if (sc >= 0) far[sc] = std::max(far[sc], hi);
            }
        }
        env_last.assign(n, 0);
        int run = 0;
        for (int s = 0; s < n_cam_active; ++s) {
            run = std::max(run, 6 * far[s] + 5);
            for (int a = 0; a < 6; ++a) env_last[6 * s + a] = run;
        }
        S_env.assign((size_t)n * n, 0.0);
    }

    int solve(const double *D, double *y, int *lin_iters) override {
        const int n = 6 * n_cam_active;
        std::vector<double> yc(std::max(n, 1), 0.0);
        *lin_iters = 0;
        if (n > 0) {
            if (o.linear_solver == ORC_DENSE_SCHUR && o.exploit_sparsity) {
                // Same arithmetic as the dense branch below on the entries that can be nonzero: the column envelope of S
                // (from camera co-visibility, closed under elimination) bounds every row range, S lives in a persistent
                // buffer of which only the envelope is cleared.  Bit-identical results (the skipped terms are exact zeros).
                envelope_init(n);
                std::vector<double> rhs(n, 0.0);
#pragma omp parallel for schedule(static) num_threads(nthreads)
                for (int j = 0; j < n; ++j) std::memset(&S_env[(size_t)j * n + j], 0, sizeof(double) * (size_t)(env_last[j] + 1 - j));
                schur_build(D, 0, Np, true, S_env.data(), rhs.data());
                *lin_iters = 1;
                int bad = 0;
#pragma omp parallel for schedule(static) reduction(| : bad) num_threads(nthreads)
                for (int j = 0; j < n; ++j)
                    for (int i = j; i <= env_last[j]; ++i) bad |= std::isnan(S_env[(size_t)j * n + i]) ? 1 : 0;
                if (bad) return 1;
                if (cholesky_inplace(n, S_env.data(), nthreads, env_last.data())) return 1;
                cholesky_solve_inplace(n, S_env.data(), rhs.data(), env_last.data());
                yc = rhs;
            } else if (o.linear_solver == ORC_DENSE_SCHUR) {
                std::vector<double> S((size_t)n * n, 0.0), rhs(n, 0.0);
                schur_build(D, 0, Np, true, S.data(), rhs.data());
                *lin_iters = 1;
                for (size_t k = 0; k < S.size(); ++k) if (std::isnan(S[k])) return 1;
                if (cholesky_inplace(n, S.data(), nthreads)) return 1;
                cholesky_solve_inplace(n, S.data(), rhs.data());
                yc = rhs;
            } else {
                Implicit im;
                implicit_setup(D, im);
                if (pcg(im, yc.data(), lin_iters)) return 1;
            }
        }
        back_substitute(D, yc.data(), y);
        return 0;
    }

    double model_cost_change(const double *step) const override {
        double total = 0.0;
#pragma omp parallel for schedule(static) reduction(+ : total) num_threads(nthreads)
        for (int i = 0; i < No; ++i) {
            const int c = obs_cam[i], p = obs_pt[i];
            const double *J = &Jp[6 * (size_t)i];
            const double *sp = step + pt_off[p];
            double mr[2];
            for (int r = 0; r < 2; ++r) mr[r] = J[r * 3 + 0] * sp[0] + J[r * 3 + 1] * sp[1] + J[r * 3 + 2] * sp[2];
            if (cam_off[c] >= 0) {
                const double *C = &Jc[12 * (size_t)i];
                const double *sc = step + cam_off[c];
                for (int r = 0; r < 2; ++r) for (int k = 0; k < 6; ++k) mr[r] += C[r * 6 + k] * sc[k];
            }
            total += mr[0] * (res[2 * (size_t)i] + mr[0] / 2.0) + mr[1] * (res[2 * (size_t)i + 1] + mr[1] / 2.0);
        }
        return -total;
    }
};

// ------------------------------------------------------------------------------------------------
// Pose-only model (one 6-vector; Ceres eliminates it as the single e-block: y = (J'J + D^2)^-1 J'r)
// ------------------------------------------------------------------------------------------------
struct PoseModel : Model {
    const orc_options &o;
    const float *K;
    int n_obs;
    const float *Xh, *uv;
    std::vector<int> idx; // active observations
    std::vector<double> res, J;
    double g[6];
    PoseModel(const orc_options &opt, const float *K_, int n, const float *Xh_, const float *uv_, const uint8_t *mask)
        : o(opt), K(K_), n_obs(n), Xh(Xh_), uv(uv_) {
        for (int i = 0; i < n; ++i) if (!mask || mask[i]) idx.push_back(i);
        num_params = 6;
        num_residuals = 2 * (int)idx.size();
        res.resize(2 * idx.size()); J.resize(12 * idx.size());
    }
    bool evaluate(const double *x, double *cost, bool with_jac) override {
        CamCache cc;
        cam_cache(x, cc);
        double total = 0;
        if (with_jac) for (int k = 0; k < 6; ++k) g[k] = 0;
        for (size_t e = 0; e < idx.size(); ++e) {
            const int i = idx[e];
            double r[2], jc[12];
            eval_pose_only_cached(cc, Xh + 4 * (size_t)i, uv + 2 * (size_t)i, K, r, with_jac ? jc : nullptr);
            total += robustify(o, r, with_jac ? jc : nullptr, 12, nullptr, 0);
            if (with_jac) {
                res[2 * e] = r[0]; res[2 * e + 1] = r[1];
                std::memcpy(&J[12 * e], jc, sizeof(jc));
                for (int k = 0; k < 6; ++k) g[k] += jc[k] * r[0] + jc[6 + k] * r[1];
            }
        }
        *cost = total;
        return std::isfinite(total);
    }
    const double *gradient() const override { return g; }
    void squared_column_norms(double *out) const override {
        for (int k = 0; k < 6; ++k) out[k] = 0;
        for (size_t e = 0; e < idx.size(); ++e)
            for (int k = 0; k < 6; ++k) out[k] += J[12 * e + k] * J[12 * e + k] + J[12 * e + 6 + k] * J[12 * e + 6 + k];
    }
    void scale_columns(const double *s) override {
        for (size_t e = 0; e < idx.size(); ++e)
            for (int k = 0; k < 6; ++k) { J[12 * e + k] *= s[k]; J[12 * e + 6 + k] *= s[k]; }
    }
    int solve(const double *D, double *y, int *lin_iters) override {
        double A[36], b[6], inv[36];
        for (int k = 0; k < 36; ++k) A[k] = 0;
        for (int k = 0; k < 6; ++k) { A[k * 6 + k] = D[k] * D[k]; b[k] = 0; }
        for (size_t e = 0; e < idx.size(); ++e) {
            const double *C = &J[12 * e];
            for (int a = 0; a < 6; ++a) {
                for (int c = 0; c < 6; ++c) A[a * 6 + c] += C[a] * C[c] + C[6 + a] * C[6 + c];
                b[a] += C[a] * res[2 * e] + C[6 + a] * res[2 * e + 1];
            }
        }
        *lin_iters = 1;
        if (!invert_spd<6>(A, inv)) return 1;
        for (int a = 0; a < 6; ++a) { double v = 0; for (int c = 0; c < 6; ++c) v += inv[a * 6 + c] * b[c]; y[a] = v; }
        return 0;
    }
    double model_cost_change(const double *step) const override {
        double total = 0;
        for (size_t e = 0; e < idx.size(); ++e) {
            const double *C = &J[12 * e];
            for (int r = 0; r < 2; ++r) {
                double mr = 0;
                for (int k = 0; k < 6; ++k) mr += C[r * 6 + k] * step[k];
                total += mr * (res[2 * e + r] + mr / 2.0);
            }
        }
        return -total;
    }
};

void copy_out(const Minimizer &mz, orc_summary *summary, orc_iteration *iters, int cap) {
    if (summary) *summary = mz.summary;
    if (iters) {
        const int n = std::min<int>(cap, (int)mz.iterations.size());
        for (int i = 0; i < n; ++i) iters[i] = mz.iterations[i];
    }
}

} // namespace

// ================================================================================================
// C API
// ================================================================================================
extern "C" {

void orc_default_options(orc_options *o) {
    std::memset(o, 0, sizeof(*o));
    o->max_num_iterations = 200;          // BAOptimizer(): m_nMaxItPerBA(200), Optimizer.h:203
    o->num_threads = 4;                   // Optimizer.cpp:88
    o->linear_solver = ORC_DENSE_SCHUR;   // Optimizer.cpp:85
    o->jacobi_scaling = 1;
    o->max_num_consecutive_invalid_steps = 5;
    o->max_linear_solver_iterations = 500;
    o->use_huber = 1;
    o->function_tolerance = 1e-6;
    o->gradient_tolerance = 1e-10;
    o->parameter_tolerance = 1e-8;
    o->initial_trust_region_radius = 1e4;
    o->max_trust_region_radius = 1e16;
    o->min_trust_region_radius = 1e-32;
    o->min_relative_decrease = 1e-3;
    o->min_lm_diagonal = 1e-6;
    o->max_lm_diagonal = 1e32;
    o->huber_delta = std::sqrt(5.991);    // Optimizer.cpp:312
    o->eta = 1e-1;
    o->cg_r_tolerance = -1.0;
}

void orc_angle_axis_to_rotation(const double w[3], double R[9]) { angle_axis_to_rotation_T<double>(w, R); }

// ceres::RotationMatrixToAngleAxis = RotationMatrixToQuaternion + QuaternionToAngleAxis (rotation.h)
void orc_rotation_to_angle_axis(const double R[9], double aa[3]) {
    auto Rm = [&](int i, int j) { return R[i + 3 * j]; };
    double q[4];
    const double trace = Rm(0, 0) + Rm(1, 1) + Rm(2, 2);
    if (trace >= 0.0) {
        double t = std::sqrt(trace + 1.0);
        q[0] = 0.5 * t;
        t = 0.5 / t;
        q[1] = (Rm(2, 1) - Rm(1, 2)) * t;
        q[2] = (Rm(0, 2) - Rm(2, 0)) * t;
        q[3] = (Rm(1, 0) - Rm(0, 1)) * t;
    } else {
        int i = 0;
        if (Rm(1, 1) > Rm(0, 0)) i = 1;
        if (Rm(2, 2) > Rm(i, i)) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double t = std::sqrt(Rm(i, i) - Rm(j, j) - Rm(k, k) + 1.0);
        q[i + 1] = 0.5 * t;
        t = 0.5 / t;
        q[0] = (Rm(k, j) - Rm(j, k)) * t;
        q[j + 1] = (Rm(j, i) + Rm(i, j)) * t;
        q[k + 1] = (Rm(k, i) + Rm(i, k)) * t;
    }
    const double q1 = q[1], q2 = q[2], q3 = q[3];
    const double sin_sq = q1 * q1 + q2 * q2 + q3 * q3;
    if (sin_sq > 0.0) {
        const double sin_theta = std::sqrt(sin_sq);
        const double cos_theta = q[0];
        const double two_theta = 2.0 * ((cos_theta < 0.0) ? std::atan2(-sin_theta, -cos_theta) : std::atan2(sin_theta, cos_theta));
        const double k = two_theta / sin_theta;
        aa[0] = q1 * k; aa[1] = q2 * k; aa[2] = q3 * k;
    } else {
        aa[0] = q1 * 2.0; aa[1] = q2 * 2.0; aa[2] = q3 * 2.0;
    }
}

// General 4x4 inverse in float32 (Eigen Matrix4f::inverse() stand-in).  Eigen's compute_inverse_size4 builds the
// adjugate from the six 2x2 minors of rows (0,1) and the six of rows (2,3); same formulation here, row-major.
void orc_inverse4x4_f32(const float m[16], float out[16]) {
    auto M = [&](int r, int c) { return m[4 * r + c]; };
    // minors of the top two rows (columns i<j) and of the bottom two rows
    const float t01 = M(0, 0) * M(1, 1) - M(1, 0) * M(0, 1);
    const float t02 = M(0, 0) * M(1, 2) - M(1, 0) * M(0, 2);
    const float t03 = M(0, 0) * M(1, 3) - M(1, 0) * M(0, 3);
    const float t12 = M(0, 1) * M(1, 2) - M(1, 1) * M(0, 2);
    const float t13 = M(0, 1) * M(1, 3) - M(1, 1) * M(0, 3);
    const float t23 = M(0, 2) * M(1, 3) - M(1, 2) * M(0, 3);
    const float b23 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3);
    const float b13 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3);
    const float b12 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2);
    const float b03 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3);
    const float b02 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2);
    const float b01 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1);
    const float det = t01 * b23 - t02 * b13 + t03 * b12 + t12 * b03 - t13 * b02 + t23 * b01;
    const float idet = 1.0f / det;
    out[0] = (M(1, 1) * b23 - M(1, 2) * b13 + M(1, 3) * b12) * idet;
    out[1] = (-M(0, 1) * b23 + M(0, 2) * b13 - M(0, 3) * b12) * idet;
    out[2] = (M(3, 1) * t23 - M(3, 2) * t13 + M(3, 3) * t12) * idet;
    out[3] = (-M(2, 1) * t23 + M(2, 2) * t13 - M(2, 3) * t12) * idet;
    out[4] = (-M(1, 0) * b23 + M(1, 2) * b03 - M(1, 3) * b02) * idet;
    out[5] = (M(0, 0) * b23 - M(0, 2) * b03 + M(0, 3) * b02) * idet;
    out[6] = (-M(3, 0) * t23 + M(3, 2) * t03 - M(3, 3) * t02) * idet;
    out[7] = (M(2, 0) * t23 - M(2, 2) * t03 + M(2, 3) * t02) * idet;
    out[8] = (M(1, 0) * b13 - M(1, 1) * b03 + M(1, 3) * b01) * idet;
    out[9] = (-M(0, 0) * b13 + M(0, 1) * b03 - M(0, 3) * b01) * idet;
    out[10] = (M(3, 0) * t13 - M(3, 1) * t03 + M(3, 3) * t01) * idet;
    out[11] = (-M(2, 0) * t13 + M(2, 1) * t03 - M(2, 3) * t01) * idet;
    out[12] = (-M(1, 0) * b12 + M(1, 1) * b02 - M(1, 2) * b01) * idet;
    out[13] = (M(0, 0) * b12 - M(0, 1) * b02 + M(0, 2) * b01) * idet;
    out[14] = (-M(3, 0) * t12 + M(3, 1) * t02 - M(3, 2) * t01) * idet;
    out[15] = (M(2, 0) * t12 - M(2, 1) * t02 + M(2, 2) * t01) * idet;
}

void orc_eval_cam_point(const double cam[6], const double pt[3], const float uv[2], const float K[9],
                        double res[2], double *Jc, double *Jp, int use_jets) {
    if (!use_jets) {
        CamCache cc;
        cam_cache(cam, cc);
        eval_cam_point_cached(cc, pt, uv, K, res, Jc, Jp);
        return;
    }
    typedef Jet<9> J9;
    J9 c[6], p[3], r[2];
    for (int i = 0; i < 6; ++i) c[i] = J9(cam[i], i);
    for (int i = 0; i < 3; ++i) p[i] = J9(pt[i], 6 + i);
    functor_cam_point<J9>(c, p, uv, K, r);
    for (int k = 0; k < 2; ++k) {
        res[k] = r[k].a;
        if (Jc) for (int i = 0; i < 6; ++i) Jc[k * 6 + i] = r[k].v[i];
        if (Jp) for (int i = 0; i < 3; ++i) Jp[k * 3 + i] = r[k].v[6 + i];
    }
}

void orc_eval_point_only(const double pt[3], const float uv[2], const float E[16], const float K[9],
                         double res[2], double *Jp, int use_jets) {
    if (!use_jets) { eval_point_only_d(pt, uv, E, K, res, Jp); return; }
    typedef Jet<3> J3;
    J3 p[3], r[2];
    for (int i = 0; i < 3; ++i) p[i] = J3(pt[i], i);
    functor_point_only<J3>(p, uv, E, K, r);
    for (int k = 0; k < 2; ++k) {
        res[k] = r[k].a;
        if (Jp) for (int i = 0; i < 3; ++i) Jp[k * 3 + i] = r[k].v[i];
    }
}

void orc_eval_pose_only(const double cam[6], const float Xh[4], const float uv[2], const float K[9],
                        double res[2], double *Jc, int use_jets) {
    if (!use_jets) {
        CamCache cc;
        cam_cache(cam, cc);
        eval_pose_only_cached(cc, Xh, uv, K, res, Jc);
        return;
    }
    typedef Jet<6> J6;
    J6 c[6], r[2];
    for (int i = 0; i < 6; ++i) c[i] = J6(cam[i], i);
    functor_pose_only<J6>(c, Xh, uv, K, r);
    for (int k = 0; k < 2; ++k) {
        res[k] = r[k].a;
        if (Jc) for (int i = 0; i < 6; ++i) Jc[k * 6 + i] = r[k].v[i];
    }
}

void orc_huber(double s, double a, double rho[3]) { huber(s, a, rho); }

int orc_evaluate(const orc_options *opt, int Nc, const double *cam, const uint8_t *cam_fixed,
                 const float *fixed_extr, const float *K, int Np, const double *pt, int No,
                 const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv,
                 double *cost, double *residuals, double *jac_cam, double *jac_pt,
                 double *grad_cam, double *grad_pt) {
    BAModel m(*opt, Nc, cam_fixed, fixed_extr, K, Np, No, obs_cam, obs_pt, obs_uv);
    std::vector<double> x;
    m.pack(cam, pt, x);
    double c = 0;
    if (!m.evaluate(x.data(), &c, true)) return 1;
    if (cost) *cost = c;
    if (residuals) std::memcpy(residuals, m.res.data(), sizeof(double) * 2 * (size_t)No);
    if (jac_cam) std::memcpy(jac_cam, m.Jc.data(), sizeof(double) * 12 * (size_t)No);
    if (jac_pt) std::memcpy(jac_pt, m.Jp.data(), sizeof(double) * 6 * (size_t)No);
    if (grad_cam) {
        std::memset(grad_cam, 0, sizeof(double) * 6 * (size_t)Nc);
        for (int c2 = 0; c2 < Nc; ++c2) if (m.cam_off[c2] >= 0) for (int k = 0; k < 6; ++k) grad_cam[6 * c2 + k] = m.grad[m.cam_off[c2] + k];
    }
    if (grad_pt) {
        std::memset(grad_pt, 0, sizeof(double) * 3 * (size_t)Np);
        for (int p = 0; p < Np; ++p) if (m.pt_off[p] >= 0) for (int k = 0; k < 3; ++k) grad_pt[3 * p + k] = m.grad[m.pt_off[p] + k];
    }
    return 0;
}

int orc_solve(const orc_options *opt, int Nc, double *cam, const uint8_t *cam_fixed,
              const float *fixed_extr, const float *K, int Np, double *pt, int No,
              const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv,
              orc_summary *summary, orc_iteration *iters, int iters_cap) {
    BAModel m(*opt, Nc, cam_fixed, fixed_extr, K, Np, No, obs_cam, obs_pt, obs_uv);
    std::vector<double> x;
    m.pack(cam, pt, x);
    Minimizer mz(*opt, m);
    mz.run(x);
    m.unpack(x, cam, pt);
    copy_out(mz, summary, iters, iters_cap);
    return 0;
}

int orc_schur_system(const orc_options *opt, double radius, int Nc, const double *cam,
                     const uint8_t *cam_fixed, const float *fixed_extr, const float *K, int Np,
                     const double *pt, int No, const int32_t *obs_cam, const int32_t *obs_pt,
                     const float *obs_uv, int pt_begin, int pt_end, double *S, double *rhs, double *cost) {
    // NOTE: for sharded use every shard must see the same Jacobi scale; this helper takes the WHOLE problem,
    // linearises it once (iteration-0 semantics) and accumulates only the landmarks in [pt_begin, pt_end).
    // The camera LM diagonal D_c^2 is added by the shard that owns landmark 0.
    // S is indexed by ALL cameras (6*Nc; inactive/fixed rows stay zero) so shards agree on the layout.
    BAModel m(*opt, Nc, cam_fixed, fixed_extr, K, Np, No, obs_cam, obs_pt, obs_uv);
    std::vector<double> x;
    m.pack(cam, pt, x);
    double c = 0;
    if (!m.evaluate(x.data(), &c, true)) return 1;
    const int n = m.num_params;
    std::vector<double> scale(n, 1.0), diag(n), D(n);
    if (opt->jacobi_scaling) {
        m.squared_column_norms(scale.data());
        for (int i = 0; i < n; ++i) scale[i] = 1.0 / (1.0 + std::sqrt(scale[i]));
        m.scale_columns(scale.data());
    }
    m.squared_column_norms(diag.data());
    for (int i = 0; i < n; ++i) D[i] = std::sqrt(std::min(std::max(diag[i], opt->min_lm_diagonal), opt->max_lm_diagonal) / radius);
    const int ns = 6 * m.n_cam_active;
    std::vector<double> Ss((size_t)ns * ns, 0.0), rs(ns, 0.0);
    m.schur_build(D.data(), pt_begin, pt_end, pt_begin == 0, Ss.data(), rs.data());
    const int N = 6 * Nc;
    std::memset(S, 0, sizeof(double) * (size_t)N * N);
    std::memset(rhs, 0, sizeof(double) * N);
    std::vector<int> slot2cam(m.n_cam_active);
    for (int cc = 0; cc < Nc; ++cc) if (m.cam_slot[cc] >= 0) slot2cam[m.cam_slot[cc]] = cc;
    for (int a = 0; a < ns; ++a) {
        const int ga = 6 * slot2cam[a / 6] + a % 6;
        rhs[ga] = rs[a];
        for (int b = 0; b < ns; ++b) S[(size_t)ga * N + 6 * slot2cam[b / 6] + b % 6] = Ss[(size_t)a * ns + b];
    }
    if (cost) {
        // partial cost of the shard
        double pc = 0;
        for (int i = 0; i < No; ++i) {
            if (obs_pt[i] < pt_begin || obs_pt[i] >= pt_end) continue;
            double r[2];
            const int cc = obs_cam[i];
            if (cam_fixed[cc]) eval_point_only_d(pt + 3 * (size_t)obs_pt[i], obs_uv + 2 * (size_t)i, fixed_extr + 16 * (size_t)cc, K + 9 * (size_t)cc, r, nullptr);
            else { CamCache ch; cam_cache(cam + 6 * (size_t)cc, ch); eval_cam_point_cached(ch, pt + 3 * (size_t)obs_pt[i], obs_uv + 2 * (size_t)i, K + 9 * (size_t)cc, r, nullptr, nullptr); }
            pc += robustify(*opt, r, nullptr, 0, nullptr, 0);
        }
        *cost = pc;
    }
    return 0;
}

int orc_solve_pose_only_batch(const orc_options *opt, int B, double *cam, const float *K,
                              const int32_t *prob_ptr, const float *Xh, const float *uv,
                              const uint8_t *obs_mask, orc_summary *summaries, orc_iteration *iters, int iters_cap) {
    const int nthreads = std::max(1, opt->num_threads);
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
    for (int b = 0; b < B; ++b) {
        const int o0 = prob_ptr[b], n = prob_ptr[b + 1] - prob_ptr[b];
        PoseModel m(*opt, K + 9 * (size_t)b, n, Xh + 4 * (size_t)o0, uv + 2 * (size_t)o0, obs_mask ? obs_mask + o0 : nullptr);
        std::vector<double> x(cam + 6 * (size_t)b, cam + 6 * (size_t)b + 6);
        Minimizer mz(*opt, m);
        mz.run(x);
        for (int k = 0; k < 6; ++k) cam[6 * (size_t)b + k] = x[k];
        copy_out(mz, summaries ? summaries + b : nullptr, iters ? iters + (size_t)b * iters_cap : nullptr, iters_cap);
    }
    return 0;
}

// BAOptimizer::pruneCorrespondences, Optimizer.cpp:6-79 (float32 throughout; the "chi square" is a plain
// pixel distance divided by 1.2^octave, compared against 5.991 -- quirk kept)
void orc_prune_frame(const float pose[16], const float K[9], int n, const float *kp_xy, const int32_t *kp_octave,
                     const float *pt_xyz, const float *pt_min_dist, const float *pt_max_dist,
                     const uint8_t *has_point, uint8_t *outlier, int consider_outlier) {
    float extr[16];
    orc_inverse4x4_f32(pose, extr);
    const float wp[3] = {pose[3], pose[7], pose[11]}; // Frame::getWorldPos(): m_pose.block(0,3,3,1)
    for (int i = 0; i < n; ++i) {
        if (!(consider_outlier || !outlier[i])) continue;
        if (!has_point[i]) continue;
        const float chiThresh = 5.991f;
        const float invSigmaP = (float)(1.0 / std::pow(1.2, kp_octave[i]));
        const float *X = pt_xyz + 3 * (size_t)i;
        float ph[4];
        for (int r = 0; r < 4; ++r) ph[r] = extr[r * 4 + 0] * X[0] + extr[r * 4 + 1] * X[1] + extr[r * 4 + 2] * X[2] + extr[r * 4 + 3] * 1.0f;
        const float cs[3] = {ph[0] / ph[3], ph[1] / ph[3], ph[2] / ph[3]};
        if (cs[2] <= 0) { outlier[i] = 1; continue; }
        const float dx = X[0] - wp[0], dy = X[1] - wp[1], dz = X[2] - wp[2];
        const float worldDist = std::sqrt(dx * dx + dy * dy + dz * dz);
        if (worldDist > pt_max_dist[i] || worldDist < pt_min_dist[i]) { outlier[i] = 1; continue; }
        float q[3];
        for (int r = 0; r < 3; ++r) q[r] = K[r * 3 + 0] * cs[0] + K[r * 3 + 1] * cs[1] + K[r * 3 + 2] * cs[2];
        const float px = q[0] / q[2], py = q[1] / q[2];
        const float ex = px - kp_xy[2 * i], ey = py - kp_xy[2 * i + 1];
        const float dist = std::sqrt(ex * ex + ey * ey);
        const float chi = dist * invSigmaP;
        if (chi > chiThresh) { outlier[i] = 1; continue; }
        outlier[i] = 0;
    }
}

// Optimizer.cpp:296-299: extr = pose.inverse() (f32); R (f32 -> f64) -> angle-axis; t f32 -> f64
void orc_pose_to_cam(const float pose[16], double cam[6], float extr_out[16]) {
    float extr[16];
    orc_inverse4x4_f32(pose, extr);
    double R[9];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[i + 3 * j] = (double)extr[i * 4 + j];
    orc_rotation_to_angle_axis(R, cam);
    for (int i = 0; i < 3; ++i) cam[3 + i] = (double)extr[i * 4 + 3];
    if (extr_out) std::memcpy(extr_out, extr, sizeof(extr));
}

// Optimizer.cpp:262-267: angle-axis -> R (f64) -> f32 4x4 extr -> inverse -> pose
void orc_cam_to_pose(const double cam[6], float pose_out[16]) {
    double R[9];
    orc_angle_axis_to_rotation(cam, R);
    float extr[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) extr[i * 4 + j] = (float)R[i + 3 * j];
        extr[i * 4 + 3] = (float)cam[3 + i];
    }
    orc_inverse4x4_f32(extr, pose_out);
}

int orc_cholesky_solve(int n, double *A, double *b, int num_threads) {
    if (cholesky_inplace(n, A, std::max(1, num_threads))) return 1;
    cholesky_solve_inplace(n, A, b);
    return 0;
}

} // extern "C"
