/*
 * ba_oracle.h -- CPU restatement of the reference's BA optimisation path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (bundleadjustment_b200/, include/ba_b200.h) never links, imports or calls it.
 *
 * PARITY UNPINNED: the arithmetic of this path lives in Ceres Solver (un-vendored,
 * version unpinned: ba_project/CMakeLists.txt:22, Dockerfile:17; inferred 2.0/2.1),
 * which is absent from /root/reference and from this image, and the reference has no
 * tests, golden vectors or fixtures for the path.  This file restates
 * Ceres' published algorithm (rotation.h, loss_function.cc, corrector.cc,
 * levenberg_marquardt_strategy.cc, trust_region_minimizer.cc, schur_eliminator_impl.h,
 * schur_complement_solver.cc, conjugate_gradients_solver.cc) anchored on the
 * reference's own call sites (ba_project/src/ba/Optimizer.h:49-194,
 * ba_project/src/ba/Optimizer.cpp:6-90, 216-333, 417-498, 500-698).
 *
 * Layout conventions shared with include/ba_b200.h:
 *   cam[6]        angle-axis (3) + translation (3), world->camera, f64
 *   K[9]          intrinsics, f32, ROW-major 3x3
 *   fixed_extr[16] world->camera 4x4, f32, ROW-major (only read where cam_fixed != 0)
 *   pt[3]         landmark xyz f64
 *   obs_uv[2]     keypoint pixel f32
 */
#ifndef BA_ORACLE_H
#define BA_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { ORC_DENSE_SCHUR = 0, ORC_ITERATIVE_SCHUR = 1 };
enum { ORC_CONVERGENCE = 0, ORC_NO_CONVERGENCE = 1, ORC_FAILURE = 2 };

/* Subset of ceres::Solver::Options the reference relies on (Optimizer.cpp:80-90);
 * everything the reference does not set is the Ceres default. */
typedef struct orc_options {
    int32_t max_num_iterations;        /* m_nMaxItPerBA, Optimizer.cpp:87 */
    int32_t num_threads;               /* 4, Optimizer.cpp:88 */
    int32_t linear_solver;             /* ORC_DENSE_SCHUR (Optimizer.cpp:85) */
    int32_t jacobi_scaling;            /* 1 */
    int32_t max_num_consecutive_invalid_steps; /* 5 */
    int32_t max_linear_solver_iterations;      /* 500 (CG) */
    int32_t use_huber;                 /* 1: HuberLoss(sqrt(5.991)), Optimizer.cpp:312 */
    int32_t exploit_sparsity;          /* 0 = factorise S densely, as Ceres' DENSE_SCHUR does; 1 = skip the structurally zero
                                          part of S (column envelope): same bits, the sparsity the CUDA tile plan exploits */
    double function_tolerance;         /* 1e-6 */
    double gradient_tolerance;         /* 1e-10 */
    double parameter_tolerance;        /* 1e-8 */
    double initial_trust_region_radius;/* 1e4 */
    double max_trust_region_radius;    /* 1e16 */
    double min_trust_region_radius;    /* 1e-32 */
    double min_relative_decrease;      /* 1e-3 */
    double min_lm_diagonal;            /* 1e-6 */
    double max_lm_diagonal;            /* 1e32 */
    double huber_delta;                /* sqrt(5.991) computed in double */
    double eta;                        /* 0.1: CG q_tolerance */
    double cg_r_tolerance;             /* -1 (Ceres: disabled); >0 enables ||r||<=tol*||b|| */
} orc_options;

typedef struct orc_iteration {
    int32_t iteration;
    int32_t step_is_valid;
    int32_t step_is_successful;
    int32_t linear_solver_iterations;
    double cost;
    double cost_change;
    double gradient_max_norm;
    double gradient_norm;
    double step_norm;
    double relative_decrease;
    double trust_region_radius;
    double eta;
} orc_iteration;

typedef struct orc_summary {
    int32_t termination_type;
    int32_t num_iterations;            /* entries written to the iteration array */
    int32_t num_successful_steps;
    int32_t num_unsuccessful_steps;
    double initial_cost;
    double final_cost;
    double seconds_total;
    double seconds_linear_solver;
    double seconds_evaluate;
    char message[128];
} orc_summary;

void orc_default_options(orc_options *o);

/* ---- geometry (ceres/rotation.h restatement) ---- */
void orc_angle_axis_to_rotation(const double w[3], double R_colmajor[9]);
void orc_rotation_to_angle_axis(const double R_colmajor[9], double w[3]);
/* general 4x4 float inverse by cofactors (Eigen Matrix4f::inverse() stand-in), row-major */
void orc_inverse4x4_f32(const float M[16], float out[16]);

/* ---- cost functors: residual + analytic Jacobians; use_jets=1 -> dual numbers ---- */
/* Jc is 2x6 row-major, Jp is 2x3 row-major; either may be NULL */
void orc_eval_cam_point(const double cam[6], const double pt[3], const float uv[2], const float K[9],
                        double res[2], double *Jc, double *Jp, int use_jets);
void orc_eval_point_only(const double pt[3], const float uv[2], const float E[16], const float K[9],
                         double res[2], double *Jp, int use_jets);
void orc_eval_pose_only(const double cam[6], const float Xh[4], const float uv[2], const float K[9],
                        double res[2], double *Jc, int use_jets);
void orc_huber(double s, double a, double rho[3]);

/* ---- whole-problem evaluation at the given parameters (robustified like Ceres) ----
 * residuals [2*No], jac_cam [12*No] (zeros for point-only rows), jac_pt [6*No]; any may be NULL.
 * gradient: grad_cam [6*Nc], grad_pt [3*Np] (J^T r on the corrected, unscaled Jacobian). */
int orc_evaluate(const orc_options *opt, int Nc, const double *cam, const uint8_t *cam_fixed,
                 const float *fixed_extr, const float *K, int Np, const double *pt, int No,
                 const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv,
                 double *cost, double *residuals, double *jac_cam, double *jac_pt,
                 double *grad_cam, double *grad_pt);

/* ---- full LM solve (ceres::Solve restatement). cam and pt are updated in place. ---- */
int orc_solve(const orc_options *opt, int Nc, double *cam, const uint8_t *cam_fixed,
              const float *fixed_extr, const float *K, int Np, double *pt, int No,
              const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv,
              orc_summary *summary, orc_iteration *iters, int iters_cap);

/* ---- one linearisation: reduced camera system of a landmark shard [pt_begin, pt_end) ----
 * Used by the world_size-2 gloo tests: sum over shards == whole.  S is (6Nc)^2 row-major, upper blocks
 * only; rhs 6Nc.  scale_cam/scale_pt may be NULL (=1).  include_diag adds U-independent D_c^2 once. */
int orc_schur_system(const orc_options *opt, double radius, int Nc, const double *cam,
                     const uint8_t *cam_fixed, const float *fixed_extr, const float *K, int Np,
                     const double *pt, int No, const int32_t *obs_cam, const int32_t *obs_pt,
                     const float *obs_uv, int pt_begin, int pt_end, double *S, double *rhs,
                     double *cost);

/* ---- frame-to-frame pose-only LM (MotionOnlyBAOptimizerAngles, Optimizer.cpp:417-498) ----
 * B independent problems; problem b owns observations [prob_ptr[b], prob_ptr[b+1]).  One ceres::Solve
 * each (the outer 4x loop + prune is host logic, see orc_prune_frame). */
int orc_solve_pose_only_batch(const orc_options *opt, int B, double *cam /*6B*/, const float *K /*9B*/,
                              const int32_t *prob_ptr, const float *Xh /*4*No*/, const float *uv /*2*No*/,
                              const uint8_t *obs_mask /*No or NULL; 0 = skipped (outlier)*/,
                              orc_summary *summaries /*B*/, orc_iteration *iters /*B*iters_cap*/,
                              int iters_cap);

/* ---- pruneCorrespondences (Optimizer.cpp:6-79), float32 semantics ----
 * pose = camera->world 4x4 f32 row-major; returns per-keypoint outlier flag updates in place. */
void orc_prune_frame(const float pose[16], const float K[9], int n, const float *kp_xy, const int32_t *kp_octave,
                     const float *pt_xyz /*3n*/, const float *pt_min_dist, const float *pt_max_dist,
                     const uint8_t *has_point, uint8_t *outlier, int consider_outlier);

/* flatten / write-back helpers (Optimizer.cpp:296-299, 262-267) */
void orc_pose_to_cam(const float pose[16] /*cam->world, row-major*/, double cam[6], float extr_out[16]);
void orc_cam_to_pose(const double cam[6], float pose_out[16]);

/* dense SPD solve used by the reduced system (Eigen LLT stand-in); returns 0 ok, 1 not PD */
int orc_cholesky_solve(int n, double *A_rowmajor_upper, double *b, int num_threads);

#ifdef __cplusplus
}
#endif
#endif
