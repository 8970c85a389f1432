/*
 * ba_b200.h -- C ABI of the B200-native bundle-adjustment solver.
 *
 * Drop-in boundary for the reference's BA optimisation entry point: the code in
 * ba_project/src/ba/Optimizer.{h,cpp} that flattens frames / keypoint observations / landmarks into a
 * ceres::Problem and calls ceres::Solve.  The reference has no FFI of its own (single C++ executable); the
 * C++ drop-in classes of the same names (bundleadjustment_b200/host/Optimizer.h) sit on top of this header,
 * and each entry point below cites the reference lines it replaces.
 *
 * Conventions (all host pointers unless the name ends in _dev):
 *   cam[6]         angle-axis (3) + translation (3), world->camera, f64      (Optimizer.cpp:296-299)
 *   K[9]           intrinsics f32 ROW-major 3x3, one per camera              (Frame::getIntrinsics)
 *   fixed_extr[16] world->camera 4x4 f32 ROW-major, read where cam_fixed!=0  (Optimizer.cpp:316)
 *   pt[3]          landmark xyz f64                                          (Optimizer.cpp:282-289)
 *   obs_uv[2]      keypoint pixel f32                                        (MapPoint.cpp:29-37)
 * Every function returns 0 on success, non-zero on error; ba_last_error() describes the last failure of the
 * calling thread.  A handle is not thread-safe; distinct handles may be used from distinct threads.
 * There is NO CPU fallback: compute entry points fail (BA_ERR_CUDA) when no sm_100 device is usable.
 */
#ifndef BA_B200_H
#define BA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BA_B200_VERSION 100

enum {
    BA_OK = 0,
    BA_ERR_INVALID = 1,   /* bad argument */
    BA_ERR_CUDA = 2,      /* CUDA / device failure (incl. no device) */
    BA_ERR_STATE = 3,     /* call order (e.g. solve before set_problem) */
    BA_ERR_NCCL = 4
};

enum { BA_DENSE_SCHUR = 0, BA_ITERATIVE_SCHUR = 1 };                  /* ceres::LinearSolverType subset */
enum { BA_CONVERGENCE = 0, BA_NO_CONVERGENCE = 1, BA_FAILURE = 2 };   /* ceres::TerminationType subset */

/* ceres::Solver::Options subset: what BAOptimizer::configureSolver sets (Optimizer.cpp:80-90) plus the Ceres
 * defaults it leaves untouched.  Field order/size identical to orc_options so tests can share initialisers. */
typedef struct ba_options {
    int32_t max_num_iterations;                 /* m_nMaxItPerBA (Optimizer.cpp:87; default 200, Optimizer.h:203) */
    int32_t num_threads;                        /* ignored (Optimizer.cpp:88 sets 4 CPU threads) */
    int32_t linear_solver;                      /* BA_DENSE_SCHUR (Optimizer.cpp:85) | BA_ITERATIVE_SCHUR */
    int32_t jacobi_scaling;                     /* 1 */
    int32_t max_num_consecutive_invalid_steps;  /* 5 */
    int32_t max_linear_solver_iterations;       /* 500 (PCG) */
    int32_t use_huber;                          /* 1: HuberLoss(sqrt(5.991)) on every block (Optimizer.cpp:312) */
    int32_t pad0;
    double function_tolerance;                  /* 1e-6 */
    double gradient_tolerance;                  /* 1e-10 */
    double parameter_tolerance;                 /* 1e-8 */
    double initial_trust_region_radius;         /* 1e4 */
    double max_trust_region_radius;             /* 1e16 */
    double min_trust_region_radius;             /* 1e-32 */
    double min_relative_decrease;               /* 1e-3 */
    double min_lm_diagonal;                     /* 1e-6 */
    double max_lm_diagonal;                     /* 1e32 */
    double huber_delta;                         /* sqrt(5.991) */
    double eta;                                 /* 0.1: PCG q-tolerance (Ceres ITERATIVE_SCHUR) */
    double cg_r_tolerance;                      /* -1: disabled; >0: also stop when |r| <= tol*|b| */
} ba_options;

/* ceres::IterationSummary subset == the "per-iteration cost trajectory" */
typedef struct ba_iteration {
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
} ba_iteration;

/* ceres::Solver::Summary subset (the reference ignores it, Optimizer.cpp:241-242; we report it anyway) */
typedef struct ba_summary {
    int32_t termination_type;
    int32_t num_iterations;          /* number of recorded iterations, iteration 0 included */
    int32_t num_successful_steps;
    int32_t num_unsuccessful_steps;
    double initial_cost;
    double final_cost;
    double seconds_total;            /* device time of the LM loop (CUDA events) */
    double seconds_linear_solver;    /* 0 unless profiling is enabled */
    double seconds_evaluate;
    char message[128];
} ba_summary;

typedef struct ba_handle ba_handle;

int ba_version(void);
const char *ba_last_error(void);
void ba_default_options(ba_options *opt);

/* ---- lifecycle --------------------------------------------------------------------------------------- */
/* One handle = one ceres::Problem + ceres::Solve replacement on one CUDA device (one stream per handle). */
int ba_create(const ba_options *opt, int device, ba_handle **out);
int ba_destroy(ba_handle *h);
int ba_set_options(ba_handle *h, const ba_options *opt);

/* ---- problem (replaces prepareConstraints' AddResidualBlock loops, Optimizer.cpp:279-333 / 612-698) ---- *
 * Observation i is an AngleReprojectionError<2,6,3> block on (cam[obs_cam[i]], pt[obs_pt[i]]) when
 * cam_fixed[obs_cam[i]]==0 and a PointOnlyReprojectionError<2,3> block with fixed_extr baked in otherwise.
 * Cameras / landmarks without observations stay untouched (they are not part of the ceres::Problem).
 * Arrays are copied to the device; the caller may free them after the call returns. */
int ba_set_problem(ba_handle *h, int32_t n_cam, const double *cam, const uint8_t *cam_fixed,
                   const float *fixed_extr, const float *K, int32_t n_pt, const double *pt, int32_t n_obs,
                   const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv);

/* Replace the parameter values only (same structure): next outer iteration after write-back/prune. */
int ba_set_parameters(ba_handle *h, const double *cam, const double *pt);
/* Restore the parameters given to ba_set_problem/ba_set_parameters, device-to-device (benchmark re-runs). */
int ba_reset_parameters(ba_handle *h);

/* ---- ceres::Solve replacement (Optimizer.cpp:242, 570).  Whole LM loop runs on the device. ------------ */
int ba_solve(ba_handle *h, ba_summary *summary);
int ba_get_cameras(ba_handle *h, double *cam /* 6*n_cam */);
int ba_get_points(ba_handle *h, double *pt /* 3*n_pt */);
int ba_get_iterations(ba_handle *h, ba_iteration *iters, int32_t cap, int32_t *n_written);

/* ---- the OUTER loop of Global / LocalBAOptimizerAngles::optimizeCamerasAndMapPoints (Optimizer.cpp:222-271, 550-603), ----
 * device-resident: n_outer x { ba_solve -> write-back (landmarks to float32, Optimizer.cpp:252-258; poses of the cameras with
 * cam_writeback != 0 -- the optimised frame list, frame ID 0 included -- to inverse(f32 extrinsics), :262-267) ->
 * pruneCorrespondences(frame, considerOutlier = false) over ALL keypoints of the problem's cameras (:270, :597-603) -> the
 * observations whose keypoint became an outlier leave the problem (:308) }.  The problem set by ba_set_problem stays on the
 * device: no re-upload and no index rebuild between outer iterations (masking, not re-sorting).
 *   pose[16 Nc]      camera->world f32 row-major of every camera of the problem, in/out
 *   frame_ptr[Nc+1]  camera c owns keypoints [frame_ptr[c], frame_ptr[c+1]) -- ALL its keypoints, as in ba_prune_frames
 *   kp_pt            problem landmark of the keypoint's map point, -1 when it has none or it is not a problem landmark
 *   kp_xyz           map point position (f32) of keypoints with has_point != 0 and kp_pt < 0 (constant during the call)
 *   kp_obs           index of the keypoint's observation in the ba_set_problem arrays, -1 when it is not an observation
 *   outlier          in/out;  pt_out[3 Np] (may be NULL): landmarks as float32;  summary: the LAST solve, seconds = all solves
 * Single GPU only.  Afterwards the handle holds the masked problem; call ba_set_problem for the next one. */
int ba_solve_outer(ba_handle *h, int32_t n_outer, const uint8_t *cam_writeback, float *pose, const int32_t *frame_ptr,
                   const float *kp_xy, const int32_t *kp_octave, const int32_t *kp_pt, const float *kp_xyz,
                   const float *pt_min_dist, const float *pt_max_dist, const uint8_t *has_point, const int32_t *kp_obs,
                   uint8_t *outlier, float *pt_out, ba_summary *summary);

/* ---- materialised evaluation at the current parameters (tests + the obs/sec metric) -------------------- *
 * Robustified exactly like Ceres' ResidualBlock::Evaluate + Corrector.  Outputs are in the CALLER's
 * observation order; any pointer may be NULL.  jac_cam rows of point-only blocks are zero. */
int ba_evaluate(ba_handle *h, double *cost, double *residuals /*2*n_obs*/, double *jac_cam /*12*n_obs*/,
                double *jac_pt /*6*n_obs*/, double *grad_cam /*6*n_cam*/, double *grad_pt /*3*n_pt*/);
/* Same evaluation, outputs left on the device (internal landmark-major order); returns kernel milliseconds. */
int ba_evaluate_device(ba_handle *h, int32_t repeats, float *ms_per_call);

/* Reduced camera system of the current linearisation at the given trust-region radius (tests, multi-GPU
 * checks): S is (6*n_cam)^2 row-major, LOWER blocks valid, rows of fixed/unobserved cameras zero. */
int ba_schur_system(ba_handle *h, double radius, double *S, double *rhs, double *cost);

/* ---- frame-to-frame "BA" pose estimation (MotionOnlyBAOptimizerAngles, Optimizer.cpp:417-498) --------- *
 * B independent PoseOnlyAngleReprojectionError<2,6> problems, one ceres::Solve each, one kernel launch for
 * all of them.  obs_mask (may be NULL): 0 = observation skipped (frame->isOutlier(j), Optimizer.cpp:472). */
int ba_solve_pose_only_batch(const ba_options *opt, int device, int32_t n_prob, double *cam /*6B in/out*/,
                             const float *K /*9B*/, const int32_t *prob_ptr /*B+1*/, const float *Xh /*4*No*/,
                             const float *uv /*2*No*/, const uint8_t *obs_mask, ba_summary *summaries /*B*/,
                             ba_iteration *iters /*B*iters_cap or NULL*/, int32_t iters_cap);

/* ---- pruneCorrespondences (Optimizer.cpp:6-79), float32 semantics, batched over frames ---------------- *
 * Frame f owns keypoints [frame_ptr[f], frame_ptr[f+1]).  pose = camera->world 4x4 f32 row-major. */
int ba_prune_frames(int device, int32_t n_frames, const float *pose /*16F*/, const float *K /*9F*/,
                    const int32_t *frame_ptr, const float *kp_xy, const int32_t *kp_octave,
                    const float *pt_xyz, const float *pt_min_dist, const float *pt_max_dist,
                    const uint8_t *has_point, uint8_t *outlier /*in/out*/, int32_t consider_outlier);

/* ---- ba_prune_frames without any allocation ----------------------------------------------------------- *
 * NO counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same two kernels and the
 * same results as ba_prune_frames, but
 *   - every array argument is a DEVICE pointer (readable / writable from `device`); n_kp = frame_ptr[n_frames] is
 *     passed by value because frame_ptr lives on the device; outlier is updated in place;
 *   - the kernels are ENQUEUED on `stream` (a cudaStream_t; NULL = the legacy default stream) and the call returns
 *     without synchronising: results are ready when the stream reaches that point;
 *   - the scratch ba_prune_frames keeps in its workspace (extr = pose.inverse(), 64 bytes per frame) lives in a
 *     file-scope `static __device__` array of BA_PRUNE_NOALLOC_MAX_FRAMES x 64 B = 256 KiB that every GPU which
 *     loads the library holds until the process exits.  n_frames beyond the maximum is rejected (BA_ERR_INVALID).
 * The call issues no cudaMalloc / cudaFree / stream or event creation.  The array's device address is looked up once
 * per device (cudaGetSymbolAddress, on the first call for that device: make that call outside a capture).
 * ALL calls on one device share the one array and it keeps the previous call's contents: each call overwrites the
 * n_frames entries it reads before reading them (nothing has to start at zero), and two calls whose kernels can
 * overlap on the GPU (different streams without an ordering between them, or several host threads) RACE on it --
 * order them on one stream or with events. */
#define BA_PRUNE_NOALLOC_MAX_FRAMES 4096
int ba_prune_frames_noalloc(int device, void *stream, int32_t n_frames, int32_t n_kp, const float *d_pose /*16F*/,
                            const float *d_K /*9F*/, const int32_t *d_frame_ptr /*F+1*/, const float *d_kp_xy,
                            const int32_t *d_kp_octave, const float *d_pt_xyz, const float *d_pt_min_dist,
                            const float *d_pt_max_dist, const uint8_t *d_has_point, uint8_t *d_outlier /*in/out*/,
                            int32_t consider_outlier);

/* ---- SfMHelper::estimatePoseUsingBA's whole loop (SfMHelper.cpp:106-117: 4 outer iterations x <= 20 LM iterations,
 * Optimizer.cpp:417-457), batched over frames and DEVICE-RESIDENT: per outer iteration the pose -> parameter conversion
 * (Optimizer.cpp:462-465), the observation list (keypoints with a map point that are not outliers, :472), one ceres::Solve
 * replacement per frame, the float32 write-back (:448-453) and pruneCorrespondences(frame, considerOutlier = true) (:455)
 * run as five kernel launches; nothing crosses PCIe between outer iterations.  Keypoint arrays as in ba_prune_frames
 * (pt_xyz = the keypoint's map point position, float32, constant).  pose and outlier are updated in place.
 * summaries (may be NULL): the LAST outer iteration's ba_summary per frame.  device_ms (may be NULL): device time of the
 * whole loop.  lm_iterations (may be NULL): LM iterations of all frames and all outer iterations (iteration 0 excluded). */
int ba_solve_pose_only_outer(const ba_options *opt, int device, int32_t n_frames, float *pose /*16F in/out*/,
                             const float *K /*9F*/, const int32_t *frame_ptr, const float *kp_xy,
                             const int32_t *kp_octave, const float *pt_xyz, const float *pt_min_dist,
                             const float *pt_max_dist, const uint8_t *has_point, uint8_t *outlier /*in/out*/,
                             int32_t n_outer, ba_summary *summaries, float *device_ms, int64_t *lm_iterations);

/* The handle-less entry points (ba_solve_pose_only_batch, ba_prune_frames, ba_solve_pose_only_outer) keep one stream, two
 * events and grow-only device buffers per calling thread and device, so that the per-frame calls of the tracker never
 * allocate.  ba_release_workspace frees the calling thread's workspaces; ba_workspace_launch_count = kernels the thread's
 * last such call launched on `device` (benchmark bookkeeping). */
int ba_release_workspace(void);
int ba_workspace_launch_count(int device);

/* ---- multi-GPU: landmark-sharded, one process per GPU, NCCL over NVLink ------------------------------- *
 * Rank r owns a contiguous landmark range and all of its observations; cameras are replicated.  Every rank
 * passes the FULL camera arrays and its LOCAL landmarks/observations (obs_pt local) to ba_set_problem after
 * ba_comm_init.  The reduced camera system (or the PCG vectors) and the LM scalars are all-reduced. */
int ba_comm_unique_id(uint8_t id[128]);                                   /* ncclGetUniqueId (rank 0) */
int ba_comm_init(ba_handle *h, int32_t n_ranks, int32_t rank, const uint8_t id[128]);

/* Number of kernels + collectives the last ba_solve enqueued (benchmark bookkeeping: "gpu_launches"). */
int ba_get_launch_count(ba_handle *h);
/* on != 0: ba_solve records CUDA events around each linearisation and each linear solve and reports the sums in
 * ba_summary.seconds_evaluate / seconds_linear_solver (event recording does not synchronise the stream). */
int ba_set_profiling(ba_handle *h, int32_t on);

/* Profiling detail of the last ba_solve (profiling on): out = {seconds inside the dominant linear-solver kernel -- the dense
 * Cholesky kernel, or the PCG matvec launches that did work -- measured with CUDA events on the solver's stream around each
 * launch; number of launches; 0; 0}.  Benchmark bookkeeping ("roofline"). */
int ba_get_kernel_profile(ba_handle *h, double out[4]);

/* Dense SPD solve A x = b with the reduced-system Cholesky (DenseSchurComplementSolver's Eigen LLT stand-in) on
 * its own: tests and the FP64 tensor-pipe micro-benchmark.  A: n x n row-major, lower triangle read.
 * ms (may be NULL): best device time of factor + solve over `repeats` runs. */
int ba_dense_spd_solve(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats);
/* The same with the Cholesky tile edge forced (64 | 128; 0 = automatic as above): tests of both kernel instantiations. */
int ba_dense_spd_solve_tile(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats,
                            int32_t tile_edge);

/* Reduced camera system of the current problem (dense Schur): out = {padded order (one 128-row tile per 21 cameras), 128 x 128 tiles of the Cholesky factor that
 * can be nonzero, tiles in the lower triangle, flops of one factorisation + solve over those tiles}.  Cameras that share
 * no landmark leave exact-zero tiles in S (DenseSchurComplementSolver fills them with zeros, ceres
 * schur_complement_solver.cc); the factorisation skips tiles that stay zero, which changes no result. */
int ba_get_reduced_system_stats(ba_handle *h, double out[4]);

/* Explicit switches of the reduced-system factorisation (tests and measurements; both default to 0): natural_order != 0
 * keeps the Cholesky tiles in camera order instead of nested dissection; dense_tiles != 0 factorises S as a fully dense
 * matrix, as Ceres' DENSE_SCHUR (Optimizer.cpp:85) does.  Results change by summation order only.  The library reads NO
 * environment variables.  Call before ba_set_problem. */
int ba_set_reduced_system_mode(ba_handle *h, int32_t natural_order, int32_t dense_tiles);
/* Tile edge of the reduced-system Cholesky: 0 = automatic (64-row tiles when co-visibility leaves at least three quarters of
 * the lower tiles of S empty, or when the system has at most 128 tile rows = 1 280 cameras -- the factorisation is then a
 * dependency chain of tile columns and shorter links win -- else 128-row tiles, which run large dense matrices closest to
 * the FP64 tensor peak), 64 or 128 = forced. */
int ba_set_reduced_system_tile(ba_handle *h, int32_t tile_edge);

/* Accumulation of the Schur-Jacobi PCG's implicit product q = S p over the landmark blocks.  1 = ORDERED: persistent CTAs fold
 * the per-observation contributions per camera in a fixed order into per-CTA private copies of q (no floating-point atomics;
 * the whole PCG path is then bit-reproducible run to run, like the dense path).  2 = ATOMIC: f64 atomicAdd into 32 copies of q
 * (results vary in the last bits between runs).  0 = automatic (default): ordered while its private copies stay L2-resident
 * (up to 48 MB), atomic beyond -- on the 30 M-observation problem the ordered kernel's copies (185 MB) cost 1.6 x the time. */
int ba_set_pcg_mode(ba_handle *h, int32_t mode);

/* ---- host-only helpers (no GPU needed) ---------------------------------------------------------------- */
/* Balanced contiguous landmark partition: cut[r]..cut[r+1] for r in [0,n_ranks); weight = sum over the
 * landmark's track length t of t (iterative) or t*(t+1)/2 (dense Schur).  obs_pt must be sorted ascending. */
int ba_partition_landmarks(int32_t n_pt, int32_t n_obs, const int32_t *obs_pt, int32_t n_ranks,
                           int32_t quadratic, int32_t *cut /* n_ranks+1 */);
/* Elimination order of the 128 x 128 Cholesky tiles for a tile adjacency (G x G bytes, lower triangle read): nested
 * dissection, so that independent parts of the reduced camera system factorise concurrently.  order[pos] = tile.
 * stats (may be NULL) = {dependency depth natural order, depth in the returned order, filled tiles natural, ordered}. */
int ba_tile_elimination_order(int32_t G, const uint8_t *adj, int32_t *order, double stats[4]);
/* Matrix4f::inverse() stand-in (float32 adjugate), ceres::AngleAxisToRotationMatrix / RotationMatrixToAngleAxis
 * (column-major 3x3) */
void ba_inverse4x4_f32(const float m[16], float out[16]);
void ba_angle_axis_to_rotation(const double w[3], double R_colmajor[9]);
void ba_rotation_to_angle_axis(const double R_colmajor[9], double w[3]);
/* Optimizer.cpp:296-299: extr = pose.inverse() (f32), rotation -> angle-axis (f64), translation -> f64 */
void ba_pose_to_cam(const float pose[16], double cam[6], float extr_out[16]);
/* Optimizer.cpp:262-267: angle-axis -> R (f64) -> f32 extrinsics -> inverse -> pose */
void ba_cam_to_pose(const double cam[6], float pose_out[16]);

#ifdef __cplusplus
}
#endif
#endif /* BA_B200_H */
