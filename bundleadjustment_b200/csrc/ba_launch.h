// ba_launch.h -- host-side launch wrappers of the kernels (internal).
#pragma once

#include "ba_solver.cuh"

#include <vector>

namespace ba {

// ba_kernels.cu
void launch_init_state(const Ctx &ctx, cudaStream_t s);
void launch_fill(double *p, size_t n, double v, cudaStream_t s);
void launch_cam_prep(const Ctx &ctx, int which, int force, cudaStream_t s);
void launch_linearize_local(const Ctx &ctx, int force, cudaStream_t s);
void launch_cam_norms(const Ctx &ctx, int force, int with_begin, cudaStream_t s);   // with_begin: + controller part 1 (the LM loop)
void launch_compute_scale(const Ctx &ctx, cudaStream_t s);
void launch_decide(const Ctx &ctx, int body, cudaStream_t s);           // accept / reject + the verdict as of the end of LM body `body` -> host ring
void launch_point_invert(const Ctx &ctx, double radius_override, cudaStream_t s);
void launch_point_invert_schur_build(const Ctx &ctx, cudaStream_t s);   // LM loop, dense Schur: point_invert + schur_init in one launch, then units + segments
void launch_schur_build(const Ctx &ctx, double radius_override, cudaStream_t s);
void launch_backsub(const Ctx &ctx, cudaStream_t s);
void launch_eval_candidate(const Ctx &ctx, double *cost_part, cudaStream_t s);
int tiles_init(const Ctx &ctx);                          // per-problem dynamic shared memory opt-in of the tile kernels
int pcg_tiles_init(const Ctx &ctx);
int launch_materialize(const Ctx &ctx, double *res, double *jc, double *jp, double *cost_part, cudaStream_t s);   // returns the number of cost partials


// ba_prep.cu -- device-side index building (ba_set_problem)
struct PrepScratch {              // grow-only temporary storage
    void *p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes);    // returns a cudaError
    void release();
};
struct PrepIn {
    int Nc, Np, No;
    const int *obs_cam, *obs_pt;          // [No] caller order (device)
    const float2 *obs_uv;                 // [No]
    const unsigned char *cam_fixed;       // [Nc]
};
struct PrepOut {
    // landmark-major
    float2 *pm_uv; int *pm_cam, *pm_pt, *pt_ptr, *perm; unsigned *pm_camk, *pm_trk;
    // cameras
    int *cnt_cam, *cnt_cam_global, *cam_kind, *cam_ptr, *cam_chunk_ptr;
    // camera-major
    float2 *cm_uv; int *cm_pt, *cm_obs, *chunk_cam, *chunk_beg, *chunk_end;
    // landmark blocks of the tile kernels
    int *blk_first_pt; int4 *blk_info; int n_blocks = 0, max_block_obs = 0;
    int *flags;                            // device [4]
    int *h_flags;                          // pinned host [8]: bad index (1-based), unsorted, Ncm, n_chunks, max block obs
    int Ncm = 0, n_chunks = 0;
};
int prep_observations(const PrepIn &in, PrepOut &out, PrepScratch &scr, cudaStream_t s);   // <0: -(1 + index of a bad observation)
int prep_blocks(const PrepIn &in, PrepOut &out, cudaStream_t s);                            // needs pt_ptr
int prep_camera_major(const PrepIn &in, PrepOut &out, PrepScratch &scr, cudaStream_t s);   // needs cnt_cam_global
int prep_chunks(const PrepIn &in, PrepOut &out, cudaStream_t s);
// per block: its observations in camera order (blk_ord word = camera << 12 | segment start << 11 | local index); cudaErrorInvalidValue
// when a camera index needs more than 20 bits or a block holds 2048+ observations
int prep_block_order(const PrepOut &out, int Nc, unsigned *blk_ord, cudaStream_t s);
int prep_pairs_count(const PrepIn &in, const PrepOut &out, PrepScratch &scr, long long *h_total, cudaStream_t s);
int prep_pairs_build(const PrepIn &in, const PrepOut &out, PairIndex &pi, PrepScratch &scr, PrepScratch &scr2, int *h_two, cudaStream_t s);
int prep_pairs_units(PairIndex &pi, cudaStream_t s);
// map[G x G] bytes (lower triangle): camera groups (cams_per_tile consecutive cameras = one Cholesky tile) that share a landmark
int prep_group_map(const PairIndex &pi, int G, int cams_per_tile, unsigned char *map, cudaStream_t s);

// ba_cholesky.cu -- dense SPD factorisation + solve of the reduced camera system (C6), one cooperative kernel
// Tile edge nb: 128 (dense / nearly dense systems) or 64 (tile-sparse systems: shorter dependency chain), chosen per plan.
int chol_init();                                        // per-device kernel attribute setup; returns cudaError
int chol_ld(int n, int nb);                             // n_pad = ceil(n / nb) * nb
size_t chol_work_doubles(int n, int nb);
// A: the plan's tile-packed storage (plan->total_doubles doubles, see the layout contract in ba_cholesky.cu): nb x nb
// tiles, each contiguous and row-major, at plan->tile_off[i * T + j]; tile row T is the right-hand side (row 0 of each tile).
// On exit y (n) holds the solution and the tiles hold L.
// state may be nullptr (standalone use); otherwise the kernel early-exits on state->done and reports breakdown in
// state->chol_fail.  fail_flag (device int, may be nullptr) is set on breakdown as well.  Returns a cudaError.
// plan: tile-level sparsity plan of this matrix (nullptr = dense).
struct CholPlan {
    int n = 0, nb = 128, T = 0, ntasks = 0;       // nb = tile edge (64 | 128)
    const int4 *task = nullptr; const int *klist = nullptr; const unsigned char *fill = nullptr;
    const unsigned char *partner = nullptr;   // [ntasks] chain fusion: off-diagonal tasks that hand their pre-TRSM tile to the row's diagonal task
    const long long *tile_off = nullptr;      // device [(T + 1) * T]: offset of tile (i, j) in doubles, -1 = not in the plan
    std::vector<long long> h_tile_off;        // host copy
    long long front_doubles = 0;     // the tiles S itself can touch are stored first: [0, front_doubles)
    long long extra_doubles = 0;     // caller-owned region right behind them (the reduced rhs in camera order)
    long long total_doubles = 0;     // size of the packed matrix buffer
    long long filled_tiles = 0;      // tiles of L (T x T lower triangle) that can be nonzero
    double flops = 0.0;              // factorisation + both substitutions over the filled tiles
    int depth = 0;                   // longest dependency chain, in tasks
    void *blob = nullptr; size_t cap = 0;   // device storage (grow-only)
};
// nz: host T x T bytes, (i, j) != 0 where tile (i, j) or (j, i) of S can be nonzero (nullptr = dense); synchronises s.
// Elimination order of the tiles from their adjacency (G x G bytes, lower triangle read): order[pos] = tile.
void chol_tile_order(int G, const unsigned char *adj, std::vector<int> &order);
int chol_plan_create(int n, int nb, const unsigned char *nz, CholPlan *plan, cudaStream_t s, size_t extra_doubles = 0);
void chol_pack_host(const CholPlan *plan, const double *A, const double *b, std::vector<double> &out);   // host matrix -> packed tiles
void chol_plan_destroy(CholPlan *plan);
// zero the packed matrix (all tiles of the plan + the extra region) before S is assembled
int chol_clear(double *A, const CholPlan *plan, cudaStream_t s);
void chol_plan_stats(int n, int nb, const unsigned char *nz, int *depth, long long *filled, double *flops);   // host only
// perm (may be nullptr): the LM loop's permutations folded into the kernel -- the right-hand side is READ from rhs_cam[row_src[i]]
// (0 where row_src[i] < 0) instead of the rhs tiles, and the solution is ALSO written to y_cam[row_src[i]]
struct CholPerm { const int *row_src; const double *rhs_cam; double *y_cam; };
int chol_factor_solve(double *A, int n, double *work, double *y, LMState *state, int *fail_flag, const CholPlan *plan, cudaStream_t s,
                      const CholPerm *perm = nullptr);

// ba_pcg.cu -- Schur-Jacobi preconditioned CG on the implicit reduced system (C6')
int pcg_matvec_grid(const Ctx &ctx);                          // private copies of q (cg_qp) for ctx.pcg_ordered
size_t pcg_ordered_copy_bytes(const Ctx &ctx);                // bytes of the ordered kernel's private copies for this problem
void launch_pcg_setup(const Ctx &ctx, cudaStream_t s);       // local block-diag(S) + rhs (all-reduce Minv, rhs after)
void launch_pcg_start(const Ctx &ctx, cudaStream_t s);       // invert blocks, x = 0, r = b
void launch_pcg_iter_a(const Ctx &ctx, cudaStream_t s, cudaEvent_t ev_before_matvec = nullptr, cudaEvent_t ev_after_matvec = nullptr);      // z = M^-1 r, p, local landmark part of q = S p
void launch_pcg_iter_b(const Ctx &ctx, cudaStream_t s);      // q += (U+D^2) p, alpha, x, r   (after all-reduce of q)
void launch_pcg_reset_a(const Ctx &ctx, cudaStream_t s);     // local landmark part of S x into cg_z
void launch_pcg_reset_b(const Ctx &ctx, cudaStream_t s);     // r = b - S x                   (after all-reduce of cg_z)
void launch_pcg_check(const Ctx &ctx, int iteration, cudaStream_t s);   // stop rules + the verdict as of the end of this CG iteration -> host ring
void launch_pcg_finish(const Ctx &ctx, cudaStream_t s);      // copies x -> yc, lin_iters

// ba_pose_batch.cu
void launch_pose_batch(const ba_options &opt, int B, double *cam, const float *K, const int *prob_ptr,
                       const float4 *Xh, const float2 *uv, const uint8_t *mask, ba_summary *summaries,
                       ba_iteration *iters, int iters_cap, cudaStream_t s);
void launch_pose_to_cam(int F, const float *pose, double *cam, cudaStream_t s);                       // Optimizer.cpp:462-465 per frame
void launch_cam_to_pose(int F, const double *cam, float *pose, float *extr, cudaStream_t s);          // write-back + extr = pose^-1
void launch_pose_inverse(int F, const float *pose, float *extr, cudaStream_t s);                      // extr = pose^-1 (f32)
void launch_pose_mask(int n_kp, const uint8_t *has_point, const uint8_t *outlier, uint8_t *mask, const float *xyz, float4 *Xh, cudaStream_t s);
// device-resident outer loop of Global / Local BA (ba_solve_outer)
void launch_outer_writeback(int Nc, const unsigned char *wb, const double *cam_cur, float *pose, double *cam_init, float *cam_extr,
                            size_t n3, const double *pt_cur, float *pt32, double *pt_init, cudaStream_t s);
void launch_outer_kp_xyz(int n_kp, const int *kp_pt, const float *pt32, float *kp_xyz, cudaStream_t s);
void launch_outer_mask(int n_kp, const int *kp_obs, const unsigned char *outlier, const int *perm, int Nc, const unsigned char *fixed, int No,
                       const int *pm_cam, unsigned char *obs_masked, int *cam_active, int *cam_kind, unsigned *pm_camk, cudaStream_t s);
void launch_prune(int n_frames, const float *pose, const float *extr, const float *K, const int *frame_ptr, const float2 *kp_xy,
                  const int *kp_octave, const float *pt_xyz, const float *pt_min, const float *pt_max,
                  const uint8_t *has_point, uint8_t *outlier, int consider_outlier, int n_kp, cudaStream_t s);

// ba_capi.cu: sets ba_last_error() of the calling thread, returns code (for entry points defined outside ba_capi.cu)
int report_error(int code, const char *text);

} // namespace ba
