// ba_solver.cuh -- device-resident state of one BA problem / LM solve (internal; the public surface is
// include/ba_b200.h).
#pragma once

#include "ba_device.cuh"
#include "../../include/ba_b200.h"

namespace ba {

// message codes (mapped to Ceres' wording on the host)
enum : int {
    MSG_NONE = 0, MSG_MAX_ITER, MSG_GRADIENT_TOL, MSG_MIN_RADIUS, MSG_PARAM_TOL, MSG_FUNCTION_TOL,
    MSG_INVALID_STEPS, MSG_EVAL_FAILED
};

// slots of the all-reduced (sum) scalar vector
enum : int {
    SUM_LIN_COST = 0,   // cost at x (from the linearisation pass)
    SUM_LIN_GSQ,        // sum over landmark params of (x - (x - g))^2
    SUM_LIN_XSQ,        // sum over active landmark params of x^2
    SUM_STEP_DXSQ,      // sum over landmark params of (x - cand)^2
    SUM_STEP_YB,        // y_p . b_p
    SUM_STEP_YHY,       // y_p^T V_p y_p + 2 a_p . y_p
    SUM_CAND_COST,      // cost at the candidate
    SUM_COUNT = 8
};
constexpr int SUM_LIN_BEGIN = 0, SUM_LIN_N = 3, SUM_STEP_BEGIN = 3, SUM_STEP_N = 4;
// The max over ranks of the landmark-side gradient norm travels in the summed message: ctx.maxs has one slot per rank, rank r
// writes slot r (ctx.maxs_loc points at it), lm_begin_iteration takes the max of the slots.
// camera-side scalars (replicated, never all-reduced)
enum : int { CS_GSQ = 0, CS_GMAX, CS_XSQ, CS_DXSQ, CS_YB, CS_YHY, CS_COUNT = 8 };

// Trust-region state (trust_region_minimizer.cc + levenberg_marquardt_strategy.cc), lives in device memory;
// a copy of `done` is mirrored to mapped host memory so the host can stop enqueueing iterations.
struct LMState {
    int done;
    int need_linearize;
    int cur;                 // which of the two parameter buffers holds x
    int num_invalid;
    int termination;
    int msg;
    int chol_fail;           // set by the dense factorisation / PCG on breakdown
    int lin_iters;
    int num_successful;
    int num_unsuccessful;
    int n_recorded;
    int cg_done;             // PCG loop control
    int cg_iter;
    int cg_bad_block;        // a 6x6 preconditioner block was not positive definite
    double radius;
    double decrease_factor;
    double x_cost;
    double x_norm;
    double initial_cost;
    // PCG scalars
    double cg_rho, cg_last_rho, cg_alpha, cg_pq, cg_Q0, cg_norm_b;
    ba_iteration it;         // summary being filled for the current iteration
};

// Camera-pair-major view of the observations (dense Schur): entry k = (observation a, observation b) of ONE landmark with
// camera(a) >= camera(b), entries sorted by (camera a, camera b); a segment = all entries of one S block, cut into
// units of <= unit_len <= PAIR_UNIT entries (one warp each).
constexpr int PAIR_UNIT = 256;
constexpr int PAIR_TARGET_UNITS = 8192;   // units wanted at least (about 3.5 warps per SM sub-partition) before a unit grows beyond 32 entries
struct PairIndex {
    long long P;             // entries
    int n_seg, n_units;
    int unit_len;            // entries per unit of THIS problem: 32 .. PAIR_UNIT (ba_prep.cu: prep_pairs_build)
    size_t seg_cap;          // capacity of the seg_* arrays
    int2 *pair_obs;          // [P]
    long long *seg_beg;      // [n_seg]
    int2 *seg_cams;          // [n_seg] (ci, cj), ci >= cj
    int *unit_ptr;           // [n_seg + 1]
    int *unit_seg;           // [n_units] entry count << 1 | (diagonal block)
    long long *unit_beg;     // [n_units]
    double *unit_part;       // [n_units][42]: 36 block entries + 6 rhs entries
};

struct Ctx {
    // sizes
    int Nc, Np, No, n;       // n = 6*Nc (reduced system order; inactive cameras are identity rows)
    int n_chunks;
    int rank, n_ranks;
    // options
    ba_options opt;
    // parameters: two buffers each, state->cur selects x, the other is the candidate
    double *cam[2];          // [Nc][6]
    double *pt[2];           // [Np][3]
    double *cam_init, *pt_init;
    // constants
    const int *cam_kind;     // [Nc] CAM_FREE / CAM_FIXED / CAM_INACTIVE
    const float *cam_K;      // [Nc][9]
    const float *cam_extr;   // [Nc][16]
    // landmark-major observations (sorted by landmark, then camera)
    const float2 *pm_uv;
    const int *pm_cam;
    const unsigned *pm_camk; // [No] camera | kind << 30 (CAM_* of that camera): what the per-observation kernels read
    const unsigned *pm_trk;  // [No] position inside the landmark's track | track length << 16
    const unsigned *blk_ord; // [No] per landmark block: its observations in camera order (PCG matvec; built on first PCG solve)
    const int *pm_pt;
    const int *pt_ptr;       // [Np+1]
    const int *blk_first_pt; // [n_blocks+1] landmark blocks (whole landmarks, <= tile_cap observations)
    const int4 *blk_info;    // [n_blocks] {first landmark, end landmark, first observation, observation count}
    int n_blocks, tile_cap;
    int masked;              // != 0: some observations carry OBS_MASKED (ba_solve_outer after a prune)
    // camera-major copy
    const float2 *cm_uv;
    const int *cm_pt;
    const int *cm_obs;       // landmark-major position of each camera-major entry
    const int *chunk_cam, *chunk_beg, *chunk_end;   // [n_chunks] work units of <= CHUNK observations
    const int *cam_chunk_ptr;                       // [Nc+1] chunk range per camera
    PairIndex pairs;         // dense Schur only (built on first use)
    // per-camera tables at x and at the candidate
    CamEntry *tab_x;         // linearisation point
    CamProj *tab_c;          // candidate point (cost only)
    // Jacobi scaling
    double *scale_cam, *scale_pt;
    // normal-equation blocks (scaled space)
    // [U | b_c | sums | max slots] are ONE contiguous buffer (the packed linearisation message, ba_capi.cu)
    double *U;               // [Nc][36]  (summed over ranks)
    double *bc;              // [Nc][6]   (summed over ranks)
    double *U_loc, *bc_loc;  // this rank's contribution (alias U / bc on a single GPU)
    double *V;               // [Np][6] sym
    double *bp;              // [Np][3]
    double *W;               // [No][18] = Jc^T Jp
    double *cam_part;        // [n_chunks][27]
    double *Vinv;            // [Np][6]
    double *wv;              // [Np][3] = Vinv * bp
    // reduced system
    // One Cholesky tile (S_nb = 128 or 64 rows) holds S_nb / 6 consecutive cameras (+ 2 or 4 identity rows); the tiles are laid
    // out in elimination order (nested dissection of the camera-group graph), so camera c's block starts at row cam_row[c].
    double *S;               // TILE-PACKED (ba_cholesky.cu layout contract): tile (i, j) of the padded lower triangle at S + S_toff[i * S_T + j],
                             // S_nb x S_nb row-major; tile row S_T = the right-hand side (row 0 of each tile)
    const long long *S_toff; // [(S_T + 1) * S_T] offsets in doubles, -1 = structurally zero tile
    int S_T;                 // tiles per side = ldS / S_nb
    int S_nb, S_shift;       // tile edge (128 | 64) and its log2
    int ldS;                 // order of the padded system = S_nb * number of camera groups
    const int *cam_row;      // [Nc]
    const int *row_src;      // [ldS] 6 c + a of the camera parameter in that row, -1 for padding rows
    double *rhs;             // [n] (becomes y_c after the solve)
    double *yc;              // [n]
    // partial sums
    double *part;            // [n_part][4]
    int n_part;
    double *sums;            // [SUM_COUNT]  (summed over ranks)
    double *maxs;            // [n_ranks] one slot per rank
    double *sums_loc, *maxs_loc;   // this rank's contribution (alias sums / maxs on a single GPU); maxs_loc = this rank's slot
    double *camsums;         // [CS_COUNT]
    double *red_part;        // [RED_SLOTS][RED_MAX_GRID][4] per-CTA partials of the multi-CTA reductions (grid_sum3_max)
    unsigned *red_count;     // [RED_SLOTS] arrival tickets (zero between launches)
    // PCG work vectors (n each) and block-Jacobi preconditioner
    double *cg_r, *cg_p, *cg_z, *cg_q, *cg_x, *Minv, *tp;
    double *cg_part;         // [1024] per-CTA partials of the cooperative CG kernels
    double *cg_qp;           // [mv_grid][n] private accumulation copies of the matvec, one per persistent CTA (all zero between products)
    int mv_grid;             // private copies of q (= persistent CTAs of the ordered matvec; 32 for the atomic one)
    int pcg_ordered;         // != 0: bit-reproducible matvec (camera-keyed ordered fold), else f64 atomics into the copies
    // control
    LMState *state;
    ba_iteration *iters;     // [max_num_iterations + 2]
    volatile int *host_flags; // mapped pinned [16]: [0] = done, [1] = cg_done (latest state), [HF_LM_RING ..] / [HF_CG_RING ..] = the
                              // verdict AS OF THE END of LM body / CG iteration number (slot), written by the publish kernels
};

// The host stops enqueueing LM bodies / CG iterations when it sees the verdict of the body / iteration it last waited for.  It
// must read the verdict AS OF THAT body, not the latest one: every rank of a sharded solve has to stop at the same body, or the
// ranks' NCCL calls no longer pair up (a faster rank could already see the flag of the following body and leave one all-reduce
// early).  Hence a ring of snapshots in the mapped flag words instead of the two "latest state" words.
constexpr int HF_LM_RING = 2, HF_LM_SLOTS = 4, HF_CG_RING = 8, HF_CG_SLOTS = 8;

// slots of red_part / red_count
enum : int { RED_LIN = 0, RED_CAM_NORMS, RED_CAM_UPDATE, RED_STEP, RED_CG_CHECK, RED_CG_START, RED_SLOTS = 8 };
__host__ __device__ inline double *red_slot(const Ctx &ctx, int slot) { return ctx.red_part + (size_t)slot * RED_MAX_GRID * 4; }

__host__ __device__ inline int cams_per_tile(int nb) { return nb / 6; }   // 21 cameras (126 rows) per 128-row tile, 10 (60 rows) per 64-row tile
// element (r, q), r >= q's tile row, of the tile-packed reduced system (the tile must be part of the plan)
__device__ __forceinline__ double *s_elem(const Ctx &ctx, int r, int q) {
    const int m = ctx.S_nb - 1;
    return ctx.S + ctx.S_toff[(size_t)(r >> ctx.S_shift) * ctx.S_T + (q >> ctx.S_shift)] + ((size_t)(r & m) << ctx.S_shift) + (q & m);
}
constexpr int CHUNK = 128;   // observations per camera-side reduction unit
constexpr int TILE_WIN = 192;   // observation window defining a landmark block of the tile kernels (ba_prep.cu)
constexpr int TILE_THREADS = 256;
constexpr int RED_THREADS = 256;   // CTA size of the multi-CTA reductions (grid_red4)

} // namespace ba
