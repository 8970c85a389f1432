// ba_capi.cu -- C ABI (include/ba_b200.h) + host orchestration of the device-resident LM loop.
//
// The host only ENQUEUES work: every accept/reject/termination decision is taken by the controller kernels
// (ba_kernels.cu) from device-resident state; kernels of iterations enqueued after termination early-exit.  The
// host looks at a mapped `done` word two iterations behind the GPU merely to stop enqueueing.
#include "ba_launch.h"

#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>

using namespace ba;

namespace {

thread_local std::string g_err;

int fail(int code, const char *fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

} // namespace
// error text of entry points that live in other translation units (ba_pose_batch.cu: ba_prune_frames_noalloc)
int ba::report_error(int code, const char *text) { return fail(code, "%s", text); }
namespace {

#define CU(expr)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (expr);                                                                    \
        if (e__ != cudaSuccess) return fail(BA_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e__));  \
    } while (0)
#define NC(expr)                                                                                     \
    do {                                                                                             \
        ncclResult_t r__ = (expr);                                                                   \
        if (r__ != ncclSuccess) return fail(BA_ERR_NCCL, "%s: %s", #expr, ncclGetErrorString(r__));  \
    } while (0)

const char *message_text(int msg) {
    switch (msg) {
        case MSG_MAX_ITER: return "Maximum number of iterations reached.";
        case MSG_GRADIENT_TOL: return "Gradient tolerance reached.";
        case MSG_MIN_RADIUS: return "Minimum trust region radius reached.";
        case MSG_PARAM_TOL: return "Parameter tolerance reached.";
        case MSG_FUNCTION_TOL: return "Function tolerance reached.";
        case MSG_INVALID_STEPS: return "Number of consecutive invalid steps more than Solver::Options::max_num_consecutive_invalid_steps.";
        case MSG_EVAL_FAILED: return "Residual and Jacobian evaluation failed.";
        default: return "";
    }
}

struct DevBuf {                      // grow-only: re-solving a problem of the same shape never touches the allocator
    void *p = nullptr;
    size_t bytes = 0, cap = 0;
    cudaError_t alloc(size_t b) {
        if (b <= cap && p) { bytes = b; return cudaSuccess; }
        free();
        if (b == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, b);
        if (e == cudaSuccess) { bytes = b; cap = b; } else p = nullptr;
        return e;
    }
    void free() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0; cap = 0;
    }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

} // namespace

struct ba_handle {
    int device = 0;
    ba_options opt;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    cudaEvent_t ev_ring[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_lm[2] = {nullptr, nullptr};       // LM-body completion markers of the enqueue loop
    bool has_problem = false;
    Ctx ctx;
    // caller order -> landmark-major position (fetched from the device on first use)
    std::vector<int> perm;           // pm position of caller observation i
    bool perm_valid = false;
    DevBuf raw_cam, raw_pt, raw_uv, raw_fixed, perm_dev, cnt_cam, cnt_cam_global, cam_ptr, prep_flags;
    DevBuf pair_obs, seg_beg, seg_cams, unit_ptr, unit_seg, unit_beg, unit_part;
    PrepScratch scr, scr2;
    PrepIn prep_in;
    PrepOut prep_out;
    bool pairs_built = false;
    bool natural_tile_order = false, dense_tiles = false;   // ba_set_reduced_system_mode (explicit API, never the environment)
    int tile_edge = 0, tile_nb = 128;                       // requested (0 = automatic) / chosen Cholesky tile edge
    int pcg_mode = 0;                                       // ba_set_pcg_mode: 0 automatic, 1 ordered (bit-reproducible), 2 atomic
    CholPlan chol_plan;              // tile-sparsity plan of the reduced camera matrix (dense Schur)
    DevBuf tile_map, cam_row, row_src, yperm;
    int *h_prep = nullptr;           // pinned [8]
    // device buffers
    DevBuf cam0, cam1, pt0, pt1, cam_init, pt_init, cam_kind, cam_K, cam_extr;
    DevBuf cm_obs, blk_first_pt, blk_info, pm_camk, pm_trk, blk_ord;
    bool ord_built = false;          // blk_ord (camera order inside each landmark block): built on the first PCG solve
    DevBuf pm_uv, pm_cam, pm_pt, pt_ptr, cm_uv, cm_pt, chunk_cam, chunk_beg, chunk_end, cam_chunk_ptr;
    DevBuf tab_x, tab_c, scale_cam, scale_pt, U, bc, V, bp, W, cam_part, Vinv, wv, S, rhs, yc;
    DevBuf part, cost_part, red_part, red_count, lin_red, lin_loc, camsums, state, iters, invL;   // lin_*: [U | b_c | sums | max slots], one all-reduce
    DevBuf cg_r, cg_p, cg_z, cg_q, cg_x, Minv, tp, cg_part, cg_qp;
    DevBuf ev_res, ev_jc, ev_jp;     // materialised evaluation outputs
    DevBuf o_pose, o_extr, o_pt32, o_wb, o_ptr, o_xy, o_oct, o_xyz, o_mn, o_mx, o_has, o_out, o_kp_pt, o_kp_obs, o_masked, o_active;   // ba_solve_outer
    int *host_flags = nullptr;       // mapped pinned
    int *host_flags_dev = nullptr;
    // NCCL
    ncclComm_t comm = nullptr;
    int n_ranks = 1, rank = 0;
    std::vector<int> cam_obs_count;  // global per-camera observation counts
    size_t lin_count = 0;            // doubles in the packed linearisation message
    int launches = 0;                // kernels/collectives enqueued by the last solve (bench bookkeeping)
    bool profile = false;            // record CUDA events around every linear solve / linearisation
    std::vector<cudaEvent_t> prof_ev;   // [4 per body]: lin begin, lin end (= solve begin), solve end, body end
    size_t prof_used = 0;
    std::vector<cudaEvent_t> dom_ev;    // event pairs around the dominant linear-solver kernel (the Cholesky dataflow kernel)
    size_t dom_used = 0;
    double dom_seconds = 0.0;
    int dom_launches = 0;
    cudaEvent_t dom_event() {
        if (dom_used == dom_ev.size()) { cudaEvent_t e; cudaEventCreate(&e); dom_ev.push_back(e); }
        return dom_ev[dom_used++];
    }
    cudaEvent_t prof_event() {
        if (prof_used == prof_ev.size()) { cudaEvent_t e; cudaEventCreate(&e); prof_ev.push_back(e); }
        return prof_ev[prof_used++];
    }

    void free_problem() {
        DevBuf *all[] = {&cam0, &cam1, &pt0, &pt1, &cam_init, &pt_init, &cam_kind, &cam_K, &cam_extr, &pm_uv, &pm_cam, &pm_pt,
                         &pt_ptr, &cm_uv, &cm_pt, &chunk_cam, &chunk_beg, &chunk_end, &cam_chunk_ptr, &tab_x, &tab_c, &scale_cam,
                         &scale_pt, &U, &bc, &V, &bp, &W, &cam_part, &Vinv, &wv, &S, &rhs, &yc, &part, &cost_part, &red_part, &red_count, &lin_red, &lin_loc,
                         &camsums, &state, &iters, &invL, &cg_r, &cg_p, &cg_z, &cg_q, &cg_x, &Minv, &tp, &cg_part, &cg_qp, &ev_res, &ev_jc, &ev_jp,
                         &raw_cam, &raw_pt, &raw_uv, &raw_fixed, &perm_dev, &cnt_cam, &cnt_cam_global, &cam_ptr, &prep_flags,
                         &pair_obs, &seg_beg, &seg_cams, &unit_ptr, &unit_seg, &unit_beg, &unit_part, &cm_obs, &blk_first_pt, &blk_info, &pm_camk, &pm_trk, &blk_ord, &tile_map, &cam_row, &row_src, &yperm,
                         &o_pose, &o_extr, &o_pt32, &o_wb, &o_ptr, &o_xy, &o_oct, &o_xyz, &o_mn, &o_mx, &o_has, &o_out, &o_kp_pt, &o_kp_obs, &o_masked, &o_active};
        for (DevBuf *b : all) b->free();
        chol_plan_destroy(&chol_plan);
        scr.release(); scr2.release();
        has_problem = false;
    }
};

extern "C" {

int ba_version(void) { return BA_B200_VERSION; }
const char *ba_last_error(void) { return g_err.c_str(); }

void ba_default_options(ba_options *o) {
    std::memset(o, 0, sizeof(*o));
    o->max_num_iterations = 200;          // BAOptimizer(): m_nMaxItPerBA(200), Optimizer.h:203
    o->num_threads = 4;                   // Optimizer.cpp:88 (unused on the GPU)
    o->linear_solver = BA_DENSE_SCHUR;    // Optimizer.cpp:85
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

static int check_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(BA_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= n) return fail(BA_ERR_INVALID, "device %d out of range (0..%d)", device, n - 1);
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(BA_ERR_CUDA, "device %d is sm_%d%d; this build targets sm_100a only", device, prop.major, prop.minor);
    CU(cudaSetDevice(device));
    return BA_OK;
}

int ba_create(const ba_options *opt, int device, ba_handle **out) {
    if (!out) return fail(BA_ERR_INVALID, "out is NULL");
    *out = nullptr;
    int rc = check_device(device);
    if (rc) return rc;
    ba_handle *h = new ba_handle();
    h->device = device;
    if (opt) h->opt = *opt; else ba_default_options(&h->opt);
    std::memset(&h->ctx, 0, sizeof(h->ctx));
    CU(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&h->ev_start));
    CU(cudaEventCreate(&h->ev_stop));
    for (auto &e : h->ev_ring) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    for (auto &e : h->ev_lm) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    CU(cudaHostAlloc((void **)&h->h_prep, 64, cudaHostAllocDefault));
    CU(cudaHostAlloc((void **)&h->host_flags, 64, cudaHostAllocMapped));
    std::memset(h->host_flags, 0, 64);
    CU(cudaHostGetDevicePointer((void **)&h->host_flags_dev, h->host_flags, 0));
    int e = chol_init();
    if (e) { g_err = std::string("chol_init: ") + cudaGetErrorString((cudaError_t)e); return BA_ERR_CUDA; }
    *out = h;
    return BA_OK;
}

int ba_destroy(ba_handle *h) {
    if (!h) return BA_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    h->free_problem();
    if (h->comm) ncclCommDestroy(h->comm);
    if (h->host_flags) cudaFreeHost(h->host_flags);
    if (h->h_prep) cudaFreeHost(h->h_prep);
    for (auto &e : h->ev_ring) if (e) cudaEventDestroy(e);
    for (auto &e : h->ev_lm) if (e) cudaEventDestroy(e);
    for (auto &e : h->prof_ev) cudaEventDestroy(e);
    for (auto &e : h->dom_ev) cudaEventDestroy(e);
    if (h->ev_start) cudaEventDestroy(h->ev_start);
    if (h->ev_stop) cudaEventDestroy(h->ev_stop);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return BA_OK;
}

int ba_set_options(ba_handle *h, const ba_options *opt) {
    if (!h || !opt) return fail(BA_ERR_INVALID, "NULL argument");
    h->opt = *opt;
    h->ctx.opt = *opt;
    if (h->has_problem) {
        // iteration log capacity follows max_num_iterations
        cudaSetDevice(h->device);
        CU(h->iters.alloc(sizeof(ba_iteration) * (size_t)(opt->max_num_iterations + 2)));
        h->ctx.iters = h->iters.as<ba_iteration>();
    }
    return BA_OK;
}

} // extern "C"

// =============================================================================================================
// problem upload
// =============================================================================================================
namespace {

template <typename T>
int upload(DevBuf &b, const T *src, size_t count, cudaStream_t s) {
    CU(b.alloc(sizeof(T) * std::max<size_t>(count, 1)));
    if (count) CU(cudaMemcpyAsync(b.p, src, sizeof(T) * count, cudaMemcpyHostToDevice, s));
    return BA_OK;
}

// automatic Cholesky tile edge: dense systems of up to this many 64-row tile rows (8 192 rows, 1 280 cameras) keep 64-row tiles
// (measured on dense patterns, profiles/r02_chol_tile_crossover_dense.txt: n = 2 048: 0.67 ms against 1.37 ms on 128-row tiles,
// 4 096: 1.71 / 2.83, 8 192: 8.05 / 8.51; beyond that the 128-row kernel's 68 % of the FP64 GEMM rate takes over)
constexpr int CHOL_SMALL_TILE_ROWS = 128;
// PCG only, once per problem: camera order inside every landmark block + the matvec's private copies of q
int ensure_pcg(ba_handle *h) {
    Ctx &ctx = h->ctx;
    if (!h->ord_built) {
        cudaStream_t s = h->stream;
        CU(h->blk_ord.alloc(4 * std::max<size_t>((size_t)ctx.No, 1)));
        const int rc = prep_block_order(h->prep_out, ctx.Nc, h->blk_ord.as<unsigned>(), s);
        if (rc == (int)cudaErrorInvalidValue) return fail(BA_ERR_INVALID, "the iterative solver supports up to 2^20 cameras and landmark blocks below 2048 observations");
        if (rc) return fail(BA_ERR_CUDA, "prep_block_order: %s", cudaGetErrorString((cudaError_t)rc));
        ctx.blk_ord = h->blk_ord.as<unsigned>();
        // Matvec accumulation (ba_set_pcg_mode): the ordered kernel is bit-reproducible (no floating-point atomics) but keeps one
        // private copy of q per persistent CTA; once those no longer fit in L2 every read-modify-write of a far camera goes to
        // DRAM (K5: 296 x 624 KB = 185 MB, 2.56 ms per launch against 1.61 ms -- profiles/r02_K5_matvec_ordered_ncu.txt), so
        // the automatic mode uses it up to 48 MB of copies and the atomic kernel beyond.
        ctx.pcg_ordered = h->pcg_mode == 1 || (h->pcg_mode == 0 && pcg_ordered_copy_bytes(ctx) <= ((size_t)48 << 20));
        if (sizeof(double) * 46 * (size_t)(ctx.tile_cap + 1) > 227 * 1024) ctx.pcg_ordered = 0;      // blocks too long for the two-stage ring
        ctx.mv_grid = pcg_matvec_grid(ctx);
        CU(h->cg_qp.alloc(8 * (size_t)ctx.mv_grid * std::max<size_t>((size_t)ctx.n, 1)));
        CU(cudaMemsetAsync(h->cg_qp.p, 0, h->cg_qp.bytes, s));
        ctx.cg_qp = h->cg_qp.as<double>();
        h->ord_built = true;
    }
    return BA_OK;
}

int ensure_dense(ba_handle *h) {
    if (!h->pairs_built) {
        // camera-pair-major index for the deterministic Schur reduction (device-side, once per problem)
        cudaStream_t s = h->stream;
        long long *h_total = reinterpret_cast<long long *>(h->h_prep + 4);
        int rc = prep_pairs_count(h->prep_in, h->prep_out, h->scr, h_total, s);
        if (rc) return fail(BA_ERR_CUDA, "prep_pairs_count: %s", cudaGetErrorString((cudaError_t)rc));
        PairIndex &pi = h->ctx.pairs;
        std::memset(&pi, 0, sizeof(pi));
        pi.P = *h_total;
        const size_t P = (size_t)std::max<long long>(pi.P, 1);
        const size_t Ncz = (size_t)h->ctx.Nc;
        pi.seg_cap = std::min<size_t>(P, Ncz * (Ncz + 1) / 2);
        CU(h->pair_obs.alloc(8 * P)); CU(h->seg_beg.alloc(8 * pi.seg_cap)); CU(h->seg_cams.alloc(8 * pi.seg_cap));
        CU(h->unit_ptr.alloc(4 * (pi.seg_cap + 1)));
        pi.pair_obs = h->pair_obs.as<int2>(); pi.seg_beg = h->seg_beg.as<long long>(); pi.seg_cams = h->seg_cams.as<int2>();
        pi.unit_ptr = h->unit_ptr.as<int>();
        rc = prep_pairs_build(h->prep_in, h->prep_out, pi, h->scr, h->scr2, h->h_prep + 6, s);
        if (rc) return fail(BA_ERR_CUDA, "prep_pairs_build: %s", cudaGetErrorString((cudaError_t)rc));
        const size_t nu = (size_t)std::max(pi.n_units, 1);
        CU(h->unit_seg.alloc(4 * nu)); CU(h->unit_beg.alloc(8 * nu)); CU(h->unit_part.alloc(8 * 42 * nu));
        pi.unit_seg = h->unit_seg.as<int>(); pi.unit_beg = h->unit_beg.as<long long>(); pi.unit_part = h->unit_part.as<double>();
        if ((rc = prep_pairs_units(pi, s))) return fail(BA_ERR_CUDA, "prep_pairs_units: %s", cudaGetErrorString((cudaError_t)rc));
        // (the sort scratch stays allocated: cudaFree + cudaMalloc of it cost up to 100 ms per ba_set_problem)
        // which camera groups share landmarks (union over ranks) -> elimination order of the tiles (nested dissection),
        // row layout of S, tile-level symbolic factorisation
        // Tile edge: 64-row tiles (10 cameras each) when the co-visibility leaves most of S empty -- the factorisation is then a
        // dependency chain along the nested-dissection separators, and the chain is much shorter in 64-row links -- else 128-row
        // tiles (21 cameras), which run a dense matrix at 67 % of the FP64 GEMM rate.  ba_set_reduced_system_tile overrides.
        int nb = h->tile_edge ? h->tile_edge : 64;
        if (h->dense_tiles && !h->tile_edge) nb = 128;
        int G = 0, n = 0;
        std::vector<unsigned char> adj;
        for (;;) {
            const int cpt = cams_per_tile(nb);
            G = (h->ctx.Nc + cpt - 1) / cpt;
            n = G * nb;
            CU(h->tile_map.alloc((size_t)G * G));
            if ((rc = prep_group_map(pi, G, cpt, h->tile_map.as<unsigned char>(), s)))
                return fail(BA_ERR_CUDA, "prep_group_map: %s", cudaGetErrorString((cudaError_t)rc));
            if (h->comm) NC(ncclAllReduce(h->tile_map.p, h->tile_map.p, (size_t)G * G, ncclUint8, ncclMax, h->comm, s));
            adj.resize((size_t)G * G);
            CU(cudaMemcpyAsync(adj.data(), h->tile_map.p, adj.size(), cudaMemcpyDeviceToHost, s));
            CU(cudaStreamSynchronize(s));
            if (nb == 128 || h->tile_edge) break;
            long long nnz = 0;
            for (int gi = 0; gi < G; ++gi)
                for (int gj = 0; gj <= gi; ++gj) nnz += (gi == gj || adj[(size_t)gi * G + gj]) ? 1 : 0;
            if (2 * nnz <= (long long)G * (G + 1) / 2 / 2) break;           // at most a quarter of the lower tiles are touched: sparse
            // SMALL and MEDIUM systems stay on 64-row tiles even when dense: up to several thousand rows the factorisation is the chain
            // potrf -> trsm -> update of the tile columns, not flops, and a 64-row link costs 17 us against 83 us for a 128-row
            // one (keyframe-BA sizes: 30 cameras 67 us against 144 us, profiles/r02_chol_chain_small_systems.txt)
            if (G <= CHOL_SMALL_TILE_ROWS) break;
            nb = 128;                                                        // (the same on every rank: adj is the union over ranks)
        }
        h->tile_nb = nb;
        std::vector<int> order;
        if (h->natural_tile_order) { order.resize(G); std::iota(order.begin(), order.end(), 0); }   // ba_set_reduced_system_mode
        else chol_tile_order(G, adj.data(), order);
        if ((int)order.size() != G) return fail(BA_ERR_STATE, "tile ordering lost nodes (internal error)");
        std::vector<int> pos(G);
        for (int p = 0; p < G; ++p) pos[order[p]] = p;
        std::vector<int> cam_row(h->ctx.Nc), row_src((size_t)n, -1);
        for (int c = 0; c < h->ctx.Nc; ++c) {
            cam_row[c] = nb * pos[c / cams_per_tile(nb)] + 6 * (c % cams_per_tile(nb));
            for (int a = 0; a < 6; ++a) row_src[(size_t)cam_row[c] + a] = 6 * c + a;
        }
        if ((rc = upload(h->cam_row, cam_row.data(), cam_row.size(), s))) return rc;
        if ((rc = upload(h->row_src, row_src.data(), row_src.size(), s))) return rc;
        std::vector<unsigned char> nz((size_t)G * G, 0);
        for (int gi = 0; gi < G; ++gi)
            for (int gj = 0; gj <= gi; ++gj)
                if (gi == gj || adj[(size_t)gi * G + gj]) nz[(size_t)std::max(pos[gi], pos[gj]) * G + std::min(pos[gi], pos[gj])] = 1;
        CU(cudaStreamSynchronize(s));                                    // the uploads read the host vectors above
        // ba_set_reduced_system_mode(dense_tiles): treat every tile as filled (what Ceres' DENSE_SCHUR does).  The plan also fixes
        // the STORAGE of S: [tiles a camera pair can touch | reduced rhs (6 Nc, camera order) | fill-in tiles | rhs tiles].
        if ((rc = chol_plan_create(n, nb, h->dense_tiles ? nullptr : nz.data(), &h->chol_plan, s, (size_t)h->ctx.n)))
            return fail(BA_ERR_CUDA, "chol_plan_create: %s", cudaGetErrorString((cudaError_t)rc));
        CU(h->S.alloc(sizeof(double) * (size_t)h->chol_plan.total_doubles));
        CU(h->invL.alloc(sizeof(double) * chol_work_doubles(n, nb)));
        CU(cudaMemsetAsync(h->invL.p, 0, sizeof(double) * chol_work_doubles(n, nb), s));    // the kernel only ever writes the lower triangles of L_jj^-1
        CU(h->yperm.alloc(sizeof(double) * (size_t)n));
        h->pairs_built = true;
    }
    h->ctx.S_nb = h->tile_nb;
    h->ctx.S_shift = h->tile_nb == 128 ? 7 : 6;
    h->ctx.S_T = h->chol_plan.T;
    h->ctx.ldS = h->chol_plan.T * h->tile_nb;
    h->ctx.S = h->S.as<double>();
    h->ctx.S_toff = h->chol_plan.tile_off;
    h->ctx.cam_row = h->cam_row.as<int>();
    h->ctx.row_src = h->row_src.as<int>();
    // the reduced right-hand side lives inside the S buffer, right behind the tiles the Schur kernels write: a sharded solve
    // all-reduces [S tiles | rhs] as ONE message out of the buffer the kernels wrote (no pack / unpack pass)
    h->ctx.rhs = h->ctx.S + h->chol_plan.front_doubles;
    return BA_OK;
}

} // namespace

extern "C" int ba_set_problem(ba_handle *h, int32_t n_cam, const double *cam, const uint8_t *cam_fixed,
                              const float *fixed_extr, const float *K, int32_t n_pt, const double *pt, int32_t n_obs,
                              const int32_t *obs_cam, const int32_t *obs_pt, const float *obs_uv) {
    if (!h) return fail(BA_ERR_INVALID, "handle is NULL");
    if (n_cam <= 0 || n_pt < 0 || n_obs < 0) return fail(BA_ERR_INVALID, "bad sizes n_cam=%d n_pt=%d n_obs=%d", n_cam, n_pt, n_obs);
    if (!cam || !cam_fixed || !K || (n_pt && !pt) || (n_obs && (!obs_cam || !obs_pt || !obs_uv)))
        return fail(BA_ERR_INVALID, "NULL array");
    CU(cudaSetDevice(h->device));
    const int Nc = n_cam, Np = n_pt, No = n_obs;
    h->has_problem = false;
    h->perm_valid = false;
    h->pairs_built = false;
    h->ord_built = false;
    cudaStream_t s = h->stream;
    Ctx &ctx = h->ctx;
    std::memset(&ctx, 0, sizeof(ctx));
    ctx.Nc = Nc; ctx.Np = Np; ctx.No = No; ctx.n = 6 * Nc;
    ctx.rank = h->rank; ctx.n_ranks = h->n_ranks;
    ctx.opt = h->opt;
    int rc;
#define UP(buf, ptr, cnt) if ((rc = upload(h->buf, ptr, (size_t)(cnt), s))) return rc
#define AL(buf, bytes) CU(h->buf.alloc(std::max<size_t>((size_t)(bytes), 8)))
    // ---- raw observations to the device; all index building happens there (ba_prep.cu) ----
    UP(raw_cam, obs_cam, No); UP(raw_pt, obs_pt, No); UP(raw_uv, obs_uv, 2 * (size_t)No); UP(raw_fixed, cam_fixed, Nc);
    const size_t max_chunks = (size_t)No / CHUNK + (size_t)Nc + 1;
    AL(pm_uv, 8 * (size_t)No); AL(pm_cam, 4 * (size_t)No); AL(pm_pt, 4 * (size_t)No); AL(pt_ptr, 4 * (size_t)(Np + 1));
    AL(perm_dev, 4 * (size_t)No); AL(cnt_cam, 4 * (size_t)Nc); AL(cnt_cam_global, 4 * (size_t)Nc); AL(cam_kind, 4 * (size_t)Nc);
    AL(cam_ptr, 4 * (size_t)(Nc + 1)); AL(cam_chunk_ptr, 4 * (size_t)(Nc + 1)); AL(cm_uv, 8 * (size_t)No); AL(cm_pt, 4 * (size_t)No);
    AL(cm_obs, 4 * (size_t)No); AL(blk_first_pt, 4 * ((size_t)No / TILE_WIN + 2)); AL(blk_info, 16 * ((size_t)No / TILE_WIN + 2)); AL(pm_camk, 4 * (size_t)No); AL(pm_trk, 4 * (size_t)No);
    AL(chunk_cam, 4 * max_chunks); AL(chunk_beg, 4 * max_chunks); AL(chunk_end, 4 * max_chunks); AL(prep_flags, 32);
    PrepIn &pin = h->prep_in;
    PrepOut &po = h->prep_out;
    pin.Nc = Nc; pin.Np = Np; pin.No = No;
    pin.obs_cam = h->raw_cam.as<int>(); pin.obs_pt = h->raw_pt.as<int>(); pin.obs_uv = h->raw_uv.as<float2>();
    pin.cam_fixed = h->raw_fixed.as<unsigned char>();
    po.pm_uv = h->pm_uv.as<float2>(); po.pm_cam = h->pm_cam.as<int>(); po.pm_pt = h->pm_pt.as<int>(); po.pt_ptr = h->pt_ptr.as<int>();
    po.perm = h->perm_dev.as<int>(); po.cnt_cam = h->cnt_cam.as<int>(); po.cnt_cam_global = h->cnt_cam_global.as<int>();
    po.cam_kind = h->cam_kind.as<int>(); po.cam_ptr = h->cam_ptr.as<int>(); po.cam_chunk_ptr = h->cam_chunk_ptr.as<int>();
    po.cm_uv = h->cm_uv.as<float2>(); po.cm_pt = h->cm_pt.as<int>(); po.cm_obs = h->cm_obs.as<int>();
    po.blk_first_pt = h->blk_first_pt.as<int>(); po.blk_info = h->blk_info.as<int4>(); po.pm_camk = h->pm_camk.as<unsigned>(); po.pm_trk = h->pm_trk.as<unsigned>();
    po.chunk_cam = h->chunk_cam.as<int>(); po.chunk_beg = h->chunk_beg.as<int>(); po.chunk_end = h->chunk_end.as<int>();
    po.flags = h->prep_flags.as<int>(); po.h_flags = h->h_prep;
    rc = prep_observations(pin, po, h->scr, s);
    if (h->comm) {
        // a rank that rejects its shard must not leave its peers blocked in the first data collective: agree on the verdict first
        int bad = rc != 0 ? 1 : 0, *flag = h->prep_flags.as<int>() + 7;
        CU(cudaMemcpyAsync(flag, &bad, 4, cudaMemcpyHostToDevice, s));
        NC(ncclAllReduce(flag, flag, 1, ncclInt, ncclMax, h->comm, s));
        CU(cudaMemcpyAsync(&bad, flag, 4, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        if (bad && rc == 0) return fail(BA_ERR_INVALID, "another rank rejected its shard of the problem (see that rank's error)");
    }
    if (rc < 0) {
        const int i = -rc - 1;
        return fail(BA_ERR_INVALID, "observation %d references camera %d / landmark %d out of range", i, obs_cam[i], obs_pt[i]);
    }
    if (rc) return fail(BA_ERR_CUDA, "prep_observations: %s", cudaGetErrorString((cudaError_t)rc));
    if ((rc = prep_blocks(pin, po, s))) return fail(BA_ERR_CUDA, "prep_blocks: %s", cudaGetErrorString((cudaError_t)rc));
    // ---- per-camera observation counts (summed over ranks) decide which cameras are part of the problem ----
    h->cam_obs_count.resize(Nc);
    if (h->comm) {
        NC(ncclAllReduce(po.cnt_cam, po.cnt_cam_global, Nc, ncclInt, ncclSum, h->comm, s));
    } else {
        CU(cudaMemcpyAsync(po.cnt_cam_global, po.cnt_cam, 4 * (size_t)Nc, cudaMemcpyDeviceToDevice, s));
    }
    if (!fixed_extr) {      // legal only if no fixed camera is observed
        CU(cudaMemcpyAsync(h->cam_obs_count.data(), po.cnt_cam_global, 4 * (size_t)Nc, cudaMemcpyDeviceToHost, s));
        CU(cudaStreamSynchronize(s));
        for (int c = 0; c < Nc; ++c)
            if (cam_fixed[c] && h->cam_obs_count[c] > 0) return fail(BA_ERR_INVALID, "fixed camera %d but fixed_extr is NULL", c);
    }
    if ((rc = prep_camera_major(pin, po, h->scr, s))) return fail(BA_ERR_CUDA, "prep_camera_major: %s", cudaGetErrorString((cudaError_t)rc));
    if ((rc = prep_chunks(pin, po, s))) return fail(BA_ERR_CUDA, "prep_chunks: %s", cudaGetErrorString((cudaError_t)rc));
    ctx.n_chunks = po.n_chunks;
    ctx.n_blocks = po.n_blocks;
    ctx.tile_cap = std::max(po.max_block_obs, 1);
    ctx.n_part = ctx.n_blocks;

    // ---- parameters and constants ----
    // one H2D copy each; the two working buffers are filled device-to-device
    UP(cam_init, cam, 6 * (size_t)Nc); UP(pt_init, pt, 3 * (size_t)Np);
    AL(cam0, 8 * 6 * (size_t)Nc); AL(cam1, 8 * 6 * (size_t)Nc); AL(pt0, 8 * 3 * (size_t)Np); AL(pt1, 8 * 3 * (size_t)Np);
    CU(cudaMemcpyAsync(h->cam0.p, h->cam_init.p, 8 * 6 * (size_t)Nc, cudaMemcpyDeviceToDevice, s));
    CU(cudaMemcpyAsync(h->cam1.p, h->cam_init.p, 8 * 6 * (size_t)Nc, cudaMemcpyDeviceToDevice, s));
    if (Np) {
        CU(cudaMemcpyAsync(h->pt0.p, h->pt_init.p, 8 * 3 * (size_t)Np, cudaMemcpyDeviceToDevice, s));
        CU(cudaMemcpyAsync(h->pt1.p, h->pt_init.p, 8 * 3 * (size_t)Np, cudaMemcpyDeviceToDevice, s));
    }
    UP(cam_K, K, 9 * (size_t)Nc);
    if (fixed_extr) { UP(cam_extr, fixed_extr, 16 * (size_t)Nc); }
    else { AL(cam_extr, 64 * (size_t)Nc); CU(cudaMemsetAsync(h->cam_extr.p, 0, 64 * (size_t)Nc, s)); }
#undef UP
    AL(tab_x, sizeof(CamEntry) * (size_t)Nc); AL(tab_c, sizeof(CamProj) * (size_t)Nc);
    AL(scale_cam, 8 * 6 * (size_t)Nc); AL(scale_pt, 8 * 3 * (size_t)Np);
    AL(V, 8 * 6 * (size_t)Np); AL(bp, 8 * 3 * (size_t)Np);
    AL(W, 8 * 18 * (size_t)No + 64);     /* + slack: rows are fetched as aligned 32-byte windows (load_row18) */ AL(cam_part, 8 * 27 * (size_t)std::max(ctx.n_chunks, 1));
    AL(Vinv, 8 * 6 * (size_t)Np + 64); AL(wv, 8 * 3 * (size_t)Np);
    AL(rhs, 8 * (size_t)ctx.n); AL(yc, 8 * (size_t)ctx.n + 64);      /* + slack: 6-double rows are gathered as 32-byte windows */
    AL(part, 8 * 4 * (size_t)ctx.n_part); AL(cost_part, 8 * (size_t)((No + 31) / 32 + 8));
    AL(camsums, 8 * CS_COUNT);
    // One contiguous message per linearisation: [U (36 Nc) | b_c (6 Nc) | sums (SUM_COUNT) | max slots (one per rank)].  A sharded
    // solve all-reduces (sum) the rank's lin_loc into lin_red in ONE call; the max over ranks of the gradient norm rides along
    // exactly: rank r writes its value into slot r (zeros elsewhere), the sum hands every rank all of them.
    const size_t lin_count = 42 * (size_t)Nc + SUM_COUNT + (size_t)std::max(h->n_ranks, 1);
    AL(lin_red, 8 * lin_count);
    CU(cudaMemsetAsync(h->lin_red.p, 0, 8 * lin_count, s));
    if (h->comm) { AL(lin_loc, 8 * lin_count); CU(cudaMemsetAsync(h->lin_loc.p, 0, 8 * lin_count, s)); }
    h->lin_count = lin_count;
    AL(red_part, 8 * 4 * (size_t)RED_MAX_GRID * RED_SLOTS); AL(red_count, 4 * RED_SLOTS);
    CU(cudaMemsetAsync(h->red_count.p, 0, 4 * RED_SLOTS, s));
    AL(state, sizeof(LMState)); AL(iters, sizeof(ba_iteration) * (size_t)(h->opt.max_num_iterations + 2));
    AL(cg_r, 8 * (size_t)ctx.n); AL(cg_p, 8 * (size_t)ctx.n + 64); AL(cg_z, 8 * (size_t)ctx.n); AL(cg_q, 8 * (size_t)ctx.n);
    AL(cg_x, 8 * (size_t)ctx.n + 64); AL(Minv, 8 * 36 * (size_t)Nc); AL(tp, 8 * 3 * (size_t)Np); AL(cg_part, 8 * 1024);
    ctx.cam[0] = h->cam0.as<double>(); ctx.cam[1] = h->cam1.as<double>();
    ctx.pt[0] = h->pt0.as<double>(); ctx.pt[1] = h->pt1.as<double>();
    ctx.cam_init = h->cam_init.as<double>(); ctx.pt_init = h->pt_init.as<double>();
    ctx.cam_kind = h->cam_kind.as<int>(); ctx.cam_K = h->cam_K.as<float>(); ctx.cam_extr = h->cam_extr.as<float>();
    ctx.pm_uv = h->pm_uv.as<float2>(); ctx.pm_cam = h->pm_cam.as<int>(); ctx.pm_pt = h->pm_pt.as<int>();
    ctx.pt_ptr = h->pt_ptr.as<int>(); ctx.cm_uv = h->cm_uv.as<float2>(); ctx.cm_pt = h->cm_pt.as<int>();
    ctx.chunk_cam = h->chunk_cam.as<int>(); ctx.chunk_beg = h->chunk_beg.as<int>(); ctx.chunk_end = h->chunk_end.as<int>();
    ctx.cam_chunk_ptr = h->cam_chunk_ptr.as<int>();
    ctx.cm_obs = h->cm_obs.as<int>(); ctx.blk_first_pt = h->blk_first_pt.as<int>(); ctx.blk_info = h->blk_info.as<int4>(); ctx.pm_camk = h->pm_camk.as<unsigned>(); ctx.pm_trk = h->pm_trk.as<unsigned>();
    ctx.tab_x = h->tab_x.as<CamEntry>(); ctx.tab_c = h->tab_c.as<CamProj>();
    ctx.scale_cam = h->scale_cam.as<double>(); ctx.scale_pt = h->scale_pt.as<double>();
    {
        double *red = h->lin_red.as<double>(), *loc = h->comm ? h->lin_loc.as<double>() : red;
        ctx.U = red; ctx.bc = red + 36 * (size_t)Nc; ctx.sums = red + 42 * (size_t)Nc; ctx.maxs = ctx.sums + SUM_COUNT;
        ctx.U_loc = loc; ctx.bc_loc = loc + 36 * (size_t)Nc; ctx.sums_loc = loc + 42 * (size_t)Nc;
        ctx.maxs_loc = ctx.sums_loc + SUM_COUNT + (h->comm ? h->rank : 0);
    }
    ctx.V = h->V.as<double>(); ctx.bp = h->bp.as<double>();
    ctx.W = h->W.as<double>(); ctx.cam_part = h->cam_part.as<double>(); ctx.Vinv = h->Vinv.as<double>(); ctx.wv = h->wv.as<double>();
    ctx.rhs = h->rhs.as<double>(); ctx.yc = h->yc.as<double>();
    ctx.part = h->part.as<double>(); ctx.camsums = h->camsums.as<double>();
    ctx.red_part = h->red_part.as<double>(); ctx.red_count = h->red_count.as<unsigned>();
    ctx.cg_r = h->cg_r.as<double>(); ctx.cg_p = h->cg_p.as<double>(); ctx.cg_z = h->cg_z.as<double>(); ctx.cg_q = h->cg_q.as<double>();
    ctx.cg_x = h->cg_x.as<double>(); ctx.Minv = h->Minv.as<double>(); ctx.tp = h->tp.as<double>(); ctx.cg_part = h->cg_part.as<double>();
    ctx.state = h->state.as<LMState>(); ctx.iters = h->iters.as<ba_iteration>();
    ctx.host_flags = h->host_flags_dev;
#undef AL
    if ((rc = tiles_init(ctx)))
        return fail(rc == (int)cudaErrorInvalidConfiguration ? BA_ERR_INVALID : BA_ERR_CUDA,
                    "a landmark block of %d observations does not fit in shared memory (longest supported track: about 1000 observations)", ctx.tile_cap);
    CU(cudaMemsetAsync(h->camsums.p, 0, 8 * CS_COUNT, s));
    launch_init_state(ctx, s);
    launch_fill(ctx.scale_cam, 6 * (size_t)Nc, 1.0, s);
    launch_fill(ctx.scale_pt, 3 * (size_t)Np, 1.0, s);
    launch_cam_prep(ctx, 0, 1, s);      // K / fixed extrinsics of both tables are written once here
    launch_cam_prep(ctx, 1, 1, s);
    CU(cudaStreamSynchronize(s));       // the caller may free / reuse its arrays after the call returns
    CU(cudaGetLastError());
    h->has_problem = true;
    return BA_OK;
}

extern "C" int ba_set_parameters(ba_handle *h, const double *cam, const double *pt) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "no problem set");
    CU(cudaSetDevice(h->device));
    const Ctx &ctx = h->ctx;
    if (cam) {
        CU(cudaMemcpyAsync(ctx.cam_init, cam, 8 * 6 * (size_t)ctx.Nc, cudaMemcpyHostToDevice, h->stream));
    }
    if (pt && ctx.Np) CU(cudaMemcpyAsync(ctx.pt_init, pt, 8 * 3 * (size_t)ctx.Np, cudaMemcpyHostToDevice, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    return ba_reset_parameters(h);
}

extern "C" int ba_reset_parameters(ba_handle *h) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "no problem set");
    CU(cudaSetDevice(h->device));
    const Ctx &ctx = h->ctx;
    cudaStream_t s = h->stream;
    for (int b = 0; b < 2; ++b) {
        CU(cudaMemcpyAsync(ctx.cam[b], ctx.cam_init, 8 * 6 * (size_t)ctx.Nc, cudaMemcpyDeviceToDevice, s));
        if (ctx.Np) CU(cudaMemcpyAsync(ctx.pt[b], ctx.pt_init, 8 * 3 * (size_t)ctx.Np, cudaMemcpyDeviceToDevice, s));
    }
    launch_init_state(ctx, s);
    return BA_OK;
}

// =============================================================================================================
// the LM loop (ceres::Solve replacement)
// =============================================================================================================
namespace {

// all-reduce helpers (no-ops on a single GPU, where the *_loc buffers alias their reduced counterparts)
int allreduce_sum(ba_handle *h, const double *src, double *dst, size_t count) {
    if (!h->comm) return BA_OK;
    NC(ncclAllReduce(src, dst, count, ncclDouble, ncclSum, h->comm, h->stream));
    h->launches++;
    return BA_OK;
}
int allreduce_sum(ba_handle *h, double *buf, size_t count) { return allreduce_sum(h, buf, buf, count); }

// sum of the ranks' contributions to the reduced system: ONE all-reduce over [tiles a camera pair can touch | rhs], in place,
// straight out of the buffer the Schur kernels wrote
int allreduce_S(ba_handle *h) {
    if (!h->comm) return BA_OK;
    return allreduce_sum(h, h->ctx.S, (size_t)(h->chol_plan.front_doubles + h->ctx.n));
}

// one linearisation at x: local blocks, then the replicated camera blocks are summed over ranks
int enqueue_linearize(ba_handle *h, int force, int with_begin = 0) {
    const Ctx &ctx = h->ctx;
    launch_linearize_local(ctx, force, h->stream);
    h->launches += 5;
    int rc;
    // ONE message [U | b_c | sums | max slots]; out-of-place so that a body that does not re-linearise reproduces the same
    // reduced blocks (the step slots of `sums` ride along and are overwritten by the step all-reduce before anyone reads them)
    if ((rc = allreduce_sum(h, h->lin_loc.as<double>(), h->lin_red.as<double>(), h->lin_count))) return rc;
    launch_cam_norms(ctx, force, with_begin, h->stream);
    h->launches += 1;
    return BA_OK;
}

// reduced system + solve for the current radius: y_c into ctx.yc
int enqueue_linear_solve(ba_handle *h) {
    const Ctx &ctx = h->ctx;
    cudaStream_t s = h->stream;
    int rc;
    if (h->opt.linear_solver == BA_DENSE_SCHUR) {
        if ((rc = chol_clear(ctx.S, &h->chol_plan, s))) return fail(BA_ERR_CUDA, "chol_clear: %s", cudaGetErrorString((cudaError_t)rc));
        launch_point_invert_schur_build(ctx, s);          // (V + D^2)^-1 and the initialisation of S share one launch
        h->launches += 4;
        if ((rc = allreduce_S(h))) return rc;
        // rhs^T is row ldS of the matrix: the forward substitution is part of the factorisation's dataflow.  The kernel reads it
        // from the camera-ordered rhs through row_src and writes the solution back in camera order (no permutation launches).
        if (h->profile) cudaEventRecord(h->dom_event(), s);
        const CholPerm perm{ctx.row_src, ctx.rhs, ctx.yc};
        const int ce = chol_factor_solve(ctx.S, ctx.ldS, h->invL.as<double>(), h->yperm.as<double>(), ctx.state, nullptr, &h->chol_plan, s, &perm);
        if (ce) return fail(BA_ERR_CUDA, "chol_factor_solve: %s", cudaGetErrorString((cudaError_t)ce));
        if (h->profile) cudaEventRecord(h->dom_event(), s);
        h->launches += 1;
    } else {
        launch_point_invert(ctx, -1.0, s);
        launch_pcg_setup(ctx, s);
        h->launches += 3;
        if ((rc = allreduce_sum(h, ctx.Minv, 36 * (size_t)ctx.Nc))) return rc;
        if ((rc = allreduce_sum(h, ctx.rhs, (size_t)ctx.n))) return rc;
        launch_pcg_start(ctx, s);
        h->launches += 2;
        const int max_it = std::max(1, h->opt.max_linear_solver_iterations);
        // The host runs R CG iterations ahead of the verdict it polls; whatever it enqueues past convergence early-exits on the
        // device but still costs full-grid launches (~0.1 ms per iteration on K5).  One iteration of a large problem takes far
        // longer than enqueueing the next, so two ahead keeps the stream fed there; small problems keep the deeper queue.
        const int R = ctx.No >= (1 << 20) ? 2 : 4;
        for (int it = 1; it <= max_it; ++it) {
            if (it > R) {
                CU(cudaEventSynchronize(h->ev_ring[it % R]));                       // iteration it - R has finished:
                if (h->host_flags[HF_CG_RING + (it - R) % HF_CG_SLOTS]) break;      // ITS verdict (the same on every rank)
            }
            if (h->profile) { cudaEvent_t e0 = h->dom_event(), e1 = h->dom_event(); launch_pcg_iter_a(ctx, s, e0, e1); }
            else launch_pcg_iter_a(ctx, s);
            if ((rc = allreduce_sum(h, ctx.cg_q, (size_t)ctx.n))) return rc;
            launch_pcg_iter_b(ctx, s);
            h->launches += 4;
            if (it % 10 == 0) {   // residual_reset_period
                launch_pcg_reset_a(ctx, s);
                if ((rc = allreduce_sum(h, ctx.cg_z, (size_t)ctx.n))) return rc;
                launch_pcg_reset_b(ctx, s);
                h->launches += 3;   // matvec, fold, reset
            }
            launch_pcg_check(ctx, it, s);
            h->launches += 1;
            CU(cudaEventRecord(h->ev_ring[it % R], s));
        }
        launch_pcg_finish(ctx, s);
        h->launches += 1;
    }
    return BA_OK;
}

} // namespace

extern "C" int ba_solve(ba_handle *h, ba_summary *summary) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "ba_solve before ba_set_problem");
    CU(cudaSetDevice(h->device));
    Ctx &ctx = h->ctx;
    ctx.opt = h->opt;
    ctx.rank = h->rank; ctx.n_ranks = h->n_ranks;
    cudaStream_t s = h->stream;
    int rc;
    if (h->iters.bytes < sizeof(ba_iteration) * (size_t)(h->opt.max_num_iterations + 2)) {
        CU(h->iters.alloc(sizeof(ba_iteration) * (size_t)(h->opt.max_num_iterations + 2)));
        ctx.iters = h->iters.as<ba_iteration>();
    }
    ctx.rhs = h->rhs.as<double>();                  // (ensure_dense moves it into the S buffer: one message [S | rhs])
    if (h->opt.linear_solver == BA_DENSE_SCHUR && (rc = ensure_dense(h))) return rc;
    if (h->opt.linear_solver != BA_DENSE_SCHUR && (rc = ensure_pcg(h))) return rc;
    h->launches = 0;
    h->prof_used = 0;
    h->dom_used = 0; h->dom_seconds = 0.0; h->dom_launches = 0;
    for (int i = 0; i < 16; ++i) h->host_flags[i] = 0;
    CU(cudaEventRecord(h->ev_start, s));
    launch_init_state(ctx, s);
    launch_fill(ctx.scale_cam, 6 * (size_t)ctx.Nc, 1.0, s);
    launch_fill(ctx.scale_pt, 3 * (size_t)ctx.Np, 1.0, s);
    h->launches += 3;
    if (h->opt.jacobi_scaling) {
        // iteration 0: column norms of the unscaled (robustified) Jacobian -> Jacobi scaling, fixed thereafter
        if ((rc = enqueue_linearize(h, 1))) return rc;
        launch_compute_scale(ctx, s);
        h->launches += 1;
    }
    cudaEvent_t *ev_lm = h->ev_lm;                 // created once per handle (ba_create): no event creation per solve
    // At most max_num_iterations + 1 bodies: body k records iteration k and computes step k + 1.
    const int bodies = h->opt.max_num_iterations + 1;
    int status = BA_OK;
    for (int k = 0; k < bodies; ++k) {
        if (k >= 2) {
            cudaError_t e = cudaEventSynchronize(ev_lm[k & 1]);    // body k-2 finished: safe to look at its verdict
            if (e != cudaSuccess) { status = fail(BA_ERR_CUDA, "LM loop: %s", cudaGetErrorString(e)); break; }
            if (h->host_flags[HF_LM_RING + (k - 2) % HF_LM_SLOTS]) break;           // the verdict of body k - 2 (the same on every rank)
        }
        if (h->profile) cudaEventRecord(h->prof_event(), s);
        if ((rc = enqueue_linearize(h, 0, 1))) { status = rc; break; }     // (the last kernel of it also runs controller part 1)
        if (h->profile) cudaEventRecord(h->prof_event(), s);
        if ((rc = enqueue_linear_solve(h))) { status = rc; break; }
        if (h->profile) cudaEventRecord(h->prof_event(), s);
        launch_backsub(ctx, s);
        launch_eval_candidate(ctx, h->cost_part.as<double>(), s);
        if ((rc = allreduce_sum(h, ctx.sums_loc + SUM_STEP_BEGIN, ctx.sums + SUM_STEP_BEGIN, SUM_STEP_N))) { status = rc; break; }
        launch_decide(ctx, k, s);
        h->launches += 5;
        cudaEventRecord(ev_lm[k & 1], s);
    }
    cudaEventRecord(h->ev_stop, s);
    cudaError_t e = cudaStreamSynchronize(s);
    if (status) return status;
    if (e != cudaSuccess) return fail(BA_ERR_CUDA, "LM loop: %s", cudaGetErrorString(e));
    CU(cudaGetLastError());
    LMState st;
    CU(cudaMemcpy(&st, ctx.state, sizeof(st), cudaMemcpyDeviceToHost));
    if (!st.done) return fail(BA_ERR_STATE, "LM loop ended without a verdict (internal error)");
    if (summary) {
        std::memset(summary, 0, sizeof(*summary));
        summary->termination_type = st.termination;
        summary->num_iterations = st.n_recorded;
        summary->num_successful_steps = st.num_successful;
        summary->num_unsuccessful_steps = st.num_unsuccessful;
        summary->initial_cost = st.initial_cost;
        // SetSummaryFinalCost (ceres solver.cc): min over the recorded iteration costs
        std::vector<ba_iteration> its(std::max(st.n_recorded, 1));
        CU(cudaMemcpy(its.data(), ctx.iters, sizeof(ba_iteration) * (size_t)st.n_recorded, cudaMemcpyDeviceToHost));
        double fc = st.initial_cost;
        for (int i = 0; i < st.n_recorded; ++i) fc = std::min(fc, its[i].cost);
        summary->final_cost = fc;
        float ms = 0;
        cudaEventElapsedTime(&ms, h->ev_start, h->ev_stop);
        summary->seconds_total = ms * 1e-3;
        if (h->profile) {
            double t_lin = 0, t_solve = 0;
            for (size_t i = 0; i + 2 < h->prof_used; i += 3) {
                float a = 0, b = 0;
                cudaEventElapsedTime(&a, h->prof_ev[i], h->prof_ev[i + 1]);
                cudaEventElapsedTime(&b, h->prof_ev[i + 1], h->prof_ev[i + 2]);
                t_lin += a; t_solve += b;
            }
            summary->seconds_evaluate = t_lin * 1e-3;
            summary->seconds_linear_solver = t_solve * 1e-3;
            // dominant kernel: every Cholesky launch (dense) / every PCG matvec launch that did work (the host runs a few CG
            // iterations ahead of the convergence verdict; what it enqueues past it early-exits in a fraction of the time)
            std::vector<float> dur;
            float longest = 0;
            for (size_t i = 0; i + 1 < h->dom_used; i += 2) {
                float a = 0;
                cudaEventElapsedTime(&a, h->dom_ev[i], h->dom_ev[i + 1]);
                dur.push_back(a);
                longest = std::max(longest, a);
            }
            const bool pcg = h->opt.linear_solver != BA_DENSE_SCHUR;
            for (float a : dur)
                if (!pcg || a >= 0.5f * longest) { h->dom_seconds += a * 1e-3; h->dom_launches++; }
        }
        std::snprintf(summary->message, sizeof(summary->message), "%s", message_text(st.msg));
    }
    return BA_OK;
}

// Explicit switches of the reduced-system factorisation (tests, measurements; defaults 0 / 0): natural_order != 0 keeps the
// Cholesky tiles in camera order instead of nested dissection; dense_tiles != 0 treats every tile as filled, i.e. factorises S
// as the dense matrix Ceres' DENSE_SCHUR would.  Neither changes a result beyond summation order.  Takes effect at the next
// ba_set_problem (the plan is built once per problem).
extern "C" int ba_set_reduced_system_mode(ba_handle *h, int32_t natural_order, int32_t dense_tiles) {
    if (!h) return fail(BA_ERR_INVALID, "handle is NULL");
    h->natural_tile_order = natural_order != 0;
    h->dense_tiles = dense_tiles != 0;
    h->pairs_built = false;
    return BA_OK;
}

// Device-resident outer loop of GlobalBAOptimizerAngles / LocalBAOptimizerAngles::optimizeCamerasAndMapPoints
// (Optimizer.cpp:222-271, 550-603).  The reference rebuilds the ceres::Problem from the object graph every outer iteration;
// here the problem set by ba_set_problem stays on the device and every step between two solves is a kernel:
//   write-back   landmarks -> float32 (Optimizer.cpp:252-258), poses of the written-back cameras -> inverse(f32 extrinsics)
//                (:262-267); the next build's float -> double / pose -> angle-axis conversion (:282-299) gives the next
//                initial parameters and the extrinsics a fixed-by-ID camera bakes in
//   prune        pruneCorrespondences(frame, considerOutlier = false) over ALL keypoints of the problem's cameras (:270, :597-603)
//   masking      observations whose keypoint became an outlier leave the problem (:308 / :640 isOutlier -> continue); cameras
//                and landmarks left without observations leave it with them.  No re-upload, no index rebuild.
extern "C" int ba_solve_outer(ba_handle *h, int32_t n_outer, const uint8_t *cam_writeback, float *pose, const int32_t *frame_ptr,
                              const float *kp_xy, const int32_t *kp_octave, const int32_t *kp_pt, const float *kp_xyz,
                              const float *pt_min_dist, const float *pt_max_dist, const uint8_t *has_point, const int32_t *kp_obs,
                              uint8_t *outlier, float *pt_out, ba_summary *summary) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "ba_solve_outer before ba_set_problem");
    if (h->comm) return fail(BA_ERR_STATE, "ba_solve_outer runs on a single GPU (the outer loop's prune walks all keypoints)");
    if (n_outer < 0 || !cam_writeback || !pose || !frame_ptr) return fail(BA_ERR_INVALID, "bad arguments");
    CU(cudaSetDevice(h->device));
    Ctx &ctx = h->ctx;
    const int Nc = ctx.Nc, Np = ctx.Np, No = ctx.No, n_kp = frame_ptr[Nc];
    for (int c = 0; c < Nc; ++c)
        if (frame_ptr[c + 1] < frame_ptr[c]) return fail(BA_ERR_INVALID, "frame_ptr not monotone at %d", c);
    if (n_kp && (!kp_xy || !kp_octave || !kp_pt || !kp_xyz || !pt_min_dist || !pt_max_dist || !has_point || !kp_obs || !outlier))
        return fail(BA_ERR_INVALID, "NULL keypoint arrays");
    for (int i = 0; i < n_kp; ++i)
        if (kp_obs[i] >= No || kp_pt[i] >= Np) return fail(BA_ERR_INVALID, "keypoint %d references observation %d / landmark %d out of range", i, kp_obs[i], kp_pt[i]);
    cudaStream_t s = h->stream;
    int rc;
    const size_t nk = (size_t)n_kp;
    if ((rc = upload(h->o_pose, pose, 16 * (size_t)Nc, s))) return rc;
    if ((rc = upload(h->o_wb, cam_writeback, (size_t)Nc, s))) return rc;
    if ((rc = upload(h->o_ptr, frame_ptr, (size_t)Nc + 1, s))) return rc;
    if ((rc = upload(h->o_xy, kp_xy, 2 * nk, s))) return rc;
    if ((rc = upload(h->o_oct, kp_octave, nk, s))) return rc;
    if ((rc = upload(h->o_xyz, kp_xyz, 3 * nk, s))) return rc;
    if ((rc = upload(h->o_mn, pt_min_dist, nk, s))) return rc;
    if ((rc = upload(h->o_mx, pt_max_dist, nk, s))) return rc;
    if ((rc = upload(h->o_has, has_point, nk, s))) return rc;
    if ((rc = upload(h->o_out, outlier, nk, s))) return rc;
    if ((rc = upload(h->o_kp_pt, kp_pt, nk, s))) return rc;
    if ((rc = upload(h->o_kp_obs, kp_obs, nk, s))) return rc;
    CU(h->o_extr.alloc(64 * (size_t)Nc));
    CU(h->o_pt32.alloc(12 * std::max<size_t>(Np, 1)));
    CU(h->o_masked.alloc(std::max<size_t>(No, 1)));
    CU(h->o_active.alloc(4 * (size_t)Nc));
    CU(cudaMemsetAsync(h->o_masked.p, 0, std::max<size_t>(No, 1), s));
    CU(cudaStreamSynchronize(s));                      // the caller's arrays may be pageable: uploads done before we go on
    ba_summary sm;
    std::memset(&sm, 0, sizeof(sm));
    int launches = 0;
    double seconds = 0.0;
    for (int it = 0; it < n_outer; ++it) {
        if (it > 0) {
            // the next problem build: masked observations, cameras without observations, parameters from the float32 state
            launch_outer_mask(n_kp, h->o_kp_obs.as<int>(), h->o_out.as<uint8_t>(), h->perm_dev.as<int>(), Nc, h->raw_fixed.as<unsigned char>(), No,
                              ctx.pm_cam, h->o_masked.as<unsigned char>(), h->o_active.as<int>(), h->cam_kind.as<int>(), h->pm_camk.as<unsigned>(), s);
            ctx.masked = 1;
            if ((rc = ba_reset_parameters(h))) return rc;                  // x <- cam_init / pt_init (written by the write-back kernels)
            launch_cam_prep(ctx, 0, 1, s);                                 // fixed cameras' extrinsics changed with their pose round trip
            launch_cam_prep(ctx, 1, 1, s);
            launches += 10;
        }
        if ((rc = ba_solve(h, &sm))) return rc;
        launches += h->launches;
        seconds += sm.seconds_total;
        LMState st;
        CU(cudaMemcpy(&st, ctx.state, sizeof(st), cudaMemcpyDeviceToHost));
        launch_outer_writeback(Nc, h->o_wb.as<unsigned char>(), ctx.cam[st.cur], h->o_pose.as<float>(), ctx.cam_init, h->cam_extr.as<float>(),
                               3 * (size_t)Np, ctx.pt[st.cur], h->o_pt32.as<float>(), ctx.pt_init, s);
        if (n_kp) {
            launch_outer_kp_xyz(n_kp, h->o_kp_pt.as<int>(), h->o_pt32.as<float>(), h->o_xyz.as<float>(), s);
            launch_pose_inverse(Nc, h->o_pose.as<float>(), h->o_extr.as<float>(), s);       // Optimizer.cpp:10 (all cameras, written back or not)
            launch_prune(Nc, h->o_pose.as<float>(), h->o_extr.as<float>(), ctx.cam_K, h->o_ptr.as<int>(), h->o_xy.as<float2>(), h->o_oct.as<int>(),
                         h->o_xyz.as<float>(), h->o_mn.as<float>(), h->o_mx.as<float>(), h->o_has.as<uint8_t>(), h->o_out.as<uint8_t>(), 0, n_kp, s);
        }
        launches += 5;
    }
    CU(cudaMemcpyAsync(pose, h->o_pose.p, 64 * (size_t)Nc, cudaMemcpyDeviceToHost, s));
    if (pt_out && Np && n_outer > 0) CU(cudaMemcpyAsync(pt_out, h->o_pt32.p, 12 * (size_t)Np, cudaMemcpyDeviceToHost, s));
    if (n_kp) CU(cudaMemcpyAsync(outlier, h->o_out.p, nk, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    h->launches = launches;
    if (summary) { *summary = sm; summary->seconds_total = seconds; }
    return BA_OK;
}

// Accumulation of the PCG matvec: 0 = automatic, 1 = ordered (camera-keyed fold, bit-reproducible), 2 = f64 atomics.
extern "C" int ba_set_pcg_mode(ba_handle *h, int32_t mode) {
    if (!h) return fail(BA_ERR_INVALID, "handle is NULL");
    if (mode < 0 || mode > 2) return fail(BA_ERR_INVALID, "mode must be 0 (automatic), 1 (ordered) or 2 (atomic)");
    h->pcg_mode = mode;
    h->ord_built = false;
    return BA_OK;
}

// Cholesky tile edge of the reduced system: 0 = automatic (64-row tiles for tile-sparse systems, 128-row tiles otherwise), 64 or
// 128 = forced.  Takes effect at the next solve (the plan is rebuilt).
extern "C" int ba_set_reduced_system_tile(ba_handle *h, int32_t tile_edge) {
    if (!h) return fail(BA_ERR_INVALID, "handle is NULL");
    if (tile_edge != 0 && tile_edge != 64 && tile_edge != 128) return fail(BA_ERR_INVALID, "tile_edge must be 0, 64 or 128");
    h->tile_edge = tile_edge;
    h->pairs_built = false;
    return BA_OK;
}

extern "C" int ba_get_launch_count(ba_handle *h) { return h ? h->launches : 0; }
extern "C" int ba_set_profiling(ba_handle *h, int32_t on) {
    if (!h) return fail(BA_ERR_INVALID, "handle is NULL");
    h->profile = on != 0;
    return BA_OK;
}

extern "C" int ba_get_cameras(ba_handle *h, double *cam) {
    if (!h || !h->has_problem || !cam) return fail(BA_ERR_STATE, "no problem / NULL output");
    CU(cudaSetDevice(h->device));
    LMState st;
    CU(cudaMemcpy(&st, h->ctx.state, sizeof(st), cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(cam, h->ctx.cam[st.cur], 8 * 6 * (size_t)h->ctx.Nc, cudaMemcpyDeviceToHost));
    return BA_OK;
}

extern "C" int ba_get_points(ba_handle *h, double *pt) {
    if (!h || !h->has_problem || (!pt && h->ctx.Np)) return fail(BA_ERR_STATE, "no problem / NULL output");
    CU(cudaSetDevice(h->device));
    LMState st;
    CU(cudaMemcpy(&st, h->ctx.state, sizeof(st), cudaMemcpyDeviceToHost));
    if (h->ctx.Np) CU(cudaMemcpy(pt, h->ctx.pt[st.cur], 8 * 3 * (size_t)h->ctx.Np, cudaMemcpyDeviceToHost));
    return BA_OK;
}

extern "C" int ba_get_iterations(ba_handle *h, ba_iteration *iters, int32_t cap, int32_t *n_written) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "no problem set");
    CU(cudaSetDevice(h->device));
    LMState st;
    CU(cudaMemcpy(&st, h->ctx.state, sizeof(st), cudaMemcpyDeviceToHost));
    const int n = std::min<int>(cap, st.n_recorded);
    if (n > 0 && iters) CU(cudaMemcpy(iters, h->ctx.iters, sizeof(ba_iteration) * (size_t)n, cudaMemcpyDeviceToHost));
    if (n_written) *n_written = n;
    return BA_OK;
}

// =============================================================================================================
// materialised evaluation, reduced system dump
// =============================================================================================================
namespace {
int ensure_eval_buffers(ba_handle *h) {
    const size_t No = (size_t)h->ctx.No;
    if (h->ev_res.bytes < 16 * std::max<size_t>(No, 1)) {
        CU(h->ev_res.alloc(16 * std::max<size_t>(No, 1)));
        CU(h->ev_jc.alloc(96 * std::max<size_t>(No, 1)));
        CU(h->ev_jp.alloc(48 * std::max<size_t>(No, 1)));
    }
    return BA_OK;
}
} // namespace

extern "C" int ba_evaluate(ba_handle *h, double *cost, double *residuals, double *jac_cam, double *jac_pt,
                           double *grad_cam, double *grad_pt) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "ba_evaluate before ba_set_problem");
    CU(cudaSetDevice(h->device));
    Ctx &ctx = h->ctx;
    ctx.opt = h->opt;
    cudaStream_t s = h->stream;
    int rc;
    if ((rc = ensure_eval_buffers(h))) return rc;
    const int No = ctx.No, Nc = ctx.Nc, Np = ctx.Np;
    launch_cam_prep(ctx, 0, 1, s);
    const int nb = launch_materialize(ctx, h->ev_res.as<double>(), h->ev_jc.as<double>(), h->ev_jp.as<double>(), h->cost_part.as<double>(), s);
    // gradient: unscaled blocks b_c, b_p of a forced linearisation with unit scale
    if (grad_cam || grad_pt) {
        launch_fill(ctx.scale_cam, 6 * (size_t)Nc, 1.0, s);
        launch_fill(ctx.scale_pt, 3 * (size_t)Np, 1.0, s);
        if ((rc = enqueue_linearize(h, 1))) return rc;
    }
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    std::vector<double> part(std::max(nb, 1));
    if (nb) CU(cudaMemcpy(part.data(), h->cost_part.p, 8 * (size_t)nb, cudaMemcpyDeviceToHost));
    if (cost) {
        double c = 0;
        for (int i = 0; i < nb; ++i) c += part[i];
        if (h->comm) {
            // sum over ranks through the device
            double *tmp = h->camsums.as<double>() + CS_COUNT - 1;      // spare slot (rewritten by every linearisation)
            CU(cudaMemcpy(tmp, &c, 8, cudaMemcpyHostToDevice));
            NC(ncclAllReduce(tmp, tmp, 1, ncclDouble, ncclSum, h->comm, s));
            CU(cudaStreamSynchronize(s));
            CU(cudaMemcpy(&c, tmp, 8, cudaMemcpyDeviceToHost));
        }
        *cost = c;
    }
    if (!h->perm_valid && No && (residuals || jac_cam || jac_pt)) {
        h->perm.resize(No);
        CU(cudaMemcpy(h->perm.data(), h->perm_dev.p, 4 * (size_t)No, cudaMemcpyDeviceToHost));
        h->perm_valid = true;
    }
    auto gather = [&](const DevBuf &src, double *dst, int width) -> int {
        std::vector<double> tmp((size_t)No * width);
        if (No) CU(cudaMemcpy(tmp.data(), src.p, 8 * tmp.size(), cudaMemcpyDeviceToHost));
        for (int i = 0; i < No; ++i) std::memcpy(dst + (size_t)i * width, tmp.data() + (size_t)h->perm[i] * width, 8 * (size_t)width);
        return BA_OK;
    };
    if (residuals && (rc = gather(h->ev_res, residuals, 2))) return rc;
    if (jac_cam && (rc = gather(h->ev_jc, jac_cam, 12))) return rc;
    if (jac_pt && (rc = gather(h->ev_jp, jac_pt, 6))) return rc;
    if (grad_cam) {
        CU(cudaMemcpy(grad_cam, ctx.bc, 8 * 6 * (size_t)Nc, cudaMemcpyDeviceToHost));
        std::vector<int> kind(Nc);
        CU(cudaMemcpy(kind.data(), ctx.cam_kind, 4 * (size_t)Nc, cudaMemcpyDeviceToHost));
        for (int c = 0; c < Nc; ++c) if (kind[c] != CAM_FREE) for (int k = 0; k < 6; ++k) grad_cam[6 * (size_t)c + k] = 0.0;
    }
    if (grad_pt && Np) CU(cudaMemcpy(grad_pt, ctx.bp, 8 * 3 * (size_t)Np, cudaMemcpyDeviceToHost));
    return BA_OK;
}

extern "C" int ba_evaluate_device(ba_handle *h, int32_t repeats, float *ms_per_call) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "ba_evaluate_device before ba_set_problem");
    CU(cudaSetDevice(h->device));
    Ctx &ctx = h->ctx;
    ctx.opt = h->opt;
    cudaStream_t s = h->stream;
    int rc;
    if ((rc = ensure_eval_buffers(h))) return rc;
    if (repeats < 1) repeats = 1;
    CU(cudaEventRecord(h->ev_start, s));
    for (int r = 0; r < repeats; ++r) {
        launch_cam_prep(ctx, 0, 1, s);
        launch_materialize(ctx, h->ev_res.as<double>(), h->ev_jc.as<double>(), h->ev_jp.as<double>(), h->cost_part.as<double>(), s);
    }
    CU(cudaEventRecord(h->ev_stop, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, h->ev_start, h->ev_stop));
    if (ms_per_call) *ms_per_call = ms / repeats;
    return BA_OK;
}

extern "C" int ba_schur_system(ba_handle *h, double radius, double *S, double *rhs, double *cost) {
    if (!h || !h->has_problem) return fail(BA_ERR_STATE, "ba_schur_system before ba_set_problem");
    if (!(radius > 0)) return fail(BA_ERR_INVALID, "radius must be positive");
    CU(cudaSetDevice(h->device));
    Ctx &ctx = h->ctx;
    ctx.opt = h->opt;
    ctx.rank = h->rank; ctx.n_ranks = h->n_ranks;
    cudaStream_t s = h->stream;
    int rc;
    if ((rc = ensure_dense(h))) return rc;
    launch_fill(ctx.scale_cam, 6 * (size_t)ctx.Nc, 1.0, s);
    launch_fill(ctx.scale_pt, 3 * (size_t)ctx.Np, 1.0, s);
    if ((rc = enqueue_linearize(h, 1))) return rc;
    if (h->opt.jacobi_scaling) {
        launch_compute_scale(ctx, s);
        if ((rc = enqueue_linearize(h, 1))) return rc;
    }
    launch_point_invert(ctx, radius, s);
    if ((rc = chol_clear(ctx.S, &h->chol_plan, s))) return fail(BA_ERR_CUDA, "chol_clear: %s", cudaGetErrorString((cudaError_t)rc));
    launch_schur_build(ctx, radius, s);
    if ((rc = allreduce_S(h))) return rc;
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    if (S) {
        // the device matrix is in elimination order (lower blocks): hand it back in camera order, lower blocks
        const size_t n = (size_t)ctx.n;
        const CholPlan &pl = h->chol_plan;
        std::vector<double> Sp((size_t)pl.total_doubles);
        std::vector<int> cam_row(ctx.Nc);
        CU(cudaMemcpy(Sp.data(), ctx.S, 8 * Sp.size(), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(cam_row.data(), ctx.cam_row, 4 * (size_t)ctx.Nc, cudaMemcpyDeviceToHost));
        const auto at = [&](size_t r, size_t q) -> double {                  // element (r, q) of the tile-packed lower triangle
            const size_t nb = (size_t)pl.nb;
            const long long o = pl.h_tile_off[(r / nb) * (size_t)pl.T + q / nb];
            return o < 0 ? 0.0 : Sp[(size_t)o + (r % nb) * nb + q % nb];
        };
        std::memset(S, 0, 8 * n * n);
        for (int ci = 0; ci < ctx.Nc; ++ci)
            for (int cj = 0; cj <= ci; ++cj) {
                const size_t r = cam_row[ci], q = cam_row[cj];
                for (int a = 0; a < 6; ++a)
                    for (int b = 0; b < 6; ++b)
                        S[(6 * (size_t)ci + a) * n + 6 * (size_t)cj + b] = (r >= q) ? at(r + a, q + b) : at(q + b, r + a);
            }
    }
    if (rhs) CU(cudaMemcpy(rhs, ctx.rhs, 8 * (size_t)ctx.n, cudaMemcpyDeviceToHost));
    if (cost) CU(cudaMemcpy(cost, ctx.sums + SUM_LIN_COST, 8, cudaMemcpyDeviceToHost));
    return BA_OK;
}

// Profiling detail of the last ba_solve (ba_set_profiling on): out = {seconds inside the dominant linear-solver kernel (CUDA
// events on the solver's stream around each launch): the dense-Cholesky kernel, memset of its flags included, or -- with the
// iterative solver -- the PCG matvec launches that did work, number of such launches, 0, 0}.
extern "C" int ba_get_kernel_profile(ba_handle *h, double out[4]) {
    if (!h || !out) return fail(BA_ERR_INVALID, "NULL argument");
    out[0] = h->dom_seconds; out[1] = (double)h->dom_launches; out[2] = 0.0; out[3] = 0.0;
    return BA_OK;
}

// Reduced-system bookkeeping of the current problem (dense Schur): out = {order n, filled tiles of L, tiles in the lower
// triangle, flops of one factorisation + solve over the filled tiles}.  Builds the pair index / plan on first use.
extern "C" int ba_get_reduced_system_stats(ba_handle *h, double out[4]) {
    if (!h || !out) return fail(BA_ERR_INVALID, "NULL argument");
    if (!h->has_problem) return fail(BA_ERR_STATE, "no problem set");
    CU(cudaSetDevice(h->device));
    int rc;
    if ((rc = ensure_dense(h))) return rc;
    const double T = (double)h->chol_plan.T;
    out[0] = (double)h->ctx.ldS; out[1] = (double)h->chol_plan.filled_tiles; out[2] = T * (T + 1) / 2; out[3] = h->chol_plan.flops;
    return BA_OK;
}

// Host-only: elimination order of the Cholesky tiles for a tile adjacency (G x G bytes, lower triangle read), as used for
// the reduced camera system.  order[pos] = tile.  stats (may be NULL) = {dependency depth in natural order, depth in the
// returned order, filled tiles natural, filled tiles ordered}.
extern "C" int ba_tile_elimination_order(int32_t G, const uint8_t *adj, int32_t *order, double stats[4]) {
    if (G <= 0 || !adj || !order) return fail(BA_ERR_INVALID, "bad arguments");
    std::vector<int> ord;
    chol_tile_order(G, adj, ord);
    if ((int)ord.size() != G) return fail(BA_ERR_STATE, "tile ordering lost nodes (internal error)");
    for (int p = 0; p < G; ++p) order[p] = ord[p];
    if (stats) {
        std::vector<int> pos(G);
        for (int p = 0; p < G; ++p) pos[ord[p]] = p;
        std::vector<unsigned char> nat((size_t)G * G, 0), per((size_t)G * G, 0);
        for (int i = 0; i < G; ++i)
            for (int j = 0; j <= i; ++j)
                if (i == j || adj[(size_t)i * G + j]) {
                    nat[(size_t)i * G + j] = 1;
                    per[(size_t)std::max(pos[i], pos[j]) * G + std::min(pos[i], pos[j])] = 1;
                }
        int d0 = 0, d1 = 0;
        long long f0 = 0, f1 = 0;
        chol_plan_stats(G * 128, 128, nat.data(), &d0, &f0, nullptr);
        chol_plan_stats(G * 128, 128, per.data(), &d1, &f1, nullptr);
        stats[0] = d0; stats[1] = d1; stats[2] = (double)f0; stats[3] = (double)f1;
    }
    return BA_OK;
}

// Dense SPD solve utility: exposes the reduced-system Cholesky on its own (tests, micro-benchmarks).
// A: n x n row-major (lower triangle read), b: n.  x: n.  ms (optional): device time of factor + solve.
static int dense_spd_solve_impl(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats, int tile_edge);
extern "C" int ba_dense_spd_solve(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats) {
    return dense_spd_solve_impl(device, n, A, b, x, ms, repeats, 0);
}
// tile_edge: 0 = automatic (by the tile sparsity of A), 64 or 128 = forced (tests exercise both instantiations of the kernel)
extern "C" int ba_dense_spd_solve_tile(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats, int32_t tile_edge) {
    if (tile_edge != 0 && tile_edge != 64 && tile_edge != 128) return fail(BA_ERR_INVALID, "tile_edge must be 0, 64 or 128");
    return dense_spd_solve_impl(device, n, A, b, x, ms, repeats, tile_edge);
}
static int dense_spd_solve_impl(int device, int32_t n, const double *A, const double *b, double *x, float *ms, int32_t repeats, int tile_edge) {
    if (n <= 0 || !A || !b || !x) return fail(BA_ERR_INVALID, "bad arguments");
    int rc = check_device(device);
    if (rc) return rc;
    int e0 = chol_init();
    if (e0) return fail(BA_ERR_CUDA, "chol_init: %s", cudaGetErrorString((cudaError_t)e0));
    // tile pattern of the caller's matrix (exact zeros only) -> the same sparsity plan (and tile-packed storage) the LM loop uses;
    // 64-row tiles when at most a quarter of the lower tiles are touched, like ensure_dense
    int nb = tile_edge == 64 || tile_edge == 128 ? tile_edge : 64;
    std::vector<unsigned char> nz;
    int np = 0;
    for (;;) {
        np = chol_ld(n, nb);
        const int T = np / nb;
        nz.assign((size_t)T * T, 0);
        for (int i = 0; i < n; ++i) {
            const double *row = A + (size_t)i * n;
            for (int j = 0; j <= i; ++j) if (row[j] != 0.0) nz[(size_t)(i / nb) * T + j / nb] = 1;
        }
        long long nnz = 0;
        for (int i = 0; i < T; ++i) for (int j = 0; j <= i; ++j) nnz += (i == j || nz[(size_t)i * T + j]) ? 1 : 0;
        if (nb == 128 || tile_edge || 2 * nnz <= (long long)T * (T + 1) / 2 / 2 || T <= CHOL_SMALL_TILE_ROWS) break;
        nb = 128;
    }
    CholPlan plan;
    if ((rc = chol_plan_create(n, nb, nz.data(), &plan, 0))) return fail(BA_ERR_CUDA, "chol_plan_create: %s", cudaGetErrorString((cudaError_t)rc));
    std::vector<double> Ah;
    chol_pack_host(&plan, A, b, Ah);
    DevBuf dA, dA0, dInv, dY, dFail;
    CU(dA.alloc(8 * Ah.size())); CU(dA0.alloc(8 * Ah.size()));
    CU(dInv.alloc(8 * chol_work_doubles(n, nb))); CU(dY.alloc(8 * (size_t)np)); CU(dFail.alloc(4));
    CU(cudaMemcpy(dA0.p, Ah.data(), 8 * Ah.size(), cudaMemcpyHostToDevice));
    CU(cudaMemset(dFail.p, 0, 4));
    CU(cudaMemset(dInv.p, 0, 8 * chol_work_doubles(n, nb)));
    cudaEvent_t e1, e2;
    CU(cudaEventCreate(&e1)); CU(cudaEventCreate(&e2));
    if (repeats < 1) repeats = 1;
    float best = 1e30f;
    for (int r = 0; r < repeats; ++r) {
        CU(cudaMemcpy(dA.p, dA0.p, 8 * Ah.size(), cudaMemcpyDeviceToDevice));
        CU(cudaEventRecord(e1, 0));
        const int ce = chol_factor_solve(dA.as<double>(), n, dInv.as<double>(), dY.as<double>(), nullptr, dFail.as<int>(), &plan, 0);
        if (ce) return fail(BA_ERR_CUDA, "chol_factor_solve: %s", cudaGetErrorString((cudaError_t)ce));
        CU(cudaEventRecord(e2, 0));
        CU(cudaDeviceSynchronize());
        float t = 0;
        CU(cudaEventElapsedTime(&t, e1, e2));
        best = std::min(best, t);
    }
    CU(cudaGetLastError());
    int failed = 0;
    CU(cudaMemcpy(&failed, dFail.p, 4, cudaMemcpyDeviceToHost));
    std::vector<double> y(np);
    CU(cudaMemcpy(y.data(), dY.p, 8 * (size_t)np, cudaMemcpyDeviceToHost));
    std::memcpy(x, y.data(), 8 * (size_t)n);
    if (ms) *ms = best;
    cudaEventDestroy(e1); cudaEventDestroy(e2);
    dA.free(); dA0.free(); dInv.free(); dY.free(); dFail.free();
    chol_plan_destroy(&plan);
    if (failed) return fail(BA_ERR_INVALID, "matrix is not positive definite");
    return BA_OK;
}

// =============================================================================================================
// batched frame-to-frame pose estimation, pruning
// =============================================================================================================
namespace {

// Persistent per-thread, per-device workspace of the handle-less entry points (pose batch, prune, pose outer loop): one stream,
// two events and grow-only device buffers, created on first use and reused by every later call of the thread -- no cudaMalloc,
// stream or event creation on the steady-state path (the tracker calls these once per frame).  Released by
// ba_release_workspace() or with the process.
struct Workspace {
    int device = -1;
    cudaStream_t stream = nullptr;
    cudaEvent_t e1 = nullptr, e2 = nullptr;
    DevBuf cam, K, ptr, Xh, uv, mask, sum, it, pose, extr, xy, oct, xyz, mn, mx, has, out;
    int launches = 0;
    void release() {
        if (device < 0) return;
        cudaSetDevice(device);
        DevBuf *all[] = {&cam, &K, &ptr, &Xh, &uv, &mask, &sum, &it, &pose, &extr, &xy, &oct, &xyz, &mn, &mx, &has, &out};
        for (DevBuf *b : all) b->free();
        if (e1) cudaEventDestroy(e1);
        if (e2) cudaEventDestroy(e2);
        if (stream) cudaStreamDestroy(stream);
        stream = nullptr; e1 = e2 = nullptr; device = -1;
    }
};
thread_local Workspace *g_ws[64];

int workspace(int device, Workspace **out) {
    int rc = check_device(device);
    if (rc) return rc;
    if (device >= 64) return fail(BA_ERR_INVALID, "device %d: at most 64 devices", device);
    Workspace *&w = g_ws[device];
    if (!w) w = new Workspace();
    if (w->device < 0) {
        CU(cudaStreamCreateWithFlags(&w->stream, cudaStreamNonBlocking));
        CU(cudaEventCreate(&w->e1));
        CU(cudaEventCreate(&w->e2));
        w->device = device;
    }
    *out = w;
    return BA_OK;
}

template <typename T>
int ws_upload(DevBuf &b, const T *src, size_t count, cudaStream_t s) { return upload(b, src, count, s); }

} // namespace

extern "C" int ba_release_workspace(void) {
    for (Workspace *&w : g_ws)
        if (w) { w->release(); delete w; w = nullptr; }
    return BA_OK;
}

extern "C" int ba_solve_pose_only_batch(const ba_options *opt_in, int device, int32_t B, double *cam, const float *K,
                                        const int32_t *prob_ptr, const float *Xh, const float *uv, const uint8_t *obs_mask,
                                        ba_summary *summaries, ba_iteration *iters, int32_t iters_cap) {
    if (B < 0 || (B && (!cam || !K || !prob_ptr))) return fail(BA_ERR_INVALID, "bad arguments");
    Workspace *w;
    int rc = workspace(device, &w);
    if (rc) return rc;
    if (B == 0) return BA_OK;
    ba_options opt;
    if (opt_in) opt = *opt_in; else { ba_default_options(&opt); opt.max_num_iterations = 20; }
    const int No = prob_ptr[B];
    for (int b = 0; b < B; ++b)
        if (prob_ptr[b + 1] < prob_ptr[b]) return fail(BA_ERR_INVALID, "prob_ptr not monotone at %d", b);
    if (No && (!Xh || !uv)) return fail(BA_ERR_INVALID, "NULL observation arrays");
    if (iters && iters_cap <= 0) return fail(BA_ERR_INVALID, "iters given but iters_cap <= 0");
    cudaStream_t s = w->stream;
    if ((rc = ws_upload(w->cam, cam, 6 * (size_t)B, s))) return rc;
    if ((rc = ws_upload(w->K, K, 9 * (size_t)B, s))) return rc;
    if ((rc = ws_upload(w->ptr, prob_ptr, (size_t)B + 1, s))) return rc;
    if ((rc = ws_upload(w->Xh, Xh, 4 * (size_t)No, s))) return rc;
    if ((rc = ws_upload(w->uv, uv, 2 * (size_t)No, s))) return rc;
    if (obs_mask && (rc = ws_upload(w->mask, obs_mask, (size_t)No, s))) return rc;
    CU(w->sum.alloc(sizeof(ba_summary) * (size_t)B));
    if (iters) {
        CU(w->it.alloc(sizeof(ba_iteration) * (size_t)B * iters_cap));
        CU(cudaMemsetAsync(w->it.p, 0, w->it.bytes, s));
    }
    CU(cudaEventRecord(w->e1, s));
    launch_pose_batch(opt, B, w->cam.as<double>(), w->K.as<float>(), w->ptr.as<int>(), w->Xh.as<float4>(), w->uv.as<float2>(),
                      obs_mask ? w->mask.as<uint8_t>() : nullptr, w->sum.as<ba_summary>(), iters ? w->it.as<ba_iteration>() : nullptr,
                      iters_cap, s);
    w->launches = 1;
    CU(cudaEventRecord(w->e2, s));
    CU(cudaMemcpyAsync(cam, w->cam.p, 48 * (size_t)B, cudaMemcpyDeviceToHost, s));
    std::vector<ba_summary> sm(B);
    CU(cudaMemcpyAsync(sm.data(), w->sum.p, sizeof(ba_summary) * (size_t)B, cudaMemcpyDeviceToHost, s));
    if (iters) CU(cudaMemcpyAsync(iters, w->it.p, w->it.bytes, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    float ms = 0;
    cudaEventElapsedTime(&ms, w->e1, w->e2);
    if (summaries) {
        for (int b = 0; b < B; ++b) {
            ba_summary &o = summaries[b];
            o = sm[b];
            const int msg = (int)(unsigned char)o.message[0];
            std::snprintf(o.message, sizeof(o.message), "%s", message_text(msg));
            o.seconds_total = ms * 1e-3;     // whole batch (one launch)
            if (iters) {
                double fc = o.initial_cost;
                const int n = std::min(o.num_iterations, iters_cap);
                for (int i = 0; i < n; ++i) fc = std::min(fc, iters[(size_t)b * iters_cap + i].cost);
                o.final_cost = fc;
            }
        }
    }
    return BA_OK;
}

namespace {
// keypoint arrays of a set of frames -> workspace (shared by ba_prune_frames and ba_solve_pose_only_outer)
int upload_keypoints(Workspace *w, int F, const float *pose, const float *K, const int32_t *frame_ptr, const float *kp_xy,
                     const int32_t *kp_octave, const float *pt_xyz, const float *pt_min_dist, const float *pt_max_dist,
                     const uint8_t *has_point, const uint8_t *outlier) {
    cudaStream_t s = w->stream;
    const size_t n = (size_t)frame_ptr[F];
    int rc;
    if ((rc = ws_upload(w->pose, pose, 16 * (size_t)F, s))) return rc;
    if ((rc = ws_upload(w->K, K, 9 * (size_t)F, s))) return rc;
    if ((rc = ws_upload(w->ptr, frame_ptr, (size_t)F + 1, s))) return rc;
    if ((rc = ws_upload(w->xy, kp_xy, 2 * n, s))) return rc;
    if ((rc = ws_upload(w->oct, kp_octave, n, s))) return rc;
    if ((rc = ws_upload(w->xyz, pt_xyz, 3 * n, s))) return rc;
    if ((rc = ws_upload(w->mn, pt_min_dist, n, s))) return rc;
    if ((rc = ws_upload(w->mx, pt_max_dist, n, s))) return rc;
    if ((rc = ws_upload(w->has, has_point, n, s))) return rc;
    if ((rc = ws_upload(w->out, outlier, n, s))) return rc;
    CU(w->extr.alloc(64 * (size_t)F));
    return BA_OK;
}
} // namespace

extern "C" int ba_prune_frames(int device, int32_t n_frames, const float *pose, const float *K, const int32_t *frame_ptr,
                               const float *kp_xy, const int32_t *kp_octave, const float *pt_xyz, const float *pt_min_dist,
                               const float *pt_max_dist, const uint8_t *has_point, uint8_t *outlier, int32_t consider_outlier) {
    if (n_frames < 0 || (n_frames && (!pose || !K || !frame_ptr))) return fail(BA_ERR_INVALID, "bad arguments");
    Workspace *w;
    int rc = workspace(device, &w);
    if (rc) return rc;
    if (n_frames == 0) return BA_OK;
    const int n_kp = frame_ptr[n_frames];
    if (n_kp == 0) return BA_OK;
    if (!kp_xy || !kp_octave || !pt_xyz || !pt_min_dist || !pt_max_dist || !has_point || !outlier) return fail(BA_ERR_INVALID, "NULL keypoint arrays");
    cudaStream_t s = w->stream;
    if ((rc = upload_keypoints(w, n_frames, pose, K, frame_ptr, kp_xy, kp_octave, pt_xyz, pt_min_dist, pt_max_dist, has_point, outlier))) return rc;
    launch_pose_inverse(n_frames, w->pose.as<float>(), w->extr.as<float>(), s);       // extr = pose.inverse() in float32 (Optimizer.cpp:10)
    launch_prune(n_frames, w->pose.as<float>(), w->extr.as<float>(), w->K.as<float>(), w->ptr.as<int>(), w->xy.as<float2>(), w->oct.as<int>(),
                 w->xyz.as<float>(), w->mn.as<float>(), w->mx.as<float>(), w->has.as<uint8_t>(), w->out.as<uint8_t>(), consider_outlier, n_kp, s);
    w->launches = 2;
    CU(cudaMemcpyAsync(outlier, w->out.p, (size_t)n_kp, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    return BA_OK;
}

// SfMHelper::estimatePoseUsingBA's loop (SfMHelper.cpp:106-117: setNbOfIterations(4), setNbOfMaxItPerBA(20)) over a batch of
// frames, device-resident: per outer iteration  pose -> (angle-axis, t)  |  observation mask = has point && !outlier  |  one
// ceres::Solve replacement per frame (pose_batch_kernel)  |  write-back pose = inverse(f32 extrinsics)  |
// pruneCorrespondences(frame, considerOutlier = true) -- five launches, nothing crosses PCIe between them.
extern "C" int ba_solve_pose_only_outer(const ba_options *opt_in, int device, int32_t n_frames, float *pose, const float *K,
                                        const int32_t *frame_ptr, const float *kp_xy, const int32_t *kp_octave,
                                        const float *pt_xyz, const float *pt_min_dist, const float *pt_max_dist,
                                        const uint8_t *has_point, uint8_t *outlier, int32_t n_outer, ba_summary *summaries,
                                        float *device_ms, int64_t *lm_iterations) {
    if (n_frames < 0 || n_outer < 0 || (n_frames && (!pose || !K || !frame_ptr))) return fail(BA_ERR_INVALID, "bad arguments");
    Workspace *w;
    int rc = workspace(device, &w);
    if (rc) return rc;
    if (device_ms) *device_ms = 0.f;
    if (lm_iterations) *lm_iterations = 0;
    if (n_frames == 0) return BA_OK;
    const int F = n_frames, n_kp = frame_ptr[F];
    for (int f = 0; f < F; ++f)
        if (frame_ptr[f + 1] < frame_ptr[f]) return fail(BA_ERR_INVALID, "frame_ptr not monotone at %d", f);
    if (n_kp && (!kp_xy || !kp_octave || !pt_xyz || !pt_min_dist || !pt_max_dist || !has_point || !outlier)) return fail(BA_ERR_INVALID, "NULL keypoint arrays");
    ba_options opt;
    if (opt_in) opt = *opt_in; else { ba_default_options(&opt); opt.max_num_iterations = 20; }
    cudaStream_t s = w->stream;
    if ((rc = upload_keypoints(w, F, pose, K, frame_ptr, kp_xy, kp_octave, pt_xyz, pt_min_dist, pt_max_dist, has_point, outlier))) return rc;
    CU(w->cam.alloc(48 * (size_t)F));
    CU(w->sum.alloc(sizeof(ba_summary) * (size_t)F * std::max(n_outer, 1)));      // one slice per outer iteration
    CU(w->mask.alloc(std::max<size_t>(n_kp, 1)));
    CU(w->Xh.alloc(16 * std::max<size_t>(n_kp, 1)));
    CU(cudaEventRecord(w->e1, s));
    w->launches = 0;
    for (int it = 0; it < n_outer; ++it) {
        launch_pose_to_cam(F, w->pose.as<float>(), w->cam.as<double>(), s);
        launch_pose_mask(n_kp, w->has.as<uint8_t>(), w->out.as<uint8_t>(), w->mask.as<uint8_t>(), w->xyz.as<float>(), it == 0 ? w->Xh.as<float4>() : nullptr, s);
        launch_pose_batch(opt, F, w->cam.as<double>(), w->K.as<float>(), w->ptr.as<int>(), w->Xh.as<float4>(), w->xy.as<float2>(),
                          w->mask.as<uint8_t>(), w->sum.as<ba_summary>() + (size_t)it * F, nullptr, 0, s);
        launch_cam_to_pose(F, w->cam.as<double>(), w->pose.as<float>(), w->extr.as<float>(), s);
        launch_prune(F, w->pose.as<float>(), w->extr.as<float>(), w->K.as<float>(), w->ptr.as<int>(), w->xy.as<float2>(), w->oct.as<int>(),
                     w->xyz.as<float>(), w->mn.as<float>(), w->mx.as<float>(), w->has.as<uint8_t>(), w->out.as<uint8_t>(), 1, n_kp, s);
        w->launches += n_kp ? 5 : 3;
    }
    CU(cudaEventRecord(w->e2, s));
    CU(cudaMemcpyAsync(pose, w->pose.p, 64 * (size_t)F, cudaMemcpyDeviceToHost, s));
    if (n_kp) CU(cudaMemcpyAsync(outlier, w->out.p, (size_t)n_kp, cudaMemcpyDeviceToHost, s));
    std::vector<ba_summary> sm((summaries || lm_iterations) ? (size_t)F * n_outer : 0);
    if (!sm.empty()) CU(cudaMemcpyAsync(sm.data(), w->sum.p, sizeof(ba_summary) * sm.size(), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    float ms = 0;
    cudaEventElapsedTime(&ms, w->e1, w->e2);
    if (device_ms) *device_ms = ms;
    if (lm_iterations)
        for (const ba_summary &x : sm) *lm_iterations += std::max(x.num_iterations - 1, 0);
    for (size_t b = 0; summaries && n_outer && b < (size_t)F; ++b) {
        ba_summary &o = summaries[b];
        o = sm[(size_t)(n_outer - 1) * F + b];
        const int msg = (int)(unsigned char)o.message[0];
        std::snprintf(o.message, sizeof(o.message), "%s", message_text(msg));
        o.seconds_total = ms * 1e-3;         // whole batch, all outer iterations
    }
    return BA_OK;
}

// kernels the last handle-less call of this thread launched on `device` (benchmark bookkeeping)
extern "C" int ba_workspace_launch_count(int device) {
    return (device >= 0 && device < 64 && g_ws[device]) ? g_ws[device]->launches : 0;
}

// =============================================================================================================
// multi-GPU plumbing: NCCL communicator per handle (one process per GPU; the unique id travels through the
// caller's own control plane, e.g. a torch.distributed broadcast)
// =============================================================================================================
extern "C" int ba_comm_unique_id(uint8_t id[128]) {
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    ncclUniqueId uid;
    NC(ncclGetUniqueId(&uid));
    std::memcpy(id, &uid, 128);
    return BA_OK;
}

extern "C" int ba_comm_init(ba_handle *h, int32_t n_ranks, int32_t rank, const uint8_t id[128]) {
    if (!h || !id || n_ranks < 1 || rank < 0 || rank >= n_ranks) return fail(BA_ERR_INVALID, "bad arguments");
    if (h->has_problem) return fail(BA_ERR_STATE, "ba_comm_init must precede ba_set_problem");
    CU(cudaSetDevice(h->device));
    if (h->comm) { ncclCommDestroy(h->comm); h->comm = nullptr; }
    h->n_ranks = n_ranks; h->rank = rank;
    if (n_ranks == 1) return BA_OK;
    ncclUniqueId uid;
    std::memcpy(&uid, id, 128);
    NC(ncclCommInitRank(&h->comm, n_ranks, uid, rank));
    return BA_OK;
}
