// ba_prep.cu -- device-side index building for ba_set_problem: what prepareConstraints' AddResidualBlock loops
// (ba_project/src/ba/Optimizer.cpp:301-331, 636-696) and Ceres' program/ordering preprocessing did on the host per
// ceres::Problem is done here with a handful of kernels + CUB radix sorts / scans (CUB = plumbing: sorting integer keys;
// every arithmetic kernel of the solver is hand-written).  Three orders of the same observations are produced:
//   landmark-major (sorted by landmark, then camera)   -> V, b_p, W, back-substitution, PCG, the sharding key
//   camera-major   (free cameras only, <=128-obs chunks) -> U, b_c by deterministic warp reductions
//   camera-PAIR-major (dense Schur only)                -> S blocks by deterministic segmented reductions, no atomics
#include "ba_launch.h"

#include <cub/cub.cuh>

namespace ba {

static inline int cdivp(long long a, int b) { return (int)((a + b - 1) / b); }

// ---- kernels ------------------------------------------------------------------------------------------------
__global__ void prep_keys_kernel(int No, int Nc, int Np, const int *__restrict__ obs_cam, const int *__restrict__ obs_pt,
                                 unsigned long long *__restrict__ key, int *__restrict__ idx, int *__restrict__ cnt_cam,
                                 int *__restrict__ cnt_pt, int *__restrict__ flags /*[0] bad index, [1] unsorted*/) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= No) return;
    const int c = obs_cam[i], p = obs_pt[i];
    if (c < 0 || c >= Nc || p < 0 || p >= Np) { atomicExch(flags, i + 1); key[i] = 0; idx[i] = i; return; }
    const unsigned long long k = (unsigned long long)p * (unsigned long long)Nc + (unsigned long long)c;
    key[i] = k;
    idx[i] = i;
    atomicAdd(cnt_cam + c, 1);
    atomicAdd(cnt_pt + p, 1);
    if (i > 0) {
        const int c0 = obs_cam[i - 1], p0 = obs_pt[i - 1];
        if (c0 >= 0 && c0 < Nc && p0 >= 0 && p0 < Np) {
            const unsigned long long k0 = (unsigned long long)p0 * (unsigned long long)Nc + (unsigned long long)c0;
            if (k0 > k) flags[1] = 1;
        }
    }
}

__global__ void prep_gather_pm_kernel(int No, int Nc, const unsigned long long *__restrict__ key, const int *__restrict__ order,
                                      const float2 *__restrict__ obs_uv, float2 *__restrict__ pm_uv, int *__restrict__ pm_cam,
                                      int *__restrict__ pm_pt, int *__restrict__ perm) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= No) return;
    const unsigned long long k = key[e];
    const int i = order[e];
    pm_pt[e] = (int)(k / (unsigned long long)Nc);
    pm_cam[e] = (int)(k % (unsigned long long)Nc);
    pm_uv[e] = obs_uv[i];
    perm[i] = e;
}

__global__ void prep_kind_kernel(int Nc, const unsigned char *__restrict__ cam_fixed, const int *__restrict__ cnt_global,
                                 const int *__restrict__ cnt_local, int *__restrict__ kind, int *__restrict__ free_cnt,
                                 int *__restrict__ chunk_cnt) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc) return;
    const int k = cam_fixed[c] ? CAM_FIXED : (cnt_global[c] > 0 ? CAM_FREE : CAM_INACTIVE);
    kind[c] = k;
    const int n = (k == CAM_FREE) ? cnt_local[c] : 0;
    free_cnt[c] = n;
    chunk_cnt[c] = (n + CHUNK - 1) / CHUNK;
}

// camera-major keys: free cameras keep their index, everything else sorts to the end (and is cut off by the count)
__global__ void prep_cam_keys_kernel(int No, int Nc, const int *__restrict__ pm_cam, const int *__restrict__ kind,
                                     unsigned int *__restrict__ key, int *__restrict__ idx) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= No) return;
    const int c = pm_cam[e];
    key[e] = (kind[c] == CAM_FREE) ? (unsigned)c : (unsigned)Nc;
    idx[e] = e;
}

__global__ void prep_gather_cm_kernel(int Ncm, const int *__restrict__ order, const float2 *__restrict__ pm_uv,
                                      const int *__restrict__ pm_pt, float2 *__restrict__ cm_uv, int *__restrict__ cm_pt,
                                      int *__restrict__ cm_obs) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= Ncm) return;
    const int e = order[d];
    cm_uv[d] = pm_uv[e];
    cm_pt[d] = pm_pt[e];
    cm_obs[d] = e;
}

// landmark blocks of the tile kernels: block b = the landmarks whose first observation index lies in
// [TILE_WIN b, TILE_WIN (b + 1)) -- whole landmarks, <= TILE_WIN + (longest track) - 1 observations, no prefix dependency
__global__ void prep_blocks_kernel(int n_blocks, int Np, const int *__restrict__ pt_ptr, int *__restrict__ first_pt) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > n_blocks) return;
    if (b == n_blocks) { first_pt[b] = Np; return; }
    const long long target = (long long)TILE_WIN * b;
    int lo = 0, hi = Np;                      // first p in [0, Np] with pt_ptr[p] >= target
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (pt_ptr[mid] >= target) hi = mid; else lo = mid + 1;
    }
    first_pt[b] = lo;
}
// + the block's record {first landmark, end landmark, first observation, observation count}: ONE 16-byte load at the top of a
// tile kernel instead of two dependent pairs of loads
__global__ void prep_blocks_max_kernel(int n_blocks, const int *__restrict__ pt_ptr, const int *__restrict__ first_pt, int *__restrict__ mx,
                                       int4 *__restrict__ info) {
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n_blocks) return;
    const int p0 = first_pt[b], p1 = first_pt[b + 1];
    const int o0 = pt_ptr[p0], n = pt_ptr[p1] - o0;
    info[b] = make_int4(p0, p1, o0, n);
    atomicMax(mx, n);
}

// position of every landmark-major observation inside its track | track length << 16: the tile kernels' per-landmark work
// is done by the thread of the track's FIRST observation (spread over all warps of the CTA) instead of a thread-per-landmark
// pass that kept two warps busy while six waited at the next barrier
__global__ void prep_trk_kernel(int No, const int *__restrict__ pm_pt, const int *__restrict__ pt_ptr, unsigned *__restrict__ pm_trk) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= No) return;
    const int p = pm_pt[e];
    const int lo = pt_ptr[p], len = pt_ptr[p + 1] - lo;
    pm_trk[e] = (unsigned)min(e - lo, 0xffff) | ((unsigned)min(len, 0xffff) << 16);
}

// camera index and camera kind of every landmark-major observation in ONE word (kind << 30 | camera): the per-observation
// kernels then need no dependent gather of cam_kind[camera] (a per-lane gather costs an L1 pass per instruction)
__global__ void prep_camk_kernel(int No, const int *__restrict__ pm_cam, const int *__restrict__ kind, unsigned *__restrict__ pm_camk) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= No) return;
    const int c = pm_cam[e];
    pm_camk[e] = (unsigned)c | ((unsigned)kind[c] << 30);
}

__global__ void prep_chunks_kernel(int Nc, const int *__restrict__ cam_ptr, const int *__restrict__ cam_chunk_ptr,
                                   int *__restrict__ chunk_cam, int *__restrict__ chunk_beg, int *__restrict__ chunk_end) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= Nc) return;
    int w = cam_chunk_ptr[c];
    for (int b = cam_ptr[c]; b < cam_ptr[c + 1]; b += CHUNK, ++w) {
        chunk_cam[w] = c;
        chunk_beg[w] = b;
        chunk_end[w] = min(b + CHUNK, cam_ptr[c + 1]);
    }
}

// ---- camera-pair index (dense Schur) ---------------------------------------------------------------------------
// Entries of landmark p: every ordered pair (a, b) of its free-camera observations with camera(a) >= camera(b).  The
// track is sorted by camera, so position order is camera order; two DIFFERENT observations by the same camera (a
// degenerate duplicate) contribute both (a, b) and (b, a) to their diagonal block.
template <typename F>
__device__ __forceinline__ void for_each_pair(int p, const int *__restrict__ pt_ptr, const int *__restrict__ pm_cam,
                                              const int *__restrict__ kind, F f) {
    const int o0 = pt_ptr[p], o1 = pt_ptr[p + 1];
    for (int a = o0; a < o1; ++a) {
        const int ca = pm_cam[a];
        if (kind[ca] != CAM_FREE) continue;
        for (int b = o0; b <= a; ++b) {
            const int cb = pm_cam[b];
            if (kind[cb] != CAM_FREE) continue;
            f(a, b, ca, cb);
            if (b != a && cb == ca) f(b, a, cb, ca);
        }
    }
}

__global__ void pair_count_kernel(int Np, const int *__restrict__ pt_ptr, const int *__restrict__ pm_cam,
                                  const int *__restrict__ kind, long long *__restrict__ npairs) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= Np) return;
    long long n = 0;
    for_each_pair(p, pt_ptr, pm_cam, kind, [&](int, int, int, int) { ++n; });
    npairs[p] = n;
}

__global__ void pair_fill_kernel(int Np, int Nc, const int *__restrict__ pt_ptr, const int *__restrict__ pm_cam,
                                 const int *__restrict__ kind, const long long *__restrict__ pair_off,
                                 unsigned long long *__restrict__ key, int *__restrict__ idx, int *__restrict__ ea, int *__restrict__ eb) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= Np) return;
    long long w = pair_off[p];
    for_each_pair(p, pt_ptr, pm_cam, kind, [&](int a, int b, int ca, int cb) {
        key[w] = (unsigned long long)ca * (unsigned long long)Nc + (unsigned long long)cb;
        idx[w] = (int)w;
        ea[w] = a;
        eb[w] = b;
        ++w;
    });
}

__global__ void pair_gather_kernel(long long P, const int *__restrict__ order, const int *__restrict__ ea, const int *__restrict__ eb,
                                   const unsigned long long *__restrict__ key, int2 *__restrict__ pair_obs, int *__restrict__ head) {
    const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= P) return;
    const int src = order[k];
    pair_obs[k] = make_int2(ea[src], eb[src]);
    head[k] = (k == 0 || key[k] != key[k - 1]) ? 1 : 0;
}

// segment s = run of equal (ci, cj); units = <= PAIR_UNIT entries of one segment
__global__ void pair_segments_kernel(long long P, int Nc, const int *__restrict__ head, const int *__restrict__ seg_id /*inclusive scan of head*/,
                                     const unsigned long long *__restrict__ key, long long *__restrict__ seg_beg, int2 *__restrict__ seg_cams) {
    const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= P) return;
    if (head[k]) {
        const int s = seg_id[k] - 1;
        seg_beg[s] = k;
        seg_cams[s] = make_int2((int)(key[k] / (unsigned long long)Nc), (int)(key[k] % (unsigned long long)Nc));
    }
}

__global__ void pair_unit_count_kernel(int n_seg, long long P, const long long *__restrict__ seg_beg, int *__restrict__ unit_cnt, int unit_len) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    const long long e = (s + 1 < n_seg) ? seg_beg[s + 1] : P;
    unit_cnt[s] = (int)((e - seg_beg[s] + unit_len - 1) / unit_len);
}

// unit record: first entry + (entry count << 1 | diagonal block), so that a unit starts with ONE round of loads
__global__ void pair_unit_fill_kernel(int n_seg, long long P, const long long *__restrict__ seg_beg, const int2 *__restrict__ seg_cams,
                                      const int *__restrict__ unit_ptr, int *__restrict__ unit_info, long long *__restrict__ unit_beg, int unit_len) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_seg) return;
    const long long e = (s + 1 < n_seg) ? seg_beg[s + 1] : P;
    const int2 c = seg_cams[s];
    const int diag = c.x == c.y;
    int u = unit_ptr[s];
    for (long long b = seg_beg[s]; b < e; b += unit_len, ++u) {
        const long long cnt = (e - b < unit_len) ? e - b : unit_len;
        unit_info[u] = ((int)cnt << 1) | diag;
        unit_beg[u] = b;
    }
}

// ---- host drivers ------------------------------------------------------------------------------------------------
#define PCU(expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) return (int)e__; } while (0)

// temp storage helper: grow-only
int PrepScratch::reserve(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return (int)e;
    cap = bytes;
    return 0;
}
void PrepScratch::release() { if (p) cudaFree(p); p = nullptr; cap = 0; }

static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

int prep_observations(const PrepIn &in, PrepOut &out, PrepScratch &scr, cudaStream_t s) {
    const int No = in.No, Nc = in.Nc, Np = in.Np;
    PCU(cudaMemsetAsync(out.cnt_cam, 0, sizeof(int) * (size_t)Nc, s));
    PCU(cudaMemsetAsync(out.pt_ptr, 0, sizeof(int) * (size_t)(Np + 1), s));
    PCU(cudaMemsetAsync(out.flags, 0, sizeof(int) * 4, s));
    if (No == 0) return 0;
    int bits = 1;
    while (bits < 64 && ((unsigned long long)Np * (unsigned long long)Nc) >> bits) ++bits;
    size_t cub_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (unsigned long long *)nullptr, (unsigned long long *)nullptr, (int *)nullptr,
                                    (int *)nullptr, No, 0, bits, s);
    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int *)nullptr, (int *)nullptr, Np + 1, s);
    cub_bytes = std::max(cub_bytes, scan_bytes);
    const size_t kb = align256(sizeof(unsigned long long) * (size_t)No), ib = align256(sizeof(int) * (size_t)No);
    const size_t cb = align256(sizeof(int) * (size_t)(Np + 1));
    int rc = scr.reserve(2 * kb + 2 * ib + cb + align256(cub_bytes));
    if (rc) return rc;
    char *base = (char *)scr.p;
    unsigned long long *key0 = (unsigned long long *)base, *key1 = (unsigned long long *)(base + kb);
    int *idx0 = (int *)(base + 2 * kb), *idx1 = (int *)(base + 2 * kb + ib);
    int *cnt_pt = (int *)(base + 2 * kb + 2 * ib);
    void *cub_tmp = base + 2 * kb + 2 * ib + cb;
    PCU(cudaMemsetAsync(cnt_pt, 0, sizeof(int) * (size_t)(Np + 1), s));
    prep_keys_kernel<<<cdivp(No, 256), 256, 0, s>>>(No, Nc, Np, in.obs_cam, in.obs_pt, key0, idx0, out.cnt_cam, cnt_pt, out.flags);
    PCU(cudaMemcpyAsync(out.h_flags, out.flags, sizeof(int) * 4, cudaMemcpyDeviceToHost, s));
    PCU(cudaStreamSynchronize(s));
    if (out.h_flags[0]) return -(out.h_flags[0]);          // negative: 1-based index of a bad observation
    const unsigned long long *key = key0;
    const int *order = idx0;
    if (out.h_flags[1]) {                                  // caller order is not landmark-major: sort
        PCU(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, key0, key1, idx0, idx1, No, 0, bits, s));
        key = key1; order = idx1;
    }
    prep_gather_pm_kernel<<<cdivp(No, 256), 256, 0, s>>>(No, Nc, key, order, in.obs_uv, out.pm_uv, out.pm_cam, out.pm_pt, out.perm);
    PCU(cub::DeviceScan::ExclusiveSum(cub_tmp, cub_bytes, cnt_pt, out.pt_ptr, Np + 1, s));
    return (int)cudaGetLastError();
}

int prep_blocks(const PrepIn &in, PrepOut &out, cudaStream_t s) {
    out.n_blocks = in.No / TILE_WIN + 1;
    PCU(cudaMemsetAsync(out.flags + 2, 0, sizeof(int), s));
    prep_blocks_kernel<<<cdivp(out.n_blocks + 1, 256), 256, 0, s>>>(out.n_blocks, in.Np, out.pt_ptr, out.blk_first_pt);
    if (in.No > 0) prep_trk_kernel<<<cdivp(in.No, 256), 256, 0, s>>>(in.No, out.pm_pt, out.pt_ptr, out.pm_trk);
    prep_blocks_max_kernel<<<cdivp(out.n_blocks, 256), 256, 0, s>>>(out.n_blocks, out.pt_ptr, out.blk_first_pt, out.flags + 2, out.blk_info);
    PCU(cudaMemcpyAsync(out.h_flags + 4, out.flags + 2, sizeof(int), cudaMemcpyDeviceToHost, s));
    PCU(cudaStreamSynchronize(s));
    out.max_block_obs = out.h_flags[4];
    return (int)cudaGetLastError();
}

int prep_camera_major(const PrepIn &in, PrepOut &out, PrepScratch &scr, cudaStream_t s) {
    const int No = in.No, Nc = in.Nc;
    int bits = 1;
    while ((Nc >> bits) != 0) ++bits;
    size_t cub_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (unsigned *)nullptr, (unsigned *)nullptr, (int *)nullptr, (int *)nullptr,
                                    std::max(No, 1), 0, bits, s);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (int *)nullptr, (int *)nullptr, Nc + 1, s);
    cub_bytes = std::max(cub_bytes, scan_bytes);
    const size_t kb = align256(sizeof(unsigned) * (size_t)std::max(No, 1)), cb = align256(sizeof(int) * (size_t)(Nc + 1));
    int rc = scr.reserve(4 * kb + 2 * cb + align256(cub_bytes));
    if (rc) return rc;
    char *base = (char *)scr.p;
    unsigned *key0 = (unsigned *)base, *key1 = (unsigned *)(base + kb);
    int *idx0 = (int *)(base + 2 * kb), *idx1 = (int *)(base + 3 * kb);
    int *free_cnt = (int *)(base + 4 * kb), *chunk_cnt = (int *)(base + 4 * kb + cb);
    void *cub_tmp = base + 4 * kb + 2 * cb;
    PCU(cudaMemsetAsync(free_cnt, 0, cb, s));
    PCU(cudaMemsetAsync(chunk_cnt, 0, cb, s));
    prep_kind_kernel<<<cdivp(Nc, 256), 256, 0, s>>>(Nc, in.cam_fixed, out.cnt_cam_global, out.cnt_cam, out.cam_kind, free_cnt, chunk_cnt);
    PCU(cub::DeviceScan::ExclusiveSum(cub_tmp, cub_bytes, free_cnt, out.cam_ptr, Nc + 1, s));
    PCU(cub::DeviceScan::ExclusiveSum(cub_tmp, cub_bytes, chunk_cnt, out.cam_chunk_ptr, Nc + 1, s));
    PCU(cudaMemcpyAsync(out.h_flags + 2, out.cam_ptr + Nc, sizeof(int), cudaMemcpyDeviceToHost, s));
    PCU(cudaMemcpyAsync(out.h_flags + 3, out.cam_chunk_ptr + Nc, sizeof(int), cudaMemcpyDeviceToHost, s));
    if (No > 0) {
        if (Nc >= (1 << 30)) return (int)cudaErrorInvalidValue;
        prep_camk_kernel<<<cdivp(No, 256), 256, 0, s>>>(No, out.pm_cam, out.cam_kind, out.pm_camk);
        prep_cam_keys_kernel<<<cdivp(No, 256), 256, 0, s>>>(No, Nc, out.pm_cam, out.cam_kind, key0, idx0);
        PCU(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, key0, key1, idx0, idx1, No, 0, bits, s));
    }
    PCU(cudaStreamSynchronize(s));
    out.Ncm = out.h_flags[2];
    out.n_chunks = out.h_flags[3];
    if (out.Ncm > 0) prep_gather_cm_kernel<<<cdivp(out.Ncm, 256), 256, 0, s>>>(out.Ncm, idx1, out.pm_uv, out.pm_pt, out.cm_uv, out.cm_pt, out.cm_obs);
    return (int)cudaGetLastError();
}

int prep_chunks(const PrepIn &in, PrepOut &out, cudaStream_t s) {
    if (out.n_chunks > 0)
        prep_chunks_kernel<<<cdivp(in.Nc, 256), 256, 0, s>>>(in.Nc, out.cam_ptr, out.cam_chunk_ptr, out.chunk_cam, out.chunk_beg, out.chunk_end);
    return (int)cudaGetLastError();
}

// Pass 1 (sizes): number of pairs.  Pass 2 (after the caller allocated PairIndex arrays): fill, sort, segment.
int prep_pairs_count(const PrepIn &in, const PrepOut &out, PrepScratch &scr, long long *h_total /*pinned*/, cudaStream_t s) {
    const int Np = in.Np;
    size_t scan_bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (long long *)nullptr, (long long *)nullptr, Np + 1, s);
    const size_t lb = align256(sizeof(long long) * (size_t)(Np + 1));
    int rc = scr.reserve(2 * lb + align256(scan_bytes));
    if (rc) return rc;
    long long *npairs = (long long *)scr.p, *off = (long long *)((char *)scr.p + lb);
    PCU(cudaMemsetAsync(npairs, 0, lb, s));
    if (Np > 0) pair_count_kernel<<<cdivp(Np, 256), 256, 0, s>>>(Np, out.pt_ptr, out.pm_cam, out.cam_kind, npairs);
    PCU(cub::DeviceScan::ExclusiveSum((char *)scr.p + 2 * lb, scan_bytes, npairs, off, Np + 1, s));
    PCU(cudaMemcpyAsync(h_total, off + Np, sizeof(long long), cudaMemcpyDeviceToHost, s));
    PCU(cudaStreamSynchronize(s));
    return 0;
}

int prep_pairs_build(const PrepIn &in, const PrepOut &out, PairIndex &pi, PrepScratch &scr, PrepScratch &scr2, int *h_two /*pinned [2]*/,
                     cudaStream_t s) {
    const int Np = in.Np, Nc = in.Nc;
    const long long P = pi.P;
    pi.n_seg = 0; pi.n_units = 0;
    if (P == 0) return 0;
    if (P > 0x7fffffffLL) return (int)cudaErrorInvalidValue;        // 2^31 pairs: far beyond what a dense S can hold anyway
    // scr still holds [npairs | off] from prep_pairs_count
    const size_t lb = align256(sizeof(long long) * (size_t)(Np + 1));
    const long long *off = (const long long *)((char *)scr.p + lb);
    int bits = 1;
    while (bits < 64 && ((unsigned long long)Nc * (unsigned long long)Nc) >> bits) ++bits;
    size_t cub_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (unsigned long long *)nullptr, (unsigned long long *)nullptr, (int *)nullptr,
                                    (int *)nullptr, (int)P, 0, bits, s);
    cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, (int *)nullptr, (int *)nullptr, (int)P, s);
    cub_bytes = std::max(cub_bytes, scan_bytes);
    const size_t kb = align256(sizeof(unsigned long long) * (size_t)P), ib = align256(sizeof(int) * (size_t)P);
    int rc = scr2.reserve(2 * kb + 6 * ib + align256(cub_bytes));
    if (rc) return rc;
    char *base = (char *)scr2.p;
    unsigned long long *key0 = (unsigned long long *)base, *key1 = (unsigned long long *)(base + kb);
    int *idx0 = (int *)(base + 2 * kb), *idx1 = idx0 + ib / 4, *ea = idx1 + ib / 4, *eb = ea + ib / 4, *head = eb + ib / 4, *segid = head + ib / 4;
    void *cub_tmp = base + 2 * kb + 6 * ib;
    pair_fill_kernel<<<cdivp(Np, 128), 128, 0, s>>>(Np, Nc, out.pt_ptr, out.pm_cam, out.cam_kind, off, key0, idx0, ea, eb);
    PCU(cub::DeviceRadixSort::SortPairs(cub_tmp, cub_bytes, key0, key1, idx0, idx1, (int)P, 0, bits, s));
    pair_gather_kernel<<<cdivp(P, 256), 256, 0, s>>>(P, idx1, ea, eb, key1, pi.pair_obs, head);
    PCU(cub::DeviceScan::InclusiveSum(cub_tmp, cub_bytes, head, segid, (int)P, s));
    PCU(cudaMemcpyAsync(h_two, segid + (P - 1), sizeof(int), cudaMemcpyDeviceToHost, s));
    PCU(cudaStreamSynchronize(s));
    pi.n_seg = h_two[0];
    if ((size_t)pi.n_seg > pi.seg_cap) return (int)cudaErrorMemoryAllocation;    // caller sized seg arrays for min(P, Nc(Nc+1)/2)
    pair_segments_kernel<<<cdivp(P, 256), 256, 0, s>>>(P, Nc, head, segid, key1, pi.seg_beg, pi.seg_cams);
    // units
    int *unit_cnt = idx0;                                 // reuse (n_seg + 1 <= P + 1 ints; idx0 has room for P)
    PCU(cudaMemsetAsync(unit_cnt, 0, sizeof(int) * (size_t)(pi.n_seg + 1), s));
    // Unit length: a unit is one warp's serial walk over its entries in batches of 32, so a problem with FEW entries per SM
    // (keyframe BA: 30 cameras, 200 k entries in 465 long segments) wants short units -- enough of them to fill the chip --
    // and a large one (K4: 5 M entries) the full PAIR_UNIT.  Fixed per problem, so the sums stay reproducible.
    {
        const long long per = (P + PAIR_TARGET_UNITS - 1) / PAIR_TARGET_UNITS;
        pi.unit_len = (int)std::min<long long>(PAIR_UNIT, std::max<long long>(32, (per + 31) / 32 * 32));
    }
    pair_unit_count_kernel<<<cdivp(pi.n_seg, 256), 256, 0, s>>>(pi.n_seg, P, pi.seg_beg, unit_cnt, pi.unit_len);
    size_t sb2 = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, sb2, (int *)nullptr, (int *)nullptr, pi.n_seg + 1, s);
    if (sb2 > cub_bytes) return (int)cudaErrorInvalidValue;
    PCU(cub::DeviceScan::ExclusiveSum(cub_tmp, cub_bytes, unit_cnt, pi.unit_ptr, pi.n_seg + 1, s));
    PCU(cudaMemcpyAsync(h_two + 1, pi.unit_ptr + pi.n_seg, sizeof(int), cudaMemcpyDeviceToHost, s));
    PCU(cudaStreamSynchronize(s));
    pi.n_units = h_two[1];
    return 0;
}

// camera groups (= Cholesky tiles) coupled through a shared landmark
__global__ void group_map_kernel(int n_seg, const int2 *__restrict__ seg_cams, int G, int cams_per_tile, unsigned char *__restrict__ map) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_seg) return;
    const int2 c = seg_cams[idx];
    const int gi = c.x / cams_per_tile, gj = c.y / cams_per_tile;
    map[(size_t)max(gi, gj) * G + min(gi, gj)] = 1;
}

int prep_group_map(const PairIndex &pi, int G, int cams_per_tile, unsigned char *map, cudaStream_t s) {
    cudaError_t e = cudaMemsetAsync(map, 0, (size_t)G * G, s);
    if (e != cudaSuccess) return (int)e;
    if (pi.n_seg > 0) group_map_kernel<<<cdivp(pi.n_seg, 256), 256, 0, s>>>(pi.n_seg, pi.seg_cams, G, cams_per_tile, map);
    return (int)cudaGetLastError();
}

// Camera-sorted order of every landmark block's observations (PCG matvec: contributions to q are folded per camera in THIS
// fixed order, one writer per (block, camera) -- no floating-point atomics, bit-reproducible).  One CTA per block: bitonic sort
// of (camera << 11 | local index) in shared memory.  Output word: camera << 12 | (first of its camera in the block) << 11 |
// local index.  Needs cameras < 2^20 and blocks < 2048 observations (checked by the caller).
__global__ void __launch_bounds__(256) block_order_kernel(const int4 *__restrict__ blk_info, const int *__restrict__ pm_cam, unsigned *__restrict__ blk_ord, int n2) {
    extern __shared__ unsigned keys[];
    const int4 bi = blk_info[blockIdx.x];
    const int o0 = bi.z, n = bi.w;
    if (n <= 0) return;
    int m = 1;
    while (m < n) m <<= 1;                                   // this block's power of two (<= n2)
    for (int i = threadIdx.x; i < m; i += blockDim.x) keys[i] = i < n ? ((unsigned)pm_cam[o0 + i] << 11) | (unsigned)i : 0xffffffffu;
    __syncthreads();
    for (int k = 2; k <= m; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < m; i += blockDim.x) {
                const int l = i ^ j;
                if (l > i) {
                    const unsigned a = keys[i], b = keys[l];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { keys[i] = b; keys[l] = a; }
                }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned key = keys[i], cam = key >> 11, idx = key & 2047u;
        const bool start = i == 0 || (keys[i - 1] >> 11) != cam;
        blk_ord[o0 + i] = (cam << 12) | (start ? 2048u : 0u) | idx;
    }
    (void)n2;
}

int prep_block_order(const PrepOut &out, int Nc, unsigned *blk_ord, cudaStream_t s) {
    if (out.n_blocks <= 0 || out.max_block_obs <= 0) return 0;
    if (Nc >= (1 << 20) || out.max_block_obs >= 2048) return (int)cudaErrorInvalidValue;
    int n2 = 1;
    while (n2 < out.max_block_obs) n2 <<= 1;
    block_order_kernel<<<out.n_blocks, 256, sizeof(unsigned) * (size_t)n2, s>>>(out.blk_info, out.pm_cam, blk_ord, n2);
    return (int)cudaGetLastError();
}

int prep_pairs_units(PairIndex &pi, cudaStream_t s) {
    if (pi.n_seg > 0)
        pair_unit_fill_kernel<<<cdivp(pi.n_seg, 256), 256, 0, s>>>(pi.n_seg, pi.P, pi.seg_beg, pi.seg_cams, pi.unit_ptr, pi.unit_seg, pi.unit_beg, pi.unit_len);
    return (int)cudaGetLastError();
}

} // namespace ba
