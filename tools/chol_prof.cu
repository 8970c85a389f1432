// chol_prof.cu -- per-CTA timing breakdown of the dataflow Cholesky (tools only; includes the product source).
// build: see tools/Makefile-less one-liner in profiles/README; run: tools/_build/chol_prof [n]
#define CHOL_PROFILE_PHASES 1
#include "../bundleadjustment_b200/csrc/ba_cholesky.cu"

#include <algorithm>
#include <cstdio>
#include <vector>

using namespace ba;

// one CTA per tile slot (i, j) of the plan's tile-packed storage: banded pseudo-random SPD-dominant entries, rhs row = 1
__global__ void fill_spd(double *A, const long long *toff, int T, int n, int band, int nb) {
    const int i = blockIdx.x / T, j = blockIdx.x % T;
    const long long o = toff[(size_t)i * T + j];
    if (o < 0) return;
    for (int idx = threadIdx.x; idx < nb * nb; idx += blockDim.x) {
        const int r = i * nb + idx / nb, c = j * nb + idx % nb;
        double v = 0.0;
        if (i == T) v = (idx / nb == 0 && c < n) ? 1.0 : 0.0;
        else if (r < n && c <= r && r - c <= band) {
            unsigned h = (unsigned)(r * 2654435761u) ^ (unsigned)(c * 40503u + 12345u);
            h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
            v = ((h & 0xffff) / 65536.0 - 0.5) * 0.01;
            if (r == c) v = 1.0 + 0.04 * sqrt((double)n);
        } else if (r == c) v = 1.0;
        A[o + idx] = v;
    }
}

int main(int argc, char **argv) {
    const int n = argc > 1 ? atoi(argv[1]) : 10200;
    const int band = argc > 2 ? atoi(argv[2]) : n;
    const int nb = argc > 3 ? atoi(argv[3]) : 128;
    const int ld = chol_ld(n, nb);
    double *A, *work, *y;
    cudaMalloc(&work, 8 * chol_work_doubles(n, nb)); cudaMemset(work, 0, 8 * chol_work_doubles(n, nb)); cudaMalloc(&y, 8 * (size_t)ld);
    cudaMalloc(&g_chol_prof, 8 * 8 * 512); cudaMemset(g_chol_prof, 0, 8 * 8 * 512);
    if (chol_init()) { printf("chol_init failed\n"); return 1; }
    // tile plan of the band (exact tile pattern of fill_spd)
    const int Tp = ld / nb;
    std::vector<unsigned char> nz((size_t)Tp * Tp, 0);
    for (int i = 0; i < Tp; ++i)
        for (int j = 0; j <= i; ++j) nz[(size_t)i * Tp + j] = (i * nb - (j * nb + nb - 1)) <= band;
    CholPlan plan;
    if (chol_plan_create(n, nb, band < n ? nz.data() : nullptr, &plan, 0)) { printf("plan failed\n"); return 1; }
    const CholPlan *plan_ptr = &plan;
    if (band < n) printf("band %d: %lld filled tiles, %d tasks, %.3f GFLOP\n", band, plan.filled_tiles, plan.ntasks, plan.flops * 1e-9);
    cudaMalloc(&A, 8 * (size_t)plan.total_doubles);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        fill_spd<<<(Tp + 1) * Tp, 256>>>(A, plan.tile_off, Tp, n, band, nb);
        cudaEventRecord(e0);
        int ce = chol_factor_solve(A, n, work, y, nullptr, nullptr, plan_ptr, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (ce || e) { printf("error %d %s\n", ce, cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        best = std::min(best, ms);
    }
    std::vector<long long> h(8 * 512);
    cudaMemcpy(h.data(), g_chol_prof, 8 * 8 * 512, cudaMemcpyDeviceToHost);
    int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int T = ld / nb, grid = std::min(sms * (nb == 64 ? 2 : 1), T * (T + 3) / 2);
    long long t0 = h[0], tend = 0;
    for (int b = 0; b < grid; ++b) { t0 = std::min(t0, h[8 * b]); tend = std::max(tend, h[8 * b + 2]); }
    double s_p1 = 0, s_w1 = 0, s_w2 = 0, s_potrf = 0, s_trsm = 0, max_p1 = 0, s_tasks = 0, potrf_tiles = T;
    for (int b = 0; b < grid; ++b) {
        const long long *o = &h[8 * b];
        s_p1 += (o[1] - o[0]); max_p1 = std::max(max_p1, (double)(o[1] - t0)); s_w1 += o[3]; s_w2 += o[4]; s_potrf += o[5]; s_trsm += o[7]; s_tasks += o[6];
    }
    printf("n=%d T=%d grid=%d: best %.3f ms  %.2f TFLOP/s\n", n, T, grid, best, ((double)n * n * n / 3 + 2.0 * n * n) / best / 1e9);
    printf("  kernel span %.3f ms; phase 1 ends (max over CTAs) at %.3f ms; mean phase-1 busy %.3f ms of which waiting %.3f ms, trsm epilogues %.3f ms\n",
           (tend - t0) * 1e-6, max_p1 * 1e-6, s_p1 / grid * 1e-6, s_w1 / grid * 1e-6, s_trsm / grid * 1e-6);
    printf("  potrf: %.1f us per diagonal tile (%d tiles); phase 2 (back substitution) %.3f ms, mean wait there %.3f ms; tasks %.0f\n",
           s_potrf / potrf_tiles * 1e-3, T, (tend - t0 - max_p1) * 1e-6, s_w2 / grid * 1e-6, s_tasks);
    long long ph[12];
    cudaMemcpyFromSymbol(ph, g_potrf_phase, sizeof(ph));
    if (ph[3]) {
        printf("  potrf phases per tile: (A) %.1f us  (B) %.1f us  (C) %.1f us  over %lld calls\n", ph[0] * 1e-3 / ph[3], ph[1] * 1e-3 / ph[3], ph[2] * 1e-3 / ph[3], ph[3]);
        const double panels = (double)ph[3] * (nb / 32);
        printf("  inside (A), SM clocks per 32 x 32 panel: load %.0f  factor sweep (warp 0; warp 1 inverts one pivot behind) %.0f  epilogue (all warps) %.0f\n",
               ph[4] / panels, ph[5] / panels, ph[7] / panels);
    }
    return 0;
}
