// chol_chain.cu -- critical path of the tile-sparse dataflow Cholesky on a banded tile pattern in nested-dissection
// order (tools only; includes the product source).  usage: chol_chain [G tiles = 81] [band in tiles = 2] [natural = 0] [tile edge = 128]
#include "../bundleadjustment_b200/csrc/ba_cholesky.cu"

#include <algorithm>
#include <cstdio>
#include <map>
#include <numeric>
#include <vector>

using namespace ba;

// one CTA per tile slot (i, j) of the plan's tile-packed storage: pseudo-random SPD-dominant entries, rhs row = 1
__global__ void fill_pattern(double *A, const long long *toff, int T, int ld, int nb) {
    const int i = blockIdx.x / T, j = blockIdx.x % T;
    const long long o = toff[(size_t)i * T + j];
    if (o < 0) return;
    for (int idx = threadIdx.x; idx < nb * nb; idx += blockDim.x) {
        const int r = i * nb + idx / nb, c = j * nb + idx % nb;
        double v = 0.0;
        if (i == T) v = (idx / nb == 0) ? 1.0 : 0.0;
        else if (c <= r) {
            unsigned h = (unsigned)(r * 2654435761u) ^ (unsigned)(c * 40503u + 12345u);
            h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
            v = ((h & 0xffff) / 65536.0 - 0.5) * 0.01;
            if (r == c) v = 1.0 + 0.04 * sqrt((double)ld);
        }
        A[o + idx] = v;
    }
}

int main(int argc, char **argv) {
    const int G = argc > 1 ? atoi(argv[1]) : 81, band = argc > 2 ? atoi(argv[2]) : 2, natural = argc > 3 ? atoi(argv[3]) : 0;
    const int nb = argc > 4 ? atoi(argv[4]) : 128;
    const int n = G * nb, ld = n, T = G;
    std::vector<unsigned char> adj((size_t)G * G, 0), nz((size_t)G * G, 0);
    for (int i = 0; i < G; ++i) for (int j = std::max(0, i - band); j <= i; ++j) adj[(size_t)i * G + j] = 1;
    std::vector<int> order(G), pos(G);
    if (natural) std::iota(order.begin(), order.end(), 0); else chol_tile_order(G, adj.data(), order);
    for (int p = 0; p < G; ++p) pos[order[p]] = p;
    for (int i = 0; i < G; ++i) for (int j = 0; j <= i; ++j) if (adj[(size_t)i * G + j]) nz[(size_t)std::max(pos[i], pos[j]) * G + std::min(pos[i], pos[j])] = 1;
    double *A, *work, *y;
    if (chol_init()) { printf("chol_init failed\n"); return 1; }
    CholPlan plan;
    if (chol_plan_create(n, nb, nz.data(), &plan, 0)) { printf("plan failed\n"); return 1; }
    cudaMalloc(&A, 8 * (size_t)plan.total_doubles); cudaMalloc(&work, 8 * chol_work_doubles(n, nb)); cudaMemset(work, 0, 8 * chol_work_doubles(n, nb)); cudaMalloc(&y, 8 * (size_t)ld);
    printf("tile edge %d: G=%d band=%d %s order: %lld filled tiles, %d tasks, depth %d, %.3f GFLOP\n", nb, G, band, natural ? "natural" : "nested-dissection",
           plan.filled_tiles, plan.ntasks, plan.depth, plan.flops * 1e-9);
    const size_t nprof = 4 * (size_t)plan.ntasks + T;
    cudaMalloc(&g_chol_task_prof, 8 * nprof);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        fill_pattern<<<(T + 1) * T, 256>>>(A, plan.tile_off, T, ld, nb);
        cudaMemset(g_chol_task_prof, 0, 8 * nprof);
        cudaEventRecord(e0);
        int ce = chol_factor_solve(A, n, work, y, nullptr, nullptr, &plan, 0);
        cudaEventRecord(e1);
        cudaError_t e = cudaDeviceSynchronize();
        if (ce || e) { printf("error %d %s\n", ce, cudaGetErrorString(e)); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        best = std::min(best, ms);
    }
    printf("best %.3f ms (with task timestamps on)\n", best);
    std::vector<long long> tp(nprof);
    cudaMemcpy(tp.data(), g_chol_task_prof, 8 * nprof, cudaMemcpyDeviceToHost);
    std::vector<int4> task(plan.ntasks + 1);
    cudaMemcpy(task.data(), plan.task, sizeof(int4) * task.size(), cudaMemcpyDeviceToHost);
    std::vector<int> klist(std::max(task[plan.ntasks].w, 1));
    cudaMemcpy(klist.data(), plan.klist, 4 * (size_t)task[plan.ntasks].w, cudaMemcpyDeviceToHost);
    long long t0 = tp[0];
    for (int t = 0; t < plan.ntasks; ++t) t0 = std::min(t0, tp[4 * (size_t)t]);
    std::map<std::pair<int, int>, int> at;
    for (int t = 0; t < plan.ntasks; ++t) at[{task[t].x, task[t].y}] = t;
    int cur = 0;
    for (int t = 0; t < plan.ntasks; ++t) if (tp[4 * (size_t)t + 3] > tp[4 * (size_t)cur + 3]) cur = t;
    printf("phase 1 ends at %.1f us; critical path backwards (times in us from kernel start):\n", (tp[4 * (size_t)cur + 3] - t0) * 1e-3);
    printf("   task (i,j)   lvl  nkb  cta | picked  consumed  diag/potrf  published | task time  since dep\n");
    while (true) {
        const int4 tk = task[cur];
        const long long *q = &tp[4 * (size_t)cur];
        int dep = -1; long long dep_end = 0;
        auto consider = [&](int a, int b) { auto it = at.find({a, b}); if (it == at.end()) return; const long long e = tp[4 * (size_t)it->second + 3]; if (e > dep_end) { dep_end = e; dep = it->second; } };
        for (int k = tk.w; k < task[cur + 1].w; ++k) { consider(tk.x, klist[k]); consider(tk.y, klist[k]); }
        if (tk.x != tk.y) consider(tk.y, tk.y);
        printf("   (%3d,%3d)  %4d %4d %4d | %7.1f  %7.1f  %7.1f  %7.1f | %7.1f  %7.1f\n", tk.x, tk.y, tk.z, task[cur + 1].w - tk.w, cur % 148,
               (q[0] - t0) * 1e-3, (q[1] - t0) * 1e-3, (q[2] - t0) * 1e-3, (q[3] - t0) * 1e-3, (q[3] - q[0]) * 1e-3, dep >= 0 ? (q[3] - dep_end) * 1e-3 : 0.0);
        if (dep < 0) break;
        cur = dep;
    }
    long long y_end = 0;
    for (int m = 0; m < T; ++m) y_end = std::max(y_end, tp[4 * (size_t)plan.ntasks + m]);
    printf("back substitution ends at %.1f us\n", (y_end - t0) * 1e-3);
    return 0;
}
