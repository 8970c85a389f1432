// dmma_bench.cu -- micro-benchmarks behind the Cholesky design (run on the GPU box):
//   1. register-only FP64 tensor-pipe peak for the mma.sync f64 shapes sm_100a accepts, and the DFMA peak
//   2. the product's gemm_nt_kernel (included from ba_cholesky.cu) on large and on K = 128 problems
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../bundleadjustment_b200/csrc -o _build/dmma_bench dmma_bench.cu
#include "../bundleadjustment_b200/csrc/ba_cholesky.cu"

#include <cstdio>
#include <vector>

using namespace ba;

template <int SHAPE>
__global__ void __launch_bounds__(256) mma_peak_kernel(double *out, int iters) {
    double c[16][4];
    for (int i = 0; i < 16; ++i) for (int j = 0; j < 4; ++j) c[i][j] = 0.0;
    double a[8], b[4];
    for (int i = 0; i < 8; ++i) a[i] = 1.0 + threadIdx.x * 1e-9 + i;
    for (int i = 0; i < 4; ++i) b[i] = 1.0 - threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (SHAPE == 0) {
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a[0]), "d"(b[0]));
            } else if (SHAPE == 1) {
                asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};\n"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
            } else if (SHAPE == 2) {
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
            } else if (SHAPE == 3) {
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};\n"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                               "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) c[i][j] = fma(a[j], b[j], c[i][j]);
            }
        }
    }
    double s = 0;
    for (int i = 0; i < 16; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int SHAPE>
void run_peak(const char *name, double flops_per_inst_per_warp, double *out) {
    const int blocks = 148 * 2, iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mma_peak_kernel<SHAPE><<<blocks, 256>>>(out, 10);
    cudaEventRecord(e0);
    mma_peak_kernel<SHAPE><<<blocks, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = (double)blocks * 8 * iters * 16 * flops_per_inst_per_warp;
    printf("%-14s %8.3f ms  %7.2f TFLOP/s  (%s)\n", name, ms, fl / ms / 1e9, cudaGetErrorString(e));
}

void run_gemm(int M, int N, int K, bool tri) {
    double *A, *C;
    cudaMalloc(&A, sizeof(double) * (size_t)M * K);
    cudaMalloc(&C, sizeof(double) * (size_t)M * N);
    cudaMemset(A, 0, sizeof(double) * (size_t)M * K);
    cudaMemset(C, 0, sizeof(double) * (size_t)M * N);
    chol_init();
    const int tm = (M + BM - 1) / BM, tn = (N + BN - 1) / BN;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0);
        if (tri) gemm_nt_kernel<MODE_SUB, true><<<tm * (tm + 1) / 2, 256, GEMM_SMEM>>>(A, K, A, K, C, N, M, N, K, 0, nullptr);
        else gemm_nt_kernel<MODE_SUB, false><<<tm * tn, 256, GEMM_SMEM>>>(A, K, A, K, C, N, M, N, K, tn, nullptr);
        cudaEventRecord(e1);
        cudaDeviceSynchronize();
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r) best = fminf(best, ms);
    }
    const double tiles = tri ? tm * (tm + 1) / 2.0 : (double)tm * tn;
    const double fl = tiles * 2.0 * BM * BN * K;
    printf("gemm_nt %s M=%d N=%d K=%d: %8.3f ms  %6.2f TFLOP/s  (%s)\n", tri ? "tri " : "full", M, N, K, best, fl / best / 1e9,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(A); cudaFree(C);
}

int main() {
    double *out;
    cudaMalloc(&out, sizeof(double) * 148 * 2 * 256);
    run_peak<0>("dmma m8n8k4", 2.0 * 8 * 8 * 4, out);
    run_peak<1>("dmma m16n8k4", 2.0 * 16 * 8 * 4, out);
    run_peak<2>("dmma m16n8k8", 2.0 * 16 * 8 * 8, out);
    run_peak<3>("dmma m16n8k16", 2.0 * 16 * 8 * 16, out);
    run_peak<4>("dfma", 2.0 * 4 * 32, out);
    cudaFuncSetAttribute(gemm_nt_kernel<MODE_SUB, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM);
    run_gemm(8192, 8192, 128, false);
    run_gemm(8192, 8192, 1024, false);
    run_gemm(8192, 8192, 8192, false);
    run_gemm(10112, 10112, 128, true);
    run_gemm(5120, 5120, 128, true);
    run_gemm(2560, 2560, 128, true);
    return 0;
}
