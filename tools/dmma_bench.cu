// dmma_bench.cu -- micro-benchmarks behind the Cholesky design (run on the GPU box):
//   1. register-only FP64 tensor-pipe peak for the mma.sync f64 shapes sm_100a accepts, and the DFMA peak
//   2. the product's gemm_nt_kernel (included from ba_cholesky.cu) on large and on K = 128 problems
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../bundleadjustment_b200/csrc -o _build/dmma_bench dmma_bench.cu
#include "../bundleadjustment_b200/csrc/ba_cholesky.cu"

#include <cstdio>
#include <vector>

using namespace ba;

template <int SHAPE>
__global__ void __launch_bounds__(512) mma_peak_kernel(double *out, int iters) {
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
void run_peak(const char *name, double flops_per_inst_per_warp, double *out, int blocks = 148 * 2, int threads = 256) {
    const int iters = 4000;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mma_peak_kernel<SHAPE><<<blocks, threads>>>(out, 10);
    cudaEventRecord(e0);
    mma_peak_kernel<SHAPE><<<blocks, threads>>>(out, iters);
    cudaEventRecord(e1);
    cudaError_t e = cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl = (double)blocks * (threads / 32) * iters * 16 * flops_per_inst_per_warp;
    printf("%-14s %8.3f ms  %7.2f TFLOP/s  (%s)\n", name, ms, fl / ms / 1e9, cudaGetErrorString(e));
}

int main() {
    double *out;
    cudaMalloc(&out, sizeof(double) * 148 * 2 * 512);
    run_peak<0>("dmma m8n8k4", 2.0 * 8 * 8 * 4, out);
    run_peak<1>("dmma m16n8k4", 2.0 * 16 * 8 * 4, out);
    run_peak<2>("dmma m16n8k8", 2.0 * 16 * 8 * 8, out);
    run_peak<3>("dmma m16n8k16", 2.0 * 16 * 8 * 16, out);
    run_peak<4>("dfma", 2.0 * 4 * 32, out);
    run_peak<0>("m8n8k4 1w/sp", 2.0 * 8 * 8 * 4, out, 148, 128);
    run_peak<0>("m8n8k4 2w/sp", 2.0 * 8 * 8 * 4, out, 148, 256);
    run_peak<0>("m8n8k4 3w/sp", 2.0 * 8 * 8 * 4, out, 148, 384);
    run_peak<0>("m8n8k4 4w/sp", 2.0 * 8 * 8 * 4, out, 148, 512);
    run_peak<3>("m16n8k16 1w/sp", 2.0 * 16 * 8 * 16, out, 148, 128);
    run_peak<3>("m16n8k16 2w/sp", 2.0 * 16 * 8 * 16, out, 148, 256);
    return 0;
}
