// Micro-benchmark: FP64 tensor (DMMA) issue rate per shape vs vector DFMA, per SM.  nvcc -arch=sm_100a dmma_rate.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int SHAPE, int NCH>
__global__ void k(double* out, int iters, double a0, double b0) {
    double a[4] = {a0, a0 + 1, a0 + 2, a0 + 3}, b[2] = {b0, b0 + 1};
    double c[NCH][4];
    for (int i = 0; i < NCH; ++i) for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x + i + j;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
            if (SHAPE == 0)
                asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                             : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a[0]), "d"(b[0]));
            else if (SHAPE == 1)
                asm volatile("mma.sync.aligned.m16n8k4.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%0,%1,%2,%3};"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3]) : "d"(a[0]), "d"(a[1]), "d"(b[0]));
            else if (SHAPE == 2)
                asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
            else {
                c[i][0] = fma(c[i][0], a[0], b[0]); c[i][1] = fma(c[i][1], a[0], b[0]);
                c[i][2] = fma(c[i][2], a[0], b[0]); c[i][3] = fma(c[i][3], a[0], b[0]);
            }
        }
    }
    double s = 0;
    for (int i = 0; i < NCH; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    if (s == 1.2345) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int SHAPE, int NCH>
void run(const char* name, double fma_per_instr, int warps) {
    double* out; cudaMalloc(&out, 148 * 1024 * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int iters = 20000;
    float best = 1e9;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0);
        k<SHAPE, NCH><<<148, warps * 32>>>(out, iters, 1e-9, 1e-9);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
    }
    double instr = (double)iters * NCH * warps;   // per SM
    double cyc = best * 1e-3 * 1.965e9;
    printf("%-10s warps=%2d chains=%d  %.2f cycles/warp-instr/SM  -> %.1f FMA/clk/SM  (%.1f TFLOP/s)  err=%s\n", name, warps, NCH,
           cyc / instr, fma_per_instr * instr / cyc, 2 * fma_per_instr * instr * 148 / (best * 1e-3) / 1e12, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out);
}

int main() {
    run<0, 4>("m8n8k4", 256, 4); run<0, 4>("m8n8k4", 256, 8); run<0, 4>("m8n8k4", 256, 16);
    run<1, 4>("m16n8k4", 512, 4); run<1, 4>("m16n8k4", 512, 8); run<1, 4>("m16n8k4", 512, 16);
    run<2, 4>("m16n8k8", 1024, 4); run<2, 4>("m16n8k8", 1024, 8); run<2, 4>("m16n8k8", 1024, 16);
    run<3, 4>("dfma x4", 128, 8); run<3, 4>("dfma x4", 128, 16); run<3, 4>("dfma x4", 128, 32);
    run<0, 1>("m8n8k4 dep", 256, 4);  run<1, 1>("m16n8k4 dep", 512, 4); run<3, 1>("dfma dep", 128, 4);
    return 0;
}
