// Micro-benchmark: dependent-issue latency and per-warp throughput of the FP64 tensor instructions (DMMA) on sm_100a.
// One CTA; cycles from clock64 inside the kernel.  nvcc -arch=sm_100a dmma_latency.cu -o dmma_latency
#include <cstdio>
#include <cuda_runtime.h>

template <int SHAPE, int NCH>
__global__ void k(double* out, long long* cyc, int iters, double a0, double b0) {
    double a[8], b[4];
    for (int i = 0; i < 8; ++i) a[i] = a0 + i;
    for (int i = 0; i < 4; ++i) b[i] = b0 + i;
    double c[NCH][4];
    for (int i = 0; i < NCH; ++i) for (int j = 0; j < 4; ++j) c[i][j] = threadIdx.x + i + j;
    __syncthreads();
    long long t0 = clock64();
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
            else
                asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, {%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                             : "+d"(c[i][0]), "+d"(c[i][1]), "+d"(c[i][2]), "+d"(c[i][3])
                             : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                               "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
        }
    }
    long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < NCH; ++i) for (int j = 0; j < 4; ++j) s += c[i][j];
    if (s == 1.2345) out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int SHAPE, int NCH>
void run(const char* name, int fma, int warps) {
    double* out; long long* cyc; cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
    int iters = 4000;
    long long h = 0;
    for (int r = 0; r < 2; ++r) { k<SHAPE, NCH><<<1, warps * 32>>>(out, cyc, iters, 1e-9, 1e-9); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    double per_round = (double)h / iters;
    printf("%-9s warps=%2d chains/warp=%d : %.1f cycles per round of %d DMMA/warp -> %.1f cycles/DMMA/warp, %.1f FMA/clk/SM %s\n",
           name, warps, NCH, per_round, NCH, per_round / NCH, (double)fma * NCH * warps / per_round, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0, 1>("m8n8k4", 256, 1); run<0, 2>("m8n8k4", 256, 1); run<0, 4>("m8n8k4", 256, 1); run<0, 8>("m8n8k4", 256, 1);
    run<0, 1>("m8n8k4", 256, 4); run<0, 2>("m8n8k4", 256, 4); run<0, 4>("m8n8k4", 256, 4); run<0, 8>("m8n8k4", 256, 4);
    run<0, 2>("m8n8k4", 256, 8); run<0, 4>("m8n8k4", 256, 8); run<0, 4>("m8n8k4", 256, 16);
    run<1, 1>("m16n8k4", 512, 1); run<1, 4>("m16n8k4", 512, 1); run<1, 4>("m16n8k4", 512, 4); run<1, 4>("m16n8k4", 512, 8);
    run<2, 1>("m16n8k8", 1024, 1); run<2, 4>("m16n8k8", 1024, 1); run<2, 4>("m16n8k8", 1024, 4); run<2, 4>("m16n8k8", 1024, 8);
    run<3, 1>("m16n8k16", 2048, 1); run<3, 2>("m16n8k16", 2048, 1); run<3, 4>("m16n8k16", 2048, 1); run<3, 2>("m16n8k16", 2048, 4); run<3, 4>("m16n8k16", 2048, 4); run<3, 4>("m16n8k16", 2048, 8);
    return 0;
}
