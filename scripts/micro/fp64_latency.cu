// Micro-benchmark: dependent-issue latency and per-warp issue interval of the FP64 vector instructions (DMUL, DADD, DFMA)
// on sm_100a, alone and right after a DMMA whose result they consume (the epilogue of the chain kernel), one warp per SM
// sub-partition as in k_chain_mma2.  nvcc -gencode arch=compute_100a,code=sm_100a fp64_latency.cu -o fp64_latency
#include <cstdio>
#include <cuda_runtime.h>

// MODE 0: one dependent chain of DMUL; 1: DADD; 2: DFMA; NCH independent chains per warp
template <int MODE, int NCH>
__global__ void k(double* out, long long* cyc, int iters, double a0) {
    double c[NCH];
    for (int i = 0; i < NCH; ++i) c[i] = 1.0 + 1e-9 * (threadIdx.x + i);
    const double m = 1.0 + a0, ad = a0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < NCH; ++i) {
                if (MODE == 0) asm volatile("mul.f64 %0, %0, %1;" : "+d"(c[i]) : "d"(m));
                else if (MODE == 1) asm volatile("add.f64 %0, %0, %1;" : "+d"(c[i]) : "d"(ad));
                else asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(c[i]) : "d"(m), "d"(ad));
            }
    }
    long long t1 = clock64();
    double s = 0;
    for (int i = 0; i < NCH; ++i) s += c[i];
    if (s == 1.2345) out[threadIdx.x] = s;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

// a DMMA, then DADD -> DMUL -> DMUL on its result, then the next DMMA on the product (the chain kernel's serial tail)
__global__ void k_tail(double* out, long long* cyc, int iters, double a0, int ndep) {
    double a = 1e-9 + a0, b = 1e-9;
    double c0 = threadIdx.x, c1 = 1.0;
    __syncthreads();
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
        if (ndep >= 1) asm volatile("add.f64 %0, %0, %1;" : "+d"(c0) : "d"(c1));
        if (ndep >= 2) asm volatile("mul.f64 %0, %0, %1;" : "+d"(c0) : "d"(a));
        if (ndep >= 3) asm volatile("mul.f64 %0, %0, %1;" : "+d"(c0) : "d"(b));
    }
    long long t1 = clock64();
    if (c0 + c1 == 1.2345) out[threadIdx.x] = c0;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE, int NCH>
void run(const char* name, int warps) {
    double* out; long long* cyc; cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
    const int iters = 2000;
    long long h = 0;
    for (int r = 0; r < 2; ++r) { k<MODE, NCH><<<1, warps * 32>>>(out, cyc, iters, 1e-12); cudaDeviceSynchronize(); }
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-5s warps=%2d chains/warp=%d : %.2f cycles per instruction per chain (%.2f per warp instruction) %s\n", name, warps, NCH,
           (double)h / iters / 8, (double)h / iters / 8 / NCH, cudaGetErrorString(cudaGetLastError()));
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0, 1>("DMUL", 1); run<1, 1>("DADD", 1); run<2, 1>("DFMA", 1);
    run<0, 4>("DMUL", 1); run<1, 4>("DADD", 1); run<2, 4>("DFMA", 1);
    run<0, 8>("DMUL", 1); run<2, 8>("DFMA", 1); run<2, 8>("DFMA", 4); run<2, 8>("DFMA", 8);
    double* out; long long* cyc; cudaMalloc(&out, 1024 * 8); cudaMalloc(&cyc, 8);
    for (int nd = 0; nd <= 3; ++nd) {
        long long h = 0;
        for (int r = 0; r < 2; ++r) { k_tail<<<1, 32>>>(out, cyc, 2000, 1e-12, nd); cudaDeviceSynchronize(); }
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
        printf("DMMA + %d dependent FP64 ops, back to back: %.1f cycles per round\n", nd, (double)h / 2000);
    }
    return 0;
}
