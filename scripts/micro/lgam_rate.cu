// Throughput of CUDA lgamma vs lgamma_pos (lgam.h) on B200; args: small (0.05..10) and large (10..5000).
#include <cstdio>
#include <cuda_runtime.h>
#include "lgam.h"
template <int WHICH>
__global__ void k(const double* __restrict__ x, double* __restrict__ y, int n, int reps) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double z = x[i], acc = 0.;
    for (int r = 0; r < reps; ++r) {
        double a = z + 1e-3 * r;
        acc += WHICH == 0 ? lgamma(a) : (WHICH == 1 ? lgamma_pos(a) : log(a));
    }
    y[i] = acc;
}
int main() {
    const int n = 148 * 2048 * 4, reps = 64;
    double *x, *y, *hx = new double[n];
    cudaMalloc(&x, n * 8); cudaMalloc(&y, n * 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int mode = 0; mode < 2; ++mode) {
        for (int i = 0; i < n; ++i) { double u = (double)rand() / RAND_MAX; hx[i] = mode == 0 ? 0.05 + 9.9 * u : 10 + 5000 * u; }
        cudaMemcpy(x, hx, n * 8, cudaMemcpyHostToDevice);
        for (int which = 0; which < 3; ++which) {
            float best = 1e9;
            for (int r = 0; r < 4; ++r) {
                cudaEventRecord(e0);
                if (which == 0) k<0><<<n / 256, 256>>>(x, y, n, reps);
                else if (which == 1) k<1><<<n / 256, 256>>>(x, y, n, reps);
                else k<2><<<n / 256, 256>>>(x, y, n, reps);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (r && ms < best) best = ms;
            }
            double calls = (double)n * reps;
            printf("%s args, %-10s : %.2f Gcalls/s  (%.1f SM-cycles*lanes per call)\n", mode == 0 ? "small" : "large",
                   which == 0 ? "lgamma" : (which == 1 ? "lgamma_pos" : "log"), calls / (best * 1e-3) / 1e9,
                   148 * 1.965e9 * 128 / (calls / (best * 1e-3)));
        }
    }
    return 0;
}
