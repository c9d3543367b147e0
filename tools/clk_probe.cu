// clk_probe.cu -- effective SM clock (clock64 / globaltimer) under a sustained FP64 load, an FP32 load and a pure
// HBM copy load.  nvidia-smi's clocks.sm is sampled too slowly to show sub-ms power throttling.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/clk_probe tools/clk_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
#include <vector>

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

template <typename T>
__global__ void load_kernel(T* out, int iters, unsigned long long* stamps) {
    const unsigned long long c0 = clock64(), t0 = gtimer();
    T a[8];
    for (int k = 0; k < 8; k++) a[k] = (T)(threadIdx.x + k) * (T)1e-3;
    const T m = (T)1.0000001, c = (T)1e-9;
    for (int i = 0; i < iters; i++)
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = a[k] * m + c;
    T s = 0;
    for (int k = 0; k < 8; k++) s += a[k];
    if (s == (T)12345.678) out[0] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        stamps[0] = clock64() - c0;
        stamps[1] = gtimer() - t0;
    }
}

__global__ void copy_kernel(const double2* __restrict__ a, double2* __restrict__ b, size_t n, unsigned long long* stamps) {
    const unsigned long long c0 = clock64(), t0 = gtimer();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        stamps[0] = clock64() - c0;
        stamps[1] = gtimer() - t0;
    }
}

int main() {
    unsigned long long* st;
    cudaMallocManaged(&st, 16);
    double* outd; cudaMalloc(&outd, 8);
    float* outf; cudaMalloc(&outf, 4);
    const size_t n = (size_t)1 << 26;   // 1 GiB each
    double2 *a, *b; cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16);
    cudaMemset(a, 0, n * 16);
    for (int mode = 0; mode < 3; mode++) {
        printf("mode %s\n", mode == 0 ? "fp64 dfma" : mode == 1 ? "fp32 ffma" : "hbm copy");
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 40; rep++) {
            const int inner = rep < 5 ? 1 : 50;       // first reps: isolated launches; then sustained back-to-back
            cudaEventRecord(e0);
            for (int k = 0; k < inner; k++) {
                if (mode == 0) load_kernel<double><<<148 * 8, 256>>>(outd, 4000, st);
                else if (mode == 1) load_kernel<float><<<148 * 8, 256>>>(outf, 60000, st);
                else copy_kernel<<<148 * 16, 256>>>(a, b, n, st);
            }
            cudaEventRecord(e1);
            cudaDeviceSynchronize();
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep < 8 || rep % 8 == 0)
                printf("  rep %2d x%2d: %.3f ms/launch, block0 effective clock %.0f MHz\n", rep, inner, ms / inner,
                       (double)st[0] / (double)st[1] * 1e3);
            if (rep < 5) { cudaDeviceSynchronize(); }
        }
    }
    return 0;
}
