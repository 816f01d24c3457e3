// Scratch micro-benchmark: dependent-issue latencies that shape the scalar chain of the march kernels.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lat_probe lat_probe.cu && ./lat_probe
#include <cstdio>
#include <cuda_runtime.h>

__global__ void k_dfma(double* out, int iters, double a, double b) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x = fma(x, b, a);
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_dadd(double* out, int iters, double a, double b) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x = x + b;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_ffma(double* out, int iters, float a, float b) {
    float x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x = fmaf(x, b, a);
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_rcp(double* out, int iters, double a) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) asm volatile("rcp.approx.ftz.f64 %0, %1;" : "=d"(x) : "d"(x));
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_lds(double* out, int iters) {
    __shared__ int s[64];
    s[threadIdx.x] = (threadIdx.x + 1) & 31;
    __syncthreads();
    int p = threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) p = ((volatile int*)s)[p];
    }
    long long t1 = clock64();
    out[threadIdx.x] = p;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_bar(double* out, int iters) {
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) __syncthreads();
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_shfl(double* out, int iters, double a) {
    double x = a + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x = __shfl_up_sync(0xffffffffu, x, 1, 8);
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 16.0);
}
__global__ void k_exp(double* out, int iters, double a) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) x = exp(-x) * 0.5;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 4.0);
}
__global__ void k_cbrt(double* out, int iters, double a) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) x = cbrt(x) + 1.0;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 4.0);
}
__global__ void k_div(double* out, int iters, double a, double b) {
    double x = a;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 4; ++j) x = b / x + 1.0;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 4.0);
}
// throughput: 4 independent DFMA chains in one warp
__global__ void k_dfma4(double* out, int iters, double a, double b) {
    double x0 = a, x1 = a + 1, x2 = a + 2, x3 = a + 3;
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { x0 = fma(x0, b, a); x1 = fma(x1, b, a); x2 = fma(x2, b, a); x3 = fma(x3, b, a); }
    }
    long long t1 = clock64();
    out[threadIdx.x] = x0 + x1 + x2 + x3;
    if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 32.0);
}

int main() {
    double* d;
    cudaMalloc(&d, 4096);
    double h;
    const int it = 2000;
#define RUN(name, thr, ...)                                            \
    name<<<1, thr>>>(d, it, ##__VA_ARGS__);                             \
    name<<<1, thr>>>(d, it, ##__VA_ARGS__);                             \
    cudaDeviceSynchronize();                                            \
    cudaMemcpy(&h, d + 64, 8, cudaMemcpyDeviceToHost);                  \
    printf("%-10s threads %3d: %.1f cycles per op\n", #name, thr, h);
    RUN(k_dfma, 32, 1.0, 0.999)
    RUN(k_dadd, 32, 1.0, 0.001)
    RUN(k_ffma, 32, 1.0f, 0.999f)
    RUN(k_rcp, 32, 1.5)
    RUN(k_lds, 32)
    RUN(k_bar, 32)
    RUN(k_bar, 64)
    RUN(k_bar, 256)
    RUN(k_shfl, 32, 1.0)
    RUN(k_exp, 32, 0.3)
    RUN(k_cbrt, 32, 0.3)
    RUN(k_div, 32, 0.3, 1.7)
    RUN(k_dfma4, 32, 1.0, 0.999)
    RUN(k_dfma4, 256, 1.0, 0.999)
    printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
