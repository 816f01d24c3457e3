// Scratch: cycles per node_rk4 call for one warp (lanes-per-wall G) — isolates the halo-exchange cost.
#include <cstdio>
#include "../heat_b200/csrc/hb_kernels.cuh"
using namespace hb;
template <int G>
__global__ void k(double* out, int iters) {
    double T[kNL], A[kNL], B[kNL];
    for (int i = 0; i < kNL; ++i) { T[i] = 20. + threadIdx.x * 0.01 + i; A[i] = 0.01 + 0.001 * i; B[i] = 0.012 + 0.001 * i; }
    long long t0 = clock64();
    for (int it = 0; it < iters; ++it) node_rk4<G, kNL>(T, A, B);
    long long t1 = clock64();
    double s = 0; for (int i = 0; i < kNL; ++i) s += T[i];
    out[threadIdx.x + blockIdx.x * blockDim.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) out[4096] = (double)(t1 - t0) / iters;
}
template <int G> void run(int threads, int blocks, double* d) {
    double h;
    k<G><<<blocks, threads>>>(d, 2000); k<G><<<blocks, threads>>>(d, 2000);
    cudaDeviceSynchronize();
    cudaMemcpy(&h, d + 4096, 8, cudaMemcpyDeviceToHost);
    printf("G=%2d threads %3d blocks %d: %.0f cycles per node_rk4\n", G, threads, blocks, h);
}
int main() {
    double* d; cudaMalloc(&d, 8 * 8192);
    run<1>(32, 1, d); run<2>(32, 1, d); run<8>(32, 1, d); run<32>(32, 1, d);
    run<8>(64, 1, d); run<8>(128, 1, d); run<8>(256, 1, d);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
