// Scratch: dependent latency of the scalar building blocks (fast_cbrt, fast_rcp, tarp_natural, exp, F2F, MUFU).
#include <cstdio>
#include "../heat_b200/csrc/hb_kernels.cuh"
using namespace hb;
#define KERN(name, body) __global__ void name(double* out, int iters, double a) { double x = a; long long t0 = clock64(); \
  for (int i = 0; i < iters; ++i) { _Pragma("unroll") for (int j = 0; j < 4; ++j) { body; } } long long t1 = clock64(); \
  out[threadIdx.x] = x; if (threadIdx.x == 0) out[64] = (double)(t1 - t0) / (iters * 4.0); }
KERN(k_fast_cbrt, x = fast_cbrt(x) + 0.5)
KERN(k_fast_rcp, x = fast_rcp(x) + 0.5)
KERN(k_tarp, x = tarp_natural(x, 20.0, 1.3, 1.5) + 17.0)
KERN(k_exp, x = exp(-x) + 0.1)
KERN(k_f2f, x = (double)((float)x) + 0.5)
KERN(k_lg2, x = (double)__log2f((float)x) + 3.5)
KERN(k_sqrt, x = sqrt(x) + 0.5)
KERN(k_dsetp_sel, x = (x > 1.0 ? x * 0.5 : x + 0.7))
KERN(k_fabs_branch, if (fabs(x) >= 1e-3) x = x * 0.999 + 0.01; else x = 0.5)
int main() {
    double* d; cudaMalloc(&d, 4096); double h;
#define RUN(name, a) name<<<1, 32>>>(d, 2000, a); name<<<1, 32>>>(d, 2000, a); cudaDeviceSynchronize(); \
    cudaMemcpy(&h, d + 64, 8, cudaMemcpyDeviceToHost); printf("%-14s %.1f cycles (incl. one DADD)\n", #name, h);
    RUN(k_fast_cbrt, 2.0) RUN(k_fast_rcp, 2.0) RUN(k_tarp, 25.0) RUN(k_exp, 0.3) RUN(k_f2f, 1.5) RUN(k_lg2, 9.0) RUN(k_sqrt, 2.0)
    RUN(k_dsetp_sel, 1.5) RUN(k_fabs_branch, 1.5)
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
}
