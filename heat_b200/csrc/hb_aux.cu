// hb_aux.cu — K3 reduction and the probe kernels.
#define HB_AUX_KERNELS
#include "hb_kernels.cuh"

namespace hb {
void launch_reduce(int grid, cudaStream_t st, const DevView& V, int n_tiles, double* out, unsigned long long* iout) {
    hb_reduce_kernel<<<grid, kReduceBlock, 0, st>>>(V, n_tiles, out, iout);
}
void launch_dfma_probe(int blocks, int threads, double* out, int iters, double seed) {
    hb_dfma_probe_kernel<<<blocks, threads>>>(out, iters, seed);
}
void launch_fill(cudaStream_t st, double* p, int64_t n, double v) { hb_fill_kernel<<<148 * 4, 256, 0, st>>>(p, n, v); }
}  // namespace hb
