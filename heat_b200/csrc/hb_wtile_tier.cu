// hb_wtile_tier.cu — instantiates hb_wtile_kernel for one lanes-per-wall tier (compile with -DHB_TIER=G).
#include <cstdio>
#include <cstdlib>
#include "hb_kernels.cuh"
#include "hb_wtile.cuh"

#define HB_CAT_(a, b) a##b
#define HB_CAT(a, b) HB_CAT_(a, b)

namespace hb {
namespace {
template <int G, int NL, int MODE, bool SIMPLE = true>
cudaError_t launch_wt(int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P, const int32_t* ids) {
    size_t dyn = (size_t)kWtWarps * wt_smem_per_warp(NL, MODE, s_max);   // <= 16 KB; NL > 16: + 0.5 KB per node slot and warp
    const int blocks = (n_tiles + kWtWarps - 1) / kWtWarps;
    if (wt_isolated(MODE, s_max, n_tiles)) {   // an SM per block (hb_dev.hpp)
        static const int kb = std::getenv("HB_WT_ISOLATE_KB") ? std::atoi(std::getenv("HB_WT_ISOLATE_KB")) : 225;
        dyn = (size_t)kb * 1024;
    }
    if (dyn > 48 * 1024) {   // opt in to more than 48 KB of dynamic shared memory: per instantiation AND per device, growing only
        static size_t configured_for[64] = {};
        int dev = 0;
        cudaError_t e = cudaGetDevice(&dev);
        if (e != cudaSuccess) return e;
        if (dev < 0 || dev >= 64 || dyn > configured_for[dev]) {
            e = cudaFuncSetAttribute(hb_wtile_kernel<G, NL, MODE, SIMPLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            if (e != cudaSuccess) return e;
            if (dev >= 0 && dev < 64) configured_for[dev] = dyn;
        }
    }
    if (!ids) {   // preload (hb_create): resolve the function now — its first launch would wait for whatever is running
        cudaFuncAttributes a;
        return cudaFuncGetAttributes(&a, hb_wtile_kernel<G, NL, MODE, SIMPLE>);
    }
#ifdef HB_PHASE_CLOCKS
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0}, o[8];
    cudaMemcpyToSymbol(hb_wt_clk, z, sizeof(z));
#endif
    if (wt_isolated(MODE, s_max, n_tiles)) {
        // pairs of blocks on the two SMs of a TPC (see the end of hb_wtile_kernel); an odd grid is padded by an empty block
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3((unsigned)wt_isolated_blocks(n_tiles));
        cfg.blockDim = dim3(32 * wt_isolated_warps());
        cfg.dynamicSmemBytes = dyn;
        cfg.stream = st;
        cudaLaunchAttribute attr{};
        attr.id = cudaLaunchAttributeClusterDimension;
        attr.val.clusterDim.x = 2;
        attr.val.clusterDim.y = 1;
        attr.val.clusterDim.z = 1;
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
        cudaError_t e = cudaLaunchKernelEx(&cfg, hb_wtile_kernel<G, NL, MODE, SIMPLE>, V, P, ids, n_tiles, s_max);
        if (e != cudaSuccess) return e;
    } else {
        hb_wtile_kernel<G, NL, MODE, SIMPLE><<<blocks, 32 * kWtWarps, dyn, st>>>(V, P, ids, n_tiles, s_max);
    }
#ifdef HB_PHASE_CLOCKS
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(o, hb_wt_clk, sizeof(o));
    if (o[7])
        fprintf(stderr, "[hb wclk] G=%d NL=%d MODE=%d tiles=%d smax=%d, cycles per sub-step of block 0: pre %.0f place %.0f rk4 %.0f "
                "gather %.0f post %.0f zone %.0f\n", G, NL, MODE, n_tiles, s_max, (double)o[0] / o[7], (double)o[1] / o[7],
                (double)o[2] / o[7], (double)o[3] / o[7], (double)o[4] / o[7], (double)o[5] / o[7]);
#endif
    return cudaGetLastError();
}
template <int G, int NL>
cudaError_t launch_wt_mode(int mode, int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P,
                           const int32_t* ids) {
    if constexpr (NL > 16) {   // (coefficients in shared memory) all-massive walls only, at most 4 lanes per wall
        if constexpr (G <= 4)
            if (mode == kFaceLean) return launch_wt<G, NL, kFaceLean>(n_tiles, s_max, st, V, P, ids);
        return cudaErrorInvalidValue;
    } else
    switch (mode) {
        case kFaceLean: return launch_wt<G, NL, kFaceLean>(n_tiles, s_max, st, V, P, ids);
        case kFaceLeanEnd: return launch_wt<G, NL, kFaceLeanEnd>(n_tiles, s_max, st, V, P, ids);
        case kFacePane:   // buildings: one zone behind all walls of a model, single panes, no-mass ends
            if constexpr (G <= 8) return launch_wt<G, NL, kFacePane, false>(n_tiles, s_max, st, V, P, ids);
            return cudaErrorInvalidValue;
        default: return cudaErrorInvalidValue;
    }
}
}  // namespace

cudaError_t HB_CAT(launch_wtile_, HB_TIER)(int mode, int nl, int n_tiles, int s_max, cudaStream_t st, const DevView& V,
                                           const SeriesDev& P, const int32_t* ids) {
    constexpr int G = HB_TIER;
    switch (nl) {
        case 8: return launch_wt_mode<G, 8>(mode, n_tiles, s_max, st, V, P, ids);
        case 10: return launch_wt_mode<G, 10>(mode, n_tiles, s_max, st, V, P, ids);
        case 12: return launch_wt_mode<G, 12>(mode, n_tiles, s_max, st, V, P, ids);
        case 16: return launch_wt_mode<G, 16>(mode, n_tiles, s_max, st, V, P, ids);
        case 24: return launch_wt_mode<G, 24>(mode, n_tiles, s_max, st, V, P, ids);
        case 32: return launch_wt_mode<G, 32>(mode, n_tiles, s_max, st, V, P, ids);
        default: return cudaErrorInvalidValue;
    }
}
}  // namespace hb
