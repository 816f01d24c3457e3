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
    size_t dyn = (size_t)kWtWarps * s_max * sizeof(SubWeather);   // <= 16 KB
    const int blocks = (n_tiles + kWtWarps - 1) / kWtWarps;
    // A handful of tiles with the longest chains of the series END it (S x n_steps sub-steps in sequence): ask for most of an
    // SM's shared memory per block, so that such a block has its SM to itself instead of time-slicing the FP64 pipes
    // with the throughput-bound bulk (HB_WT_ISOLATE=0 switches it off).
    static const bool isolate = !(std::getenv("HB_WT_ISOLATE") && std::getenv("HB_WT_ISOLATE")[0] == '0');
    if (isolate && MODE != kFaceLean && s_max >= 96 && n_tiles <= 64) {
        dyn = 225 * 1024;
        static bool configured = false;   // per instantiation
        if (!configured) {
            cudaError_t e = cudaFuncSetAttribute(hb_wtile_kernel<G, NL, MODE, SIMPLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
            if (e != cudaSuccess) return e;
            configured = true;
        }
    }
#ifdef HB_PHASE_CLOCKS
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0}, o[8];
    cudaMemcpyToSymbol(hb_wt_clk, z, sizeof(z));
#endif
    hb_wtile_kernel<G, NL, MODE, SIMPLE><<<blocks, 32 * kWtWarps, dyn, st>>>(V, P, ids, n_tiles, s_max);
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
#ifdef HB_WT_BIG_NL
        case 20: return launch_wt_mode<G, 20>(mode, n_tiles, s_max, st, V, P, ids);
        case 24: return launch_wt_mode<G, 24>(mode, n_tiles, s_max, st, V, P, ids);
#endif
        default: return cudaErrorInvalidValue;
    }
}
}  // namespace hb
