// hb_wtile_tier.cu — instantiates hb_wtile_kernel for one lanes-per-wall tier (compile with -DHB_TIER=G).
#include <cstdio>
#include "hb_kernels.cuh"
#include "hb_wtile.cuh"

#define HB_CAT_(a, b) a##b
#define HB_CAT(a, b) HB_CAT_(a, b)

namespace hb {
namespace {
template <int G, int NL, int MODE>
cudaError_t launch_wt(int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P, const int32_t* ids) {
    const size_t dyn = (size_t)s_max * sizeof(SubWeather);   // <= 4 KB
    hb_wtile_kernel<G, NL, MODE, true><<<n_tiles, 32, dyn, st>>>(V, P, ids);
    return cudaGetLastError();
}
template <int G, int NL>
cudaError_t launch_wt_mode(int mode, int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P,
                           const int32_t* ids) {
    switch (mode) {
        case kFaceLean: return launch_wt<G, NL, kFaceLean>(n_tiles, s_max, st, V, P, ids);
        case kFaceLeanEnd: return launch_wt<G, NL, kFaceLeanEnd>(n_tiles, s_max, st, V, P, ids);
        default: return cudaErrorInvalidValue;
    }
}
}  // namespace

cudaError_t HB_CAT(launch_wtile_, HB_TIER)(int mode, int nl, int n_tiles, int s_max, cudaStream_t st, const DevView& V,
                                           const SeriesDev& P, const int32_t* ids) {
    constexpr int G = HB_TIER;
    switch (nl) {
        case 8: return launch_wt_mode<G, 8>(mode, n_tiles, s_max, st, V, P, ids);
        case 12: return launch_wt_mode<G, 12>(mode, n_tiles, s_max, st, V, P, ids);
        case 16: return launch_wt_mode<G, 16>(mode, n_tiles, s_max, st, V, P, ids);
        default: return cudaErrorInvalidValue;
    }
}
}  // namespace hb
