// hb_march_tier.cu — instantiates hb_march_kernel for one lanes-per-wall tier (compile with -DHB_TIER=G).
#include "hb_kernels.cuh"

#define HB_CAT_(a, b) a##b
#define HB_CAT(a, b) HB_CAT_(a, b)

namespace hb {
namespace {
template <int G, int W, int VAR>
cudaError_t launch_variant(int n_tiles, cudaStream_t st, const DevView& V, const SeriesDev& P, const int32_t* ids) {
    size_t dyn = 0;
    if (VAR >= kVarExt) {
        dyn = (size_t)3 * W * (G * kNL + 1) * sizeof(double);
        static bool configured = false;  // per instantiation
        if (!configured) {
            cudaError_t e = cudaFuncSetAttribute(hb_march_kernel<G, W, VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)dyn);
            if (e != cudaSuccess) return e;
            configured = true;
        }
    }
    hb_march_kernel<G, W, VAR><<<n_tiles, G * W, dyn, st>>>(V, P, ids);
    return cudaGetLastError();
}
}  // namespace

cudaError_t HB_CAT(launch_march_, HB_TIER)(int var, int n_tiles, cudaStream_t st, const DevView& V, const SeriesDev& P,
                                           const int32_t* ids) {
    constexpr int G = HB_TIER, W = walls_of_group(HB_TIER);
    switch (var) {
        case kVarPlain: return launch_variant<G, W, kVarPlain>(n_tiles, st, V, P, ids);
        case kVarExt: return launch_variant<G, W, kVarExt>(n_tiles, st, V, P, ids);
        default: return launch_variant<G, W, kVarDist>(n_tiles, st, V, P, ids);
    }
}
}  // namespace hb
