// hb_march_tier.cu — instantiates hb_march_kernel for one register tier (compile with -DHB_TIER=N).
#include "hb_kernels.cuh"

#define HB_CAT_(a, b) a##b
#define HB_CAT(a, b) HB_CAT_(a, b)

namespace hb {
void HB_CAT(launch_march_, HB_TIER)(bool dist, int n_tiles, cudaStream_t st, const DevView& V, const SeriesDev& P,
                                    const int32_t* ids) {
    if (dist)
        hb_march_kernel<HB_TIER, true, (HB_TIER > 32)><<<n_tiles, kBlock, 0, st>>>(V, P, ids);
    else
        hb_march_kernel<HB_TIER, false, (HB_TIER > 32)><<<n_tiles, kBlock, 0, st>>>(V, P, ids);
}
}  // namespace hb
