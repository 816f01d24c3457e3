// hb_march_tier.cu — instantiates hb_march_kernel for one lanes-per-wall tier (compile with -DHB_TIER=G).
#include <cstdio>
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
    if (!ids) {   // preload (hb_create): resolve the function now — its first launch would wait for whatever is running
        cudaFuncAttributes a;
        return cudaFuncGetAttributes(&a, hb_march_kernel<G, W, VAR>);
    }
    hb_march_kernel<G, W, VAR><<<n_tiles, G * W, dyn, st>>>(V, P, ids);
    return cudaGetLastError();
}
template <int G, int W, int END, bool SIMPLE, int NL = kNL>
cudaError_t launch_face(int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P,
                        const int32_t* ids) {
    const size_t dyn = (size_t)2 * s_max * sizeof(SubWeather);   // <= 8 KB: no opt-in needed
    if (!ids) {   // preload (hb_create)
        cudaFuncAttributes a;
        return cudaFuncGetAttributes(&a, hb_face_kernel<G, W, END, SIMPLE, NL>);
    }
#ifdef HB_PHASE_CLOCKS
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0}, o[8];
    cudaMemcpyToSymbol(hb_phase_clk, z, sizeof(z));
#endif
    hb_face_kernel<G, W, END, SIMPLE, NL><<<n_tiles, face_threads(G, W), dyn, st>>>(V, P, ids);
#ifdef HB_PHASE_CLOCKS
    cudaStreamSynchronize(st);
    cudaMemcpyFromSymbol(o, hb_phase_clk, sizeof(o));
    if (o[7])
        fprintf(stderr, "[hb clk] G=%d W=%d MODE=%d%s per sub-step of S>=100 tiles: waitB1 %.0f node %.0f waitB2 %.0f post %.0f "
                "zone %.0f pre %.0f\n", G, W, (int)END, SIMPLE ? "s" : "", (double)o[0] / o[7], (double)o[1] / o[7], (double)o[2] / o[7],
                (double)o[3] / o[7], (double)o[4] / o[7], (double)o[5] / o[7]);
#endif
    return cudaGetLastError();
}

template <int G, int END, bool SIMPLE>
cudaError_t launch_face_w(int W, int nl, int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P,
                          const int32_t* ids) {
    constexpr int W0 = pref_walls_of_group(G), W1 = walls_of_group(G);
    static_assert(W1 == W0 || W1 == 2 * W0, "two tile widths per tier");
    if constexpr (G == kMaxG) {   // walls beyond 254 nodes: 16 or 32 slots per lane
        if (nl == kNLBig && W == W0) return launch_face<G, W0, END, SIMPLE, kNLBig>(n_tiles, s_max, st, V, P, ids);
        if (nl == kNLBig && W == W1) return launch_face<G, W1, END, SIMPLE, kNLBig>(n_tiles, s_max, st, V, P, ids);
        if (nl == kNLHuge && W == W0) return launch_face<G, W0, END, SIMPLE, kNLHuge>(n_tiles, s_max, st, V, P, ids);
        if (nl == kNLHuge && W == W1) return launch_face<G, W1, END, SIMPLE, kNLHuge>(n_tiles, s_max, st, V, P, ids);
    }
    if (nl != kNL) return cudaErrorInvalidValue;
    if (W == W0) return launch_face<G, W0, END, SIMPLE>(n_tiles, s_max, st, V, P, ids);
    if (W == W1) return launch_face<G, W1, END, SIMPLE>(n_tiles, s_max, st, V, P, ids);
    return cudaErrorInvalidValue;
}
}  // namespace

cudaError_t HB_CAT(launch_march_, HB_TIER)(int var, int W, int nl, int n_tiles, int s_max, cudaStream_t st,
                                           const DevView& V, const SeriesDev& P, const int32_t* ids) {
    constexpr int G = HB_TIER, WL = walls_of_group(HB_TIER);
    switch (var) {
        case kVarPlain: return launch_face_w<G, kFaceLean, false>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarEnd: return launch_face_w<G, kFaceLeanEnd, false>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarGen: return launch_face_w<G, kFaceGen, false>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarPane: return launch_face_w<G, kFacePane, false>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarPlainSimple: return launch_face_w<G, kFaceLean, true>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarEndSimple: return launch_face_w<G, kFaceLeanEnd, true>(W, nl, n_tiles, s_max, st, V, P, ids);
        case kVarExt: return W == WL && nl == kNL ? launch_variant<G, WL, kVarExt>(n_tiles, st, V, P, ids) : cudaErrorInvalidValue;
        default: return W == WL && nl == kNL ? launch_variant<G, WL, kVarDist>(n_tiles, st, V, P, ids) : cudaErrorInvalidValue;
    }
}
}  // namespace hb
