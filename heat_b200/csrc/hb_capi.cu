// hb_capi.cu — the C ABI declared in include/heat_b200.h: handle, device buffers, launches.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include <cuda_runtime.h>
#include <dlfcn.h>
#include <math_constants.h>

#include "hb_dev.hpp"

namespace {

thread_local std::string g_err;

// The launch groups of a series run concurrently on streams of their own; with the default of 8 hardware work queues
// streams share queues and groups queue up behind each other.  Takes effect when the library is loaded before the
// process creates its CUDA context (never overrides a value the user set).
struct EnvInit {
    EnvInit() { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", 0); }
} g_env_init;

struct CudaError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

#define HB_CUDA(expr)                                                                              \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            throw CudaError(std::string(#expr) + ": " + cudaGetErrorString(_e));                   \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    void alloc(size_t count) {
        free();
        n = count;
        if (count) HB_CUDA(cudaMalloc(&p, count * sizeof(T)));
    }
    void upload(const std::vector<T>& h, cudaStream_t st) {
        alloc(h.size());
        if (!h.empty()) HB_CUDA(cudaMemcpyAsync(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, st));
    }
    void free() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
    ~DevBuf() { free(); }
};

// ---- NCCL, resolved at run time (dlopen): the library carries no link-time dependency on it, and a process that
// already loaded a libnccl.so.2 (torch bundles one) shares that copy ----
struct Nccl {
    using Comm = void*;
    typedef struct { char internal[128]; } UniqueId;
    int (*GetUniqueId)(UniqueId*) = nullptr;
    int (*CommInitRank)(Comm*, int, UniqueId, int) = nullptr;
    int (*CommDestroy)(Comm) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, Comm, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    enum { kFloat64 = 8, kSum = 0, kMax = 2, kMin = 3 };   // ncclDataType_t / ncclRedOp_t values of nccl.h
    static Nccl& get() {
        static Nccl n = [] {
            Nccl x;
            // A process holds ONE libnccl.so.2 (the loader matches by soname): whoever loads it first decides which.
            // HEAT_B200_NCCL names the copy to take when this library is first — heat_b200/_lib.py points it at the
            // NCCL that ships with torch, so that a later `import torch` still finds the symbols it was built against.
            void* lib = nullptr;
            if (const char* path = std::getenv("HEAT_B200_NCCL"))
                if (path[0]) lib = dlopen(path, RTLD_NOW | RTLD_GLOBAL);
            if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
            if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
            if (!lib) throw CudaError(std::string("cannot load libnccl.so.2: ") + dlerror());
            auto sym = [&](const char* name) {
                void* p = dlsym(lib, name);
                if (!p) throw CudaError(std::string("libnccl has no symbol ") + name);
                return p;
            };
            x.GetUniqueId = reinterpret_cast<decltype(x.GetUniqueId)>(sym("ncclGetUniqueId"));
            x.CommInitRank = reinterpret_cast<decltype(x.CommInitRank)>(sym("ncclCommInitRank"));
            x.CommDestroy = reinterpret_cast<decltype(x.CommDestroy)>(sym("ncclCommDestroy"));
            x.AllReduce = reinterpret_cast<decltype(x.AllReduce)>(sym("ncclAllReduce"));
            x.GetErrorString = reinterpret_cast<decltype(x.GetErrorString)>(sym("ncclGetErrorString"));
            return x;
        }();
        return n;
    }
    void check(int rc, const char* what) const {
        if (rc != 0) throw CudaError(std::string(what) + ": " + (GetErrorString ? GetErrorString(rc) : "NCCL error"));
    }
};

}  // namespace

struct hb_handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    hb::BatchPlan plan;
    int64_t launches = 0;

    // device buffers (uploaded verbatim from hb::HostArrays)
    DevBuf<int32_t> n_nodes, flags, zone_f, zone_b, site, desc_index, nm_off, nm_pairs, cav_off;
    DevBuf<hb::CavityDev> cav;
    DevBuf<double> area, nx, ny, cos_tilt, forced_c, eps_f, eps_b, amb_f, amb_b, hs_ov_f, hs_ov_b, af0, abL, v0, vL, u_f, u_b, u_f2, u_b2, ax_f, ax_b, wa1f, wa1b, wa2f, wa2b;
    DevBuf<double> nodeT, coefA, coefB, ext_u, ext_mass, ext_af, ext_ab, sc, energy_f, energy_b;
    DevBuf<unsigned long long> nm_iters;
    DevBuf<int32_t> status;
    DevBuf<double> zoneT, zvol, zq, zinf_v, zinf_t, zvent_v, zvent_t;
    DevBuf<int32_t> z_surf_begin, z_surf_end;
    DevBuf<hb::Tile> tiles;
    DevBuf<double> red_out;
    DevBuf<unsigned long long> red_iout;
    DevBuf<double> comm_buf;        // staging of hb_reduce_all: 7 sums | 2 minima | 2 + 8 maxima
    bool zone_has_inf_v = false, zone_has_vent_v = false;   // volumes given with hb_set_zone_inputs
    void* comm = nullptr;           // ncclComm_t (hb_comm_init)
    int comm_world = 1;
    DevBuf<double> flush;

    // one launch per (lanes-per-wall, kernel variant); each on its own stream so that small groups overlap
    struct Group {
        int G, W, var;
        int NL = hb::kNL;   // slots per lane of the group's tiles
        int n_tiles;
        int s_max = 1;   // most sub-steps of a tile of the group
        DevBuf<int32_t> ids;   // tile ids, heaviest first
        cudaStream_t stream = nullptr;
        cudaEvent_t done = nullptr;
        Group() = default;
        Group(const Group&) = delete;
        Group& operator=(const Group&) = delete;
        ~Group() {   // a create that throws half-way (cudaMalloc OOM on a large shard) must not leak streams / events
            if (stream) {
                cudaStreamSynchronize(stream);
                cudaStreamDestroy(stream);
            }
            if (done) cudaEventDestroy(done);
        }
    };
    std::vector<std::unique_ptr<Group>> groups;
    cudaEvent_t fork = nullptr;
    // "block has started" flags of the launches that take an SM per block (hb::wt_isolated): pinned host memory the
    // kernels write and launch_series polls
    static constexpr int kStartedSlots = 256;
    volatile int32_t* started = nullptr;
    std::vector<int32_t> host_site;  // per device surface slot (rebuilt when model_site changes)
    std::vector<int32_t> model_site_cache;
    bool single_site = true;

    hb_handle() = default;
    hb_handle(const hb_handle&) = delete;
    hb_handle& operator=(const hb_handle&) = delete;
    ~hb_handle() {
        if (device < 0) return;
        cudaSetDevice(device);
        if (comm) Nccl::get().CommDestroy(comm);
        groups.clear();   // joins and destroys the group streams / events
        if (fork) cudaEventDestroy(fork);
        if (started) cudaFreeHost(const_cast<int32_t*>(started));
        if (stream) {
            cudaStreamSynchronize(stream);
            cudaStreamDestroy(stream);
        }
    }

    hb::DevView view() {
        hb::DevView V{};
        V.n_nodes = n_nodes.p; V.flags = flags.p; V.zone_f = zone_f.p; V.zone_b = zone_b.p; V.site = site.p;
        V.desc_index = desc_index.p; V.area = area.p; V.nx = nx.p; V.ny = ny.p; V.cos_tilt = cos_tilt.p;
        V.forced_c = forced_c.p; V.eps_f = eps_f.p; V.eps_b = eps_b.p; V.amb_f = amb_f.p; V.amb_b = amb_b.p;
        V.hs_ov_f = hs_ov_f.p; V.hs_ov_b = hs_ov_b.p; V.af0 = af0.p; V.abL = abL.p; V.v0 = v0.p; V.vL = vL.p; V.u_f = u_f.p; V.u_b = u_b.p; V.u_f2 = u_f2.p; V.u_b2 = u_b2.p; V.ax_f = ax_f.p; V.ax_b = ax_b.p;
        V.wa1f = wa1f.p; V.wa1b = wa1b.p; V.wa2f = wa2f.p; V.wa2b = wa2b.p;
        V.nm_off = nm_off.p; V.nm_pairs = nm_pairs.p; V.cav_off = cav_off.p; V.cav = cav.p;
        V.nodeT = nodeT.p; V.coefA = coefA.p; V.coefB = coefB.p;
        V.ext_u = ext_u.p; V.ext_mass = ext_mass.p; V.ext_af = ext_af.p; V.ext_ab = ext_ab.p;
        V.sc = sc.p; V.n_dev_surfaces = plan.n_dev_surfaces; V.energy_f = energy_f.p; V.energy_b = energy_b.p;
        V.nm_iters = nm_iters.p; V.status = status.p;
        V.zoneT = zoneT.p; V.zvol = zvol.p; V.zq = zq.p; V.zinf_v = zinf_v.p; V.zinf_t = zinf_t.p;
        V.zvent_v = zvent_v.p; V.zvent_t = zvent_t.p;
        V.z_surf_begin = z_surf_begin.p; V.z_surf_end = z_surf_end.p; V.n_dev_zones = plan.n_dev_zones;
        V.tiles = tiles.p;
        return V;
    }
};

namespace {


void need_device(const hb_handle* h) {
    if (h->device < 0)
        throw CudaError("handle was created with HB_DEVICE_NONE (host planning only); heat_b200 has no CPU fallback");
    HB_CUDA(cudaSetDevice(h->device));
}

template <class F>
hb_status guarded(F&& f) {
    try {
        f();
        return HB_OK;
    } catch (const CudaError& e) {
        g_err = e.what();
        return HB_ERR_CUDA;
    } catch (const std::exception& e) {
        g_err = e.what();
        const bool unsupported = g_err.find("not implemented") != std::string::npos;
        return unsupported ? HB_ERR_UNSUPPORTED : HB_ERR_INVALID;
    }
}

void build_device(hb_handle* h) {
    const hb::BatchPlan& B = h->plan;
    hb::HostArrays H;
    hb::pack_arrays(B, H);
    cudaStream_t st = h->stream;
    h->n_nodes.upload(H.n_nodes, st); h->flags.upload(H.flags, st); h->zone_f.upload(H.zone_f, st);
    h->zone_b.upload(H.zone_b, st); h->site.upload(H.site, st); h->desc_index.upload(H.desc_index, st);
    h->nm_off.upload(H.nm_off, st); h->nm_pairs.upload(H.nm_pairs, st); h->cav_off.upload(H.cav_off, st);
    h->cav.upload(H.cav, st);
    h->area.upload(H.area, st); h->nx.upload(H.nx, st); h->ny.upload(H.ny, st); h->cos_tilt.upload(H.cos_tilt, st);
    h->forced_c.upload(H.forced_c, st); h->eps_f.upload(H.eps_f, st); h->eps_b.upload(H.eps_b, st);
    h->amb_f.upload(H.amb_f, st); h->amb_b.upload(H.amb_b, st); h->hs_ov_f.upload(H.hs_ov_f, st);
    h->hs_ov_b.upload(H.hs_ov_b, st); h->af0.upload(H.af0, st); h->abL.upload(H.abL, st); h->v0.upload(H.v0, st);
    h->vL.upload(H.vL, st); h->u_f.upload(H.u_f, st); h->u_b.upload(H.u_b, st); h->u_f2.upload(H.u_f2, st); h->u_b2.upload(H.u_b2, st); h->ax_f.upload(H.ax_f, st); h->ax_b.upload(H.ax_b, st);
    h->wa1f.upload(H.wa1f, st); h->wa1b.upload(H.wa1b, st); h->wa2f.upload(H.wa2f, st); h->wa2b.upload(H.wa2b, st);
    h->nodeT.upload(H.nodeT, st); h->coefA.upload(H.coefA, st); h->coefB.upload(H.coefB, st);
    h->ext_u.upload(H.ext_u, st); h->ext_mass.upload(H.ext_mass, st); h->ext_af.upload(H.ext_af, st);
    h->ext_ab.upload(H.ext_ab, st); h->sc.upload(H.sc, st);
    h->energy_f.upload(H.energy_f, st); h->energy_b.upload(H.energy_b, st);
    h->nm_iters.upload(H.nm_iters, st); h->status.upload(H.status, st);
    h->zoneT.upload(H.zoneT, st); h->zvol.upload(H.zvol, st); h->zq.upload(H.zq, st); h->zinf_v.upload(H.zinf_v, st);
    h->zinf_t.upload(H.zinf_t, st); h->zvent_v.upload(H.zvent_v, st); h->zvent_t.upload(H.zvent_t, st);
    h->z_surf_begin.upload(H.z_surf_begin, st); h->z_surf_end.upload(H.z_surf_end, st);
    h->tiles.upload(B.tiles, st);
    h->red_out.alloc(8);
    h->red_iout.alloc(4);
    h->host_site.assign(B.n_dev_surfaces, 0);

    // launch groups: (lanes per wall, kernel variant); inside a group the heaviest tiles are scheduled first
    // A tile's run time is proportional to its sub-step count S (2 … 118), and a series is only as short as its
    // longest tile: tiles are binned by S and the bins with the most sub-steps are launched first, on streams of
    // higher priority, so that across ALL groups the long tiles start at once and the short ones fill in behind.
    // (three bins: with more than 32 launch groups — the number of hardware work queues, CUDA_DEVICE_MAX_CONNECTIONS —
    // late groups queue up behind the longest ones they share a queue with)
    auto s_class = [](int S) { return S >= 56 ? 0 : S >= 20 ? 1 : 2; };
    std::map<std::tuple<int, int, int, int, int, int>, std::vector<int32_t>> by;   // (S bin, fast?, -NL, -G, variant, W)
    for (size_t ti = 0; ti < B.tiles.size(); ++ti) {
        const hb::Tile& t = B.tiles[ti];
        const int var = (t.flags & hb::TILE_WARP) ? (!(t.flags & hb::TILE_SIMPLE) ? 10 : (t.flags & hb::TILE_END) ? 9 : 8)   // warp tiles
                        : (t.flags & hb::TILE_DIST) ? 2 : (t.flags & hb::TILE_EXT) ? 1 : (t.flags & hb::TILE_GEN) ? 4 : (t.flags & hb::TILE_PANE) ? 5
                        : (t.flags & hb::TILE_END) ? ((t.flags & hb::TILE_SIMPLE) ? 7 : 3) : ((t.flags & hb::TILE_SIMPLE) ? 6 : 0);
        // slow variants (shared-memory mirror, fixed-point loops) before fast ones inside a bin
        by[{s_class(t.S), var == 1 || var == 2 ? 0 : 1, -t.NL, -t.G, var, t.W}].push_back((int32_t)ti);
    }
    int prio_lo = 0, prio_hi = 0;   // numerically lower = higher priority
    HB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    HB_CUDA(cudaEventCreateWithFlags(&h->fork, cudaEventDisableTiming));
    {
        void* p = nullptr;
        HB_CUDA(cudaHostAlloc(&p, hb_handle::kStartedSlots * sizeof(int32_t), cudaHostAllocMapped | cudaHostAllocPortable));
        std::memset(p, 0, hb_handle::kStartedSlots * sizeof(int32_t));
        h->started = static_cast<volatile int32_t*>(p);
    }
    for (auto& kv : by) {
        auto grp = std::make_unique<hb_handle::Group>();
        grp->NL = -std::get<2>(kv.first);
        grp->G = -std::get<3>(kv.first);
        grp->var = std::get<4>(kv.first);
        grp->W = std::get<5>(kv.first);
        std::stable_sort(kv.second.begin(), kv.second.end(),
                         [&](int32_t a, int32_t b) { return B.tiles[a].S > B.tiles[b].S; });
        grp->n_tiles = (int)kv.second.size();
        grp->s_max = B.tiles[kv.second.front()].S;
        grp->ids.upload(kv.second, st);
        h->groups.push_back(std::move(grp));
    }
    // Launch order and stream priority: first the few no-mass-end tiles that take an SM per block (they need empty SMs),
    // then by class of chain length (longest first: they bound the series), and inside a class the kernels compiled
    // for 255 registers before those for 168 — on streams of decreasing priority in the same order.  A sub-partition's
    // register file holds 2 warps of 255 registers or 3 of 168, but only 1 + 1 of a mix (a sixth of its nodes fewer in
    // flight): with the 255-register blocks ahead in every class, SMs fill with one kind and go over to the other
    // only when its queue is empty, instead of mixing them at random (run-to-run spread of a series: 6 % -> see
    // profiles/r2_summary.md).
    auto mode_of = [](const hb_handle::Group& g) { return g.var == 10 ? 3 : g.var == 9 ? 1 : 0; };
    auto rank = [&](const hb_handle::Group& g) { return g.var >= 8 && hb::wt_isolated(mode_of(g), g.s_max, g.n_tiles) ? 1 : 0; };
    auto heavy = [](const hb_handle::Group& g) { return g.var >= 8 ? (g.NL > 12 || g.var != 8 ? 1 : 0) : 1; };
    static const bool by_regs = !(std::getenv("HB_PRIO_BY_REGS") && std::getenv("HB_PRIO_BY_REGS")[0] == '0');
    std::stable_sort(h->groups.begin(), h->groups.end(),
                     [&](const std::unique_ptr<hb_handle::Group>& a, const std::unique_ptr<hb_handle::Group>& b) {
                         if (rank(*a) != rank(*b)) return rank(*a) > rank(*b);
                         if (by_regs && s_class(a->s_max) != s_class(b->s_max)) return s_class(a->s_max) < s_class(b->s_max);
                         if (by_regs && heavy(*a) != heavy(*b)) return heavy(*a) > heavy(*b);
                         return a->s_max > b->s_max;
                     });
    for (auto& grp : h->groups) {
        const int cls = s_class(grp->s_max);
        const int level = by_regs ? (rank(*grp) ? 0 : 2 * cls + (heavy(*grp) ? 0 : 1)) : cls;
        HB_CUDA(cudaStreamCreateWithPriority(&grp->stream, cudaStreamNonBlocking, std::max(prio_hi, std::min(prio_lo, prio_hi + level))));
        HB_CUDA(cudaEventCreateWithFlags(&grp->done, cudaEventDisableTiming));
    }
    HB_CUDA(cudaStreamSynchronize(st));
}

}  // namespace

namespace hb {
// one translation unit per lanes-per-wall tier (hb_march_tier.cu compiled with -DHB_TIER=G) so they build in parallel
#define HB_DECL_TIER(N) \
    cudaError_t launch_march_##N(int var, int W, int nl, int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P, const int32_t* ids);
HB_DECL_TIER(1) HB_DECL_TIER(2) HB_DECL_TIER(4) HB_DECL_TIER(8) HB_DECL_TIER(16) HB_DECL_TIER(32)
#undef HB_DECL_TIER
// warp tiles: mode = hb_face_kernel's lean modes (0 all-massive, 1 no-mass ends)
#define HB_DECL_WTIER(N) \
    cudaError_t launch_wtile_##N(int mode, int nl, int n_tiles, int s_max, cudaStream_t st, const DevView& V, const SeriesDev& P, const int32_t* ids);
HB_DECL_WTIER(1) HB_DECL_WTIER(2) HB_DECL_WTIER(4) HB_DECL_WTIER(8) HB_DECL_WTIER(16) HB_DECL_WTIER(32)
#undef HB_DECL_WTIER
void launch_reduce(int grid, cudaStream_t st, const DevView& V, int n_tiles, double* out, unsigned long long* iout);
void launch_dfma_probe(int blocks, int threads, double* out, int iters, double seed);
void launch_fill(cudaStream_t st, double* p, int64_t n, double v);
}  // namespace hb

namespace {

// The launcher of a group's kernel (ids == nullptr: resolve the kernel only, nothing is launched).
cudaError_t launch_group(const hb_handle::Group& g, int var, cudaStream_t st, const hb::DevView& V, const hb::SeriesDev& P,
                         const int32_t* ids) {
    if (var >= 8) {   // warp tiles (per-site weather is handled inside the kernel)
        const int mode = var == 10 ? 3 : var == 9 ? 1 : 0;   // 3: buildings (one zone per model, panes, no-mass ends)
        switch (g.G) {
            case 1: return hb::launch_wtile_1(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
            case 2: return hb::launch_wtile_2(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
            case 4: return hb::launch_wtile_4(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
            case 8: return hb::launch_wtile_8(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
            case 16: return hb::launch_wtile_16(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
            default: return hb::launch_wtile_32(mode, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        }
    }
    switch (g.G) {
        case 1: return hb::launch_march_1(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        case 2: return hb::launch_march_2(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        case 4: return hb::launch_march_4(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        case 8: return hb::launch_march_8(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        case 16: return hb::launch_march_16(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
        default: return hb::launch_march_32(var, g.W, g.NL, g.n_tiles, g.s_max, st, V, P, ids);
    }
}

// CUDA resolves a kernel at its first launch, and that waits for the kernels already running: the first series of a
// process ran its ~30 launch groups one after the other (11.5 s against 1.6 s for a year of configs[2]).  Resolve every
// kernel a handle can launch when the handle is made — the per-site variant of the lean block-tile kernels included.
void preload_kernels(hb_handle* h) {
    const hb::DevView V = h->view();
    const hb::SeriesDev P{};
    for (auto& gp : h->groups) {
        const hb_handle::Group& g = *gp;
        HB_CUDA(launch_group(g, g.var, nullptr, V, P, nullptr));
        if (g.var == 0 || g.var == 3 || (g.var >= 5 && g.var < 8)) HB_CUDA(launch_group(g, 4, nullptr, V, P, nullptr));
    }
}

// Fork the handle's stream into one stream per launch group, join back: the groups run concurrently and the
// handle's stream (and CUDA events recorded on it) sees the whole series.
void launch_series(hb_handle* h, hb::SeriesDev P) {
    const hb::DevView V = h->view();
    P.single_site = h->single_site ? 1 : 0;
    HB_CUDA(cudaEventRecord(h->fork, h->stream));
    // HB_GROUP_TIMES=1 (diagnostic): run the launch groups one after the other and print each one's duration
    static const bool group_times = std::getenv("HB_GROUP_TIMES") != nullptr;
    // HB_GROUP_TIMELINE=1 (diagnostic): start and end of every group relative to the fork, groups left concurrent
    // (the timing events hold the hardware queues: use it with CUDA_DEVICE_MAX_CONNECTIONS=32)
    static const bool timeline = std::getenv("HB_GROUP_TIMELINE") != nullptr;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> tl;
    cudaEvent_t tl0 = nullptr;
    if (timeline) {
        HB_CUDA(cudaEventCreate(&tl0));
        HB_CUDA(cudaEventRecord(tl0, h->stream));
    }
    // The groups that take an SM per block come first (create() sorts them so).  Blocks are not placed in launch order
    // across hardware queues: if the bulk fills the SMs first, such a block waits until one DRAINS — measured: 40–110 ms
    // of a 410 ms series for the last of four such groups.  So every one of their blocks reports when it begins (a flag
    // in pinned host memory) and the bulk is launched once all have, or after 2 ms (the series goes on regardless).
    int n_started = 0;
    bool gate_open = false;
    for (int i = 0; i < hb_handle::kStartedSlots; ++i) h->started[i] = 0;
    auto wait_for_isolated = [&] {
        gate_open = true;
        if (n_started == 0 || group_times) return;
        const auto t_begin = std::chrono::steady_clock::now(), t_end = t_begin + std::chrono::milliseconds(2);
        int seen = 0;
        do {
            seen = 0;
            for (int i = 0; i < n_started; ++i) seen += h->started[i] != 0;
        } while (seen < n_started && std::chrono::steady_clock::now() < t_end);
        if (timeline)
            std::fprintf(stderr, "[hb tl] %d of %d SM-per-block blocks had begun after %.0f us\n", seen, n_started,
                         std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t_begin).count());
    };
    for (auto& gp : h->groups) {
        hb_handle::Group& g = *gp;
        hb::SeriesDev Pg = P;
        {
            const int mode = g.var == 10 ? 3 : g.var == 9 ? 1 : 0;
            const int blocks = hb::wt_isolated_blocks(g.n_tiles);
            if (!gate_open && g.var >= 8 && hb::wt_isolated(mode, g.s_max, g.n_tiles) &&
                n_started + blocks <= hb_handle::kStartedSlots) {
                HB_CUDA(cudaHostGetDevicePointer((void**)&Pg.started, const_cast<int32_t*>(h->started) + n_started, 0));
                n_started += blocks;
            } else if (!gate_open) {
                wait_for_isolated();
            }
        }
        HB_CUDA(cudaStreamWaitEvent(g.stream, h->fork, 0));
        cudaEvent_t t0 = nullptr, t1 = nullptr;
        if (group_times) {
            HB_CUDA(cudaEventCreate(&t0));
            HB_CUDA(cudaEventCreate(&t1));
            HB_CUDA(cudaEventRecord(t0, g.stream));
        }
        if (timeline) {
            cudaEvent_t a, b;
            HB_CUDA(cudaEventCreate(&a));
            HB_CUDA(cudaEventCreate(&b));
            HB_CUDA(cudaEventRecord(a, g.stream));
            tl.emplace_back(a, b);
        }
        cudaError_t e = cudaSuccess;
        // the lean face kernel (variant 0) reads one weather table per tile: per-site weather takes the general one
        const int var = (g.var == 0 || g.var == 3 || (g.var >= 5 && g.var < 8)) && !P.single_site ? 4 : g.var;
        // A group's blocks march the whole series one after the other on the SM slots they get: a series is a list
        // schedule of (tiles) jobs on (resident warps) slots, and its tail — slots idle while the last jobs end — is
        // about half a job.  The groups of few sub-steps are the filler that runs last (lowest stream priority): their
        // jobs are cut into segments of about kSegmentMs, launched back to back on the group's stream (the state goes
        // through HBM in between: 40 B per node, nothing).  Long chains are never cut: a segment would wait for a slot.
        int segments = 1;
        static const int seg_smax = std::getenv("HB_WT_SEGMENT_SMAX") ? std::atoi(std::getenv("HB_WT_SEGMENT_SMAX")) : 20;
        if (var >= 8 && !Pg.started && g.s_max < seg_smax && P.record_every == 0 && P.ir_mode != HB_IR_NET_GAIN && !group_times) {
            static const double seg_ms = std::getenv("HB_WT_SEGMENT_MS") ? std::atof(std::getenv("HB_WT_SEGMENT_MS")) : 20.;
            const double job_ms = (double)P.n_steps * g.s_max * 3800. / 1.9e6;   // ~3800 cycles per sub-step when the SM is full
            if (seg_ms > 0.) segments = (int)std::min(32., std::max(1., std::floor(job_ms / seg_ms)));
        }
        Pg.step_begin = 0;
        Pg.step_end = P.n_steps;
        for (int seg = 0; seg < segments && e == cudaSuccess; ++seg) {
        if (segments > 1) {
            Pg.step_begin = (int32_t)((int64_t)P.n_steps * seg / segments);
            Pg.step_end = (int32_t)((int64_t)P.n_steps * (seg + 1) / segments);
        }
        e = launch_group(g, var, g.stream, V, Pg, g.ids.p);
        h->launches += 1;
        }
        HB_CUDA(e);
        if (group_times) {
            HB_CUDA(cudaEventRecord(t1, g.stream));
            HB_CUDA(cudaEventSynchronize(t1));
            float ms = 0.f;
            HB_CUDA(cudaEventElapsedTime(&ms, t0, t1));
            std::fprintf(stderr, "[hb] group G=%d NL=%d W=%d var=%d tiles=%d: %.3f ms\n", g.G, g.NL, g.W, g.var, g.n_tiles, ms);
            cudaEventDestroy(t0);
            cudaEventDestroy(t1);
        }
        if (timeline) HB_CUDA(cudaEventRecord(tl.back().second, g.stream));
        HB_CUDA(cudaEventRecord(g.done, g.stream));
        HB_CUDA(cudaStreamWaitEvent(h->stream, g.done, 0));
    }
    if (timeline) {
        HB_CUDA(cudaStreamSynchronize(h->stream));
        for (size_t i = 0; i < tl.size(); ++i) {
            const hb_handle::Group& g = *h->groups[i];
            float a = 0.f, b = 0.f;
            cudaEventElapsedTime(&a, tl0, tl[i].first);
            cudaEventElapsedTime(&b, tl0, tl[i].second);
            std::fprintf(stderr, "[hb tl] G=%d NL=%d W=%d var=%d tiles=%d smax=%d: %.2f -> %.2f ms\n", g.G, g.NL, g.W, g.var,
                         g.n_tiles, g.s_max, a, b);
            cudaEventDestroy(tl[i].first);
            cudaEventDestroy(tl[i].second);
        }
        cudaEventDestroy(tl0);
        if (n_started > 0) {   // when each of those blocks began (~µs, GPU clock), relative to the first
            int32_t lo = INT32_MAX;
            for (int i = 0; i < n_started; ++i) if (h->started[i]) lo = std::min(lo, (int32_t)(h->started[i] & 0xfffff));
            std::fprintf(stderr, "[hb tl] SM-per-block blocks: SM id @ begin (us):");
            for (int i = 0; i < n_started; ++i)
                std::fprintf(stderr, " %d@%d", (int)((h->started[i] >> 20) & 0x3ff), h->started[i] ? (int)((h->started[i] & 0xfffff) - lo) : -1);
            std::fprintf(stderr, "\n");
        }
    }
}

void validate_series(const hb_handle* h, const hb_series* s) {
    if (!s) throw std::runtime_error("null series");
    if (s->n_steps < 0) throw std::runtime_error("n_steps must be >= 0");
    if (s->n_sites < 1) throw std::runtime_error("n_sites must be >= 1");
    if (!s->weather) throw std::runtime_error("Trying to march on Thermal Model, but dry bulb temperature was not provided");
    const hb::BatchPlan& B = h->plan;
    for (const auto& t : B.tiles)
        if (t.S > hb::kMaxSubsteps) throw std::runtime_error("more sub-steps per main step than the supported maximum (128)");
    int64_t need = s->n_steps;
    if (s->weather_mode == HB_WEATHER_LINEAR) need = (int64_t)s->n_steps + 1;
    if (s->weather_mode == HB_WEATHER_PER_SUBSTEP) {
        // all models of a site must share S
        std::vector<int> site_S(s->n_sites, -1);
        int64_t max_rows = 0;
        for (size_t m = 0; m < B.models.size(); ++m) {
            const int site = s->model_site ? s->model_site[m] : 0;
            if (site < 0 || site >= s->n_sites) throw std::runtime_error("model_site out of range");
            if (site_S[site] < 0) site_S[site] = B.models[m].dt_subdivisions;
            else if (site_S[site] != B.models[m].dt_subdivisions)
                throw std::runtime_error("HB_WEATHER_PER_SUBSTEP needs one dt_subdivisions per site");
            max_rows = std::max<int64_t>(max_rows, (int64_t)s->n_steps * B.models[m].dt_subdivisions);
        }
        need = max_rows;
    } else if (s->weather_mode != HB_WEATHER_STEP_CONSTANT && s->weather_mode != HB_WEATHER_LINEAR) {
        throw std::runtime_error("unknown weather_mode");
    }
    if (s->weather_rows < need) throw std::runtime_error("weather table has too few rows for this series");
    if (s->input_mode != HB_INPUT_STATE && !s->inputs) throw std::runtime_error("inputs missing for this input_mode");
    if (s->input_mode < HB_INPUT_STATE || s->input_mode > HB_INPUT_PER_SURFACE) throw std::runtime_error("unknown input_mode");
    if (s->model_site)
        for (size_t m = 0; m < B.models.size(); ++m)
            if (s->model_site[m] < 0 || s->model_site[m] >= s->n_sites) throw std::runtime_error("model_site out of range");
}

void update_sites(hb_handle* h, const hb_series* s) {
    const hb::BatchPlan& B = h->plan;
    std::vector<int32_t> ms(B.models.size(), 0);
    if (s->model_site) ms.assign(s->model_site, s->model_site + B.models.size());
    if (ms == h->model_site_cache) return;
    h->model_site_cache = ms;
    h->single_site = std::all_of(ms.begin(), ms.end(), [](int32_t v) { return v == 0; });
    std::fill(h->host_site.begin(), h->host_site.end(), 0);
    for (size_t gd = 0; gd < B.dev_surface_of.size(); ++gd) h->host_site[B.dev_surface_of[gd]] = ms[B.surface_model[gd]];
    HB_CUDA(cudaMemcpyAsync(h->site.p, h->host_site.data(), h->host_site.size() * sizeof(int32_t),
                            cudaMemcpyHostToDevice, h->stream));
}

int64_t input_count(const hb_handle* h, const hb_series* s) {
    if (s->input_mode == HB_INPUT_SITE) return (int64_t)s->n_sites * s->n_steps * 4;
    if (s->input_mode == HB_INPUT_PER_SURFACE) return (int64_t)s->n_steps * (int64_t)h->plan.surfaces.size() * 4;
    return 0;
}

void clear_status(hb_handle* h) {
    if (h->status.n) HB_CUDA(cudaMemsetAsync(h->status.p, 0, h->status.n * sizeof(int32_t), h->stream));
}

void check_status(hb_handle* h) {
    // a non-zero per-surface status word is the reference's panic! — of THIS call: the words are cleared at the start of
    // every march (clear_status), so a handle stays usable after a reported failure (hb_set_state + march again)
    std::vector<int32_t> st(h->plan.n_dev_surfaces);
    HB_CUDA(cudaMemcpyAsync(st.data(), h->status.p, st.size() * sizeof(int32_t), cudaMemcpyDeviceToHost, h->stream));
    HB_CUDA(cudaStreamSynchronize(h->stream));
    for (size_t g = 0; g < st.size(); ++g) {
        if (!st[g]) continue;
        std::string what;
        if (st[g] & hb::ST_NAN_H) what += " NaN convection coefficients;";
        if (st[g] & hb::ST_NAN_ERR) what += " NaN no-mass error;";
        if (st[g] & hb::ST_NAN_ZONE) what += " Future temperatures is NaN;";
        if (st[g] & hb::ST_ITER_CAP) what += " Excessive number of iterations;";
        if (st[g] & hb::ST_BAD_GAMMA) what += " cavity tilt out of range;";
        // name the surface the way the caller knows it (descriptor order), not only the device slot
        int64_t surf = -1, model = -1;
        for (size_t gd = 0; gd < h->plan.dev_surface_of.size(); ++gd)
            if (h->plan.dev_surface_of[gd] == (int32_t)g) { surf = (int64_t)gd; model = h->plan.surface_model[gd]; break; }
        throw std::domain_error("physics failure on surface " + std::to_string(surf) + " of the batch (model " +
                                std::to_string(model) + ", device surface slot " + std::to_string(g) + "):" + what);
    }
}

hb_status physics_status(hb_handle* h) {
    try {
        check_status(h);
    } catch (const CudaError& e) {
        g_err = e.what();
        return HB_ERR_CUDA;
    } catch (const std::exception& e) {
        g_err = e.what();
        return HB_ERR_PHYSICS;
    }
    return HB_OK;
}

}  // namespace

extern "C" {

const char* hb_last_error(void) { return g_err.c_str(); }
const char* hb_version(void) { return "heat_b200 0.1 (sm_100a)"; }

hb_status hb_create(const hb_batch_desc* desc, int32_t device, hb_handle** out) {
    if (!out) return HB_ERR_INVALID;
    *out = nullptr;
    std::unique_ptr<hb_handle> h(new hb_handle());
    hb_status rc = guarded([&] {
        hb::plan_batch(desc, h->plan, device != HB_DEVICE_NONE);
        if (device == HB_DEVICE_NONE) {  // host-only planning handle: layout queries work, marching does not
            h->device = HB_DEVICE_NONE;
            return;
        }
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count <= 0)
            throw CudaError(std::string("no usable CUDA device (heat_b200 has no CPU fallback): ") +
                            (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
        if (device < 0 || device >= count) throw CudaError("device index out of range");
        h->device = device;
        HB_CUDA(cudaSetDevice(device));
        HB_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
        build_device(h.get());
        preload_kernels(h.get());
    });
    if (rc == HB_OK) *out = h.release();
    return rc;
}

void hb_destroy(hb_handle* h) { delete h; }   // ~hb_handle releases streams, events and device buffers

int32_t hb_n_models(const hb_handle* h) { return (int32_t)h->plan.models.size(); }
int64_t hb_n_surfaces_total(const hb_handle* h) { return (int64_t)h->plan.surfaces.size(); }
int64_t hb_n_zones_total(const hb_handle* h) { return (int64_t)h->plan.zone_volume.size(); }
int64_t hb_state_len_total(const hb_handle* h) { return h->plan.state_len_total; }
int64_t hb_nodes_total(const hb_handle* h) { return h->plan.nodes_total; }
int64_t hb_updates_per_step(const hb_handle* h) { return h->plan.updates_per_step; }
int64_t hb_launch_count(const hb_handle* h) { return h->launches; }

hb_status hb_get_model_info(const hb_handle* h, int32_t model, hb_model_info* out) {
    return guarded([&] {
        if (model < 0 || model >= (int32_t)h->plan.models.size()) throw std::runtime_error("model index out of range");
        const hb::ModelPlan& M = h->plan.models[model];
        out->n_zones = M.n_zones;
        out->n_surfaces = M.n_surfaces;
        out->zone_begin = M.zone_begin;
        out->surface_begin = M.surface_begin;
        out->dt_subdivisions = M.dt_subdivisions;
        out->dt = M.dt;
        out->state_offset = M.state_offset;
        out->state_len = M.state_len;
        out->nodes = M.nodes;
    });
}

hb_status hb_get_surface_info(const hb_handle* h, int64_t s, hb_surface_info* out) {
    return guarded([&] {
        if (s < 0 || s >= (int64_t)h->plan.plans.size()) throw std::runtime_error("surface index out of range");
        const hb::SurfacePlan& P = h->plan.plans[s];
        out->n_nodes = P.n_nodes;
        out->n_layers = (int32_t)P.n_elements.size();
        out->tstep_subdivision = P.tstep_subdivision;
        out->n_massive_chunks = (int32_t)P.massive_chunks.size() / 2;
        out->n_nomass_chunks = (int32_t)P.nomass_chunks.size() / 2;
        out->state_offset = h->plan.surface_state_offset[s];
        out->front_emissivity = P.front_emissivity;
        out->back_emissivity = P.back_emissivity;
        out->wind_speed_modifier = P.wind_speed_modifier;
        out->cos_tilt = P.cos_tilt;
    });
}

hb_status hb_surface_rk4_number(const hb_handle* h, int64_t s, double film, double* out) {
    return guarded([&] {
        if (s < 0 || s >= (int64_t)h->plan.plans.size()) throw std::runtime_error("surface index out of range");
        if (!out || !(film >= 0.)) throw std::runtime_error("hb_surface_rk4_number: film must be >= 0 and out non-NULL");
        *out = hb::rk4_number(h->plan.plans[s], h->plan.models[h->plan.surface_model[s]].dt, film);
    });
}

hb_status hb_get_surface_arrays(const hb_handle* h, int64_t s, int32_t* n_elements, double* mass, double* u,
                                double* front_alpha, double* back_alpha, int32_t* massive_chunks,
                                int32_t* nomass_chunks) {
    return guarded([&] {
        if (s < 0 || s >= (int64_t)h->plan.plans.size()) throw std::runtime_error("surface index out of range");
        const hb::SurfacePlan& P = h->plan.plans[s];
        if (n_elements) std::copy(P.n_elements.begin(), P.n_elements.end(), n_elements);
        if (mass) std::copy(P.mass.begin(), P.mass.end(), mass);
        if (u) std::copy(P.u.begin(), P.u.end(), u);
        if (front_alpha) std::copy(P.front_alpha.begin(), P.front_alpha.end(), front_alpha);
        if (back_alpha) std::copy(P.back_alpha.begin(), P.back_alpha.end(), back_alpha);
        if (massive_chunks) std::copy(P.massive_chunks.begin(), P.massive_chunks.end(), massive_chunks);
        if (nomass_chunks) std::copy(P.nomass_chunks.begin(), P.nomass_chunks.end(), nomass_chunks);
    });
}

// Scatter a reference-layout state vector into the SoA device buffers.
hb_status hb_set_state(hb_handle* h, const double* state) {
    return guarded([&] {
        need_device(h);
        const hb::BatchPlan& B = h->plan;
        // ghost / padding slots are rewritten by the kernels before they are read: a zero-filled slab is fine
        std::vector<double> nodeT(B.node_slab_len, 0.), sc(8 * B.n_dev_surfaces, 0.), zoneT(B.n_dev_zones, 0.);
        hb::scatter_state(B, state, nodeT.data(), sc.data(), zoneT.data());
        HB_CUDA(cudaMemcpyAsync(h->nodeT.p, nodeT.data(), nodeT.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        HB_CUDA(cudaMemcpyAsync(h->sc.p, sc.data(), sc.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        HB_CUDA(cudaMemcpyAsync(h->zoneT.p, zoneT.data(), zoneT.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        clear_status(h);   // a new state starts with a clean record (the reference's error belongs to one march call)
        HB_CUDA(cudaStreamSynchronize(h->stream));
    });
}

using hb::gather_state;

hb_status hb_get_state(hb_handle* h, double* state) {
    return guarded([&] {
        need_device(h);
        const hb::BatchPlan& B = h->plan;
        std::vector<double> nodeT(B.node_slab_len), sc(8 * B.n_dev_surfaces), zoneT(B.n_dev_zones);
        HB_CUDA(cudaMemcpyAsync(nodeT.data(), h->nodeT.p, nodeT.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaMemcpyAsync(sc.data(), h->sc.p, sc.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaMemcpyAsync(zoneT.data(), h->zoneT.p, zoneT.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));
        gather_state(B, nodeT.data(), sc.data(), zoneT.data(), state);
    });
}

hb_status hb_set_zone_inputs(hb_handle* h, const double* q_internal, const double* infiltration_volume,
                             const double* infiltration_temperature, const double* ventilation_volume,
                             const double* ventilation_temperature) {
    return guarded([&] {
        need_device(h);
        const hb::BatchPlan& B = h->plan;
        const int64_t NZ = B.n_dev_zones;
        auto scatter = [&](const double* src, double dflt, DevBuf<double>& dst) {
            std::vector<double> v(NZ, dflt);
            if (src)
                for (size_t z = 0; z < B.dev_zone_of.size(); ++z) v[B.dev_zone_of[z]] = src[z];
            HB_CUDA(cudaMemcpyAsync(dst.p, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
            HB_CUDA(cudaStreamSynchronize(h->stream));
        };
        if ((infiltration_temperature && !infiltration_volume) || (ventilation_temperature && !ventilation_volume))
            throw std::runtime_error("Space has infiltration/ventilation temperature but not volume");
        h->zone_has_inf_v = infiltration_volume != nullptr;
        h->zone_has_vent_v = ventilation_volume != nullptr;
        scatter(q_internal, 0., h->zq);
        scatter(infiltration_volume, 0., h->zinf_v);
        scatter(infiltration_temperature, NAN, h->zinf_t);
        scatter(ventilation_volume, 0., h->zvent_v);
        scatter(ventilation_temperature, NAN, h->zvent_t);
    });
}

// hb_series zone schedules (host, [n_steps][zones of the description]) -> device tables in device zone order.
struct ZoneSchedules {
    DevBuf<double> gain, inf_v, inf_t, vent_v, vent_t;
};
static void upload_zone_schedules(hb_handle* h, const hb_series* s, ZoneSchedules& d, hb::SeriesDev& P) {
    const hb::BatchPlan& B = h->plan;
    P.n_dev_zones = B.n_dev_zones;
    if (s->n_steps <= 0) return;
    if ((s->zone_infiltration_temperature && !s->zone_infiltration_volume && !h->zone_has_inf_v) ||
        (s->zone_ventilation_temperature && !s->zone_ventilation_volume && !h->zone_has_vent_v))
        throw std::runtime_error("Space has infiltration/ventilation temperature but not volume");
    const int64_t nz = (int64_t)B.dev_zone_of.size();
    auto up = [&](const double* src, double pad, DevBuf<double>& dst, const double*& out) {
        if (!src) return;
        std::vector<double> v((size_t)s->n_steps * B.n_dev_zones, pad);
        for (int64_t t = 0; t < s->n_steps; ++t)
            for (int64_t z = 0; z < nz; ++z) v[(size_t)t * B.n_dev_zones + B.dev_zone_of[z]] = src[t * nz + z];
        dst.alloc(v.size());
        HB_CUDA(cudaMemcpyAsync(dst.p, v.data(), v.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));   // `v` goes out of scope
        out = dst.p;
    };
    up(s->zone_gain, 0., d.gain, P.zone_gain);
    up(s->zone_infiltration_volume, 0., d.inf_v, P.zone_inf_v);
    up(s->zone_infiltration_temperature, NAN, d.inf_t, P.zone_inf_t);
    up(s->zone_ventilation_volume, 0., d.vent_v, P.zone_vent_v);
    up(s->zone_ventilation_temperature, NAN, d.vent_t, P.zone_vent_t);
}

hb_status hb_march_series(hb_handle* h, const hb_series* s) {
    hb_status rc = guarded([&] {
        need_device(h);
        validate_series(h, s);
        if (s->n_steps == 0) return;
        update_sites(h, s);
        clear_status(h);
        const hb::BatchPlan& B = h->plan;
        DevBuf<double> d_weather, d_inputs, d_zone_rec, d_node_rec, d_sc_rec;
        ZoneSchedules d_sched;
        const int64_t wcount = (int64_t)s->n_sites * s->weather_rows * 3;
        d_weather.alloc(wcount);
        HB_CUDA(cudaMemcpyAsync(d_weather.p, s->weather, wcount * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        const int64_t icount = input_count(h, s);
        if (icount) {
            d_inputs.alloc(icount);
            HB_CUDA(cudaMemcpyAsync(d_inputs.p, s->inputs, icount * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        }
        const int n_rec = s->record_every > 0 ? s->n_steps / s->record_every : 0;
        hb::SeriesDev P{};
        P.n_steps = s->n_steps;
        P.weather_mode = s->weather_mode;
        P.input_mode = s->input_mode;
        P.ir_mode = s->ir_mode;
        P.weather = d_weather.p;
        P.weather_rows = s->weather_rows;
        P.inputs = d_inputs.p;
        P.n_desc_surfaces = (int64_t)B.surfaces.size();
        P.record_every = n_rec > 0 && (s->zone_t_record || s->state_record) ? s->record_every : 0;
        P.slab_len = B.node_slab_len;
        upload_zone_schedules(h, s, d_sched, P);
        if (P.record_every) {
            d_zone_rec.alloc((size_t)n_rec * B.n_dev_zones);
            P.zone_rec = d_zone_rec.p;
            if (s->state_record) {
                d_node_rec.alloc((size_t)n_rec * B.node_slab_len);
                d_sc_rec.alloc((size_t)n_rec * 8 * B.n_dev_surfaces);
                P.node_rec = d_node_rec.p;
                P.sc_rec = d_sc_rec.p;
            }
        }
        launch_series(h, P);
        if (P.record_every) {
            std::vector<double> zr((size_t)n_rec * B.n_dev_zones);
            HB_CUDA(cudaMemcpyAsync(zr.data(), d_zone_rec.p, zr.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            std::vector<double> nr, sr;
            if (s->state_record) {
                nr.resize((size_t)n_rec * B.node_slab_len);
                sr.resize((size_t)n_rec * 8 * B.n_dev_surfaces);
                HB_CUDA(cudaMemcpyAsync(nr.data(), d_node_rec.p, nr.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
                HB_CUDA(cudaMemcpyAsync(sr.data(), d_sc_rec.p, sr.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
            }
            HB_CUDA(cudaStreamSynchronize(h->stream));
            const int64_t nz_total = (int64_t)B.zone_volume.size();
            for (int r = 0; r < n_rec; ++r) {
                const double* z = zr.data() + (size_t)r * B.n_dev_zones;
                if (s->zone_t_record)
                    for (int64_t k = 0; k < nz_total; ++k) s->zone_t_record[(int64_t)r * nz_total + k] = z[B.dev_zone_of[k]];
                if (s->state_record)
                    gather_state(B, nr.data() + (size_t)r * B.node_slab_len, sr.data() + (size_t)r * 8 * B.n_dev_surfaces, z,
                                 s->state_record + (int64_t)r * B.state_len_total);
            }
        }
        HB_CUDA(cudaStreamSynchronize(h->stream));
        HB_CUDA(cudaGetLastError());
    });
    if (rc != HB_OK) return rc;
    return physics_status(h);
}

hb_status hb_read_epw(const char* path, double* weather, double* inputs, int64_t capacity, int64_t* n_rows) {
    return guarded([&] {
        if (!path || !n_rows || capacity < 0) throw std::invalid_argument("hb_read_epw: path, n_rows and capacity >= 0 are required");
        *n_rows = hb::read_epw(path, weather, inputs, capacity);
    });
}

hb_status hb_upload(hb_handle* h, const double* host, int64_t count, double** dev_out) {
    return guarded([&] {
        need_device(h);
        double* p = nullptr;
        HB_CUDA(cudaMalloc(&p, (size_t)std::max<int64_t>(count, 1) * sizeof(double)));
        if (count) HB_CUDA(cudaMemcpyAsync(p, host, (size_t)count * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));
        *dev_out = p;
    });
}

hb_status hb_free_dev(hb_handle* h, double* dev) {
    return guarded([&] {
        need_device(h);
        if (dev) HB_CUDA(cudaFree(dev));
    });
}

hb_status hb_march_series_dev(hb_handle* h, const hb_series* s, float* elapsed_ms) {
    hb_status rc = guarded([&] {
        need_device(h);
        validate_series(h, s);
        update_sites(h, s);
        clear_status(h);
        hb::SeriesDev P{};
        P.n_steps = s->n_steps;
        P.weather_mode = s->weather_mode;
        P.input_mode = s->input_mode;
        P.ir_mode = s->ir_mode;
        P.weather = s->weather;
        P.weather_rows = s->weather_rows;
        P.inputs = s->inputs;
        P.n_desc_surfaces = (int64_t)h->plan.surfaces.size();
        P.record_every = 0;
        P.slab_len = h->plan.node_slab_len;
        ZoneSchedules d_sched;
        upload_zone_schedules(h, s, d_sched, P);
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (elapsed_ms) {
            HB_CUDA(cudaEventCreate(&e0));
            HB_CUDA(cudaEventCreate(&e1));
            HB_CUDA(cudaEventRecord(e0, h->stream));
        }
        if (s->n_steps > 0) launch_series(h, P);
        if (elapsed_ms) {
            HB_CUDA(cudaEventRecord(e1, h->stream));
            HB_CUDA(cudaEventSynchronize(e1));
            HB_CUDA(cudaEventElapsedTime(elapsed_ms, e0, e1));
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        } else {
            HB_CUDA(cudaStreamSynchronize(h->stream));
        }
        HB_CUDA(cudaGetLastError());
    });
    if (rc != HB_OK) return rc;
    return physics_status(h);   // after the timed events: a poisoned run must not return HB_OK
}

hb_status hb_reduce(hb_handle* h, hb_ensemble_stats* out) {
    return guarded([&] {
        need_device(h);
        const double init[8] = {0., INFINITY, -INFINITY, 0., INFINITY, -INFINITY, 0., 0.};
        HB_CUDA(cudaMemcpyAsync(h->red_out.p, init, sizeof(init), cudaMemcpyHostToDevice, h->stream));
        HB_CUDA(cudaMemsetAsync(h->red_iout.p, 0, 4 * sizeof(unsigned long long), h->stream));
        const int n_tiles = (int)h->plan.tiles.size();
        const int grid = std::min(n_tiles, 148 * 8);
        hb::launch_reduce(grid, h->stream, h->view(), n_tiles, h->red_out.p, h->red_iout.p);
        HB_CUDA(cudaGetLastError());
        h->launches += 1;
        double o[8];
        unsigned long long io[4];
        HB_CUDA(cudaMemcpyAsync(o, h->red_out.p, sizeof(o), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaMemcpyAsync(io, h->red_iout.p, sizeof(io), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));
        out->sum_zone_t = o[0]; out->min_zone_t = o[1]; out->max_zone_t = o[2];
        out->sum_node_t = o[3]; out->min_node_t = o[4]; out->max_node_t = o[5];
        out->front_conv_energy_j = o[6]; out->back_conv_energy_j = o[7];
        out->n_zones = (int64_t)io[0]; out->n_nodes = (int64_t)io[1];
        out->nomass_iterations = (int64_t)io[2]; out->status_flags = (int64_t)io[3];
    });
}

hb_status hb_comm_unique_id(void* id_out) {
    return guarded([&] {
        if (!id_out) throw std::invalid_argument("hb_comm_unique_id: null buffer");
        Nccl& n = Nccl::get();
        Nccl::UniqueId id;
        n.check(n.GetUniqueId(&id), "ncclGetUniqueId");
        static_assert(sizeof(id) == HB_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
        std::memcpy(id_out, &id, sizeof(id));
    });
}

hb_status hb_comm_init(hb_handle* h, const void* id, int32_t rank, int32_t world) {
    return guarded([&] {
        need_device(h);
        if (!id || world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("hb_comm_init: bad id / rank / world");
        Nccl& n = Nccl::get();
        if (h->comm) {
            n.CommDestroy(h->comm);
            h->comm = nullptr;
        }
        Nccl::UniqueId uid;
        std::memcpy(&uid, id, sizeof(uid));
        n.check(n.CommInitRank(&h->comm, world, uid, rank), "ncclCommInitRank");
        h->comm_world = world;
        h->comm_buf.alloc(7 + 2 + 10);
    });
}

hb_status hb_comm_destroy(hb_handle* h) {
    return guarded([&] {
        if (h->comm) {
            need_device(h);
            Nccl::get().CommDestroy(h->comm);
            h->comm = nullptr;
            h->comm_world = 1;
        }
    });
}

hb_status hb_reduce_all(hb_handle* h, hb_ensemble_stats* out) {
    hb_status rc = hb_reduce(h, out);
    if (rc != HB_OK || !h->comm) return rc;   // (a one-rank communicator still goes through NCCL)
    return guarded([&] {
        Nccl& n = Nccl::get();
        constexpr int kBits = 8;   // status words: ST_* of hb_dev.hpp, OR-ed as a maximum per bit
        double v[7 + 2 + 2 + kBits] = {out->sum_zone_t, out->sum_node_t, out->front_conv_energy_j, out->back_conv_energy_j,
                                       (double)out->n_zones, (double)out->n_nodes, (double)out->nomass_iterations,
                                       out->min_zone_t, out->min_node_t, out->max_zone_t, out->max_node_t};
        for (int k = 0; k < kBits; ++k) v[11 + k] = (double)((out->status_flags >> k) & 1);
        double* d = h->comm_buf.p;
        HB_CUDA(cudaMemcpyAsync(d, v, sizeof(v), cudaMemcpyHostToDevice, h->stream));
        n.check(n.AllReduce(d, d, 7, Nccl::kFloat64, Nccl::kSum, h->comm, h->stream), "ncclAllReduce(sum)");
        n.check(n.AllReduce(d + 7, d + 7, 2, Nccl::kFloat64, Nccl::kMin, h->comm, h->stream), "ncclAllReduce(min)");
        n.check(n.AllReduce(d + 9, d + 9, 2 + kBits, Nccl::kFloat64, Nccl::kMax, h->comm, h->stream), "ncclAllReduce(max)");
        HB_CUDA(cudaMemcpyAsync(v, d, sizeof(v), cudaMemcpyDeviceToHost, h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));
        out->sum_zone_t = v[0]; out->sum_node_t = v[1]; out->front_conv_energy_j = v[2]; out->back_conv_energy_j = v[3];
        out->n_zones = (int64_t)std::llround(v[4]); out->n_nodes = (int64_t)std::llround(v[5]);
        out->nomass_iterations = (int64_t)std::llround(v[6]);
        out->min_zone_t = v[7]; out->min_node_t = v[8]; out->max_zone_t = v[9]; out->max_node_t = v[10];
        int64_t flags = 0;
        for (int k = 0; k < kBits; ++k)
            if (v[11 + k] > 0.5) flags |= (int64_t)1 << k;
        out->status_flags = flags;
    });
}

hb_status hb_reset_energy(hb_handle* h) {
    return guarded([&] {
        need_device(h);
        HB_CUDA(cudaMemsetAsync(h->energy_f.p, 0, h->energy_f.n * sizeof(double), h->stream));
        HB_CUDA(cudaMemsetAsync(h->energy_b.p, 0, h->energy_b.n * sizeof(double), h->stream));
        HB_CUDA(cudaMemsetAsync(h->nm_iters.p, 0, h->nm_iters.n * sizeof(unsigned long long), h->stream));
        HB_CUDA(cudaStreamSynchronize(h->stream));
    });
}

hb_status hb_measure_fp64_peak(int32_t device, double* tflops) {
    return guarded([&] {
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) throw CudaError("no usable CUDA device");
        HB_CUDA(cudaSetDevice(device));
        cudaDeviceProp prop;
        HB_CUDA(cudaGetDeviceProperties(&prop, device));
        const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 1 << 15;
        double* out = nullptr;
        HB_CUDA(cudaMalloc(&out, (size_t)blocks * threads * sizeof(double)));
        cudaEvent_t e0, e1;
        HB_CUDA(cudaEventCreate(&e0));
        HB_CUDA(cudaEventCreate(&e1));
        double best = 0.;
        for (int rep = 0; rep < 5; ++rep) {
            HB_CUDA(cudaEventRecord(e0));
            hb::launch_dfma_probe(blocks, threads, out, iters, 1.0 + rep);
            HB_CUDA(cudaEventRecord(e1));
            HB_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            HB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            const double flops = 2.0 * 8.0 * (double)iters * (double)blocks * threads;
            if (rep > 0) best = std::max(best, flops / (ms * 1e-3) / 1e12);
        }
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(out);
        *tflops = best;
    });
}

hb_status hb_flush_l2(hb_handle* h, int64_t bytes) {
    return guarded([&] {
        need_device(h);
        const size_t n = (size_t)bytes / sizeof(double);
        if (h->flush.n < n) h->flush.alloc(n);
        hb::launch_fill(h->stream, h->flush.p, (int64_t)n, 1.0);
        HB_CUDA(cudaGetLastError());
        HB_CUDA(cudaStreamSynchronize(h->stream));
    });
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// hb_reduce_noalloc — the K3 reduction for callers that cannot allocate (CUDA graph capture, serving loops).
// No counterpart in the reference (see include/heat_b200.h).
//
// Variant of hb_reduce_kernel (hb_kernels.cuh) without atomics and without the handle's accumulator buffers: stage 1
// writes one 12-word partial per tile into a file-scope device array, stage 2 (one block) folds the partials in tile
// order — the sums are the same from call to call.  The array lives in every GPU that loads the library, until exit.
// ------------------------------------------------------------------------------------------------------------------
namespace {

constexpr int kNaWords = 12;                 // 8 doubles (sums / extrema / energies) + 4 integers per tile
constexpr int kNaBlock = 128;                // threads of a stage-1 block: one block per tile (G·W <= 128 lanes)
constexpr int kNaFoldBlock = 1024;
static_assert(sizeof(hb_ensemble_stats) == kNaWords * 8, "hb_ensemble_stats is 8 doubles + 4 int64");

// [HB_REDUCE_NOALLOC_MAX_TILES][12] partials, then the folded hb_ensemble_stats.  In HOST code this name is only a
// stand-in for the symbol — never a device address: hb_reduce_noalloc asks cudaGetSymbolAddress once per device.
static __device__ __align__(256) unsigned long long hb_na_scratch[((size_t)HB_REDUCE_NOALLOC_MAX_TILES + 1) * kNaWords];

__global__ void __launch_bounds__(kNaBlock) hb_na_tile_kernel(hb::DevView V, int n_tiles, unsigned long long* scratch) {
    __shared__ unsigned long long s_part[kNaBlock / 32][kNaWords];
    const int ti = blockIdx.x, t = threadIdx.x;
    if (ti >= n_tiles) return;
    double sz = 0., mnz = CUDART_INF, mxz = -CUDART_INF, sn = 0., mnn = CUDART_INF, mxn = -CUDART_INF, ef = 0., eb = 0.;
    unsigned long long nz = 0, nn = 0, it = 0, st = 0;
    const hb::Tile tile = V.tiles[ti];
    const int TPB = tile.G * tile.W;
    for (int th = t; th < TPB; th += kNaBlock) {
        const int wl = th / tile.G, part = th % tile.G;
        const int n = V.n_nodes[(int64_t)tile.surface_begin + wl];
        for (int i = 0; i < tile.NL; ++i) {
            const int slot = part * tile.NL + i;
            // block tiles: slots 1..n hold the nodes (0 and n+1 are ghosts); warp tiles: slots 0..n-1
            if ((tile.flags & hb::TILE_WARP) ? slot < n : (slot >= 1 && slot <= n)) {
                const double v = V.nodeT[tile.node_offset + (int64_t)i * TPB + th];
                sn += v;
                mnn = fmin(mnn, v);
                mxn = fmax(mxn, v);
                nn += 1;
            }
        }
    }
    if (t < tile.W) {
        const int64_t g = (int64_t)tile.surface_begin + t;
        if (V.n_nodes[g] > 0) {
            ef += V.energy_f[g];
            eb += V.energy_b[g];
            it += V.nm_iters[g];
            st |= (unsigned long long)V.status[g];
        }
        if (t < tile.n_zones) {
            const double v = V.zoneT[(int64_t)tile.zone_begin + t];
            sz += v;
            mnz = fmin(mnz, v);
            mxz = fmax(mxz, v);
            nz += 1;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        sz += __shfl_down_sync(0xffffffffu, sz, o);
        sn += __shfl_down_sync(0xffffffffu, sn, o);
        ef += __shfl_down_sync(0xffffffffu, ef, o);
        eb += __shfl_down_sync(0xffffffffu, eb, o);
        mnz = fmin(mnz, __shfl_down_sync(0xffffffffu, mnz, o));
        mxz = fmax(mxz, __shfl_down_sync(0xffffffffu, mxz, o));
        mnn = fmin(mnn, __shfl_down_sync(0xffffffffu, mnn, o));
        mxn = fmax(mxn, __shfl_down_sync(0xffffffffu, mxn, o));
        nz += __shfl_down_sync(0xffffffffu, nz, o);
        nn += __shfl_down_sync(0xffffffffu, nn, o);
        it += __shfl_down_sync(0xffffffffu, it, o);
        st |= __shfl_down_sync(0xffffffffu, st, o);
    }
    if ((t & 31) == 0) {
        unsigned long long* p = s_part[t >> 5];
        p[0] = __double_as_longlong(sz); p[1] = __double_as_longlong(mnz); p[2] = __double_as_longlong(mxz);
        p[3] = __double_as_longlong(sn); p[4] = __double_as_longlong(mnn); p[5] = __double_as_longlong(mxn);
        p[6] = __double_as_longlong(ef); p[7] = __double_as_longlong(eb);
        p[8] = nz; p[9] = nn; p[10] = it; p[11] = st;
    }
    __syncthreads();
    if (t == 0) {   // the block's warps in order: every word of the tile's partial is (re)written, nothing is accumulated
        for (int w = 1; w < kNaBlock / 32; ++w) {
            const unsigned long long* p = s_part[w];
            sz += __longlong_as_double(p[0]); mnz = fmin(mnz, __longlong_as_double(p[1])); mxz = fmax(mxz, __longlong_as_double(p[2]));
            sn += __longlong_as_double(p[3]); mnn = fmin(mnn, __longlong_as_double(p[4])); mxn = fmax(mxn, __longlong_as_double(p[5]));
            ef += __longlong_as_double(p[6]); eb += __longlong_as_double(p[7]);
            nz += p[8]; nn += p[9]; it += p[10]; st |= p[11];
        }
        unsigned long long* o = scratch + (size_t)ti * kNaWords;
        o[0] = __double_as_longlong(sz); o[1] = __double_as_longlong(mnz); o[2] = __double_as_longlong(mxz);
        o[3] = __double_as_longlong(sn); o[4] = __double_as_longlong(mnn); o[5] = __double_as_longlong(mxn);
        o[6] = __double_as_longlong(ef); o[7] = __double_as_longlong(eb);
        o[8] = nz; o[9] = nn; o[10] = it; o[11] = st;
    }
}

// word k of the statistics: 0,3,6,7 sums; 1,4 minima; 2,5 maxima; 8-10 integer sums; 11 OR
__device__ __forceinline__ unsigned long long na_fold(int k, unsigned long long a, unsigned long long b) {
    if (k >= 8) return k == 11 ? (a | b) : a + b;
    const double x = __longlong_as_double(a), y = __longlong_as_double(b);
    return __double_as_longlong(k == 1 || k == 4 ? fmin(x, y) : k == 2 || k == 5 ? fmax(x, y) : x + y);
}

__global__ void __launch_bounds__(kNaFoldBlock) hb_na_fold_kernel(int n_tiles, unsigned long long* scratch,
                                                                   hb_ensemble_stats* out_dev) {
    __shared__ unsigned long long s_acc[kNaFoldBlock / 32][kNaWords];
    const int t = threadIdx.x;
    const unsigned long long inf = __double_as_longlong(CUDART_INF), ninf = __double_as_longlong(-CUDART_INF);
    unsigned long long acc[kNaWords];
#pragma unroll
    for (int k = 0; k < kNaWords; ++k) acc[k] = k == 1 || k == 4 ? inf : k == 2 || k == 5 ? ninf : 0ull;   // 0.0 is all-zero bits
    for (int ti = t; ti < n_tiles; ti += kNaFoldBlock) {
#pragma unroll
        for (int k = 0; k < kNaWords; ++k) acc[k] = na_fold(k, acc[k], scratch[(size_t)ti * kNaWords + k]);
    }
#pragma unroll
    for (int k = 0; k < kNaWords; ++k)
        for (int o = 16; o > 0; o >>= 1) acc[k] = na_fold(k, acc[k], __shfl_down_sync(0xffffffffu, acc[k], o));
    if ((t & 31) == 0)
        for (int k = 0; k < kNaWords; ++k) s_acc[t >> 5][k] = acc[k];
    __syncthreads();
    if (t < kNaWords) {
        unsigned long long v = s_acc[0][t];
        for (int w = 1; w < kNaFoldBlock / 32; ++w) v = na_fold(t, v, s_acc[w][t]);
        scratch[(size_t)HB_REDUCE_NOALLOC_MAX_TILES * kNaWords + t] = v;
        if (out_dev) reinterpret_cast<unsigned long long*>(out_dev)[t] = v;   // same 12-word layout as hb_ensemble_stats
    }
}

// device address of hb_na_scratch, per device: resolved once, with that device current
unsigned long long* na_scratch_of(int device) {
    constexpr int kMaxDevices = 64;
    static unsigned long long* ptr[kMaxDevices] = {};
    static std::mutex mu;
    if (device < 0 || device >= kMaxDevices) throw std::invalid_argument("hb_reduce_noalloc: device index beyond 63");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        HB_CUDA(cudaGetSymbolAddress(&p, hb_na_scratch));   // the handle's device is current (need_device)
        ptr[device] = static_cast<unsigned long long*>(p);
    }
    return ptr[device];
}

}  // namespace

extern "C" hb_status hb_reduce_noalloc(hb_handle* h, hb_ensemble_stats* out_host, hb_ensemble_stats* out_dev) {
    return guarded([&] {
        need_device(h);
        if (!out_host && !out_dev) throw std::invalid_argument("hb_reduce_noalloc: both outputs are NULL");
        const int64_t n_tiles = (int64_t)h->plan.tiles.size();
        if (n_tiles > HB_REDUCE_NOALLOC_MAX_TILES)
            throw std::invalid_argument("hb_reduce_noalloc: the batch has " + std::to_string(n_tiles) +
                                        " tiles, the static scratch holds HB_REDUCE_NOALLOC_MAX_TILES = " +
                                        std::to_string(HB_REDUCE_NOALLOC_MAX_TILES) + "; use hb_reduce");
        unsigned long long* scratch = na_scratch_of(h->device);
        if (n_tiles > 0) hb_na_tile_kernel<<<(int)n_tiles, kNaBlock, 0, h->stream>>>(h->view(), (int)n_tiles, scratch);
        hb_na_fold_kernel<<<1, kNaFoldBlock, 0, h->stream>>>((int)n_tiles, scratch, out_dev);
        HB_CUDA(cudaGetLastError());
        h->launches += n_tiles > 0 ? 2 : 1;
        if (out_host) {
            HB_CUDA(cudaMemcpyAsync(out_host, scratch + (size_t)HB_REDUCE_NOALLOC_MAX_TILES * kNaWords, sizeof(*out_host),
                                    cudaMemcpyDeviceToHost, h->stream));
            HB_CUDA(cudaStreamSynchronize(h->stream));
        }
    });
}
