// multi.cuh -- the multi-GPU part of the boundary, below the C ABI:
//
//   * stepth_comm: ONE rank of a row-band group = a context + an NCCL communicator.  `stepth_cuda_dev_depth_dense_band`
//     enqueues this rank's share of ONE giant pair: ring search of its band, all-reduce MAX of the normaliser
//     (depth_image.rs:124), 3-row halo exchange of the raw distance plane for the unit-scale Gaussian (:133) across the seams,
//     finish, all-reduce SUM of the 256-bin histogram + depth_split + mask/apply when the tail is fused (:162-245), and an
//     all-gather of the depth bands when the assembled map is wanted on every GPU.  One process per GPU (bench.py under
//     torchrun, the ranks share a ncclUniqueId) or one thread per GPU (stepth_multi, below) drive the same function.
//   * stepth_multi: ONE process driving several GPUs -- what a Rust host links: a context per device, a host thread per
//     device per call.  Pair batches are dealt to the devices with no collective at all (configs[3]); one giant pair runs in
//     row bands (configs[4]) with the bands going straight back into the caller's buffer.
//
// NCCL is resolved at run time (dlopen "libnccl.so.2": the copy already loaded by the process, torch's, when there is one):
// a box without NCCL still loads the library and every single-GPU entry point; the multi-GPU ones then fail loudly.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <thread>

#include <unistd.h>

#include "ctx.cuh"
#include "peer.cuh"

#define STEPTH_ERR_NCCL 101

namespace stepth {

struct NcclApi {
    void *lib = nullptr;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclBroadcast) Broadcast = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    std::string why;
};

inline NcclApi &nccl_api() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            api.lib = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (api.lib) break;
        }
        if (!api.lib) {
            api.why = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : "?");
            return;
        }
        bool ok = true;
        auto sym = [&](const char *n) {
            void *p = dlsym(api.lib, n);
            if (!p) {
                ok = false;
                api.why = std::string("NCCL symbol missing: ") + n;
            }
            return p;
        };
        api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
        api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
        api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
        api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
        api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
        api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
        api.Send = (decltype(api.Send))sym("ncclSend");
        api.Recv = (decltype(api.Recv))sym("ncclRecv");
        api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
        if (!ok) {
            dlclose(api.lib);
            api.lib = nullptr;
        }
    });
    return api;
}

}  // namespace stepth

// one rank of a row-band group
struct stepth_comm {
    stepth_ctx *ctx = nullptr;
    ncclComm_t nccl = nullptr;   // null when n_ranks == 1
    int n_ranks = 1, rank = 0;
    bool owns_nccl = true;
    // scratch of the band path (device): raw distances of the band with 3 halo rows above and below, small state
    stepth_ctx::Buf d8h, small;
    void *h_small = nullptr;     // pinned mirror of `small`
    cudaEvent_t marks[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // start | search | exchanges | finish | tail + gather
    // peer window (peer.cuh): the exchanges as stores into the other ranks' HBM over NVLink
    int peer_mode = -1;               // -1 not tried yet, 0 = not available (the exchanges go through NCCL), 1 = windows mapped
    void *win = nullptr;              // this rank's window
    size_t win_bytes = 0;
    uint8_t *peer_base[8] = {};       // every rank's window as seen from this device ([rank] = win)
    bool peer_opened[8] = {};         // mapped through a cudaIpc handle (another process): closed before the owner frees it
    stepth_ctx::Buf xchg;             // bootstrap scratch: handles in, handles out
    uint32_t seq = 0;                 // band calls so far (the same on every rank)
    size_t off_d8h = 0, off_full = 0, halo_zone = 0;
    bool win_has_full = false;
    // small state layout (u32 words): [0] max  [1] status  [2] n_out  [3] pad  [4..260) hist  then 512 bytes of ranges
    static constexpr size_t SMALL_BYTES = 4 * 260 + 512;
};

namespace stepth {

constexpr uint32_t BAND_HALO = 3;   // rows of the unit-scale Gaussian's support on each side (image 0.23.14: support 3.0)

#define STEPTH_NCCL_TRY(ctx, expr)                                                                                     \
    do {                                                                                                               \
        ncclResult_t r__ = (expr);                                                                                     \
        if (r__ != ncclSuccess)                                                                                        \
            return stepth::set_err((ctx), STEPTH_ERR_NCCL, "%s failed: %s (%s:%d)", #expr, nccl_api().GetErrorString(r__), \
                                   __FILE__, __LINE__);                                                                \
    } while (0)

// main rows [y0, y1) of rank r of n (multiples of 2 rows: the two rows that share an origin row stay together) and the
// additional rows [ay0, ay1) they can reach: origin rows (y+1)/2, widened by max_rings - 1 on both sides
inline void band_plan(uint32_t h, uint32_t add_h, uint32_t max_rings, uint32_t n, uint32_t r, uint32_t out[4]) {
    const uint64_t units = ((uint64_t)h + 1) / 2;
    const uint64_t u0 = units * r / n, u1 = units * (r + 1) / n;
    const uint32_t y0 = (uint32_t)std::min<uint64_t>(h, u0 * 2), y1 = (uint32_t)std::min<uint64_t>(h, u1 * 2);
    out[0] = y0;
    out[1] = y1 > y0 ? y1 : y0;
    out[2] = out[3] = 0;
    if (y1 <= y0) return;
    const int64_t o0 = ((int64_t)y0 + 1) / 2, o1 = (int64_t)y1 / 2;
    const int64_t a0 = std::max<int64_t>(0, o0 - ((int64_t)max_rings - 1));
    const int64_t a1 = std::min<int64_t>(add_h, o1 + ((int64_t)max_rings - 1) + 1);
    out[2] = (uint32_t)a0;
    out[3] = (uint32_t)std::max<int64_t>(a0, a1);
}

}  // namespace stepth
