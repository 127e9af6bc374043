// peer.cuh -- the row-band exchanges over NVLink peer memory (multi_api.cuh drives it).
//
// Every rank of a band group owns one WINDOW in its HBM and maps the windows of all the others (cudaIpc handles between
// processes, plain peer access between the threads of one process).  The three exchanges of a band call are then stores into
// the peers' windows by this library's own kernels -- no NCCL on the data path:
//
//   exchange 1+2  peer_exchange_kernel: after the ring search a rank stores its maximum into every peer's mailbox and its first /
//                 last three raw rows into the neighbours' halo zones, raises its flag in every window (release, system scope),
//                 waits for the n flags in its own, folds the maxima (depth_image.rs:124) and moves the halo rows next to
//                 its band for the Gaussian (:133).  One launch where NCCL needed an all-reduce and a send/recv group.
//   exchange 4    the depth band is stored into every rank's assembled map BY THE KERNEL THAT PRODUCES IT (finish_horz_kernel's
//                 mirror stores): the all-gather rides on the finish pass tile by tile instead of following it; a flag round
//                 (peer_barrier_kernel) tells every rank that all bands have landed.
//
// Window layout (identical on every rank; all offsets from the base):
//   [0, 4096)   control: flags[PW_PHASES][8] u32 (sequence numbers, monotone), maxima[2][8] u32, a CTA counter
//   halo zones  [2 (parity of the call)][2 (from above, from below)][3 rows]
//   d8h         the band's raw distances with three halo rows above and below (the search writes here)
//   full        the assembled depth map (when a caller asks for it)
// Mailboxes and halo zones are double-buffered by the parity of the call number: a rank can enter call k+1 while a slow
// peer still reads what call k left, but not call k+2 (its exchange of call k+1 waits for every rank's flag of call k+1).
#pragma once
#include "common.cuh"

namespace stepth {

constexpr int PW_MAX_RANKS = 8;
constexpr int PW_PHASES = 2;             // 0: maxima + halos, 1: assembled map complete
constexpr size_t PW_OFF_FLAGS = 0;       // [PW_PHASES][8] u32
constexpr size_t PW_OFF_MAX = 256;       // [2][8] u32
constexpr size_t PW_OFF_COUNT = 512;     // u32: CTAs of the exchange kernel that have finished their stores
constexpr size_t PW_OFF_HALO = 4096;

struct PeerView {
    uint8_t *base[PW_MAX_RANKS];   // every rank's window as seen from this device; base[rank] is its own
    int n, rank;
};

#if defined(__CUDACC__)

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
constexpr unsigned long long PW_TIMEOUT_NS = 20ull * 1000 * 1000 * 1000;   // a rank that never arrives must not hang the others' GPUs

// waits until *flag has reached seq; false (and *timed_out = 1) when that takes longer than PW_TIMEOUT_NS
__device__ __forceinline__ bool wait_flag(const uint32_t *flag, uint32_t seq, uint32_t *timed_out) {
    const unsigned long long t0 = global_ns();
    while ((int32_t)(ld_acquire_sys(flag) - seq) < 0) {
        __nanosleep(64);
        if (global_ns() - t0 > PW_TIMEOUT_NS) {
            *timed_out = 1u;
            return false;
        }
    }
    return true;
}

// bytes from src to dst by the whole grid (16-byte words when both are aligned)
__device__ __forceinline__ void grid_copy(uint8_t *dst, const uint8_t *src, size_t bytes, uint32_t tid, uint32_t nthreads) {
    if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src) | bytes) & 15) == 0) {
        const uint4 *s = reinterpret_cast<const uint4 *>(src);
        uint4 *d = reinterpret_cast<uint4 *>(dst);
        for (size_t i = tid; i < bytes / 16; i += nthreads) d[i] = s[i];
    } else {
        for (size_t i = tid; i < bytes; i += nthreads) dst[i] = src[i];
    }
}

struct PeerExchangeArgs {
    PeerView pv;
    uint32_t seq;          // number of this band call (the same on every rank), > 0
    size_t halo_zone;      // bytes of one halo zone (3 rows, rounded up to 16)
    size_t halo_bytes;     // 3 * w
    uint8_t *band;         // this rank's raw distances, row 0 of the band (inside its own window)
    uint32_t rows, w;
    uint32_t *max_io;      // in: the band's maximum; out: the maximum over all bands
    uint32_t *timed_out;   // set to 1 when a peer's flag did not arrive in PW_TIMEOUT_NS
};

// exchanges 1 + 2 (see the file comment).  Any small grid; the CTA that finishes its stores last does the waiting.
__global__ void __launch_bounds__(256) peer_exchange_kernel(PeerExchangeArgs a) {
    __shared__ uint32_t s_last;
    const int n = a.pv.n, r = a.pv.rank;
    const uint32_t par = a.seq & 1u;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    uint8_t *own = a.pv.base[r];
    // my first three rows are the rows BELOW the band of rank r - 1, my last three the rows ABOVE the band of rank r + 1
    if (r > 0) grid_copy(a.pv.base[r - 1] + PW_OFF_HALO + (par * 2 + 1) * a.halo_zone, a.band, a.halo_bytes, tid, nth);
    if (r < n - 1)
        grid_copy(a.pv.base[r + 1] + PW_OFF_HALO + (par * 2 + 0) * a.halo_zone, a.band + (size_t)(a.rows - 3) * a.w, a.halo_bytes, tid, nth);
    if (tid < (uint32_t)n) reinterpret_cast<uint32_t *>(a.pv.base[tid] + PW_OFF_MAX)[par * 8 + r] = *a.max_io;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t done = atomicAdd(reinterpret_cast<uint32_t *>(own + PW_OFF_COUNT), 1u);
        s_last = done == gridDim.x - 1;
    }
    __syncthreads();
    if (!s_last) return;
    // every store of this rank has been issued and fenced: raise the flag everywhere, then wait for everybody's
    if (threadIdx.x < (uint32_t)n) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(a.pv.base[threadIdx.x] + PW_OFF_FLAGS) + r, a.seq);
        wait_flag(reinterpret_cast<const uint32_t *>(own + PW_OFF_FLAGS) + threadIdx.x, a.seq, a.timed_out);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        *reinterpret_cast<uint32_t *>(own + PW_OFF_COUNT) = 0;   // for the next call (ordered by the stream)
        const volatile uint32_t *mx = reinterpret_cast<const volatile uint32_t *>(own + PW_OFF_MAX) + par * 8;
        uint32_t m = 0;
        for (int q = 0; q < n; ++q) m = max(m, mx[q]);
        *a.max_io = m;
    }
    // the halo rows the neighbours left in my zones move next to the band: above it (rows -3 .. -1), below it (rows .. rows + 2)
    if (r > 0) grid_copy(a.band - a.halo_bytes, own + PW_OFF_HALO + (par * 2 + 0) * a.halo_zone, a.halo_bytes, threadIdx.x, blockDim.x);
    if (r < n - 1)
        grid_copy(a.band + (size_t)a.rows * a.w, own + PW_OFF_HALO + (par * 2 + 1) * a.halo_zone, a.halo_bytes, threadIdx.x, blockDim.x);
}

// a flag round: returns when every rank has reached it (so everything every rank stored before it has landed)
__global__ void __launch_bounds__(32) peer_barrier_kernel(PeerView pv, int phase, uint32_t seq, uint32_t *timed_out) {
    const int r = pv.rank;
    if (threadIdx.x < (uint32_t)pv.n) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(pv.base[threadIdx.x] + PW_OFF_FLAGS) + phase * 8 + r, seq);
        wait_flag(reinterpret_cast<const uint32_t *>(pv.base[r] + PW_OFF_FLAGS) + phase * 8 + threadIdx.x, seq, timed_out);
    }
}

// the band into every rank's assembled map (fallback when the finish pass that ran could not mirror its stores)
__global__ void __launch_bounds__(256) peer_scatter_kernel(PeerView pv, size_t off_full_band, const uint8_t *band, size_t bytes) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (int q = 0; q < pv.n; ++q) grid_copy(pv.base[q] + off_full_band, band, bytes, tid, nth);
}

#endif  // __CUDACC__

}  // namespace stepth
