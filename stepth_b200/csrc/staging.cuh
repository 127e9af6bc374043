// staging.cuh -- host buffers that are NOT page-locked, moved through a pinned ring.
//
// The drop-in caller is Rust: its images are `Vec<u8>` inside `ImageBuffer` (src/depth_image.rs:8-9), ordinary pageable
// memory.  cudaMemcpyAsync on pageable memory is staged by the driver and synchronous for the host, which would turn the
// batch path's copy-in / compute / copy-out pipeline into a sequence.  Instead the library keeps its own page-locked ring:
//
//   host -> device   a pool of copy threads memcpy's a piece (<= PIECE bytes) of the caller's buffer into a ring slot, the
//                    calling thread enqueues one cudaMemcpyAsync per piece on the copy-in stream; while the DMA of piece k
//                    runs, the threads fill piece k+1.  A slot is reused once its DMA event has completed.
//   device -> host   one cudaMemcpyAsync per piece into a ring slot on the copy-out stream; the calling thread drains
//                    completed pieces (event query) into the caller's buffer with the same pool, opportunistically
//                    while it stages later inputs and completely before the call returns.
//
// Page-locked caller buffers (cudaHostAlloc / cudaHostRegister: what bench.py's pinned arm and any CUDA-aware host use)
// skip all of this: the DMA reads and writes them directly.  No batched-memcpy API is involved: one copy per piece.
#pragma once
#include <atomic>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

namespace stepth {

// ---- a small pool of threads that run memcpy slices ----------------------------------------------------------------
class CopyPool {
  public:
    struct Slice {
        uint8_t *dst;
        const uint8_t *src;
        size_t row_bytes, rows, dst_pitch, src_pitch;   // rows x row_bytes; pitch == row_bytes for a linear copy
    };

    explicit CopyPool(int n_threads) {
        n_ = n_threads < 1 ? 1 : n_threads;
        for (int i = 0; i < n_ - 1; ++i) workers_.emplace_back([this] { loop(); });   // the caller is the n-th copier
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto &t : workers_) t.join();
    }
    int threads() const { return n_; }

    // rows x row_bytes from src (pitch src_pitch) to dst (pitch dst_pitch); returns when every byte has been copied.
    // The work is cut into at most `threads()` slices of whole rows (or byte ranges of a linear copy); the caller copies one.
    void copy2d(void *dst, size_t dst_pitch, const void *src, size_t src_pitch, size_t row_bytes, size_t rows) {
        if (rows == 0 || row_bytes == 0) return;
        const bool linear = dst_pitch == row_bytes && src_pitch == row_bytes;
        const size_t total = row_bytes * rows;
        size_t parts = (size_t)n_;
        if (total < parts * (256u << 10)) parts = total / (256u << 10) ? total / (256u << 10) : 1;   // >= 256 KiB per slice
        std::vector<Slice> slices;
        if (linear) {
            const size_t per = ((total + parts - 1) / parts + 63) & ~(size_t)63;
            for (size_t o = 0; o < total; o += per)
                slices.push_back({(uint8_t *)dst + o, (const uint8_t *)src + o, per < total - o ? per : total - o, 1, 0, 0});
        } else {
            if (parts > rows) parts = rows;
            const size_t per = (rows + parts - 1) / parts;
            for (size_t r = 0; r < rows; r += per)
                slices.push_back({(uint8_t *)dst + r * dst_pitch, (const uint8_t *)src + r * src_pitch, row_bytes,
                                  per < rows - r ? per : rows - r, dst_pitch, src_pitch});
        }
        if (slices.size() == 1) {
            run(slices[0]);
            return;
        }
        std::atomic<int> left((int)slices.size() - 1);
        {
            std::lock_guard<std::mutex> lk(m_);
            for (size_t i = 1; i < slices.size(); ++i) q_.push_back({slices[i], &left});
        }
        cv_.notify_all();
        run(slices[0]);
        // the tail of the queue may still hold slices of this call when all workers are busy: help out
        for (;;) {
            Task t;
            {
                std::lock_guard<std::mutex> lk(m_);
                if (q_.empty()) break;
                t = q_.front();
                q_.pop_front();
            }
            run(t.s);
            t.left->fetch_sub(1, std::memory_order_acq_rel);
        }
        while (left.load(std::memory_order_acquire) > 0) std::this_thread::yield();
    }
    void copy(void *dst, const void *src, size_t bytes) { copy2d(dst, bytes, src, bytes, bytes, 1); }

  private:
    struct Task {
        Slice s;
        std::atomic<int> *left;
    };
    static void run(const Slice &s) {
        if (s.rows == 1) {
            memcpy(s.dst, s.src, s.row_bytes);
            return;
        }
        for (size_t r = 0; r < s.rows; ++r) memcpy(s.dst + r * s.dst_pitch, s.src + r * s.src_pitch, s.row_bytes);
    }
    void loop() {
        for (;;) {
            Task t;
            {
                std::unique_lock<std::mutex> lk(m_);
                cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
                if (stop_ && q_.empty()) return;
                t = q_.front();
                q_.pop_front();
            }
            run(t.s);
            t.left->fetch_sub(1, std::memory_order_acq_rel);
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Task> q_;
    bool stop_ = false;
};

// ---- a FIFO of page-locked slots ---------------------------------------------------------------------------------------
// Slots are handed out in order and freed in order.  A slot is busy until its event has completed (host->device ring) or
// until it has been drained into the caller's buffer (device->host ring).
struct PinnedRing {
    struct Slot {
        size_t off = 0, bytes = 0;
        cudaEvent_t ev = nullptr;   // the DMA that reads (in-ring) or fills (out-ring) the slot
        // out-ring only: where the bytes go once the DMA is over (a 2-D window of the caller's buffer)
        uint8_t *user = nullptr;
        size_t user_pitch = 0, row_bytes = 0, rows = 0;
    };
    uint8_t *base = nullptr;
    size_t cap = 0, head = 0;       // next allocation starts at head (wraps to 0 when the tail end is too short)
    std::deque<Slot> live;          // oldest first
    std::vector<cudaEvent_t> spare; // events of freed slots

    cudaError_t init(size_t bytes) {
        if (base && cap >= bytes) return cudaSuccess;
        if (base) cudaFreeHost(base);
        base = nullptr;
        cap = head = 0;
        cudaError_t e = cudaMallocHost((void **)&base, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void destroy() {
        for (auto &s : live)
            if (s.ev) cudaEventDestroy(s.ev);
        for (auto e : spare) cudaEventDestroy(e);
        live.clear();
        spare.clear();
        if (base) cudaFreeHost(base);
        base = nullptr;
        cap = head = 0;
    }
    // room for `bytes` contiguous bytes at head without touching a live slot?
    bool fits(size_t bytes, size_t &off) const {
        if (live.empty()) {
            off = 0;
            return bytes <= cap;
        }
        const size_t tail = live.front().off;           // oldest live byte
        if (head > tail) {   // live region is [tail, head): free = [head, cap) + [0, tail)
            if (cap - head >= bytes) { off = head; return true; }
            if (tail >= bytes) { off = 0; return true; }
            return false;
        }
        // live region wraps: [tail, cap) + [0, head): free = [head, tail)   (head == tail: the ring is full)
        if (head < tail && tail - head >= bytes) { off = head; return true; }
        return false;
    }
    cudaEvent_t take_event() {
        if (!spare.empty()) {
            cudaEvent_t e = spare.back();
            spare.pop_back();
            return e;
        }
        cudaEvent_t e = nullptr;
        cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        return e;
    }
    Slot &push(size_t off, size_t bytes) {
        Slot s;
        s.off = off;
        s.bytes = bytes;
        s.ev = take_event();
        live.push_back(s);
        head = off + bytes;
        return live.back();
    }
    void pop() {
        spare.push_back(live.front().ev);
        live.pop_front();
        if (live.empty()) head = 0;
    }
};

constexpr size_t STAGE_PIECE = 4u << 20;      // bytes per staged piece
constexpr size_t STAGE_IN_RING = 64u << 20;   // page-locked bytes of the host->device ring
constexpr size_t STAGE_OUT_RING = 192u << 20; // page-locked bytes of the device->host ring

struct Stager {
    std::unique_ptr<CopyPool> pool;
    PinnedRing in, out;
    uint64_t staged_in = 0, staged_out = 0;   // bytes that went through the rings (diagnostics)

    cudaError_t ensure(int threads) {
        if (!pool) pool.reset(new CopyPool(threads));
        cudaError_t e = in.init(STAGE_IN_RING);
        if (e != cudaSuccess) return e;
        return out.init(STAGE_OUT_RING);
    }
    void destroy() {
        in.destroy();
        out.destroy();
        pool.reset();
    }
    // after a failed call whose streams have been drained: forget every slot (nothing is copied to the caller any more)
    void abandon() {
        while (!in.live.empty()) in.pop();
        while (!out.live.empty()) out.pop();
    }

    // frees the in-ring slots whose DMA is over; block: wait for the oldest one
    cudaError_t reap_in(bool block) {
        while (!in.live.empty()) {
            cudaError_t q = block ? cudaEventSynchronize(in.live.front().ev) : cudaEventQuery(in.live.front().ev);
            if (q == cudaErrorNotReady) return cudaSuccess;
            if (q != cudaSuccess) return q;
            in.pop();
            block = false;
        }
        return cudaSuccess;
    }
    // copies the out-ring slots whose DMA is over into the caller's buffers; block: wait for the oldest one; all: every one
    cudaError_t drain_out(bool block, bool all = false) {
        while (!out.live.empty()) {
            PinnedRing::Slot &s = out.live.front();
            cudaError_t q = (block || all) ? cudaEventSynchronize(s.ev) : cudaEventQuery(s.ev);
            if (q == cudaErrorNotReady) return cudaSuccess;
            if (q != cudaSuccess) return q;
            pool->copy2d(s.user, s.user_pitch, out.base + s.off, s.row_bytes, s.row_bytes, s.rows);
            out.pop();
            block = false;
        }
        return cudaSuccess;
    }

    // host (pageable) -> device: rows x row_bytes with pitch src_pitch -> contiguous device bytes
    cudaError_t h2d(void *dev, const void *host, size_t src_pitch, size_t row_bytes, size_t rows, cudaStream_t st) {
        if (src_pitch == row_bytes || rows == 1) {   // linear: cut into pieces of bytes, not of rows
            const size_t total = row_bytes * rows;
            row_bytes = total < STAGE_PIECE ? total : STAGE_PIECE;
            // pieces are `row_bytes` long except the last: walk them as rows of a pitch == row_bytes image
            const size_t full = total / row_bytes, rest = total - full * row_bytes;
            cudaError_t e = h2d_rows(dev, host, row_bytes, row_bytes, full, st);
            if (e != cudaSuccess || rest == 0) return e;
            return h2d_rows((uint8_t *)dev + full * row_bytes, (const uint8_t *)host + full * row_bytes, rest, rest, 1, st);
        }
        return h2d_rows(dev, host, src_pitch, row_bytes, rows, st);
    }

    cudaError_t h2d_rows(void *dev, const void *host, size_t src_pitch, size_t row_bytes, size_t rows, cudaStream_t st) {
        const size_t rows_per_piece = row_bytes >= STAGE_PIECE ? 1 : STAGE_PIECE / row_bytes;
        if (row_bytes > STAGE_IN_RING / 2) return cudaErrorInvalidValue;   // a single row wider than the ring: not staged
        for (size_t r = 0; r < rows; r += rows_per_piece) {
            const size_t nr = rows - r < rows_per_piece ? rows - r : rows_per_piece, bytes = nr * row_bytes;
            size_t off = 0;
            cudaError_t e = reap_in(false);
            if (e != cudaSuccess) return e;
            while (!in.fits(bytes, off)) {
                if ((e = reap_in(true)) != cudaSuccess) return e;
            }
            PinnedRing::Slot &s = in.push(off, bytes);
            pool->copy2d(in.base + off, row_bytes, (const uint8_t *)host + r * src_pitch, src_pitch, row_bytes, nr);
            if ((e = cudaMemcpyAsync((uint8_t *)dev + r * row_bytes, in.base + off, bytes, cudaMemcpyHostToDevice, st)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(s.ev, st)) != cudaSuccess) return e;
            staged_in += bytes;
            if ((e = drain_out(false)) != cudaSuccess) return e;   // results that have arrived meanwhile
        }
        return cudaSuccess;
    }

    // device -> host (pageable): `bytes` contiguous device bytes -> the caller's buffer; completes in drain_out()
    cudaError_t d2h(void *host, const void *dev, size_t bytes, cudaStream_t st) {
        for (size_t o = 0; o < bytes; o += STAGE_PIECE) {
            const size_t nb = bytes - o < STAGE_PIECE ? bytes - o : STAGE_PIECE;
            size_t off = 0;
            cudaError_t e;
            while (!out.fits(nb, off)) {
                if ((e = drain_out(true)) != cudaSuccess) return e;
            }
            PinnedRing::Slot &s = out.push(off, nb);
            s.user = (uint8_t *)host + o;
            s.user_pitch = s.row_bytes = nb;
            s.rows = 1;
            if ((e = cudaMemcpyAsync(out.base + off, (const uint8_t *)dev + o, nb, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(s.ev, st)) != cudaSuccess) return e;
            staged_out += nb;
        }
        return cudaSuccess;
    }
};

// page-locked (cudaHostAlloc / cudaHostRegister / managed) or ordinary pageable memory?
inline bool host_ptr_is_pageable(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();   // older drivers report unregistered memory as an error: pageable
        return true;
    }
    return a.type == cudaMemoryTypeUnregistered;
}

}  // namespace stepth
