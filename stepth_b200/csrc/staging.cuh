// staging.cuh -- host buffers that are NOT page-locked, moved through a pinned ring.
//
// The drop-in caller is Rust: its images are `Vec<u8>` inside `ImageBuffer` (src/depth_image.rs:8-9), ordinary pageable
// memory.  cudaMemcpyAsync on pageable memory is staged by the driver and synchronous for the host, which would turn the
// batch path's copy-in / compute / copy-out pipeline into a sequence.  Instead the library keeps its own page-locked rings:
//
//   host -> device   the caller's buffer is cut into pieces (<= STAGE_PIECE bytes).  A pool of copy threads memcpy's piece
//                    k+1, k+2 .. into ring slots while the calling thread enqueues the cudaMemcpyAsync of piece k on the
//                    copy-in stream: the threads never wait for the DMA and the DMA never waits for more than one piece.
//                    A slot is reused once the event behind its DMA has completed.
//   device -> host   one cudaMemcpyAsync per piece into a ring slot on the copy-out stream; whenever the calling thread
//                    passes by (while it stages later inputs, and before the call returns) the pieces whose DMA has
//                    completed are handed to the same threads, which copy them into the caller's buffer.
//
// Page-locked caller buffers (cudaHostAlloc / cudaHostRegister: what bench.py's pinned arm and any CUDA-aware host use)
// skip all of this: the DMA reads and writes them directly.  No batched-memcpy API is involved: one copy per piece.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

extern "C" void stepth_stream_copy(void *dst, const void *src, size_t n);   // hostcopy.cpp: non-temporal stores

namespace stepth {

// ---- a small pool of threads that run memcpy slices ----------------------------------------------------------------
class CopyPool {
  public:
    struct Job {
        std::atomic<int> left{0};
        bool done() const { return left.load(std::memory_order_acquire) == 0; }
    };
    using JobPtr = std::shared_ptr<Job>;

    explicit CopyPool(int n_threads) {
        n_ = n_threads < 1 ? 1 : n_threads;
        for (int i = 0; i < n_ - 1; ++i) workers_.emplace_back([this] { loop(); });   // the caller helps while it waits
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_.store(true);
        }
        cv_.notify_all();
        for (auto &t : workers_) t.join();
    }
    int threads() const { return n_; }

    // rows x row_bytes from src (pitch src_pitch) to dst (pitch dst_pitch), cut into at most threads() slices of whole rows
    // (byte ranges for a linear copy).  Returns at once; wait() blocks until every byte has been copied.
    JobPtr submit(void *dst, size_t dst_pitch, const void *src, size_t src_pitch, size_t row_bytes, size_t rows) {
        JobPtr job = std::make_shared<Job>();
        if (rows == 0 || row_bytes == 0) return job;
        const bool linear = (dst_pitch == row_bytes && src_pitch == row_bytes) || rows == 1;
        const size_t total = row_bytes * rows;
        size_t parts = (size_t)n_;
        if (total < parts * MIN_SLICE) parts = total / MIN_SLICE ? total / MIN_SLICE : 1;
        std::vector<Task> tasks;
        if (linear) {
            const size_t per = ((total + parts - 1) / parts + 63) & ~(size_t)63;
            for (size_t o = 0; o < total; o += per)
                tasks.push_back({(uint8_t *)dst + o, (const uint8_t *)src + o, per < total - o ? per : total - o, 1, 0, 0, job});
        } else {
            if (parts > rows) parts = rows;
            const size_t per = (rows + parts - 1) / parts;
            for (size_t r = 0; r < rows; r += per)
                tasks.push_back({(uint8_t *)dst + r * dst_pitch, (const uint8_t *)src + r * src_pitch, row_bytes,
                                 per < rows - r ? per : rows - r, dst_pitch, src_pitch, job});
        }
        job->left.store((int)tasks.size(), std::memory_order_release);
        {
            std::lock_guard<std::mutex> lk(m_);
            for (auto &t : tasks) q_.push_back(t);
            queued_.fetch_add((int)tasks.size(), std::memory_order_release);
        }
        if (sleepers_.load(std::memory_order_acquire) > 0) cv_.notify_all();
        return job;
    }
    // the caller copies slices itself (of any job: the queue is FIFO) until `job` is complete
    void wait(const JobPtr &job) {
        while (!job->done()) {
            Task t;
            if (pop(t)) run(t);
            else std::this_thread::yield();
        }
    }
    void copy2d(void *dst, size_t dst_pitch, const void *src, size_t src_pitch, size_t row_bytes, size_t rows) {
        wait(submit(dst, dst_pitch, src, src_pitch, row_bytes, rows));
    }

  private:
    static constexpr size_t MIN_SLICE = 256u << 10;   // bytes: below this a slice costs more to hand over than to copy
    struct Task {
        uint8_t *dst;
        const uint8_t *src;
        size_t row_bytes, rows, dst_pitch, src_pitch;
        JobPtr job;
    };
    static void run(const Task &t) {
        if (t.rows == 1) stepth_stream_copy(t.dst, t.src, t.row_bytes);
        else
            for (size_t r = 0; r < t.rows; ++r) stepth_stream_copy(t.dst + r * t.dst_pitch, t.src + r * t.src_pitch, t.row_bytes);
        t.job->left.fetch_sub(1, std::memory_order_acq_rel);
    }
    bool pop(Task &t) {
        if (queued_.load(std::memory_order_acquire) <= 0) return false;
        std::lock_guard<std::mutex> lk(m_);
        if (q_.empty()) return false;
        t = std::move(q_.front());
        q_.pop_front();
        queued_.fetch_sub(1, std::memory_order_release);
        return true;
    }
    void loop() {
        for (;;) {
            Task t;
            if (pop(t)) {
                run(t);
                continue;
            }
            // between the pieces of a call the next slice is microseconds away: spin for a while before sleeping
            const auto t0 = std::chrono::steady_clock::now();
            bool got = false;
            while (std::chrono::steady_clock::now() - t0 < std::chrono::microseconds(200)) {
                if ((got = pop(t))) break;
                if (stop_.load(std::memory_order_relaxed)) return;
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            if (got) {
                run(t);
                continue;
            }
            std::unique_lock<std::mutex> lk(m_);
            sleepers_.fetch_add(1, std::memory_order_acq_rel);
            cv_.wait(lk, [this] { return stop_.load() || !q_.empty(); });
            sleepers_.fetch_sub(1, std::memory_order_acq_rel);
            if (stop_.load() && q_.empty()) return;
        }
    }
    int n_;
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_;
    std::deque<Task> q_;
    std::atomic<int> queued_{0}, sleepers_{0};
    std::atomic<bool> stop_{false};
};

// ---- a FIFO of page-locked slots ---------------------------------------------------------------------------------------
// Slots are handed out in order and freed in order.
struct PinnedRing {
    struct Slot {
        size_t off = 0, bytes = 0;
        cudaEvent_t ev = nullptr;   // the DMA that reads (in-ring) or fills (out-ring) the slot
        CopyPool::JobPtr job;       // in-ring: the threads filling the slot;  out-ring: the threads emptying it (once submitted)
        bool dma_issued = false;    // in-ring: the cudaMemcpyAsync out of this slot has been enqueued
        uint8_t *dev = nullptr;     // in-ring: where the DMA puts the bytes
        uint8_t *user = nullptr;    // out-ring: where the bytes go once the DMA is over
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
    // room for `bytes` contiguous bytes without touching a live slot?
    bool fits(size_t bytes, size_t &off) const {
        if (live.empty()) {
            off = 0;
            return bytes <= cap;
        }
        const size_t tail = live.front().off;   // oldest live byte
        if (head > tail) {                      // live region is [tail, head): free = [head, cap) + [0, tail)
            if (cap - head >= bytes) { off = head; return true; }
            if (tail >= bytes) { off = 0; return true; }
            return false;
        }
        // live region wraps: [tail, cap) + [0, head): free = [head, tail)   (head == tail: the ring is full)
        if (head < tail && tail - head >= bytes) { off = head; return true; }
        return false;
    }
    Slot &push(size_t off, size_t bytes) {
        Slot s;
        s.off = off;
        s.bytes = bytes;
        if (!spare.empty()) {
            s.ev = spare.back();
            spare.pop_back();
        } else {
            cudaEventCreateWithFlags(&s.ev, cudaEventDisableTiming);
        }
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

constexpr size_t STAGE_PIECE = 4u << 20;       // bytes per staged piece
constexpr size_t STAGE_IN_RING = 64u << 20;    // page-locked bytes of the host->device ring
constexpr size_t STAGE_OUT_RING = 192u << 20;  // page-locked bytes of the device->host ring
constexpr size_t STAGE_AHEAD = 3;              // pieces the copy threads may be ahead of the DMA queue

struct Stager {
    std::unique_ptr<CopyPool> pool;
    PinnedRing in, out;
    cudaStream_t in_stream = nullptr;         // the stream the pending in-slots belong to
    uint64_t staged_in = 0, staged_out = 0;   // bytes that went through the rings (diagnostics)

    cudaError_t ensure(int threads) {
        if (!pool) pool.reset(new CopyPool(threads));
        cudaError_t e = in.init(STAGE_IN_RING);
        if (e != cudaSuccess) return e;
        return out.init(STAGE_OUT_RING);
    }
    void destroy() {
        if (pool) abandon();
        in.destroy();
        out.destroy();
        pool.reset();
    }
    // after a failed call whose streams have been drained: forget every slot (nothing is copied to the caller any more)
    void abandon() {
        for (auto &s : in.live)
            if (s.job) pool->wait(s.job);
        for (auto &s : out.live)
            if (s.job) pool->wait(s.job);
        while (!in.live.empty()) in.pop();
        while (!out.live.empty()) out.pop();
    }

    // ---- host -> device -------------------------------------------------------------------------------------------------
    // enqueues the DMA of the filled slots that have none yet, oldest first; the newest `keep_ahead` may go on filling
    cudaError_t issue_in(size_t keep_ahead) {
        size_t waiting = 0;
        for (auto &s : in.live) waiting += s.dma_issued ? 0 : 1;
        for (auto &s : in.live) {
            if (s.dma_issued) continue;
            if (waiting <= keep_ahead && !s.job->done()) break;
            pool->wait(s.job);
            cudaError_t e = cudaMemcpyAsync(s.dev, in.base + s.off, s.bytes, cudaMemcpyHostToDevice, in_stream);
            if (e != cudaSuccess) return e;
            if ((e = cudaEventRecord(s.ev, in_stream)) != cudaSuccess) return e;
            s.dma_issued = true;
            s.job.reset();
            --waiting;
        }
        return cudaSuccess;
    }
    // every staged piece handed in so far is on its way: call before recording an event the consumers wait on
    cudaError_t flush_in() { return in.live.empty() ? cudaSuccess : issue_in(0); }
    // frees the in-ring slots whose DMA is over; block: wait for the oldest one
    cudaError_t reap_in(bool block) {
        while (!in.live.empty() && in.live.front().dma_issued) {
            cudaError_t q = block ? cudaEventSynchronize(in.live.front().ev) : cudaEventQuery(in.live.front().ev);
            if (q == cudaErrorNotReady) return cudaSuccess;
            if (q != cudaSuccess) return q;
            in.pop();
            block = false;
        }
        return cudaSuccess;
    }
    // rows x row_bytes with pitch src_pitch (pageable) -> contiguous device bytes, ordered on `st`.  The last pieces may still
    // be filling when this returns: flush_in() before anything on the device is told that the upload is complete.
    cudaError_t h2d(void *dev, const void *host, size_t src_pitch, size_t row_bytes, size_t rows, cudaStream_t st) {
        cudaError_t e;
        if (in_stream != st) {
            if ((e = flush_in()) != cudaSuccess) return e;
            in_stream = st;
        }
        if (src_pitch == row_bytes || rows == 1) {   // linear: pieces of bytes
            const size_t total = row_bytes * rows;
            for (size_t o = 0; o < total; o += STAGE_PIECE) {
                const size_t nb = total - o < STAGE_PIECE ? total - o : STAGE_PIECE;
                if ((e = stage_piece((uint8_t *)dev + o, (const uint8_t *)host + o, nb, nb, 1)) != cudaSuccess) return e;
            }
            return cudaSuccess;
        }
        if (row_bytes > STAGE_IN_RING / 4) return cudaErrorInvalidValue;   // a single row wider than the ring: not staged
        const size_t rows_per_piece = row_bytes >= STAGE_PIECE ? 1 : STAGE_PIECE / row_bytes;
        for (size_t r = 0; r < rows; r += rows_per_piece) {
            const size_t nr = rows - r < rows_per_piece ? rows - r : rows_per_piece;
            if ((e = stage_piece((uint8_t *)dev + r * row_bytes, (const uint8_t *)host + r * src_pitch, src_pitch, row_bytes, nr)) != cudaSuccess)
                return e;
        }
        return cudaSuccess;
    }
    cudaError_t stage_piece(uint8_t *dev, const uint8_t *host, size_t src_pitch, size_t row_bytes, size_t rows) {
        const size_t bytes = row_bytes * rows;
        size_t off = 0;
        cudaError_t e = reap_in(false);
        if (e != cudaSuccess) return e;
        while (!in.fits(bytes, off)) {
            if ((e = issue_in(0)) != cudaSuccess) return e;   // the oldest slots need a DMA before they can be waited for
            if ((e = reap_in(true)) != cudaSuccess) return e;
        }
        PinnedRing::Slot &s = in.push(off, bytes);
        s.dev = dev;
        s.job = pool->submit(in.base + off, row_bytes, host, src_pitch, row_bytes, rows);
        staged_in += bytes;
        if ((e = issue_in(STAGE_AHEAD)) != cudaSuccess) return e;
        return drain_out(false);   // results that have arrived meanwhile
    }

    // ---- device -> host -------------------------------------------------------------------------------------------------
    // hands the out-ring slots whose DMA is over to the copy threads and frees the ones they have emptied.
    // block: wait for the oldest slot;  all: for every slot
    cudaError_t drain_out(bool block, bool all = false) {
        bool first = true;
        for (auto &s : out.live) {
            if (!s.job) {
                cudaError_t q = (all || (block && first)) ? cudaEventSynchronize(s.ev) : cudaEventQuery(s.ev);
                if (q == cudaErrorNotReady) break;
                if (q != cudaSuccess) return q;
                s.job = pool->submit(s.user, s.bytes, out.base + s.off, s.bytes, s.bytes, 1);
            }
            first = false;
        }
        while (!out.live.empty() && out.live.front().job) {
            if (!out.live.front().job->done()) {
                if (!(block || all)) break;
                pool->wait(out.live.front().job);
            }
            out.pop();
            block = false;
        }
        return cudaSuccess;
    }
    // `bytes` contiguous device bytes -> the caller's (pageable) buffer; completes in drain_out()
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
            if ((e = cudaMemcpyAsync(out.base + off, (const uint8_t *)dev + o, nb, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
            if ((e = cudaEventRecord(s.ev, st)) != cudaSuccess) return e;
            staged_out += nb;
        }
        return cudaSuccess;
    }
    // before the call returns: every input consumed, every result in the caller's buffers
    cudaError_t finish() {
        cudaError_t e = flush_in();
        if (e != cudaSuccess) return e;
        while (!in.live.empty())
            if ((e = reap_in(true)) != cudaSuccess) return e;
        return drain_out(true, true);
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
