// hostcopy.cpp -- the memcpy of the staging threads (staging.cuh), compiled by the host compiler.
//
// A staged piece is written once and read once, by somebody else (the DMA engine, or the caller much later): plain stores
// would first pull every destination line into the cache (read-for-ownership), i.e. move three bytes per byte copied, and
// with a dozen copy threads beside two DMA streams the host's memory bandwidth is what the pageable path runs out of.
// Non-temporal stores write the lines straight to memory: two bytes moved per byte copied.
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#if defined(__x86_64__)
#include <immintrin.h>

__attribute__((target("avx2"))) static void stream_copy_avx2(uint8_t *dst, const uint8_t *src, size_t n) {
    // head: up to the first 64-byte boundary of the destination
    const size_t head = (64 - (reinterpret_cast<uintptr_t>(dst) & 63)) & 63;
    if (head) {
        const size_t h = head < n ? head : n;
        memcpy(dst, src, h);
        dst += h; src += h; n -= h;
    }
    size_t i = 0;
    for (; i + 128 <= n; i += 128) {
        const __m256i a = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i));
        const __m256i b = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 32));
        const __m256i c = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 64));
        const __m256i d = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(src + i + 96));
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i), a);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 32), b);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 64), c);
        _mm256_stream_si256(reinterpret_cast<__m256i *>(dst + i + 96), d);
    }
    _mm_sfence();
    if (i < n) memcpy(dst + i, src + i, n - i);
}
#endif

// dst and src do not overlap
extern "C" void stepth_stream_copy(void *dst, const void *src, size_t n) {
#if defined(__x86_64__)
    static const bool avx2 = __builtin_cpu_supports("avx2") && !(getenv("STEPTH_STAGE_NT") && atoi(getenv("STEPTH_STAGE_NT")) == 0);
    if (avx2 && n >= 4096) {
        stream_copy_avx2(static_cast<uint8_t *>(dst), static_cast<const uint8_t *>(src), n);
        return;
    }
#endif
    memcpy(dst, src, n);
}
