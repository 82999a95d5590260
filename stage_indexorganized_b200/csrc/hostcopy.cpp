// hostcopy.cpp — leaf images into the pinned staging buffers with non-temporal stores (compiled by g++ -mavx2; the rest
// of the library goes through nvcc).
//
// The flattening copies every 64 KiB leaf once from the tree's memory into a pinned buffer that the DMA engine then
// reads. glibc's memcpy uses ordinary stores at this size: every destination line is first read into the cache (read for
// ownership) and written back later — three memory transfers per byte copied, for data the CPU never touches again.
// Streaming stores write the lines straight to memory (two transfers per byte).
#include <immintrin.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

extern "C" int stage_host_has_avx2(void) { return __builtin_cpu_supports("avx2") ? 1 : 0; }

// dst 32-byte aligned, n a multiple of 128
extern "C" void stage_host_stream_copy(uint8_t *dst, const uint8_t *src, size_t n) {
    for (size_t i = 0; i < n; i += 128) {
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
}
