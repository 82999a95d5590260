// keys.cuh — order-preserving ("sortable") key form and comparators.
//
// The reference orders keys with BaseNode::KeyCompare (include/vstore/b_tree.h:116-134): for keys
// shorter than 16 bytes a byte-wise difference of *signed* chars starting at byte 0 (my_memcmp,
// b_tree.h:99-106), otherwise libc memcmp (unsigned bytes). Both are byte-lexicographic orders, so
// packing the key bytes big-endian into 32-bit words — after flipping the sign bit of every byte in
// the signed case — gives words whose unsigned lexicographic order IS the reference order. All
// device-side search (fences, leaf keys) works on this form; the inverse is applied when keys are
// emitted. YCSB int keys (4 raw little-endian bytes of the row id) become bswap32(k) ^ 0x80808080.
#pragma once
#include <stdint.h>
#include "node_layout.h"

namespace stage {

// per-leaf flags kept in the `cont` column
constexpr uint8_t kLeafHop = 1;          // the leaf's largest visible key equals its upper fence: the iterator hops on (Q6)
constexpr uint8_t kLeafSlotOrdered = 2;  // slot order == key order inside the leaf: the slot-order truncation (Q5) is a no-op

template <int NW>
struct SKey {
    uint32_t w[NW];
};

STAGE_HD uint32_t bswap32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(x, 0, 0x0123);
#else
    return __builtin_bswap32(x);
#endif
}

// xormask: 0x80808080 for the signed-byte order, 0 for memcmp order.
template <int NW>
STAGE_HD SKey<NW> to_sortable(const uint8_t *key, uint32_t key_size, uint32_t xormask) {
    SKey<NW> s;
    if ((key_size & 3u) == 0 && ((uintptr_t) key & 3u) == 0) {
        const uint32_t *k32 = reinterpret_cast<const uint32_t *>(key);
#pragma unroll
        for (int i = 0; i < NW; ++i) s.w[i] = (uint32_t) (i * 4) < key_size ? (bswap32(k32[i]) ^ xormask) : 0u;
    } else {
#pragma unroll
        for (int i = 0; i < NW; ++i) {
            uint32_t v = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                uint32_t idx = (uint32_t) (i * 4 + b);
                uint32_t byte = idx < key_size ? ((uint32_t) key[idx] ^ (xormask & 0xFFu)) : 0u;
                v = (v << 8) | byte;
            }
            s.w[i] = v;
        }
    }
    return s;
}

// Variable-length keys (STAGE_KEY_VARLEN). KeyCompare orders two keys by their first min(l1, l2) bytes and then by
// length (b_tree.h:116-134): "prefix first, shorter first". Sortable form of a key of `len` bytes in a table whose keys
// have at most slot_size - 1 bytes: the transformed bytes, zero padding (below every transformed byte value or equal
// to the smallest), and the length as the LAST byte — a key that is a proper prefix of another, or equals it up to
// bytes of the minimum value, is decided by that byte exactly as the reference decides by length. Outside the library
// such a key travels as a fixed slot of slot_size bytes, = from_sortable of this form (stage_gpu_key_pack in stage_gpu.h),
// so every kernel treats the table as a fixed-length-key table; only the leaf transform sees real lengths.
template <int NW>
STAGE_HD SKey<NW> varlen_sortable(const uint8_t *key, uint32_t len, uint32_t slot_size, uint32_t xormask) {
    SKey<NW> s;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        uint32_t v = 0;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const uint32_t idx = (uint32_t) (i * 4 + b);
            const uint32_t byte = idx < len ? ((uint32_t) key[idx] ^ (xormask & 0xFFu)) : (idx + 1 == slot_size ? len : 0u);
            v = (v << 8) | byte;
        }
        s.w[i] = v;
    }
    return s;
}

// inverse: write the original key bytes
template <int NW>
STAGE_HD void from_sortable(const SKey<NW> &s, uint8_t *key, uint32_t key_size, uint32_t xormask) {
    if (key_size == (uint32_t) NW * 4 && ((uintptr_t) key & 3u) == 0) {
        uint32_t *k32 = reinterpret_cast<uint32_t *>(key);
#pragma unroll
        for (int i = 0; i < NW; ++i) k32[i] = bswap32(s.w[i] ^ xormask);
        return;
    }
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        uint32_t v = s.w[i] ^ xormask;
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            uint32_t idx = (uint32_t) (i * 4 + b);
            if (idx < key_size) key[idx] = (uint8_t) (v >> (24 - 8 * b));
        }
    }
}

template <int NW>
STAGE_HD bool skey_less(const SKey<NW> &a, const SKey<NW> &b) {
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        if (a.w[i] != b.w[i]) return a.w[i] < b.w[i];
    }
    return false;
}

template <int NW>
STAGE_HD bool skey_equal(const SKey<NW> &a, const SKey<NW> &b) {
    bool e = true;
#pragma unroll
    for (int i = 0; i < NW; ++i) e = e && (a.w[i] == b.w[i]);
    return e;
}

template <int NW>
STAGE_HD SKey<NW> skey_inf() {
    SKey<NW> s;
#pragma unroll
    for (int i = 0; i < NW; ++i) s.w[i] = 0xFFFFFFFFu;
    return s;
}

#if defined(__CUDACC__)
// vector load of one sortable key from a key column (NW words per key, 4*NW-byte aligned)
template <int NW>
__device__ __forceinline__ SKey<NW> load_skey(const uint32_t *col, uint64_t idx) {
    SKey<NW> s;
    if constexpr (NW == 1) {
        s.w[0] = __ldg(col + idx);
    } else if constexpr (NW == 2) {
        uint2 v = __ldg(reinterpret_cast<const uint2 *>(col) + idx);
        s.w[0] = v.x; s.w[1] = v.y;
    } else if constexpr (NW == 4) {
        uint4 v = __ldg(reinterpret_cast<const uint4 *>(col) + idx);
        s.w[0] = v.x; s.w[1] = v.y; s.w[2] = v.z; s.w[3] = v.w;
    } else {
        const uint4 *p = reinterpret_cast<const uint4 *>(col + idx * NW);
#pragma unroll
        for (int i = 0; i < NW / 4; ++i) {
            uint4 v = __ldg(p + i);
            s.w[4 * i] = v.x; s.w[4 * i + 1] = v.y; s.w[4 * i + 2] = v.z; s.w[4 * i + 3] = v.w;
        }
    }
    return s;
}

// same, through a generic pointer (shared-memory copy of a level)
template <int NW>
__device__ __forceinline__ SKey<NW> load_skey_generic(const uint32_t *col, uint64_t idx) {
    SKey<NW> s;
#pragma unroll
    for (int i = 0; i < NW; ++i) s.w[i] = col[idx * NW + i];
    return s;
}

template <int NW>
__device__ __forceinline__ void store_skey(uint32_t *col, uint64_t idx, const SKey<NW> &s) {
#pragma unroll
    for (int i = 0; i < NW; ++i) col[idx * NW + i] = s.w[i];
}
#endif

}  // namespace stage
