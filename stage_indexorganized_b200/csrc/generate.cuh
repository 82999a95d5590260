// generate.cuh — on-device construction of the YCSB int table in flattened form, for table sizes whose
// host tree cannot be materialised.
//
// Row function restated from LoadYCSBRows (benchmark/ycsb/ycsb_loader.cpp:144-159): key = uint32 rowid
// stored as 4 raw bytes, payload = payload_size x (uint8) rowid, commit id INVALID_CID (0).
// Rows are laid out in REFERENCE key order (signed bytes, LSB first — keys.cuh), i.e. ascending
// sortable(k) = bswap32(k) ^ 0x80808080, so row r needs the r-th smallest sortable value among
// k in [0, N): computed in closed form digit by digit (no sort, no N-sized temporary).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "keys.cuh"

namespace stage {

struct GenArgs {
    uint64_t n_total, first_rank, n_local, n_leaves;
    uint32_t rows_per_leaf, payload_size;
    uint32_t *skey;
    uint16_t *slot;
    uint32_t *ts, *chain;
    uint8_t *payload;
    uint32_t *leaf_off, *sep;
    uint8_t *cont;
};

// number of t' < t (t in 0..256) whose byte b = t' ^ 0x80 satisfies b <= al
STAGE_HD uint32_t gen_le_count(uint32_t t, uint32_t al) {
    // t' in [0,128) -> b in [128,256); t' in [128,256) -> b in [0,128)
    uint32_t c = 0;
    const uint32_t t_hi = t < 128 ? t : 128;                    // b = 128 .. 127 + t_hi
    if (al >= 128 && t_hi > 0) { uint32_t top = al < 127 + t_hi ? al : 127 + t_hi; c += top - 128 + 1; }
    if (t > 128) { uint32_t last = t - 129; c += (al < last ? al : last) + 1; }    // b = 0 .. t-129
    return c;
}

// r-th (0-based) key of {0..N-1} in reference order
STAGE_HD uint32_t ycsb_rank_to_key(uint64_t N, uint64_t r) {
    uint64_t low = 0;
    const uint64_t M = N - 1;
    for (int d = 0; d < 4; ++d) {
        const int shift = 8 * d;
        const uint64_t a = (M - low) >> shift;                  // byte values b <= a are possible
        const uint64_t ah = a >> 8;
        const uint32_t al = (uint32_t) (a & 255);
        // rows before digit t: t*ah + gen_le_count(t, al); find the largest t with that <= r
        uint32_t lo = 0, hi = 255;
        while (lo < hi) {
            const uint32_t mid = (lo + hi + 1) >> 1;
            const uint64_t before = (uint64_t) mid * ah + gen_le_count(mid, al);
            if (before <= r) lo = mid; else hi = mid - 1;
        }
        r -= (uint64_t) lo * ah + gen_le_count(lo, al);
        low |= (uint64_t) (lo ^ 0x80u) << shift;
    }
    return (uint32_t) low;
}

#if defined(__CUDACC__)
// ---- bulk load from device-resident columns (stage_gpu_snapshot_alloc_sorted / seal_sorted) -------------------
// raw keys (n x key_size bytes, already in reference key order) -> sortable key column; consolidated leaves of
// rows_per_leaf records (slot = position in the leaf, no chains).
template <int NW>
__global__ void bulk_keys_kernel(const uint8_t *raw, uint64_t n, uint32_t key_size, uint32_t xormask, uint32_t rows_per_leaf,
                                 uint32_t *skey, uint16_t *slot, uint32_t *chain) {
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const SKey<NW> k = to_sortable<NW>(raw + i * key_size, key_size, xormask);
#pragma unroll
        for (int w = 0; w < NW; ++w) skey[i * NW + w] = k.w[w];
        slot[i] = (uint16_t) (i % rows_per_leaf);
        chain[i] = 0;
    }
}
// leaf tables of a consolidated key column + the order check (keys must be strictly ascending: they are unique
// in the reference, LeafNode::CheckUnique, src/vstore/b_tree.cpp:1395-1417); err = 3 on a violation.
template <int NW>
__global__ void bulk_leaves_kernel(const uint32_t *skey, uint64_t n, uint64_t n_leaves, uint32_t rows_per_leaf,
                                   uint32_t *leaf_off, uint32_t *sep, uint8_t *cont, uint32_t *err) {
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i + 1 < n; i += stride)
        if (!skey_less<NW>(load_skey<NW>(skey, i), load_skey<NW>(skey, i + 1))) atomicExch(err, 3u);
    for (uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; j <= n_leaves; j += stride) {
        const uint64_t off = j * rows_per_leaf < n ? j * rows_per_leaf : n;
        leaf_off[j] = (uint32_t) off;
        if (j < n_leaves) {
            const uint64_t end = (j + 1) * rows_per_leaf < n ? (j + 1) * rows_per_leaf : n;
            const bool hop = j + 1 < n_leaves && end > off;      // separator of a consolidated leaf = its largest key
#pragma unroll
            for (int w = 0; w < NW; ++w) sep[j * NW + w] = hop ? skey[(end - 1) * NW + w] : 0xFFFFFFFFu;
            cont[j] = (hop ? kLeafHop : 0) | kLeafSlotOrdered;          // bulk-loaded leaves: slot order == key order
        }
    }
}

__global__ void generate_ycsb_meta_kernel(GenArgs g) {
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < g.n_local; i += stride) {
        const uint32_t k = ycsb_rank_to_key(g.n_total, g.first_rank + i);
        g.skey[i] = bswap32(k) ^ 0x80808080u;
        g.slot[i] = (uint16_t) (i % g.rows_per_leaf);
        g.ts[i] = 0;
        g.chain[i] = 0;
    }
    for (uint64_t j = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; j <= g.n_leaves; j += stride) {
        const uint64_t off = j * g.rows_per_leaf < g.n_local ? j * g.rows_per_leaf : g.n_local;
        g.leaf_off[j] = (uint32_t) off;
        if (j < g.n_leaves) {
            const uint64_t end = (j + 1) * g.rows_per_leaf < g.n_local ? (j + 1) * g.rows_per_leaf : g.n_local;
            const bool last = j + 1 == g.n_leaves;
            // separator of a consolidated leaf = its largest key (split rule, src/vstore/b_tree.cpp:1651-1657)
            uint32_t s = 0xFFFFFFFFu;
            if (!last && end > off) s = bswap32(ycsb_rank_to_key(g.n_total, g.first_rank + end - 1)) ^ 0x80808080u;
            g.sep[j] = s;
            g.cont[j] = ((!last && end > off) ? kLeafHop : 0) | kLeafSlotOrdered;
        }
    }
}

__global__ void generate_ycsb_payload_kernel(GenArgs g) {
    const uint32_t P = g.payload_size;
    const uint64_t total = g.n_local * (uint64_t) P;
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    if ((P & 7) == 0) {
        const uint64_t words = total / 8, wpr = P / 8;
        uint64_t *out = reinterpret_cast<uint64_t *>(g.payload);
        for (uint64_t w = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; w < words; w += stride) {
            const uint64_t row = w / wpr;
            const uint64_t b = ((g.skey[row] >> 24) ^ 0x80u) & 0xFFu;      // (uint8) rowid
            out[w] = b * 0x0101010101010101ull;
        }
    } else {
        for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
            g.payload[i] = (uint8_t) (((g.skey[i / P] >> 24) ^ 0x80u) & 0xFFu);
    }
}
#endif

}  // namespace stage
