// transform.cuh — record-layout transform kernel: BzTree leaf images -> flattened columns.
//
// Input : raw LeafNode images (include/vstore/b_tree.h:571-740; records grow downward from the node
//         tail, src/vstore/b_tree.cpp:863-898; RecordMetadata array grows upward from byte 40).
// Output: for every leaf, its visible records sorted in reference key order (keys.cuh), as dense
//         columns: sortable key, original slot index (needed to reproduce the slot-order truncation
//         of LeafNode::RangeScanBySize, b_tree.cpp:1276-1302), commit id (meta & 0xFFFFFFFF),
//         version-chain head, payload row; plus per leaf whether its largest visible key equals the
//         parent separator (the iterator's leaf-hop rule, b_tree.h:913-941).
// One CTA per leaf: decode + filter in slot order, bitonic sort in shared memory, cooperative
// vectorised payload gather. HBM-bound: 64 KiB read + ~n*(P+K+10) written per leaf.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "keys.cuh"
#include "node_layout.h"

namespace stage {

struct TransformArgs {
    const uint8_t *leaves;      // staged images of this chunk, leaf_size bytes each
    uint32_t n_chunk_leaves;
    uint32_t leaf_size;
    uint64_t base_leaf;         // index of the chunk's first leaf in the snapshot
    const uint32_t *leaf_index; // optional [n_chunk_leaves]: snapshot index of every staged leaf (scattered leaves:
                                // structural refresh); base_leaf + j when null
    const uint32_t *leaf_off;   // [n_leaves+1] row offset of every leaf (device)
    const uint32_t *sep;        // [n_leaves] sortable upper fence of every leaf
    const uint8_t *has_sep;     // [n_leaves]
    uint32_t key_size, payload_size, xormask;
    uint32_t varlen;            // STAGE_KEY_VARLEN: key_size is the slot size, records carry keys of 0..key_size-1 bytes
    uint32_t sort_cap;          // power of two >= max record_count in the chunk
    uint32_t *skey;
    uint16_t *slot;
    uint32_t *ts;
    uint32_t *chain;
    uint8_t *payload;
    uint8_t *cont;              // [n_leaves] 1 if max visible key == upper fence
    uint32_t *err;              // sticky error flag (count mismatch / in-flight record)
};

template <typename V>
__device__ __forceinline__ void warp_copy_vec(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst,
                                              uint32_t nbytes, int lane) {
    const V *s = reinterpret_cast<const V *>(src);
    V *d = reinterpret_cast<V *>(dst);
    uint32_t n = nbytes / (uint32_t) sizeof(V);
    for (uint32_t i = lane; i < n; i += 32) d[i] = s[i];
}

// generic warp copy; alignment of both pointers decides the vector width
__device__ __forceinline__ void warp_copy_bytes(const uint8_t *src, uint8_t *dst, uint32_t nbytes, int lane) {
    uintptr_t a = (uintptr_t) src | (uintptr_t) dst | nbytes;
    if ((a & 15) == 0) warp_copy_vec<uint4>(src, dst, nbytes, lane);
    else if ((a & 7) == 0) warp_copy_vec<uint2>(src, dst, nbytes, lane);
    else if ((a & 3) == 0) warp_copy_vec<uint32_t>(src, dst, nbytes, lane);
    else warp_copy_vec<uint8_t>(src, dst, nbytes, lane);
}

template <int NW>
__global__ void __launch_bounds__(128) leaf_transform_kernel(TransformArgs a) {
    extern __shared__ __align__(16) uint8_t smem_raw[];
    uint32_t *s_key = reinterpret_cast<uint32_t *>(smem_raw);                       // [sort_cap][NW]
    uint16_t *s_slot = reinterpret_cast<uint16_t *>(s_key + (size_t) a.sort_cap * NW);   // [sort_cap]
    __shared__ uint32_t s_n;

    const uint32_t j = blockIdx.x;
    if (j >= a.n_chunk_leaves) return;
    const uint64_t gj = a.leaf_index ? (uint64_t) a.leaf_index[j] : a.base_leaf + j;
    const uint8_t *leaf = a.leaves + (size_t) j * a.leaf_size;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;

    const uint64_t status = *reinterpret_cast<const uint64_t *>(leaf + kOffStatus);
    const uint32_t count = status_record_count(status);
    if (tid == 0) s_n = 0;
    for (uint32_t i = tid; i < a.sort_cap; i += blockDim.x) {
#pragma unroll
        for (int w = 0; w < NW; ++w) s_key[(size_t) i * NW + w] = 0xFFFFFFFFu;
        s_slot[i] = 0xFFFFu;
    }
    __syncthreads();

    // 1. decode + filter (IsVisible, record_meta.h:99-102). Order does not matter: sorted next.
    for (uint32_t i = tid; i < count; i += blockDim.x) {
        const uint64_t m = *reinterpret_cast<const uint64_t *>(leaf + kOffMeta + (size_t) i * kMetaSize);
        if (!meta_visible(m)) continue;
        if (m & kMetaControl) atomicExch(a.err, 1u);            // in-flight record: tree not quiesced
        const uint32_t pos = atomicAdd(&s_n, 1u);
        if (pos < a.sort_cap) {
            SKey<NW> k;
            if (a.varlen) {
                const uint32_t len = meta_key_len(m);
                // longer than the slot allows, or a short key in a table ordered by memcmp (KeyCompare would switch to
                // the signed-byte order for that pair, b_tree.h:118-124): not representable
                if (len + 1 > a.key_size || (a.xormask == 0 && len < 16)) atomicExch(a.err, 3u);
                k = varlen_sortable<NW>(leaf + meta_offset(m), min(len, a.key_size - 1), a.key_size, a.xormask);
            } else {
                k = to_sortable<NW>(leaf + meta_offset(m), a.key_size, a.xormask);
            }
#pragma unroll
            for (int w = 0; w < NW; ++w) s_key[(size_t) pos * NW + w] = k.w[w];
            s_slot[pos] = (uint16_t) i;
        }
    }
    __syncthreads();
    const uint32_t n = s_n;
    const uint32_t row0 = a.leaf_off[gj];
    if (tid == 0 && (n != a.leaf_off[gj + 1] - row0 || n > a.sort_cap)) atomicExch(a.err, 2u);

    // 2. bitonic sort of (key, slot) in shared memory, ascending in reference order
    for (uint32_t k = 2; k <= a.sort_cap; k <<= 1) {
        for (uint32_t jj = k >> 1; jj > 0; jj >>= 1) {
            for (uint32_t i = tid; i < a.sort_cap; i += blockDim.x) {
                const uint32_t p = i ^ jj;
                if (p > i) {
                    SKey<NW> x, y;
#pragma unroll
                    for (int w = 0; w < NW; ++w) { x.w[w] = s_key[(size_t) i * NW + w]; y.w[w] = s_key[(size_t) p * NW + w]; }
                    const bool up = (i & k) == 0;
                    const bool gt = skey_less<NW>(y, x) || (skey_equal<NW>(x, y) && s_slot[p] < s_slot[i]);
                    if (gt == up) {
#pragma unroll
                        for (int w = 0; w < NW; ++w) { s_key[(size_t) i * NW + w] = y.w[w]; s_key[(size_t) p * NW + w] = x.w[w]; }
                        const uint16_t t = s_slot[i]; s_slot[i] = s_slot[p]; s_slot[p] = t;
                    }
                }
            }
            __syncthreads();
        }
    }

    // 3. key / slot / commit-id / chain columns
    for (uint32_t i = tid; i < n; i += blockDim.x) {
        const uint32_t sl = s_slot[i];
        const uint8_t *mp = leaf + kOffMeta + (size_t) sl * kMetaSize;
        const uint64_t m = *reinterpret_cast<const uint64_t *>(mp);
        const uint64_t nx = *reinterpret_cast<const uint64_t *>(mp + kMetaNextPtr);   // patched: 1-based version idx
        const uint64_t r = (uint64_t) row0 + i;
#pragma unroll
        for (int w = 0; w < NW; ++w) a.skey[r * NW + w] = s_key[(size_t) i * NW + w];
        a.slot[r] = (uint16_t) sl;
        a.ts[r] = meta_commit_id(m);
        a.chain[r] = (uint32_t) nx;
    }
    if (tid == 0) {
        uint8_t c = 0;
        if (n > 0 && a.has_sep[gj]) {
            bool eq = true;
#pragma unroll
            for (int w = 0; w < NW; ++w) eq = eq && (s_key[(size_t) (n - 1) * NW + w] == a.sep[gj * NW + w]);
            c = eq ? 1 : 0;
        }
        bool ordered = true;                                     // slots ascending in key order?
        for (uint32_t i = 1; i < n; ++i) ordered = ordered && s_slot[i - 1] < s_slot[i];
        a.cont[gj] = c | (ordered ? kLeafSlotOrdered : 0);
    }

    // 4. payload gather: sorted row i <- record at slot s_slot[i]; one warp per row
    const uint32_t P = a.payload_size;
    for (uint32_t i = warp; i < n; i += nwarps) {
        const uint32_t sl = s_slot[i];
        const uint64_t m = *reinterpret_cast<const uint64_t *>(leaf + kOffMeta + (size_t) sl * kMetaSize);
        const uint8_t *src = leaf + meta_offset(m) + pad_key_len(meta_key_len(m));
        uint8_t *dst = a.payload + ((uint64_t) row0 + i) * P;
        warp_copy_bytes(src, dst, P, lane);
    }
}

// Structural refresh: rows of the leaves that did not change since the previous snapshot are carried over on the
// device (old columns -> new columns at the leaf's new row offset); only changed leaves are re-staged from the host.
struct LeafCarryArgs {
    const uint32_t *src_leaf;     // [n_leaves] leaf index in the previous snapshot, kNoLeaf = re-transformed
    uint32_t n_leaves;
    const uint32_t *old_off, *new_off;
    const uint32_t *old_skey; uint32_t *new_skey;
    const uint16_t *old_slot; uint16_t *new_slot;
    const uint32_t *old_ts; uint32_t *new_ts;
    const uint32_t *old_chain; uint32_t *new_chain;
    const uint8_t *old_payload; uint8_t *new_payload;
    const uint8_t *old_cont; uint8_t *new_cont;
    uint32_t nw, payload_size;
};
constexpr uint32_t kNoLeaf = 0xFFFFFFFFu;

template <typename V>
__device__ __forceinline__ void cta_copy_vec(const uint8_t *__restrict__ src, uint8_t *__restrict__ dst, uint64_t nbytes) {
    const V *s = reinterpret_cast<const V *>(src);
    V *d = reinterpret_cast<V *>(dst);
    const uint64_t n = nbytes / sizeof(V);
    for (uint64_t i = threadIdx.x; i < n; i += blockDim.x) d[i] = s[i];
}
__device__ __forceinline__ void cta_copy_bytes(const uint8_t *src, uint8_t *dst, uint64_t nbytes) {
    const uintptr_t a = (uintptr_t) src | (uintptr_t) dst | (uintptr_t) nbytes;
    if ((a & 15) == 0) cta_copy_vec<uint4>(src, dst, nbytes);
    else if ((a & 7) == 0) cta_copy_vec<uint2>(src, dst, nbytes);
    else if ((a & 3) == 0) cta_copy_vec<uint32_t>(src, dst, nbytes);
    else if ((a & 1) == 0) cta_copy_vec<uint16_t>(src, dst, nbytes);
    else cta_copy_vec<uint8_t>(src, dst, nbytes);
}

__global__ void __launch_bounds__(256) leaf_carry_kernel(LeafCarryArgs a) {
    for (uint32_t j = blockIdx.x; j < a.n_leaves; j += gridDim.x) {
        const uint32_t o = a.src_leaf[j];
        if (o == kNoLeaf) continue;
        const uint64_t r0 = a.old_off[o], n = a.old_off[o + 1] - r0, d0 = a.new_off[j];
        cta_copy_bytes(reinterpret_cast<const uint8_t *>(a.old_skey + r0 * a.nw), reinterpret_cast<uint8_t *>(a.new_skey + d0 * a.nw), n * a.nw * 4);
        cta_copy_bytes(reinterpret_cast<const uint8_t *>(a.old_slot + r0), reinterpret_cast<uint8_t *>(a.new_slot + d0), n * 2);
        cta_copy_bytes(reinterpret_cast<const uint8_t *>(a.old_ts + r0), reinterpret_cast<uint8_t *>(a.new_ts + d0), n * 4);
        cta_copy_bytes(reinterpret_cast<const uint8_t *>(a.old_chain + r0), reinterpret_cast<uint8_t *>(a.new_chain + d0), n * 4);
        // in-place refresh: the payload column is the previous snapshot's buffer, its rows are shifted separately
        if (a.new_payload) cta_copy_bytes(a.old_payload + r0 * a.payload_size, a.new_payload + d0 * a.payload_size, n * a.payload_size);
        if (threadIdx.x == 0) a.new_cont[j] = a.old_cont[o];
    }
}

// sampled fence level: dst[i] = src[(i+1)*kFan - 1] for i < n_dst, all-ones pad keys for n_dst <= i < n_pad
// (kFan = 4, probe.cuh)
template <int NW>
__global__ void sample_level_kernel(const uint32_t *__restrict__ src, uint32_t *__restrict__ dst, uint64_t n_dst,
                                    uint64_t n_pad) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pad) return;
    uint64_t s = (i + 1) * 4 - 1;
#pragma unroll
    for (int w = 0; w < NW; ++w) dst[i * NW + w] = i < n_dst ? src[s * NW + w] : 0xFFFFFFFFu;
}

}  // namespace stage
