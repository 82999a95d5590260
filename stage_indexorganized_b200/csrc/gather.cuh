// gather.cuh — payload row gather: the kernel that moves >98 % of the bytes of both read operators.
//
// Replaces the per-row heap allocation + two memcpys of Record::New (include/vstore/b_tree.h:407-428)
// and the executors' sizeof(T) copies (include/execute/executor.h:397, 476, 637): output row slot o
// receives the payload row named by src_row[o] (a row of the flattened table, or — top bit set — a row
// of the staged version table; kNoRow = nothing to copy, e.g. a missed probe or the unused tail of a
// scan's output slab).
//
// HBM-bound streaming copy. One CTA handles kGatherRows consecutive output slots; each warp copies
// kU rows at a time, issuing every load of the batch before the first store so that ~4 KB per warp
// (32 KB per CTA, ~200 KB per SM) is in flight. Loads/stores are as wide as the row alignment allows
// (YCSB rows are 1000 B => 8-byte aligned => 64-bit accesses, 256 B contiguous per warp instruction).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace stage {

constexpr uint32_t kNoRow = 0xFFFFFFFFu;
constexpr uint32_t kVersionRowFlag = 0x80000000u;   // src_row refers to ver_payload
constexpr int kGatherRows = 64;                     // output slots per CTA
constexpr int kGatherThreads = 256;

struct GatherArgs {
    const uint32_t *src_row;     // [n_slots]
    uint64_t n_slots;
    const uint8_t *payload;      // flattened table rows
    const uint8_t *ver_payload;  // staged version rows
    uint8_t *dst;                // [n_slots][P]
    uint32_t payload_size;
    // compact destination (STAGE_IO_COMPACT): slot o = scan o / out_stride, position o % out_stride goes to row
    // dst_base[scan] + position instead of row o; nullptr = dense slab
    const uint32_t *dst_base;
    uint32_t out_stride;
};

__device__ __forceinline__ uint64_t gather_dst_row(const GatherArgs &a, uint64_t slot) {
    if (!a.dst_base) return slot;
    const uint64_t scan = slot / a.out_stride;
    return (uint64_t) __ldg(a.dst_base + scan) + (slot - scan * a.out_stride);
}

template <typename V, int U>
__global__ void __launch_bounds__(kGatherThreads) gather_rows_kernel(GatherArgs a) {
    __shared__ uint32_t s_row[kGatherRows];
    __shared__ uint64_t s_dst[kGatherRows];
    const uint64_t slot0 = (uint64_t) blockIdx.x * kGatherRows;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kGatherThreads / 32;
    if (tid < kGatherRows) {
        const uint64_t o = slot0 + tid;
        const uint32_t r = o < a.n_slots ? __ldg(a.src_row + o) : kNoRow;
        s_row[tid] = r;
        s_dst[tid] = r == kNoRow ? 0 : gather_dst_row(a, o);
    }
    __syncthreads();
    const uint32_t P = a.payload_size;
    const uint32_t nvec = P / (uint32_t) sizeof(V);
    // warp w takes rows w*U .. w*U+U-1, then (w+kWarps)*U ..., so that one warp's U rows are adjacent
    for (int r0 = warp * U; r0 < kGatherRows; r0 += kWarps * U) {
        const V *src[U];
        V *dst[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t row = s_row[r0 + u];
            const uint8_t *base = (row & kVersionRowFlag) ? a.ver_payload : a.payload;
            src[u] = row == kNoRow ? nullptr
                                   : reinterpret_cast<const V *>(base + (uint64_t) (row & ~kVersionRowFlag) * P);
            dst[u] = reinterpret_cast<V *>(a.dst + s_dst[r0 + u] * P);
        }
        // 4 x U loads per lane are issued before the first store (K = 4 covers a 1000-byte row at 8 B/lane)
        constexpr int K = 4;
        for (uint32_t c0 = lane; c0 < nvec; c0 += 32 * K) {
            V v[U][K];
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (src[u] && c0 + 32 * k < nvec) v[u][k] = __ldcs(src[u] + c0 + 32 * k);
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int k = 0; k < K; ++k)
                    if (src[u] && c0 + 32 * k < nvec) __stcs(dst[u] + c0 + 32 * k, v[u][k]);
        }
    }
}

}  // namespace stage

// ---- column projection (operator push-down) -------------------------------------------------
// The TPC-C callers of the path consume one small field of each row they read: RunStockLevel tests S_QUANTITY
// against a threshold (benchmark/tpcc/tpcc_stock_level.cpp:167-176), CH-Q2* takes MIN(S_QUANTITY)
// (benchmark/tpcc/tpcc_new_order.cpp:808-866). gather_cols_kernel moves only payload bytes
// [col_off, col_off + col_len) of every source row: one 32-byte sector per row instead of the whole row, and
// col_len instead of P bytes back to the caller — which lifts the PCIe / NVLink bound of the full-row form.
// Rows without a source (kNoRow) get zero bytes. One thread per output slot.
namespace stage {

struct GatherColsArgs {
    const uint32_t *src_row;
    uint64_t n_slots;
    const uint8_t *payload, *ver_payload;
    uint8_t *dst;                // [n_slots][col_len]
    uint32_t payload_size, col_off, col_len;
};

__global__ void __launch_bounds__(256) gather_cols_kernel(GatherColsArgs a) {
    const uint64_t o = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= a.n_slots) return;
    const uint32_t row = __ldg(a.src_row + o);
    uint8_t *d = a.dst + o * a.col_len;
    if (row == kNoRow) {
        for (uint32_t b = 0; b < a.col_len; ++b) d[b] = 0;
        return;
    }
    const uint8_t *base = (row & kVersionRowFlag) ? a.ver_payload : a.payload;
    const uint8_t *s = base + (uint64_t) (row & ~kVersionRowFlag) * a.payload_size + a.col_off;
    if ((((uintptr_t) s | (uintptr_t) d | a.col_len) & 3u) == 0) {
        for (uint32_t b = 0; b < a.col_len; b += 4) *reinterpret_cast<uint32_t *>(d + b) = __ldg(reinterpret_cast<const uint32_t *>(s + b));
    } else {
        for (uint32_t b = 0; b < a.col_len; ++b) d[b] = __ldg(s + b);
    }
}

}  // namespace stage

// ---- predicate / aggregate push-down ---------------------------------------------------------
// What the TPC-C callers finally compute from the rows they probe: RunStockLevel counts the items whose S_QUANTITY is
// below a threshold (benchmark/tpcc/tpcc_stock_level.cpp:167-176), CH-Q2* takes MIN(S_QUANTITY)
// (benchmark/tpcc/tpcc_new_order.cpp:808-866). The whole batch answers with one 24-byte record: nothing per probe
// crosses PCIe. Column = little-endian int32 at payload byte col_off (any alignment).
namespace stage {

struct AggI32 {
    unsigned long long found, below;
    int min_value, max_value;
};

__global__ void __launch_bounds__(256) aggregate_i32_kernel(const uint32_t *__restrict__ src_row, uint64_t n_slots,
                                                            const uint8_t *payload, const uint8_t *ver_payload,
                                                            uint32_t payload_size, uint32_t col_off, int threshold, AggI32 *out) {
    unsigned found = 0, below = 0;
    int mn = 0x7FFFFFFF, mx = (int) 0x80000000;
    for (uint64_t o = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; o < n_slots; o += (uint64_t) gridDim.x * blockDim.x) {
        const uint32_t row = __ldg(src_row + o);
        if (row == kNoRow) continue;
        const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
        const uint8_t *p = base + (uint64_t) (row & ~kVersionRowFlag) * payload_size + col_off;
        const int v = (int) ((uint32_t) __ldg(p) | ((uint32_t) __ldg(p + 1) << 8) | ((uint32_t) __ldg(p + 2) << 16) |
                             ((uint32_t) __ldg(p + 3) << 24));
        ++found;
        below += v < threshold ? 1u : 0u;
        mn = min(mn, v);
        mx = max(mx, v);
    }
    found = __reduce_add_sync(0xFFFFFFFFu, found);
    below = __reduce_add_sync(0xFFFFFFFFu, below);
    mn = __reduce_min_sync(0xFFFFFFFFu, mn);
    mx = __reduce_max_sync(0xFFFFFFFFu, mx);
    if ((threadIdx.x & 31) == 0 && found) {
        atomicAdd(&out->found, (unsigned long long) found);
        atomicAdd(&out->below, (unsigned long long) below);
        atomicMin(&out->min_value, mn);
        atomicMax(&out->max_value, mx);
    }
}

// One record per GROUP of consecutive probes (CH-Q2* loops over suppliers and reduces the ~300 stock rows of each,
// benchmark/tpcc/tpcc_new_order.cpp:803-866). min_index = first probe of the group that attains the minimum (what
// "the stock tuple having min. stock level" means); last_below_index = the last found probe whose value is below the
// threshold — what the reference's loop literally keeps, because its running minimum is never lowered (min_qty stays
// INT16_MAX, :810-860), with threshold = 32767. Indices count from the start of the batch; 0xFFFFFFFF = none.
struct GroupAggI32 {
    uint32_t found, below;
    int min_value;
    uint32_t min_index;
    int max_value;
    uint32_t last_below_index;
};

__global__ void __launch_bounds__(256) group_aggregate_i32_kernel(const uint32_t *__restrict__ src_row,
                                                                  const uint64_t *__restrict__ group_off, uint64_t n_groups,
                                                                  const uint8_t *payload, const uint8_t *ver_payload,
                                                                  uint32_t payload_size, uint32_t col_off, int threshold,
                                                                  GroupAggI32 *out) {
    const int lane = threadIdx.x & 31;
    const uint64_t warps = ((uint64_t) gridDim.x * blockDim.x) >> 5;
    for (uint64_t g = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; g < n_groups; g += warps) {
        const uint64_t lo = __ldg(group_off + g), hi = __ldg(group_off + g + 1);
        unsigned found = 0, below = 0, mi = 0xFFFFFFFFu, lb = 0;      // lb: index + 1, 0 = none
        int mn = 0x7FFFFFFF, mx = (int) 0x80000000;
        for (uint64_t o = lo + lane; o < hi; o += 32) {
            const uint32_t row = __ldg(src_row + o);
            if (row == kNoRow) continue;
            const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
            const uint8_t *p = base + (uint64_t) (row & ~kVersionRowFlag) * payload_size + col_off;
            const int v = (int) ((uint32_t) __ldg(p) | ((uint32_t) __ldg(p + 1) << 8) | ((uint32_t) __ldg(p + 2) << 16) |
                                 ((uint32_t) __ldg(p + 3) << 24));
            ++found;
            if (v < threshold) {
                ++below;
                lb = (unsigned) o + 1;                    // o ascends within a lane
            }
            if (v < mn) {                                 // strict: the first index wins within a lane
                mn = v;
                mi = (unsigned) o;
            }
            mx = max(mx, v);
        }
        found = __reduce_add_sync(0xFFFFFFFFu, found);
        below = __reduce_add_sync(0xFFFFFFFFu, below);
        const int wmn = __reduce_min_sync(0xFFFFFFFFu, mn);
        mi = __reduce_min_sync(0xFFFFFFFFu, (found && mn == wmn) ? mi : 0xFFFFFFFFu);
        mx = __reduce_max_sync(0xFFFFFFFFu, mx);
        lb = __reduce_max_sync(0xFFFFFFFFu, lb);
        if (lane == 0) out[g] = GroupAggI32{found, below, wmn, mi, mx, lb - 1u};
    }
}

}  // namespace stage

// ---- mid-size rows (TPC-C stock: 400 B) ---------------------------------------------------------------------------
// gather_rows_kernel is shaped for ~1 KB rows: a warp covers a row with 4 unrolled accesses per lane, so a 400-byte row
// (25 x 16 B) leaves 7 lanes idle and issues ONE of the four loads — 1.6 KB in flight per warp instead of 4 KB
// (0.47 of the HBM peak on `stock` probes). Here the output is a flat stream of 16-byte vectors:
// thread t moves vectors t, t+256, t+512, ... of its CTA's span (kMidVecs independent loads before the first store),
// consecutive threads write consecutive vectors (fully coalesced stores) and read consecutive vectors of a source row.
namespace stage {

constexpr uint32_t kMidRowMax = 512;         // rows of 272..512 bytes, 16-byte aligned

template <int kMidVecs>
__global__ void __launch_bounds__(kGatherThreads) gather_rows_mid_kernel(GatherArgs a) {
    __shared__ uint32_t s_row[kGatherRows];
    __shared__ uint64_t s_dst[kGatherRows];
    const uint64_t slot0 = (uint64_t) blockIdx.x * kGatherRows;
    const int tid = threadIdx.x;
    if (tid < kGatherRows) {
        const uint64_t o = slot0 + tid;
        const uint32_t r = o < a.n_slots ? __ldg(a.src_row + o) : kNoRow;
        s_row[tid] = r;
        s_dst[tid] = r == kNoRow ? 0 : gather_dst_row(a, o);
    }
    __syncthreads();
    const uint32_t P = a.payload_size, vpr = P / 16;          // vectors per row
    const uint32_t total = kGatherRows * vpr;                 // vectors of this CTA's 64 output rows
    for (uint32_t v0 = tid; v0 < total; v0 += kGatherThreads * kMidVecs) {
        uint4 val[kMidVecs];
        uint4 *dst[kMidVecs];
#pragma unroll
        for (int j = 0; j < kMidVecs; ++j) {
            const uint32_t v = v0 + j * kGatherThreads;
            dst[j] = nullptr;
            if (v >= total) continue;
            const uint32_t r = v / vpr, k = v - r * vpr;
            const uint32_t row = s_row[r];
            if (row == kNoRow) continue;
            const uint8_t *base = (row & kVersionRowFlag) ? a.ver_payload : a.payload;
            val[j] = __ldcs(reinterpret_cast<const uint4 *>(base + (uint64_t) (row & ~kVersionRowFlag) * P) + k);
            dst[j] = reinterpret_cast<uint4 *>(a.dst + s_dst[r] * P) + k;
        }
#pragma unroll
        for (int j = 0; j < kMidVecs; ++j)
            if (dst[j]) __stcs(dst[j], val[j]);
    }
}

}  // namespace stage

// ---- compact layout of keys / commit ids (STAGE_IO_COMPACT) -------------------------------------------------------
// The plan kernels write a scan's keys and commit ids at [scan * out_stride + position]; this moves them to
// [dst_base[scan] + position], next to each other, so that one DMA copy per array returns exactly the rows that exist.
namespace stage {

__global__ void __launch_bounds__(256) compact_keys_ts_kernel(const uint32_t *__restrict__ row_counts, const uint32_t *__restrict__ dst_base,
                                                              uint64_t n_slots, uint32_t out_stride, uint32_t key_size,
                                                              const uint8_t *key_in, uint8_t *key_out, const uint32_t *ts_in,
                                                              uint32_t *ts_out) {
    const uint64_t o = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= n_slots) return;
    const uint64_t scan = o / out_stride;
    const uint32_t pos = (uint32_t) (o - scan * out_stride);
    if (pos >= __ldg(row_counts + scan)) return;
    const uint64_t d = (uint64_t) __ldg(dst_base + scan) + pos;
    if (ts_out) ts_out[d] = ts_in[o];
    if (key_out) {
        if ((key_size & 3u) == 0)
            for (uint32_t b = 0; b < key_size; b += 4)
                *reinterpret_cast<uint32_t *>(key_out + d * key_size + b) = *reinterpret_cast<const uint32_t *>(key_in + o * key_size + b);
        else
            for (uint32_t b = 0; b < key_size; ++b) key_out[d * key_size + b] = key_in[o * key_size + b];
    }
}

}  // namespace stage

// ---- short rows (TPC-C order_line: 60 B) ------------------------------------------------------------------------
// A warp per row would use 15 of 32 lanes on a 60-byte row. Here every thread moves one 32-bit word of the output:
// consecutive threads write consecutive words (fully coalesced stores) and read consecutive words of a source row.
namespace stage {

constexpr int kSmallWordsPerThread = 4;      // independent loads per thread, each warp access still fully coalesced

// I = uint32_t when the batch has fewer than 2^32 words (32-bit index arithmetic: the divide is the hot instruction)
template <typename I>
__global__ void __launch_bounds__(256) gather_rows_small_kernel(GatherArgs a, uint32_t words_per_row, uint64_t n_words) {
    const I base_w = (I) blockIdx.x * (I) (256 * kSmallWordsPerThread) + (I) threadIdx.x;
    uint32_t val[kSmallWordsPerThread];
    uint32_t *dst[kSmallWordsPerThread];
#pragma unroll
    for (int j = 0; j < kSmallWordsPerThread; ++j) {
        const I w = base_w + (I) (j * 256);
        dst[j] = nullptr;
        if ((uint64_t) w >= n_words) continue;
        const I slot = w / (I) words_per_row;
        const uint32_t row = __ldg(a.src_row + slot);
        if (row == kNoRow) continue;
        const uint32_t k = (uint32_t) (w - slot * (I) words_per_row);
        const uint8_t *base = (row & kVersionRowFlag) ? a.ver_payload : a.payload;
        val[j] = __ldg(reinterpret_cast<const uint32_t *>(base + (uint64_t) (row & ~kVersionRowFlag) * a.payload_size) + k);
        dst[j] = reinterpret_cast<uint32_t *>(a.dst) + (a.dst_base ? gather_dst_row(a, (uint64_t) slot) * words_per_row + k : (uint64_t) w);
    }
#pragma unroll
    for (int j = 0; j < kSmallWordsPerThread; ++j)
        if (dst[j]) *dst[j] = val[j];
}

}  // namespace stage
