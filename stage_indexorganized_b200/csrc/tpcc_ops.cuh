// tpcc_ops.cuh — RunStockLevel's read side as one device-side operator (the remaining half).
//
// Reference (benchmark/tpcc/tpcc_stock_level.cpp:37-196): for each of the 20 orders below the district's next order id
//   1. IndexScanExecutor range scan of 10 order_line rows from {w, d, o, 5} (MVCC range rule); the predicate keeps
//      OL_I_ID of the returned rows that still belong to (w, d, o)                                   (:100-150)
//   2. if any: point lookup of stock {w, FIRST such OL_I_ID} (MVCC point rule, chains walked)        (:152-186)
//   3. if found and S_QUANTITY < threshold: insert S_I_ID into a set                                 (:187-190)
// and the answer is the size of the set. Through the plain ABI that is three calls with item ids round-tripping
// through the caller; here the start keys are built, the scans planned (scan_short.cuh), the first matching row of
// every scan picked, stock probed through its directory (probe.cuh) and the distinct count taken without anything
// but the 4-byte count per transaction leaving the device.
//
// Layouts (benchmark/tpcc/tpcc_record.h): order_line key {int64 OL_W_ID, OL_D_ID, OL_O_ID, OL_NUMBER} (32 B, :694-739),
// payload starts with int32 OL_I_ID (:752); stock key {int64 S_W_ID, S_I_ID} (16 B, :438-459), payload starts with
// int32 S_QUANTITY (:471). Keys >= 16 bytes are memcmp-ordered raw little-endian bytes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"

namespace stage {

constexpr int kStockLevelOrders = 20;      // tpcc_stock_level.cpp:97-98
constexpr int kStockLevelScan = 10;        // :114
constexpr int64_t kStockLevelFirstLine = 5;  // start key {w, d, o, 5}, :101

struct StockLevelReq {                     // == stage_stock_level_req_t
    int64_t w_id, d_id, next_o_id;
    int32_t threshold, reserved;
};

// scan q = txn * 20 + k starts at {w, d, next_o_id - 20 + k, 5}
__global__ void __launch_bounds__(256) stock_level_keys_kernel(const StockLevelReq *__restrict__ reqs, uint64_t n_txn, int64_t *start_keys,
                                                               uint32_t *sizes) {
    const uint64_t q = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_txn * kStockLevelOrders) return;
    const StockLevelReq r = reqs[q / kStockLevelOrders];
    int64_t *k = start_keys + q * 4;
    k[0] = r.w_id;
    k[1] = r.d_id;
    k[2] = r.next_o_id - kStockLevelOrders + (int64_t) (q % kStockLevelOrders);
    k[3] = kStockLevelFirstLine;
    sizes[q] = kStockLevelScan;
}

// first returned row of scan q that belongs to (w, d, o) -> stock probe key {w, OL_I_ID}; {-1, -1} (never a row) if none
__global__ void __launch_bounds__(256) stock_level_pick_kernel(SnapshotView ol, const int64_t *__restrict__ start_keys,
                                                               const uint32_t *__restrict__ row_counts, const uint32_t *__restrict__ src_row,
                                                               uint64_t n_scans, int64_t *stock_keys) {
    const uint64_t q = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_scans) return;
    const SKey<8> want = to_sortable<8>(reinterpret_cast<const uint8_t *>(start_keys + q * 4), 32, ol.xormask);
    const uint32_t c = min(row_counts[q], (uint32_t) kStockLevelScan);
    int64_t item = -1;
    for (uint32_t p = 0; p < c; ++p) {
        const uint32_t row = src_row[q * kStockLevelScan + p];
        const uint4 a = __ldg(reinterpret_cast<const uint4 *>(ol.level[0] + (uint64_t) row * 8));
        const uint2 b = __ldg(reinterpret_cast<const uint2 *>(ol.level[0] + (uint64_t) row * 8 + 4));
        if (a.x == want.w[0] && a.y == want.w[1] && a.z == want.w[2] && a.w == want.w[3] && b.x == want.w[4] && b.y == want.w[5]) {
            item = (int64_t) * reinterpret_cast<const int32_t *>(ol.payload + (uint64_t) row * ol.payload_size);
            break;
        }
    }
    stock_keys[q * 2] = item < 0 ? -1 : start_keys[q * 4];
    stock_keys[q * 2 + 1] = item;
}

// one thread per transaction: S_QUANTITY of the visible stock versions against the threshold, distinct item count
__global__ void __launch_bounds__(128) stock_level_count_kernel(const StockLevelReq *__restrict__ reqs, uint64_t n_txn,
                                                                const int64_t *__restrict__ stock_keys, const uint32_t *__restrict__ stock_src,
                                                                const uint8_t *payload, const uint8_t *ver_payload, uint32_t payload_size,
                                                                uint32_t *distinct_out) {
    const uint64_t t = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_txn) return;
    const int32_t threshold = reqs[t].threshold;
    int64_t seen[kStockLevelOrders];
    uint32_t n = 0;
    for (int k = 0; k < kStockLevelOrders; ++k) {
        const uint64_t q = t * kStockLevelOrders + k;
        const uint32_t row = stock_src[q];
        if (row == kNoRow) continue;
        const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
        const int32_t qty = *reinterpret_cast<const int32_t *>(base + (uint64_t) (row & ~kVersionRowFlag) * payload_size);
        if (qty >= threshold) continue;
        const int64_t item = stock_keys[q * 2 + 1];
        bool dup = false;
        for (uint32_t i = 0; i < n; ++i) dup = dup || seen[i] == item;
        if (!dup) seen[n++] = item;
    }
    distinct_out[t] = n;
}

}  // namespace stage
