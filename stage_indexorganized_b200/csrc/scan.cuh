// scan.cuh — batched range scan (replaces BTree::RangeScanBySize -> Iterator::GetNext,
// include/vstore/b_tree.h:830-941, LeafNode::RangeScanBySize, src/vstore/b_tree.cpp:1261-1315, and the
// executor loops include/execute/executor.h:455-535 / :620-638).
//
// Normative behaviour reproduced exactly:
//   * start leaf = first leaf whose upper fence >= begin (GetChildIndex with get_le, b_tree.cpp:664-702);
//   * a leaf contributes its visible rows with key >= bound; the reference collects them in SLOT order
//     and stops once it holds to_scan+1 rows, THEN sorts (b_tree.cpp:1276-1309) — so when a leaf has more
//     than to_scan+1 qualifying rows, the rows returned are the to_scan smallest keys among the
//     to_scan+1 lowest-slot qualifying rows;
//   * the iterator moves to the next leaf only when its buffer runs down to one row with budget left,
//     by re-descending with le_child=false from that row's key (b_tree.h:913-926): it reaches the right
//     sibling iff that key equals the leaf's upper fence (cont[leaf]), otherwise it lands in the same
//     leaf, sees the same record first and ends the scan (b_tree.h:930-938). An empty leaf ends it too.
//
// Because the flattened leaves are contiguous in key order, those rules collapse to: the scan emits the
// contiguous row window [q0, q0+S) of the table, cut at the end of the first leaf whose cont flag is 0,
// with the slot-order selection applied only inside the leaf where the budget runs out. One CTA plans
// one scan: it stages the window's commit-id / chain-head columns in shared memory with TMA bulk copies
// (cp.async.bulk + mbarrier), evaluates the MVCC predicate one row per thread, compacts with
// ballot/prefix sums and writes the scan's source-row list, keys and commit ids; gather_rows_kernel
// (gather.cuh) then streams the payload rows of all scans at once, perfectly load-balanced.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"

namespace stage {

struct ScanArgs {
    const uint8_t *start_keys;
    const uint32_t *scan_sizes;
    uint64_t n;
    uint32_t out_stride;
    uint32_t read_ts;
    int apply_visibility;
    uint32_t *row_counts;
    uint8_t *key_out;
    uint32_t *ts_out;
    uint32_t *src_row;          // [n][out_stride] rows to gather, kNoRow beyond row_counts[i]
    const uint32_t *n_dev;      // optional: the number of scans lives on the device (min(*n_dev, n) are run)
    unsigned long long *stats;  // optional: [0] += scans whose rows were selected by the slot-order rule (Q5)
    const uint32_t *read_ts_each;    // optional [n]: read id per scan (coalesced single operations); nullptr = read_ts
    const uint8_t *visibility_each;  // optional [n]: apply_visibility per scan; nullptr = apply_visibility
};

// MVCC filter of a scanned row, exactly as the reference's executors apply it. TableScanExecutor: none
// (executor.h:620-638). IndexScanExecutor range branch (executor.h:455-535): a row is returned iff
// read_id >= the commit id of its in-leaf version. The branch that would walk the version chain of a too-new row
// reads RecordMeta::next_tuple_ptr, which is 0 for every record the iterator produces (RecordMeta's constructor
// zeroes it, record_meta.h:240; LeafNode::RangeScanBySize never sets it, b_tree.cpp:1290-1297; only BTree::Read
// does, b_tree.cpp:2099/2116) — so in a range scan a row whose latest version is newer than the reader is
// DROPPED, never replaced by an older version. Checked against the unmodified executor in
// tests/cpp/dropin_test.cpp ("range over chains"). Point probes do walk the chain (probe.cuh).
__device__ __forceinline__ bool scan_row_visible(int apply_visibility, uint32_t read_ts, uint32_t t) {
    return !apply_visibility || read_ts >= t;
}
// the filter of scan `sid`: (apply_visibility, read id)
__device__ __forceinline__ int scan_vis_of(const ScanArgs &a, uint64_t sid) {
    return a.visibility_each ? (int) a.visibility_each[sid] : a.apply_visibility;
}
__device__ __forceinline__ uint32_t scan_ts_of(const ScanArgs &a, uint64_t sid) {
    return a.read_ts_each ? a.read_ts_each[sid] : a.read_ts;
}

constexpr int kScanThreads = 256;
constexpr int kScanRowsPerThread = 4;
constexpr int kScanTile = kScanThreads * kScanRowsPerThread;     // rows staged per TMA tile

// ---- mbarrier / TMA bulk copy (PTX) -----------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (TMA, SASS UBLKCP); 16-byte aligned addresses, size multiple of 16
__device__ __forceinline__ void tma_load_1d(void *smem_dst, const void *gmem_src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// block-wide exclusive prefix sum of small per-thread counts; *total = block sum
__device__ __forceinline__ uint32_t block_exclusive_sum(uint32_t x, uint32_t *s_warp, uint32_t *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = x;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += y;
    }
    __syncthreads();                       // s_warp may still be read by the previous call
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < kScanThreads / 32; ++w) {
        const uint32_t s = s_warp[w];
        if (w < warp) base += s;
        tot += s;
    }
    *total = tot;
    return base + incl - x;
}

template <int NW>
__global__ void __launch_bounds__(kScanThreads) scan_plan_kernel(SnapshotView v, ScanArgs a) {
    __shared__ __align__(16) uint32_t s_ts[kScanTile + 8];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_min;
    extern __shared__ __align__(16) uint16_t s_slots[];          // [max_leaf_rows] for the slot-rank rule

    const int tid = threadIdx.x;
    const uint32_t K = v.key_size;
    const uint32_t L = (uint32_t) v.n_leaves;
    uint32_t parity = 0;
    if (tid == 0) mbar_init(&s_bar, 1);
    __syncthreads();

    const uint64_t n_scans = a.n_dev ? min((uint64_t) __ldcg(a.n_dev), a.n) : a.n;
    for (uint64_t sid = blockIdx.x; sid < n_scans; sid += gridDim.x) {
        const uint32_t S = min(a.scan_sizes[sid], a.out_stride);   // a scan never writes past its slab
        const SKey<NW> begin = to_sortable<NW>(a.start_keys + sid * K, K, v.xormask);
        const uint64_t out_base = sid * (uint64_t) a.out_stride;
        const int vis_on = scan_vis_of(a, sid);
        const uint32_t rts = scan_ts_of(a, sid);
        uint32_t cnt = 0;

        // emits one row (already selected) at output position `pos` of this scan
        auto emit = [&](uint32_t row, uint32_t src, uint32_t stamp, uint32_t pos) {
            const uint64_t o = out_base + pos;
            a.src_row[o] = src;
            if (a.ts_out) a.ts_out[o] = stamp;
            if (a.key_out) from_sortable<NW>(load_skey<NW>(v.level[0], row), a.key_out + o * K, K, v.xormask);
        };

        if (S > 0 && L > 0) {
            // ---- start leaf: number of upper fences < begin (fences of leaves 0..L-2) ------------------
            // CTA-wide k-ary search: every round the 256 threads test 256 evenly spaced fences of the remaining
            // range at once (3 dependent loads for 16 M leaves instead of 24 for a binary search by one thread).
            uint32_t j0 = 0;
            for (uint32_t n = L - 1; n > 0;) {
                const uint32_t step = (n + kScanThreads - 1) / kScanThreads;
                const uint64_t idx = (uint64_t) j0 + (uint64_t) (tid + 1) * step - 1;
                const bool lt = idx < (uint64_t) j0 + n && skey_less<NW>(load_skey<NW>(v.sep, idx), begin);
                const uint32_t cnt = __syncthreads_count(lt);
                const uint32_t adv = cnt * step;                 // fences j0 .. j0+adv-1 are all < begin
                n = min(step - 1, n - min(n, adv));
                j0 += adv;
            }
            if (tid == 0) s_min = 0xFFFFFFFFu;
            __syncthreads();
            const uint32_t lo0 = v.leaf_off[j0], hi0 = v.leaf_off[j0 + 1];
            uint32_t below = 0;                                   // rows of the start leaf with key < begin
            for (uint32_t base = 0; base < hi0 - lo0; base += kScanThreads) {
                const uint32_t i = base + tid;
                const bool lt = i < hi0 - lo0 && skey_less<NW>(load_skey<NW>(v.level[0], (uint64_t) lo0 + i), begin);
                below += __syncthreads_count(lt);
            }
            const uint32_t q0 = lo0 + below;
            if (hi0 > q0) {                                       // an empty first buffer ends the scan at once
                // ---- last leaf touched: first j >= j0 that stops the hop chain or exhausts the budget --
                const uint64_t need_end = (uint64_t) q0 + S;
                for (uint32_t base = j0;; base += kScanThreads) {
                    const uint32_t j = base + tid;
                    if (j < L && (!(v.cont[j] & kLeafHop) || (uint64_t) v.leaf_off[j + 1] >= need_end || j == L - 1)) atomicMin(&s_min, j);
                    __syncthreads();
                    const uint32_t m = s_min;
                    __syncthreads();
                    if (m != 0xFFFFFFFFu) break;                 // leaf L-1 always qualifies, so this terminates
                }
                const uint32_t jl = s_min;
                const uint32_t row_end = v.leaf_off[jl + 1];
                const uint32_t total = (uint64_t) row_end >= need_end ? S : row_end - q0;
                const uint32_t start_l = max(q0, v.leaf_off[jl]);
                const uint32_t T = S - (start_l - q0);            // budget left when the iterator reaches leaf jl
                const uint32_t nql = row_end - start_l;           // qualifying rows of leaf jl
                const bool quirk = nql > T + 1 && !(v.cont[jl] & kLeafSlotOrdered);   // slot-order early break bites (Q5)
                const uint32_t main_n = quirk ? start_l - q0 : total;

                // ---- rows [q0, q0 + main_n): straight window, TMA-staged commit ids / chain heads -------
                for (uint32_t off = 0; off < main_n; off += kScanTile) {
                    const uint32_t r0 = q0 + off;
                    const uint32_t n = min((uint32_t) kScanTile, main_n - off);
                    const uint32_t a0 = r0 & ~3u;                 // 16-byte aligned start of the staged window
                    const uint32_t bytes = ((r0 + n - a0) * 4u + 15u) & ~15u;
                    if (tid == 0) {
                        mbar_expect_tx(&s_bar, bytes);
                        tma_load_1d(s_ts, v.ts + a0, bytes, &s_bar);
                    }
                    mbar_wait(&s_bar, parity);
                    parity ^= 1;
                    uint32_t src[kScanRowsPerThread], stamp[kScanRowsPerThread], mine = 0;
#pragma unroll
                    for (int k = 0; k < kScanRowsPerThread; ++k) {
                        const uint32_t i = tid * kScanRowsPerThread + k;
                        src[k] = kNoRow;
                        stamp[k] = 0;
                        if (i < n) {
                            const uint32_t row = r0 + i, t = s_ts[row - a0];
                            if (scan_row_visible(vis_on, rts, t)) { src[k] = row; stamp[k] = t; }
                            mine += src[k] != kNoRow;
                        }
                    }
                    uint32_t tot;
                    uint32_t pos = cnt + block_exclusive_sum(mine, s_warp, &tot);
#pragma unroll
                    for (int k = 0; k < kScanRowsPerThread; ++k)
                        if (src[k] != kNoRow) emit(r0 + tid * kScanRowsPerThread + k, src[k], stamp[k], pos++);
                    cnt += tot;
                    __syncthreads();                              // tile fully consumed before it is overwritten
                }

                // ---- the leaf where the budget runs out, when the slot-order rule changes the result -----
                if (quirk) {
                    if (tid == 0 && a.stats) atomicAdd(a.stats, 1ull);
                    for (uint32_t i = tid; i < nql; i += kScanThreads) s_slots[i] = v.slot[start_l + i];
                    __syncthreads();
                    uint32_t sel_before = 0;
                    for (uint32_t base = 0; base < nql; base += kScanThreads) {
                        const uint32_t i = base + tid;
                        bool sel = false;
                        if (i < nql) {
                            const uint16_t me = s_slots[i];
                            uint32_t rank = 0;
                            for (uint32_t x = 0; x < nql; ++x) rank += s_slots[x] < me;
                            sel = rank < T + 1;                   // among the T+1 lowest-slot qualifying rows
                        }
                        uint32_t tot;
                        const uint32_t pre = block_exclusive_sum(sel ? 1u : 0u, s_warp, &tot);
                        sel = sel && sel_before + pre < T;        // first T of them in key order
                        sel_before += tot;
                        uint32_t src = kNoRow, stamp = 0;
                        const uint32_t row = start_l + i;
                        if (sel) {
                            const uint32_t t = __ldg(v.ts + row);
                            if (scan_row_visible(vis_on, rts, t)) { src = row; stamp = t; }
                        }
                        const uint32_t p2 = block_exclusive_sum(src != kNoRow ? 1u : 0u, s_warp, &tot);
                        if (src != kNoRow) emit(row, src, stamp, cnt + p2);
                        cnt += tot;
                    }
                }
            }
        }
        // unused tail of the scan's slab: nothing to gather
        for (uint32_t i = cnt + tid; i < a.out_stride; i += kScanThreads) a.src_row[out_base + i] = kNoRow;
        if (tid == 0) a.row_counts[sid] = cnt;
        __syncthreads();
    }
}

}  // namespace stage

// ---- physical whole-tree scan (TableScanExecutor with scan_sz == -1, include/execute/executor.h:581-619) --------
// The reference walks the tree physically: leaves in tree order, record slots in SLOT order (insertion order), and
// copies every slot. In the flattened table the rows of a leaf are in key order with their slot index kept, so the
// physical order is a per-leaf permutation: phys[leaf_off[j] + rank] = row, rank = number of rows of the leaf with a
// smaller slot. (Invisible / deleted slots, whose bytes the reference copies too — zeroed metadata, i.e. host
// garbage — are not part of a snapshot and are skipped.)
namespace stage {

__global__ void __launch_bounds__(128) physical_order_kernel(SnapshotView v, uint32_t *phys) {
    extern __shared__ uint16_t s_sl[];
    const uint32_t j = blockIdx.x;
    if (j >= v.n_leaves) return;
    const uint32_t lo = v.leaf_off[j], n = v.leaf_off[j + 1] - lo;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) s_sl[i] = v.slot[lo + i];
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) {
        const uint16_t me = s_sl[i];
        uint32_t rank = 0;
        for (uint32_t x = 0; x < n; ++x) rank += s_sl[x] < me;
        phys[lo + rank] = lo + i;
    }
}

// keys and commit ids of physical rows [first, first + n)
template <int NW>
__global__ void physical_emit_kernel(SnapshotView v, const uint32_t *phys, uint64_t first, uint64_t n, uint8_t *key_out,
                                     uint32_t *ts_out) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t row = phys[first + i];
    if (ts_out) ts_out[i] = __ldg(v.ts + row);
    if (key_out) from_sortable<NW>(load_skey<NW>(v.level[0], row), key_out + i * v.key_size, v.key_size, v.xormask);
}

}  // namespace stage

// ---- short scans: one WARP per scan ---------------------------------------------------------------------------
// RunStockLevel issues scans of 10 rows (benchmark/tpcc/tpcc_stock_level.cpp:114); a 256-thread CTA per such scan
// idles 96 % of its lanes and pays a dozen CTA barriers. scan_plan_warp_kernel runs the SAME algorithm as
// scan_plan_kernel — start leaf, rows below the bound, last leaf touched, straight window with the MVCC predicate,
// slot-order selection in the leaf where the budget runs out — with warp ballots in place of CTA barriers and
// direct loads in place of the TMA-staged tile. Chosen by the host when out_stride is small.
namespace stage {

constexpr int kScanWarpThreads = 256;          // 8 scans in flight per CTA

template <int NW>
__global__ void __launch_bounds__(kScanWarpThreads) scan_plan_warp_kernel(SnapshotView v, ScanArgs a) {
    extern __shared__ __align__(16) uint16_t s_slots_all[];      // [warps per CTA][max_leaf_rows]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr uint32_t FULL = 0xFFFFFFFFu;
    uint16_t *s_slots = s_slots_all + (size_t) warp * ((v.max_leaf_rows + 7u) & ~7u);
    const uint32_t K = v.key_size;
    const uint32_t L = (uint32_t) v.n_leaves;
    const uint64_t n_warps = (uint64_t) gridDim.x * (kScanWarpThreads / 32);
    const uint32_t lt_mask = (1u << lane) - 1u;

    const uint64_t n_scans = a.n_dev ? min((uint64_t) __ldcg(a.n_dev), a.n) : a.n;
    for (uint64_t sid = (uint64_t) blockIdx.x * (kScanWarpThreads / 32) + warp; sid < n_scans; sid += n_warps) {
        const uint32_t S = min(a.scan_sizes[sid], a.out_stride);   // a scan never writes past its slab
        const SKey<NW> begin = to_sortable<NW>(a.start_keys + sid * K, K, v.xormask);
        const uint64_t out_base = sid * (uint64_t) a.out_stride;
        const int vis_on = scan_vis_of(a, sid);
        const uint32_t rts = scan_ts_of(a, sid);
        uint32_t cnt = 0;
        auto emit = [&](uint32_t row, uint32_t src, uint32_t stamp, uint32_t pos) {
            const uint64_t o = out_base + pos;
            a.src_row[o] = src;
            if (a.ts_out) a.ts_out[o] = stamp;
            if (a.key_out) from_sortable<NW>(load_skey<NW>(v.level[0], row), a.key_out + o * K, K, v.xormask);
        };
        if (S > 0 && L > 0) {
            // A start key that IS a row (the usual case: the callers scan from keys they know) is found through the
            // point-probe directory: its row is the scan's first row and its leaf is the one the fences would select
            // (a leaf's keys lie in (sep[j-1], sep[j]]; an equal separator routes left). All lanes look the same slot up
            // (broadcast loads). Then only a 32-ary search over the u32 leaf offsets is left, instead of ~5 KB of keys.
            bool have_row = false;
            uint32_t row0 = 0;
            if (v.dir) {
                const SKey<NW> kk[1] = {begin};
                uint32_t cc[1];
                bool hh[1];
                dir_lookup<NW, 1>(v, kk, cc, hh);
                have_row = hh[0];
                row0 = cc[0];
            }
            uint32_t j0 = 0, lo0, hi0, below = 0;
            if (have_row) {
                // start leaf: number of leaves that end at or before row0
                for (uint32_t n = L; n > 0;) {
                    const uint32_t step = (n + 31) / 32;
                    const uint64_t idx = (uint64_t) j0 + (uint64_t) (lane + 1) * step - 1;
                    const bool lt = idx < (uint64_t) j0 + n && __ldg(v.leaf_off + idx + 1) <= row0;
                    const uint32_t adv = __popc(__ballot_sync(FULL, lt)) * step;
                    n = min(step - 1, n - min(n, adv));
                    j0 += adv;
                }
                lo0 = v.leaf_off[j0];
                hi0 = v.leaf_off[j0 + 1];
                below = row0 - lo0;
            } else {
                // start leaf: number of upper fences < begin, 32-ary search
                for (uint32_t n = L - 1; n > 0;) {
                    const uint32_t step = (n + 31) / 32;
                    const uint64_t idx = (uint64_t) j0 + (uint64_t) (lane + 1) * step - 1;
                    const bool lt = idx < (uint64_t) j0 + n && skey_less<NW>(load_skey<NW>(v.sep, idx), begin);
                    const uint32_t adv = __popc(__ballot_sync(FULL, lt)) * step;
                    n = min(step - 1, n - min(n, adv));
                    j0 += adv;
                }
                lo0 = v.leaf_off[j0];
                hi0 = v.leaf_off[j0 + 1];
                // rows of the start leaf with key < begin (sorted: 32-ary search)
                for (uint32_t n = hi0 - lo0; n > 0;) {
                    const uint32_t step = (n + 31) / 32;
                    const uint64_t idx = (uint64_t) below + (uint64_t) (lane + 1) * step - 1;
                    const bool lt = idx < (uint64_t) below + n && skey_less<NW>(load_skey<NW>(v.level[0], (uint64_t) lo0 + idx), begin);
                    const uint32_t adv = __popc(__ballot_sync(FULL, lt)) * step;
                    n = min(step - 1, n - min(n, adv));
                    below += adv;
                }
            }
            const uint32_t q0 = lo0 + below;
            if (hi0 > q0) {
                const uint64_t need_end = (uint64_t) q0 + S;
                uint32_t jl = L - 1;
                for (uint32_t base = j0;; base += 32) {
                    const uint32_t j = base + lane;
                    const bool stop = j < L && (!(v.cont[j] & kLeafHop) || (uint64_t) v.leaf_off[j + 1] >= need_end || j == L - 1);
                    const uint32_t m = __ballot_sync(FULL, stop);
                    if (m) { jl = base + (uint32_t) __ffs(m) - 1; break; }
                }
                const uint32_t row_end = v.leaf_off[jl + 1];
                const uint32_t total = (uint64_t) row_end >= need_end ? S : row_end - q0;
                const uint32_t start_l = max(q0, v.leaf_off[jl]);
                const uint32_t T = S - (start_l - q0);
                const uint32_t nql = row_end - start_l;
                const bool quirk = nql > T + 1 && !(v.cont[jl] & kLeafSlotOrdered);
                const uint32_t main_n = quirk ? start_l - q0 : total;
                for (uint32_t off = 0; off < main_n; off += 32) {
                    const uint32_t i = off + lane;
                    uint32_t src = kNoRow, stamp = 0;
                    const uint32_t row = q0 + i;
                    if (i < main_n) {
                        const uint32_t t = __ldg(v.ts + row);
                        if (scan_row_visible(vis_on, rts, t)) { src = row; stamp = t; }
                    }
                    const uint32_t m = __ballot_sync(FULL, src != kNoRow);
                    if (src != kNoRow) emit(row, src, stamp, cnt + __popc(m & lt_mask));
                    cnt += __popc(m);
                }
                if (quirk) {
                    if (lane == 0 && a.stats) atomicAdd(a.stats, 1ull);
                    for (uint32_t i = lane; i < nql; i += 32) s_slots[i] = v.slot[start_l + i];
                    __syncwarp();
                    // slot value of the (T+1)-th lowest slot (slots are unique): bisection on the 16-bit value, a warp
                    // count per step — O(16 n / 32) instead of ranking every row against every other
                    uint32_t vlo = 0, vhi = 0xFFFFu;
                    while (vlo < vhi) {
                        const uint32_t mid = (vlo + vhi) >> 1;
                        uint32_t c = 0;
                        for (uint32_t i = lane; i < nql; i += 32) c += s_slots[i] <= mid;
                        c = __reduce_add_sync(FULL, c);
                        if (c >= T + 1) vhi = mid; else vlo = mid + 1;
                    }
                    uint32_t sel_before = 0;
                    for (uint32_t base = 0; base < nql; base += 32) {
                        const uint32_t i = base + lane;
                        const bool in_first = i < nql && s_slots[i] <= vlo;      // among the T+1 lowest-slot qualifying rows
                        bool sel = in_first;
                        const uint32_t ms = __ballot_sync(FULL, sel);
                        sel = sel && sel_before + __popc(ms & lt_mask) < T;
                        sel_before += __popc(ms);
                        uint32_t src = kNoRow, stamp = 0;
                        const uint32_t row = start_l + i;
                        if (sel) {
                            const uint32_t t = __ldg(v.ts + row);
                            if (scan_row_visible(vis_on, rts, t)) { src = row; stamp = t; }
                        }
                        const uint32_t m2 = __ballot_sync(FULL, src != kNoRow);
                        if (src != kNoRow) emit(row, src, stamp, cnt + __popc(m2 & lt_mask));
                        cnt += __popc(m2);
                    }
                    __syncwarp();
                }
            }
        }
        for (uint32_t i = cnt + lane; i < a.out_stride; i += 32) a.src_row[out_base + i] = kNoRow;
        if (lane == 0) a.row_counts[sid] = cnt;
    }
}

}  // namespace stage
