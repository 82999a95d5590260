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
//     sibling iff that key equals the leaf's upper fence, otherwise it lands in the same leaf, sees the
//     same record first and ends the scan (b_tree.h:930-938). An empty leaf ends the scan too.
// One CTA per scan walks the flattened leaves from the start leaf: per leaf it evaluates the qualify /
// slot-rank / MVCC predicates one row per thread, compacts with ballot + prefix sums, and the warps copy
// the selected payload rows (coalesced vector loads/stores) into the scan's output slab.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
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
    uint8_t *payload_out;
};

constexpr int kScanThreads = 256;

// block-wide exclusive prefix sum of a 0/1 flag; returns this thread's prefix, *total = block sum
__device__ __forceinline__ uint32_t block_prefix(bool flag, uint32_t *s_warp, uint32_t *total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, flag);
    const uint32_t within = __popc(m & ((1u << lane) - 1u));
    __syncthreads();                       // protect s_warp reuse
    if (lane == 0) s_warp[warp] = __popc(m);
    __syncthreads();
    uint32_t base = 0, tot = 0;
    const int nw = blockDim.x >> 5;
    for (int w = 0; w < nw; ++w) {
        const uint32_t x = s_warp[w];
        if (w < warp) base += x;
        tot += x;
    }
    *total = tot;
    return base + within;
}

template <int NW>
__global__ void __launch_bounds__(kScanThreads) scan_kernel(SnapshotView v, ScanArgs a) {
    __shared__ uint32_t s_warp[kScanThreads / 32];
    __shared__ uint32_t s_leaf;
    __shared__ const uint8_t *s_src[kScanThreads];
    extern __shared__ __align__(16) uint16_t s_slots[];          // [max_leaf_rows] for the slot-rank quirk

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
    const uint32_t K = v.key_size, P = v.payload_size;
    const uintptr_t align = (uintptr_t) v.payload | (uintptr_t) v.ver_payload | (uintptr_t) a.payload_out | P;

    for (uint64_t sid = blockIdx.x; sid < a.n; sid += gridDim.x) {
        const uint32_t S = a.scan_sizes[sid];
        const SKey<NW> begin = to_sortable<NW>(a.start_keys + sid * K, K, v.xormask);
        const uint64_t out_base = sid * (uint64_t) a.out_stride;

        // locate the start leaf: number of upper fences < begin (fences of leaves 0..L-2)
        if (tid == 0) {
            uint64_t lo = 0, hi = v.n_leaves ? v.n_leaves - 1 : 0;
            while (lo < hi) {
                const uint64_t mid = (lo + hi) >> 1;
                if (skey_less<NW>(load_skey<NW>(v.sep, mid), begin)) lo = mid + 1; else hi = mid;
            }
            s_leaf = (uint32_t) lo;
        }
        __syncthreads();
        uint32_t j = s_leaf;
        uint32_t remaining = S, cnt = 0;
        bool first = true;

        while (remaining > 0 && j < v.n_leaves) {
            const uint32_t lo = v.leaf_off[j], hi = v.leaf_off[j + 1];
            const uint32_t nleaf = hi - lo;
            // rows of this leaf with key >= bound. Later leaves: every key is > the previous leaf's keys.
            uint32_t q = 0;
            if (first) {
                uint32_t below = 0;
                for (uint32_t base = 0; base < nleaf; base += blockDim.x) {
                    const uint32_t i = base + tid;
                    const bool lt = i < nleaf && skey_less<NW>(load_skey<NW>(v.level[0], (uint64_t) lo + i), begin);
                    below += __syncthreads_count(lt);
                }
                q = below;
            }
            const uint32_t q0 = lo + q, nq = hi - q0, T = remaining;
            if (nq == 0) break;                                   // empty buffer: every later GetNext is null
            const uint32_t take = nq < T ? nq : T;
            const bool quirk = nq > T + 1;                        // slot-order early break bites

            if (quirk) {
                for (uint32_t i = tid; i < nq; i += blockDim.x) s_slots[i] = v.slot[q0 + i];
                __syncthreads();
            }
            // emit `take` rows in key order, one row per thread per pass
            uint32_t emitted_sel = 0;                              // selected rows seen so far (quirk path)
            for (uint32_t base = 0; base < nq; base += blockDim.x) {
                const uint32_t i = base + tid;
                bool sel = i < nq;
                if (quirk) {
                    if (sel) {
                        const uint16_t me = s_slots[i];
                        uint32_t rank = 0;
                        for (uint32_t x = 0; x < nq; ++x) rank += s_slots[x] < me;
                        sel = rank < T + 1;                        // among the T+1 lowest-slot qualifying rows
                    }
                    uint32_t tot;
                    const uint32_t pre = block_prefix(sel, s_warp, &tot);
                    sel = sel && (emitted_sel + pre) < T;          // first T of them in key order
                    emitted_sel += tot;
                } else {
                    sel = sel && i < take;
                }
                // MVCC per selected row
                const uint8_t *src = nullptr;
                uint32_t stamp = 0;
                const uint32_t row = q0 + i;
                if (sel) {
                    if (a.apply_visibility) {
                        src = visible_version(v, row, a.read_ts, &stamp);
                    } else {
                        stamp = __ldg(v.ts + row);
                        src = v.payload + (uint64_t) row * P;
                    }
                }
                uint32_t tot;
                const uint32_t pre = block_prefix(src != nullptr, s_warp, &tot);
                if (src) {
                    const uint64_t o = out_base + cnt + pre;
                    if (a.ts_out) a.ts_out[o] = stamp;
                    if (a.key_out) {
                        const SKey<NW> kk = load_skey<NW>(v.level[0], row);
                        from_sortable<NW>(kk, a.key_out + o * K, K, v.xormask);
                    }
                    s_src[pre] = src;
                }
                __syncthreads();
                if (a.payload_out) {
                    for (uint32_t r = warp; r < tot; r += nwarps) {
                        const uint8_t *s1[1] = { s_src[r] };
                        uint8_t *d1[1] = { a.payload_out + (out_base + cnt + r) * P };
                        if ((align & 15) == 0) warp_copy_rows<uint4, 1>(s1, d1, P, lane);
                        else if ((align & 7) == 0) warp_copy_rows<uint2, 1>(s1, d1, P, lane);
                        else if ((align & 3) == 0) warp_copy_rows<uint32_t, 1>(s1, d1, P, lane);
                        else warp_copy_rows<uint8_t, 1>(s1, d1, P, lane);
                    }
                }
                cnt += tot;
                __syncthreads();
                if (!quirk && base + blockDim.x >= take) break;
            }
            remaining -= take;
            // hop: only if the whole leaf was consumed and its max visible key is the parent separator
            if (take < nq || !v.cont[j]) break;
            ++j;
            first = false;
        }
        if (tid == 0) a.row_counts[sid] = cnt;
        __syncthreads();
    }
}

}  // namespace stage
