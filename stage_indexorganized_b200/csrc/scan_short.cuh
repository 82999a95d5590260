// scan_short.cuh — short range scans of short rows, one kernel: plan AND rows.
//
// RunStockLevel issues 20 scans of 10 order_line rows per transaction (benchmark/tpcc/tpcc_stock_level.cpp:100-126;
// order_line: 32-byte key, 60-byte payload). For such scans the two-kernel form (scan_plan_warp_kernel, then
// gather_rows_small_kernel) is bound by the plan kernel's chain of dependent loads, not by bytes: ~10 dependent loads per
// scan with one scan per warp in flight (plan 0.638 ms + gather 0.320 ms per 2^20 scans = 0.32 of the
// HBM peak). This kernel
//   * runs one scan per TILE of 16 (or 32) lanes, so twice as many scans are in flight per SM for scans of <= 16 rows;
//   * cuts the chain: a start key that is a row is found through the point-probe directory and its leaf through a
//     row -> leaf table (one load instead of a 32-ary search over the leaf offsets); the commit ids of the window are
//     requested as soon as the first row is known, in parallel with the leaf bookkeeping;
//   * emits the result itself — keys, commit ids and payload rows of the selected rows — so no source-row list is written
//     and read back, and no second launch is needed.
// The scan semantics are those of scan_plan_warp_kernel line by line (start leaf, hop chain, slot-order truncation in
// the leaf where the budget runs out, MVCC range rule); tests compare both against the CPU checker.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"
#include "scan.cuh"

namespace stage {

constexpr int kShortThreads = 256;
constexpr uint32_t kShortMaxRow = 256;       // payload bytes; rows are moved word by word

// row -> leaf table (built once per snapshot, on first use): row_leaf[r] = j with leaf_off[j] <= r < leaf_off[j+1]
__global__ void __launch_bounds__(128) row_leaf_build_kernel(const uint32_t *__restrict__ leaf_off, uint32_t n_leaves, uint32_t *row_leaf) {
    for (uint32_t j = blockIdx.x; j < n_leaves; j += gridDim.x) {
        const uint32_t lo = leaf_off[j], hi = leaf_off[j + 1];
        for (uint32_t r = lo + threadIdx.x; r < hi; r += blockDim.x) row_leaf[r] = j;
    }
}

struct ShortScanOut {
    uint8_t *payload_out;        // [n][out_stride][P] or nullptr
    const uint32_t *row_leaf;    // [n_rows]
};

template <int NW, int TW, int MINB>
__global__ void __launch_bounds__(kShortThreads, MINB) scan_short_fused_kernel(SnapshotView v, ScanArgs a, ShortScanOut o) {
    constexpr int kTiles = kShortThreads / TW;                   // scans in flight per CTA
    extern __shared__ __align__(16) uint16_t s_slots_all[];      // [kTiles][max_leaf_rows] for the slot-order rule
    __shared__ uint32_t s_src[kTiles][TW], s_tsv[kTiles][TW];
    const int lane = threadIdx.x & 31;
    const int tl = lane % TW, tbase = lane - tl;                 // lane inside the tile, first lane of the tile
    const int tile = threadIdx.x / TW;
    const uint32_t tmask = TW == 32 ? 0xFFFFFFFFu : (((1u << TW) - 1u) << tbase);
    const uint32_t lt_mask = (1u << tl) - 1u;
    auto tballot = [&](bool p) -> uint32_t { return (__ballot_sync(tmask, p) >> tbase) & (TW == 32 ? 0xFFFFFFFFu : ((1u << TW) - 1u)); };
    uint16_t *s_slots = s_slots_all + (size_t) tile * ((v.max_leaf_rows + 7u) & ~7u);
    uint32_t *my_src = s_src[tile], *my_ts = s_tsv[tile];
    const uint32_t K = v.key_size, P = v.payload_size, wpr = P / 4;
    const uint32_t L = (uint32_t) v.n_leaves;
    const uint64_t n_tiles = (uint64_t) gridDim.x * kTiles;
    const uint32_t *pay32 = reinterpret_cast<const uint32_t *>(v.payload);

    const uint64_t n_scans = a.n_dev ? min((uint64_t) __ldcg(a.n_dev), a.n) : a.n;
    for (uint64_t sid = (uint64_t) blockIdx.x * kTiles + tile; sid < n_scans; sid += n_tiles) {
        const uint32_t S = min(a.scan_sizes[sid], a.out_stride);
        SKey<NW> begin;                            // K == 4 NW and 4-byte aligned keys (checked by the host): word path only
        {
            const uint32_t *k32 = reinterpret_cast<const uint32_t *>(a.start_keys + sid * K);
#pragma unroll
            for (int i = 0; i < NW; ++i) begin.w[i] = bswap32(__ldg(k32 + i)) ^ v.xormask;
        }
        const uint64_t out_base = sid * (uint64_t) a.out_stride;
        const int vis_on = scan_vis_of(a, sid);
        const uint32_t rts = scan_ts_of(a, sid);
        uint32_t cnt = 0;
        if (S > 0 && L > 0) {
            bool have_row = false;
            uint32_t row0 = 0, j0 = 0;
            uint32_t t_spec = 0;                  // commit id of row q0 + tl, requested as soon as the first row is known
            if (v.dir) {
                // directory probe (probe.cuh dir_lookup) with the follow-up loads issued together: on a tag match the key
                // to verify, the row's leaf and the window's commit ids are requested in one round
                uint32_t h = __umulhi(skey_hash<NW>(begin, 0x9E3779B9u), v.dir_cap);
                const uint32_t tag = skey_tag<NW>(begin);
                for (;;) {
                    const uint2 e = __ldg(v.dir + h);
                    if (e.y == kNoRow) break;
                    if (e.x == tag) {
                        const uint32_t jj = __ldg(o.row_leaf + e.y);
                        const uint32_t tt = (uint64_t) e.y + tl < v.n_rows ? __ldg(v.ts + e.y + tl) : 0u;
                        if (NW == 1 || skey_equal<NW>(load_skey<NW>(v.level[0], e.y), begin)) {
                            have_row = true;
                            row0 = e.y;
                            j0 = jj;
                            t_spec = tt;
                            break;
                        }
                    }
                    h = h + 1 == v.dir_cap ? 0 : h + 1;
                }
            }
            uint32_t q0 = row0;
            bool nonempty = true;                 // the start leaf holds a row >= begin
            if (!have_row) {
                uint32_t below = 0;
                for (uint32_t n = L - 1; n > 0;) {                  // start leaf: number of upper fences < begin
                    const uint32_t step = (n + TW - 1) / TW;
                    const uint64_t idx = (uint64_t) j0 + (uint64_t) (tl + 1) * step - 1;
                    const bool lt = idx < (uint64_t) j0 + n && skey_less<NW>(load_skey<NW>(v.sep, idx), begin);
                    const uint32_t adv = __popc(tballot(lt)) * step;
                    n = min(step - 1, n - min(n, adv));
                    j0 += adv;
                }
                const uint32_t lo0 = v.leaf_off[j0], hi0 = v.leaf_off[j0 + 1];
                for (uint32_t n = hi0 - lo0; n > 0;) {              // rows of the start leaf with key < begin
                    const uint32_t step = (n + TW - 1) / TW;
                    const uint64_t idx = (uint64_t) below + (uint64_t) (tl + 1) * step - 1;
                    const bool lt = idx < (uint64_t) below + n && skey_less<NW>(load_skey<NW>(v.level[0], (uint64_t) lo0 + idx), begin);
                    const uint32_t adv = __popc(tballot(lt)) * step;
                    n = min(step - 1, n - min(n, adv));
                    below += adv;
                }
                q0 = lo0 + below;
                nonempty = hi0 > q0;
            }
            if (nonempty) {
                const uint64_t need_end = (uint64_t) q0 + S;
                // leaves j0 .. j0+TW-1 in one round: offsets and flags per lane, the last leaf touched by ballot, its
                // bounds by shuffle (a window of <= TW rows rarely spans more leaves; then the loop goes on)
                uint32_t jl = L - 1, row_end = 0, lo_jl = 0, cont_jl = 0;
                for (uint32_t base = j0;; base += TW) {
                    const uint32_t j = base + tl;
                    const uint32_t lo_t = j < L ? __ldg(v.leaf_off + j) : 0u, hi_t = j < L ? __ldg(v.leaf_off + j + 1) : 0u;
                    const uint32_t c_t = j < L ? (uint32_t) __ldg(v.cont + j) : 0u;
                    const bool stop = j < L && (!(c_t & kLeafHop) || (uint64_t) hi_t >= need_end || j == L - 1);
                    const uint32_t m = tballot(stop);
                    if (m) {
                        const int src = tbase + __ffs(m) - 1;
                        jl = base + (uint32_t) __ffs(m) - 1;
                        row_end = __shfl_sync(tmask, hi_t, src);
                        lo_jl = __shfl_sync(tmask, lo_t, src);
                        cont_jl = __shfl_sync(tmask, c_t, src);
                        break;
                    }
                }
                (void) jl;
                const uint32_t total = (uint64_t) row_end >= need_end ? S : row_end - q0;
                const uint32_t start_l = max(q0, lo_jl);
                const uint32_t T = S - (start_l - q0);
                const uint32_t nql = row_end - start_l;
                const bool quirk = nql > T + 1 && !(cont_jl & kLeafSlotOrdered);
                const uint32_t main_n = quirk ? start_l - q0 : total;        // <= S <= TW: one pass
                {
                    uint32_t src = kNoRow, stamp = 0;
                    const uint32_t row = q0 + tl;
                    if ((uint32_t) tl < main_n) {
                        const uint32_t t = have_row ? t_spec : __ldg(v.ts + row);
                        if (scan_row_visible(vis_on, rts, t)) { src = row; stamp = t; }
                    }
                    const uint32_t m = tballot(src != kNoRow);
                    if (src != kNoRow) {
                        const uint32_t pos = __popc(m & lt_mask);
                        my_src[pos] = src;
                        my_ts[pos] = stamp;
                    }
                    cnt = __popc(m);
                }
                if (quirk) {
                    if (tl == 0 && a.stats) atomicAdd(a.stats, 1ull);
                    for (uint32_t i = tl; i < nql; i += TW) s_slots[i] = v.slot[start_l + i];
                    __syncwarp(tmask);
                    uint32_t vlo = 0, vhi = 0xFFFFu;              // slot value of the (T+1)-th lowest slot: bisection
                    while (vlo < vhi) {
                        const uint32_t mid = (vlo + vhi) >> 1;
                        uint32_t c = 0;
                        for (uint32_t i = tl; i < nql; i += TW) c += s_slots[i] <= mid;
#pragma unroll
                        for (int d = TW / 2; d > 0; d >>= 1) c += __shfl_xor_sync(tmask, c, d);
                        if (c >= T + 1) vhi = mid; else vlo = mid + 1;
                    }
                    uint32_t sel_before = 0;
                    for (uint32_t base = 0; base < nql; base += TW) {
                        const uint32_t i = base + tl;
                        bool sel = i < nql && s_slots[i] <= vlo;
                        const uint32_t ms = tballot(sel);
                        sel = sel && sel_before + __popc(ms & lt_mask) < T;
                        sel_before += __popc(ms);
                        uint32_t src = kNoRow, stamp = 0;
                        const uint32_t row = start_l + i;
                        if (sel) {
                            const uint32_t t = __ldg(v.ts + row);
                            if (scan_row_visible(vis_on, rts, t)) { src = row; stamp = t; }
                        }
                        const uint32_t m2 = tballot(src != kNoRow);
                        if (src != kNoRow) {
                            const uint32_t pos = cnt + __popc(m2 & lt_mask);
                            my_src[pos] = src;
                            my_ts[pos] = stamp;
                        }
                        cnt += __popc(m2);
                    }
                }
            }
        }
        __syncwarp(tmask);
        // ---- emit: count, commit ids, source rows (when a caller wants them), keys, payload rows ---------------
        if (tl == 0) a.row_counts[sid] = cnt;
        if (a.ts_out && (uint32_t) tl < cnt) a.ts_out[out_base + tl] = my_ts[tl];
        if (a.src_row && (uint32_t) tl < a.out_stride) a.src_row[out_base + tl] = (uint32_t) tl < cnt ? my_src[tl] : kNoRow;
        // The kernel is ISSUE-bound, not memory-bound (ncu r2: 76 % issue-slot utilisation, DRAM at 22 %; the newest orders of
        // a district are L2-resident), so the emit is written for few instructions: when no row was dropped the selected
        // rows are consecutive in the table, and keys / payload are flat runs copied with one index (no per-row address
        // arithmetic). (Holding a whole scan's words in registers before the first store was measured and is slower.)
        const bool contig = cnt > 0 && my_src[cnt - 1] - my_src[0] == cnt - 1;        // rows ascend: consecutive iff so
        if (a.key_out) {
            if (contig && NW % 4 == 0) {
                const uint4 *kin = reinterpret_cast<const uint4 *>(v.level[0] + (uint64_t) my_src[0] * NW);
                uint4 *ko4 = reinterpret_cast<uint4 *>(reinterpret_cast<uint32_t *>(a.key_out) + out_base * NW);
                for (uint32_t it = tl; it < cnt * (NW / 4); it += TW) {
                    uint4 q = __ldg(kin + it);
                    q.x = bswap32(q.x ^ v.xormask); q.y = bswap32(q.y ^ v.xormask);
                    q.z = bswap32(q.z ^ v.xormask); q.w = bswap32(q.w ^ v.xormask);
                    ko4[it] = q;
                }
            } else {
                uint32_t *ko = reinterpret_cast<uint32_t *>(a.key_out) + out_base * NW;
                for (uint32_t it = tl; it < cnt * NW; it += TW) {
                    const uint32_t r = it / NW, w = it % NW;
                    ko[it] = bswap32(__ldg(v.level[0] + (uint64_t) my_src[r] * NW + w) ^ v.xormask);
                }
            }
        }
        if (o.payload_out) {
            uint32_t *po = reinterpret_cast<uint32_t *>(o.payload_out) + out_base * wpr;
            if (contig) {
                const uint32_t *pin = pay32 + (uint64_t) my_src[0] * wpr;
                const uint32_t n_words = cnt * wpr;
                for (uint32_t it = tl; it < n_words; it += 4 * TW) {
                    uint32_t val[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (it + u * TW < n_words) val[u] = __ldg(pin + it + u * TW);
#pragma unroll
                    for (int u = 0; u < 4; ++u)
                        if (it + u * TW < n_words) po[it + u * TW] = val[u];
                }
            } else {
                for (uint32_t r0 = 0; r0 < cnt; r0 += 4) {
                    for (uint32_t w0 = tl; w0 < wpr; w0 += TW) {
                        uint32_t val[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (r0 + u < cnt) val[u] = __ldg(pay32 + (uint64_t) my_src[r0 + u] * wpr + w0);
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (r0 + u < cnt) po[(uint64_t) (r0 + u) * wpr + w0] = val[u];
                    }
                }
            }
        }
        __syncwarp(tmask);                                        // my_src / my_ts are reused by the next scan
    }
}

}  // namespace stage
