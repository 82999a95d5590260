// probe.cuh — batched point probe (replaces BTree::Read, src/vstore/b_tree.cpp:2066-2129, plus the
// point branch of IndexScanExecutor::Execute, include/execute/executor.h:374-454).
//
// The reference descends root->leaf with a binary search per inner node (InternalNode::GetChildIndex,
// b_tree.cpp:664-702) and then scans the leaf's slots linearly (BaseNode::SearchRecordMeta,
// b_tree.cpp:18-122). Keys are unique (LeafNode::CheckUnique, b_tree.cpp:1395-1417) and every leaf's
// keys lie inside its fence range, so "the first visible slot with an equal key in the leaf the fences
// route to" is exactly "the equal key in the globally sorted key column".
//
// Search structure = implicit 4-ary fence levels over the key column: level l+1 holds every 4th key of
// level l (the last key of each group), down to a top level small enough for shared memory. One lane
// per probe: the top level is binary-searched in shared memory (the warp's 32 probes proceed
// together, no global traffic), then every lower level costs ONE 16-byte load — the 4 keys of the
// group the level above selected — and 4 compares. (A warp-cooperative 32-ary descent was measured
// first: it is issue-bound, ~130 warp instructions per probe, 0.19 ms per 2^20 probes; profiles/.)
// Then the MVCC rule (executor.h:383-446) picks the in-leaf version or walks the staged chain. The
// kernel emits found / commit id / key and the source row of every probe; gather_rows_kernel
// (gather.cuh) then streams the payload rows.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"

namespace stage {

constexpr int kMaxLevels = 16;
constexpr int kFan = 4;             // keys per fence group: 4 x 4-byte keys = one 16-byte load
constexpr int kProbeTopBytes = 12 * 1024;   // top fence level kept in shared memory (8 CTAs/SM stay resident)

struct SnapshotView {
    const uint32_t *level[kMaxLevels];   // level[0] = key column; level[l+1][i] = level[l][kFan*(i+1)-1]
    uint32_t level_n[kMaxLevels];
    int n_levels;                        // level[n_levels-1] is the top level (cached in shared memory)
    uint32_t smem_words;                 // 32-bit words of the top level
    const uint16_t *slot;
    const uint32_t *ts;
    const uint32_t *chain;
    const uint8_t *payload;
    const uint32_t *leaf_off;
    const uint32_t *sep;
    const uint8_t *cont;
    const uint32_t *ver_begin, *ver_commit, *ver_next;
    const uint8_t *ver_payload;
    uint64_t n_rows, n_leaves;
    uint32_t key_size, payload_size, xormask, max_leaf_rows;
    const uint2 *dir;                    // point-probe directory (below), nullptr = fence search only
    uint32_t dir_cap;
};

struct ProbeArgs {
    const uint8_t *keys;
    uint64_t n;
    uint32_t read_ts;
    uint8_t *found;
    uint32_t *ts_out;
    uint8_t *key_out;
    uint32_t *src_row;   // [n] row to gather (gather.cuh encoding), kNoRow on a miss
    const uint32_t *read_ts_each;   // optional [n]: a read id per probe (coalesced single operations of different
                                    // transactions, csrc/stage_gpu.cu "batching front-end"); nullptr = read_ts for all
};

// MVCC visibility (executor.h:383-446): the version of `row` visible at read_ts, in gather.cuh's
// source-row encoding (kNoRow if none), and its stamp. `t` = commit id of the in-leaf version.
__device__ __forceinline__ uint32_t visible_version(const SnapshotView &v, uint32_t row, uint32_t t, uint32_t read_ts,
                                                    uint32_t *stamp) {
    if (read_ts >= t) {
        *stamp = t;
        return row;
    }
    uint32_t c = __ldg(v.chain + row);            // 1-based index into the staged version table
    while (c != 0) {
        const uint32_t b = __ldg(v.ver_begin + (c - 1)), e = __ldg(v.ver_commit + (c - 1));
        if (b == 0 || e == 0) return kNoRow;      // INVALID_CID stops the walk (executor.h:418)
        if (read_ts >= b && read_ts <= e) {
            *stamp = b;
            return (c - 1) | kVersionRowFlag;
        }
        c = __ldg(v.ver_next + (c - 1));
    }
    *stamp = 0;
    return kNoRow;
}

// One fence group: the kFan keys [grp*kFan, grp*kFan+kFan) of a level. Levels (and the key column) are
// padded with all-ones keys, so no bounds checks are needed: cnt = keys < k, eq = some key == k.
template <int NW>
__device__ __forceinline__ uint32_t group_count_less(const uint32_t *lvl, uint32_t grp, const SKey<NW> &k, bool *eq) {
    uint32_t cnt;
    if constexpr (NW == 1) {
        const uint4 q = __ldg(reinterpret_cast<const uint4 *>(lvl) + grp);
        const uint32_t kk = k.w[0];
        cnt = (q.x < kk) + (q.y < kk) + (q.z < kk) + (q.w < kk);
        *eq = (q.x == kk) | (q.y == kk) | (q.z == kk) | (q.w == kk);
    } else {
        cnt = 0;
        bool e = false;
#pragma unroll
        for (int i = 0; i < kFan; ++i) {
            const SKey<NW> x = load_skey<NW>(lvl, (uint64_t) grp * kFan + i);
            cnt += skey_less<NW>(x, k);
            e = e || skey_equal<NW>(x, k);
        }
        *eq = e;
    }
    return cnt;
}

constexpr int kProbeThreads = 256;

// ---- point-probe directory --------------------------------------------------------------------------------
// Keys are unique and a snapshot's key column never changes after the seal, so a point probe does not need the
// ordered descent at all: an open-addressing table (linear probing, load <= 0.5) of (tag, row) pairs over the key
// column answers it in ~1.5 dependent loads instead of one per fence level (6 global levels at 10 M rows, 9 at
// 100 M). For 4-byte keys the tag IS the sortable key (no verification read); longer keys carry a 32-bit tag and
// are verified against the key column on a tag match. The ordered fences stay: scans need them, and routing a
// probe to "the leaf the fences select" and to "the row holding the equal key" is the same row (see the header).
// Built once at seal time by dir_build_kernel; 16 B per row of HBM.
template <int NW>
__device__ __forceinline__ uint32_t skey_hash(const SKey<NW> &k, uint32_t seed) {
    uint32_t h = seed;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
        h ^= k.w[w];
        h *= 0x85EBCA6Bu;
        h ^= h >> 15;
    }
    h *= 0xC2B2AE35u;
    h ^= h >> 16;
    return h;
}
template <int NW>
__device__ __forceinline__ uint32_t skey_tag(const SKey<NW> &k) {
    if constexpr (NW == 1) return k.w[0];
    else return skey_hash<NW>(k, 0x27D4EB2Fu);
}
constexpr unsigned long long kDirEmpty = ~0ull;          // row half == kNoRow

template <int NW>
__global__ void dir_build_kernel(const uint32_t *__restrict__ skey, uint64_t n_rows, unsigned long long *dir, uint32_t cap) {
    const uint64_t r = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const SKey<NW> k = load_skey<NW>(skey, r);
    uint32_t h = __umulhi(skey_hash<NW>(k, 0x9E3779B9u), cap);
    const unsigned long long mine = ((unsigned long long) (uint32_t) r << 32) | skey_tag<NW>(k);
    while (atomicCAS(dir + h, kDirEmpty, mine) != kDirEmpty) h = h + 1 == cap ? 0 : h + 1;
}

// kProbePerLane independent lookups of one lane, interleaved round by round; same contract as fence_search
// except that c[p] is only meaningful where hit[p].
template <int NW, int kProbePerLane>
__device__ __forceinline__ void dir_lookup(const SnapshotView &v, const SKey<NW> (&k)[kProbePerLane],
                                           uint32_t (&c)[kProbePerLane], bool (&hit)[kProbePerLane]) {
    uint32_t h[kProbePerLane], tag[kProbePerLane];
    bool open[kProbePerLane];
#pragma unroll
    for (int p = 0; p < kProbePerLane; ++p) {
        h[p] = __umulhi(skey_hash<NW>(k[p], 0x9E3779B9u), v.dir_cap);
        tag[p] = skey_tag<NW>(k[p]);
        open[p] = true;
        hit[p] = false;
        c[p] = 0;
    }
    bool any = true;
    while (any) {
        uint2 e[kProbePerLane];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p)
            if (open[p]) e[p] = __ldg(v.dir + h[p]);
        any = false;
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            if (!open[p]) continue;
            if (e[p].y == kNoRow) open[p] = false;                               // empty slot: the key is absent
            else if (e[p].x == tag[p] && (NW == 1 || skey_equal<NW>(load_skey<NW>(v.level[0], e[p].y), k[p]))) {
                hit[p] = true;
                c[p] = e[p].y;
                open[p] = false;
            } else {
                h[p] = h[p] + 1 == v.dir_cap ? 0 : h[p] + 1;
                any = true;
            }
        }
    }
}

// Copies the top fence level into shared memory (all threads of the CTA; ends with a barrier).
__device__ __forceinline__ void load_top_level(const SnapshotView &v, uint32_t *s_top, int nw) {
    const int top = v.n_levels - 1;
    const uint32_t words = v.level_n[top] * (uint32_t) nw;
    for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) s_top[i] = __ldg(v.level[top] + i);
    __syncthreads();
}

// The search proper, for kProbePerLane independent probes of one lane, interleaved step by step (ILP):
// on return c[p] = number of keys of the key column < k[p] (the row an equal key would occupy) and
// hit[p] = an equal key sits there (pad keys excluded).
template <int NW, int kProbePerLane>
__device__ __forceinline__ void fence_search(const SnapshotView &v, const uint32_t *s_top,
                                             const SKey<NW> (&k)[kProbePerLane], uint32_t (&c)[kProbePerLane],
                                             bool (&hit)[kProbePerLane]) {
    const int top = v.n_levels - 1;
    const uint32_t n_top = v.level_n[top];
#pragma unroll
    for (int p = 0; p < kProbePerLane; ++p) {
        c[p] = 0;
        hit[p] = false;
    }
    // top level, in shared memory: branch-free lower bound, c = number of keys < k
    if (n_top > 0) {
        uint32_t n = n_top;
        while (n > 1) {
            const uint32_t half = n >> 1;
#pragma unroll
            for (int p = 0; p < kProbePerLane; ++p)
                if (skey_less<NW>(load_skey_generic<NW>(s_top, c[p] + half - 1), k[p])) c[p] += half;
            n -= half;
        }
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            const SKey<NW> x = load_skey_generic<NW>(s_top, c[p]);
            if (top == 0) hit[p] = skey_equal<NW>(x, k[p]);
            c[p] += skey_less<NW>(x, k[p]);
        }
    }
    for (int l = top - 1; l >= 0; --l) {
        const uint32_t *lvl = v.level[l];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            bool eq;
            const uint32_t cnt = group_count_less<NW>(lvl, c[p], k[p], &eq);
            c[p] = c[p] * kFan + cnt;
            hit[p] = eq;
        }
    }
    const uint32_t n_rows = (uint32_t) v.n_rows;
#pragma unroll
    for (int p = 0; p < kProbePerLane; ++p) hit[p] = hit[p] && c[p] < n_rows;      // an equal PAD key is not a row
}

// kProbePerLane independent probes per lane; kDir: locate rows through the directory instead of the fences
template <int NW, int kProbePerLane, bool kDir = false>
__global__ void __launch_bounds__(kProbeThreads) probe_kernel(SnapshotView v, ProbeArgs a) {
    extern __shared__ __align__(16) uint32_t s_top[];
    if constexpr (!kDir) load_top_level(v, s_top, NW);

    const uint32_t K = v.key_size;
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    const uint64_t per_iter = stride * kProbePerLane;
    for (uint64_t g0 = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; g0 < a.n; g0 += per_iter) {
        SKey<NW> k[kProbePerLane];
        uint32_t c[kProbePerLane];
        bool hit[kProbePerLane];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            const uint64_t gi = g0 + p * stride;
            k[p] = gi < a.n ? to_sortable<NW>(a.keys + gi * K, K, v.xormask) : skey_inf<NW>();
        }
        if constexpr (kDir) {
            dir_lookup<NW, kProbePerLane>(v, k, c, hit);
#pragma unroll
            for (int p = 0; p < kProbePerLane; ++p) hit[p] = hit[p] && g0 + p * stride < a.n;
        } else {
            fence_search<NW, kProbePerLane>(v, s_top, k, c, hit);
        }
        // MVCC + outputs
        uint32_t t[kProbePerLane];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) t[p] = hit[p] ? __ldg(v.ts + c[p]) : 0u;
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            const uint64_t gi = g0 + p * stride;
            if (gi >= a.n) continue;
            uint32_t src = kNoRow, stamp = 0;
            if (hit[p]) src = visible_version(v, c[p], t[p], a.read_ts_each ? a.read_ts_each[gi] : a.read_ts, &stamp);
            if (a.found) a.found[gi] = src != kNoRow ? 1 : 0;
            if (a.ts_out) a.ts_out[gi] = stamp;
            if (a.key_out && src != kNoRow) from_sortable<NW>(k[p], a.key_out + gi * K, K, v.xormask);
            if (a.src_row) a.src_row[gi] = src;
        }
    }
}

}  // namespace stage
