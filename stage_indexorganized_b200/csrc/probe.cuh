// probe.cuh — batched point probe (replaces BTree::Read, src/vstore/b_tree.cpp:2066-2129, plus the
// point branch of IndexScanExecutor::Execute, include/execute/executor.h:374-454).
//
// The reference descends root->leaf with a binary search per inner node (InternalNode::GetChildIndex,
// b_tree.cpp:664-702) and then scans the leaf's slots linearly (BaseNode::SearchRecordMeta,
// b_tree.cpp:18-122). Keys are unique (LeafNode::CheckUnique, b_tree.cpp:1395-1417) and every leaf's
// keys lie inside its fence range, so "the first visible slot with an equal key in the leaf the fences
// route to" is exactly "the equal key in the globally sorted key column". The flattened column is
// searched with a warp-cooperative 32-ary descent over sampled levels (level l+1 holds every 32nd
// key of level l): each step is one coalesced 32-key load + ballot/popc. The top levels live in
// shared memory; G probes are in flight per warp to overlap the L2/HBM latency of the lower levels.
// Then the MVCC rule (executor.h:383-446) picks the in-leaf version or walks the staged chain. The
// kernel emits found / commit id / key and the source row of every probe; gather_rows_kernel
// (gather.cuh) then streams the payload rows.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"

namespace stage {

constexpr int kMaxLevels = 8;
constexpr int kFan = 32;

struct SnapshotView {
    const uint32_t *level[kMaxLevels];   // level[0] = key column; level[l+1][i] = level[l][32(i+1)-1]
    uint32_t level_n[kMaxLevels];
    int n_levels;
    int smem_levels;                     // how many of the top levels are cached in shared memory
    uint32_t smem_words;                 // total 32-bit words of those levels
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
};

struct ProbeArgs {
    const uint8_t *keys;
    uint64_t n;
    uint32_t read_ts;
    uint8_t *found;
    uint32_t *ts_out;
    uint8_t *key_out;
    uint32_t *src_row;   // [n] row to gather (gather.cuh encoding), kNoRow on a miss
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

template <int NW, int G>
__global__ void __launch_bounds__(256) probe_kernel(SnapshotView v, ProbeArgs a) {
    extern __shared__ __align__(16) uint32_t s_levels[];
    // stage the top levels in shared memory
    const uint32_t *lvl[kMaxLevels];
    {
        uint32_t off = 0;
        for (int l = v.n_levels - 1; l >= 0; --l) {
            if (l >= v.n_levels - v.smem_levels) {
                const uint32_t words = v.level_n[l] * NW;
                for (uint32_t i = threadIdx.x; i < words; i += blockDim.x) s_levels[off + i] = v.level[l][i];
                lvl[l] = s_levels + off;
                off += words;
            } else {
                lvl[l] = v.level[l];
            }
        }
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const uint32_t warps_per_block = blockDim.x >> 5;
    const uint64_t warp_global = (uint64_t) blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    const uint64_t n_warps = (uint64_t) gridDim.x * warps_per_block;
    const uint64_t n_tiles = (a.n + 31) / 32;
    const uint32_t K = v.key_size;
    const int first_smem = v.n_levels - v.smem_levels;

    for (uint64_t tile = warp_global; tile < n_tiles; tile += n_warps) {
        const uint64_t gi = tile * 32 + lane;
        const bool active = gi < a.n;
        SKey<NW> myk = skey_inf<NW>();
        if (active) myk = to_sortable<NW>(a.keys + gi * K, K, v.xormask);

        uint32_t mypos = 0;
        bool myhit = false;
        for (int g = 0; g < 32; g += G) {
            SKey<NW> kq[G];
            uint32_t c[G];
            bool hit[G];
#pragma unroll
            for (int q = 0; q < G; ++q) {
#pragma unroll
                for (int w = 0; w < NW; ++w) kq[q].w[w] = __shfl_sync(0xFFFFFFFFu, myk.w[w], g + q);
                c[q] = 0;
                hit[q] = false;
            }
            for (int l = v.n_levels - 1; l >= 0; --l) {
                const uint32_t nl = v.level_n[l];
                SKey<NW> e[G];
                bool valid[G];
#pragma unroll
                for (int q = 0; q < G; ++q) {
                    const uint64_t idx = (uint64_t) c[q] * kFan + lane;
                    valid[q] = idx < nl;
                    if (valid[q]) e[q] = (l >= first_smem) ? load_skey_generic<NW>(lvl[l], idx) : load_skey<NW>(lvl[l], idx);
                }
#pragma unroll
                for (int q = 0; q < G; ++q) {
                    const bool lt = valid[q] && skey_less<NW>(e[q], kq[q]);
                    const uint32_t m = __ballot_sync(0xFFFFFFFFu, lt);
                    if (l == 0) hit[q] = __ballot_sync(0xFFFFFFFFu, valid[q] && skey_equal<NW>(e[q], kq[q])) != 0;
                    c[q] = c[q] * kFan + __popc(m);
                }
            }
#pragma unroll
            for (int q = 0; q < G; ++q)
                if (lane == g + q) { mypos = c[q]; myhit = hit[q]; }
        }

        // MVCC: one lane per probe
        if (active) {
            uint32_t src = kNoRow, stamp = 0;
            if (myhit) src = visible_version(v, mypos, __ldg(v.ts + mypos), a.read_ts, &stamp);
            if (a.found) a.found[gi] = src != kNoRow ? 1 : 0;
            if (a.ts_out) a.ts_out[gi] = stamp;
            if (a.key_out && src != kNoRow) {
                for (uint32_t b = 0; b < K; ++b) a.key_out[gi * K + b] = a.keys[gi * K + b];
            }
            if (a.src_row) a.src_row[gi] = src;
        }
    }
}

}  // namespace stage
