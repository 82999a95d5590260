// refresh.cuh — incremental snapshot refresh: committed single-row updates applied to the flattened table.
//
// Installs, on the device, exactly the state SSNTransactionManager::CommitTransaction leaves behind for an
// updated record (src/concurrency/transaction_manager.cpp:610-676): the old version [payload] is retired into a
// version slot with begin = the old commit id, commit = the committer's id, next = the previous chain head; the
// row is overwritten in place, its commit id becomes the committer's id and its chain head the new slot. The
// update itself has the shape of the drivers' column updates (PointUpdateExecutor, ycsb_mixed.cpp:38-80): a patch
// of `patch_len` bytes at `patch_off` of the payload (the whole row when patch_off = 0, patch_len = P).
//
// A batch is applied in order. Updates of the same row must be chained one after the other, so the batch is
// stable-sorted by row (cub radix sort, on the host side of the call) and ONE warp walks each run of equal rows
// sequentially; different rows proceed in parallel. Update i always takes version slot v0 + i, so the version
// numbering does not depend on the schedule.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"

namespace stage {

struct RefreshArgs {
    const uint8_t *keys;          // [n][K]
    const uint8_t *patches;       // [n][patch_len]
    const uint32_t *commit_ids;   // [n]
    uint64_t n;
    uint32_t patch_off, patch_len;
    uint32_t v0;                  // versions already staged: update i -> version index v0 + i (0-based)
    // snapshot columns (writable)
    uint32_t *ts, *chain;
    uint8_t *payload;
    uint32_t *ver_begin, *ver_commit, *ver_next;
    uint8_t *ver_payload;
    uint32_t payload_size;
    // scratch
    uint32_t *row;                // [n] row of update i (kNoRow: key absent, update ignored)
    uint32_t *order;              // [n] update indices sorted by (row, i)
    uint32_t *row_sorted;         // [n]
    uint32_t *applied;            // number of updates applied
};

// row of every update key (same search as a probe; no visibility: the writer addresses the latest version)
template <int NW>
__global__ void __launch_bounds__(kProbeThreads) refresh_locate_kernel(SnapshotView v, RefreshArgs a) {
    extern __shared__ __align__(16) uint32_t s_top[];
    load_top_level(v, s_top, NW);
    const uint64_t stride = (uint64_t) gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride) {
        SKey<NW> k[1];
        uint32_t c[1];
        bool hit[1];
        k[0] = to_sortable<NW>(a.keys + i * v.key_size, v.key_size, v.xormask);
        fence_search<NW, 1>(v, s_top, k, c, hit);
        a.row[i] = hit[0] ? c[0] : kNoRow;
        a.order[i] = (uint32_t) i;
    }
}

// one warp per sorted position; the warp at the head of a run applies the whole run in batch order
__global__ void __launch_bounds__(256) refresh_apply_kernel(RefreshArgs a) {
    const uint64_t j = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (j >= a.n) return;
    const uint32_t row = a.row_sorted[j];
    if (row == kNoRow || (j > 0 && a.row_sorted[j - 1] == row)) return;
    const uint32_t P = a.payload_size;
    uint8_t *cur = a.payload + (uint64_t) row * P;
    uint32_t t = a.ts[row], head = a.chain[row], done = 0;
    for (uint64_t q = j; q < a.n && a.row_sorted[q] == row; ++q) {
        const uint32_t i = a.order[q];
        const uint32_t v = a.v0 + i;
        uint8_t *old = a.ver_payload + (uint64_t) v * P;
        if (((uintptr_t) cur | (uintptr_t) old | P) % 8 == 0) {
            for (uint32_t b = lane * 8; b < P; b += 256) *reinterpret_cast<uint2 *>(old + b) = *reinterpret_cast<const uint2 *>(cur + b);
        } else {
            for (uint32_t b = lane; b < P; b += 32) old[b] = cur[b];
        }
        __syncwarp();
        const uint8_t *patch = a.patches + (uint64_t) i * a.patch_len;
        for (uint32_t b = lane; b < a.patch_len; b += 32) cur[a.patch_off + b] = patch[b];
        __syncwarp();
        if (lane == 0) {
            a.ver_begin[v] = t;
            a.ver_commit[v] = a.commit_ids[i];
            a.ver_next[v] = head;
        }
        t = a.commit_ids[i];
        head = v + 1;                                   // chain heads are 1-based
        ++done;
    }
    if (lane == 0) {
        a.ts[row] = t;
        a.chain[row] = head;
        atomicAdd(a.applied, done);
    }
}

// version slots of updates whose key is absent stay unreachable; give them inert headers
__global__ void refresh_clear_unused_kernel(RefreshArgs a) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < a.n && a.row[i] == kNoRow) {
        a.ver_begin[a.v0 + i] = 0;
        a.ver_commit[a.v0 + i] = 0;
        a.ver_next[a.v0 + i] = 0;
    }
}

// ---- version-table compaction -----------------------------------------------------------------------------------
// Every refresh that re-transforms a leaf stages the chains of its rows again (append-only), so the copies the previous
// snapshot held for that leaf stay behind unreferenced. Mark what the chain column still reaches, renumber, move.
__global__ void versions_mark_kernel(const uint32_t *__restrict__ chain, uint64_t n_rows, const uint32_t *__restrict__ ver_next,
                                     uint64_t n_versions, uint32_t *__restrict__ live) {
    for (uint64_t r = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (uint64_t) gridDim.x * blockDim.x) {
        uint32_t v = chain[r];                               // 1-based, 0 = no chain
        uint32_t hops = 0;
        while (v != 0 && v <= n_versions && hops++ < (1u << 20)) {
            if (atomicExch(live + (v - 1), 1u)) break;       // somebody else walks the rest of this chain
            v = ver_next[v - 1];
        }
    }
}

// new_index = exclusive prefix sum of live[]; copies the surviving versions to their new places (one warp per version)
__global__ void versions_move_kernel(const uint32_t *__restrict__ live, const uint32_t *__restrict__ new_index, uint64_t n_versions,
                                     const uint32_t *__restrict__ ob, const uint32_t *__restrict__ oc, const uint32_t *__restrict__ on,
                                     const uint8_t *__restrict__ op, uint32_t *__restrict__ nb, uint32_t *__restrict__ nc,
                                     uint32_t *__restrict__ nn, uint8_t *__restrict__ np, uint32_t payload_size) {
    const int lane = threadIdx.x & 31;
    const uint64_t warps = ((uint64_t) gridDim.x * blockDim.x) >> 5;
    for (uint64_t v = ((uint64_t) blockIdx.x * blockDim.x + threadIdx.x) >> 5; v < n_versions; v += warps) {
        if (!live[v]) continue;
        const uint32_t d = new_index[v];
        if (lane == 0) {
            nb[d] = ob[v];
            nc[d] = oc[v];
            const uint32_t nx = on[v];
            nn[d] = (nx != 0 && nx <= n_versions && live[nx - 1]) ? new_index[nx - 1] + 1 : 0;
        }
        const uint8_t *src = op + v * (uint64_t) payload_size;
        uint8_t *dst = np + d * (uint64_t) payload_size;
        if ((payload_size & 3u) == 0) {
            for (uint32_t b = lane * 4; b < payload_size; b += 128)
                *reinterpret_cast<uint32_t *>(dst + b) = *reinterpret_cast<const uint32_t *>(src + b);
        } else {
            for (uint32_t b = lane; b < payload_size; b += 32) dst[b] = src[b];
        }
    }
}

__global__ void versions_remap_chain_kernel(uint32_t *__restrict__ chain, uint64_t n_rows, const uint32_t *__restrict__ live,
                                            const uint32_t *__restrict__ new_index, uint64_t n_versions) {
    for (uint64_t r = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; r < n_rows; r += (uint64_t) gridDim.x * blockDim.x) {
        const uint32_t v = chain[r];
        if (v != 0) chain[r] = (v <= n_versions && live[v - 1]) ? new_index[v - 1] + 1 : 0;
    }
}

}  // namespace stage
