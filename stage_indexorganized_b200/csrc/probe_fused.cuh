// probe_fused.cuh — point probe, search and payload movement in ONE warp-specialised kernel.
//
// probe_kernel + gather_rows_kernel (probe.cuh, gather.cuh) run back to back: the search is a chain of dependent
// fence loads (latency-bound, ~15 % of the step) during which HBM idles, then the gather saturates HBM. Here every
// CTA runs both at once: the producer warps search tile i+1 — 8 probes interleaved per lane, same
// fence_search / visible_version as probe_kernel — while the 8 consumer warps stream the payload rows of tile i,
// handing the source-row lists over through a ring of shared-memory slots guarded by mbarriers. The search
// latency disappears behind the row traffic of the same SM; outputs are bit-identical to the two-kernel path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"
#include "scan.cuh"   // mbarrier helpers

namespace stage {

constexpr int kFusedPPL = 4;                    // probes per producer lane, searched interleaved
constexpr int kFusedProducerWarps = 2;          // enough probes in flight to cover the LOADED memory latency
constexpr int kFusedTile = 32 * kFusedPPL * kFusedProducerWarps;      // probes per tile
constexpr int kFusedConsumerWarps = 6;
constexpr int kFusedStages = 4;                 // source-row slots between producers and consumers
constexpr int kFusedThreads = 32 * (kFusedProducerWarps + kFusedConsumerWarps);

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int NW, typename V>
__global__ void __launch_bounds__(kFusedThreads) probe_fused_kernel(SnapshotView v, ProbeArgs a, uint8_t *payload_out) {
    extern __shared__ __align__(16) uint32_t s_top[];                 // top fence level (dynamic)
    __shared__ uint32_t s_row[kFusedStages][kFusedTile];
    __shared__ __align__(8) uint64_t s_full[kFusedStages], s_empty[kFusedStages];
    if (!v.dir) load_top_level(v, s_top, NW);
    if (threadIdx.x == 0) {
        for (int i = 0; i < kFusedStages; ++i) {
            mbar_init(&s_full[i], kFusedProducerWarps);
            mbar_init(&s_empty[i], kFusedConsumerWarps);
        }
    }
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t K = v.key_size, P = v.payload_size;
    const uint64_t tiles = (a.n + kFusedTile - 1) / kFusedTile;

    if (warp < kFusedProducerWarps) {
        // ---------------- producer: search + MVCC, scalar outputs, source rows into the slot ----------------
        uint32_t it = 0;
        for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const uint32_t buf = it % kFusedStages, use = it / kFusedStages;
            if (use >= 1) mbar_wait(&s_empty[buf], (use - 1) & 1);           // consumers released this slot
            SKey<NW> k[kFusedPPL];
            uint32_t c[kFusedPPL];
            bool hit[kFusedPPL];
            const uint32_t l0 = warp * 32 * kFusedPPL + lane;             // this lane's first probe of the tile
            const uint64_t g0 = tile * kFusedTile + l0;
#pragma unroll
            for (int p = 0; p < kFusedPPL; ++p) {
                const uint64_t gi = g0 + 32 * p;
                k[p] = gi < a.n ? to_sortable<NW>(a.keys + gi * K, K, v.xormask) : skey_inf<NW>();
            }
            if (v.dir) {
                dir_lookup<NW, kFusedPPL>(v, k, c, hit);
#pragma unroll
                for (int p = 0; p < kFusedPPL; ++p) hit[p] = hit[p] && g0 + 32 * p < a.n;
            } else {
                fence_search<NW, kFusedPPL>(v, s_top, k, c, hit);
            }
            uint32_t t[kFusedPPL];
#pragma unroll
            for (int p = 0; p < kFusedPPL; ++p) t[p] = hit[p] ? __ldg(v.ts + c[p]) : 0u;
#pragma unroll
            for (int p = 0; p < kFusedPPL; ++p) {
                const uint64_t gi = g0 + 32 * p;
                uint32_t src = kNoRow, stamp = 0;
                if (gi < a.n) {
                    if (hit[p]) src = visible_version(v, c[p], t[p], a.read_ts, &stamp);
                    if (a.found) a.found[gi] = src != kNoRow ? 1 : 0;
                    if (a.ts_out) a.ts_out[gi] = stamp;
                    if (a.key_out && src != kNoRow) from_sortable<NW>(k[p], a.key_out + gi * K, K, v.xormask);
                }
                s_row[buf][l0 + 32 * p] = src;
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_full[buf]);
        }
    } else {
        // ---------------- consumers: payload rows of the tile the producer finished --------------------------
        const int cw = warp - kFusedProducerWarps;
        constexpr int U = 4, KK = 4;
        const uint32_t nvec = P / (uint32_t) sizeof(V);
        uint32_t it = 0;
        for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const uint32_t buf = it % kFusedStages;
            mbar_wait(&s_full[buf], (it / kFusedStages) & 1);
            const uint64_t slot0 = tile * kFusedTile;
            for (int r0 = cw * U; r0 < kFusedTile; r0 += kFusedConsumerWarps * U) {
                const V *src[U];
                V *dst[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const uint32_t row = s_row[buf][r0 + u];
                    const uint8_t *base = (row & kVersionRowFlag) ? v.ver_payload : v.payload;
                    src[u] = row == kNoRow ? nullptr : reinterpret_cast<const V *>(base + (uint64_t) (row & ~kVersionRowFlag) * P);
                    dst[u] = reinterpret_cast<V *>(payload_out + (slot0 + r0 + u) * P);
                }
                for (uint32_t c0 = lane; c0 < nvec; c0 += 32 * KK) {
                    V val[U][KK];
#pragma unroll
                    for (int u = 0; u < U; ++u)
#pragma unroll
                        for (int q = 0; q < KK; ++q)
                            if (src[u] && c0 + 32 * q < nvec) val[u][q] = __ldcs(src[u] + c0 + 32 * q);
#pragma unroll
                    for (int u = 0; u < U; ++u)
#pragma unroll
                        for (int q = 0; q < KK; ++q)
                            if (src[u] && c0 + 32 * q < nvec) __stcs(dst[u] + c0 + 32 * q, val[u][q]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[buf]);
        }
    }
}

}  // namespace stage
