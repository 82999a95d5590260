// probe_rows.cuh — point probe with the payload rows moved by the warp that found them (directory lookups only).
//
// probe_kernel + gather run back to back: a latency-bound lookup kernel (0.045-0.058 ms per 2^20 probes, HBM nearly idle)
// in front of the bandwidth-bound row mover; for 400-byte `stock` rows the lookup is 28 % of the step. Here a warp takes
// 32 probes: every lane looks its key up in the directory (~1.5 dependent loads, + the key check for keys > 4 B) and
// applies the MVCC rule, then the warp copies the 32 rows cooperatively as ONE flat stream of vectors (the destination
// rows of 32 consecutive probes are contiguous; the source row of vector i comes from the lane that found it, by
// shuffle). Other warps of the SM are in the other phase meanwhile, so lookups hide behind row traffic without any
// hand-off structure (the warp-specialised form of round 1, probe_fused.cuh, lost to its own producer / consumer
// imbalance). Same outputs as the two-kernel path; opt-in or default per row size, see launch_probe_rows in stage_gpu.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"

namespace stage {

constexpr int kProbeRowsThreads = 256;

template <int NW, typename V, int U>
__global__ void __launch_bounds__(kProbeRowsThreads) probe_rows_kernel(SnapshotView v, ProbeArgs a, uint8_t *payload_out) {
    const int lane = threadIdx.x & 31;
    const uint32_t K = v.key_size, P = v.payload_size;
    const uint32_t vpr = P / (uint32_t) sizeof(V);                           // vectors per row
    const uint64_t n_warps = (uint64_t) gridDim.x * (kProbeRowsThreads / 32);
    const uint64_t n_chunks = (a.n + 31) / 32;
    for (uint64_t ch = (uint64_t) blockIdx.x * (kProbeRowsThreads / 32) + (threadIdx.x >> 5); ch < n_chunks; ch += n_warps) {
        const uint64_t gi = ch * 32 + lane;
        uint32_t src = kNoRow, stamp = 0;
        SKey<NW> k[1];
        if (gi < a.n) {
            k[0] = to_sortable<NW>(a.keys + gi * K, K, v.xormask);
            uint32_t c[1];
            bool hit[1];
            dir_lookup<NW, 1>(v, k, c, hit);
            if (hit[0]) {
                const uint32_t t = __ldg(v.ts + c[0]);
                src = visible_version(v, c[0], t, a.read_ts_each ? a.read_ts_each[gi] : a.read_ts, &stamp);
            }
            if (a.found) a.found[gi] = src != kNoRow ? 1 : 0;
            if (a.ts_out) a.ts_out[gi] = stamp;
            if (a.key_out && src != kNoRow) from_sortable<NW>(k[0], a.key_out + gi * K, K, v.xormask);
            if (a.src_row) a.src_row[gi] = src;
        }
        // ---- the 32 rows of this chunk as one flat stream of vectors: vector i = row i / vpr, word i % vpr ------------
        const uint32_t rows_here = (uint32_t) min((uint64_t) 32, a.n - ch * 32);
        const uint32_t total = rows_here * vpr;
        V *dst = reinterpret_cast<V *>(payload_out + ch * 32 * (uint64_t) P);
        uint32_t r = (uint32_t) lane / vpr, w = (uint32_t) lane % vpr;      // position of vector `lane`
        const uint32_t dr = 32 / vpr, dw = 32 % vpr;                          // advance by 32 vectors
        for (uint32_t b0 = 0; b0 < total; b0 += 32 * U) {                    // warp-uniform bound: the shuffles need every lane
            const uint32_t i0 = b0 + lane;
            V val[U];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t i = i0 + 32 * u;
                const uint32_t row = __shfl_sync(0xFFFFFFFFu, src, (int) (r & 31));
                ok[u] = i < total && row != kNoRow;
                if (ok[u]) {
                    const uint8_t *base = (row & kVersionRowFlag) ? v.ver_payload : v.payload;
                    val[u] = __ldcs(reinterpret_cast<const V *>(base + (uint64_t) (row & ~kVersionRowFlag) * P) + w);
                }
                r += dr;
                w += dw;
                if (w >= vpr) {
                    w -= vpr;
                    ++r;
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (ok[u]) __stcs(dst + i0 + 32 * u, val[u]);
        }
    }
}

}  // namespace stage
