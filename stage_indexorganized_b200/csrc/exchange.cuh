// exchange.cuh — routed point probes over peer memory: the probe kernels fused with their exchange.
//
// Range partitioning: shard g owns a contiguous range of the table in reference key order;
// a probe must run on the GPU that owns its key. Instead of all-to-all collectives around the local kernels
// (keys out, results back, then a permutation back into request order), every GPU maps every other GPU's
// exchange arena (NVLink 5 / NVSwitch peer memory) and the kernels address it directly:
//
//   post   (requester)  route each key (owner = #splitters < key), claim a slot of the owner's inbox and store
//                       the key and the request index there                        -- peer stores, 8 B / probe
//   serve  (owner)      search + MVCC over its inboxes (fence_search, probe.cuh), then stream every found
//                       payload row from local HBM straight into the REQUESTER's result slab at the row of
//                       the original request                                       -- peer stores, P B / hit
//
// Results therefore land in request order with no return collective and no restore-order pass. The two
// hand-offs (posts visible before serving, rows visible before the requester reads them) are either a
// caller-side barrier or — STAGE_XCHG_DEVICE_SYNC — epoch flags written into the peers' arenas with
// system-scope release stores and polled locally with acquire loads (bounded spin, sticky error flag).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "gather.cuh"
#include "keys.cuh"
#include "probe.cuh"

namespace stage {

constexpr int kMaxExchangeRanks = 16;
constexpr uint32_t kDstRankShift = 28;                 // dst_slot = (requester rank << 28) | request index
constexpr uint32_t kDstIndexMask = (1u << kDstRankShift) - 1;

// Offsets inside one rank's arena; the layout is identical on every rank.
struct ExchangeLayout {
    uint64_t off_count;     // u32 [world]           requests posted by each source rank (this batch)
    uint64_t off_posted;    // u32 [world]           epoch flag: source rank's posts are complete
    uint64_t off_served;    // u32 [world]           epoch flag: owner rank has written all my rows
    uint64_t off_keys;      // u8  [world][cap][kpad] inbox: keys
    uint64_t off_req;       // u32 [world][cap]       inbox: request index at the source
    uint64_t off_found;     // u8  [cap]              result slab of THIS rank's requests
    uint64_t off_ts;        // u32 [cap]
    uint64_t off_payload;   // u8  [cap*rows][pstride] rows padded to 128 B so that peer stores are line-aligned
    uint64_t off_sizes;     // u32 [world][cap]       inbox: scan sizes (scan exchange only)
    uint64_t off_counts;    // u32 [cap]              result: rows returned per scan (scan exchange only)
    uint64_t off_keyout;    // u8  [cap*rows][kpad]   result: keys of the scanned rows (scan exchange only)
    uint64_t off_agg;       // AggI32 [world]         result: every owner's partial aggregate of MY requests (probe exchange)
    uint64_t bytes;
    uint32_t cap, kpad, pstride;
    uint32_t rows;          // result rows per request: 1 for probes, the scan stride for scans
};

struct ExchangeView {
    uint8_t *arena[kMaxExchangeRanks];   // arena of every rank as mapped in this process (arena[rank] = local)
    ExchangeLayout lay;
    uint32_t world, rank;
    uint32_t *local_cnt;                 // [world] slots claimed per owner while posting (local scratch)
    uint32_t *err;                       // sticky: 1 inbox overflow, 2 device-sync timeout
};

struct ExchangeSplit {
    uint32_t w[(kMaxExchangeRanks - 1) * 8];   // splitters in sortable form, NW words each
};

// ---- system-scope flag helpers ------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// Thread 0 of the CTA waits until flags[0..world) all carry `epoch`; everyone leaves through the barrier.
__device__ __forceinline__ void cta_wait_flags(const uint32_t *flags, uint32_t world, uint32_t epoch, uint32_t *err) {
    if (threadIdx.x == 0) {
        const uint64_t t0 = global_ns();
        for (uint32_t s = 0; s < world; ++s) {
            while (ld_acquire_sys(flags + s) != epoch) {
                if (global_ns() - t0 > 4000000000ull) {          // 4 s: a peer died; fail loudly, do not hang the box
                    atomicExch(err, 2u);
                    break;
                }
                __nanosleep(200);
            }
        }
    }
    __syncthreads();
}

// ---- post ---------------------------------------------------------------------------------------------
// One tile of 256 keys per iteration: owner of every key, slot claim (shared-memory counters, then one
// atomic per owner per tile on the local running counters), peer store of key + request index.
template <int NW>
__global__ void __launch_bounds__(256) exchange_post_kernel(ExchangeView x, ExchangeSplit sp, const uint8_t *keys, uint64_t n,
                                                            uint32_t key_size, uint32_t xormask, const uint32_t *sizes) {
    __shared__ uint32_t s_cnt[kMaxExchangeRanks], s_base[kMaxExchangeRanks];
    const int tid = threadIdx.x;
    const uint64_t tiles = (n + 255) / 256;
    for (uint64_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        if (tid < kMaxExchangeRanks) s_cnt[tid] = 0;
        __syncthreads();
        const uint64_t i = tile * 256 + tid;
        uint32_t owner = 0, pos = 0;
        if (i < n) {
            const SKey<NW> k = to_sortable<NW>(keys + i * key_size, key_size, xormask);
            uint32_t lo = 0, hi = x.world - 1;                   // owner = number of splitters < key
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                SKey<NW> s;
#pragma unroll
                for (int w = 0; w < NW; ++w) s.w[w] = sp.w[mid * NW + w];
                if (skey_less<NW>(s, k)) lo = mid + 1; else hi = mid;
            }
            owner = lo;
            pos = atomicAdd(&s_cnt[owner], 1u);
        }
        __syncthreads();
        if (tid < (int) x.world && s_cnt[tid]) s_base[tid] = atomicAdd(x.local_cnt + tid, s_cnt[tid]);
        __syncthreads();
        if (i < n) {
            const uint32_t p = s_base[owner] + pos;
            if (p < x.lay.cap) {
                uint8_t *inbox = x.arena[owner];
                const uint64_t slot = (uint64_t) x.rank * x.lay.cap + p;
                uint8_t *kd = inbox + x.lay.off_keys + slot * x.lay.kpad;
                const uint8_t *ks = keys + i * key_size;
                if ((key_size & 3u) == 0 && ((uintptr_t) ks & 3u) == 0) {
                    for (uint32_t b = 0; b < key_size; b += 4)
                        *reinterpret_cast<uint32_t *>(kd + b) = *reinterpret_cast<const uint32_t *>(ks + b);
                } else {
                    for (uint32_t b = 0; b < key_size; ++b) kd[b] = ks[b];
                }
                reinterpret_cast<uint32_t *>(inbox + x.lay.off_req)[slot] = (uint32_t) i;
                if (sizes) reinterpret_cast<uint32_t *>(inbox + x.lay.off_sizes)[slot] = min(sizes[i], x.lay.rows);
            } else {
                atomicExch(x.err, 1u);
            }
        }
        __syncthreads();
    }
}

// After the posts (same stream): publish the per-owner counts in the owners' arenas, reset the local counters
// for the next batch and — device sync — raise this rank's "posted" flag everywhere.
__global__ void exchange_post_finish_kernel(ExchangeView x, uint32_t epoch, int device_sync) {
    const uint32_t t = threadIdx.x;
    if (t < x.world) {
        const uint32_t c = min(x.local_cnt[t], x.lay.cap);
        x.local_cnt[t] = 0;
        uint8_t *peer = x.arena[t];
        reinterpret_cast<uint32_t *>(peer + x.lay.off_count)[x.rank] = c;
        if (device_sync) {
            __threadfence_system();
            st_release_sys(reinterpret_cast<uint32_t *>(peer + x.lay.off_posted) + x.rank, epoch);
        }
    }
}

// ---- serve --------------------------------------------------------------------------------------------
struct ServeArgs {
    uint32_t read_ts;
    uint32_t epoch;
    int device_sync;
    uint32_t *src_row;      // [world*cap] local scratch: row to gather per inbox entry (compact over sources)
    uint32_t *dst_slot;     // [world*cap] (requester << 28) | request index
    int tile_order;         // 0: permuted (t * mul mod tiles); 1: all-to-all shift schedule (TilePlan)
    uint32_t seg_mul;       // result rows per inbox entry (1: probes; scan stride: scans)
};

// prefix offsets of the per-source inbox counts, into shared memory (ends with a barrier)
__device__ __forceinline__ void load_segments(const ExchangeView &x, uint32_t *s_off, uint32_t seg_mul = 1) {
    if (threadIdx.x == 0) {
        const uint32_t *cnt = reinterpret_cast<const uint32_t *>(x.arena[x.rank] + x.lay.off_count);
        uint32_t acc = 0;
        for (int s = 0; s < kMaxExchangeRanks; ++s) {
            s_off[s] = acc;
            if (s < (int) x.world) acc += min(__ldcg(cnt + s), x.lay.cap) * seg_mul;
        }
        s_off[kMaxExchangeRanks] = acc;
    }
    __syncthreads();
}

// Search + MVCC over everything posted to this rank. found / commit id go straight to the requester's slab;
// the payload row to move and its destination are handed to exchange_gather_kernel.
template <int NW, int kProbePerLane, bool kDir = false>
__global__ void __launch_bounds__(kProbeThreads) exchange_serve_kernel(SnapshotView v, ExchangeView x, ServeArgs a) {
    extern __shared__ __align__(16) uint32_t s_top[];
    if constexpr (!kDir) load_top_level(v, s_top, NW);
    if (a.device_sync)
        cta_wait_flags(reinterpret_cast<const uint32_t *>(x.arena[x.rank] + x.lay.off_posted), x.world, a.epoch, x.err);
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off);
    const uint32_t total = off[kMaxExchangeRanks];
    const uint8_t *inbox = x.arena[x.rank];
    const uint32_t *req = reinterpret_cast<const uint32_t *>(inbox + x.lay.off_req);

    const uint32_t stride = gridDim.x * blockDim.x;
    for (uint32_t g0 = blockIdx.x * blockDim.x + threadIdx.x; g0 < total; g0 += stride * kProbePerLane) {
        SKey<NW> k[kProbePerLane];
        uint32_t c[kProbePerLane], dst[kProbePerLane];
        bool hit[kProbePerLane];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            const uint32_t g = g0 + p * stride;
            k[p] = skey_inf<NW>();
            dst[p] = 0;
            if (g < total) {
                uint32_t s = 0;
                for (uint32_t q = 1; q < x.world; ++q) s += g >= off[q] ? 1u : 0u;
                const uint64_t slot = (uint64_t) s * x.lay.cap + (g - off[s]);
                k[p] = to_sortable<NW>(inbox + x.lay.off_keys + slot * x.lay.kpad, v.key_size, v.xormask);
                dst[p] = (s << kDstRankShift) | (req[slot] & kDstIndexMask);
            }
        }
        if constexpr (kDir) {
            dir_lookup<NW, kProbePerLane>(v, k, c, hit);
#pragma unroll
            for (int p = 0; p < kProbePerLane; ++p) hit[p] = hit[p] && g0 + p * stride < total;
        } else {
            fence_search<NW, kProbePerLane>(v, s_top, k, c, hit);
        }
        uint32_t t[kProbePerLane];
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) t[p] = hit[p] ? __ldg(v.ts + c[p]) : 0u;
#pragma unroll
        for (int p = 0; p < kProbePerLane; ++p) {
            const uint32_t g = g0 + p * stride;
            if (g >= total) continue;
            uint32_t src = kNoRow, stamp = 0;
            if (hit[p]) src = visible_version(v, c[p], t[p], a.read_ts, &stamp);
            uint8_t *peer = x.arena[dst[p] >> kDstRankShift];
            const uint32_t r = dst[p] & kDstIndexMask;
            (peer + x.lay.off_found)[r] = src != kNoRow ? 1 : 0;
            reinterpret_cast<uint32_t *>(peer + x.lay.off_ts)[r] = stamp;
            a.src_row[g] = src;
            a.dst_slot[g] = dst[p];
        }
    }
}

// Inbox entries are grouped by requester, so walking the tiles in order would first move all rows of requester 0,
// then requester 1, ...: the local (HBM-bound) and the remote (NVLink-bound) rows would be serialised. Tiles are
// therefore visited in the order t -> t * mul mod tiles (mul coprime with tiles: a bijection), which spreads every
// requester's tiles over the whole kernel so both paths stay busy at once.
__device__ __forceinline__ uint32_t tile_multiplier(uint32_t tiles) {
    const uint32_t primes[4] = {7919u, 7927u, 7933u, 7937u};
#pragma unroll
    for (int i = 0; i < 4; ++i)
        if (tiles % primes[i] != 0) return primes[i];
    return 1u;
}

// Order 1 (all-to-all shift schedule): tiles never straddle two requesters; owner r walks the REMOTE requesters in
// the rotated order r+1, r+2, ... (mod world), each requester's tiles in inbox order, with the tiles of its local
// segment spread evenly in between. At any moment every requester's slab is then written by (about) one owner —
// no incast on a single GPU's NVLink ingress — and the CTAs running together write a narrow window of that slab.
struct TilePlan {
    uint32_t tiles, mul, nloc, nrem_segs;
    uint32_t rstart[kMaxExchangeRanks + 1];   // prefix of tile counts over the rotated remote segments
    uint32_t rseg[kMaxExchangeRanks];
};
__device__ __forceinline__ void make_tile_plan(TilePlan *p, const uint32_t *off, uint32_t world, uint32_t rank, int order) {
    if (threadIdx.x == 0) {
        const uint32_t total = off[kMaxExchangeRanks];
        if (order == 0) {
            p->tiles = (total + kGatherRows - 1) / kGatherRows;
            p->mul = tile_multiplier(p->tiles);
        } else {
            p->nloc = (off[rank + 1] - off[rank] + kGatherRows - 1) / kGatherRows;
            uint32_t acc = 0, k = 0;
            for (uint32_t d = 1; d < world; ++d) {
                const uint32_t sgm = (rank + d) % world;
                p->rseg[k] = sgm;
                p->rstart[k] = acc;
                acc += (off[sgm + 1] - off[sgm] + kGatherRows - 1) / kGatherRows;
                ++k;
            }
            p->rstart[k] = acc;
            p->nrem_segs = k;
            p->tiles = p->nloc + acc;
        }
    }
    __syncthreads();
}
// entries [*g0, *gend) of visit number t (at most kGatherRows of them are taken)
__device__ __forceinline__ void tile_span(const TilePlan *p, const uint32_t *off, uint32_t rank, int order, uint32_t t,
                                          uint32_t *g0, uint32_t *gend) {
    if (order == 0) {
        *g0 = (uint32_t) (((uint64_t) t * p->mul) % p->tiles) * kGatherRows;
        *gend = off[kMaxExchangeRanks];
        return;
    }
    const uint32_t L = p->nloc, T = p->tiles;
    const uint32_t before = (uint32_t) (((uint64_t) t * L) / T), after = (uint32_t) (((uint64_t) (t + 1) * L) / T);
    if (after > before) {
        *g0 = off[rank] + before * kGatherRows;
        *gend = off[rank + 1];
        return;
    }
    const uint32_t r = t - before;
    uint32_t k = 0;
    while (k + 1 < p->nrem_segs && r >= p->rstart[k + 1]) ++k;
    const uint32_t sgm = p->rseg[k];
    *g0 = off[sgm] + (r - p->rstart[k]) * kGatherRows;
    *gend = off[sgm + 1];
}

// Payload rows: local HBM -> the requesters' result slabs (peer stores over NVLink for remote requesters).
// Persistent CTAs over tiles of kGatherRows inbox entries; same in-flight structure as gather_rows_kernel.
template <typename V, int U>
__global__ void __launch_bounds__(kGatherThreads) exchange_gather_kernel(ExchangeView x, ServeArgs a, const uint8_t *payload,
                                                                        const uint8_t *ver_payload, uint32_t P) {
    __shared__ uint32_t s_row[kGatherRows], s_dst[kGatherRows];
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, a.seg_mul);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kGatherThreads / 32;
    const uint32_t nvec = P / (uint32_t) sizeof(V);
    __shared__ TilePlan plan;
    make_tile_plan(&plan, off, x.world, x.rank, a.tile_order);
    const uint32_t tiles = plan.tiles;
    for (uint32_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        uint32_t g0, gend;
        tile_span(&plan, off, x.rank, a.tile_order, t, &g0, &gend);
        __syncthreads();
        if (tid < kGatherRows) {
            const uint32_t g = g0 + tid;
            s_row[tid] = g < gend ? a.src_row[g] : kNoRow;
            s_dst[tid] = g < gend ? a.dst_slot[g] : 0u;
        }
        __syncthreads();
        for (int r0 = warp * U; r0 < kGatherRows; r0 += kWarps * U) {
            const V *src[U];
            V *dst[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t row = s_row[r0 + u], d = s_dst[r0 + u];
                const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
                src[u] = row == kNoRow ? nullptr : reinterpret_cast<const V *>(base + (uint64_t) (row & ~kVersionRowFlag) * P);
                dst[u] = reinterpret_cast<V *>(x.arena[d >> kDstRankShift] + x.lay.off_payload +
                                               (uint64_t) (d & kDstIndexMask) * x.lay.pstride);
            }
            constexpr int K = 4;
            for (uint32_t c0 = lane; c0 < nvec; c0 += 32 * K) {
                V val[U][K];
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        if (src[u] && c0 + 32 * k < nvec) val[u][k] = __ldcs(src[u] + c0 + 32 * k);
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int k = 0; k < K; ++k)
                        if (src[u] && c0 + 32 * k < nvec) dst[u][c0 + 32 * k] = val[u][k];
            }
        }
    }
}

// Same, for rows whose size is a multiple of 8 (8-byte aligned sources, e.g. the 1000-byte YCSB rows): each lane
// moves 16-byte chunks — two 8-byte loads, ONE 16-byte store — so a warp store instruction covers 512 contiguous,
// 128-byte-aligned bytes of the destination row (the slab rows are padded to 128 B): full NVLink write packets.
template <int U>
__global__ void __launch_bounds__(kGatherThreads) exchange_gather16_kernel(ExchangeView x, ServeArgs a, const uint8_t *payload,
                                                                          const uint8_t *ver_payload, uint32_t P) {
    __shared__ uint32_t s_row[kGatherRows], s_dst[kGatherRows];
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, a.seg_mul);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kGatherThreads / 32;
    const uint32_t nfull = P / 16, tail8 = (P & 15u) >> 3;       // 16-byte chunks, then one optional 8-byte chunk
    __shared__ TilePlan plan;
    make_tile_plan(&plan, off, x.world, x.rank, a.tile_order);
    const uint32_t tiles = plan.tiles;
    for (uint32_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        uint32_t g0, gend;
        tile_span(&plan, off, x.rank, a.tile_order, t, &g0, &gend);
        __syncthreads();
        if (tid < kGatherRows) {
            const uint32_t g = g0 + tid;
            s_row[tid] = g < gend ? a.src_row[g] : kNoRow;
            s_dst[tid] = g < gend ? a.dst_slot[g] : 0u;
        }
        __syncthreads();
        for (int r0 = warp * U; r0 < kGatherRows; r0 += kWarps * U) {
            const uint2 *src[U];
            uint8_t *dst[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const uint32_t row = s_row[r0 + u], d = s_dst[r0 + u];
                const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
                src[u] = row == kNoRow ? nullptr : reinterpret_cast<const uint2 *>(base + (uint64_t) (row & ~kVersionRowFlag) * P);
                dst[u] = x.arena[d >> kDstRankShift] + x.lay.off_payload + (uint64_t) (d & kDstIndexMask) * x.lay.pstride;
            }
            constexpr int K = 2;
            for (uint32_t c0 = lane; c0 < nfull + tail8; c0 += 32 * K) {
                uint2 lo[U][K], hi[U][K];
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int k = 0; k < K; ++k) {
                        const uint32_t c = c0 + 32 * k;
                        if (src[u] && c < nfull + tail8) lo[u][k] = __ldcs(src[u] + 2 * c);
                        if (src[u] && c < nfull) hi[u][k] = __ldcs(src[u] + 2 * c + 1);
                    }
#pragma unroll
                for (int u = 0; u < U; ++u)
#pragma unroll
                    for (int k = 0; k < K; ++k) {
                        const uint32_t c = c0 + 32 * k;
                        if (src[u] && c < nfull)
                            *reinterpret_cast<uint4 *>(dst[u] + 16 * c) = make_uint4(lo[u][k].x, lo[u][k].y, hi[u][k].x, hi[u][k].y);
                        else if (src[u] && c < nfull + tail8)
                            *reinterpret_cast<uint2 *>(dst[u] + 16 * c) = lo[u][k];
                    }
            }
        }
    }
}

// ---- TMA row mover (opt-in, STAGE_XCHG_TMA=1) ---------------------------------------------------------------------
// Same job as exchange_gather16_kernel, with the store side handed to the TMA unit: a warp loads kTmaRows rows from local
// HBM into shared memory (8-byte loads: 1000-byte rows are only 8-byte aligned), then ONE lane issues one
// cp.async.bulk shared -> global per row — a single bulk write of align16(P) bytes into the requester's slab row (slab
// rows are padded to 128 B, so the few pad bytes land in the padding) instead of 63 lanes x 16-byte stores. Two groups of
// kTmaRows buffers per warp: the next group is loaded while the previous group's bulk stores drain.
constexpr int kTmaRows = 4;
constexpr int kTmaGroups = 2;
constexpr uint32_t kTmaRowBytes = 1024;       // shared-memory slot per row (P <= 1016)

__device__ __forceinline__ void tma_store_1d(void *gmem_dst, const void *smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"((uint32_t) __cvta_generic_to_shared(smem_src)),
                 "r"(bytes)
                 : "memory");
}

__global__ void __launch_bounds__(kGatherThreads) exchange_gather_tma_kernel(ExchangeView x, ServeArgs a, const uint8_t *payload,
                                                                            const uint8_t *ver_payload, uint32_t P) {
    extern __shared__ __align__(128) uint8_t s_buf[];     // [warps][kTmaGroups][kTmaRows][kTmaRowBytes]
    __shared__ uint32_t s_row[kGatherRows], s_dst[kGatherRows];
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, a.seg_mul);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kGatherThreads / 32;
    const uint32_t n8 = P / 8, bulk = (P + 15u) & ~15u;
    uint8_t *wbuf = s_buf + (size_t) warp * kTmaGroups * kTmaRows * kTmaRowBytes;
    __shared__ TilePlan plan;
    make_tile_plan(&plan, off, x.world, x.rank, a.tile_order);
    const uint32_t tiles = plan.tiles;
    uint32_t grp = 0;
    for (uint32_t t = blockIdx.x; t < tiles; t += gridDim.x) {
        uint32_t g0, gend;
        tile_span(&plan, off, x.rank, a.tile_order, t, &g0, &gend);
        __syncthreads();
        if (tid < kGatherRows) {
            const uint32_t g = g0 + tid;
            s_row[tid] = g < gend ? a.src_row[g] : kNoRow;
            s_dst[tid] = g < gend ? a.dst_slot[g] : 0u;
        }
        __syncthreads();
        for (int r0 = warp * kTmaRows; r0 < kGatherRows; r0 += kWarps * kTmaRows) {
            uint8_t *gbuf = wbuf + (size_t) (grp & 1) * kTmaRows * kTmaRowBytes;
            // the bulk stores issued two groups ago have finished READING this group's buffers
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kTmaGroups - 1) : "memory");
            __syncwarp();
            uint2 v[kTmaRows][4];
#pragma unroll
            for (int u = 0; u < kTmaRows; ++u) {
                const uint32_t row = s_row[r0 + u];
                if (row == kNoRow) continue;
                const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
                const uint2 *src = reinterpret_cast<const uint2 *>(base + (uint64_t) (row & ~kVersionRowFlag) * P);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (lane + 32 * k < n8) v[u][k] = __ldcs(src + lane + 32 * k);
            }
#pragma unroll
            for (int u = 0; u < kTmaRows; ++u) {
                if (s_row[r0 + u] == kNoRow) continue;
                uint2 *dst = reinterpret_cast<uint2 *>(gbuf + (size_t) u * kTmaRowBytes);
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (lane + 32 * k < n8) dst[lane + 32 * k] = v[u][k];
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> visible to the TMA unit
            __syncwarp();
            if (lane == 0) {
#pragma unroll
                for (int u = 0; u < kTmaRows; ++u) {
                    if (s_row[r0 + u] == kNoRow) continue;
                    const uint32_t d = s_dst[r0 + u];
                    uint8_t *dst = x.arena[d >> kDstRankShift] + x.lay.off_payload + (uint64_t) (d & kDstIndexMask) * x.lay.pstride;
                    tma_store_1d(dst, gbuf + (size_t) u * kTmaRowBytes, bulk);
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
            ++grp;
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // all rows written before the kernel ends
}

// ---- routed push-down ("return 4-16 B per probe instead of 416-1004 B, lifts the NVLink bound") ------
// Column projection: after exchange_serve_kernel has located and MVCC-filtered every inbox entry, only payload bytes
// [col_off, col_off + col_len) of each found row cross NVLink, into the requester's slab at stride col_stride (rows of
// missed probes get zeros). One thread per inbox entry.
__global__ void __launch_bounds__(256) exchange_cols_kernel(ExchangeView x, ServeArgs a, const uint8_t *payload, const uint8_t *ver_payload,
                                                            uint32_t P, uint32_t col_off, uint32_t col_len, uint32_t col_stride) {
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, 1);
    const uint32_t total = off[kMaxExchangeRanks];
    for (uint32_t g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
        const uint32_t row = a.src_row[g], d = a.dst_slot[g];
        uint8_t *dst = x.arena[d >> kDstRankShift] + x.lay.off_payload + (uint64_t) (d & kDstIndexMask) * col_stride;
        if (row == kNoRow) {
            for (uint32_t b = 0; b < col_len; ++b) dst[b] = 0;
            continue;
        }
        const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
        const uint8_t *src = base + (uint64_t) (row & ~kVersionRowFlag) * P + col_off;
        if ((((uintptr_t) src | col_len | col_stride) & 3u) == 0)
            for (uint32_t b = 0; b < col_len; b += 4) *reinterpret_cast<uint32_t *>(dst + b) = __ldg(reinterpret_cast<const uint32_t *>(src + b));
        else
            for (uint32_t b = 0; b < col_len; ++b) dst[b] = __ldg(src + b);
    }
}

// Aggregate push-down: every owner reduces an int32 column over the entries of each requester's segment (count found,
// count below the threshold, MIN, MAX — what RunStockLevel and CH-Q2* compute from their stock probes) into local
// partials; exchange_agg_publish_kernel then stores ONE 24-byte record per requester over NVLink. blockIdx.y = segment.
__global__ void __launch_bounds__(256) exchange_agg_kernel(ExchangeView x, ServeArgs a, const uint8_t *payload, const uint8_t *ver_payload,
                                                           uint32_t P, uint32_t col_off, int threshold, AggI32 *partial /* [world] */) {
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, 1);
    const uint32_t sgm = blockIdx.y;
    if (sgm >= x.world) return;
    unsigned found = 0, below = 0;
    int mn = 0x7FFFFFFF, mx = (int) 0x80000000;
    for (uint32_t g = off[sgm] + blockIdx.x * blockDim.x + threadIdx.x; g < off[sgm + 1]; g += gridDim.x * blockDim.x) {
        const uint32_t row = a.src_row[g];
        if (row == kNoRow) continue;
        const uint8_t *base = (row & kVersionRowFlag) ? ver_payload : payload;
        const uint8_t *p = base + (uint64_t) (row & ~kVersionRowFlag) * P + col_off;
        const int v = (int) ((uint32_t) __ldg(p) | ((uint32_t) __ldg(p + 1) << 8) | ((uint32_t) __ldg(p + 2) << 16) | ((uint32_t) __ldg(p + 3) << 24));
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
        atomicAdd(&partial[sgm].found, (unsigned long long) found);
        atomicAdd(&partial[sgm].below, (unsigned long long) below);
        atomicMin(&partial[sgm].min_value, mn);
        atomicMax(&partial[sgm].max_value, mx);
    }
}
// partial[t] -> requester t's arena, slot [my rank]; the local partials are reset for the next batch
__global__ void exchange_agg_publish_kernel(ExchangeView x, AggI32 *partial) {
    const uint32_t t = threadIdx.x;
    if (t < x.world) {
        AggI32 *dst = reinterpret_cast<AggI32 *>(x.arena[t] + x.lay.off_agg) + x.rank;
        *dst = partial[t];
        partial[t] = AggI32{0ull, 0ull, 0x7FFFFFFF, (int) 0x80000000};
    }
}
// requester side, after the wait: combine the owners' partials of my batch
__global__ void exchange_agg_combine_kernel(ExchangeView x, AggI32 *out) {
    if (threadIdx.x == 0) {
        const AggI32 *p = reinterpret_cast<const AggI32 *>(x.arena[x.rank] + x.lay.off_agg);
        AggI32 r{0ull, 0ull, 0x7FFFFFFF, (int) 0x80000000};
        for (uint32_t o = 0; o < x.world; ++o) {
            r.found += p[o].found;
            r.below += p[o].below;
            r.min_value = min(r.min_value, p[o].min_value);
            r.max_value = max(r.max_value, p[o].max_value);
        }
        *out = r;
    }
}

// Device sync, owner side: all rows of this batch are written -> raise "served" at every requester.
__global__ void exchange_serve_finish_kernel(ExchangeView x, uint32_t epoch) {
    const uint32_t t = threadIdx.x;
    if (t < x.world) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<uint32_t *>(x.arena[t] + x.lay.off_served) + x.rank, epoch);
    }
}
// Device sync, requester side: every owner has served this batch -> the result slab is complete.
__global__ void exchange_wait_served_kernel(ExchangeView x, uint32_t epoch) {
    cta_wait_flags(reinterpret_cast<const uint32_t *>(x.arena[x.rank] + x.lay.off_served), x.world, epoch, x.err);
}

// ---- scans over the exchange ------------------------------------------------------------------------------
// Scan requests (start key, size) are posted like probe keys. The owner plans them with the ordinary scan plan
// kernels, one launch per source rank over that rank's inbox (the number of requests is read on the device), into
// local scratch in inbox order; exchange_scan_deliver_kernel then hands every scan to its requester: row count,
// keys and commit ids are stored into the requester's slab at the request's index, and the compact
// (source row, destination row) lists for the payload mover — the same exchange_gather kernels as for probes — are
// written. Result rows of request q live at rows [q*rows, q*rows + count) of the requester's slab.
__global__ void exchange_wait_posted_kernel(ExchangeView x, uint32_t epoch) {
    cta_wait_flags(reinterpret_cast<const uint32_t *>(x.arena[x.rank] + x.lay.off_posted), x.world, epoch, x.err);
}

struct ScanDeliverArgs {
    const uint32_t *row_counts;   // [world*cap]           plan outputs, inbox order
    const uint8_t *keys;          // [world*cap*rows][K]
    const uint32_t *ts;           // [world*cap*rows]
    const uint32_t *src_row;      // [world*cap*rows]
    uint32_t *src_row_c;          // compact over (source, entry, pos): input of the payload mover
    uint32_t *dst_slot_c;
    uint32_t key_size;
};

__global__ void __launch_bounds__(256) exchange_scan_deliver_kernel(ExchangeView x, ScanDeliverArgs d) {
    __shared__ uint32_t off[kMaxExchangeRanks + 1];
    load_segments(x, off, 1);                                  // entries, not rows
    const uint32_t total = off[kMaxExchangeRanks], R = x.lay.rows, K = d.key_size;
    const uint8_t *inbox = x.arena[x.rank];
    const uint32_t *req = reinterpret_cast<const uint32_t *>(inbox + x.lay.off_req);
    for (uint32_t e = blockIdx.x; e < total; e += gridDim.x) {
        uint32_t sgm = 0;
        for (uint32_t q = 1; q < x.world; ++q) sgm += e >= off[q] ? 1u : 0u;
        const uint64_t slot = (uint64_t) sgm * x.lay.cap + (e - off[sgm]);
        const uint32_t q = req[slot], cnt = d.row_counts[slot];
        uint8_t *peer = x.arena[sgm];
        if (threadIdx.x == 0) reinterpret_cast<uint32_t *>(peer + x.lay.off_counts)[q] = cnt;
        for (uint32_t pos = threadIdx.x; pos < R; pos += blockDim.x) {
            const uint64_t in = slot * R + pos, g = (uint64_t) e * R + pos, out = (uint64_t) q * R + pos;
            const uint32_t src = pos < cnt ? d.src_row[in] : kNoRow;
            d.src_row_c[g] = src;
            d.dst_slot_c[g] = (sgm << kDstRankShift) | (uint32_t) out;
            if (pos < cnt) {
                reinterpret_cast<uint32_t *>(peer + x.lay.off_ts)[out] = d.ts[in];
                for (uint32_t b = 0; b < K; b += 4)
                    *reinterpret_cast<uint32_t *>(peer + x.lay.off_keyout + out * x.lay.kpad + b) =
                        *reinterpret_cast<const uint32_t *>(d.keys + in * K + b);
            }
        }
    }
}

}  // namespace stage
