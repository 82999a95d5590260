// stage_gpu.cu — host side of libstage_gpu.so: snapshot construction (tree walker, staging,
// transform launch), batch entry points, C ABI (include/stage_gpu.h).
//
// Reference anchors (paths relative to the reference repository):
//   tree walk order        include/execute/executor.h:592-609 (ScanInnerNode), b_tree.h:336-345
//   leaf / meta decoding   include/vstore/record_meta.h:66-93, version_store.h:176-217
//   version chains         include/vstore/version_store.h:28-46, src/concurrency/transaction_manager.cpp:610-676
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <stdlib.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <functional>
#include <mutex>
#include <thread>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/stage_gpu.h"
#include "keys.cuh"
#include "node_layout.h"
#include "probe.cuh"
#include "scan.cuh"
#include "scan_short.cuh"
#include "transform.cuh"
#include "gather.cuh"
#include "generate.cuh"
#include "exchange.cuh"
#include "probe_fused.cuh"
#include "probe_rows.cuh"
#include "refresh.cuh"
#include "tpcc_ops.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

using namespace stage;

// ------------------------------------------------------------------------------------------------
// error plumbing
// ------------------------------------------------------------------------------------------------
static thread_local std::string g_last_error;
static std::atomic<uint64_t> g_launches{0};
static int g_device = 0;

#define CUDA_TRY(expr)                                                                       \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            g_last_error = std::string(#expr) + ": " + cudaGetErrorString(_e);               \
            return STAGE_ERR_CUDA;                                                           \
        }                                                                                    \
    } while (0)

#define LAUNCH_CHECK()                                                                       \
    do {                                                                                     \
        g_launches.fetch_add(1, std::memory_order_relaxed);                                  \
        CUDA_TRY(cudaGetLastError());                                                        \
    } while (0)

static int g_sm_count = 0;
static int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

// ------------------------------------------------------------------------------------------------
// snapshot
// ------------------------------------------------------------------------------------------------
struct DevBuf {
    void *p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return STAGE_OK;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        CUDA_TRY(cudaMalloc(&p, want));
        cap = want;
        return STAGE_OK;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

// Where a read call runs: the stream it launches on and the scratch its plan kernel hands to its gather kernel.
struct Lane {
    cudaStream_t st = nullptr;
    DevBuf *srcrow = nullptr;
    bool timing = false;            // record the snapshot's timing events around the launches (device-I/O calls only)
};

// One in-flight STAGE_IO_HOST call: its own streams and scratch, so that concurrent callers (the reference's N pinned
// worker threads, ycsb_workload.cpp:522-529) overlap instead of queueing behind one snapshot-wide lock.
struct HostCtx {
    cudaStream_t st = nullptr, cp = nullptr;        // kernels + H2D; D2H copies of the staged mode
    cudaEvent_t done[2] = {nullptr, nullptr};       // chunk planned + gathered (on st)
    cudaEvent_t drained[2] = {nullptr, nullptr};    // chunk copied out (on cp)
    DevBuf keys, sizes, srcrow;
    DevBuf counts[2], found[2], ts[2], keyout[2], payload[2];
    DevBuf kstr, tstr, off, cubtmp;                 // compact mode: strided keys / commit ids of the plan, row offsets, scan temp
    DevBuf tse, vise;                               // per-operation read ids / visibility flags (batching front-end)
    uint32_t *h_counts = nullptr;                   // pinned: row counts of the chunk in flight (per-scan copies)
    size_t h_counts_cap = 0;
    uint8_t *h_small = nullptr, *d_small = nullptr; // pinned slab (and its device alias) of the small-call path
    bool busy = false;
    void release_all() {
        keys.release(); sizes.release(); srcrow.release();
        kstr.release(); tstr.release(); off.release(); cubtmp.release(); tse.release(); vise.release();
        for (int i = 0; i < 2; ++i) {
            counts[i].release(); found[i].release(); ts[i].release(); keyout[i].release(); payload[i].release();
            if (done[i]) cudaEventDestroy(done[i]);
            if (drained[i]) cudaEventDestroy(drained[i]);
        }
        if (h_counts) cudaFreeHost(h_counts);
        if (h_small) cudaFreeHost(h_small);
        if (st) cudaStreamDestroy(st);
        if (cp) cudaStreamDestroy(cp);
    }
};

struct stage_snapshot {
    stage_schema_t schema{};
    int nw = 1;
    uint32_t xormask = 0;
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    bool sealed = false;
    std::mutex mu;

    uint64_t n_rows = 0, n_leaves = 0, n_versions = 0;
    uint64_t cap_rows = 0, cap_leaves = 0;
    uint32_t max_leaf_rows = 0;
    size_t device_bytes = 0;

    // columns
    uint32_t *d_skey = nullptr;
    uint16_t *d_slot = nullptr;
    uint32_t *d_ts = nullptr, *d_chain = nullptr;
    uint8_t *d_payload = nullptr;
    uint32_t *d_leaf_off = nullptr, *d_sep = nullptr;
    uint8_t *d_cont = nullptr, *d_has_sep = nullptr;
    uint32_t *d_levels = nullptr;               // inside the d_skey allocation
    size_t key_blob_bytes = 0;
    uint32_t *d_ver_begin = nullptr, *d_ver_commit = nullptr, *d_ver_next = nullptr;
    uint8_t *d_ver_payload = nullptr;
    uint32_t *d_err = nullptr;
    uint2 *d_dir = nullptr;                     // point-probe directory (probe.cuh), built at seal time
    uint32_t dir_cap = 0;

    // streaming build state
    std::vector<uint32_t> h_leaf_off;           // [n_leaves+1]
    std::vector<uint32_t> h_sep;                // sortable, n_leaves * nw
    std::vector<uint8_t> h_has_sep;
    std::vector<uint64_t> h_leaf_addr, h_leaf_fp;   // from_tree / refresh_from_tree: host address and fingerprint per leaf
    std::vector<stage_version_t> h_versions;
    std::vector<uint8_t> h_ver_payload;
    uint32_t *d_phys = nullptr;                 // physical (leaf, slot) order -> row, built on first use
    uint8_t *ext_ver_payload = nullptr;         // malloc'd version rows handed over by a caller (instead of h_ver_payload)
    bool versions_on_device = false;            // the version table was uploaded before the seal (from_tree)
    uint8_t *h_stage[2] = {nullptr, nullptr};   // pinned
    uint8_t *d_stage[2] = {nullptr, nullptr};
    cudaEvent_t stage_ev[2] = {nullptr, nullptr};
    uint32_t chunk_leaves = 0;
    uint32_t stage_fill = 0, stage_cur = 0, stage_maxcount = 0;
    uint64_t staged_base_leaf = 0;              // first leaf index of the chunk being filled
    uint64_t rows_staged = 0;

    SnapshotView view{};

    // per-call kernel timing (CUDA events on the snapshot stream): [0] before plan, [1] between, [2] after gather
    cudaEvent_t tev[3] = {nullptr, nullptr, nullptr};
    bool timing = false, timed_gather = false;

    uint64_t ver_cap = 0;                       // capacity (versions) of the d_ver_* arrays
    DevBuf rf_keys, rf_patch, rf_cid, rf_row, rf_order, rf_row_sorted, rf_order_sorted, rf_tmp, rf_applied;   // refresh scratch
    uint8_t *d_raw_keys = nullptr;              // bulk load: caller-filled raw keys, released by seal_sorted
    uint32_t bulk_rows_per_leaf = 0;

    // scratch for STAGE_IO_HOST batches of the push-down calls (serialised on `mu`)
    DevBuf sc_keys, sc_found, sc_ts, sc_keyout, sc_payload, sc_sizes, sc_counts;
    DevBuf sc_srcrow;                           // source-row lists handed from the plan kernels to the gather
    DevBuf sl_req, sl_keys, sl_sizes, sl_counts, sl_src, sl_skeys, sl_psrc, sl_out;   // stage_gpu_tpcc_stock_level scratch
    // contexts of the host-buffer probe / scan calls (one per concurrent caller, created on demand)
    std::mutex ctx_mu;
    std::condition_variable ctx_cv;
    std::vector<HostCtx *> ctx_pool;
    uint32_t *d_row_leaf = nullptr;             // row -> leaf table of the fused short-scan kernel, built on first use
    std::mutex aux_mu;
    unsigned long long *d_stats = nullptr;      // [0] scans in which the slot-order truncation (Q5) selected the rows
};

template <typename T>
static int dev_alloc(stage_snapshot *s, T **p, size_t n) {
    size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    bytes = (bytes + 255) & ~size_t(255);
    CUDA_TRY(cudaMalloc((void **) p, bytes + 256));   // +256: vector loads may touch the padding
    s->device_bytes += bytes + 256;
    return STAGE_OK;
}

static int key_words(uint32_t key_size) {
    uint32_t w = (key_size + 3) / 4;
    if (w <= 1) return 1;
    if (w <= 2) return 2;
    if (w <= 4) return 4;
    if (w <= 8) return 8;
    return 0;
}

static int schema_check(const stage_schema_t *sc, int *nw, uint32_t *xormask) {
    if (!sc || sc->key_size == 0 || sc->payload_size == 0 || sc->leaf_node_size < 64 ||
        sc->leaf_node_size > 65536 || (sc->leaf_node_size & 7))
        return STAGE_ERR_ARG;
    *nw = key_words(sc->key_size);
    if (*nw == 0) return STAGE_ERR_ARG;
    const bool varlen = (sc->key_kind & STAGE_KEY_VARLEN) != 0;
    uint32_t kind = sc->key_kind & ~(uint32_t) STAGE_KEY_VARLEN;
    // variable-length keys: key_size is the slot size, the longest key has key_size - 1 bytes
    const uint32_t longest = varlen ? sc->key_size - 1 : sc->key_size;
    if (varlen && longest == 0) return STAGE_ERR_ARG;
    if (kind == STAGE_KEY_AUTO) kind = longest < 16 ? STAGE_KEY_SIGNED_BYTES : STAGE_KEY_MEMCMP;
    if (kind != STAGE_KEY_SIGNED_BYTES && kind != STAGE_KEY_MEMCMP) return STAGE_ERR_ARG;
    // KeyCompare picks the byte order per PAIR (min length < 16: signed bytes, b_tree.h:118-124): a table whose lengths
    // straddle 16 has no single order. Signed tables hold keys of <= 15 bytes; memcmp tables keys of >= 16 (checked per
    // record by the transform).
    if (varlen && kind == STAGE_KEY_SIGNED_BYTES && longest > 15) return STAGE_ERR_ARG;
    *xormask = kind == STAGE_KEY_SIGNED_BYTES ? 0x80808080u : 0u;
    return STAGE_OK;
}
static bool schema_varlen(const stage_schema_t *sc) { return (sc->key_kind & STAGE_KEY_VARLEN) != 0; }

// stage_gpu_key_pack: key bytes, padding that sorts lowest, length byte — the inverse image of varlen_sortable (keys.cuh)
static void pack_key_slot(const uint8_t *key, uint32_t len, uint32_t slot_size, uint32_t xormask, uint8_t *slot) {
    const uint8_t mask = (uint8_t) (xormask & 0xFFu);
    for (uint32_t i = 0; i + 1 < slot_size; ++i) slot[i] = i < len ? key[i] : mask;
    slot[slot_size - 1] = (uint8_t) len ^ mask;
}

extern "C" int stage_gpu_key_pack(const stage_schema_t *schema, const void *key, uint32_t len, void *slot) {
    int nw;
    uint32_t xm;
    if (!slot || (len && !key) || schema_check(schema, &nw, &xm) || !schema_varlen(schema) || len + 1 > schema->key_size ||
        (xm == 0 && len < 16))
        return STAGE_ERR_ARG;
    pack_key_slot(static_cast<const uint8_t *>(key), len, schema->key_size, xm, static_cast<uint8_t *>(slot));
    return STAGE_OK;
}

extern "C" int stage_gpu_key_unpack(const stage_schema_t *schema, const void *slot, void *key_out, uint32_t *len_out) {
    int nw;
    uint32_t xm;
    if (!slot || !len_out || schema_check(schema, &nw, &xm) || !schema_varlen(schema)) return STAGE_ERR_ARG;
    const uint8_t *p = static_cast<const uint8_t *>(slot);
    const uint32_t len = p[schema->key_size - 1] ^ (uint8_t) (xm & 0xFFu);
    if (len + 1 > schema->key_size) return STAGE_ERR_ARG;
    *len_out = len;
    if (key_out && len) memcpy(key_out, p, len);
    return STAGE_OK;
}

template <int NW>
static void host_sortable(const uint8_t *key, uint32_t K, uint32_t xormask, uint32_t *out) {
    SKey<NW> k = to_sortable<NW>(key, K, xormask);
    for (int i = 0; i < NW; ++i) out[i] = k.w[i];
}
static void host_sortable_nw(int nw, const uint8_t *key, uint32_t K, uint32_t xormask, uint32_t *out) {
    switch (nw) {
        case 1: host_sortable<1>(key, K, xormask, out); break;
        case 2: host_sortable<2>(key, K, xormask, out); break;
        case 4: host_sortable<4>(key, K, xormask, out); break;
        default: host_sortable<8>(key, K, xormask, out); break;
    }
}

static void snapshot_free(stage_snapshot *s) {
    cudaSetDevice(s->device);
    void *ptrs[] = {s->d_skey, s->d_slot, s->d_ts, s->d_chain, s->d_payload, s->d_leaf_off, s->d_sep, s->d_cont,
                    s->d_has_sep, s->d_ver_begin, s->d_ver_commit, s->d_ver_next, s->d_ver_payload,
                    s->d_err, s->d_stage[0], s->d_stage[1], s->d_raw_keys, s->d_phys, s->d_dir, s->d_stats, s->d_row_leaf};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    free(s->ext_ver_payload);
    for (int i = 0; i < 2; ++i) {
        if (s->h_stage[i]) cudaFreeHost(s->h_stage[i]);
        if (s->stage_ev[i]) cudaEventDestroy(s->stage_ev[i]);
    }
    for (HostCtx *c : s->ctx_pool) {
        c->release_all();
        delete c;
    }
    s->ctx_pool.clear();
    s->sc_keys.release(); s->sc_found.release(); s->sc_ts.release(); s->sc_keyout.release();
    s->sc_payload.release(); s->sc_sizes.release(); s->sc_counts.release(); s->sc_srcrow.release();
    s->sl_req.release(); s->sl_keys.release(); s->sl_sizes.release(); s->sl_counts.release(); s->sl_src.release();
    s->sl_skeys.release(); s->sl_psrc.release(); s->sl_out.release();
    s->rf_keys.release(); s->rf_patch.release(); s->rf_cid.release(); s->rf_row.release(); s->rf_order.release();
    s->rf_row_sorted.release(); s->rf_order_sorted.release(); s->rf_tmp.release(); s->rf_applied.release();
    for (int i = 0; i < 3; ++i)
        if (s->tev[i]) cudaEventDestroy(s->tev[i]);
    if (s->own_stream && s->stream) cudaStreamDestroy(s->stream);
    delete s;
}

// ------------------------------------------------------------------------------------------------
// build: begin / add_versions / add_leaves / seal
// ------------------------------------------------------------------------------------------------
static int snapshot_new(const stage_schema_t *schema, stage_snapshot **out) {
    int nw;
    uint32_t xm;
    int rc = schema_check(schema, &nw, &xm);
    if (rc) return rc;
    CUDA_TRY(cudaSetDevice(g_device));
    auto *s = new stage_snapshot();
    s->schema = *schema;
    s->nw = nw;
    s->xormask = xm;
    s->device = g_device;
    cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        g_last_error = cudaGetErrorString(e);
        delete s;
        return STAGE_ERR_CUDA;
    }
    s->own_stream = true;
    *out = s;
    return STAGE_OK;
}

static int alloc_columns(stage_snapshot *s, uint64_t cap_leaves, uint64_t cap_rows, uint8_t *adopt_payload = nullptr) {
    // source rows travel as u32 with bit 31 = "row of the version table" (gather.cuh kVersionRowFlag): a shard holds
    // fewer than 2^31 rows (include/stage_gpu.h, stage_schema_t)
    if (cap_rows >= (uint64_t) kVersionRowFlag) return STAGE_ERR_ARG;
    s->cap_leaves = cap_leaves;
    s->cap_rows = cap_rows;
    int rc;
    // key column and its fence levels share one allocation (the search structure is one address range)
    {
        const uint64_t key_words = (((cap_rows + 8) * s->nw + 63) & ~uint64_t(63));       // + pad keys
        uint64_t lvl_words = 64;
        for (uint64_t n = cap_rows / kFan; n > 0; n /= kFan) lvl_words += ((n + 8) & ~uint64_t(3)) * s->nw;
        if ((rc = dev_alloc(s, &s->d_skey, key_words + lvl_words))) return rc;
        s->d_levels = s->d_skey + key_words;
        s->key_blob_bytes = (key_words + lvl_words) * 4;
    }
    if ((rc = dev_alloc(s, &s->d_slot, cap_rows))) return rc;
    if ((rc = dev_alloc(s, &s->d_ts, cap_rows))) return rc;
    if ((rc = dev_alloc(s, &s->d_chain, cap_rows))) return rc;
    if (adopt_payload) {                       // in-place refresh: the previous snapshot's payload buffer, same capacity
        s->d_payload = adopt_payload;
        s->device_bytes += cap_rows * (uint64_t) s->schema.payload_size + 256;
    } else if ((rc = dev_alloc(s, &s->d_payload, cap_rows * (uint64_t) s->schema.payload_size))) {
        return rc;
    }
    if ((rc = dev_alloc(s, &s->d_leaf_off, cap_leaves + 1))) return rc;
    if ((rc = dev_alloc(s, &s->d_sep, (cap_leaves + 1) * s->nw))) return rc;
    if ((rc = dev_alloc(s, &s->d_cont, cap_leaves + 1))) return rc;
    if ((rc = dev_alloc(s, &s->d_has_sep, cap_leaves + 1))) return rc;
    if ((rc = dev_alloc(s, &s->d_err, 1))) return rc;
    CUDA_TRY(cudaMemsetAsync(s->d_err, 0, 4, s->stream));
    if ((rc = dev_alloc(s, &s->d_stats, 8))) return rc;
    CUDA_TRY(cudaMemsetAsync(s->d_stats, 0, 64, s->stream));
    return STAGE_OK;
}

static int alloc_staging(stage_snapshot *s) {
    const uint32_t ls = s->schema.leaf_node_size;
    s->chunk_leaves = std::max<uint32_t>(1, (32u << 20) / ls);       // 32 MiB per staging buffer
    for (int i = 0; i < 2; ++i) {
        CUDA_TRY(cudaHostAlloc((void **) &s->h_stage[i], (size_t) s->chunk_leaves * ls, cudaHostAllocDefault));
        CUDA_TRY(cudaMalloc((void **) &s->d_stage[i], (size_t) s->chunk_leaves * ls));
        CUDA_TRY(cudaEventCreateWithFlags(&s->stage_ev[i], cudaEventDisableTiming));
    }
    return STAGE_OK;
}

static int snapshot_begin_impl(const stage_schema_t *schema, uint64_t n_leaves_hint, uint64_t n_rows_hint, uint8_t *adopt_payload,
                               stage_snapshot_t **out) {
    if (!out) return STAGE_ERR_ARG;
    stage_snapshot *s = nullptr;
    int rc = snapshot_new(schema, &s);
    if (rc) return rc;
    if ((rc = alloc_columns(s, std::max<uint64_t>(n_leaves_hint, 1), n_rows_hint, adopt_payload)) || (rc = alloc_staging(s))) {
        if (adopt_payload) s->d_payload = nullptr;       // still owned by the snapshot it came from
        snapshot_free(s);
        return rc;
    }
    s->h_leaf_off.push_back(0);
    *out = s;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_begin(const stage_schema_t *schema, uint64_t n_leaves_hint, uint64_t n_rows_hint,
                                        stage_snapshot_t **out) {
    return snapshot_begin_impl(schema, n_leaves_hint, n_rows_hint, nullptr, out);
}

// Row capacity kept free in snapshots built from a host tree (fraction of the rows), so that a later
// stage_gpu_snapshot_refresh_in_place has room for inserted rows. Process-wide, default 0.
static std::atomic<uint32_t> g_row_slack_ppm{0};
extern "C" int stage_gpu_set_row_slack(double fraction) {
    if (!(fraction >= 0.0) || fraction > 4.0) return STAGE_ERR_ARG;
    g_row_slack_ppm.store((uint32_t) (fraction * 1e6));
    return STAGE_OK;
}
static uint64_t with_row_slack(uint64_t rows) {
    const uint64_t extra = (uint64_t) ((double) rows * g_row_slack_ppm.load() * 1e-6);
    return std::min<uint64_t>(rows + extra, (uint64_t) kVersionRowFlag - 1);
}

extern "C" int stage_gpu_snapshot_add_versions(stage_snapshot_t *s, const stage_version_t *versions,
                                               const void *payloads, uint64_t n) {
    if (!s || s->sealed || (n && (!versions || !payloads))) return STAGE_ERR_ARG;
    s->h_versions.insert(s->h_versions.end(), versions, versions + n);
    const uint8_t *p = static_cast<const uint8_t *>(payloads);
    s->h_ver_payload.insert(s->h_ver_payload.end(), p, p + n * (uint64_t) s->schema.payload_size);
    return STAGE_OK;
}

template <int NW>
static int launch_transform(stage_snapshot *s, int buf, uint32_t n_chunk, uint64_t base_leaf, uint32_t maxcount,
                            const uint32_t *leaf_index = nullptr) {
    TransformArgs a{};
    a.leaf_index = leaf_index;
    a.leaves = s->d_stage[buf];
    a.n_chunk_leaves = n_chunk;
    a.leaf_size = s->schema.leaf_node_size;
    a.base_leaf = base_leaf;
    a.leaf_off = s->d_leaf_off;
    a.sep = s->d_sep;
    a.has_sep = s->d_has_sep;
    a.key_size = s->schema.key_size;
    a.payload_size = s->schema.payload_size;
    a.xormask = s->xormask;
    a.varlen = schema_varlen(&s->schema) ? 1u : 0u;
    uint32_t cap = 32;
    while (cap < maxcount) cap <<= 1;
    a.sort_cap = cap;
    a.skey = s->d_skey; a.slot = s->d_slot; a.ts = s->d_ts; a.chain = s->d_chain; a.payload = s->d_payload;
    a.cont = s->d_cont;
    a.err = s->d_err;
    size_t smem = (size_t) cap * (NW * 4 + 2);
    if (smem > 48 * 1024)
        CUDA_TRY(cudaFuncSetAttribute(leaf_transform_kernel<NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    leaf_transform_kernel<NW><<<n_chunk, 128, smem, s->stream>>>(a);
    LAUNCH_CHECK();
    return STAGE_OK;
}

// push the chunk being filled: upload its fences/offsets, H2D the images, run the transform
static int flush_chunk(stage_snapshot *s) {
    const uint32_t n = s->stage_fill;
    if (n == 0) return STAGE_OK;
    const int buf = s->stage_cur;
    const uint64_t base = s->staged_base_leaf;
    const uint32_t ls = s->schema.leaf_node_size;
    CUDA_TRY(cudaMemcpyAsync(s->d_leaf_off + base, s->h_leaf_off.data() + base, (n + 1) * sizeof(uint32_t),
                             cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_sep + base * s->nw, s->h_sep.data() + base * s->nw, (size_t) n * s->nw * 4,
                             cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_has_sep + base, s->h_has_sep.data() + base, n, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_stage[buf], s->h_stage[buf], (size_t) n * ls, cudaMemcpyHostToDevice, s->stream));
    int rc;
    switch (s->nw) {
        case 1: rc = launch_transform<1>(s, buf, n, base, s->stage_maxcount); break;
        case 2: rc = launch_transform<2>(s, buf, n, base, s->stage_maxcount); break;
        case 4: rc = launch_transform<4>(s, buf, n, base, s->stage_maxcount); break;
        default: rc = launch_transform<8>(s, buf, n, base, s->stage_maxcount); break;
    }
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(s->stage_ev[buf], s->stream));
    s->stage_cur ^= 1;
    s->stage_fill = 0;
    s->stage_maxcount = 0;
    s->staged_base_leaf = base + n;
    // the other buffer may still be in flight from two chunks ago
    CUDA_TRY(cudaEventSynchronize(s->stage_ev[s->stage_cur]));
    return STAGE_OK;
}

// In-flight (overwritten) versions, SURVEY 8f f2: a record whose update has not committed (control bit set,
// next_ptr = location of its undo copy in the EphemeralPool, b_tree.cpp:1122-1159) is staged as a reader sees it
// (BTree::Read, b_tree.cpp:2087-2113): the payload of the undo copy, the unchanged commit id, and the chain the
// undo header saved. Without an undo copy (an in-flight insert, or the copy is gone) the read misses: the row is
// dropped from the snapshot.
struct InflightPatch {
    uint32_t slot;
    const uint8_t *old_payload;      // nullptr = drop the row
};
struct LeafExtras {
    const uint32_t *heads = nullptr;           // [record_count] 1-based version index per slot (0 = none)
    const InflightPatch *patches = nullptr;    // in-flight records of this leaf, resolved by the caller
    uint32_t n_patches = 0;
    const uint64_t *expect_fp = nullptr;       // live trees: fingerprint of header + record metadata seen by the discovery pass
};
static uint64_t leaf_fingerprint(const uint8_t *leaf, const uint8_t *fence, uint32_t K, uint32_t leaf_size);
constexpr int kStageRetry = -100;              // internal: the leaf moved between the discovery pass and its copy

// Copy one leaf image into its pinned staging slot, count its visible records and replace the host TupleHeader
// pointers in RecordMetadata.next_ptr by version indices. Thread-safe: touches only `dst`.
static int stage_copy_leaf(const stage_snapshot *s, uint8_t *dst, const uint8_t *leaf, const LeafExtras &ex,
                           uint32_t *vis_out, uint32_t *count_out) {
    const uint32_t ls = s->schema.leaf_node_size;
    if (!leaf[kOffIsLeaf]) return STAGE_ERR_LAYOUT;
    uint64_t status;
    memcpy(&status, leaf + kOffStatus, 8);
    if (status & kStatusFrozen) return STAGE_ERR_NOT_QUIESCED;
    const uint32_t count = status_record_count(status);
    if ((uint64_t) kOffMeta + (uint64_t) count * kMetaSize > ls) return STAGE_ERR_LAYOUT;
    memcpy(dst, leaf, ls);
    if (ex.expect_fp) {
        // writers are running: what was copied (and what the leaf holds right after the copy) must still be what the
        // discovery pass saw — every begin / commit / abort of an update rewrites a metadata word or the chain pointer
        // (b_tree.cpp:1122-1159, transaction_manager.cpp:640-664). Otherwise the caller re-discovers and re-copies.
        if (leaf_fingerprint(dst, nullptr, 0, ls) != *ex.expect_fp || leaf_fingerprint(leaf, nullptr, 0, ls) != *ex.expect_fp)
            return kStageRetry;
    }
    for (uint32_t p = 0; p < ex.n_patches; ++p) {
        const InflightPatch &ip = ex.patches[p];
        if (ip.slot >= count) return STAGE_ERR_LAYOUT;
        uint8_t *mp = dst + kOffMeta + (size_t) ip.slot * kMetaSize;
        uint64_t m;
        memcpy(&m, mp, 8);
        m &= ~kMetaControl;
        if (!ip.old_payload) m &= ~kMetaVisible;
        else {
            if (meta_offset(m) + pad_key_len(meta_key_len(m)) + s->schema.payload_size > ls) return STAGE_ERR_LAYOUT;
            memcpy(dst + meta_offset(m) + pad_key_len(meta_key_len(m)), ip.old_payload, s->schema.payload_size);
        }
        memcpy(mp, &m, 8);
    }
    uint32_t vis = 0;
    for (uint32_t i = 0; i < count; ++i) {
        uint8_t *mp = dst + kOffMeta + (size_t) i * kMetaSize;
        uint64_t m;
        memcpy(&m, mp, 8);
        uint64_t head = 0;
        if (meta_visible(m)) {
            if (m & kMetaControl) return STAGE_ERR_NOT_QUIESCED;
            if (meta_offset(m) + pad_key_len(meta_key_len(m)) + s->schema.payload_size > ls) return STAGE_ERR_LAYOUT;
            ++vis;
            if (ex.heads) head = ex.heads[i];
        }
        memcpy(mp + kMetaNextPtr, &head, 8);   // pointer translation: host TupleHeader* -> version index
    }
    *vis_out = vis;
    *count_out = count;
    return STAGE_OK;
}

static unsigned staging_threads() {
    static const unsigned n = [] {
        if (const char *e = getenv("STAGE_STAGING_THREADS")) return (unsigned) std::max(1, atoi(e));
        return std::min(32u, std::max(1u, std::thread::hardware_concurrency()));
    }();
    return n;
}

// Host threads that copy leaf images / version rows into the pinned staging buffers. They live for the process (a
// flattening call used to spawn and join its threads once per 32 MiB chunk: ~4000 thread creations for the 10 M-row
// table); work is handed out in grains through one atomic counter, the calling thread takes part.
class WorkPool {
public:
    static WorkPool &get() {
        static WorkPool pool(staging_threads() > 1 ? staging_threads() - 1 : 0);
        return pool;
    }
    void run(uint64_t n, uint64_t grain, const std::function<void(uint64_t, uint64_t)> &f) {
        if (n == 0) return;
        grain = std::max<uint64_t>(grain, 1);
        if (th_.empty() || n <= grain) {
            f(0, n);
            return;
        }
        std::lock_guard<std::mutex> one(run_mu_);          // one parallel region at a time
        {
            std::lock_guard<std::mutex> lk(mu_);
            job_ = &f;
            n_ = n;
            grain_ = grain;
            next_.store(0);
            active_ = (unsigned) th_.size();
            ++gen_;
        }
        cv_start_.notify_all();
        drain();
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [&] { return active_ == 0; });
        job_ = nullptr;
    }
    ~WorkPool() {
        {
            std::lock_guard<std::mutex> lk(mu_);
            quit_ = true;
        }
        cv_start_.notify_all();
        for (auto &t : th_) t.join();
    }

private:
    explicit WorkPool(unsigned workers) {
        for (unsigned i = 0; i < workers; ++i) th_.emplace_back([this] { loop(); });
    }
    void drain() {
        for (;;) {
            const uint64_t lo = next_.fetch_add(grain_);
            if (lo >= n_) break;
            (*job_)(lo, std::min(n_, lo + grain_));
        }
    }
    void loop() {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(mu_);
        for (;;) {
            cv_start_.wait(lk, [&] { return quit_ || gen_ != seen; });
            if (quit_) return;
            seen = gen_;
            lk.unlock();
            drain();
            lk.lock();
            if (--active_ == 0) cv_done_.notify_one();
        }
    }
    std::vector<std::thread> th_;
    std::mutex mu_, run_mu_;
    std::condition_variable cv_start_, cv_done_;
    const std::function<void(uint64_t, uint64_t)> *job_ = nullptr;
    uint64_t n_ = 0, grain_ = 1, gen_ = 0;
    std::atomic<uint64_t> next_{0};
    unsigned active_ = 0;
    bool quit_ = false;
};

// Stage n leaves (tree order). The 64 KiB image copies — the bulk of the host work — run on a few threads, one
// staging chunk at a time; bookkeeping (row offsets, fences) and the chunk flush stay on the calling thread.
// get(i, &leaf, &fence, &extras) supplies leaf i.
struct NoRedo {
    int operator()(uint64_t) const { return STAGE_ERR_NOT_QUIESCED; }
};
// redo(i): re-run the discovery pass for leaf i (serial, on the calling thread) after its copy found the leaf changed.
template <class Get, class Redo = NoRedo>
static int stage_leaves(stage_snapshot *s, uint64_t n, Get get, Redo redo = Redo()) {
    const uint32_t ls = s->schema.leaf_node_size;
    std::vector<uint32_t> vis, cnt;
    std::vector<int> rcs;
    for (uint64_t base = 0; base < n;) {
        const uint32_t room = s->chunk_leaves - s->stage_fill;
        const uint32_t m = (uint32_t) std::min<uint64_t>(room, n - base);
        if (s->n_leaves + m > s->cap_leaves) return STAGE_ERR_ARG;
        vis.assign(m, 0); cnt.assign(m, 0); rcs.assign(m, 0);
        uint8_t *slot0 = s->h_stage[s->stage_cur] + (size_t) s->stage_fill * ls;
        auto work = [&](uint32_t lo, uint32_t hi) {
            for (uint32_t j = lo; j < hi; ++j) {
                const uint8_t *leaf, *fence;
                LeafExtras ex;
                get(base + j, &leaf, &fence, &ex);
                rcs[j] = stage_copy_leaf(s, slot0 + (size_t) j * ls, leaf, ex, &vis[j], &cnt[j]);
            }
        };
        WorkPool::get().run(m, 4, [&](uint64_t lo, uint64_t hi) { work((uint32_t) lo, (uint32_t) hi); });
        for (uint32_t j = 0; j < m; ++j) {
            for (int attempt = 0; rcs[j] == kStageRetry; ++attempt) {      // a writer touched the leaf under the copy
                if (attempt >= 64) return STAGE_ERR_NOT_QUIESCED;
                int rr = redo(base + j);
                if (rr) return rr;
                const uint8_t *leaf2, *fence2;
                LeafExtras ex2;
                get(base + j, &leaf2, &fence2, &ex2);
                rcs[j] = stage_copy_leaf(s, slot0 + (size_t) j * ls, leaf2, ex2, &vis[j], &cnt[j]);
            }
            if (rcs[j]) return rcs[j];
            const uint8_t *leaf, *fence;
            LeafExtras ex;
            get(base + j, &leaf, &fence, &ex);
            if (s->rows_staged + vis[j] > s->cap_rows) return STAGE_ERR_ARG;
            s->rows_staged += vis[j];
            s->h_leaf_off.push_back((uint32_t) s->rows_staged);
            s->h_sep.resize((s->n_leaves + 1) * s->nw, 0xFFFFFFFFu);
            if (fence) host_sortable_nw(s->nw, fence, s->schema.key_size, s->xormask, s->h_sep.data() + s->n_leaves * s->nw);
            s->h_has_sep.push_back(fence ? 1 : 0);
            s->n_leaves++;
            s->stage_fill++;
            s->stage_maxcount = std::max(s->stage_maxcount, cnt[j]);
            s->max_leaf_rows = std::max(s->max_leaf_rows, vis[j]);
        }
        base += m;
        if (s->stage_fill == s->chunk_leaves) {
            int rc = flush_chunk(s);
            if (rc) return rc;
        }
    }
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_add_leaves(stage_snapshot_t *s, const void *const *leaf_images,
                                             const void *upper_fences, const uint8_t *has_upper_fence,
                                             const uint32_t *const *chain_heads, uint64_t n) {
    if (!s || s->sealed || (n && !leaf_images)) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(s->device));
    const uint8_t *f = static_cast<const uint8_t *>(upper_fences);
    const uint32_t K = s->schema.key_size;
    return stage_leaves(s, n, [&](uint64_t i, const uint8_t **leaf, const uint8_t **fence, LeafExtras *ex) {
        *leaf = static_cast<const uint8_t *>(leaf_images[i]);
        const bool hf = f && (!has_upper_fence || has_upper_fence[i]);
        *fence = hf ? f + i * K : nullptr;
        ex->heads = chain_heads ? chain_heads[i] : nullptr;
    });
}

template <int NW>
static int build_levels(stage_snapshot *s) {
    // implicit kFan-ary fence levels over the key column; the top level must fit in shared memory.
    // Every level (and the key column) is followed by >= kFan all-ones pad keys: fence groups never
    // need bounds checks.
    const uint64_t top_max = (kProbeTopBytes) / (NW * 4);
    std::vector<uint64_t> sizes{s->n_rows};
    while (sizes.back() > top_max) sizes.push_back(sizes.back() / kFan);
    if ((int) sizes.size() > kMaxLevels) return STAGE_ERR_ARG;
    SnapshotView &v = s->view;
    v.n_levels = (int) sizes.size();
    v.level[0] = s->d_skey;
    v.level_n[0] = (uint32_t) sizes[0];
    CUDA_TRY(cudaMemsetAsync(s->d_skey + s->n_rows * NW, 0xFF, 8 * NW * 4, s->stream));   // key-column pad
    uint64_t off = 0;
    for (size_t l = 1; l < sizes.size(); ++l) {
        uint32_t *dst = s->d_levels + off * NW;
        const uint64_t nd = sizes[l], npad = (nd + 8) & ~uint64_t(3);
        sample_level_kernel<NW><<<(unsigned) ((npad + 255) / 256), 256, 0, s->stream>>>(v.level[l - 1], dst, nd, npad);
        LAUNCH_CHECK();
        v.level[l] = dst;
        v.level_n[l] = (uint32_t) nd;
        off += npad;
    }
    v.smem_words = (uint32_t) (sizes.back() * NW);
    return STAGE_OK;
}

static void fill_view(stage_snapshot *s) {
    SnapshotView &v = s->view;
    v.slot = s->d_slot; v.ts = s->d_ts; v.chain = s->d_chain; v.payload = s->d_payload;
    v.leaf_off = s->d_leaf_off; v.sep = s->d_sep; v.cont = s->d_cont;
    v.ver_begin = s->d_ver_begin; v.ver_commit = s->d_ver_commit; v.ver_next = s->d_ver_next;
    v.ver_payload = s->d_ver_payload;
    v.n_rows = s->n_rows; v.n_leaves = s->n_leaves;
    v.key_size = s->schema.key_size; v.payload_size = s->schema.payload_size; v.xormask = s->xormask;
    v.max_leaf_rows = s->max_leaf_rows;
    v.dir = s->d_dir; v.dir_cap = s->dir_cap;
}

// point-probe directory over the sealed key column (probe.cuh); STAGE_PROBE_DIR=0 keeps the fence search
static int build_directory(stage_snapshot *s) {
    static const bool off = getenv("STAGE_PROBE_DIR") && atoi(getenv("STAGE_PROBE_DIR")) == 0;
    if (off || s->n_rows == 0 || s->n_rows > 0x7FFFFF00ull) return STAGE_OK;      // 2n + 64 must fit 32 bits
    if (s->d_dir) { cudaFree(s->d_dir); s->d_dir = nullptr; }
    const uint32_t cap = (uint32_t) (2 * s->n_rows + 64);
    if (dev_alloc(s, &s->d_dir, cap) != STAGE_OK) {          // optional structure: fall back to the fences
        cudaGetLastError();
        s->d_dir = nullptr;
        return STAGE_OK;
    }
    s->dir_cap = cap;
    CUDA_TRY(cudaMemsetAsync(s->d_dir, 0xFF, (size_t) cap * 8, s->stream));
    const unsigned blocks = (unsigned) ((s->n_rows + 255) / 256);
    unsigned long long *d = reinterpret_cast<unsigned long long *>(s->d_dir);
    switch (s->nw) {
        case 1: dir_build_kernel<1><<<blocks, 256, 0, s->stream>>>(s->d_skey, s->n_rows, d, cap); break;
        case 2: dir_build_kernel<2><<<blocks, 256, 0, s->stream>>>(s->d_skey, s->n_rows, d, cap); break;
        case 4: dir_build_kernel<4><<<blocks, 256, 0, s->stream>>>(s->d_skey, s->n_rows, d, cap); break;
        default: dir_build_kernel<8><<<blocks, 256, 0, s->stream>>>(s->d_skey, s->n_rows, d, cap); break;
    }
    LAUNCH_CHECK();
    return STAGE_OK;
}

static int finish_seal(stage_snapshot *s) {
    int rc;
    switch (s->nw) {
        case 1: rc = build_levels<1>(s); break;
        case 2: rc = build_levels<2>(s); break;
        case 4: rc = build_levels<4>(s); break;
        default: rc = build_levels<8>(s); break;
    }
    if (rc) return rc;
    if ((rc = build_directory(s))) return rc;
    fill_view(s);
    uint32_t err = 0;
    CUDA_TRY(cudaMemcpyAsync(&err, s->d_err, 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (err == 1) return STAGE_ERR_NOT_QUIESCED;
    if (err) return STAGE_ERR_LAYOUT;
    s->sealed = true;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_seal(stage_snapshot_t *s) {
    if (!s || s->sealed) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(s->device));
    int rc = flush_chunk(s);
    if (rc) return rc;
    s->n_rows = s->rows_staged;
    // versions (from_tree uploads them itself through the staging buffers: versions_on_device)
    if (!s->versions_on_device) s->n_versions = s->h_versions.size();
    if (s->n_versions && !s->versions_on_device) {
        const uint64_t V = s->n_versions;
        if ((rc = dev_alloc(s, &s->d_ver_begin, V)) || (rc = dev_alloc(s, &s->d_ver_commit, V)) ||
            (rc = dev_alloc(s, &s->d_ver_next, V)) ||
            (rc = dev_alloc(s, &s->d_ver_payload, V * (uint64_t) s->schema.payload_size)))
            return rc;
        s->ver_cap = V;
        std::vector<uint32_t> b(V), c(V), nx(V);
        for (uint64_t i = 0; i < V; ++i) {
            b[i] = s->h_versions[i].begin_id; c[i] = s->h_versions[i].commit_id; nx[i] = s->h_versions[i].next;
            if (nx[i] > V) return STAGE_ERR_ARG;
        }
        CUDA_TRY(cudaMemcpyAsync(s->d_ver_begin, b.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->d_ver_commit, c.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->d_ver_next, nx.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->d_ver_payload, s->ext_ver_payload ? s->ext_ver_payload : s->h_ver_payload.data(),
                                 V * (uint64_t) s->schema.payload_size, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    rc = finish_seal(s);
    // staging no longer needed
    for (int i = 0; i < 2; ++i) {
        if (s->h_stage[i]) { cudaFreeHost(s->h_stage[i]); s->h_stage[i] = nullptr; }
        if (s->d_stage[i]) { cudaFree(s->d_stage[i]); s->d_stage[i] = nullptr; }
    }
    std::vector<uint8_t>().swap(s->h_ver_payload);
    std::vector<stage_version_t>().swap(s->h_versions);
    free(s->ext_ver_payload);
    s->ext_ver_payload = nullptr;
    return rc;
}

// ------------------------------------------------------------------------------------------------
// host tree walker: root -> leaves in tree order, fences, version chains
// ------------------------------------------------------------------------------------------------
struct WalkOut {
    std::vector<const uint8_t *> leaves;
    std::vector<const uint8_t *> fences;   // fence[i] = separator to the right of leaf i (nullptr for the last)
    uint64_t rows_upper = 0;
    // variable-length keys: separators are packed into slots here (fences[] then points into this store)
    bool varlen = false;
    uint32_t xormask = 0;
    std::deque<std::array<uint8_t, 32>> slots;
    void key_form(const stage_schema_t *sc, uint32_t xm) {
        varlen = schema_varlen(sc);
        xormask = xm;
    }
};

static int walk(const uint8_t *node, uint32_t key_size, WalkOut &w, int depth) {
    if (!node || depth > 32) return STAGE_ERR_LAYOUT;
    if (node[kOffIsLeaf]) {
        uint64_t status;
        memcpy(&status, node + kOffStatus, 8);
        w.rows_upper += status_record_count(status);
        w.leaves.push_back(node);
        w.fences.push_back(nullptr);
        return STAGE_OK;
    }
    uint32_t sorted;
    memcpy(&sorted, node + kOffSortedCount, 4);
    for (uint32_t i = 0; i < sorted; ++i) {
        uint64_t m;
        memcpy(&m, node + kOffMeta + (size_t) i * kMetaSize, 8);
        const uint8_t *rec = node + meta_offset(m);
        if (i > 0) {
            // separator i sits between child i-1 and child i: upper fence of the last leaf emitted
            if (w.leaves.empty()) return STAGE_ERR_LAYOUT;
            if (w.varlen) {
                const uint32_t len = meta_key_len(m);
                if (len + 1 > key_size || (w.xormask == 0 && len < 16)) return STAGE_ERR_LAYOUT;
                w.slots.emplace_back();
                pack_key_slot(rec, len, key_size, w.xormask, w.slots.back().data());
                w.fences.back() = w.slots.back().data();
            } else {
                if (meta_key_len(m) != key_size) return STAGE_ERR_LAYOUT;
                w.fences.back() = rec;
            }
        }
        uint64_t child;
        memcpy(&child, rec + pad_key_len(meta_key_len(m)), 8);      // GetChildByMetaIndex, b_tree.h:336-345
        int rc = walk(reinterpret_cast<const uint8_t *>(child), key_size, w, depth + 1);
        if (rc) return rc;
    }
    return STAGE_OK;
}

// Version chains of a host tree: RecordMetadata.next_ptr -> TupleHeader (version_store.h:28-46), slot = [key][payload].
// Stages every header reachable from the chain heads once; indices are 1-based and start after `base` versions that
// already exist on the device (structural refresh keeps the previous version table as a prefix).
struct ChainStager {
    std::unordered_map<uint64_t, uint32_t> seen;   // TupleHeader* -> 1-based version index
    std::vector<stage_version_t> vers;
    std::vector<uint64_t> vslot;                   // TupleHeader.tuple_slot_ of every staged version
    std::vector<uint64_t> path;
    uint32_t base = 0;
    uint32_t add_chain(uint64_t head) {
        path.clear();
        uint32_t tail = 0;
        const uint32_t first = base + (uint32_t) vers.size() + 1;
        for (uint64_t cur = head; cur != 0 && cur != UINT64_MAX;) {
            auto it = seen.find(cur);
            if (it != seen.end()) { tail = it->second; break; }      // joins a chain already staged (or a cycle)
            seen.emplace(cur, first + (uint32_t) path.size());
            path.push_back(cur);
            memcpy(&cur, reinterpret_cast<const uint8_t *>(cur) + kThNext, 8);
        }
        for (size_t k = 0; k < path.size(); ++k) {
            const uint8_t *hd = reinterpret_cast<const uint8_t *>(path[k]);
            stage_version_t ver{};
            uint64_t slot;
            memcpy(&ver.begin_id, hd + kThBegin, 4);
            memcpy(&ver.commit_id, hd + kThCommit, 4);
            memcpy(&slot, hd + kThSlot, 8);
            ver.next = k + 1 < path.size() ? first + (uint32_t) k + 1 : tail;
            vers.push_back(ver);
            vslot.push_back(slot);
        }
        return path.empty() ? tail : first;
    }
    // dense version rows of everything staged (malloc'd; nullptr when out of memory)
    uint8_t *payload_rows(uint32_t K, uint32_t P) const {
        const size_t V = vers.size();
        uint8_t *vbuf = static_cast<uint8_t *>(malloc(std::max<size_t>(V, 1) * (size_t) P));
        if (!vbuf) return nullptr;
        auto work = [&](size_t lo, size_t hi) {
            for (size_t i = lo; i < hi; ++i) {
                if (vslot[i] != 0 && vslot[i] != UINT64_MAX) memcpy(vbuf + i * P, reinterpret_cast<const uint8_t *>(vslot[i]) + K, P);
                else memset(vbuf + i * P, 0, P);
            }
        };
        WorkPool::get().run(V, 1024, [&](uint64_t lo, uint64_t hi) { work((size_t) lo, (size_t) hi); });
        return vbuf;
    }
};

// Change detector of the structural refresh: a leaf whose header (sizes, sorted count, status word), record
// metadata array (visibility, offsets, commit ids, chain heads) and upper fence hash to the same value at the same
// address holds the same rows as in the previous snapshot (every committed insert, update and delete rewrites a
// metadata word: b_tree.cpp:863-898 insert, transaction_manager.cpp:600-766 commit install).
static uint64_t mix64(uint64_t h, uint64_t v) {
    h ^= v + 0x9E3779B97F4A7C15ull + (h << 6) + (h >> 2);
    h *= 0xFF51AFD7ED558CCDull;
    return h ^ (h >> 32);
}
static uint64_t leaf_fingerprint(const uint8_t *leaf, const uint8_t *fence, uint32_t K, uint32_t leaf_size) {
    uint64_t status;
    memcpy(&status, leaf + kOffStatus, 8);
    const uint32_t count = status_record_count(status);
    uint64_t h = 0x5354414745ull;
    size_t end = (size_t) kOffMeta + (size_t) count * kMetaSize;
    if (end > leaf_size) end = leaf_size;
    for (size_t o = kOffIsLeaf; o + 8 <= end; o += 8) {
        uint64_t v;
        memcpy(&v, leaf + o, 8);
        h = mix64(h, v);
    }
    if (fence) {
        for (uint32_t b = 0; b < K; ++b) h = mix64(h, fence[b] + 256ull * b);
    } else {
        h = mix64(h, 0xFEull << 32);
    }
    return h;
}

// Discovery pass over the leaves that will be staged (serial, metadata only): the version chains hanging off the records,
// the in-flight records resolved through the caller's undo pool — their old payload is COPIED here, the pool slot may be
// recycled before the copy pass runs — and, for live trees, the fingerprint of header + record metadata that the copy
// pass validates against (stage_copy_leaf). discover() can be repeated for one leaf: its blocks are appended again and
// the previous ones are simply no longer referenced.
struct LeafDiscovery {
    const stage_schema_t *schema = nullptr;
    stage_undo_resolver_t resolve = nullptr;
    void *ctx = nullptr;
    ChainStager cs;
    std::vector<uint32_t> heads_flat;
    std::vector<int64_t> heads_at;                            // per leaf: offset into heads_flat, -1 = no chain
    std::vector<InflightPatch> patches;                       // blocks per leaf, slot order
    std::vector<std::pair<size_t, uint32_t>> patch_at;
    std::vector<uint32_t> dropped;                            // in-flight rows without an undo copy
    std::vector<uint64_t> fp;
    std::deque<std::vector<uint8_t>> undo_copies;             // stable addresses
    std::vector<uint32_t> heads;                              // scratch
    std::vector<InflightPatch> tmp;

    void init(size_t n) {
        heads_at.assign(n, -1);
        patch_at.assign(n, {0, 0});
        dropped.assign(n, 0);
        fp.assign(resolve ? n : 0, 0);
    }
    int discover(size_t idx, const uint8_t *leaf) {
        const uint32_t ls = schema->leaf_node_size, P = schema->payload_size;
        for (int spin = 0;; ++spin) {
            if (spin > 100000) return STAGE_ERR_NOT_QUIESCED;
            const uint64_t f = resolve ? leaf_fingerprint(leaf, nullptr, 0, ls) : 0;
            uint64_t status;
            memcpy(&status, leaf + kOffStatus, 8);
            const uint32_t count = status_record_count(status);
            if ((uint64_t) kOffMeta + (uint64_t) count * kMetaSize > ls) return STAGE_ERR_LAYOUT;
            heads.assign(count, 0);
            tmp.clear();
            bool any = false, moving = false;
            uint32_t drop = 0;
            for (uint32_t i = 0; i < count && !moving; ++i) {
                uint64_t m, nx;
                memcpy(&m, leaf + kOffMeta + (size_t) i * kMetaSize, 8);
                memcpy(&nx, leaf + kOffMeta + (size_t) i * kMetaSize + kMetaNextPtr, 8);
                if (!meta_visible(m)) continue;
                if ((m & kMetaControl) && !resolve) return STAGE_ERR_NOT_QUIESCED;   // next_ptr is an undo copy, not a chain
                const bool has_next = nx != 0 && nx != UINT64_MAX;
                if (m & kMetaControl) {
                    // in flight: the reader's version is the undo copy [key][payload]; its header saved the chain
                    uint64_t saved_next = 0;
                    const void *old_rec = nullptr;
                    const bool have = has_next && resolve(ctx, nx, &saved_next, &old_rec) == 1 && old_rec;
                    const uint8_t *pay = nullptr;
                    if (!have && spin < 2000) {
                        // control bit set but no undo copy behind next_ptr: either an insert in flight (stays like this
                        // until its transaction ends) or an update caught between its two compare-and-swaps
                        // (LeafNode::Update, b_tree.cpp:1122-1148: the copy is published a moment later). Look again.
                        moving = true;
                        continue;
                    }
                    if (have) {
                        undo_copies.emplace_back(static_cast<const uint8_t *>(old_rec) + meta_key_len(m),
                                                 static_cast<const uint8_t *>(old_rec) + meta_key_len(m) + P);
                        pay = undo_copies.back().data();
                    } else {
                        ++drop;
                    }
                    tmp.push_back({i, pay});
                    nx = have ? saved_next : 0;
                } else if (resolve && has_next) {
                    // committed record of a live tree: between the two compare-and-swaps of CommitTransaction
                    // (transaction_manager.cpp:651-664) next_ptr still names the undo copy, not a TupleHeader
                    uint64_t sn = 0;
                    const void *rec = nullptr;
                    if (resolve(ctx, nx, &sn, &rec) == 1) moving = true;
                }
                if (moving || nx == 0 || nx == UINT64_MAX) continue;
                heads[i] = cs.add_chain(nx);
                any = true;
            }
            if (moving || (resolve && leaf_fingerprint(leaf, nullptr, 0, ls) != f)) {            // re-read the leaf
                if (spin > 16) std::this_thread::yield();
                continue;
            }
            if (resolve) fp[idx] = f;
            heads_at[idx] = -1;
            if (any) {
                heads_at[idx] = (int64_t) heads_flat.size();
                heads_flat.insert(heads_flat.end(), heads.begin(), heads.end());
            }
            patch_at[idx] = {patches.size(), (uint32_t) tmp.size()};
            patches.insert(patches.end(), tmp.begin(), tmp.end());
            dropped[idx] = drop;
            return STAGE_OK;
        }
    }
    void fill(size_t idx, LeafExtras *ex) const {
        ex->heads = heads_at[idx] >= 0 ? heads_flat.data() + heads_at[idx] : nullptr;
        if (patch_at[idx].second) {
            ex->patches = patches.data() + patch_at[idx].first;
            ex->n_patches = patch_at[idx].second;
        }
        ex->expect_fp = resolve ? &fp[idx] : nullptr;
    }
};

// Discovery over n leaves. Quiesced trees: contiguous ranges of leaves are discovered by the staging threads, each with
// its own chain stager (version chains belong to one record each, so ranges never share a TupleHeader), then the parts are
// concatenated and their version indices shifted. Live trees stay serial: the caller's resolver runs on the calling thread.
template <class LeafAt>
static int discover_all(LeafDiscovery &d, size_t n, LeafAt leaf_at) {
    const unsigned T = (d.resolve || n < 4096) ? 1u : std::min<unsigned>(staging_threads(), 64u);
    if (T <= 1) {
        for (size_t i = 0; i < n; ++i) {
            const int rc = d.discover(i, leaf_at(i));
            if (rc) return rc;
        }
        return STAGE_OK;
    }
    std::vector<LeafDiscovery> part(T);
    std::vector<int> rcs(T, STAGE_OK);
    WorkPool::get().run(T, 1, [&](uint64_t tlo, uint64_t thi) {
        for (uint64_t t = tlo; t < thi; ++t) {
            const size_t lo = n * t / T, hi = n * (t + 1) / T;
            LeafDiscovery &p = part[t];
            p.schema = d.schema;
            p.cs.base = d.cs.base;
            p.init(hi - lo);
            for (size_t i = lo; i < hi && !rcs[t]; ++i) rcs[t] = p.discover(i - lo, leaf_at(i));
        }
    });
    for (unsigned t = 0; t < T; ++t)
        if (rcs[t]) return rcs[t];
    // concatenate the parts (offsets first, then every part copies itself into place in parallel)
    std::vector<size_t> voff(T + 1, d.cs.vers.size()), hoff(T + 1, d.heads_flat.size());
    for (unsigned t = 0; t < T; ++t) {
        voff[t + 1] = voff[t] + part[t].cs.vers.size();
        hoff[t + 1] = hoff[t] + part[t].heads_flat.size();
    }
    d.cs.vers.resize(voff[T]);
    d.cs.vslot.resize(voff[T]);
    d.heads_flat.resize(hoff[T]);
    WorkPool::get().run(T, 1, [&](uint64_t tlo, uint64_t thi) {
        for (uint64_t t = tlo; t < thi; ++t) {
            const size_t lo = n * t / T, hi = n * (t + 1) / T;
            LeafDiscovery &p = part[t];
            const uint32_t shift = (uint32_t) (voff[t] - voff[0]);       // versions staged by the parts before this one
            for (size_t i = 0; i < p.cs.vers.size(); ++i) {
                stage_version_t v = p.cs.vers[i];
                if (v.next) v.next += shift;
                d.cs.vers[voff[t] + i] = v;
                d.cs.vslot[voff[t] + i] = p.cs.vslot[i];
            }
            for (size_t i = 0; i < p.heads_flat.size(); ++i) d.heads_flat[hoff[t] + i] = p.heads_flat[i] ? p.heads_flat[i] + shift : 0u;
            for (size_t i = lo; i < hi; ++i) {
                d.heads_at[i] = p.heads_at[i - lo] >= 0 ? p.heads_at[i - lo] + (int64_t) hoff[t] : -1;
                d.dropped[i] = p.dropped[i - lo];
            }
        }
    });
    return STAGE_OK;
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// Version table of a freshly walked tree -> device: headers in three small copies, the retired rows [key][payload] ->
// dense payload rows gathered by the staging threads into the pinned staging buffers and copied chunk by chunk (DMA at
// link speed; a pageable 1 GB source used to go through the driver's bounce buffers).
static int upload_versions_staged(stage_snapshot *s, const ChainStager &cs) {
    const uint64_t V = cs.vers.size();
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    int rc = flush_chunk(s);                       // the last (partial) chunk of leaf images leaves the buffers first
    if (rc) return rc;
    if ((rc = dev_alloc(s, &s->d_ver_begin, V)) || (rc = dev_alloc(s, &s->d_ver_commit, V)) || (rc = dev_alloc(s, &s->d_ver_next, V)) ||
        (rc = dev_alloc(s, &s->d_ver_payload, V * (uint64_t) P)))
        return rc;
    s->ver_cap = V;
    std::vector<uint32_t> b(V), c(V), nx(V);
    for (uint64_t i = 0; i < V; ++i) {
        b[i] = cs.vers[i].begin_id; c[i] = cs.vers[i].commit_id; nx[i] = cs.vers[i].next;
        if (nx[i] > V) return STAGE_ERR_ARG;
    }
    CUDA_TRY(cudaMemcpyAsync(s->d_ver_begin, b.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_ver_commit, c.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
    CUDA_TRY(cudaMemcpyAsync(s->d_ver_next, nx.data(), V * 4, cudaMemcpyHostToDevice, s->stream));
    const uint64_t per = std::max<uint64_t>(1, (uint64_t) s->chunk_leaves * s->schema.leaf_node_size / P);
    for (uint64_t v0 = 0; v0 < V; v0 += per) {
        const uint64_t m = std::min(per, V - v0);
        const int buf = s->stage_cur;
        CUDA_TRY(cudaEventSynchronize(s->stage_ev[buf]));
        uint8_t *dst = s->h_stage[buf];
        WorkPool::get().run(m, 512, [&](uint64_t lo, uint64_t hi) {
            for (uint64_t i = lo; i < hi; ++i) {
                const uint64_t slot = cs.vslot[v0 + i];
                if (slot != 0 && slot != UINT64_MAX) memcpy(dst + i * P, reinterpret_cast<const uint8_t *>(slot) + K, P);
                else memset(dst + i * P, 0, P);
            }
        });
        CUDA_TRY(cudaMemcpyAsync(s->d_ver_payload + v0 * P, dst, m * P, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaEventRecord(s->stage_ev[buf], s->stream));
        s->stage_cur ^= 1;
    }
    CUDA_TRY(cudaStreamSynchronize(s->stream));    // b / c / nx are pageable locals
    s->n_versions = V;
    s->versions_on_device = true;
    return STAGE_OK;
}

static int from_tree_impl(const void *root, const stage_schema_t *schema, stage_undo_resolver_t resolve, void *resolve_ctx,
                          stage_snapshot_t **out) {
    if (!root || !out) return STAGE_ERR_ARG;
    static const bool dbg = getenv("STAGE_DEBUG_TIMING") != nullptr;
    double t0 = now_s(), t1;
#define PHASE(name) do { if (dbg) { t1 = now_s(); fprintf(stderr, "[from_tree] %-22s %.3f s\n", name, t1 - t0); t0 = t1; } } while (0)
    WalkOut w;
    int nw;
    uint32_t xm;
    int rc = schema_check(schema, &nw, &xm);
    if (rc) return rc;
    // variable-length keys: quiesced trees without version chains (the version store's slots are [key][payload] of the
    // catalog's fixed key length, version_store.h:445-481; no variable-length table goes through it)
    if (schema_varlen(schema) && resolve) return STAGE_ERR_ARG;
    w.key_form(schema, xm);
    if ((rc = walk(static_cast<const uint8_t *>(root), schema->key_size, w, 0))) return rc;

    PHASE("walk");
    stage_snapshot *s = nullptr;
    if ((rc = stage_gpu_snapshot_begin(schema, w.leaves.size(), with_row_slack(w.rows_upper), &s))) return rc;
    PHASE("alloc");

    const uint32_t K = schema->key_size, P = schema->payload_size;
    LeafDiscovery d;
    d.schema = schema;
    d.resolve = resolve;
    d.ctx = resolve_ctx;
    d.init(w.leaves.size());
    ChainStager &cs = d.cs;
    std::vector<stage_version_t> &vers = cs.vers;
    // pass 1 (metadata only): version indices in leaf / slot order, in-flight records
    rc = discover_all(d, w.leaves.size(), [&](size_t li) { return w.leaves[li]; });
    PHASE("chain discovery");
    // pass 2: leaf images, copied by the staging threads (live trees: validated against pass 1, re-discovered on a miss)
    if (!rc)
        rc = stage_leaves(s, w.leaves.size(),
                          [&](uint64_t li, const uint8_t **leaf, const uint8_t **fence, LeafExtras *ex) {
                              *leaf = w.leaves[li];
                              *fence = w.fences[li];
                              d.fill(li, ex);
                          },
                          [&](uint64_t li) { return d.discover(li, w.leaves[li]); });
    PHASE("stage + transform");
    if (!rc && !vers.empty() && schema_varlen(schema)) rc = STAGE_ERR_LAYOUT;
    if (!rc && !vers.empty()) rc = upload_versions_staged(s, cs);
    if (!rc) {       // fingerprints for a later stage_gpu_snapshot_refresh_from_tree
        s->h_leaf_addr.resize(w.leaves.size());
        s->h_leaf_fp.resize(w.leaves.size());
        WorkPool::get().run(w.leaves.size(), 512, [&](uint64_t lo, uint64_t hi) {
            for (uint64_t li = lo; li < hi; ++li) {
                s->h_leaf_addr[li] = (uint64_t) (uintptr_t) w.leaves[li];
                s->h_leaf_fp[li] = leaf_fingerprint(w.leaves[li], w.fences[li], K, schema->leaf_node_size);
            }
        });
    }
    PHASE("version rows");
    if (!rc) rc = stage_gpu_snapshot_seal(s);
    PHASE("seal");
#undef PHASE
    if (rc) {
        snapshot_free(s);
        return rc;
    }
    *out = s;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_from_tree(const void *root, const stage_schema_t *schema, stage_snapshot_t **out) {
    return from_tree_impl(root, schema, nullptr, nullptr, out);
}
extern "C" int stage_gpu_snapshot_from_tree_live(const void *root, const stage_schema_t *schema,
                                                 stage_undo_resolver_t resolve, void *resolve_ctx, stage_snapshot_t **out) {
    if (!resolve) return STAGE_ERR_ARG;
    return from_tree_impl(root, schema, resolve, resolve_ctx, out);
}

// ------------------------------------------------------------------------------------------------
// structural refresh: a new snapshot of the same host tree after committed inserts, updates,
// deletes and the leaf splits they caused. Unchanged leaves (same address, same fingerprint) are carried over on
// the device; only changed and new leaves cross PCIe and go through the transform kernel again.
// ------------------------------------------------------------------------------------------------
// In-place refresh: rows of the carried leaves move inside the payload buffer. Consecutive carried leaves whose rows
// shift by the same distance form a run; runs keep their order (a B-tree never reorders untouched leaves), so moving the
// left-shifting runs in ascending order and then the right-shifting runs in descending order never overwrites a row
// that has not been moved yet. A run whose source and destination overlap goes through a bounce buffer, chunk by chunk
// in the direction of the shift.
struct RowRun {
    uint64_t src, dst, n;      // rows
};
static int shift_payload_runs(uint8_t *payload, uint32_t P, const std::vector<RowRun> &runs, cudaStream_t st) {
    constexpr uint64_t kBounce = 64ull << 20;
    uint8_t *tmp = nullptr;
    auto move_run = [&](const RowRun &r) -> int {
        const uint64_t sb = r.src * P, db = r.dst * P, len = r.n * P;
        if (sb == db || len == 0) return STAGE_OK;
        const uint64_t dist = sb > db ? sb - db : db - sb;
        const bool left = db < sb;
        if (dist >= len) {                                   // disjoint: one copy
            CUDA_TRY(cudaMemcpyAsync(payload + db, payload + sb, len, cudaMemcpyDeviceToDevice, st));
            return STAGE_OK;
        }
        const bool direct = dist >= kBounce;                 // a chunk never reaches into its own source
        if (!direct && !tmp) CUDA_TRY(cudaMalloc((void **) &tmp, kBounce));
        const uint64_t n_chunks = (len + kBounce - 1) / kBounce;
        for (uint64_t k = 0; k < n_chunks; ++k) {
            const uint64_t c = left ? k : n_chunks - 1 - k;  // ascending for a left shift, descending for a right shift
            const uint64_t off = c * kBounce, m = std::min(kBounce, len - off);
            if (direct) {
                CUDA_TRY(cudaMemcpyAsync(payload + db + off, payload + sb + off, m, cudaMemcpyDeviceToDevice, st));
            } else {
                CUDA_TRY(cudaMemcpyAsync(tmp, payload + sb + off, m, cudaMemcpyDeviceToDevice, st));
                CUDA_TRY(cudaMemcpyAsync(payload + db + off, tmp, m, cudaMemcpyDeviceToDevice, st));
            }
        }
        return STAGE_OK;
    };
    int rc = STAGE_OK;
    for (size_t i = 0; i < runs.size() && !rc; ++i)
        if (runs[i].dst < runs[i].src) rc = move_run(runs[i]);
    for (size_t i = runs.size(); i-- > 0 && !rc;)
        if (runs[i].dst > runs[i].src) rc = move_run(runs[i]);
    if (tmp) {
        cudaStreamSynchronize(st);
        cudaFree(tmp);
    }
    return rc;
}

static int refresh_impl(stage_snapshot_t *old, const void *root, stage_undo_resolver_t resolve, void *resolve_ctx,
                        stage_snapshot_t **out, stage_refresh_stats_t *stats, bool in_place = false) {
    if (!old || !root || !out) return STAGE_ERR_ARG;
    if (!old->sealed || old->h_leaf_addr.size() != old->n_leaves || old->h_leaf_off.size() != old->n_leaves + 1)
        return STAGE_ERR_STATE;                       // only snapshots that came from a host tree carry fingerprints
    std::lock_guard<std::mutex> lock(old->mu);
    const stage_schema_t *schema = &old->schema;
    const uint32_t K = schema->key_size, P = schema->payload_size, ls = schema->leaf_node_size;
    if (schema_varlen(schema) && resolve) return STAGE_ERR_ARG;
    WalkOut w;
    w.key_form(schema, old->xormask);
    int rc;
    if ((rc = walk(static_cast<const uint8_t *>(root), K, w, 0))) return rc;
    const size_t L = w.leaves.size();
    if (L >= kNoLeaf) return STAGE_ERR_ARG;

    // 1. classify: carried over or re-transformed; row offsets of the new snapshot
    std::unordered_map<uint64_t, uint32_t> old_at;
    old_at.reserve(old->n_leaves * 2);
    for (uint64_t i = 0; i < old->n_leaves; ++i) old_at.emplace(old->h_leaf_addr[i], (uint32_t) i);
    std::vector<uint32_t> src_leaf(L, kNoLeaf), vis(L, 0), changed;
    std::vector<uint64_t> fp(L);
    {
        auto work = [&](size_t lo, size_t hi) {
            for (size_t j = lo; j < hi; ++j) {
                const uint8_t *leaf = w.leaves[j];
                fp[j] = leaf_fingerprint(leaf, w.fences[j], K, ls);
                auto it = old_at.find((uint64_t) (uintptr_t) leaf);
                if (it != old_at.end() && old->h_leaf_fp[it->second] == fp[j]) {
                    src_leaf[j] = it->second;
                    vis[j] = old->h_leaf_off[it->second + 1] - old->h_leaf_off[it->second];
                    continue;
                }
                uint64_t status;
                memcpy(&status, leaf + kOffStatus, 8);
                const uint32_t count = status_record_count(status);
                uint32_t v = 0;
                if ((uint64_t) kOffMeta + (uint64_t) count * kMetaSize <= ls)
                    for (uint32_t i = 0; i < count; ++i) {
                        uint64_t m;
                        memcpy(&m, leaf + kOffMeta + (size_t) i * kMetaSize, 8);
                        v += meta_visible(m) ? 1u : 0u;
                    }
                vis[j] = v;
            }
        };
        WorkPool::get().run(L, 256, [&](uint64_t lo, uint64_t hi) { work((size_t) lo, (size_t) hi); });
    }
    for (size_t j = 0; j < L; ++j)
        if (src_leaf[j] == kNoLeaf) changed.push_back((uint32_t) j);
    rc = STAGE_OK;
    // 1b. changed leaves: version chains (new indices follow the previous table) and in-flight records
    LeafDiscovery d;
    d.schema = schema;
    d.resolve = resolve;
    d.ctx = resolve_ctx;
    d.init(changed.size());
    ChainStager &cs = d.cs;
    cs.base = (uint32_t) old->n_versions;
    std::vector<uint32_t> &dropped = d.dropped;
    rc = discover_all(d, changed.size(), [&](size_t c) { return w.leaves[changed[c]]; });
    if (rc) return rc;
    uint64_t n_rows = 0;
    for (size_t c = 0; c < changed.size(); ++c) vis[changed[c]] -= dropped[c];
    for (size_t j = 0; j < L; ++j) n_rows += vis[j];

    stage_snapshot *s = nullptr;
    bool rows_moved = false, versions_adopted = false;
    if (in_place) {
        if (n_rows > old->cap_rows) {
            g_last_error = "in-place refresh: the snapshot has no room for the new rows (stage_gpu_set_row_slack before it is built)";
            return STAGE_ERR_NOMEM;
        }
        uint32_t prev = 0;
        bool any = false;
        for (size_t j = 0; j < L; ++j) {                    // carried leaves keep their order
            if (src_leaf[j] == kNoLeaf) continue;
            if (any && src_leaf[j] <= prev) return STAGE_ERR_STATE;
            prev = src_leaf[j];
            any = true;
        }
        CUDA_TRY(cudaStreamSynchronize(old->stream));
        rc = snapshot_begin_impl(schema, L, old->cap_rows, old->d_payload, &s);
    } else {
        rc = stage_gpu_snapshot_begin(schema, L, with_row_slack(n_rows), &s);
    }
    if (rc) return rc;
    uint32_t *d_src_leaf = nullptr, *d_changed = nullptr;
    auto fail = [&](int code) {
        cudaStreamSynchronize(s->stream);
        if (d_src_leaf) cudaFree(d_src_leaf);
        if (d_changed) cudaFree(d_changed);
        if (in_place) {
            s->d_payload = nullptr;                         // belongs to `old`
            if (versions_adopted) s->d_ver_begin = s->d_ver_commit = s->d_ver_next = nullptr, s->d_ver_payload = nullptr;
            if (rows_moved) old->sealed = false;            // its rows are no longer where its columns say: unusable
        }
        snapshot_free(s);
        return code;
    };
    s->h_leaf_off.resize(L + 1);
    s->h_sep.assign(L * s->nw, 0xFFFFFFFFu);
    s->h_has_sep.resize(L);
    s->h_leaf_addr.resize(L);
    s->h_leaf_fp = fp;
    for (size_t j = 0; j < L; ++j) {
        s->h_leaf_off[j + 1] = s->h_leaf_off[j] + vis[j];
        if (w.fences[j]) host_sortable_nw(s->nw, w.fences[j], K, s->xormask, s->h_sep.data() + j * s->nw);
        s->h_has_sep[j] = w.fences[j] ? 1 : 0;
        s->h_leaf_addr[j] = (uint64_t) (uintptr_t) w.leaves[j];
        s->max_leaf_rows = std::max(s->max_leaf_rows, vis[j]);
    }
    s->n_leaves = L;
    s->rows_staged = n_rows;
    cudaStream_t st = s->stream;
#define RF_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { g_last_error = cudaGetErrorString(e_); return fail(STAGE_ERR_CUDA); } } while (0)
    RF_TRY(cudaMemcpyAsync(s->d_leaf_off, s->h_leaf_off.data(), (L + 1) * 4, cudaMemcpyHostToDevice, st));
    RF_TRY(cudaMemcpyAsync(s->d_sep, s->h_sep.data(), L * s->nw * 4, cudaMemcpyHostToDevice, st));
    RF_TRY(cudaMemcpyAsync(s->d_has_sep, s->h_has_sep.data(), L, cudaMemcpyHostToDevice, st));

    // 2. carried-over leaves: old columns -> new columns, on the device
    RF_TRY(cudaMalloc((void **) &d_src_leaf, std::max<size_t>(L, 1) * 4));
    RF_TRY(cudaMemcpyAsync(d_src_leaf, src_leaf.data(), L * 4, cudaMemcpyHostToDevice, st));
    RF_TRY(cudaStreamSynchronize(old->stream));      // the previous snapshot's columns are final (refresh kernels done)
    if (L > changed.size()) {
        LeafCarryArgs c{};
        c.src_leaf = d_src_leaf; c.n_leaves = (uint32_t) L;
        c.old_off = old->d_leaf_off; c.new_off = s->d_leaf_off;
        c.old_skey = old->d_skey; c.new_skey = s->d_skey;
        c.old_slot = old->d_slot; c.new_slot = s->d_slot;
        c.old_ts = old->d_ts; c.new_ts = s->d_ts;
        c.old_chain = old->d_chain; c.new_chain = s->d_chain;
        c.old_payload = old->d_payload; c.new_payload = in_place ? nullptr : s->d_payload;
        c.old_cont = old->d_cont; c.new_cont = s->d_cont;
        c.nw = (uint32_t) s->nw; c.payload_size = P;
        leaf_carry_kernel<<<(unsigned) std::min<uint64_t>(L, (uint64_t) sm_count() * 8), 256, 0, st>>>(c);
        if (cudaGetLastError() != cudaSuccess) return fail(STAGE_ERR_CUDA);
        g_launches.fetch_add(1);
        if (in_place) {
            std::vector<RowRun> runs;
            for (size_t j = 0; j < L; ++j) {
                if (src_leaf[j] == kNoLeaf || vis[j] == 0) continue;
                const uint64_t src = old->h_leaf_off[src_leaf[j]], dst = s->h_leaf_off[j];
                if (!runs.empty() && runs.back().src + runs.back().n == src && runs.back().dst + runs.back().n == dst) runs.back().n += vis[j];
                else runs.push_back(RowRun{src, dst, vis[j]});
            }
            rows_moved = true;
            if ((rc = shift_payload_runs(s->d_payload, P, runs, st))) return fail(rc);
        }
    }

    // 3. changed leaves: images, transform
    if (in_place && !changed.empty()) rows_moved = true;        // the transform writes into the previous snapshot's buffer
    if (!rc && !changed.empty()) {
        if (cudaMalloc((void **) &d_changed, changed.size() * 4) != cudaSuccess) rc = STAGE_ERR_NOMEM;
        else if (cudaMemcpyAsync(d_changed, changed.data(), changed.size() * 4, cudaMemcpyHostToDevice, st) != cudaSuccess)
            rc = STAGE_ERR_CUDA;
    }
    std::vector<uint32_t> cvis, ccnt;
    std::vector<int> rcs;
    bool used[2] = {false, false};
    for (size_t c0 = 0; c0 < changed.size() && !rc; c0 += s->chunk_leaves) {
        const uint32_t m = (uint32_t) std::min<size_t>(s->chunk_leaves, changed.size() - c0);
        const int buf = s->stage_cur;
        if (used[buf] && cudaEventSynchronize(s->stage_ev[buf]) != cudaSuccess) { rc = STAGE_ERR_CUDA; break; }
        cvis.assign(m, 0); ccnt.assign(m, 0); rcs.assign(m, 0);
        auto work = [&](uint32_t lo, uint32_t hi) {
            for (uint32_t j = lo; j < hi; ++j) {
                const size_t c = c0 + j;
                LeafExtras ex;
                d.fill(c, &ex);
                rcs[j] = stage_copy_leaf(s, s->h_stage[buf] + (size_t) j * ls, w.leaves[changed[c]], ex, &cvis[j], &ccnt[j]);
            }
        };
        WorkPool::get().run(m, 4, [&](uint64_t lo, uint64_t hi) { work((uint32_t) lo, (uint32_t) hi); });
        uint32_t maxcount = 0;
        for (uint32_t j = 0; j < m && !rc; ++j) {
            for (int attempt = 0; rcs[j] == kStageRetry && !rc; ++attempt) {    // a writer touched the leaf under the copy
                const size_t c = c0 + j;
                if (attempt >= 64 || (rc = d.discover(c, w.leaves[changed[c]]))) {
                    if (!rc) rc = STAGE_ERR_NOT_QUIESCED;
                    break;
                }
                LeafExtras ex;
                d.fill(c, &ex);
                rcs[j] = stage_copy_leaf(s, s->h_stage[buf] + (size_t) j * ls, w.leaves[changed[c]], ex, &cvis[j], &ccnt[j]);
            }
            if (rc) break;
            if (rcs[j]) rc = rcs[j];
            else if (cvis[j] != vis[changed[c0 + j]]) rc = STAGE_ERR_NOT_QUIESCED;      // the tree moved under the walk
            maxcount = std::max(maxcount, ccnt[j]);
        }
        if (rc) break;
        if (cudaMemcpyAsync(s->d_stage[buf], s->h_stage[buf], (size_t) m * ls, cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = STAGE_ERR_CUDA; break; }
        switch (s->nw) {
            case 1: rc = launch_transform<1>(s, buf, m, 0, maxcount, d_changed + c0); break;
            case 2: rc = launch_transform<2>(s, buf, m, 0, maxcount, d_changed + c0); break;
            case 4: rc = launch_transform<4>(s, buf, m, 0, maxcount, d_changed + c0); break;
            default: rc = launch_transform<8>(s, buf, m, 0, maxcount, d_changed + c0); break;
        }
        if (rc) break;
        if (cudaEventRecord(s->stage_ev[buf], st) != cudaSuccess) { rc = STAGE_ERR_CUDA; break; }
        used[buf] = true;
        s->stage_cur ^= 1;
    }

    // 4. version table: the previous one (device copy) followed by the chains of the changed leaves
    const uint64_t V0 = old->n_versions, V1 = cs.vers.size(), V = V0 + V1;
    uint8_t *vbuf = nullptr;
    if (!rc && in_place && V1 == 0 && V0) {           // nothing to append: the previous version table as it is
        s->d_ver_begin = old->d_ver_begin; s->d_ver_commit = old->d_ver_commit; s->d_ver_next = old->d_ver_next;
        s->d_ver_payload = old->d_ver_payload;
        s->ver_cap = old->ver_cap;
        s->n_versions = V0;
        versions_adopted = true;
    } else if (!rc && V) {
        if ((rc = dev_alloc(s, &s->d_ver_begin, V)) || (rc = dev_alloc(s, &s->d_ver_commit, V)) ||
            (rc = dev_alloc(s, &s->d_ver_next, V)) || (rc = dev_alloc(s, &s->d_ver_payload, V * (uint64_t) P))) {
        } else {
            s->ver_cap = V;
            s->n_versions = V;
            cudaError_t e = cudaSuccess;
            if (V0) {
                e = cudaMemcpyAsync(s->d_ver_begin, old->d_ver_begin, V0 * 4, cudaMemcpyDeviceToDevice, st);
                if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_commit, old->d_ver_commit, V0 * 4, cudaMemcpyDeviceToDevice, st);
                if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_next, old->d_ver_next, V0 * 4, cudaMemcpyDeviceToDevice, st);
                if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_payload, old->d_ver_payload, V0 * (uint64_t) P, cudaMemcpyDeviceToDevice, st);
            }
            if (e == cudaSuccess && V1) {
                std::vector<uint32_t> b(V1), c(V1), nx(V1);
                for (uint64_t i = 0; i < V1; ++i) { b[i] = cs.vers[i].begin_id; c[i] = cs.vers[i].commit_id; nx[i] = cs.vers[i].next; }
                vbuf = cs.payload_rows(K, P);
                if (!vbuf) rc = STAGE_ERR_NOMEM;
                else {
                    e = cudaMemcpyAsync(s->d_ver_begin + V0, b.data(), V1 * 4, cudaMemcpyHostToDevice, st);
                    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_commit + V0, c.data(), V1 * 4, cudaMemcpyHostToDevice, st);
                    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_next + V0, nx.data(), V1 * 4, cudaMemcpyHostToDevice, st);
                    if (e == cudaSuccess) e = cudaMemcpyAsync(s->d_ver_payload + V0 * (uint64_t) P, vbuf, V1 * (uint64_t) P, cudaMemcpyHostToDevice, st);
                    if (e == cudaSuccess) e = cudaStreamSynchronize(st);     // b / c / nx / vbuf are pageable locals
                }
            }
            if (e != cudaSuccess) { g_last_error = cudaGetErrorString(e); rc = STAGE_ERR_CUDA; }
        }
    }
    free(vbuf);
    if (!rc) {
        s->n_rows = n_rows;
        rc = finish_seal(s);
    } else {
        cudaStreamSynchronize(st);
    }
    for (int i = 0; i < 2; ++i) {
        if (s->h_stage[i]) { cudaFreeHost(s->h_stage[i]); s->h_stage[i] = nullptr; }
        if (s->d_stage[i]) { cudaFree(s->d_stage[i]); s->d_stage[i] = nullptr; }
    }
#undef RF_TRY
    if (rc) return fail(rc);
    cudaFree(d_src_leaf);
    if (d_changed) cudaFree(d_changed);
    if (stats) {
        stats->leaves_total = L;
        stats->leaves_carried = L - changed.size();
        stats->leaves_transformed = changed.size();
        stats->rows = n_rows;
        stats->versions_added = V1;
        stats->host_bytes_staged = (uint64_t) changed.size() * ls + V1 * (uint64_t) P;
    }
    if (in_place) {                                     // the buffers now belong to the new snapshot
        old->d_payload = nullptr;
        if (versions_adopted) old->d_ver_begin = old->d_ver_commit = old->d_ver_next = nullptr, old->d_ver_payload = nullptr;
        old->sealed = false;
    }
    *out = s;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_refresh_in_place(stage_snapshot_t **snap, const void *root, stage_refresh_stats_t *stats) {
    if (!snap || !*snap) return STAGE_ERR_ARG;
    stage_snapshot *old = *snap, *next = nullptr;
    const int rc = refresh_impl(old, root, nullptr, nullptr, &next, stats, true);
    if (rc) return rc;
    stage_gpu_snapshot_destroy(old);
    *snap = next;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_refresh_from_tree(stage_snapshot_t *old, const void *root, stage_snapshot_t **out,
                                                    stage_refresh_stats_t *stats) {
    return refresh_impl(old, root, nullptr, nullptr, out, stats);
}
extern "C" int stage_gpu_snapshot_refresh_from_tree_live(stage_snapshot_t *old, const void *root,
                                                         stage_undo_resolver_t resolve, void *resolve_ctx,
                                                         stage_snapshot_t **out, stage_refresh_stats_t *stats) {
    if (!resolve) return STAGE_ERR_ARG;
    return refresh_impl(old, root, resolve, resolve_ctx, out, stats);
}

// ------------------------------------------------------------------------------------------------
// on-device YCSB table generation (consolidated leaves)
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_snapshot_generate_ycsb(const stage_schema_t *schema, uint64_t n_rows_total, uint64_t first_rank,
                                                uint64_t n_local, uint32_t rows_per_leaf, stage_snapshot_t **out) {
    if (!out || !schema || schema->key_size != 4 || rows_per_leaf == 0 || first_rank + n_local > n_rows_total ||
        n_rows_total > 0xFFFFFFFFull || n_local >= (uint64_t) kVersionRowFlag)
        return STAGE_ERR_ARG;
    stage_snapshot *s = nullptr;
    int rc = snapshot_new(schema, &s);
    if (rc) return rc;
    const uint64_t L = (n_local + rows_per_leaf - 1) / rows_per_leaf;
    if ((rc = alloc_columns(s, std::max<uint64_t>(L, 1), n_local))) { snapshot_free(s); return rc; }
    s->n_rows = n_local;
    s->n_leaves = std::max<uint64_t>(L, 1);
    s->max_leaf_rows = (uint32_t) std::min<uint64_t>(rows_per_leaf, n_local);
    GenArgs g{};
    g.n_total = n_rows_total; g.first_rank = first_rank; g.n_local = n_local; g.rows_per_leaf = rows_per_leaf;
    g.n_leaves = s->n_leaves; g.payload_size = schema->payload_size;
    g.skey = s->d_skey; g.slot = s->d_slot; g.ts = s->d_ts; g.chain = s->d_chain; g.payload = s->d_payload;
    g.leaf_off = s->d_leaf_off; g.sep = s->d_sep; g.cont = s->d_cont;
    auto fail = [&](int code) { snapshot_free(s); return code; };
    {
        const unsigned blocks = (unsigned) std::min<uint64_t>((n_local + 255) / 256 + 1, 148ull * 32);
        generate_ycsb_meta_kernel<<<blocks, 256, 0, s->stream>>>(g);
        g_launches++;
        if (cudaGetLastError() != cudaSuccess) return fail(STAGE_ERR_CUDA);
        generate_ycsb_payload_kernel<<<148 * 8, 256, 0, s->stream>>>(g);
        g_launches++;
        if (cudaGetLastError() != cudaSuccess) return fail(STAGE_ERR_CUDA);
    }
    if ((rc = finish_seal(s))) return fail(rc);
    *out = s;
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// bulk load from device-resident sorted columns
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_snapshot_alloc_sorted(const stage_schema_t *schema, uint64_t n_rows, uint32_t rows_per_leaf,
                                               stage_snapshot_t **out, void **raw_keys_dev, void **payload_dev,
                                               uint32_t **ts_dev) {
    if (!out || !schema || n_rows == 0 || rows_per_leaf == 0 || rows_per_leaf > 0xFFFFu) return STAGE_ERR_ARG;
    stage_snapshot *s = nullptr;
    int rc = snapshot_new(schema, &s);
    if (rc) return rc;
    const uint64_t L = (n_rows + rows_per_leaf - 1) / rows_per_leaf;
    if ((rc = alloc_columns(s, L, n_rows)) || (rc = dev_alloc(s, &s->d_raw_keys, n_rows * (uint64_t) schema->key_size))) {
        snapshot_free(s);
        return rc;
    }
    if (cudaMemsetAsync(s->d_ts, 0, n_rows * 4, s->stream) != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) {
        snapshot_free(s);
        return STAGE_ERR_CUDA;
    }
    s->n_rows = n_rows;
    s->n_leaves = L;
    s->max_leaf_rows = (uint32_t) std::min<uint64_t>(rows_per_leaf, n_rows);
    s->bulk_rows_per_leaf = rows_per_leaf;
    if (raw_keys_dev) *raw_keys_dev = s->d_raw_keys;
    if (payload_dev) *payload_dev = s->d_payload;
    if (ts_dev) *ts_dev = s->d_ts;
    *out = s;
    return STAGE_OK;
}

template <int NW>
static int launch_bulk(stage_snapshot *s) {
    const unsigned blocks = (unsigned) std::min<uint64_t>((s->n_rows + 255) / 256 + 1, 148ull * 32);
    bulk_keys_kernel<NW><<<blocks, 256, 0, s->stream>>>(s->d_raw_keys, s->n_rows, s->schema.key_size, s->xormask,
                                                        s->bulk_rows_per_leaf, s->d_skey, s->d_slot, s->d_chain);
    LAUNCH_CHECK();
    bulk_leaves_kernel<NW><<<blocks, 256, 0, s->stream>>>(s->d_skey, s->n_rows, s->n_leaves, s->bulk_rows_per_leaf,
                                                          s->d_leaf_off, s->d_sep, s->d_cont, s->d_err);
    LAUNCH_CHECK();
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_seal_sorted(stage_snapshot_t *s) {
    if (!s || s->sealed || !s->d_raw_keys) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaDeviceSynchronize());          // the caller filled the columns on its own streams
    int rc;
    switch (s->nw) {
        case 1: rc = launch_bulk<1>(s); break;
        case 2: rc = launch_bulk<2>(s); break;
        case 4: rc = launch_bulk<4>(s); break;
        default: rc = launch_bulk<8>(s); break;
    }
    if (rc) return rc;
    uint32_t err = 0;
    CUDA_TRY(cudaMemcpyAsync(&err, s->d_err, 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    if (err == 3) {
        g_last_error = "bulk load: keys are not strictly ascending in reference key order";
        return STAGE_ERR_ARG;
    }
    const size_t raw_bytes = ((s->n_rows * (uint64_t) s->schema.key_size + 255) & ~size_t(255)) + 256;
    cudaFree(s->d_raw_keys);
    s->d_raw_keys = nullptr;
    s->device_bytes -= std::min(s->device_bytes, raw_bytes);
    return finish_seal(s);
}

// ------------------------------------------------------------------------------------------------
// incremental refresh: committed updates (refresh.cuh)
// ------------------------------------------------------------------------------------------------
static int grow_versions(stage_snapshot *s, uint64_t need) {
    if (need <= s->ver_cap) return STAGE_OK;
    const uint64_t cap = std::max<uint64_t>(need, s->ver_cap * 2);
    const uint64_t V = s->n_versions, P = s->schema.payload_size;
    uint32_t *nb = nullptr, *nc = nullptr, *nn = nullptr;
    uint8_t *np_ = nullptr;
    int rc;
    if ((rc = dev_alloc(s, &nb, cap)) || (rc = dev_alloc(s, &nc, cap)) || (rc = dev_alloc(s, &nn, cap)) ||
        (rc = dev_alloc(s, &np_, cap * P)))
        return rc;
    if (V) {
        CUDA_TRY(cudaMemcpyAsync(nb, s->d_ver_begin, V * 4, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(nc, s->d_ver_commit, V * 4, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(nn, s->d_ver_next, V * 4, cudaMemcpyDeviceToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(np_, s->d_ver_payload, V * P, cudaMemcpyDeviceToDevice, s->stream));
    }
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    void *old[] = {s->d_ver_begin, s->d_ver_commit, s->d_ver_next, s->d_ver_payload};
    for (void *p : old)
        if (p) cudaFree(p);
    s->d_ver_begin = nb; s->d_ver_commit = nc; s->d_ver_next = nn; s->d_ver_payload = np_;
    s->ver_cap = cap;
    return STAGE_OK;
}

// Version-table compaction: drops the versions no chain reaches any more (the copies left behind by refreshes that
// re-transformed their leaves) and renumbers the rest. Exclusive like apply_updates: no read in flight.
extern "C" int stage_gpu_snapshot_compact_versions(stage_snapshot_t *s, uint64_t *kept_out, uint64_t *dropped_out) {
    if (!s) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    const uint64_t V = s->n_versions, P = s->schema.payload_size;
    if (kept_out) *kept_out = V;
    if (dropped_out) *dropped_out = 0;
    if (V == 0) return STAGE_OK;
    cudaStream_t st = s->stream;
    uint32_t *live = nullptr, *idx = nullptr, *nb = nullptr, *nc = nullptr, *nn = nullptr;
    uint8_t *np_ = nullptr;
    void *tmp = nullptr;
    auto cleanup = [&](int code) {
        cudaStreamSynchronize(st);
        for (void *p : {(void *) live, (void *) idx, (void *) nb, (void *) nc, (void *) nn, (void *) np_, tmp})
            if (p) cudaFree(p);
        return code;
    };
#define CV_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { g_last_error = cudaGetErrorString(e_); return cleanup(STAGE_ERR_CUDA); } } while (0)
    CV_TRY(cudaMalloc((void **) &live, (V + 1) * 4));
    CV_TRY(cudaMalloc((void **) &idx, (V + 1) * 4));
    CV_TRY(cudaMemsetAsync(live, 0, (V + 1) * 4, st));
    const unsigned rb = (unsigned) std::min<uint64_t>((s->n_rows + 255) / 256 + 1, (uint64_t) sm_count() * 16);
    versions_mark_kernel<<<rb, 256, 0, st>>>(s->d_chain, s->n_rows, s->d_ver_next, V, live);
    CV_TRY(cudaGetLastError());
    size_t tmp_bytes = 0;
    CV_TRY(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, live, idx, V + 1, st));
    CV_TRY(cudaMalloc(&tmp, tmp_bytes));
    CV_TRY(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, live, idx, V + 1, st));      // idx[V] = number of live versions
    uint32_t kept32 = 0;
    CV_TRY(cudaMemcpyAsync(&kept32, idx + V, 4, cudaMemcpyDeviceToHost, st));
    CV_TRY(cudaStreamSynchronize(st));
    const uint64_t kept = kept32;
    if (kept == V) return cleanup(STAGE_OK);                 // nothing to drop
    const uint64_t cap = std::max<uint64_t>(kept, 1);
    CV_TRY(cudaMalloc((void **) &nb, cap * 4));
    CV_TRY(cudaMalloc((void **) &nc, cap * 4));
    CV_TRY(cudaMalloc((void **) &nn, cap * 4));
    CV_TRY(cudaMalloc((void **) &np_, cap * P + 256));
    const unsigned vb = (unsigned) std::min<uint64_t>((V + 7) / 8, (uint64_t) sm_count() * 16);
    versions_move_kernel<<<vb, 256, 0, st>>>(live, idx, V, s->d_ver_begin, s->d_ver_commit, s->d_ver_next, s->d_ver_payload, nb, nc, nn,
                                             np_, (uint32_t) P);
    CV_TRY(cudaGetLastError());
    versions_remap_chain_kernel<<<rb, 256, 0, st>>>(s->d_chain, s->n_rows, live, idx, V);
    CV_TRY(cudaGetLastError());
    CV_TRY(cudaStreamSynchronize(st));
#undef CV_TRY
    g_launches.fetch_add(3);
    void *old[] = {s->d_ver_begin, s->d_ver_commit, s->d_ver_next, s->d_ver_payload};
    for (void *p : old)
        if (p) cudaFree(p);
    s->d_ver_begin = nb; s->d_ver_commit = nc; s->d_ver_next = nn; s->d_ver_payload = np_;
    nb = nc = nn = nullptr;
    np_ = nullptr;
    s->n_versions = kept;
    s->ver_cap = cap;
    fill_view(s);
    if (kept_out) *kept_out = kept;
    if (dropped_out) *dropped_out = V - kept;
    return cleanup(STAGE_OK);
}

extern "C" int stage_gpu_snapshot_apply_updates(stage_snapshot_t *s, const void *keys, const void *patches,
                                                uint32_t patch_off, uint32_t patch_len, const uint32_t *commit_ids,
                                                uint64_t n, int flags, uint64_t *applied_out) {
    if (!s || (n && (!keys || !patches || !commit_ids))) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if (patch_len == 0 || (uint64_t) patch_off + patch_len > P || n > 0x7FFFFFFFull ||
        s->n_versions + n >= (uint64_t) kVersionRowFlag)
        return STAGE_ERR_ARG;
    if (applied_out) *applied_out = 0;
    if (n == 0) return STAGE_OK;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    int rc;
    rc = grow_versions(s, s->n_versions + n);
    fill_view(s);                 // the version arrays may have moved: never leave the kernels' view dangling
    if (rc) return rc;
    RefreshArgs a{};
    if (flags & STAGE_IO_DEVICE) {
        a.keys = static_cast<const uint8_t *>(keys);
        a.patches = static_cast<const uint8_t *>(patches);
        a.commit_ids = commit_ids;
    } else {
        if ((rc = s->rf_keys.ensure(n * K)) || (rc = s->rf_patch.ensure(n * (uint64_t) patch_len)) || (rc = s->rf_cid.ensure(n * 4)))
            return rc;
        CUDA_TRY(cudaMemcpyAsync(s->rf_keys.p, keys, n * K, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->rf_patch.p, patches, n * (uint64_t) patch_len, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(s->rf_cid.p, commit_ids, n * 4, cudaMemcpyHostToDevice, s->stream));
        a.keys = static_cast<const uint8_t *>(s->rf_keys.p);
        a.patches = static_cast<const uint8_t *>(s->rf_patch.p);
        a.commit_ids = static_cast<const uint32_t *>(s->rf_cid.p);
    }
    if ((rc = s->rf_row.ensure(n * 4)) || (rc = s->rf_order.ensure(n * 4)) || (rc = s->rf_row_sorted.ensure(n * 4)) ||
        (rc = s->rf_order_sorted.ensure(n * 4)) || (rc = s->rf_applied.ensure(4)))
        return rc;
    a.n = n;
    a.patch_off = patch_off;
    a.patch_len = patch_len;
    a.v0 = (uint32_t) s->n_versions;
    a.ts = s->d_ts; a.chain = s->d_chain; a.payload = s->d_payload;
    a.ver_begin = s->d_ver_begin; a.ver_commit = s->d_ver_commit; a.ver_next = s->d_ver_next; a.ver_payload = s->d_ver_payload;
    a.payload_size = P;
    a.row = static_cast<uint32_t *>(s->rf_row.p);
    a.order = static_cast<uint32_t *>(s->rf_order.p);
    a.applied = static_cast<uint32_t *>(s->rf_applied.p);
    CUDA_TRY(cudaMemsetAsync(a.applied, 0, 4, s->stream));
    {
        const size_t smem = (size_t) std::max<uint32_t>(s->view.smem_words, 4) * 4;
        const unsigned blocks = (unsigned) std::min<uint64_t>((n + kProbeThreads - 1) / kProbeThreads, (uint64_t) sm_count() * 8);
        switch (s->nw) {
            case 1: refresh_locate_kernel<1><<<blocks, kProbeThreads, smem, s->stream>>>(s->view, a); break;
            case 2: refresh_locate_kernel<2><<<blocks, kProbeThreads, smem, s->stream>>>(s->view, a); break;
            case 4: refresh_locate_kernel<4><<<blocks, kProbeThreads, smem, s->stream>>>(s->view, a); break;
            default: refresh_locate_kernel<8><<<blocks, kProbeThreads, smem, s->stream>>>(s->view, a); break;
        }
        LAUNCH_CHECK();
    }
    // stable sort of (row, update index) by row: runs of equal rows keep the batch order
    size_t tmp_bytes = 0;
    uint32_t *row_sorted = static_cast<uint32_t *>(s->rf_row_sorted.p), *order_sorted = static_cast<uint32_t *>(s->rf_order_sorted.p);
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, a.row, row_sorted, a.order, order_sorted, (int) n, 0, 32, s->stream));
    if ((rc = s->rf_tmp.ensure(tmp_bytes))) return rc;
    CUDA_TRY(cub::DeviceRadixSort::SortPairs(s->rf_tmp.p, tmp_bytes, a.row, row_sorted, a.order, order_sorted, (int) n, 0, 32, s->stream));
    g_launches.fetch_add(1, std::memory_order_relaxed);
    a.row_sorted = row_sorted;
    a.order = order_sorted;
    refresh_apply_kernel<<<(unsigned) ((n * 32 + 255) / 256), 256, 0, s->stream>>>(a);
    LAUNCH_CHECK();
    refresh_clear_unused_kernel<<<(unsigned) ((n + 255) / 256), 256, 0, s->stream>>>(a);
    LAUNCH_CHECK();
    uint32_t applied = 0;
    CUDA_TRY(cudaMemcpyAsync(&applied, a.applied, 4, cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    s->n_versions += n;
    fill_view(s);
    if (applied_out) *applied_out = applied;
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// snapshot serialisation: the flattened arrays, streamed through a pinned buffer
// ------------------------------------------------------------------------------------------------
struct SnapFileHeader {
    char magic[8];                  // "STGSNAP1"
    stage_schema_t schema;
    uint32_t nw, xormask, max_leaf_rows, reserved;
    uint64_t n_rows, n_leaves, n_versions;
};

static int stream_io(stage_snapshot *s, FILE *f, void *dev, size_t bytes, bool save, uint8_t *pinned, size_t chunk) {
    uint8_t *d = static_cast<uint8_t *>(dev);
    for (size_t off = 0; off < bytes; off += chunk) {
        const size_t m = std::min(chunk, bytes - off);
        if (save) {
            CUDA_TRY(cudaMemcpyAsync(pinned, d + off, m, cudaMemcpyDeviceToHost, s->stream));
            CUDA_TRY(cudaStreamSynchronize(s->stream));
            if (fwrite(pinned, 1, m, f) != m) return STAGE_ERR_ARG;
        } else {
            if (fread(pinned, 1, m, f) != m) return STAGE_ERR_LAYOUT;
            CUDA_TRY(cudaMemcpyAsync(d + off, pinned, m, cudaMemcpyHostToDevice, s->stream));
            CUDA_TRY(cudaStreamSynchronize(s->stream));
        }
    }
    return STAGE_OK;
}

static int snapshot_arrays_io(stage_snapshot *s, FILE *f, bool save) {
    const size_t chunk = 64u << 20;
    uint8_t *pinned = nullptr;
    CUDA_TRY(cudaHostAlloc((void **) &pinned, chunk, cudaHostAllocDefault));
    const uint64_t R = s->n_rows, L = s->n_leaves, V = s->n_versions, P = s->schema.payload_size, NW = s->nw;
    struct { void *p; size_t bytes; } arr[] = {
        {s->d_skey, R * NW * 4}, {s->d_slot, R * 2}, {s->d_ts, R * 4}, {s->d_chain, R * 4}, {s->d_payload, R * P},
        {s->d_leaf_off, (L + 1) * 4}, {s->d_sep, L * NW * 4}, {s->d_cont, L},
        {s->d_ver_begin, V * 4}, {s->d_ver_commit, V * 4}, {s->d_ver_next, V * 4}, {s->d_ver_payload, V * P}};
    int rc = STAGE_OK;
    for (auto &a : arr)
        if (a.bytes && (rc = stream_io(s, f, a.p, a.bytes, save, pinned, chunk))) break;
    cudaFreeHost(pinned);
    return rc;
}

extern "C" int stage_gpu_snapshot_save(stage_snapshot_t *s, const char *path) {
    if (!s || !path) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    FILE *f = fopen(path, "wb");
    if (!f) { g_last_error = std::string("cannot open ") + path; return STAGE_ERR_ARG; }
    SnapFileHeader h{};
    memcpy(h.magic, "STGSNAP1", 8);
    h.schema = s->schema; h.nw = (uint32_t) s->nw; h.xormask = s->xormask; h.max_leaf_rows = s->max_leaf_rows;
    h.n_rows = s->n_rows; h.n_leaves = s->n_leaves; h.n_versions = s->n_versions;
    int rc = fwrite(&h, sizeof(h), 1, f) == 1 ? STAGE_OK : STAGE_ERR_ARG;
    if (!rc) rc = snapshot_arrays_io(s, f, true);
    if (fclose(f) != 0 && !rc) rc = STAGE_ERR_ARG;
    return rc;
}

extern "C" int stage_gpu_snapshot_load(const char *path, stage_snapshot_t **out) {
    if (!path || !out) return STAGE_ERR_ARG;
    FILE *f = fopen(path, "rb");
    if (!f) { g_last_error = std::string("cannot open ") + path; return STAGE_ERR_ARG; }
    SnapFileHeader h{};
    int nw;
    uint32_t xm;
    if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, "STGSNAP1", 8) != 0 || schema_check(&h.schema, &nw, &xm) ||
        (uint32_t) nw != h.nw || xm != h.xormask || h.n_leaves == 0) {
        fclose(f);
        return STAGE_ERR_LAYOUT;
    }
    stage_snapshot *s = nullptr;
    int rc = snapshot_new(&h.schema, &s);
    if (rc) { fclose(f); return rc; }
    auto fail = [&](int code) { fclose(f); snapshot_free(s); return code; };
    if ((rc = alloc_columns(s, h.n_leaves, h.n_rows))) return fail(rc);
    s->n_rows = h.n_rows; s->n_leaves = h.n_leaves; s->n_versions = h.n_versions; s->max_leaf_rows = h.max_leaf_rows;
    if (h.n_versions) {
        const uint64_t V = h.n_versions;
        if ((rc = dev_alloc(s, &s->d_ver_begin, V)) || (rc = dev_alloc(s, &s->d_ver_commit, V)) ||
            (rc = dev_alloc(s, &s->d_ver_next, V)) || (rc = dev_alloc(s, &s->d_ver_payload, V * (uint64_t) h.schema.payload_size)))
            return fail(rc);
        s->ver_cap = V;
    }
    if ((rc = snapshot_arrays_io(s, f, false))) return fail(rc);
    if ((rc = finish_seal(s))) return fail(rc);
    fclose(f);
    *out = s;
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// info / lifecycle
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_init(int device) {
    int n = 0;
    CUDA_TRY(cudaGetDeviceCount(&n));
    if (device < 0 || device >= n) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(device));
    g_device = device;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_info(const stage_snapshot_t *s, stage_snapshot_info_t *info) {
    if (!s || !info) return STAGE_ERR_ARG;
    info->n_rows = s->n_rows;
    info->n_leaves = s->n_leaves;
    info->n_versions = s->n_versions;
    info->device_bytes = s->device_bytes;
    info->key_words = (uint32_t) s->nw;
    info->n_levels = (uint32_t) s->view.n_levels;
    info->max_leaf_rows = s->max_leaf_rows;
    info->device = (uint32_t) s->device;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_schema(const stage_snapshot_t *s, stage_schema_t *schema) {
    if (!s || !schema) return STAGE_ERR_ARG;
    *schema = s->schema;
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_destroy(stage_snapshot_t *s) {
    if (!s) return STAGE_ERR_ARG;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    snapshot_free(s);
    return STAGE_OK;
}

extern "C" int stage_gpu_snapshot_set_stream(stage_snapshot_t *s, void *cuda_stream) {
    if (!s) return STAGE_ERR_ARG;
    if (s->own_stream && s->stream) {
        cudaStreamSynchronize(s->stream);
        cudaStreamDestroy(s->stream);
    }
    s->stream = static_cast<cudaStream_t>(cuda_stream);
    s->own_stream = false;
    return STAGE_OK;
}
extern "C" int stage_gpu_snapshot_enable_timing(stage_snapshot_t *s, int on) {
    if (!s) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(s->device));
    if (on && !s->tev[0])
        for (int i = 0; i < 3; ++i) CUDA_TRY(cudaEventCreate(&s->tev[i]));
    s->timing = on != 0;
    return STAGE_OK;
}
extern "C" int stage_gpu_snapshot_last_timing(stage_snapshot_t *s, float *plan_ms, float *gather_ms) {
    if (!s || !s->timing || !s->tev[0]) return STAGE_ERR_STATE;
    CUDA_TRY(cudaEventSynchronize(s->tev[2]));
    float p = 0.f, g = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&p, s->tev[0], s->tev[1]));
    CUDA_TRY(cudaEventElapsedTime(&g, s->tev[1], s->tev[2]));
    if (plan_ms) *plan_ms = p;
    if (gather_ms) *gather_ms = s->timed_gather ? g : 0.f;
    return STAGE_OK;
}
extern "C" void *stage_gpu_snapshot_get_stream(const stage_snapshot_t *s) { return s ? (void *) s->stream : nullptr; }

// ------------------------------------------------------------------------------------------------
// reads
// ------------------------------------------------------------------------------------------------

template <typename V>
static int launch_gather_v(cudaStream_t st, const GatherArgs &g) {
    const unsigned blocks = (unsigned) ((g.n_slots + kGatherRows - 1) / kGatherRows);
    gather_rows_kernel<V, 4><<<blocks, kGatherThreads, 0, st>>>(g);
    LAUNCH_CHECK();
    return STAGE_OK;
}

// payload rows named by src_row[0..n_slots) -> dst[n_slots][P]. dst may be device memory or the device alias of
// pinned host memory (the gather then streams the rows straight across PCIe).
static int launch_gather(stage_snapshot *s, cudaStream_t st, const uint32_t *src_row, uint64_t n_slots, uint8_t *dst,
                         const uint32_t *dst_base = nullptr, uint32_t out_stride = 1) {
    if (n_slots == 0 || !dst) return STAGE_OK;
    if ((n_slots + kGatherRows - 1) / kGatherRows > 0x7FFFFFFFull) return STAGE_ERR_ARG;
    GatherArgs g{};
    g.src_row = src_row;
    g.n_slots = n_slots;
    g.payload = s->d_payload;
    g.ver_payload = s->d_ver_payload;
    g.dst = dst;
    g.payload_size = s->schema.payload_size;
    g.dst_base = dst_base;
    g.out_stride = std::max<uint32_t>(out_stride, 1);
    const uintptr_t align = (uintptr_t) g.payload | (uintptr_t) g.ver_payload | (uintptr_t) dst | g.payload_size;
    if (g.payload_size <= 256 && (align & 3) == 0) {          // short rows: one thread per 32-bit word
        const uint32_t wpr = g.payload_size / 4;
        const uint64_t words = n_slots * wpr;
        const uint64_t per_cta = 256 * kSmallWordsPerThread;
        const uint64_t ctas = (words + per_cta - 1) / per_cta;
        if (ctas > 0x7FFFFFFFull) return STAGE_ERR_ARG;
        if (words + per_cta < 0xFFFFFFFFull) gather_rows_small_kernel<uint32_t><<<(unsigned) ctas, 256, 0, st>>>(g, wpr, words);
        else gather_rows_small_kernel<uint64_t><<<(unsigned) ctas, 256, 0, st>>>(g, wpr, words);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    if (g.payload_size <= kMidRowMax && (align & 15) == 0) {  // mid-size rows (TPC-C stock: 400 B): flat 16-byte vector stream
        static const int vecs = getenv("STAGE_MID_VECS") ? atoi(getenv("STAGE_MID_VECS")) : 4;
        const unsigned blocks = (unsigned) ((n_slots + kGatherRows - 1) / kGatherRows);
        if (vecs >= 7) gather_rows_mid_kernel<7><<<blocks, kGatherThreads, 0, st>>>(g);
        else if (vecs <= 2) gather_rows_mid_kernel<2><<<blocks, kGatherThreads, 0, st>>>(g);
        else gather_rows_mid_kernel<4><<<blocks, kGatherThreads, 0, st>>>(g);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    if ((align & 15) == 0) return launch_gather_v<uint4>(st, g);
    if ((align & 7) == 0) return launch_gather_v<uint2>(st, g);
    if ((align & 3) == 0) return launch_gather_v<uint32_t>(st, g);
    return launch_gather_v<uint8_t>(st, g);
}

template <int NW>
static int launch_probe(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a) {
    static const int ppl = getenv("STAGE_PROBE_PPL") ? atoi(getenv("STAGE_PROBE_PPL")) : 4;
    const uint64_t want_blocks = (a.n + kProbeThreads * ppl - 1) / (kProbeThreads * ppl);
    if (s->view.dir) {
        // directory lookup: no shared-memory top level
        static int occd[9] = {0};
        if (!occd[NW]) {
            int o = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, probe_kernel<NW, 4, true>, kProbeThreads, 0));
            occd[NW] = std::max(o, 1);
        }
        const unsigned blocks = (unsigned) std::min<uint64_t>(want_blocks, (uint64_t) sm_count() * occd[NW]);
        if (ppl == 8) probe_kernel<NW, 8, true><<<blocks, kProbeThreads, 0, st>>>(s->view, a);
        else if (ppl == 4) probe_kernel<NW, 4, true><<<blocks, kProbeThreads, 0, st>>>(s->view, a);
        else probe_kernel<NW, 2, true><<<blocks, kProbeThreads, 0, st>>>(s->view, a);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    const size_t smem = (size_t) std::max<uint32_t>(s->view.smem_words, 4) * 4;
    // persistent grid: the CTAs that are resident at once; threads stride over the probes
    static int occ[9] = {0};
    if (!occ[NW]) {
        int o = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, probe_kernel<NW, 4>, kProbeThreads, kProbeTopBytes));
        occ[NW] = std::max(o, 1);
    }
    const unsigned blocks = (unsigned) std::min<uint64_t>(want_blocks, (uint64_t) sm_count() * occ[NW]);
    if (ppl == 8) probe_kernel<NW, 8><<<blocks, kProbeThreads, smem, st>>>(s->view, a);
    else if (ppl == 4) probe_kernel<NW, 4><<<blocks, kProbeThreads, smem, st>>>(s->view, a);
    else probe_kernel<NW, 2><<<blocks, kProbeThreads, smem, st>>>(s->view, a);
    LAUNCH_CHECK();
    return STAGE_OK;
}
static int launch_probe_nw(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a) {
    switch (s->nw) {
        case 1: return launch_probe<1>(s, st, a);
        case 2: return launch_probe<2>(s, st, a);
        case 4: return launch_probe<4>(s, st, a);
        default: return launch_probe<8>(s, st, a);
    }
}

template <int NW, typename V>
static int launch_probe_fused_v(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a, uint8_t *payload_out) {
    const size_t smem = (size_t) std::max<uint32_t>(s->view.smem_words, 4) * 4;
    static int occ = 0;
    if (!occ) {
        int o = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, probe_fused_kernel<NW, V>, kFusedThreads, kProbeTopBytes));
        occ = std::max(o, 1);
    }
    const uint64_t tiles = (a.n + kFusedTile - 1) / kFusedTile;
    const unsigned blocks = (unsigned) std::min<uint64_t>(tiles, (uint64_t) sm_count() * occ);
    probe_fused_kernel<NW, V><<<blocks, kFusedThreads, smem, st>>>(s->view, a, payload_out);
    LAUNCH_CHECK();
    return STAGE_OK;
}
template <int NW>
static int launch_probe_fused(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a, uint8_t *payload_out) {
    const uintptr_t align = (uintptr_t) s->d_payload | (uintptr_t) s->d_ver_payload | (uintptr_t) payload_out |
                            s->schema.payload_size;
    if ((align & 15) == 0) return launch_probe_fused_v<NW, uint4>(s, st, a, payload_out);
    if ((align & 7) == 0) return launch_probe_fused_v<NW, uint2>(s, st, a, payload_out);
    if ((align & 3) == 0) return launch_probe_fused_v<NW, uint32_t>(s, st, a, payload_out);
    return launch_probe_fused_v<NW, uint8_t>(s, st, a, payload_out);
}

constexpr bool kProbeRowsDefault = false;            // set by measurement, see profiles/README.md
constexpr uint32_t kProbeRowsDefaultMaxRow = 512;
// lookup + rows in one kernel, the warp that found 32 rows copies them (probe_rows.cuh). STAGE_PROBE_ROWS=1 forces it,
// =0 forbids it; default: see probe_device.
template <int NW, typename V, int U>
static int launch_probe_rows_v(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a, uint8_t *payload_out) {
    static int occ = 0;
    if (!occ) {
        int o = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, probe_rows_kernel<NW, V, U>, kProbeRowsThreads, 0));
        occ = std::max(o, 1);
    }
    const uint64_t want = (a.n + kProbeRowsThreads - 1) / kProbeRowsThreads;
    const unsigned blocks = (unsigned) std::min<uint64_t>(want, (uint64_t) sm_count() * occ);
    probe_rows_kernel<NW, V, U><<<blocks, kProbeRowsThreads, 0, st>>>(s->view, a, payload_out);
    LAUNCH_CHECK();
    return STAGE_OK;
}
template <int NW>
static int launch_probe_rows(stage_snapshot *s, cudaStream_t st, const ProbeArgs &a, uint8_t *payload_out) {
    const uintptr_t align = (uintptr_t) s->d_payload | (uintptr_t) s->d_ver_payload | (uintptr_t) payload_out | s->schema.payload_size;
    if ((align & 15) == 0) return launch_probe_rows_v<NW, uint4, 5>(s, st, a, payload_out);
    return launch_probe_rows_v<NW, uint2, 8>(s, st, a, payload_out);
}
static bool probe_rows_eligible(const stage_snapshot *s, const uint8_t *payload_out) {
    const uintptr_t align = (uintptr_t) s->d_payload | (uintptr_t) s->d_ver_payload | (uintptr_t) payload_out | s->schema.payload_size;
    return s->view.dir != nullptr && payload_out != nullptr && (align & 7) == 0 && s->schema.payload_size >= 64;
}

// search + MVCC + payload rows. Default: probe_kernel then gather_rows_kernel (0.398 ms per 2^20 YCSB probes);
// STAGE_PROBE_FUSED=1 selects the warp-specialised single kernel of probe_fused.cuh (measured 0.437 ms: under a
// saturated memory system the producers' dependent loads are too slow to keep the consumers fed — profiles/README.md).
// All pointers are device-visible (device memory, or device aliases of pinned host memory for the outputs).
static int probe_device(stage_snapshot *s, const Lane &ln, ProbeArgs a, uint8_t *payload_out) {
    if (a.n == 0) return STAGE_OK;
    int rc;
    static const bool fused = getenv("STAGE_PROBE_FUSED") && atoi(getenv("STAGE_PROBE_FUSED")) != 0;
    if (fused && payload_out) {
        if (ln.timing) {
            CUDA_TRY(cudaEventRecord(s->tev[0], ln.st));
            CUDA_TRY(cudaEventRecord(s->tev[1], ln.st));
        }
        switch (s->nw) {
            case 1: rc = launch_probe_fused<1>(s, ln.st, a, payload_out); break;
            case 2: rc = launch_probe_fused<2>(s, ln.st, a, payload_out); break;
            case 4: rc = launch_probe_fused<4>(s, ln.st, a, payload_out); break;
            default: rc = launch_probe_fused<8>(s, ln.st, a, payload_out); break;
        }
        if (ln.timing) {
            CUDA_TRY(cudaEventRecord(s->tev[2], ln.st));
            s->timed_gather = true;
        }
        return rc;
    }
    {
        const char *e = getenv("STAGE_PROBE_ROWS");
        const int mode = e ? atoi(e) : -1;                                   // -1 = default rule
        const bool use = mode == 1 || (mode == -1 && kProbeRowsDefault && s->schema.payload_size <= kProbeRowsDefaultMaxRow);
        if (use && probe_rows_eligible(s, payload_out)) {
            if (ln.timing) {
                CUDA_TRY(cudaEventRecord(s->tev[0], ln.st));
                CUDA_TRY(cudaEventRecord(s->tev[1], ln.st));
            }
            switch (s->nw) {
                case 1: rc = launch_probe_rows<1>(s, ln.st, a, payload_out); break;
                case 2: rc = launch_probe_rows<2>(s, ln.st, a, payload_out); break;
                case 4: rc = launch_probe_rows<4>(s, ln.st, a, payload_out); break;
                default: rc = launch_probe_rows<8>(s, ln.st, a, payload_out); break;
            }
            if (ln.timing) {
                CUDA_TRY(cudaEventRecord(s->tev[2], ln.st));
                s->timed_gather = true;
            }
            return rc;
        }
    }
    if (payload_out) {
        if ((rc = ln.srcrow->ensure(a.n * 4))) return rc;
        a.src_row = static_cast<uint32_t *>(ln.srcrow->p);
    }
    if (ln.timing) CUDA_TRY(cudaEventRecord(s->tev[0], ln.st));
    if ((rc = launch_probe_nw(s, ln.st, a))) return rc;
    if (ln.timing) CUDA_TRY(cudaEventRecord(s->tev[1], ln.st));
    rc = launch_gather(s, ln.st, a.src_row, a.n, payload_out);
    if (ln.timing) {
        CUDA_TRY(cudaEventRecord(s->tev[2], ln.st));
        s->timed_gather = payload_out != nullptr;
    }
    return rc;
}

// ---- host-buffer calls: contexts, device aliases of pinned memory ---------------------------------------------
static HostCtx *ctx_acquire(stage_snapshot *s, int *rc_out) {
    constexpr size_t kMaxCtx = 16;
    std::unique_lock<std::mutex> lk(s->ctx_mu);
    for (;;) {
        for (HostCtx *c : s->ctx_pool)
            if (!c->busy) {
                c->busy = true;
                return c;
            }
        if (s->ctx_pool.size() < kMaxCtx) break;
        s->ctx_cv.wait(lk);
    }
    auto *c = new HostCtx();
    cudaError_t e = cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&c->cp, cudaStreamNonBlocking);
    for (int i = 0; i < 2 && e == cudaSuccess; ++i) {
        e = cudaEventCreateWithFlags(&c->done[i], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->drained[i], cudaEventDisableTiming);
    }
    if (e != cudaSuccess) {
        g_last_error = std::string("host-call context: ") + cudaGetErrorString(e);
        c->release_all();
        delete c;
        *rc_out = STAGE_ERR_CUDA;
        return nullptr;
    }
    c->busy = true;
    s->ctx_pool.push_back(c);
    return c;
}
struct CtxGuard {
    stage_snapshot *s;
    HostCtx *c;
    ~CtxGuard() {
        if (!c) return;
        {
            std::lock_guard<std::mutex> lk(s->ctx_mu);
            c->busy = false;
        }
        s->ctx_cv.notify_one();
    }
};

// Device alias of a host pointer the kernels may store to directly (cudaHostAlloc / cudaHostRegister memory under
// unified addressing), or nullptr for pageable memory.
static void *mapped_alias(const void *p) {
    if (!p) return nullptr;
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return at.type == cudaMemoryTypeHost ? at.devicePointer : nullptr;
}

// How STAGE_IO_HOST results cross PCIe. Default ("staged"): results are produced into double-buffered HBM scratch and
// leave through the DMA engine as one large copy per output array and chunk, while the next chunk is planned and
// gathered; works with any host memory, fastest with pinned memory. STAGE_HOST_PATH=zerocopy: the kernels store keys,
// commit ids and payload rows straight into pinned caller buffers (one pass, no slab) — lowest latency for small batches,
// but on this pool SM stores into mapped host memory reach 44-53 GB/s against 57.3 GB/s for the DMA engine
// (profiles/r2_pcie_ceiling.txt), so it is opt-in.
enum { kHostAuto = 0, kHostStaged = 1, kHostZeroCopy = 2 };
static int host_path_mode() {
    static const int m = [] {
        const char *e = getenv("STAGE_HOST_PATH");
        if (e && !strcmp(e, "staged")) return (int) kHostStaged;
        if (e && !strcmp(e, "zerocopy")) return (int) kHostZeroCopy;
        return (int) kHostAuto;
    }();
    return m;
}

static int probe_host_zerocopy(stage_snapshot *s, HostCtx *c, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *d_found,
                               uint32_t *d_ts, uint8_t *d_key, uint8_t *d_pay) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    int rc;
    if ((rc = c->keys.ensure(n * K))) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->keys.p, keys, n * K, cudaMemcpyHostToDevice, c->st));
    const Lane ln{c->st, &c->srcrow, false};
    const uint64_t chunk = 1ull << 22;                       // bounds the source-row scratch (16 MiB)
    for (uint64_t b = 0; b < n; b += chunk) {
        ProbeArgs a{};
        a.n = std::min(chunk, n - b);
        a.read_ts = read_ts;
        a.keys = static_cast<const uint8_t *>(c->keys.p) + b * K;
        a.found = d_found ? d_found + b : nullptr;
        a.ts_out = d_ts ? d_ts + b : nullptr;
        a.key_out = d_key ? d_key + b * K : nullptr;
        if ((rc = probe_device(s, ln, a, d_pay ? d_pay + b * (uint64_t) P : nullptr))) return rc;
    }
    CUDA_TRY(cudaStreamSynchronize(c->st));
    return STAGE_OK;
}

static int probe_host_staged(stage_snapshot *s, HostCtx *c, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                             uint32_t *ts, void *key_out, void *payload_out, const uint32_t *read_ts_each = nullptr) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    // ~64 MiB of result rows per chunk, two chunks in flight: chunk i is copied out (copy stream) while chunk i+1 is
    // searched and gathered (kernel stream)
    const uint64_t chunk = std::max<uint64_t>(1024, std::min<uint64_t>(1ull << 20, (64ull << 20) / std::max<uint32_t>(P, 1)));
    const uint64_t cm = std::min(chunk, n);
    int rc;
    if ((rc = c->keys.ensure(n * K))) return rc;
    for (int i = 0; i < 2; ++i) {
        if (found && (rc = c->found[i].ensure(cm))) return rc;
        if (ts && (rc = c->ts[i].ensure(cm * 4))) return rc;
        if (key_out && (rc = c->keyout[i].ensure(cm * K))) return rc;
        if (payload_out && (rc = c->payload[i].ensure(cm * (uint64_t) P))) return rc;
    }
    CUDA_TRY(cudaMemcpyAsync(c->keys.p, keys, n * K, cudaMemcpyHostToDevice, c->st));
    if (read_ts_each) {
        if ((rc = c->tse.ensure(n * 4))) return rc;
        CUDA_TRY(cudaMemcpyAsync(c->tse.p, read_ts_each, n * 4, cudaMemcpyHostToDevice, c->st));
    }
    const Lane ln{c->st, &c->srcrow, false};
    uint64_t i = 0;
    for (uint64_t b = 0; b < n; b += chunk, ++i) {
        const int buf = (int) (i & 1);
        const uint64_t m = std::min(chunk, n - b);
        if (i >= 2) CUDA_TRY(cudaStreamWaitEvent(c->st, c->drained[buf], 0));
        ProbeArgs a{};
        a.n = m;
        a.read_ts = read_ts;
        a.read_ts_each = read_ts_each ? static_cast<const uint32_t *>(c->tse.p) + b : nullptr;
        a.keys = static_cast<const uint8_t *>(c->keys.p) + b * K;
        a.found = found ? static_cast<uint8_t *>(c->found[buf].p) : nullptr;
        a.ts_out = ts ? static_cast<uint32_t *>(c->ts[buf].p) : nullptr;
        a.key_out = key_out ? static_cast<uint8_t *>(c->keyout[buf].p) : nullptr;
        // rows of missed probes come back zero-filled in this mode
        if (key_out) CUDA_TRY(cudaMemsetAsync(c->keyout[buf].p, 0, m * K, c->st));
        if (payload_out) CUDA_TRY(cudaMemsetAsync(c->payload[buf].p, 0, m * (uint64_t) P, c->st));
        if ((rc = probe_device(s, ln, a, payload_out ? static_cast<uint8_t *>(c->payload[buf].p) : nullptr))) return rc;
        CUDA_TRY(cudaEventRecord(c->done[buf], c->st));
        CUDA_TRY(cudaStreamWaitEvent(c->cp, c->done[buf], 0));
        if (found) CUDA_TRY(cudaMemcpyAsync(found + b, c->found[buf].p, m, cudaMemcpyDeviceToHost, c->cp));
        if (ts) CUDA_TRY(cudaMemcpyAsync(ts + b, c->ts[buf].p, m * 4, cudaMemcpyDeviceToHost, c->cp));
        if (key_out)
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(key_out) + b * K, c->keyout[buf].p, m * K, cudaMemcpyDeviceToHost, c->cp));
        if (payload_out)
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(payload_out) + b * (uint64_t) P, c->payload[buf].p, m * (uint64_t) P,
                                     cudaMemcpyDeviceToHost, c->cp));
        CUDA_TRY(cudaEventRecord(c->drained[buf], c->cp));
    }
    CUDA_TRY(cudaStreamSynchronize(c->cp));
    CUDA_TRY(cudaStreamSynchronize(c->st));
    return STAGE_OK;
}

constexpr int kNotSmall = 1;                     // small-call path: "does not fit", the caller takes the staged path
static int probe_host_small(stage_snapshot *s, HostCtx *c, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                            uint32_t *ts, void *key_out, void *payload_out);          // "small host calls" below

extern "C" int stage_gpu_probe_batch(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                                     uint32_t *ts, void *key_out, void *payload_out, int flags) {
    if (!s || (n && !keys)) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(s->device));
    if (flags & STAGE_IO_DEVICE) {
        std::lock_guard<std::mutex> lock(s->mu);        // one enqueue at a time on the snapshot stream (shared scratch)
        ProbeArgs a{};
        a.n = n;
        a.read_ts = read_ts;
        a.keys = static_cast<const uint8_t *>(keys);
        a.found = found; a.ts_out = ts;
        a.key_out = static_cast<uint8_t *>(key_out);
        return probe_device(s, Lane{s->stream, &s->sc_srcrow, s->timing}, a, static_cast<uint8_t *>(payload_out));
    }
    if (n == 0) return STAGE_OK;
    int rc = STAGE_OK;
    CtxGuard g{s, ctx_acquire(s, &rc)};
    if (!g.c) return rc;
    if (host_path_mode() == kHostAuto) {
        rc = probe_host_small(s, g.c, keys, n, read_ts, found, ts, key_out, payload_out);
        if (rc != kNotSmall) return rc;
    }
    if (host_path_mode() == kHostZeroCopy) {
        uint8_t *df = static_cast<uint8_t *>(mapped_alias(found)), *dk = static_cast<uint8_t *>(mapped_alias(key_out)),
                *dp = static_cast<uint8_t *>(mapped_alias(payload_out));
        uint32_t *dt = static_cast<uint32_t *>(mapped_alias(ts));
        if ((!found || df) && (!ts || dt) && (!key_out || dk) && (!payload_out || dp))
            return probe_host_zerocopy(s, g.c, keys, n, read_ts, df, dt, dk, dp);
        g_last_error = "STAGE_HOST_PATH=zerocopy needs pinned output buffers (stage_gpu_host_alloc / stage_gpu_host_register)";
        return STAGE_ERR_ARG;
    }
    return probe_host_staged(s, g.c, keys, n, read_ts, found, ts, key_out, payload_out);
}

template <int NW>
static int launch_scan(stage_snapshot *s, const ScanArgs &a, cudaStream_t st) {
    // short scans (stock-level: 10 rows): one warp per scan instead of one CTA per scan
    static const uint32_t warp_max = getenv("STAGE_SCAN_WARP_MAX") ? (uint32_t) atoi(getenv("STAGE_SCAN_WARP_MAX")) : 128;
    if (a.out_stride <= warp_max) {
        constexpr int wpc = kScanWarpThreads / 32;
        const size_t wsmem = (size_t) wpc * ((s->view.max_leaf_rows + 7u) & ~7u) * 2 + 16;
        if (wsmem > 48 * 1024)
            CUDA_TRY(cudaFuncSetAttribute(scan_plan_warp_kernel<NW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) wsmem));
        const unsigned wblocks = (unsigned) std::min<uint64_t>((a.n + wpc - 1) / wpc, (uint64_t) sm_count() * 16);
        scan_plan_warp_kernel<NW><<<wblocks, kScanWarpThreads, wsmem, st>>>(s->view, a);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    const size_t smem = (size_t) std::max<uint32_t>(s->view.max_leaf_rows, 8) * 2;
    const unsigned blocks = (unsigned) std::min<uint64_t>(a.n, (uint64_t) sm_count() * 64);
    scan_plan_kernel<NW><<<blocks, kScanThreads, smem, st>>>(s->view, a);
    LAUNCH_CHECK();
    return STAGE_OK;
}
static int launch_scan_nw(stage_snapshot *s, const ScanArgs &a, cudaStream_t st) {
    switch (s->nw) {
        case 1: return launch_scan<1>(s, a, st);
        case 2: return launch_scan<2>(s, a, st);
        case 4: return launch_scan<4>(s, a, st);
        default: return launch_scan<8>(s, a, st);
    }
}

// ---- fused short scans (scan_short.cuh) ---------------------------------------------------------------------
static int ensure_row_leaf(stage_snapshot *s) {
    std::lock_guard<std::mutex> lk(s->aux_mu);
    if (s->d_row_leaf) return STAGE_OK;
    uint32_t *p = nullptr;
    int rc = dev_alloc(s, &p, s->n_rows);
    if (rc) return rc;
    row_leaf_build_kernel<<<(unsigned) std::min<uint64_t>(std::max<uint64_t>(s->n_leaves, 1), (uint64_t) sm_count() * 16), 128, 0, s->stream>>>(
        s->d_leaf_off, (uint32_t) s->n_leaves, p);
    LAUNCH_CHECK();
    CUDA_TRY(cudaStreamSynchronize(s->stream));       // other streams (host-call contexts) read it next
    s->d_row_leaf = p;
    return STAGE_OK;
}

static bool short_scan_eligible(const stage_snapshot *s, const ScanArgs &a, const uint8_t *payload_out) {
    static const bool off = getenv("STAGE_SCAN_FUSED") && atoi(getenv("STAGE_SCAN_FUSED")) == 0;
    const uint32_t P = s->schema.payload_size, K = s->schema.key_size;
    return !off && a.out_stride >= 1 && a.out_stride <= 32 && (P & 3u) == 0 && P <= kShortMaxRow && K == (uint32_t) s->nw * 4 &&
           (((uintptr_t) payload_out | (uintptr_t) s->d_payload | (uintptr_t) a.start_keys) & 3u) == 0 &&
           ((uintptr_t) a.key_out & 15u) == 0 && s->n_rows > 0;
}

template <int NW, int TW, int MINB>
static int launch_short_scan_v(stage_snapshot *s, const ScanArgs &a, const ShortScanOut &o, cudaStream_t st) {
    constexpr int tiles = kShortThreads / TW;
    const size_t smem = (size_t) ((s->view.max_leaf_rows + 7u) & ~7u) * 2 * tiles + 16;
    if (smem > 48 * 1024)
        CUDA_TRY(cudaFuncSetAttribute(scan_short_fused_kernel<NW, TW, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    const unsigned blocks = (unsigned) std::min<uint64_t>((a.n + tiles - 1) / tiles, (uint64_t) sm_count() * 16);
    scan_short_fused_kernel<NW, TW, MINB><<<blocks, kShortThreads, smem, st>>>(s->view, a, o);
    LAUNCH_CHECK();
    return STAGE_OK;
}
template <int NW>
static int launch_short_scan(stage_snapshot *s, const ScanArgs &a, uint8_t *payload_out, cudaStream_t st) {
    const ShortScanOut o{payload_out, s->d_row_leaf};
    // MINB = resident CTAs per SM the register allocation is held to (8 = all 2048 threads at 32 registers, with spills;
    // 5 = 48 registers, no spills: measured fastest, 0.399 vs 0.460 ms per 10^6 order_line scans)
    static const int minb = getenv("STAGE_SHORT_MINB") ? atoi(getenv("STAGE_SHORT_MINB")) : 5;
    if (a.out_stride <= 16)
        return minb >= 8 ? launch_short_scan_v<NW, 16, 8>(s, a, o, st)
                         : minb <= 4 ? launch_short_scan_v<NW, 16, 4>(s, a, o, st) : launch_short_scan_v<NW, 16, 5>(s, a, o, st);
    return minb >= 8 ? launch_short_scan_v<NW, 32, 8>(s, a, o, st)
                     : minb <= 4 ? launch_short_scan_v<NW, 32, 4>(s, a, o, st) : launch_short_scan_v<NW, 32, 5>(s, a, o, st);
}

// plan (scan_plan_kernel) then payload gather; all pointers are device-visible. dst_base (compact layout): row offset of
// every scan in payload_out, computed by the caller between the two launches (scan_plan_launch / scan_gather_launch).
static int scan_plan_launch(stage_snapshot *s, const Lane &ln, ScanArgs &a) {
    const uint64_t slots = a.n * (uint64_t) a.out_stride;
    int rc;
    if ((rc = ln.srcrow->ensure(std::max<uint64_t>(slots, 1) * 4))) return rc;
    a.src_row = static_cast<uint32_t *>(ln.srcrow->p);
    a.stats = s->d_stats;
    if (ln.timing) CUDA_TRY(cudaEventRecord(s->tev[0], ln.st));
    if ((rc = launch_scan_nw(s, a, ln.st))) return rc;
    if (ln.timing) CUDA_TRY(cudaEventRecord(s->tev[1], ln.st));
    return STAGE_OK;
}
static int scan_gather_launch(stage_snapshot *s, const Lane &ln, const ScanArgs &a, uint8_t *payload_out,
                              const uint32_t *dst_base = nullptr) {
    const int rc = launch_gather(s, ln.st, a.src_row, a.n * (uint64_t) a.out_stride, payload_out, dst_base, a.out_stride);
    if (ln.timing) {
        CUDA_TRY(cudaEventRecord(s->tev[2], ln.st));
        s->timed_gather = payload_out != nullptr;
    }
    return rc;
}
static int scan_device(stage_snapshot *s, const Lane &ln, ScanArgs a, uint8_t *payload_out) {
    if (a.n == 0) return STAGE_OK;
    int rc;
    if (short_scan_eligible(s, a, payload_out)) {      // short scans of short rows: one kernel plans and emits the rows
        if (!s->d_row_leaf && (rc = ensure_row_leaf(s))) return rc;
        a.src_row = nullptr;
        a.stats = s->d_stats;
        if (ln.timing) CUDA_TRY(cudaEventRecord(s->tev[0], ln.st));
        switch (s->nw) {
            case 1: rc = launch_short_scan<1>(s, a, payload_out, ln.st); break;
            case 2: rc = launch_short_scan<2>(s, a, payload_out, ln.st); break;
            case 4: rc = launch_short_scan<4>(s, a, payload_out, ln.st); break;
            default: rc = launch_short_scan<8>(s, a, payload_out, ln.st); break;
        }
        if (ln.timing) {
            CUDA_TRY(cudaEventRecord(s->tev[1], ln.st));
            CUDA_TRY(cudaEventRecord(s->tev[2], ln.st));
            s->timed_gather = false;
        }
        return rc;
    }
    if ((rc = scan_plan_launch(s, ln, a))) return rc;
    return scan_gather_launch(s, ln, a, payload_out);
}

static int scan_host_zerocopy(stage_snapshot *s, HostCtx *c, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                              uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *d_counts, uint8_t *d_key,
                              uint32_t *d_ts, uint8_t *d_pay) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    int rc;
    if ((rc = c->keys.ensure(n * K)) || (rc = c->sizes.ensure(n * 4))) return rc;
    CUDA_TRY(cudaMemcpyAsync(c->keys.p, start_keys, n * K, cudaMemcpyHostToDevice, c->st));
    CUDA_TRY(cudaMemcpyAsync(c->sizes.p, scan_sizes, n * 4, cudaMemcpyHostToDevice, c->st));
    const Lane ln{c->st, &c->srcrow, false};
    // bounds the source-row scratch: <= 64 Mi result slots per launch pair
    const uint64_t chunk = std::max<uint64_t>(1, (1ull << 26) / std::max<uint32_t>(out_stride, 1));
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t rb = b * out_stride;
        ScanArgs a{};
        a.out_stride = out_stride;
        a.read_ts = read_ts;
        a.apply_visibility = apply_visibility;
        a.n = std::min(chunk, n - b);
        a.start_keys = static_cast<const uint8_t *>(c->keys.p) + b * K;
        a.scan_sizes = static_cast<const uint32_t *>(c->sizes.p) + b;
        a.row_counts = d_counts + b;
        a.key_out = d_key ? d_key + rb * K : nullptr;
        a.ts_out = d_ts ? d_ts + rb : nullptr;
        if ((rc = scan_device(s, ln, a, d_pay ? d_pay + rb * (uint64_t) P : nullptr))) return rc;
    }
    CUDA_TRY(cudaStreamSynchronize(c->st));
    return STAGE_OK;
}

// Results through HBM scratch and the DMA engine, two chunks in flight: chunk i is copied out on the copy stream while
// chunk i+1 is planned and gathered on the kernel stream. Measured on this pool (profiles/r2_pcie_ceiling.txt): one
// large cudaMemcpyAsync reaches 57.3 GB/s of the Gen5 x16 link, one copy per scan 46 GB/s, kernel stores into mapped
// pinned memory 44-53 GB/s — so every chunk leaves as ONE copy per output array. compact = STAGE_IO_COMPACT: the rows
// of all scans are packed back to back on the device first, so the copies carry exactly the rows that exist.
static int scan_host_staged(stage_snapshot *s, HostCtx *c, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                            uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
                            uint32_t *ts_out, void *payload_out, bool compact, const uint32_t *read_ts_each = nullptr,
                            const uint8_t *vis_each = nullptr) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    const uint64_t slot_bytes = (uint64_t) P + K + 4;
    // ~128 MiB of result slots per chunk; fewer for small batches so that the first copy starts early
    uint64_t chunk = std::max<uint64_t>(1, (128ull << 20) / std::max<uint64_t>(1, slot_bytes * out_stride));
    if (n < 4 * chunk) chunk = std::max<uint64_t>(1, (n + 3) / 4);
    chunk = std::min<uint64_t>(chunk, n);
    if (chunk * out_stride >= 0xFFFFFFFFull) return STAGE_ERR_ARG;
    int rc;
    if ((rc = c->keys.ensure(n * K)) || (rc = c->sizes.ensure(n * 4))) return rc;
    for (int i = 0; i < 2; ++i) {
        if ((rc = c->counts[i].ensure(chunk * 4))) return rc;
        if (key_out && (rc = c->keyout[i].ensure(chunk * out_stride * K))) return rc;
        if (ts_out && (rc = c->ts[i].ensure(chunk * out_stride * 4))) return rc;
        if (payload_out && (rc = c->payload[i].ensure(chunk * out_stride * (uint64_t) P))) return rc;
    }
    size_t cub_bytes = 0;
    if (compact) {
        if (key_out && (rc = c->kstr.ensure(chunk * out_stride * K))) return rc;
        if (ts_out && (rc = c->tstr.ensure(chunk * out_stride * 4))) return rc;
        if ((rc = c->off.ensure(chunk * 4))) return rc;
        CUDA_TRY(cub::DeviceScan::ExclusiveSum(nullptr, cub_bytes, (const uint32_t *) nullptr, (uint32_t *) nullptr, (int) chunk, c->st));
        if ((rc = c->cubtmp.ensure(cub_bytes))) return rc;
    }
    if (c->h_counts_cap < 2 * chunk) {
        if (c->h_counts) cudaFreeHost(c->h_counts);
        c->h_counts = nullptr;
        c->h_counts_cap = 0;
        CUDA_TRY(cudaHostAlloc((void **) &c->h_counts, 2 * chunk * 4, cudaHostAllocDefault));
        c->h_counts_cap = 2 * chunk;
    }
    CUDA_TRY(cudaMemcpyAsync(c->keys.p, start_keys, n * K, cudaMemcpyHostToDevice, c->st));
    CUDA_TRY(cudaMemcpyAsync(c->sizes.p, scan_sizes, n * 4, cudaMemcpyHostToDevice, c->st));
    if (read_ts_each) {
        if ((rc = c->tse.ensure(n * 4))) return rc;
        CUDA_TRY(cudaMemcpyAsync(c->tse.p, read_ts_each, n * 4, cudaMemcpyHostToDevice, c->st));
    }
    if (vis_each) {
        if ((rc = c->vise.ensure(n))) return rc;
        CUDA_TRY(cudaMemcpyAsync(c->vise.p, vis_each, n, cudaMemcpyHostToDevice, c->st));
    }
    const Lane ln{c->st, &c->srcrow, false};
    uint64_t i = 0, row_base = 0;                 // row_base: rows already returned (compact layout)
    for (uint64_t b = 0; b < n; b += chunk, ++i) {
        const int buf = (int) (i & 1);
        const uint64_t m = std::min(chunk, n - b), slots = m * out_stride;
        if (i >= 2) CUDA_TRY(cudaStreamWaitEvent(c->st, c->drained[buf], 0));
        uint8_t *d_key = key_out ? static_cast<uint8_t *>(c->keyout[buf].p) : nullptr;
        uint32_t *d_ts = ts_out ? static_cast<uint32_t *>(c->ts[buf].p) : nullptr;
        uint8_t *d_pay = payload_out ? static_cast<uint8_t *>(c->payload[buf].p) : nullptr;
        if (!compact) {        // slots beyond a scan's row count come back zero-filled
            if (d_key) CUDA_TRY(cudaMemsetAsync(d_key, 0, slots * K, c->st));
            if (d_ts) CUDA_TRY(cudaMemsetAsync(d_ts, 0, slots * 4, c->st));
            if (d_pay) CUDA_TRY(cudaMemsetAsync(d_pay, 0, slots * (uint64_t) P, c->st));
        }
        ScanArgs a{};
        a.out_stride = out_stride;
        a.read_ts = read_ts;
        a.apply_visibility = apply_visibility;
        a.n = m;
        a.start_keys = static_cast<const uint8_t *>(c->keys.p) + b * K;
        a.scan_sizes = static_cast<const uint32_t *>(c->sizes.p) + b;
        a.row_counts = static_cast<uint32_t *>(c->counts[buf].p);
        a.read_ts_each = read_ts_each ? static_cast<const uint32_t *>(c->tse.p) + b : nullptr;
        a.visibility_each = vis_each ? static_cast<const uint8_t *>(c->vise.p) + b : nullptr;
        a.key_out = compact ? (key_out ? static_cast<uint8_t *>(c->kstr.p) : nullptr) : d_key;
        a.ts_out = compact ? (ts_out ? static_cast<uint32_t *>(c->tstr.p) : nullptr) : d_ts;
        if ((rc = scan_plan_launch(s, ln, a))) return rc;
        uint32_t *hc = c->h_counts + (size_t) buf * chunk;
        CUDA_TRY(cudaMemcpyAsync(hc, a.row_counts, m * 4, cudaMemcpyDeviceToHost, c->st));
        if (compact) {
            uint32_t *d_off = static_cast<uint32_t *>(c->off.p);
            size_t tb = c->cubtmp.cap;
            CUDA_TRY(cub::DeviceScan::ExclusiveSum(c->cubtmp.p, tb, a.row_counts, d_off, (int) m, c->st));
            g_launches.fetch_add(1, std::memory_order_relaxed);
            if ((rc = scan_gather_launch(s, ln, a, d_pay, d_off))) return rc;
            if (d_key || d_ts) {
                compact_keys_ts_kernel<<<(unsigned) ((slots + 255) / 256), 256, 0, c->st>>>(a.row_counts, d_off, slots, out_stride, K,
                                                                                           a.key_out, d_key, a.ts_out, d_ts);
                LAUNCH_CHECK();
            }
        } else if ((rc = scan_gather_launch(s, ln, a, d_pay))) {
            return rc;
        }
        CUDA_TRY(cudaEventRecord(c->done[buf], c->st));
        // the row counts size the copies of the compact layout; the copy stream is still busy with chunk i-1 meanwhile
        CUDA_TRY(cudaEventSynchronize(c->done[buf]));
        memcpy(row_counts + b, hc, m * 4);
        uint64_t rows = slots, dst_row = b * out_stride;
        if (compact) {
            rows = 0;
            for (uint64_t j = 0; j < m; ++j) rows += hc[j];
            dst_row = row_base;
            row_base += rows;
        }
        CUDA_TRY(cudaStreamWaitEvent(c->cp, c->done[buf], 0));
        if (rows) {
            if (key_out)
                CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(key_out) + dst_row * K, d_key, rows * K, cudaMemcpyDeviceToHost, c->cp));
            if (ts_out) CUDA_TRY(cudaMemcpyAsync(ts_out + dst_row, d_ts, rows * 4, cudaMemcpyDeviceToHost, c->cp));
            if (payload_out)
                CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(payload_out) + dst_row * (uint64_t) P, d_pay, rows * (uint64_t) P,
                                         cudaMemcpyDeviceToHost, c->cp));
        }
        CUDA_TRY(cudaEventRecord(c->drained[buf], c->cp));
    }
    CUDA_TRY(cudaStreamSynchronize(c->cp));
    return STAGE_OK;
}

// ---- small host calls -----------------------------------------------------------------------------------------
// A call with a few operations (the reference's drivers issue ONE per executor) is all latency: the staged path costs
// two H2D copies, two fills, two launches, four D2H copies and their events. Below kSmallHostBytes of inputs + results
// the kernels instead read the keys from, and store the results into, a pinned slab of the call context through its
// device alias: memcpy in, two launches, one stream synchronisation, memcpy out — any host memory, either layout.
constexpr size_t kSmallHostBytes = 192u << 10;

static int ctx_small_slab(HostCtx *c) {
    if (c->h_small) return c->d_small ? STAGE_OK : kNotSmall;
    CUDA_TRY(cudaHostAlloc((void **) &c->h_small, kSmallHostBytes, cudaHostAllocDefault));
    c->d_small = static_cast<uint8_t *>(mapped_alias(c->h_small));
    return c->d_small ? STAGE_OK : kNotSmall;
}

struct SlabCarve {
    size_t off = 0;
    size_t take(size_t bytes) {
        const size_t o = off;
        off = (off + bytes + 15) & ~size_t(15);
        return o;
    }
};

static int probe_host_small(stage_snapshot *s, HostCtx *c, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                            uint32_t *ts, void *key_out, void *payload_out) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if (n > kSmallHostBytes) return kNotSmall;
    SlabCarve cv;
    const size_t o_keys = cv.take(n * K), o_found = cv.take(n), o_ts = cv.take(n * 4), o_kout = cv.take(key_out ? n * K : 0),
                 o_pay = cv.take(payload_out ? n * (size_t) P : 0);
    if (cv.off > kSmallHostBytes) return kNotSmall;
    int rc;
    if ((rc = ctx_small_slab(c))) return rc;
    uint8_t *h = c->h_small, *d = c->d_small;
    memcpy(h + o_keys, keys, n * K);
    ProbeArgs a{};
    a.n = n;
    a.read_ts = read_ts;
    a.keys = d + o_keys;
    a.found = d + o_found;
    a.ts_out = reinterpret_cast<uint32_t *>(d + o_ts);
    a.key_out = key_out ? d + o_kout : nullptr;
    if ((rc = probe_device(s, Lane{c->st, &c->srcrow, false}, a, payload_out ? d + o_pay : nullptr))) return rc;
    CUDA_TRY(cudaStreamSynchronize(c->st));
    if (found) memcpy(found, h + o_found, n);
    if (ts) memcpy(ts, h + o_ts, n * 4);
    for (uint64_t i = 0; i < n; ++i) {               // rows of missed probes stay untouched
        if (!h[o_found + i]) continue;
        if (key_out) memcpy(static_cast<uint8_t *>(key_out) + i * K, h + o_kout + i * K, K);
        if (payload_out) memcpy(static_cast<uint8_t *>(payload_out) + i * (size_t) P, h + o_pay + i * (size_t) P, P);
    }
    return STAGE_OK;
}

static int scan_host_small(stage_snapshot *s, HostCtx *c, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                           uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
                           uint32_t *ts_out, void *payload_out, bool compact) {
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if (n > kSmallHostBytes || (uint64_t) n * out_stride > kSmallHostBytes) return kNotSmall;
    const size_t slots = (size_t) n * out_stride;
    SlabCarve cv;
    const size_t o_keys = cv.take(n * K), o_sizes = cv.take(n * 4), o_counts = cv.take(n * 4), o_ts = cv.take(ts_out ? slots * 4 : 0),
                 o_kout = cv.take(key_out ? slots * K : 0), o_pay = cv.take(payload_out ? slots * P : 0);
    if (cv.off > kSmallHostBytes) return kNotSmall;
    int rc;
    if ((rc = ctx_small_slab(c))) return rc;
    uint8_t *h = c->h_small, *d = c->d_small;
    memcpy(h + o_keys, start_keys, n * K);
    memcpy(h + o_sizes, scan_sizes, n * 4);
    ScanArgs a{};
    a.n = n;
    a.out_stride = out_stride;
    a.read_ts = read_ts;
    a.apply_visibility = apply_visibility;
    a.start_keys = d + o_keys;
    a.scan_sizes = reinterpret_cast<const uint32_t *>(d + o_sizes);
    a.row_counts = reinterpret_cast<uint32_t *>(d + o_counts);
    a.key_out = key_out ? d + o_kout : nullptr;
    a.ts_out = ts_out ? reinterpret_cast<uint32_t *>(d + o_ts) : nullptr;
    if ((rc = scan_device(s, Lane{c->st, &c->srcrow, false}, a, payload_out ? d + o_pay : nullptr))) return rc;
    CUDA_TRY(cudaStreamSynchronize(c->st));
    const uint32_t *cnt = reinterpret_cast<const uint32_t *>(h + o_counts);
    memcpy(row_counts, cnt, n * 4);
    size_t packed = 0;
    for (uint64_t i = 0; i < n; ++i) {               // row slots beyond row_counts[i] stay untouched
        const size_t src = (size_t) i * out_stride, dst = compact ? packed : src, m = cnt[i];
        packed += m;
        if (!m) continue;
        if (ts_out) memcpy(ts_out + dst, h + o_ts + src * 4, m * 4);
        if (key_out) memcpy(static_cast<uint8_t *>(key_out) + dst * K, h + o_kout + src * K, m * K);
        if (payload_out) memcpy(static_cast<uint8_t *>(payload_out) + dst * P, h + o_pay + src * P, m * P);
    }
    return STAGE_OK;
}

extern "C" int stage_gpu_scan_batch(stage_snapshot_t *s, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                                    uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts,
                                    void *key_out, uint32_t *ts_out, void *payload_out, int flags) {
    if (!s || (n && (!start_keys || !scan_sizes || !row_counts))) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(s->device));
    if (flags & STAGE_IO_DEVICE) {
        if (flags & STAGE_IO_COMPACT) return STAGE_ERR_ARG;      // the compact layout is a host-buffer (PCIe) feature
        std::lock_guard<std::mutex> lock(s->mu);        // one enqueue at a time on the snapshot stream (shared scratch)
        ScanArgs a{};
        a.out_stride = out_stride;
        a.read_ts = read_ts;
        a.apply_visibility = apply_visibility;
        a.start_keys = static_cast<const uint8_t *>(start_keys);
        a.scan_sizes = scan_sizes;                      // sizes above out_stride are clamped by the plan kernels
        a.n = n;
        a.row_counts = row_counts;
        a.key_out = static_cast<uint8_t *>(key_out);
        a.ts_out = ts_out;
        return scan_device(s, Lane{s->stream, &s->sc_srcrow, s->timing}, a, static_cast<uint8_t *>(payload_out));
    }
    for (uint64_t i = 0; i < n; ++i)
        if (scan_sizes[i] > out_stride) return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    int rc = STAGE_OK;
    CtxGuard g{s, ctx_acquire(s, &rc)};
    if (!g.c) return rc;
    const bool compact = (flags & STAGE_IO_COMPACT) != 0;
    if (host_path_mode() == kHostAuto) {
        rc = scan_host_small(s, g.c, start_keys, scan_sizes, n, out_stride, read_ts, apply_visibility, row_counts, key_out, ts_out,
                             payload_out, compact);
        if (rc != kNotSmall) return rc;
    }
    if (host_path_mode() == kHostZeroCopy && !compact) {
        uint32_t *dc = static_cast<uint32_t *>(mapped_alias(row_counts)), *dt = static_cast<uint32_t *>(mapped_alias(ts_out));
        uint8_t *dk = static_cast<uint8_t *>(mapped_alias(key_out)), *dp = static_cast<uint8_t *>(mapped_alias(payload_out));
        if (dc && (!ts_out || dt) && (!key_out || dk) && (!payload_out || dp))
            return scan_host_zerocopy(s, g.c, start_keys, scan_sizes, n, out_stride, read_ts, apply_visibility, dc, dk, dt, dp);
        g_last_error = "STAGE_HOST_PATH=zerocopy needs pinned output buffers (stage_gpu_host_alloc / stage_gpu_host_register)";
        return STAGE_ERR_ARG;
    }
    return scan_host_staged(s, g.c, start_keys, scan_sizes, n, out_stride, read_ts, apply_visibility, row_counts, key_out, ts_out,
                            payload_out, compact);
}

#include "batcher.inc"

// ------------------------------------------------------------------------------------------------
// column projection (push-down): plan kernels + gather_cols_kernel
// ------------------------------------------------------------------------------------------------
static int launch_gather_cols(stage_snapshot *s, const uint32_t *src_row, uint64_t n_slots, uint32_t col_off,
                              uint32_t col_len, uint8_t *dst) {
    if (n_slots == 0 || !dst) return STAGE_OK;
    GatherColsArgs g{};
    g.src_row = src_row; g.n_slots = n_slots; g.payload = s->d_payload; g.ver_payload = s->d_ver_payload; g.dst = dst;
    g.payload_size = s->schema.payload_size; g.col_off = col_off; g.col_len = col_len;
    gather_cols_kernel<<<(unsigned) ((n_slots + 255) / 256), 256, 0, s->stream>>>(g);
    LAUNCH_CHECK();
    return STAGE_OK;
}

extern "C" int stage_gpu_probe_project(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                                       uint32_t *ts, uint32_t col_off, uint32_t col_len, void *col_out, int flags) {
    if (!s || (n && (!keys || !col_out))) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if (col_len == 0 || (uint64_t) col_off + col_len > P) return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    int rc;
    const bool dev = flags & STAGE_IO_DEVICE;
    const uint64_t chunk = dev ? n : std::min<uint64_t>(n, 1ull << 22);
    if ((rc = s->sc_srcrow.ensure(chunk * 4))) return rc;
    if (!dev) {
        if ((rc = s->sc_keys.ensure(chunk * K)) || (rc = s->sc_payload.ensure(chunk * (uint64_t) col_len))) return rc;
        if (found && (rc = s->sc_found.ensure(chunk))) return rc;
        if (ts && (rc = s->sc_ts.ensure(chunk * 4))) return rc;
    }
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t m = std::min(chunk, n - b);
        ProbeArgs a{};
        a.n = m;
        a.read_ts = read_ts;
        a.src_row = static_cast<uint32_t *>(s->sc_srcrow.p);
        uint8_t *cdst;
        if (dev) {
            a.keys = static_cast<const uint8_t *>(keys);
            a.found = found; a.ts_out = ts;
            cdst = static_cast<uint8_t *>(col_out);
        } else {
            CUDA_TRY(cudaMemcpyAsync(s->sc_keys.p, static_cast<const uint8_t *>(keys) + b * K, m * K, cudaMemcpyHostToDevice, s->stream));
            a.keys = static_cast<const uint8_t *>(s->sc_keys.p);
            a.found = found ? static_cast<uint8_t *>(s->sc_found.p) : nullptr;
            a.ts_out = ts ? static_cast<uint32_t *>(s->sc_ts.p) : nullptr;
            cdst = static_cast<uint8_t *>(s->sc_payload.p);
        }
        if ((rc = launch_probe_nw(s, s->stream, a))) return rc;
        if ((rc = launch_gather_cols(s, a.src_row, m, col_off, col_len, cdst))) return rc;
        if (!dev) {
            if (found) CUDA_TRY(cudaMemcpyAsync(found + b, s->sc_found.p, m, cudaMemcpyDeviceToHost, s->stream));
            if (ts) CUDA_TRY(cudaMemcpyAsync(ts + b, s->sc_ts.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(col_out) + b * col_len, s->sc_payload.p, m * (uint64_t) col_len,
                                     cudaMemcpyDeviceToHost, s->stream));
            CUDA_TRY(cudaStreamSynchronize(s->stream));
        }
    }
    return STAGE_OK;
}

extern "C" int stage_gpu_probe_aggregate_i32(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts,
                                             uint32_t col_off, int32_t threshold, stage_agg_i32_t *out, int flags) {
    if (!s || !out || (n && !keys)) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if ((uint64_t) col_off + 4 > P) return STAGE_ERR_ARG;
    static_assert(sizeof(AggI32) == sizeof(stage_agg_i32_t), "aggregate record layout");
    AggI32 init{0, 0, 0x7FFFFFFF, (int) 0x80000000};
    if (n == 0) {
        memcpy(out, &init, sizeof(init));
        return STAGE_OK;
    }
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    int rc;
    const bool dev = flags & STAGE_IO_DEVICE;
    const uint64_t chunk = dev ? n : std::min<uint64_t>(n, 1ull << 22);
    if ((rc = s->sc_srcrow.ensure(chunk * 4)) || (rc = s->sc_counts.ensure(sizeof(AggI32)))) return rc;
    if (!dev && (rc = s->sc_keys.ensure(chunk * K))) return rc;
    AggI32 *d_out = static_cast<AggI32 *>(s->sc_counts.p);
    CUDA_TRY(cudaMemcpyAsync(d_out, &init, sizeof(init), cudaMemcpyHostToDevice, s->stream));
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t m = std::min(chunk, n - b);
        ProbeArgs a{};
        a.n = m;
        a.read_ts = read_ts;
        a.src_row = static_cast<uint32_t *>(s->sc_srcrow.p);
        if (dev) a.keys = static_cast<const uint8_t *>(keys);
        else {
            CUDA_TRY(cudaMemcpyAsync(s->sc_keys.p, static_cast<const uint8_t *>(keys) + b * K, m * K, cudaMemcpyHostToDevice, s->stream));
            a.keys = static_cast<const uint8_t *>(s->sc_keys.p);
        }
        if ((rc = launch_probe_nw(s, s->stream, a))) return rc;
        const unsigned blocks = (unsigned) std::min<uint64_t>((m + 255) / 256, (uint64_t) sm_count() * 8);
        aggregate_i32_kernel<<<blocks, 256, 0, s->stream>>>(a.src_row, m, s->d_payload, s->d_ver_payload, P, col_off, threshold, d_out);
        LAUNCH_CHECK();
        if (!dev) CUDA_TRY(cudaStreamSynchronize(s->stream));    // sc_keys is reused by the next chunk
    }
    CUDA_TRY(cudaMemcpyAsync(out, d_out, sizeof(AggI32), cudaMemcpyDeviceToHost, s->stream));
    CUDA_TRY(cudaStreamSynchronize(s->stream));
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of stage_gpu_probe_aggregate_i32 (no counterpart in the reference).
// Callers that cannot call cudaMalloc — code under CUDA graph capture, tight serving loops, harnesses
// that forbid it — need an entry point that allocates nothing. The ordinary call keeps the source row of every probe in
// a scratch buffer it grows with cudaMalloc; here that scratch is a file-scope device array of a fixed maximum size
// (kNoAllocMaxProbes source rows = 4 MiB, held by every GPU that loads the library until exit).
// The array's name is a host-side stand-in in host code, not a device address: its address is taken once per device,
// with that device current, through cudaGetSymbolAddress and cached. Every call shares the array (it keeps the last
// call's contents; entries [0, n) are rewritten by the probe kernel before they are read, the caller's result record is
// initialised by a kernel): calls that overlap on the GPU race.
// ------------------------------------------------------------------------------------------------
constexpr uint64_t kNoAllocMaxProbes = 1ull << 20;
static __device__ __align__(256) uint32_t g_noalloc_src_row[kNoAllocMaxProbes];

__global__ void noalloc_agg_init_kernel(AggI32 *out) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *out = AggI32{0ull, 0ull, 0x7FFFFFFF, (int) 0x80000000};
}

constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static uint32_t *g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

// device address of the scratch array on `device` (which must be current)
static int noalloc_scratch(int device, uint32_t **out) {
    if (device < 0 || device >= kNoAllocMaxDevices) return STAGE_ERR_ARG;
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        uint32_t *p = nullptr;
        CUDA_TRY(cudaGetSymbolAddress((void **) &p, g_noalloc_src_row));
        if (!p) {
            g_last_error = "cudaGetSymbolAddress returned a null address for the no-alloc scratch array";
            return STAGE_ERR_CUDA;
        }
        g_noalloc_ptr[device] = p;
    }
    *out = g_noalloc_ptr[device];
    return STAGE_OK;
}

extern "C" int stage_gpu_probe_aggregate_i32_noalloc(stage_snapshot_t *s, const void *keys_dev, uint64_t n, uint32_t read_ts,
                                                     uint32_t col_off, int32_t threshold, stage_agg_i32_t *out_dev) {
    if (!s || !out_dev || (n && !keys_dev)) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    if ((uint64_t) col_off + 4 > s->schema.payload_size) return STAGE_ERR_ARG;
    if (n > kNoAllocMaxProbes) {
        g_last_error = "stage_gpu_probe_aggregate_i32_noalloc: n = " + std::to_string(n) + " probes exceed the fixed scratch of " +
                       std::to_string(kNoAllocMaxProbes) + " (split the batch or use stage_gpu_probe_aggregate_i32)";
        return STAGE_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(s->mu);          // one enqueue at a time on the snapshot's stream
    CUDA_TRY(cudaSetDevice(s->device));
    uint32_t *src_row = nullptr;
    int rc;
    if ((rc = noalloc_scratch(s->device, &src_row))) return rc;
    AggI32 *d_out = reinterpret_cast<AggI32 *>(out_dev);
    noalloc_agg_init_kernel<<<1, 32, 0, s->stream>>>(d_out);
    LAUNCH_CHECK();
    if (n == 0) return STAGE_OK;
    ProbeArgs a{};
    a.n = n;
    a.read_ts = read_ts;
    a.src_row = src_row;
    a.keys = static_cast<const uint8_t *>(keys_dev);
    if ((rc = launch_probe_nw(s, s->stream, a))) return rc;
    const unsigned blocks = (unsigned) std::min<uint64_t>((n + 255) / 256, (uint64_t) sm_count() * 8);
    aggregate_i32_kernel<<<blocks, 256, 0, s->stream>>>(src_row, n, s->d_payload, s->d_ver_payload, s->schema.payload_size, col_off,
                                                        threshold, d_out);
    LAUNCH_CHECK();
    return STAGE_OK;
}

extern "C" int stage_gpu_probe_group_aggregate_i32(stage_snapshot_t *s, const void *keys, uint64_t n, const uint64_t *group_offsets,
                                                   uint64_t n_groups, uint32_t read_ts, uint32_t col_off, int32_t threshold,
                                                   stage_group_agg_i32_t *out, int flags) {
    if (!s || (n && !keys) || (n_groups && (!group_offsets || !out))) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if ((uint64_t) col_off + 4 > P || n > (1ull << 26) || n_groups > (1ull << 26)) return STAGE_ERR_ARG;
    static_assert(sizeof(GroupAggI32) == sizeof(stage_group_agg_i32_t), "group aggregate record layout");
    if (n_groups == 0) return STAGE_OK;
    const bool dev = flags & STAGE_IO_DEVICE;
    if (!dev) {
        if (group_offsets[0] != 0 || group_offsets[n_groups] != n) return STAGE_ERR_ARG;
        for (uint64_t g = 0; g < n_groups; ++g)
            if (group_offsets[g] > group_offsets[g + 1]) return STAGE_ERR_ARG;
    }
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    int rc;
    if ((rc = s->sc_srcrow.ensure(std::max<uint64_t>(n, 1) * 4))) return rc;
    const uint64_t *d_off = group_offsets;
    GroupAggI32 *d_out = reinterpret_cast<GroupAggI32 *>(out);
    ProbeArgs a{};
    a.n = n;
    a.read_ts = read_ts;
    a.src_row = static_cast<uint32_t *>(s->sc_srcrow.p);
    a.keys = static_cast<const uint8_t *>(keys);
    if (!dev) {
        // scratch: keys | offsets | records
        const size_t o_off = (n * K + 15) & ~size_t(15), o_out = o_off + (n_groups + 1) * 8;
        if ((rc = s->sc_keys.ensure(o_out + n_groups * sizeof(GroupAggI32)))) return rc;
        uint8_t *base = static_cast<uint8_t *>(s->sc_keys.p);
        if (n) CUDA_TRY(cudaMemcpyAsync(base, keys, n * K, cudaMemcpyHostToDevice, s->stream));
        CUDA_TRY(cudaMemcpyAsync(base + o_off, group_offsets, (n_groups + 1) * 8, cudaMemcpyHostToDevice, s->stream));
        a.keys = base;
        d_off = reinterpret_cast<const uint64_t *>(base + o_off);
        d_out = reinterpret_cast<GroupAggI32 *>(base + o_out);
    }
    if (n && (rc = launch_probe_nw(s, s->stream, a))) return rc;
    const unsigned blocks = (unsigned) std::min<uint64_t>((n_groups + 7) / 8, (uint64_t) sm_count() * 8);
    group_aggregate_i32_kernel<<<blocks, 256, 0, s->stream>>>(a.src_row, d_off, n_groups, s->d_payload, s->d_ver_payload, P, col_off,
                                                              threshold, d_out);
    LAUNCH_CHECK();
    if (!dev) {
        CUDA_TRY(cudaMemcpyAsync(out, d_out, n_groups * sizeof(GroupAggI32), cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    return STAGE_OK;
}

extern "C" int stage_gpu_scan_project(stage_snapshot_t *s, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                                      uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts,
                                      void *key_out, uint32_t col_off, uint32_t col_len, void *col_out, int flags) {
    if (!s || (n && (!start_keys || !scan_sizes || !row_counts || !col_out))) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    if (col_len == 0 || (uint64_t) col_off + col_len > P) return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    const bool dev = flags & STAGE_IO_DEVICE;
    if (!dev)
        for (uint64_t i = 0; i < n; ++i)
            if (scan_sizes[i] > out_stride) return STAGE_ERR_ARG;
    int rc;
    const uint64_t chunk = dev ? n : std::min<uint64_t>(n, std::max<uint64_t>(1, (1ull << 26) / std::max<uint32_t>(out_stride, 1)));
    if ((rc = s->sc_srcrow.ensure(chunk * out_stride * 4))) return rc;
    if (!dev) {
        if ((rc = s->sc_keys.ensure(chunk * K)) || (rc = s->sc_sizes.ensure(chunk * 4)) || (rc = s->sc_counts.ensure(chunk * 4)) ||
            (rc = s->sc_payload.ensure(chunk * out_stride * (uint64_t) col_len)))
            return rc;
        if (key_out && (rc = s->sc_keyout.ensure(chunk * out_stride * K))) return rc;
    }
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t m = std::min(chunk, n - b), rows = m * out_stride;
        ScanArgs a{};
        a.n = m;
        a.out_stride = out_stride;
        a.read_ts = read_ts;
        a.apply_visibility = apply_visibility;
        a.src_row = static_cast<uint32_t *>(s->sc_srcrow.p);
        uint8_t *cdst;
        if (dev) {
            a.start_keys = static_cast<const uint8_t *>(start_keys);
            a.scan_sizes = scan_sizes;
            a.row_counts = row_counts;
            a.key_out = static_cast<uint8_t *>(key_out);
            cdst = static_cast<uint8_t *>(col_out);
        } else {
            CUDA_TRY(cudaMemcpyAsync(s->sc_keys.p, static_cast<const uint8_t *>(start_keys) + b * K, m * K, cudaMemcpyHostToDevice, s->stream));
            CUDA_TRY(cudaMemcpyAsync(s->sc_sizes.p, scan_sizes + b, m * 4, cudaMemcpyHostToDevice, s->stream));
            if (key_out) CUDA_TRY(cudaMemsetAsync(s->sc_keyout.p, 0, rows * K, s->stream));
            a.start_keys = static_cast<const uint8_t *>(s->sc_keys.p);
            a.scan_sizes = static_cast<const uint32_t *>(s->sc_sizes.p);
            a.row_counts = static_cast<uint32_t *>(s->sc_counts.p);
            a.key_out = key_out ? static_cast<uint8_t *>(s->sc_keyout.p) : nullptr;
            cdst = static_cast<uint8_t *>(s->sc_payload.p);
        }
        a.stats = s->d_stats;
        if ((rc = launch_scan_nw(s, a, s->stream))) return rc;
        if ((rc = launch_gather_cols(s, a.src_row, rows, col_off, col_len, cdst))) return rc;
        if (!dev) {
            CUDA_TRY(cudaMemcpyAsync(row_counts + b, s->sc_counts.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
            if (key_out)
                CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(key_out) + b * out_stride * K, s->sc_keyout.p, rows * K,
                                         cudaMemcpyDeviceToHost, s->stream));
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(col_out) + b * out_stride * col_len, s->sc_payload.p,
                                     rows * (uint64_t) col_len, cudaMemcpyDeviceToHost, s->stream));
            CUDA_TRY(cudaStreamSynchronize(s->stream));
        }
    }
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// RunStockLevel's read side as one operator (tpcc_ops.cuh)
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_tpcc_stock_level(stage_snapshot_t *order_line, stage_snapshot_t *stock, const stage_stock_level_req_t *reqs,
                                          uint64_t n, uint32_t read_ts, uint32_t *distinct_items, int flags) {
    static_assert(sizeof(StockLevelReq) == sizeof(stage_stock_level_req_t), "request layout");
    if (!order_line || !stock || (n && (!reqs || !distinct_items))) return STAGE_ERR_ARG;
    if (!order_line->sealed || !stock->sealed) return STAGE_ERR_STATE;
    stage_snapshot *ol = order_line;
    if (ol->schema.key_size != 32 || stock->schema.key_size != 16 || (ol->schema.payload_size & 3u) || ol->schema.payload_size < 4 ||
        (stock->schema.payload_size & 3u) || stock->schema.payload_size < 4 || ol->device != stock->device || ol->xormask || stock->xormask)
        return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    std::lock_guard<std::mutex> lock(ol->mu);
    CUDA_TRY(cudaSetDevice(ol->device));
    cudaStream_t st = ol->stream;
    const bool dev = flags & STAGE_IO_DEVICE;
    const uint64_t chunk = std::min<uint64_t>(n, 1ull << 16);           // transactions per pass: ~130 MB of scratch
    const uint64_t cs = chunk * kStockLevelOrders;
    int rc;
    if ((rc = ol->sl_keys.ensure(cs * 32)) || (rc = ol->sl_sizes.ensure(cs * 4)) || (rc = ol->sl_counts.ensure(cs * 4)) ||
        (rc = ol->sl_src.ensure(cs * kStockLevelScan * 4)) || (rc = ol->sl_skeys.ensure(cs * 16)) || (rc = ol->sl_psrc.ensure(cs * 4)))
        return rc;
    if (!dev && ((rc = ol->sl_req.ensure(chunk * sizeof(StockLevelReq))) || (rc = ol->sl_out.ensure(chunk * 4)))) return rc;
    if (!ol->d_row_leaf && (rc = ensure_row_leaf(ol))) return rc;
    if (dev) CUDA_TRY(cudaStreamSynchronize(stock->stream));            // the stock snapshot's own pending work (updates)
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t m = std::min(chunk, n - b), ms = m * kStockLevelOrders;
        const StockLevelReq *d_req;
        uint32_t *d_out;
        if (dev) {
            d_req = reinterpret_cast<const StockLevelReq *>(reqs) + b;
            d_out = distinct_items + b;
        } else {
            CUDA_TRY(cudaMemcpyAsync(ol->sl_req.p, reqs + b, m * sizeof(StockLevelReq), cudaMemcpyHostToDevice, st));
            d_req = static_cast<const StockLevelReq *>(ol->sl_req.p);
            d_out = static_cast<uint32_t *>(ol->sl_out.p);
        }
        int64_t *d_keys = static_cast<int64_t *>(ol->sl_keys.p);
        stock_level_keys_kernel<<<(unsigned) ((ms + 255) / 256), 256, 0, st>>>(d_req, m, d_keys, static_cast<uint32_t *>(ol->sl_sizes.p));
        LAUNCH_CHECK();
        ScanArgs a{};
        a.start_keys = static_cast<const uint8_t *>(ol->sl_keys.p);
        a.scan_sizes = static_cast<const uint32_t *>(ol->sl_sizes.p);
        a.n = ms;
        a.out_stride = kStockLevelScan;
        a.read_ts = read_ts;
        a.apply_visibility = 1;
        a.row_counts = static_cast<uint32_t *>(ol->sl_counts.p);
        a.src_row = static_cast<uint32_t *>(ol->sl_src.p);
        a.stats = ol->d_stats;
        if (short_scan_eligible(ol, a, nullptr)) rc = launch_short_scan<8>(ol, a, nullptr, st);
        else rc = launch_scan_nw(ol, a, st);
        if (rc) return rc;
        int64_t *d_skeys = static_cast<int64_t *>(ol->sl_skeys.p);
        stock_level_pick_kernel<<<(unsigned) ((ms + 255) / 256), 256, 0, st>>>(ol->view, d_keys, a.row_counts, a.src_row, ms, d_skeys);
        LAUNCH_CHECK();
        ProbeArgs pa{};
        pa.keys = reinterpret_cast<const uint8_t *>(d_skeys);
        pa.n = ms;
        pa.read_ts = read_ts;
        pa.src_row = static_cast<uint32_t *>(ol->sl_psrc.p);
        if ((rc = launch_probe_nw(stock, st, pa))) return rc;
        stock_level_count_kernel<<<(unsigned) ((m + 127) / 128), 128, 0, st>>>(d_req, m, d_skeys, pa.src_row, stock->d_payload,
                                                                              stock->d_ver_payload, stock->schema.payload_size, d_out);
        LAUNCH_CHECK();
        if (!dev) {
            CUDA_TRY(cudaMemcpyAsync(distinct_items + b, d_out, m * 4, cudaMemcpyDeviceToHost, st));
            CUDA_TRY(cudaStreamSynchronize(st));
        }
    }
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// physical whole-tree scan (TableScanExecutor, scan_sz == -1)
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_scan_physical(stage_snapshot_t *s, uint64_t first_row, uint64_t n, void *key_out, uint32_t *ts_out,
                                       void *payload_out, int flags) {
    if (!s) return STAGE_ERR_ARG;
    if (!s->sealed) return STAGE_ERR_STATE;
    if (first_row > s->n_rows || n > s->n_rows - first_row) return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    std::lock_guard<std::mutex> lock(s->mu);
    CUDA_TRY(cudaSetDevice(s->device));
    const uint32_t K = s->schema.key_size, P = s->schema.payload_size;
    int rc;
    if (!s->d_phys) {
        if ((rc = dev_alloc(s, &s->d_phys, s->n_rows))) return rc;
        physical_order_kernel<<<(unsigned) s->n_leaves, 128, (size_t) std::max<uint32_t>(s->max_leaf_rows, 8) * 2, s->stream>>>(s->view, s->d_phys);
        LAUNCH_CHECK();
    }
    auto emit = [&](uint64_t first, uint64_t m, uint8_t *k, uint32_t *t) -> int {
        if (!k && !t) return STAGE_OK;
        const unsigned blocks = (unsigned) ((m + 255) / 256);
        switch (s->nw) {
            case 1: physical_emit_kernel<1><<<blocks, 256, 0, s->stream>>>(s->view, s->d_phys, first, m, k, t); break;
            case 2: physical_emit_kernel<2><<<blocks, 256, 0, s->stream>>>(s->view, s->d_phys, first, m, k, t); break;
            case 4: physical_emit_kernel<4><<<blocks, 256, 0, s->stream>>>(s->view, s->d_phys, first, m, k, t); break;
            default: physical_emit_kernel<8><<<blocks, 256, 0, s->stream>>>(s->view, s->d_phys, first, m, k, t); break;
        }
        LAUNCH_CHECK();
        return STAGE_OK;
    };
    if (flags & STAGE_IO_DEVICE) {
        if ((rc = emit(first_row, n, static_cast<uint8_t *>(key_out), ts_out))) return rc;
        return launch_gather(s, s->stream, s->d_phys + first_row, n, static_cast<uint8_t *>(payload_out));
    }
    const uint64_t chunk = std::max<uint64_t>(1, std::min<uint64_t>(n, (512ull << 20) / (P + K + 4)));
    if (key_out && (rc = s->sc_keyout.ensure(chunk * K))) return rc;
    if (ts_out && (rc = s->sc_ts.ensure(chunk * 4))) return rc;
    if (payload_out && (rc = s->sc_payload.ensure(chunk * (uint64_t) P))) return rc;
    for (uint64_t b = 0; b < n; b += chunk) {
        const uint64_t m = std::min(chunk, n - b);
        if ((rc = emit(first_row + b, m, key_out ? static_cast<uint8_t *>(s->sc_keyout.p) : nullptr,
                       ts_out ? static_cast<uint32_t *>(s->sc_ts.p) : nullptr)))
            return rc;
        if (payload_out && (rc = launch_gather(s, s->stream, s->d_phys + first_row + b, m, static_cast<uint8_t *>(s->sc_payload.p)))) return rc;
        if (key_out) CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(key_out) + b * K, s->sc_keyout.p, m * K, cudaMemcpyDeviceToHost, s->stream));
        if (ts_out) CUDA_TRY(cudaMemcpyAsync(ts_out + b, s->sc_ts.p, m * 4, cudaMemcpyDeviceToHost, s->stream));
        if (payload_out)
            CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t *>(payload_out) + b * (uint64_t) P, s->sc_payload.p, m * (uint64_t) P,
                                     cudaMemcpyDeviceToHost, s->stream));
        CUDA_TRY(cudaStreamSynchronize(s->stream));
    }
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// routing (host): owner = number of splitters < key, in reference key order
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_route(const stage_schema_t *schema, const void *splitters, uint32_t n_shards, const void *keys,
                               uint64_t n, uint32_t *owner) {
    int nw;
    uint32_t xm;
    int rc = schema_check(schema, &nw, &xm);
    if (rc) return rc;
    if (n_shards == 0 || (n && (!keys || !owner)) || (n_shards > 1 && !splitters)) return STAGE_ERR_ARG;
    const uint32_t K = schema->key_size;
    std::vector<uint32_t> sp((size_t) (n_shards - 1) * nw);
    for (uint32_t i = 0; i + 1 < n_shards; ++i)
        host_sortable_nw(nw, static_cast<const uint8_t *>(splitters) + (size_t) i * K, K, xm, sp.data() + (size_t) i * nw);
    std::vector<uint32_t> k(nw);
    for (uint64_t i = 0; i < n; ++i) {
        host_sortable_nw(nw, static_cast<const uint8_t *>(keys) + i * K, K, xm, k.data());
        uint32_t lo = 0, hi = n_shards - 1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            const uint32_t *sm = sp.data() + (size_t) mid * nw;
            bool less = false;                       // splitter[mid] < key ?
            for (int w = 0; w < nw; ++w)
                if (sm[w] != k[w]) { less = sm[w] < k[w]; break; }
            if (less) lo = mid + 1; else hi = mid;
        }
        owner[i] = lo;
    }
    return STAGE_OK;
}

// device-side routing: one thread per key, splitters (<= 63) passed by value in sortable form
struct RouteArgs {
    uint32_t split[64 * 8];
    uint32_t n_split;
};
template <int NW>
__global__ void route_kernel(RouteArgs r, const uint8_t *keys, uint64_t n, uint32_t key_size, uint32_t xormask,
                             uint32_t *owner) {
    const uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const SKey<NW> k = to_sortable<NW>(keys + i * key_size, key_size, xormask);
    uint32_t lo = 0, hi = r.n_split;                 // owner = number of splitters < key
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        SKey<NW> sp;
#pragma unroll
        for (int w = 0; w < NW; ++w) sp.w[w] = r.split[mid * NW + w];
        if (skey_less<NW>(sp, k)) lo = mid + 1; else hi = mid;
    }
    owner[i] = lo;
}

extern "C" int stage_gpu_route_device(const stage_schema_t *schema, const void *splitters, uint32_t n_shards,
                                      const void *keys_dev, uint64_t n, uint32_t *owner_dev, void *cuda_stream) {
    int nw;
    uint32_t xm;
    int rc = schema_check(schema, &nw, &xm);
    if (rc) return rc;
    if (n_shards == 0 || n_shards > 64 || (n && (!keys_dev || !owner_dev)) || (n_shards > 1 && !splitters))
        return STAGE_ERR_ARG;
    if (n == 0) return STAGE_OK;
    RouteArgs r{};
    r.n_split = n_shards - 1;
    for (uint32_t i = 0; i + 1 < n_shards; ++i)
        host_sortable_nw(nw, static_cast<const uint8_t *>(splitters) + (size_t) i * schema->key_size, schema->key_size, xm,
                         r.split + (size_t) i * nw);
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const unsigned blocks = (unsigned) ((n + 255) / 256);
    const uint8_t *k = static_cast<const uint8_t *>(keys_dev);
    switch (nw) {
        case 1: route_kernel<1><<<blocks, 256, 0, st>>>(r, k, n, schema->key_size, xm, owner_dev); break;
        case 2: route_kernel<2><<<blocks, 256, 0, st>>>(r, k, n, schema->key_size, xm, owner_dev); break;
        case 4: route_kernel<4><<<blocks, 256, 0, st>>>(r, k, n, schema->key_size, xm, owner_dev); break;
        default: route_kernel<8><<<blocks, 256, 0, st>>>(r, k, n, schema->key_size, xm, owner_dev); break;
    }
    LAUNCH_CHECK();
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// routed probes over peer memory (exchange.cuh)
// ------------------------------------------------------------------------------------------------
struct stage_exchange {
    ExchangeView view{};
    stage_schema_t schema{};
    int nw = 1;
    uint32_t xormask = 0;
    int device = 0;
    uint8_t *arena = nullptr;
    bool opened[kMaxExchangeRanks] = {false};
    uint32_t *d_local_cnt = nullptr, *d_err = nullptr;
    uint32_t *d_src_row = nullptr, *d_dst_slot = nullptr;
    // scan exchange: plan outputs in inbox order (row counts, keys, commit ids, source rows)
    uint32_t *d_sc_counts = nullptr, *d_sc_ts = nullptr, *d_sc_src = nullptr;
    uint8_t *d_sc_keys = nullptr;
    uint32_t epoch_post = 0, epoch_serve = 0, epoch_wait = 0;
    bool attached = false;
    AggI32 *d_agg = nullptr;                     // [kMaxExchangeRanks] partials per requester + [1] combined result
    uint32_t col_stride = 0;                     // stride of the projected column in the result slab (last serve_project)
    // per-phase CUDA events: [0] before post, [1] after post, [2] after the serve search, [3] after the row gather,
    // [4] after the wait kernel
    cudaEvent_t tev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    bool timing = false;
};
#define XCHG_MARK(x, i, st) do { if ((x)->timing) CUDA_TRY(cudaEventRecord((x)->tev[i], st)); } while (0)

static uint64_t align_up(uint64_t v, uint64_t a) { return (v + a - 1) / a * a; }

static int exchange_create(uint32_t world, uint32_t rank, uint64_t capacity, uint32_t rows, const stage_schema_t *schema,
                           stage_exchange_t **out) {
    if (!out || world == 0 || world > (uint32_t) kMaxExchangeRanks || rank >= world || capacity == 0 || rows == 0 ||
        capacity * rows > kDstIndexMask)
        return STAGE_ERR_ARG;
    int nw;
    uint32_t xm;
    int rc = schema_check(schema, &nw, &xm);
    if (rc) return rc;
    if (rows > 1 && (schema->key_size & 3u)) return STAGE_ERR_ARG;   // scan inboxes feed the plan kernels directly
    CUDA_TRY(cudaSetDevice(g_device));
    auto *x = new stage_exchange();
    x->schema = *schema;
    x->nw = nw;
    x->xormask = xm;
    x->device = g_device;
    ExchangeLayout &l = x->view.lay;
    l.cap = (uint32_t) capacity;
    l.kpad = (schema->key_size + 3u) & ~3u;
    uint64_t off = 0;
    l.off_count = off;   off = align_up(off + 4ull * world, 256);
    l.off_posted = off;  off = align_up(off + 4ull * world, 256);
    l.off_served = off;  off = align_up(off + 4ull * world, 256);
    l.off_keys = off;    off = align_up(off + (uint64_t) world * capacity * l.kpad, 256);
    l.off_req = off;     off = align_up(off + (uint64_t) world * capacity * 4, 256);
    l.off_found = off;   off = align_up(off + capacity, 256);
    l.rows = rows;
    const bool scans = rows > 1;
    l.off_ts = off;      off = align_up(off + capacity * rows * 4, 256);
    l.pstride = (uint32_t) align_up(schema->payload_size, 128);
    l.off_payload = off; off = align_up(off + capacity * rows * (uint64_t) l.pstride + 256, 256);
    l.off_sizes = off;   off = align_up(off + (scans ? (uint64_t) world * capacity * 4 : 0), 256);
    l.off_counts = off;  off = align_up(off + (scans ? capacity * 4 : 0), 256);
    l.off_keyout = off;  off = align_up(off + (scans ? capacity * rows * (uint64_t) l.kpad : 0), 256);
    l.off_agg = off;     off = align_up(off + sizeof(AggI32) * kMaxExchangeRanks, 256);
    l.bytes = off;
    x->view.world = world;
    x->view.rank = rank;
    auto fail = [&](int code) { stage_gpu_exchange_destroy(x); return code; };
    if (cudaMalloc((void **) &x->arena, l.bytes) != cudaSuccess || cudaMemset(x->arena, 0, l.bytes) != cudaSuccess ||
        cudaMalloc((void **) &x->d_local_cnt, 4 * kMaxExchangeRanks) != cudaSuccess ||
        cudaMemset(x->d_local_cnt, 0, 4 * kMaxExchangeRanks) != cudaSuccess ||
        cudaMalloc((void **) &x->d_err, 4) != cudaSuccess || cudaMemset(x->d_err, 0, 4) != cudaSuccess ||
        cudaMalloc((void **) &x->d_agg, sizeof(AggI32) * (kMaxExchangeRanks + 1)) != cudaSuccess ||
        cudaMalloc((void **) &x->d_src_row, (uint64_t) world * capacity * rows * 4) != cudaSuccess ||
        cudaMalloc((void **) &x->d_dst_slot, (uint64_t) world * capacity * rows * 4) != cudaSuccess ||
        (scans && (cudaMalloc((void **) &x->d_sc_counts, (uint64_t) world * capacity * 4) != cudaSuccess ||
                   cudaMalloc((void **) &x->d_sc_ts, (uint64_t) world * capacity * rows * 4) != cudaSuccess ||
                   cudaMalloc((void **) &x->d_sc_src, (uint64_t) world * capacity * rows * 4) != cudaSuccess ||
                   cudaMalloc((void **) &x->d_sc_keys, (uint64_t) world * capacity * rows * l.kpad) != cudaSuccess))) {
        g_last_error = std::string("exchange allocation: ") + cudaGetErrorString(cudaGetLastError());
        return fail(STAGE_ERR_NOMEM);
    }
    {
        AggI32 init[kMaxExchangeRanks + 1];
        for (auto &a : init) a = AggI32{0ull, 0ull, 0x7FFFFFFF, (int) 0x80000000};
        CUDA_TRY(cudaMemcpy(x->d_agg, init, sizeof(init), cudaMemcpyHostToDevice));
    }
    CUDA_TRY(cudaDeviceSynchronize());
    x->view.arena[rank] = x->arena;
    x->view.local_cnt = x->d_local_cnt;
    x->view.err = x->d_err;
    x->attached = world == 1;
    *out = x;
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_create(uint32_t world, uint32_t rank, uint64_t capacity, const stage_schema_t *schema,
                                         stage_exchange_t **out) {
    return exchange_create(world, rank, capacity, 1, schema, out);
}
extern "C" int stage_gpu_exchange_create_scan(uint32_t world, uint32_t rank, uint64_t max_scans, uint32_t max_scan_size,
                                              const stage_schema_t *schema, stage_exchange_t **out) {
    if (max_scan_size < 2) max_scan_size = 2;                    // rows == 1 denotes the probe exchange
    return exchange_create(world, rank, max_scans, max_scan_size, schema, out);
}

extern "C" int stage_gpu_exchange_ipc_handle(stage_exchange_t *x, void *handle_out) {
    if (!x || !handle_out) return STAGE_ERR_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == STAGE_IPC_HANDLE_BYTES, "IPC handle size");
    CUDA_TRY(cudaSetDevice(x->device));
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, x->arena));
    memcpy(handle_out, &h, sizeof(h));
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_attach_ipc(stage_exchange_t *x, const void *handles) {
    if (!x || !handles) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    for (uint32_t r = 0; r < x->view.world; ++r) {
        if (r == x->view.rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const uint8_t *>(handles) + (size_t) r * STAGE_IPC_HANDLE_BYTES, sizeof(h));
        void *p = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        x->view.arena[r] = static_cast<uint8_t *>(p);
        x->opened[r] = true;
    }
    x->attached = true;
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_attach_ptrs(stage_exchange_t *x, void *const *arenas) {
    if (!x || !arenas) return STAGE_ERR_ARG;
    for (uint32_t r = 0; r < x->view.world; ++r)
        if (r != x->view.rank) {
            if (!arenas[r]) return STAGE_ERR_ARG;
            x->view.arena[r] = static_cast<uint8_t *>(arenas[r]);
        }
    x->attached = true;
    return STAGE_OK;
}

extern "C" void *stage_gpu_exchange_arena(stage_exchange_t *x) { return x ? x->arena : nullptr; }

static int exchange_post(stage_exchange_t *x, const void *splitters, const void *keys_dev, const uint32_t *sizes_dev,
                         uint64_t n, int flags, void *cuda_stream) {
    if (!x || (n && !keys_dev) || (x->view.world > 1 && !splitters) || n > x->view.lay.cap) return STAGE_ERR_ARG;
    if (!x->attached) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    ExchangeSplit sp{};
    XCHG_MARK(x, 0, st);
    for (uint32_t i = 0; i + 1 < x->view.world; ++i)
        host_sortable_nw(x->nw, static_cast<const uint8_t *>(splitters) + (size_t) i * x->schema.key_size,
                         x->schema.key_size, x->xormask, sp.w + (size_t) i * x->nw);
    if (n) {
        // few CTAs on purpose: inbox slots are claimed tile by tile, so the fewer tiles are in flight at once, the
        // closer the inbox order follows the request order (and the narrower the window of the requester's slab
        // that the owner's concurrent CTAs write to)
        static const int post_ctas = getenv("STAGE_XCHG_POST_CTAS") ? atoi(getenv("STAGE_XCHG_POST_CTAS")) : sm_count();
        const unsigned blocks = (unsigned) std::min<uint64_t>((n + 255) / 256, post_ctas > 0 ? (uint64_t) post_ctas : (uint64_t) sm_count() * 8);
        const uint8_t *k = static_cast<const uint8_t *>(keys_dev);
        switch (x->nw) {
            case 1: exchange_post_kernel<1><<<blocks, 256, 0, st>>>(x->view, sp, k, n, x->schema.key_size, x->xormask, sizes_dev); break;
            case 2: exchange_post_kernel<2><<<blocks, 256, 0, st>>>(x->view, sp, k, n, x->schema.key_size, x->xormask, sizes_dev); break;
            case 4: exchange_post_kernel<4><<<blocks, 256, 0, st>>>(x->view, sp, k, n, x->schema.key_size, x->xormask, sizes_dev); break;
            default: exchange_post_kernel<8><<<blocks, 256, 0, st>>>(x->view, sp, k, n, x->schema.key_size, x->xormask, sizes_dev); break;
        }
        LAUNCH_CHECK();
    }
    ++x->epoch_post;
    exchange_post_finish_kernel<<<1, 32, 0, st>>>(x->view, x->epoch_post, flags & STAGE_XCHG_DEVICE_SYNC ? 1 : 0);
    LAUNCH_CHECK();
    XCHG_MARK(x, 1, st);
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_post(stage_exchange_t *x, const void *splitters, const void *keys_dev, uint64_t n,
                                       int flags, void *cuda_stream) {
    if (x && x->view.lay.rows != 1) return STAGE_ERR_STATE;
    return exchange_post(x, splitters, keys_dev, nullptr, n, flags, cuda_stream);
}
extern "C" int stage_gpu_exchange_post_scans(stage_exchange_t *x, const void *splitters, const void *start_keys_dev,
                                             const uint32_t *scan_sizes_dev, uint64_t n, int flags, void *cuda_stream) {
    if (x && x->view.lay.rows == 1) return STAGE_ERR_STATE;
    if (n && !scan_sizes_dev) return STAGE_ERR_ARG;
    return exchange_post(x, splitters, start_keys_dev, scan_sizes_dev, n, flags, cuda_stream);
}

template <int NW>
static int launch_serve(stage_exchange *x, stage_snapshot *s, const ServeArgs &a, cudaStream_t st) {
    // persistent grid: the inbox sizes are only known on the device
    if (s->view.dir) {
        static int occd[9] = {0};
        if (!occd[NW]) {
            int o = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, exchange_serve_kernel<NW, 4, true>, kProbeThreads, 0));
            occd[NW] = std::max(o, 1);
        }
        exchange_serve_kernel<NW, 4, true><<<sm_count() * occd[NW], kProbeThreads, 0, st>>>(s->view, x->view, a);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    const size_t smem = (size_t) std::max<uint32_t>(s->view.smem_words, 4) * 4;
    static int occ[9] = {0};
    if (!occ[NW]) {
        int o = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, exchange_serve_kernel<NW, 4>, kProbeThreads, kProbeTopBytes));
        occ[NW] = std::max(o, 1);
    }
    exchange_serve_kernel<NW, 4><<<sm_count() * occ[NW], kProbeThreads, smem, st>>>(s->view, x->view, a);
    LAUNCH_CHECK();
    return STAGE_OK;
}

template <typename V>
static int launch_exchange_gather(stage_exchange *x, stage_snapshot *s, const ServeArgs &a, cudaStream_t st) {
    static int occ = 0;
    if (!occ) {
        int o = 0;
        CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, exchange_gather_kernel<V, 4>, kGatherThreads, 0));
        occ = std::max(o, 1);
    }
    exchange_gather_kernel<V, 4><<<sm_count() * occ, kGatherThreads, 0, st>>>(x->view, a, s->d_payload, s->d_ver_payload,
                                                                             s->schema.payload_size);
    LAUNCH_CHECK();
    return STAGE_OK;
}

// payload rows of the planned (source row, destination slot) lists -> the requesters' slabs
static int exchange_move_rows(stage_exchange *x, stage_snapshot *s, const ServeArgs &a, cudaStream_t st) {
    int rc = STAGE_OK;
    // vector width: every arena and the payload offset are 256-byte aligned, so the row stride decides
    const uintptr_t align = (uintptr_t) s->d_payload | (uintptr_t) s->d_ver_payload | s->schema.payload_size;
    static const bool no16 = getenv("STAGE_XCHG_NO16") != nullptr;
    const char *tma = getenv("STAGE_XCHG_TMA");                 // read per call: tools/xchg_sweep.py flips it between runs
    const char *cps = getenv("STAGE_XCHG_CTAS_PER_SM");
    // 0 = as many as fit. From four ranks on, two CTAs per SM keep a narrower window of every requester's slab in flight
    // and were 2-4 % faster in the row phase (profiles/r2_xchg_sweep_n8.txt); at two ranks full occupancy is (slightly) better
    const int ctas_per_sm = cps ? std::max(1, atoi(cps)) : (x->view.world >= 4 ? 2 : 0);
    if (tma && atoi(tma) != 0 && (align & 7) == 0 && s->schema.payload_size <= 1016 && x->view.lay.pstride >= ((s->schema.payload_size + 15u) & ~15u)) {
        const size_t smem = (size_t) (kGatherThreads / 32) * kTmaGroups * kTmaRows * kTmaRowBytes;
        static int occt = 0;
        if (!occt) {
            CUDA_TRY(cudaFuncSetAttribute(exchange_gather_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
            int o = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, exchange_gather_tma_kernel, kGatherThreads, smem));
            occt = std::max(o, 1);
        }
        const int per_sm = ctas_per_sm ? std::min(ctas_per_sm, occt) : occt;
        exchange_gather_tma_kernel<<<sm_count() * per_sm, kGatherThreads, smem, st>>>(x->view, a, s->d_payload, s->d_ver_payload,
                                                                                   s->schema.payload_size);
        LAUNCH_CHECK();
        return STAGE_OK;
    }
    if ((align & 15) == 0) rc = launch_exchange_gather<uint4>(x, s, a, st);
    else if ((align & 7) == 0 && !no16) {
        static int occ16 = 0;
        if (!occ16) {
            int o = 0;
            CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, exchange_gather16_kernel<4>, kGatherThreads, 0));
            occ16 = std::max(o, 1);
        }
        exchange_gather16_kernel<4><<<sm_count() * (ctas_per_sm ? std::min(ctas_per_sm, occ16) : occ16), kGatherThreads, 0, st>>>(
            x->view, a, s->d_payload, s->d_ver_payload, s->schema.payload_size);
        LAUNCH_CHECK();
    }
    else if ((align & 7) == 0) rc = launch_exchange_gather<uint2>(x, s, a, st);
    else if ((align & 3) == 0) rc = launch_exchange_gather<uint32_t>(x, s, a, st);
    else rc = launch_exchange_gather<uint8_t>(x, s, a, st);
    return rc;
}

// what the owner sends back for every probe of its inboxes: the payload row, one column of it, or nothing per probe
// (one aggregate record per requester)
enum { kServeRows = 0, kServeProject = 1, kServeAggregate = 2 };

static int exchange_serve_common(stage_exchange_t *x, stage_snapshot_t *s, uint32_t read_ts, int flags, void *cuda_stream, int mode,
                                 uint32_t col_off, uint32_t col_len, int32_t threshold) {
    if (!x || !s) return STAGE_ERR_ARG;
    if (!x->attached || !s->sealed) return STAGE_ERR_STATE;
    if (s->schema.key_size != x->schema.key_size || s->schema.payload_size != x->schema.payload_size || s->nw != x->nw ||
        s->device != x->device)
        return STAGE_ERR_ARG;
    if (mode != kServeRows && (col_len == 0 || (uint64_t) col_off + col_len > s->schema.payload_size)) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    ServeArgs a{};
    a.read_ts = read_ts;
    a.epoch = ++x->epoch_serve;
    a.device_sync = flags & STAGE_XCHG_DEVICE_SYNC ? 1 : 0;
    a.src_row = x->d_src_row;
    a.dst_slot = x->d_dst_slot;
    a.seg_mul = 1;
    if (x->view.lay.rows != 1) return STAGE_ERR_STATE;
    a.tile_order = getenv("STAGE_XCHG_ORDER") ? atoi(getenv("STAGE_XCHG_ORDER")) : 1;             // 1 = shift schedule (exchange.cuh)
    int rc;
    switch (x->nw) {
        case 1: rc = launch_serve<1>(x, s, a, st); break;
        case 2: rc = launch_serve<2>(x, s, a, st); break;
        case 4: rc = launch_serve<4>(x, s, a, st); break;
        default: rc = launch_serve<8>(x, s, a, st); break;
    }
    if (rc) return rc;
    XCHG_MARK(x, 2, st);
    if (mode == kServeRows) {
        if ((rc = exchange_move_rows(x, s, a, st))) return rc;
    } else if (mode == kServeProject) {
        x->col_stride = (col_len + 3u) & ~3u;
        exchange_cols_kernel<<<sm_count() * 8, 256, 0, st>>>(x->view, a, s->d_payload, s->d_ver_payload, s->schema.payload_size, col_off,
                                                            col_len, x->col_stride);
        LAUNCH_CHECK();
    } else {
        exchange_agg_kernel<<<dim3((unsigned) sm_count(), x->view.world), 256, 0, st>>>(x->view, a, s->d_payload, s->d_ver_payload,
                                                                                        s->schema.payload_size, col_off, threshold, x->d_agg);
        LAUNCH_CHECK();
        exchange_agg_publish_kernel<<<1, 32, 0, st>>>(x->view, x->d_agg);
        LAUNCH_CHECK();
    }
    if (a.device_sync) {
        exchange_serve_finish_kernel<<<1, 32, 0, st>>>(x->view, a.epoch);
        LAUNCH_CHECK();
    }
    XCHG_MARK(x, 3, st);
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_serve(stage_exchange_t *x, stage_snapshot_t *s, uint32_t read_ts, int flags, void *cuda_stream) {
    return exchange_serve_common(x, s, read_ts, flags, cuda_stream, kServeRows, 0, 0, 0);
}
extern "C" int stage_gpu_exchange_serve_project(stage_exchange_t *x, stage_snapshot_t *s, uint32_t read_ts, uint32_t col_off,
                                                uint32_t col_len, int flags, void *cuda_stream) {
    return exchange_serve_common(x, s, read_ts, flags, cuda_stream, kServeProject, col_off, col_len, 0);
}
extern "C" int stage_gpu_exchange_serve_aggregate_i32(stage_exchange_t *x, stage_snapshot_t *s, uint32_t read_ts, uint32_t col_off,
                                                      int32_t threshold, int flags, void *cuda_stream) {
    return exchange_serve_common(x, s, read_ts, flags, cuda_stream, kServeAggregate, col_off, 4, threshold);
}
extern "C" int stage_gpu_exchange_project_results(stage_exchange_t *x, void **found, void **ts, void **col, uint32_t *col_stride) {
    if (!x) return STAGE_ERR_ARG;
    if (x->view.lay.rows != 1 || !x->col_stride) return STAGE_ERR_STATE;
    if (found) *found = x->arena + x->view.lay.off_found;
    if (ts) *ts = x->arena + x->view.lay.off_ts;
    if (col) *col = x->arena + x->view.lay.off_payload;
    if (col_stride) *col_stride = x->col_stride;
    return STAGE_OK;
}
extern "C" int stage_gpu_exchange_aggregate_result(stage_exchange_t *x, stage_agg_i32_t *out, void *cuda_stream) {
    if (!x || !out) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    exchange_agg_combine_kernel<<<1, 32, 0, st>>>(x->view, x->d_agg + kMaxExchangeRanks);
    LAUNCH_CHECK();
    CUDA_TRY(cudaMemcpyAsync(out, x->d_agg + kMaxExchangeRanks, sizeof(AggI32), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return STAGE_OK;
}

// Scans posted with stage_gpu_exchange_post_scans: plan every source rank's inbox with the ordinary scan plan
// kernels, deliver counts / keys / commit ids, then move the payload rows with the same kernels as the probes.
extern "C" int stage_gpu_exchange_serve_scans(stage_exchange_t *x, stage_snapshot_t *s, uint32_t read_ts,
                                              int apply_visibility, int flags, void *cuda_stream) {
    if (!x || !s) return STAGE_ERR_ARG;
    if (!x->attached || !s->sealed || x->view.lay.rows == 1) return STAGE_ERR_STATE;
    if (s->schema.key_size != x->schema.key_size || s->schema.payload_size != x->schema.payload_size || s->nw != x->nw ||
        s->device != x->device)
        return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    const ExchangeLayout &l = x->view.lay;
    ServeArgs a{};
    a.read_ts = read_ts;
    a.epoch = ++x->epoch_serve;
    a.device_sync = flags & STAGE_XCHG_DEVICE_SYNC ? 1 : 0;
    a.src_row = x->d_src_row;
    a.dst_slot = x->d_dst_slot;
    a.seg_mul = l.rows;
    a.tile_order = 1;
    if (a.device_sync) {
        exchange_wait_posted_kernel<<<1, 32, 0, st>>>(x->view, a.epoch);
        LAUNCH_CHECK();
    }
    int rc = STAGE_OK;
    for (uint32_t src = 0; src < x->view.world && !rc; ++src) {
        const uint64_t slot0 = (uint64_t) src * l.cap;
        ScanArgs sa{};
        sa.start_keys = x->arena + l.off_keys + slot0 * l.kpad;
        sa.scan_sizes = reinterpret_cast<const uint32_t *>(x->arena + l.off_sizes) + slot0;
        sa.n = l.cap;
        sa.n_dev = reinterpret_cast<const uint32_t *>(x->arena + l.off_count) + src;
        sa.out_stride = l.rows;
        sa.read_ts = read_ts;
        sa.apply_visibility = apply_visibility;
        sa.row_counts = x->d_sc_counts + slot0;
        sa.key_out = x->d_sc_keys + slot0 * l.rows * l.kpad;
        sa.ts_out = x->d_sc_ts + slot0 * l.rows;
        sa.src_row = x->d_sc_src + slot0 * l.rows;
        sa.stats = s->d_stats;
        rc = launch_scan_nw(s, sa, st);
    }
    if (rc) return rc;
    ScanDeliverArgs d{};
    d.row_counts = x->d_sc_counts;
    d.keys = x->d_sc_keys;
    d.ts = x->d_sc_ts;
    d.src_row = x->d_sc_src;
    d.src_row_c = x->d_src_row;
    d.dst_slot_c = x->d_dst_slot;
    d.key_size = x->schema.key_size;
    exchange_scan_deliver_kernel<<<sm_count() * 8, 256, 0, st>>>(x->view, d);
    LAUNCH_CHECK();
    XCHG_MARK(x, 2, st);
    if ((rc = exchange_move_rows(x, s, a, st))) return rc;
    if (a.device_sync) {
        exchange_serve_finish_kernel<<<1, 32, 0, st>>>(x->view, a.epoch);
        LAUNCH_CHECK();
    }
    XCHG_MARK(x, 3, st);
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_scan_results(stage_exchange_t *x, void **row_counts, void **key_out, void **ts,
                                               void **payload, uint32_t *rows_per_scan, uint32_t *key_stride,
                                               uint32_t *payload_stride) {
    if (!x) return STAGE_ERR_ARG;
    if (x->view.lay.rows == 1) return STAGE_ERR_STATE;
    if (rows_per_scan) *rows_per_scan = x->view.lay.rows;
    if (key_stride) *key_stride = x->view.lay.kpad;
    if (payload_stride) *payload_stride = x->view.lay.pstride;
    if (row_counts) *row_counts = x->arena + x->view.lay.off_counts;
    if (key_out) *key_out = x->arena + x->view.lay.off_keyout;
    if (ts) *ts = x->arena + x->view.lay.off_ts;
    if (payload) *payload = x->arena + x->view.lay.off_payload;
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_wait(stage_exchange_t *x, void *cuda_stream) {
    if (!x) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    exchange_wait_served_kernel<<<1, 32, 0, static_cast<cudaStream_t>(cuda_stream)>>>(x->view, ++x->epoch_wait);
    LAUNCH_CHECK();
    XCHG_MARK(x, 4, static_cast<cudaStream_t>(cuda_stream));
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_results(stage_exchange_t *x, void **found, void **ts, void **payload,
                                          uint32_t *payload_stride) {
    if (!x) return STAGE_ERR_ARG;
    if (payload_stride) *payload_stride = x->view.lay.pstride;
    if (found) *found = x->arena + x->view.lay.off_found;
    if (ts) *ts = x->arena + x->view.lay.off_ts;
    if (payload) *payload = x->arena + x->view.lay.off_payload;
    return STAGE_OK;
}

// Measurement helper: ONE cudaMemcpyAsync of `bytes` from this rank's result slab into rank dst_rank's result slab (peer
// memory over NVLink, copy engine). bench.py / tools/xchg_sweep.py time it on all ranks at once to get the link ceiling
// for the exchange's traffic pattern on the box at hand; the slab contents are garbage afterwards (until the next batch).
extern "C" int stage_gpu_exchange_peer_copy(stage_exchange_t *x, uint32_t dst_rank, uint64_t dst_offset, uint64_t bytes, void *cuda_stream) {
    if (!x || dst_rank >= x->view.world || !x->attached) return STAGE_ERR_ARG;
    const uint64_t cap_bytes = (uint64_t) x->view.lay.cap * x->view.lay.rows * x->view.lay.pstride;
    if (dst_offset + bytes > cap_bytes || bytes > cap_bytes) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    CUDA_TRY(cudaMemcpyAsync(x->view.arena[dst_rank] + x->view.lay.off_payload + dst_offset, x->arena + x->view.lay.off_payload, bytes,
                             cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(cuda_stream)));
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_error(stage_exchange_t *x, void *cuda_stream, uint32_t *err_out) {
    if (!x || !err_out) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
    CUDA_TRY(cudaMemcpyAsync(err_out, x->d_err, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_enable_timing(stage_exchange_t *x, int on) {
    if (!x) return STAGE_ERR_ARG;
    CUDA_TRY(cudaSetDevice(x->device));
    if (on && !x->tev[0])
        for (int i = 0; i < 5; ++i) CUDA_TRY(cudaEventCreate(&x->tev[i]));
    x->timing = on != 0;
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_last_timing(stage_exchange_t *x, float *ms4) {
    if (!x || !ms4 || !x->timing) return STAGE_ERR_STATE;
    CUDA_TRY(cudaEventSynchronize(x->tev[4]));
    for (int i = 0; i < 4; ++i) CUDA_TRY(cudaEventElapsedTime(ms4 + i, x->tev[i], x->tev[i + 1]));
    return STAGE_OK;
}

extern "C" int stage_gpu_exchange_destroy(stage_exchange_t *x) {
    if (!x) return STAGE_ERR_ARG;
    cudaSetDevice(x->device);
    cudaDeviceSynchronize();
    for (int i = 0; i < 5; ++i)
        if (x->tev[i]) cudaEventDestroy(x->tev[i]);
    for (uint32_t r = 0; r < x->view.world; ++r)
        if (x->opened[r] && x->view.arena[r]) cudaIpcCloseMemHandle(x->view.arena[r]);
    void *ptrs[] = {x->arena, x->d_local_cnt, x->d_err, x->d_src_row, x->d_dst_slot, x->d_agg,
                    x->d_sc_counts, x->d_sc_ts, x->d_sc_src, x->d_sc_keys};
    for (void *p : ptrs)
        if (p) cudaFree(p);
    delete x;
    return STAGE_OK;
}

// ------------------------------------------------------------------------------------------------
// utilities
// ------------------------------------------------------------------------------------------------
extern "C" int stage_gpu_snapshot_counters(stage_snapshot_t *s, stage_counters_t *out, int reset) {
    if (!s || !out) return STAGE_ERR_ARG;
    if (!s->d_stats) return STAGE_ERR_STATE;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaDeviceSynchronize());
    unsigned long long h[8] = {0};
    CUDA_TRY(cudaMemcpy(h, s->d_stats, sizeof(h), cudaMemcpyDeviceToHost));
    out->slot_order_scans = h[0];
    if (reset) CUDA_TRY(cudaMemset(s->d_stats, 0, sizeof(h)));
    return STAGE_OK;
}

extern "C" int stage_gpu_host_register(void *p, size_t bytes) {
    if (!p || !bytes) return STAGE_ERR_ARG;
    CUDA_TRY(cudaHostRegister(p, bytes, cudaHostRegisterDefault));
    return STAGE_OK;
}
extern "C" int stage_gpu_host_unregister(void *p) {
    if (!p) return STAGE_ERR_ARG;
    CUDA_TRY(cudaHostUnregister(p));
    return STAGE_OK;
}
extern "C" void *stage_gpu_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
    return p;
}
extern "C" void stage_gpu_host_free(void *p) {
    if (p) cudaFreeHost(p);
}
extern "C" uint64_t stage_gpu_launch_count(void) { return g_launches.load(); }
extern "C" const char *stage_gpu_strerror(int status) {
    switch (status) {
        case STAGE_OK: return "ok";
        case STAGE_ERR_ARG: return "bad argument or unsupported schema";
        case STAGE_ERR_CUDA: return "CUDA error";
        case STAGE_ERR_NOT_QUIESCED: return "tree not quiesced (frozen node or in-flight record)";
        case STAGE_ERR_STATE: return "call order";
        case STAGE_ERR_NOMEM: return "out of memory";
        case STAGE_ERR_LAYOUT: return "node bytes are not a Stage BzTree node";
        default: return "unknown";
    }
}
extern "C" const char *stage_gpu_last_error(void) { return g_last_error.c_str(); }
