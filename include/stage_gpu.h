/* stage_gpu.h — C ABI of libstage_gpu.so: the B200-native index-organized read path for Stage.
 *
 * This is the drop-in boundary. The reference has no FFI layer: its read interface
 * is in-process C++ (paths relative to the reference repository):
 *
 *   std::unique_ptr<Record>   BTree::Read(const char *key, uint16_t key_size, cid_t commit_id,
 *                                         bool is_for_update)            include/vstore/b_tree.h:812-813
 *   std::unique_ptr<Iterator> BTree::RangeScanBySize(const char *key1, uint16_t size1,
 *                                         uint32_t scan_size)            include/vstore/b_tree.h:830-835
 *   std::unique_ptr<Record>   Iterator::GetNext()                        include/vstore/b_tree.h:899-941
 *   IndexScanExecutor<Key,T>::Execute / GetResults / GetRecord           include/execute/executor.h:342-570
 *   TableScanExecutor<Key,T>::Execute / GetResults / GetFailNum          include/execute/executor.h:572-661
 *
 * The entry points below are what a cgo-style binding of that interface would bind: plain
 * pointers and sizes, no C++ types, no exceptions, caller-allocated outputs. The header-only C++
 * adapters in include/stage_gpu_executors.h put the reference's own class shapes back on top.
 *
 * All functions return 0 on success or a negative stage_status code; stage_gpu_strerror() names it.
 * A snapshot is immutable once sealed. Threading (the reference's callers are N pinned worker threads sharing one
 * BTree, benchmark/ycsb/ycsb_workload.cpp:522-529): any number of threads may call the *_batch reads on one snapshot
 * at once. STAGE_IO_HOST calls each run on a context of their own (streams + scratch, up to 16 in flight, then
 * callers queue); STAGE_IO_DEVICE calls are enqueued one after the other on the snapshot's stream and return
 * without waiting for the GPU.
 *
 * Limits: keys are fixed-length per table, 1..32 bytes; a snapshot (one range shard) holds fewer than 2^31 rows and
 * fewer than 2^31 staged versions (source rows travel as 32-bit indices whose top bit selects the version table);
 * leaf_node_size <= 65536; the routed-scan exchange additionally needs key_size % 4 == 0.
 */
#ifndef STAGE_GPU_H_
#define STAGE_GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct stage_snapshot stage_snapshot_t;

enum stage_status {
    STAGE_OK = 0,
    STAGE_ERR_ARG = -1,          /* bad argument / unsupported schema */
    STAGE_ERR_CUDA = -2,         /* CUDA runtime error (text via stage_gpu_last_error) */
    STAGE_ERR_NOT_QUIESCED = -3, /* tree has a frozen node or an in-flight (IsInserting) record */
    STAGE_ERR_STATE = -4,        /* call order (e.g. probe before seal) */
    STAGE_ERR_NOMEM = -5,
    STAGE_ERR_LAYOUT = -6        /* node bytes do not look like a Stage BzTree node */
};

/* key_kind: how BaseNode::KeyCompare orders keys of this table (include/vstore/b_tree.h:116-134) */
enum stage_key_kind {
    STAGE_KEY_AUTO = 0,         /* key_size < 16 -> signed bytes, else memcmp */
    STAGE_KEY_SIGNED_BYTES = 1, /* my_memcmp: signed char difference, byte 0 first (b_tree.h:99-106) */
    STAGE_KEY_MEMCMP = 2,       /* libc memcmp: unsigned bytes */
    STAGE_KEY_VARLEN = 0x100    /* flag, OR-ed in: keys of different lengths in one table (see stage_gpu_key_pack) */
};

/* Table schema = the part of ParameterSet (include/vstore/b_tree.h:23-40) and Catalog
 * (include/common/catalog.h:69) the read path depends on. Keys are fixed-length per table, as in
 * every driver table (YCSB int 4 B, YCSB string 16 B, TPC-C 8/16/32 B), unless key_kind carries
 * STAGE_KEY_VARLEN (below). */
typedef struct stage_schema {
    uint32_t key_size;       /* bytes of key actually passed to Read/Insert (YCSB int: 4); 1..32 */
    uint32_t payload_size;   /* ParameterSet.payload_size */
    uint32_t leaf_node_size; /* ParameterSet.leaf_node_size (64 KiB in both drivers) */
    uint32_t key_kind;       /* enum stage_key_kind */
} stage_schema_t;

/* Variable-length keys (key_kind | STAGE_KEY_VARLEN). BaseNode::KeyCompare orders two keys by their first
 * min(l1, l2) bytes and then by length (include/vstore/b_tree.h:116-134) — the order test/testing_btree.cpp's
 * std::to_string keys rely on. Such a table is described by key_size = SLOT size = longest key + 1 (<= 32), and every key
 * crosses this interface (probe keys, scan start keys, separators, key_out) as a slot of key_size bytes:
 *   bytes [0, len) the key, bytes [len, key_size - 1) padding that sorts lowest (0x80 for the signed-byte order, 0x00
 *   for memcmp), last byte the length (XOR 0x80 for the signed-byte order)
 * so that comparing two slots as fixed-length keys of the table's byte order IS KeyCompare on the keys (a proper prefix
 * sorts first, keys that differ only by trailing minimum bytes are decided by the length byte). stage_gpu_key_pack /
 * stage_gpu_key_unpack convert; the adapters in stage_gpu_executors.h do it for their callers. Because KeyCompare picks the
 * byte order per pair (signed bytes when the shorter key has < 16 bytes), a table has ONE order only if its keys are
 * all <= 15 bytes (signed; what STAGE_KEY_AUTO picks for key_size <= 16) or all >= 16 bytes (memcmp): other tables are
 * rejected (schema: STAGE_ERR_ARG, a record outside the range: STAGE_ERR_LAYOUT). Quiesced trees without version
 * chains only (from_tree, refresh_from_tree, begin / add_leaves / seal): the version store keeps [key][payload] slots of
 * the catalog's fixed key length (version_store.h:445-481), no variable-length table goes through it. */
int stage_gpu_key_pack(const stage_schema_t *schema, const void *key, uint32_t len, void *slot);
int stage_gpu_key_unpack(const stage_schema_t *schema, const void *slot, void *key_out, uint32_t *len_out);

/* One retired version = one TupleHeader (include/vstore/version_store.h:28-46) in index form. */
typedef struct stage_version {
    uint32_t begin_id;  /* TupleHeader.begin_id_  */
    uint32_t commit_id; /* TupleHeader.comm_id_   */
    uint32_t next;      /* 1-based index of the next older version in this snapshot, 0 = end of chain */
    uint32_t reserved;
} stage_version_t;

typedef struct stage_snapshot_info {
    uint64_t n_rows;        /* visible records flattened */
    uint64_t n_leaves;
    uint64_t n_versions;
    uint64_t device_bytes;  /* HBM held by the snapshot */
    uint32_t key_words;     /* 32-bit words per sortable key */
    uint32_t n_levels;      /* sampled search levels above the key column */
    uint32_t max_leaf_rows;
    uint32_t device;        /* CUDA device ordinal */
} stage_snapshot_info_t;

/* flags for *_batch */
enum {
    STAGE_IO_HOST = 0,   /* all in/out pointers are host memory (pinned or pageable) */
    STAGE_IO_DEVICE = 1, /* all in/out pointers are device memory on the snapshot's GPU */
    STAGE_IO_COMPACT = 2 /* stage_gpu_scan_batch with host buffers: pack the result rows (see there) */
};
/* STAGE_IO_HOST results are produced into double-buffered HBM scratch and copied out by the DMA engine (one large
 * copy per output array and chunk) while the next chunk is planned and gathered; pinned buffers (stage_gpu_host_alloc,
 * stage_gpu_host_register) make those copies run at link speed. STAGE_HOST_PATH=zerocopy lets the kernels store into
 * pinned outputs directly instead (lower latency for small batches, lower bandwidth for large ones). Calls whose keys
 * and results fit 192 KiB (a handful of operations: what one executor of the reference issues) skip the staging: the
 * kernels work on a pinned slab of the call context through its device alias — memcpy in, two launches, one stream
 * synchronisation, memcpy out. */

/* ---- lifecycle --------------------------------------------------------------------------- */

/* Bind the calling process to a CUDA device (one process per GPU). Optional; default device 0. */
int stage_gpu_init(int device);

/* Flatten a live, quiesced reference tree. `root` is BTree::GetRootNodeSafe() (b_tree.h:872-875).
 * Walks inner nodes in tree order (the order TableScanExecutor::ScanInnerNode uses,
 * executor.h:592-609), stages the leaf images and the TupleHeader chains hanging off
 * RecordMetadata.next_ptr, and runs the record-layout transform kernel. Replaces the pointer
 * chasing of BTree::TraverseToLeaf (src/vstore/b_tree.cpp:1804-1846) for all later reads. */
int stage_gpu_snapshot_from_tree(const void *root, const stage_schema_t *schema, stage_snapshot_t **out);

/* Snapshot of a tree with writers in flight. A record whose update has not committed carries the
 * control bit and next_ptr = the location of its undo copy in the EphemeralPool (LeafNode::Update,
 * src/vstore/b_tree.cpp:1122-1159). BTree::Read (b_tree.cpp:2087-2113) hands readers that copy — old payload, the
 * unchanged commit id, and the version chain saved in the copy's header — and the snapshot stages exactly that, so
 * probes answer as the reference answers them while the writer is in flight (the uncommitted bytes never reach the
 * device). `resolve` is called on the calling thread for every in-flight record with RecordMetadata.next_ptr and
 * returns 1 after filling *next_tuple_header (EphemeralPool::GetNext) and *old_record (the copy: [key_len bytes of
 * key][payload]), or 0 when there is no copy (in-flight insert; copy already released): BTree::Read misses then and
 * the row is left out. Range scans return the same committed rows; the reference's iterator reads the uncommitted
 * in-place bytes there (LeafNode::RangeScanBySize takes no such care, b_tree.cpp:1276-1302), which is not reproduced.
 * Node structure must still be stable during the call (no split in progress: frozen nodes are refused). */
typedef int (*stage_undo_resolver_t)(void *ctx, uint64_t copy_location, uint64_t *next_tuple_header,
                                     const void **old_record);
int stage_gpu_snapshot_from_tree_live(const void *root, const stage_schema_t *schema, stage_undo_resolver_t resolve,
                                      void *resolve_ctx, stage_snapshot_t **out);

/* Structural refresh: a NEW snapshot of the same host tree after committed inserts, updates and
 * deletes — including the leaf splits / consolidations they caused (src/vstore/b_tree.cpp:1558-1690) — at the cost
 * of the leaves that changed. `old` must come from stage_gpu_snapshot_from_tree or a previous refresh of this tree
 * (it carries every leaf's host address and a fingerprint of its header, record metadata and upper fence). Leaves
 * found unchanged are carried over device-to-device; changed and new leaves are staged and run through the
 * record-layout transform again; the previous version table is kept as a prefix and the chains of the changed
 * leaves are appended. The host tree is the truth: device-only changes made to `old` with
 * stage_gpu_snapshot_apply_updates survive only in leaves the host tree did not touch. `old` stays valid and
 * unchanged (readers keep using it until the caller swaps and destroys it); both tables are resident meanwhile.
 * The tree must be quiesced during the call, as for stage_gpu_snapshot_from_tree. */
typedef struct {
    uint64_t leaves_total, leaves_carried, leaves_transformed;
    uint64_t rows, versions_added;
    uint64_t host_bytes_staged;     /* leaf images + version rows that crossed PCIe */
} stage_refresh_stats_t;
int stage_gpu_snapshot_refresh_from_tree(stage_snapshot_t *old, const void *root, stage_snapshot_t **out,
                                         stage_refresh_stats_t *stats /* may be NULL */);
/* Same, with writers in flight: in-flight records of the changed leaves are staged from their undo copies
 * (stage_gpu_snapshot_from_tree_live above). Together: a snapshot kept current without ever stopping the update
 * transactions — only structural changes (splits) must not be running during the call. */
int stage_gpu_snapshot_refresh_from_tree_live(stage_snapshot_t *old, const void *root, stage_undo_resolver_t resolve,
                                              void *resolve_ctx, stage_snapshot_t **out,
                                              stage_refresh_stats_t *stats /* may be NULL */);

/* The same refresh for a shard that fills its GPU (two copies of the table do not fit): *snap is REPLACED. The
 * payload rows of the unchanged leaves are shifted inside the existing buffer to their new offsets (runs of leaves, a
 * bounce buffer of 64 MiB where source and destination overlap), the rows of the changed leaves are transformed into
 * the gaps; only the small columns (keys, slots, commit ids, chain heads: 14-42 B per row), the search levels and the
 * directory are built anew, and the version table is kept as it is unless chains have to be appended. Needs room in
 * the buffer for the new row count: stage_gpu_set_row_slack(fraction) BEFORE the snapshot is built keeps that fraction
 * of its rows free (STAGE_ERR_NOMEM here otherwise, *snap untouched and still usable). No read may be in flight on
 * *snap during the call and every handle derived from it (batcher) must be gone; on any other error after the rows
 * started moving the snapshot is unusable (reads return STAGE_ERR_STATE) and must be destroyed. Quiesced trees.
 * No counterpart in the reference (its tree is updated in place by the writers, b_tree.cpp:1558-1690 for splits). */
int stage_gpu_snapshot_refresh_in_place(stage_snapshot_t **snap, const void *root, stage_refresh_stats_t *stats /* may be NULL */);
int stage_gpu_set_row_slack(double fraction);

/* Both refresh forms stage the chains of every re-transformed leaf again (the version table is append-only), so the
 * copies the previous snapshot held for those leaves stay behind unreferenced. This drops every version no chain head
 * reaches any more and renumbers the rest (mark from the chain column, prefix sum, move); reads give the same answers
 * before and after. No read may be in flight during the call. kept / dropped may be NULL. (The reference reclaims
 * retired versions through VersionStore::CollectRetiredVersionSet / CleanCandidateSlot, version_store.h:629-640; this
 * is the same duty for the device-side copy of the chains.) */
int stage_gpu_snapshot_compact_versions(stage_snapshot_t *s, uint64_t *kept, uint64_t *dropped);

/* Streaming construction for callers that own the walk (same result as from_tree):
 *   begin -> add_versions* -> add_leaves* -> seal.
 * leaf_images[i] points at leaf_node_size bytes of an unmodified LeafNode, leaves in tree order.
 * upper_fences holds n keys of key_size bytes: the separator the parent stores to the RIGHT of
 * each leaf (ignored for the last leaf of the tree; pass has_upper_fence[i] = 0 there).
 * chain_heads: for every leaf, an array of record_count 1-based version indices (0 = no chain)
 * replacing the host pointers in RecordMetadata.next_ptr; may be NULL if the tree has no chains. */
int stage_gpu_snapshot_begin(const stage_schema_t *schema, uint64_t n_leaves_hint, uint64_t n_rows_hint,
                             stage_snapshot_t **out);
int stage_gpu_snapshot_add_versions(stage_snapshot_t *s, const stage_version_t *versions,
                                    const void *payloads /* n x payload_size */, uint64_t n);
int stage_gpu_snapshot_add_leaves(stage_snapshot_t *s, const void *const *leaf_images, const void *upper_fences,
                                  const uint8_t *has_upper_fence, const uint32_t *const *chain_heads, uint64_t n);
int stage_gpu_snapshot_seal(stage_snapshot_t *s);

/* Build the YCSB int table (benchmark/ycsb/ycsb_loader.cpp:144-159: key = uint32 rowid as 4 raw
 * bytes, payload = payload_size x (uint8)rowid, commit id 0) directly on the device in consolidated
 * form: rows [first_rank, first_rank + n_local) of the table of n_rows_total keys in reference key
 * order, rows_per_leaf records per leaf (every leaf sorted, separator = its largest key).
 * Used where the host tree cannot be materialised. */
int stage_gpu_snapshot_generate_ycsb(const stage_schema_t *schema, uint64_t n_rows_total, uint64_t first_rank,
                                     uint64_t n_local, uint32_t rows_per_leaf, stage_snapshot_t **out);

/* Bulk load from device-resident columns (tables whose host tree cannot be materialised, or that are produced
 * on the GPU): alloc_sorted allocates the snapshot's own columns and hands out device pointers to fill —
 * raw_keys n_rows x key_size bytes, payload n_rows x payload_size bytes, ts n_rows commit ids (zero-initialised) —
 * with rows in REFERENCE key order (BaseNode::KeyCompare, b_tree.h:116-134); seal_sorted converts the keys to the
 * sortable form, verifies they are strictly ascending (STAGE_ERR_ARG otherwise), lays out consolidated leaves of
 * rows_per_leaf records (every leaf sorted, separator = its largest key, as after the reference's consolidation)
 * and builds the search levels. No version chains. */
int stage_gpu_snapshot_alloc_sorted(const stage_schema_t *schema, uint64_t n_rows, uint32_t rows_per_leaf,
                                    stage_snapshot_t **out, void **raw_keys_dev, void **payload_dev, uint32_t **ts_dev);
int stage_gpu_snapshot_seal_sorted(stage_snapshot_t *s);

/* Incremental refresh: apply n committed single-row updates, in batch order, to a sealed
 * snapshot. Update i overwrites payload bytes [patch_off, patch_off + patch_len) of the row with key keys[i] by
 * patches[i] (patch_len bytes each; the whole row when patch_off = 0 and patch_len = payload_size) and commits with
 * commit_ids[i]. The state installed is the one SSNTransactionManager::CommitTransaction leaves
 * (src/concurrency/transaction_manager.cpp:610-676): the old version is retired into the version table with
 * begin = old commit id, commit = commit_ids[i], next = previous chain head; the row takes commit id
 * commit_ids[i] and the new chain head. Updates of absent keys are ignored (*applied_out counts the rest).
 * Not concurrent with reads on the same snapshot: reads issued after the call returns see the new state.
 * Inserts and deletes change leaf membership and need a re-transform of the leaf (not covered). */
int stage_gpu_snapshot_apply_updates(stage_snapshot_t *s, const void *keys, const void *patches, uint32_t patch_off,
                                     uint32_t patch_len, const uint32_t *commit_ids, uint64_t n, int flags,
                                     uint64_t *applied_out);

/* Snapshot serialisation: dump / restore the flattened arrays (keys, slots, commit ids, chain
 * heads, payload rows, leaf tables, version table) so that a table need not be re-flattened after a restart (the
 * reference itself has no checkpointing, README.md:144-146). The search levels are rebuilt at load. The file is
 * tied to this library's layout ("STGSNAP1"), little-endian. */
int stage_gpu_snapshot_save(stage_snapshot_t *s, const char *path);
int stage_gpu_snapshot_load(const char *path, stage_snapshot_t **out);

int stage_gpu_snapshot_info(const stage_snapshot_t *s, stage_snapshot_info_t *info);
int stage_gpu_snapshot_schema(const stage_snapshot_t *s, stage_schema_t *schema);
int stage_gpu_snapshot_destroy(stage_snapshot_t *s);

/* The CUDA stream (cudaStream_t) the snapshot launches on; set it to share a caller's stream. */
int stage_gpu_snapshot_set_stream(stage_snapshot_t *s, void *cuda_stream);
void *stage_gpu_snapshot_get_stream(const stage_snapshot_t *s);

/* Per-call kernel timing with CUDA events recorded on the snapshot's stream around the two launches of
 * a *_batch call with STAGE_IO_DEVICE (plan kernel: search / scan planning + MVCC; gather kernel: payload
 * rows). last_timing() waits for the last call and returns its two durations in milliseconds. */
int stage_gpu_snapshot_enable_timing(stage_snapshot_t *s, int on);
int stage_gpu_snapshot_last_timing(stage_snapshot_t *s, float *plan_ms, float *gather_ms);

/* ---- reads -------------------------------------------------------------------------------- */

/* Batched BTree::Read + IndexScanExecutor point branch (b_tree.cpp:2066-2129, executor.h:374-454).
 * keys: n x key_size bytes. For probe i:
 *   found[i]   1 if a version visible at read_ts exists (latest in-leaf version if
 *              read_ts >= its commit id, else the first chain version with begin <= read_ts <= commit)
 *   ts[i]      commit id of the version returned (in-leaf: meta & 0xFFFFFFFF; chain: begin_id_)
 *   key_out    n x key_size, payload_out n x payload_size: the key / payload of the version returned where
 *              found[i] == 1; rows of missed probes are either left untouched (device buffers, pinned host
 *              buffers) or zero-filled (pageable host buffers) — pre-fill if the difference matters
 * Any output pointer may be NULL. */
int stage_gpu_probe_batch(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                          uint32_t *ts, void *key_out, void *payload_out, int flags);

/* Batched BTree::RangeScanBySize + the executor loop over Iterator::GetNext
 * (b_tree.h:830-941, b_tree.cpp:1261-1315; executor.h:455-535 with apply_visibility = 1,
 * executor.h:620-638 with apply_visibility = 0). Scan i starts at start_keys[i] (inclusive) and asks
 * for scan_sizes[i] rows; its rows are written, in the reference's order, to row slots
 * [i*out_stride, i*out_stride + row_counts[i]) of key_out / ts_out / payload_out
 * (out_stride >= max scan size: STAGE_ERR_ARG otherwise with host buffers; with device buffers, where the sizes are
 * not visible to the host, a larger size is clamped to out_stride). Row slots beyond row_counts[i] are either left
 * untouched or zero-filled. With STAGE_IO_COMPACT (host buffers only) the rows of all scans are packed back to back
 * instead: scan i occupies row slots [sum of row_counts[0..i), + row_counts[i]) — exactly the rows that exist cross
 * PCIe; the buffers must still hold n * out_stride rows in the worst case. Reproduces the iterator's slot-order truncation and
 * leaf-hop termination rules exactly. apply_visibility = 1 is the executor's
 * range rule: a row is returned iff read_ts >= the commit id of its in-leaf version; a newer row is
 * dropped — the range branch never reaches a version chain, because the iterator's records carry no
 * chain pointer (record_meta.h:240; only BTree::Read fills RecordMeta::next_tuple_ptr). */
int stage_gpu_scan_batch(stage_snapshot_t *s, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                         uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts,
                         void *key_out, uint32_t *ts_out, void *payload_out, int flags);

/* Operator push-down: the same probe / scan, but only payload bytes [col_off, col_off + col_len)
 * of every row are returned — what the TPC-C callers of the path consume (S_QUANTITY against a threshold,
 * benchmark/tpcc/tpcc_stock_level.cpp:167-176; MIN(S_QUANTITY), benchmark/tpcc/tpcc_new_order.cpp:808-866).
 * col_out: n x col_len (probe) or n x out_stride x col_len (scan); entries without a row are zero.
 * Row selection, order, found / ts / row_counts / key_out are exactly those of the full-row calls. */
int stage_gpu_probe_project(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts, uint8_t *found,
                            uint32_t *ts, uint32_t col_off, uint32_t col_len, void *col_out, int flags);
int stage_gpu_scan_project(stage_snapshot_t *s, const void *start_keys, const uint32_t *scan_sizes, uint64_t n,
                           uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts,
                           void *key_out, uint32_t col_off, uint32_t col_len, void *col_out, int flags);

/* Predicate / aggregate push-down: probe the batch and reduce one little-endian int32 payload column (at byte col_off,
 * any alignment) of the visible rows on the device. found = probes that hit a visible version, below = those whose
 * column is < threshold (RunStockLevel's S_QUANTITY test, benchmark/tpcc/tpcc_stock_level.cpp:167-176), min / max over
 * the found rows (CH-Q2*'s MIN(S_QUANTITY), benchmark/tpcc/tpcc_new_order.cpp:808-866; INT32_MAX / INT32_MIN when
 * nothing was found). 24 bytes come back for the whole batch. `out` is a host pointer; keys follow `flags`. */
typedef struct {
    uint64_t found, below;
    int32_t min_value, max_value;
} stage_agg_i32_t;
int stage_gpu_probe_aggregate_i32(stage_snapshot_t *s, const void *keys, uint64_t n, uint32_t read_ts, uint32_t col_off,
                                  int32_t threshold, stage_agg_i32_t *out, int flags);

/* stage_gpu_probe_aggregate_i32 for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). No counterpart in the reference.
 * The call allocates nothing (no cudaMalloc / cudaFree / allocating library routine) and does not synchronise: its
 * scratch — the source row of every probe — is a fixed device array inside the library (4 MiB per GPU that loads the
 * library, held until exit), so n <= 1 048 576 probes per call (STAGE_ERR_ARG with a message in
 * stage_gpu_last_error() beyond that). keys_dev (n x key_size) and out_dev (one record, 8-byte aligned) are DEVICE
 * memory on the snapshot's GPU; the three kernels are enqueued on the snapshot's stream
 * (stage_gpu_snapshot_set_stream) and out_dev holds the result once they have run — same record, same values as
 * stage_gpu_probe_aggregate_i32. The array is shared by every call on that GPU and keeps the last call's contents
 * (what a call reads it has written before; the record is initialised on the device): calls that overlap on the GPU
 * — two snapshots, two streams — race; serialise them. The array's device address is looked up once per device on the
 * first call there (cudaGetSymbolAddress): make one call before starting a graph capture. */
int stage_gpu_probe_aggregate_i32_noalloc(stage_snapshot_t *s, const void *keys_dev, uint64_t n, uint32_t read_ts,
                                          uint32_t col_off, int32_t threshold, stage_agg_i32_t *out_dev);

/* The same reduction per GROUP of consecutive probes: CH-Q2* loops over the suppliers and reduces the ~300 stock rows
 * of each (benchmark/tpcc/tpcc_new_order.cpp:803-866; the (w, i) lists come from supp_stock_map,
 * benchmark/tpcc/tpcc_workload.cpp:581-585). Group g = probes [group_offsets[g], group_offsets[g+1]); group_offsets has
 * n_groups + 1 ascending entries, [0] = 0 and [n_groups] = n <= 2^26. One 24-byte record per group comes back:
 *   found, below            probes of the group that hit a visible version / whose column is < threshold
 *   min_value, max_value    over the found rows (INT32_MAX / INT32_MIN when none)
 *   min_index               batch index of the FIRST probe attaining min_value ("the stock tuple having min. stock
 *                           level"), 0xFFFFFFFF when none
 *   last_below_index        batch index of the LAST found probe whose column is < threshold, 0xFFFFFFFF when none — with
 *                           threshold = 32767 this is the row the reference's loop literally keeps: its running minimum
 *                           is never lowered (min_qty stays INT16_MAX, tpcc_new_order.cpp:810-860), so every row below it
 *                           overwrites the pick
 * keys, group_offsets and out all follow `flags` (host or device). */
typedef struct {
    uint32_t found, below;
    int32_t min_value;
    uint32_t min_index;
    int32_t max_value;
    uint32_t last_below_index;
} stage_group_agg_i32_t;
int stage_gpu_probe_group_aggregate_i32(stage_snapshot_t *s, const void *keys, uint64_t n, const uint64_t *group_offsets,
                                        uint64_t n_groups, uint32_t read_ts, uint32_t col_off, int32_t threshold,
                                        stage_group_agg_i32_t *out, int flags);

/* RunStockLevel's read side as ONE device-side operator (benchmark/tpcc/tpcc_stock_level.cpp:37-196).
 * Per request: for each of the 20 orders o in [next_o_id - 20, next_o_id) a range scan of 10 order_line rows from
 * {w_id, d_id, o, 5} under the executor's range visibility rule (:100-150); the FIRST returned row that still belongs
 * to (w_id, d_id, o) names an item (:152-155); stock {w_id, item} is probed under the point visibility rule (version
 * chains walked, :156-186); items whose visible S_QUANTITY is below `threshold` are counted once (:187-190).
 * distinct_items[i] = that count. Nothing but 4 bytes per transaction leaves the device. Table layouts are TPC-C's
 * (benchmark/tpcc/tpcc_record.h): order_line key {int64 w, d, o, number} (32 B), payload starts with int32 OL_I_ID;
 * stock key {int64 w, i} (16 B), payload starts with int32 S_QUANTITY. reqs / distinct_items follow `flags`
 * (STAGE_IO_HOST: the call returns when the counts are in host memory; STAGE_IO_DEVICE: enqueued on order_line's
 * stream). */
typedef struct {
    int64_t w_id, d_id;
    int64_t next_o_id;   /* D_NEXT_O_ID of the district (the caller's first read, tpcc_stock_level.cpp:58-90) */
    int32_t threshold;
    int32_t reserved;
} stage_stock_level_req_t;
int stage_gpu_tpcc_stock_level(stage_snapshot_t *order_line, stage_snapshot_t *stock, const stage_stock_level_req_t *reqs,
                               uint64_t n, uint32_t read_ts, uint32_t *distinct_items, int flags);

/* TableScanExecutor with scan_sz == -1 (include/execute/executor.h:581-619): the PHYSICAL whole-tree walk — leaves in
 * tree order, record slots in slot (insertion) order, no visibility test. Returns rows [first_row, first_row + n) of
 * that order (n_rows of stage_gpu_snapshot_info in total): latest in-leaf versions, keys, commit ids, payload rows.
 * Slots that are invisible / deleted in the tree (the reference copies their stale bytes too) are not part of a
 * snapshot and are skipped. Any output pointer may be NULL. */
int stage_gpu_scan_physical(stage_snapshot_t *s, uint64_t first_row, uint64_t n, void *key_out, uint32_t *ts_out,
                            void *payload_out, int flags);

/* Range partitioning across GPUs (one process per GPU): splitters = the largest key of every shard
 * but the last, in reference key order. owner[i] = shard that owns keys[i]. Host pointers. */
int stage_gpu_route(const stage_schema_t *schema, const void *splitters, uint32_t n_shards, const void *keys,
                    uint64_t n, uint32_t *owner);

/* Same on the device: keys_dev / owner_dev are device pointers, splitters stay on the host (n_shards <= 64).
 * Launches on cuda_stream (cudaStream_t, may be NULL = default stream). */
int stage_gpu_route_device(const stage_schema_t *schema, const void *splitters, uint32_t n_shards, const void *keys_dev,
                           uint64_t n, uint32_t *owner_dev, void *cuda_stream);

/* ---- routed probes over peer memory (N GPUs, one process per GPU) ----------------------------- */

/* The probe exchange: every rank owns an "arena" in its HBM — an inbox per source rank and the
 * result slab of its own requests — that all other ranks map (CUDA IPC over NVLink / NVSwitch). A routed batch is
 *   post   : route this rank's keys and store (key, request index) straight into the owners' inboxes
 *   serve  : search + MVCC over this rank's inboxes on its shard, payload rows written straight into the
 *            requesters' result slabs at the row of the original request (request order, no return collective).
 * Hand-offs: with STAGE_XCHG_DEVICE_SYNC the kernels signal / wait through epoch flags in the arenas (every rank
 * must call post, serve and wait once per batch, in that order, on one stream); without it the caller places a
 * barrier across ranks between post and serve and another after serve (e.g. a stream-ordered NCCL all-reduce).
 * capacity = the largest batch (probes per rank per call). world <= 16. */
typedef struct stage_exchange stage_exchange_t;
enum { STAGE_IPC_HANDLE_BYTES = 64, STAGE_XCHG_DEVICE_SYNC = 1 };

int stage_gpu_exchange_create(uint32_t world, uint32_t rank, uint64_t capacity, const stage_schema_t *schema,
                              stage_exchange_t **out);
/* 64-byte CUDA IPC handle of this rank's arena, to be all-gathered by the caller's control plane. */
int stage_gpu_exchange_ipc_handle(stage_exchange_t *x, void *handle_out);
/* handles: world x STAGE_IPC_HANDLE_BYTES, indexed by rank (the own entry is ignored). */
int stage_gpu_exchange_attach_ipc(stage_exchange_t *x, const void *handles);
/* Same-process peers (several ranks driven by one process, peer access already enabled): arena pointers by rank. */
int stage_gpu_exchange_attach_ptrs(stage_exchange_t *x, void *const *arenas);
void *stage_gpu_exchange_arena(stage_exchange_t *x);
/* splitters: (world-1) x key_size host bytes, the largest key of every shard but the last (as stage_gpu_route).
 * keys_dev: n x key_size device bytes, n <= capacity. cuda_stream: cudaStream_t (NULL = default stream). */
int stage_gpu_exchange_post(stage_exchange_t *x, const void *splitters, const void *keys_dev, uint64_t n, int flags,
                            void *cuda_stream);
int stage_gpu_exchange_serve(stage_exchange_t *x, stage_snapshot_t *shard, uint32_t read_ts, int flags, void *cuda_stream);
/* STAGE_XCHG_DEVICE_SYNC only: stream-ordered wait until every owner has served this rank's batch. */
int stage_gpu_exchange_wait(stage_exchange_t *x, void *cuda_stream);
/* Device pointers (local) of this rank's result slab, rows in request order: found u8[capacity], ts u32[capacity],
 * payload rows of payload_size bytes every *payload_stride bytes (payload_size rounded up to 128, so that the
 * owners' peer stores are line-aligned). Rows of missed probes keep their previous content. */
int stage_gpu_exchange_results(stage_exchange_t *x, void **found, void **ts, void **payload, uint32_t *payload_stride);

/* Routed push-down: the same post / serve / wait sequence, but the owner sends back only what the
 * caller consumes instead of the payload row — what lifts the NVLink bound of the full-row form (4-16 B instead of
 * 416-1004 B per probe). serve_project: payload bytes [col_off, col_off + col_len) of every found row (zeros for a miss)
 * into the requester's slab, found / ts as usual; project_results returns the device pointers (column every
 * *col_stride bytes = col_len rounded up to 4). serve_aggregate_i32: every owner reduces the little-endian int32
 * column at col_off over each requester's probes (found, below threshold, MIN, MAX — RunStockLevel's S_QUANTITY test,
 * benchmark/tpcc/tpcc_stock_level.cpp:167-190; CH-Q2*'s MIN(S_QUANTITY) over a supplier's (w, i) keys that span all
 * warehouses, i.e. all shards, benchmark/tpcc/tpcc_new_order.cpp:808-866) and ships ONE 24-byte record per requester;
 * aggregate_result (after the wait) combines the owners' records of this rank's batch into *out (host pointer). */
int stage_gpu_exchange_serve_project(stage_exchange_t *x, stage_snapshot_t *shard, uint32_t read_ts, uint32_t col_off,
                                     uint32_t col_len, int flags, void *cuda_stream);
int stage_gpu_exchange_project_results(stage_exchange_t *x, void **found, void **ts, void **col, uint32_t *col_stride);
int stage_gpu_exchange_serve_aggregate_i32(stage_exchange_t *x, stage_snapshot_t *shard, uint32_t read_ts, uint32_t col_off,
                                           int32_t threshold, int flags, void *cuda_stream);
int stage_gpu_exchange_aggregate_result(stage_exchange_t *x, stage_agg_i32_t *out, void *cuda_stream);

/* Routed range scans over the same arenas ("scans run shard-local followed by a gather"). The scan
 * (start key, size) is posted to the owner of its START key; the owner plans it on its shard with the ordinary scan
 * plan kernels (same iterator quirks and MVCC filter as stage_gpu_scan_batch) and writes the row count, keys, commit
 * ids and payload rows straight into the requester's slab, request q at rows [q*rows_per_scan, +count). A scan does
 * not continue on the next shard: shards are expected to overlap their right neighbour by a halo of rows (the
 * caller's sharding decides how many: enough for the longest scan including the rows the filter drops).
 * Replaces: the per-thread loop over BTree::RangeScanBySize / Iterator::GetNext (include/vstore/b_tree.h:830-835,
 * 883-941) inside IndexScanExecutor / TableScanExecutor::Execute (include/execute/executor.h:455-535, 620-638) on a
 * range-partitioned table. key_size must be a multiple of 4.
 * Call order per batch as for probes: post_scans, serve_scans, wait (device sync) on one stream. */
int stage_gpu_exchange_create_scan(uint32_t world, uint32_t rank, uint64_t max_scans, uint32_t max_scan_size,
                                   const stage_schema_t *schema, stage_exchange_t **out);
/* start_keys_dev: n x key_size, scan_sizes_dev: u32[n] (clamped to max_scan_size), device memory. */
int stage_gpu_exchange_post_scans(stage_exchange_t *x, const void *splitters, const void *start_keys_dev,
                                  const uint32_t *scan_sizes_dev, uint64_t n, int flags, void *cuda_stream);
int stage_gpu_exchange_serve_scans(stage_exchange_t *x, stage_snapshot_t *shard, uint32_t read_ts, int apply_visibility,
                                   int flags, void *cuda_stream);
/* Device pointers (local) of this rank's scan result slab: row_counts u32[max_scans]; per scan *rows_per_scan rows
 * of keys (every *key_stride bytes), commit ids (u32) and payload rows (every *payload_stride bytes). Rows beyond
 * row_counts[q] keep their previous content. */
int stage_gpu_exchange_scan_results(stage_exchange_t *x, void **row_counts, void **key_out, void **ts, void **payload,
                                    uint32_t *rows_per_scan, uint32_t *key_stride, uint32_t *payload_stride);
/* Measurement helper: one cudaMemcpyAsync of `bytes` from this rank's result slab into rank dst_rank's slab at byte
 * dst_offset (copy engine over NVLink). Timed on all ranks at once it gives the box's link ceiling for the exchange's
 * traffic pattern. The slab contents are undefined afterwards, until the next batch. */
int stage_gpu_exchange_peer_copy(stage_exchange_t *x, uint32_t dst_rank, uint64_t dst_offset, uint64_t bytes, void *cuda_stream);
/* Sticky device-side error of the exchange: 0 ok, 1 inbox overflow, 2 device-sync timeout. Synchronises the stream. */
int stage_gpu_exchange_error(stage_exchange_t *x, void *cuda_stream, uint32_t *err_out);
/* Per-phase timing of the last post/serve/wait sequence (STAGE_XCHG_DEVICE_SYNC), CUDA events on the caller's stream:
 * ms4 = { post (route + inbox stores), serve search (incl. waiting for the peers' posts), payload rows (local +
 * peer stores), wait for the owners }. */
int stage_gpu_exchange_enable_timing(stage_exchange_t *x, int on);
int stage_gpu_exchange_last_timing(stage_exchange_t *x, float *ms4);
int stage_gpu_exchange_destroy(stage_exchange_t *x);

/* ---- batching front-end for per-operation callers ------------------------------------------- */

/* The reference's drivers run ONE operation per executor on N pinned threads (benchmark/ycsb/ycsb_workload.cpp:522-529,
 * ycsb_mixed.cpp:83-93: IndexScanExecutor per key). A batcher coalesces the operations that concurrent callers have in
 * flight into one batched launch: the first caller of a forming batch leads it, waits up to max_wait_us (or until
 * max_batch operations joined), runs the batch with every operation's own read id, and all callers pick up their
 * results. The calls block until the caller's operation is done and are safe from any number of threads. Results are
 * those of stage_gpu_probe_batch / stage_gpu_scan_batch for the same (key, read id). The batcher must be destroyed
 * before its snapshot, with no call in flight. */
typedef struct stage_batcher stage_batcher_t;
int stage_gpu_batcher_create(stage_snapshot_t *s, uint32_t max_batch, uint32_t max_wait_us, stage_batcher_t **out);
int stage_gpu_batcher_probe(stage_batcher_t *b, const void *key, uint32_t read_ts, uint8_t *found, uint32_t *ts, void *key_out,
                            void *payload_out);
/* key_out / ts_out / payload_out hold scan_size rows; *row_count rows are written */
int stage_gpu_batcher_scan(stage_batcher_t *b, const void *start_key, uint32_t scan_size, uint32_t read_ts, int apply_visibility,
                           uint32_t *row_count, void *key_out, uint32_t *ts_out, void *payload_out);
int stage_gpu_batcher_stats(stage_batcher_t *b, uint64_t *batches, uint64_t *operations);
int stage_gpu_batcher_destroy(stage_batcher_t *b);

/* ---- utilities ---------------------------------------------------------------------------- */
void *stage_gpu_host_alloc(size_t bytes); /* pinned host memory for STAGE_IO_HOST batches */
void stage_gpu_host_free(void *p);
/* Pin memory the caller already owns (e.g. the std::vector<T> storage a TableScanExecutor-style caller returns rows
 * in, include/execute/executor.h:644-649) so that results can be stored into it directly. */
int stage_gpu_host_register(void *p, size_t bytes);
int stage_gpu_host_unregister(void *p);
/* Counters kept on the device since the snapshot was sealed (or last reset). slot_order_scans: range scans in which
 * the reference's slot-order early break (LeafNode::RangeScanBySize, src/vstore/b_tree.cpp:1276-1309; SURVEY.md §3.4
 * Q5) decided which rows of the last leaf were returned — 0 on consolidated tables. Synchronises the device. */
typedef struct {
    uint64_t slot_order_scans;
    uint64_t reserved[7];
} stage_counters_t;
int stage_gpu_snapshot_counters(stage_snapshot_t *s, stage_counters_t *out, int reset);
/* Kernels launched by this library since process start (for bench.py's gpu_launches). */
uint64_t stage_gpu_launch_count(void);
const char *stage_gpu_strerror(int status);
const char *stage_gpu_last_error(void); /* text of the last CUDA error seen on this thread */

#ifdef __cplusplus
}
#endif
#endif /* STAGE_GPU_H_ */
