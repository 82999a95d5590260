// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never shipped, never linked into libstage_gpu.so).
//
// Our own C-API harness over the UNMODIFIED reference library (compiled by oracle/Makefile.ref
// from the sources where they lie under /root/reference into oracle/_ref/libstage_ref.so).
// It drives the reference's real code path for the hot path:
//   * table construction   : BTree::Insert + BTree::FinalizeInsert, like LoadYCSBRows
//                            (/root/reference/benchmark/ycsb/ycsb_loader.cpp:93-173)
//   * point probe          : BTree::Read (src/vstore/b_tree.cpp:2066-2129) + the visibility rule of
//                            IndexScanExecutor (include/execute/executor.h:383-446)
//   * range scan, MVCC     : the reference's own IndexScanExecutor range branch (executor.h:455-535) run
//                            with a read-only TransactionContext of the requested read id; rows are captured
//                            in its predicate callback (its result vector holds pointers into freed Records)
//   * range scan, no filter: BTree::RangeScanBySize -> Iterator::GetNext (include/vstore/b_tree.h:830-941)
//                            + the TableScanExecutor loop (executor.h:626-638), restated here because that
//                            executor copies a compile-time struct; tests/cpp/dropin_test.cpp pins the
//                            restatement against the real TableScanExecutor
// The point executor's chain-hit branch is undefined behaviour at HEAD (it copies into a null tuple,
// SURVEY.md §3.4 Q10), so for probes the harness applies the visibility rule itself on the Record that
// BTree::Read returns: "emit the matching version once, then stop".
//
// Used by tests/ (as the checker) and by bench.py's cpu_baseline / --impl reference legs only.

#include <cstdint>
#include <cstring>
#include <vector>
#include <thread>
#include <chrono>
#include <atomic>
#include <limits>

#include <cstddef>
#include "execute/executor.h"
#include "vstore/b_tree.h"
#include "vstore/valen_buffer.h"
#include "vstore/version_store.h"
#include "common/catalog.h"
#include "common/ephemeral_pool.h"

using namespace mvstore;

struct ref_table {
    BTree *tree = nullptr;
    Catalog *catalog = nullptr;
    DramBlockPool *leaf_pool = nullptr;
    InnerNodeBuffer *inner_pool = nullptr;
    EphemeralPool *undo_pool = nullptr;
    uint32_t payload_size = 0;
    uint32_t leaf_node_size = 0;
    // version-chain storage created by ref_commit_update (TupleHeader + slot bytes)
    std::vector<char *> chain_allocs;
};

extern "C" {

// ---------------------------------------------------------------------------------------------
// construction
// ---------------------------------------------------------------------------------------------
ref_table *ref_table_create(uint32_t catalog_key_len, uint32_t payload_size, uint32_t split_threshold,
                            uint32_t merge_threshold, uint32_t leaf_node_size, uint64_t leaf_pool_blocks) {
    auto *t = new ref_table();
    t->payload_size = payload_size;
    t->leaf_node_size = leaf_node_size;
    t->catalog = new Catalog();
    uint8_t *bitmap = new uint8_t[2];
    bitmap[0] = 1;
    bitmap[1] = 0;
    t->catalog->init("table0", 2, bitmap, catalog_key_len, 1, false);
    t->catalog->add_col("Key", catalog_key_len, "VARCHAR");
    t->catalog->add_col("payload", payload_size, "VARCHAR");
    ParameterSet param(split_threshold, merge_threshold, leaf_node_size, payload_size);
    t->leaf_pool = new DramBlockPool(leaf_pool_blocks, leaf_pool_blocks);
    RecordBufferPool *p1 = new RecordBufferPool(5000000, 5000000);
    t->inner_pool = new InnerNodeBuffer(p1);
    RecordBufferPool *p2 = new RecordBufferPool(5000000, 5000000);
    UndoBuffer *undo = new UndoBuffer(p2);
    t->undo_pool = new EphemeralPool(undo);
    t->tree = new BTree(param, *t->catalog, t->leaf_pool, t->inner_pool, t->undo_pool);
    SSNTransactionManager::GetInstance()->Init(t->undo_pool);        // the executors ask the singleton (PerformRead)
    return t;
}

// Insert n (key, payload) pairs in order and finalize each with commit_id, exactly like the loaders do.
// Returns number of rows actually inserted (duplicates are skipped like the loader's KeyExists branch).
int64_t ref_insert(ref_table *t, const void *keys, uint32_t key_size, const void *payloads, uint64_t n,
                   uint32_t commit_id) {
    const char *k = static_cast<const char *>(keys);
    const char *p = static_cast<const char *>(payloads);
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; ++i) {
        RecordMeta meta;
        ReturnCode rc = t->tree->Insert(k + i * key_size, (uint16_t) key_size, p + i * (uint64_t) t->payload_size,
                                        &meta, commit_id);
        if (!rc.IsOk()) continue;
        RecordMetadata *meta_ptr =
                reinterpret_cast<RecordMetadata *>(meta.GetMetaData().GetLocationPtr()->record_meta_ptr);
        t->tree->FinalizeInsert(meta_ptr, commit_id);
        t->tree->IncreaseTupleCount(1);
        ++ok;
    }
    return ok;
}

// Variable-length keys (reference test fixture uses std::to_string(i) keys): key i occupies
// keys[off[i] .. off[i+1]).
int64_t ref_insert_varlen(ref_table *t, const void *keys, const uint32_t *off, const void *payloads, uint64_t n,
                          uint32_t commit_id) {
    const char *k = static_cast<const char *>(keys);
    const char *p = static_cast<const char *>(payloads);
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; ++i) {
        RecordMeta meta;
        ReturnCode rc = t->tree->Insert(k + off[i], (uint16_t) (off[i + 1] - off[i]),
                                        p + i * (uint64_t) t->payload_size, &meta, commit_id);
        if (!rc.IsOk()) continue;
        RecordMetadata *meta_ptr =
                reinterpret_cast<RecordMetadata *>(meta.GetMetaData().GetLocationPtr()->record_meta_ptr);
        t->tree->FinalizeInsert(meta_ptr, commit_id);
        ++ok;
    }
    return ok;
}

// YCSB int loader body (ycsb_loader.cpp:144-172): key = 4 LE bytes of rowid, payload = 1000 x (uint8)rowid.
int64_t ref_load_ycsb_int(ref_table *t, uint64_t begin_rowid, uint64_t end_rowid) {
    std::vector<char> payload(t->payload_size);
    int64_t ok = 0;
    for (uint64_t rowid = begin_rowid; rowid < end_rowid; ++rowid) {
        memset(payload.data(), (int) rowid, payload.size());
        uint32_t k_ = (uint32_t) rowid;
        RecordMeta meta;
        ReturnCode rc = t->tree->Insert(reinterpret_cast<const char *>(&k_), sizeof(uint32_t), payload.data(),
                                        &meta, INVALID_CID);
        if (!rc.IsOk()) continue;
        RecordMetadata *meta_ptr =
                reinterpret_cast<RecordMetadata *>(meta.GetMetaData().GetLocationPtr()->record_meta_ptr);
        t->tree->FinalizeInsert(meta_ptr, INVALID_CID);
        t->tree->IncreaseTupleCount(1);
        ++ok;
    }
    return ok;
}

// Delete via BTree::Delete (is_for_update=true path zeroes the slot meta immediately,
// src/vstore/b_tree.cpp:1210-1221) — used to create leaves whose max key is not the separator (Q6).
int ref_delete(ref_table *t, const void *key, uint32_t key_size) {
    RecordMeta meta;
    TupleHeader hdr;
    ReturnCode rc = t->tree->Delete(static_cast<const char *>(key), (uint16_t) key_size, &meta, &hdr, true, 0);
    return rc.IsOk() ? 0 : -(int) rc.rc;
}

// Post-commit state of a committed single-row update, constructed with the reference's own types
// exactly as SSNTransactionManager::CommitTransaction installs it
// (src/concurrency/transaction_manager.cpp:610-676): the old version [key][payload] goes into a
// TupleHeader slot with begin_id_ = old cstamp, comm_id_ = committer stamp, next_ = previous chain
// head; the leaf record is overwritten in place, its meta takes the new commit id and next_ptr
// points at the new TupleHeader. (Going through the full Index-SSN machinery is out of scope and
// nondeterministic because of its GC thread; the bytes the read path sees are these.)
int ref_commit_update(ref_table *t, const void *key, uint32_t key_size, const void *new_payload,
                      uint32_t commit_id) {
    LeafNode *leaf = t->tree->TraverseToLeaf(nullptr, static_cast<const char *>(key), (uint16_t) key_size);
    if (!leaf) return -1;
    RecordMetadata *meta_ptr = nullptr;
    ReturnCode rc = leaf->Read(static_cast<const char *>(key), (uint16_t) key_size, t->payload_size, &meta_ptr,
                               false);
    if (!rc.IsOk() || !meta_ptr) return -2;
    char *rec_key = nullptr, *rec_payload = nullptr;
    leaf->GetRawRecord(*meta_ptr, &rec_key, &rec_payload);
    uint32_t klen = meta_ptr->GetKeyLength();
    char *mem = new char[sizeof(TupleHeader) + klen + t->payload_size];
    t->chain_allocs.push_back(mem);
    TupleHeader *hd = new (mem) TupleHeader();
    char *slot = mem + sizeof(TupleHeader);
    hd->init(hd, reinterpret_cast<uint64_t>(slot), (oid_t) t->chain_allocs.size());
    memcpy(slot, rec_key, klen);
    memcpy(slot + klen, rec_payload, t->payload_size);
    hd->SetBeginId(meta_ptr->GetTxnCommitId());
    hd->SetCommitId(commit_id);
    hd->SetNextHeaderPointer(meta_ptr->GetNextPointer());
    memcpy(rec_payload, new_payload, t->payload_size);
    meta_ptr->FinalizeForUpdate(commit_id);
    meta_ptr->SetNextPointer(reinterpret_cast<uint64_t>(hd));
    return 0;
}

// n committed single-row updates: update i rewrites payload bytes [100,200) with (uint8)(first_cid+i) and
// commits with id first_cid+i (the shape of the YCSB update transaction, ycsb_mixed.cpp:38-80, after commit).
int64_t ref_commit_update_batch(ref_table *t, const void *keys, uint32_t key_size, uint64_t n, uint32_t first_cid) {
    const char *k = static_cast<const char *>(keys);
    const uint32_t P = t->payload_size;
    std::vector<char> np_(P);
    int64_t ok = 0;
    for (uint64_t i = 0; i < n; ++i) {
        LeafNode *leaf = t->tree->TraverseToLeaf(nullptr, k + i * key_size, (uint16_t) key_size);
        if (!leaf) continue;
        RecordMetadata *meta_ptr = nullptr;
        ReturnCode rc = leaf->Read(k + i * key_size, (uint16_t) key_size, P, &meta_ptr, false);
        if (!rc.IsOk() || !meta_ptr) continue;
        char *rk = nullptr, *rp = nullptr;
        leaf->GetRawRecord(*meta_ptr, &rk, &rp);
        memcpy(np_.data(), rp, P);
        uint32_t lo = P > 100 ? 100 : 0, hi = P > 200 ? 200 : P;
        memset(np_.data() + lo, (int) ((first_cid + i) & 0xFF), hi - lo);
        ok += ref_commit_update(t, k + i * key_size, key_size, np_.data(), first_cid + (uint32_t) i) == 0;
    }
    return ok;
}

// An update left IN FLIGHT: LeafNode::Update (b_tree.cpp:1061-1163) sets the control bit, parks the
// old record in the EphemeralPool, points next_ptr at that copy and overwrites the payload in place — and nothing
// finalizes it. new_payload replaces column 1. Returns the ReturnCode byte (0 = ok).
int ref_begin_update(ref_table *t, const void *key, uint32_t key_size, const void *new_payload, uint32_t txn_id) {
    std::vector<oid_t> columns{1};
    RecordMeta meta;
    ReturnCode rc = t->tree->Update(static_cast<const char *>(key), (uint16_t) key_size,
                                    static_cast<const char *>(new_payload), columns, &meta, false, txn_id);
    return rc.IsOk() ? 0 : (int) rc.rc;
}
// One whole update transaction on the record of `key`, with the reference's own steps and their order: LeafNode::Update
// (control bit set by compare-and-swap, old record parked in the EphemeralPool, next_ptr -> that copy, payload overwritten
// in place, b_tree.cpp:1061-1163), then the commit install of SSNTransactionManager::CommitTransaction
// (transaction_manager.cpp:610-676): the parked copy becomes a TupleHeader slot (begin = old commit stamp, commit =
// committer stamp, next = the chain the copy saved), the record's meta word is finalized by compare-and-swap and THEN
// next_ptr is swung to the new header by a second compare-and-swap. Used by the tests that snapshot a tree while
// writers are really running. Returns 0, or the ReturnCode byte of a refused update.
int ref_update_txn(ref_table *t, const void *key, uint32_t key_size, const void *new_payload, uint32_t commit_id) {
    std::vector<oid_t> columns{1};
    RecordMeta rm;
    ReturnCode rc = t->tree->Update(static_cast<const char *>(key), (uint16_t) key_size, static_cast<const char *>(new_payload),
                                    columns, &rm, false, commit_id);
    if (!rc.IsOk()) return (int) rc.rc ? (int) rc.rc : -1;
    RecordMetadata *meta_ptr = reinterpret_cast<RecordMetadata *>(rm.GetMetaData().GetLocationPtr()->record_meta_ptr);
    const uint64_t copy_location = meta_ptr->GetNextPointer();
    const uint32_t klen = meta_ptr->GetKeyLength();
    char *mem = new char[sizeof(TupleHeader) + klen + t->payload_size];
    t->chain_allocs.push_back(mem);
    TupleHeader *hd = new (mem) TupleHeader();
    char *slot = mem + sizeof(TupleHeader);
    hd->init(hd, reinterpret_cast<uint64_t>(slot), (oid_t) t->chain_allocs.size());
    memcpy(slot, reinterpret_cast<const char *>(copy_location), klen + t->payload_size);
    hd->SetBeginId(meta_ptr->GetTxnCommitId());
    hd->SetCommitId(commit_id);
    hd->SetNextHeaderPointer(t->undo_pool->GetNext(copy_location));
    RecordMetadata finl = *meta_ptr, fresh = finl;
    fresh.FinalizeForUpdate(commit_id);
    fresh.SetNextPointer(reinterpret_cast<uint64_t>(hd));
    if (!__atomic_compare_exchange_n(&meta_ptr->meta, &finl.meta, fresh.meta, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) return -2;
    if (!__atomic_compare_exchange_n(&meta_ptr->next_ptr, &finl.next_ptr, fresh.next_ptr, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) return -3;
    return 0;
}

// A writer: update transactions over keys[0..n) (YCSB rows: payload = (uint8) rowid, column [100,200) <- (uint8) commit id,
// commit ids first_cid + i) until *stop becomes non-zero. committed[i] = 1 for the transactions that went through (an
// update whose new column equals the old one is refused as "not needed", b_tree.cpp:1087-1089). Returns how many were tried.
uint64_t ref_update_loop(ref_table *t, const void *keys, uint32_t key_size, uint64_t n, uint32_t first_cid, volatile int *stop,
                         uint8_t *committed) {
    const char *k = static_cast<const char *>(keys);
    const uint32_t P = t->payload_size;
    std::vector<char> pay(P);
    uint64_t i = 0;
    for (; i < n && !*stop; ++i) {
        uint32_t rowid;
        memcpy(&rowid, k + i * key_size, 4);
        memset(pay.data(), (int) (rowid & 0xFF), P);
        uint32_t lo = P > 100 ? 100 : 0, hi = P > 200 ? 200 : P;
        memset(pay.data() + lo, (int) ((first_cid + i) & 0xFF), hi - lo);
        const int rc = ref_update_txn(t, k + i * key_size, key_size, pay.data(), first_cid + (uint32_t) i);
        if (committed) committed[i] = rc == 0;
        if (rc != 0 && rc != (int) ReturnCode::RetNotNeededUpdate) break;
    }
    return i;
}

// stage_undo_resolver_t over the table's EphemeralPool (what BTree::Read does, b_tree.cpp:2087-2103); ctx = ref_table*
int ref_resolve_undo(void *ctx, uint64_t copy_location, uint64_t *next_tuple_header, const void **old_record) {
    ref_table *t = static_cast<ref_table *>(ctx);
    auto hd = t->undo_pool->GetOversionHeader(copy_location);
    if (hd == nullptr) return 0;
    *next_tuple_header = t->undo_pool->GetNext(copy_location);
    *old_record = reinterpret_cast<const void *>(copy_location);
    return 1;
}

const void *ref_root(ref_table *t) { return t->tree->GetRootNodeSafe(); }
uint64_t ref_tuple_count(ref_table *t) { return t->tree->GetTupleCount(); }

// ---------------------------------------------------------------------------------------------
// reads
// ---------------------------------------------------------------------------------------------
// returns 1 if a visible version was found; writes commit id, key (key_size bytes), payload.
static inline int visible_version(Record *rec, uint32_t key_size, uint32_t payload_size, uint32_t read_ts,
                                  uint32_t *ts_out, char *key_out, char *payload_out) {
    uint32_t cid = rec->meta.meta_data.GetTxnCommitId();
    if (read_ts >= cid) {                                   // executor.h:385 / :469
        if (ts_out) *ts_out = cid;
        if (key_out) memcpy(key_out, rec->GetKey(), key_size);
        if (payload_out) memcpy(payload_out, rec->GetPayload(), payload_size);
        return 1;
    }
    // The executors follow RecordMeta::next_tuple_ptr (executor.h:407 point, :489 range). BTree::Read fills it
    // (b_tree.cpp:2099/2116); the records Iterator::GetNext hands out leave it 0 (record_meta.h:240), so a range
    // scan never walks a chain: a row whose in-leaf version is newer than the reader is dropped.
    uint64_t next = rec->meta.next_tuple_ptr;
    while (next != 0 && next != std::numeric_limits<uint64_t>::max()) {
        TupleHeader *hd = reinterpret_cast<TupleHeader *>(next);
        cid_t begn = hd->GetBeginId(), comm = hd->GetEndCommitId();
        // executor.h:418/:502 stop on INVALID_CID (==0). NB a loader-created version has begin 0, so
        // the reference can never return the original (commit id 0) version through a chain.
        if (begn == INVALID_CID || comm == INVALID_CID) return 0;
        if (read_ts >= begn && read_ts <= comm) {           // executor.h:422 / :505
            const char *slot = reinterpret_cast<const char *>(hd->GetTupleSlot());
            if (ts_out) *ts_out = begn;
            if (key_out) memcpy(key_out, slot, key_size);
            if (payload_out) memcpy(payload_out, slot + key_size, payload_size);
            return 1;
        }
        next = hd->GetNextHeaderPointer();
    }
    return 0;
}

int ref_probe(ref_table *t, const void *keys, uint32_t key_size, uint64_t n, uint32_t read_ts, uint8_t *found,
              uint32_t *ts, void *key_out, void *payload_out) {
    const char *k = static_cast<const char *>(keys);
    const uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; ++i) {
        std::unique_ptr<Record> rec = t->tree->Read(k + i * key_size, (uint16_t) key_size, read_ts, false);
        int f = 0;
        if (rec != nullptr) {
            f = visible_version(rec.get(), key_size, P, read_ts, ts ? ts + i : nullptr,
                                key_out ? static_cast<char *>(key_out) + i * key_size : nullptr,
                                payload_out ? static_cast<char *>(payload_out) + i * (uint64_t) P : nullptr);
            delete[] reinterpret_cast<char *>(rec.release());
        }
        if (found) found[i] = (uint8_t) f;
        if (!f) {
            if (ts) ts[i] = 0;
            if (key_out) memset(static_cast<char *>(key_out) + i * key_size, 0, key_size);
            if (payload_out) memset(static_cast<char *>(payload_out) + i * (uint64_t) P, 0, P);
        }
    }
    return 0;
}

// One scan: executor loop over Iterator::GetNext. Rows go to out row index base .. base+count.
// MVCC scan through the UNMODIFIED IndexScanExecutor (range branch). The predicate sees every row the executor
// returns, while the Record behind it is still alive.
static inline uint32_t scan_one_executor(ref_table *t, const char *start_key, uint32_t key_size, uint32_t scan_size,
                                         uint32_t read_ts, uint32_t *ts_out, char *key_out, char *payload_out) {
    const uint32_t P = t->payload_size;
    TransactionContext txn(0, IsolationLevelType::SERIALIZABLE, read_ts, read_ts);
    txn.SetReadOnly();                                       // PerformRead returns at once (transaction_manager.cpp:364)
    uint32_t cnt = 0;
    const char *k = start_key;
    IndexScanExecutor<const char *, char> ex(
        t->tree, (uint16_t) key_size, k, scan_size, false,
        [&](const char *data, bool &) {
            // data == Record::GetData() == tuple_data_ + sizeof(cid_t) (b_tree.h:496-498)
            const Record *r = reinterpret_cast<const Record *>(data - sizeof(cid_t) - offsetof(Record, tuple_data_));
            if (ts_out) ts_out[cnt] = r->meta.meta_data.GetTxnCommitId();
            if (key_out) memcpy(key_out + (uint64_t) cnt * key_size, data, key_size);
            if (payload_out) memcpy(payload_out + (uint64_t) cnt * P, data + r->meta.meta_data.GetPaddedKeyLength(), P);
            ++cnt;
            return true;
        },
        false, &txn, nullptr);
    ex.Execute();
    return cnt;
}

static inline uint32_t scan_one(ref_table *t, const char *start_key, uint32_t key_size, uint32_t scan_size,
                                uint32_t read_ts, int apply_visibility, uint32_t *ts_out, char *key_out,
                                char *payload_out) {
    const uint32_t P = t->payload_size;
    auto iter = t->tree->RangeScanBySize(start_key, (uint16_t) key_size, scan_size);
    uint32_t cnt = 0;
    for (uint32_t scanned = 0; scanned < scan_size; ++scanned) {
        std::unique_ptr<Record> rec = iter->GetNext();
        if (rec == nullptr) continue;                        // executor.h:461 / :629
        int f;
        if (apply_visibility) {
            f = visible_version(rec.get(), key_size, P, read_ts, ts_out ? ts_out + cnt : nullptr,
                                key_out ? key_out + (uint64_t) cnt * key_size : nullptr,
                                payload_out ? payload_out + (uint64_t) cnt * P : nullptr);
        } else {                                             // TableScanExecutor: no visibility test
            if (ts_out) ts_out[cnt] = rec->meta.meta_data.GetTxnCommitId();
            if (key_out) memcpy(key_out + (uint64_t) cnt * key_size, rec->GetKey(), key_size);
            if (payload_out) memcpy(payload_out + (uint64_t) cnt * P, rec->GetPayload(), P);
            f = 1;
        }
        cnt += (uint32_t) f;
        delete[] reinterpret_cast<char *>(rec.release());
    }
    return cnt;
}

int ref_scan(ref_table *t, const void *start_keys, uint32_t key_size, const uint32_t *scan_sizes, uint64_t n,
             uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
             uint32_t *ts_out, void *payload_out) {
    const char *k = static_cast<const char *>(start_keys);
    const uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; ++i) {
        uint64_t base = i * (uint64_t) out_stride;
        uint32_t *to = ts_out ? ts_out + base : nullptr;
        char *ko = key_out ? static_cast<char *>(key_out) + base * key_size : nullptr;
        char *po = payload_out ? static_cast<char *>(payload_out) + base * P : nullptr;
        row_counts[i] = apply_visibility && scan_sizes[i]
                            ? scan_one_executor(t, k + i * key_size, key_size, scan_sizes[i], read_ts, to, ko, po)
                            : scan_one(t, k + i * key_size, key_size, scan_sizes[i], read_ts, apply_visibility, to, ko, po);
    }
    return 0;
}

// Tables whose keys differ in length (test/testing_btree.cpp's std::to_string keys): key i = keys[off[i] .. off[i+1]).
// Tree-level calls only — BTree::Read and RangeScanBySize / Iterator::GetNext take the length per call
// (b_tree.h:812-835); the executors are typed on a fixed-size key struct.
int ref_probe_varlen(ref_table *t, const void *keys, const uint32_t *off, uint64_t n, uint32_t read_ts, uint8_t *found,
                     uint32_t *ts, void *payload_out) {
    const char *k = static_cast<const char *>(keys);
    const uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; ++i) {
        const uint32_t len = off[i + 1] - off[i];
        std::unique_ptr<Record> rec = t->tree->Read(k + off[i], (uint16_t) len, read_ts, false);
        int f = 0;
        if (rec != nullptr) {
            f = visible_version(rec.get(), len, P, read_ts, ts ? ts + i : nullptr, nullptr,
                                payload_out ? static_cast<char *>(payload_out) + i * (uint64_t) P : nullptr);
            delete[] reinterpret_cast<char *>(rec.release());
        }
        found[i] = (uint8_t) f;
    }
    return 0;
}

int ref_scan_varlen(ref_table *t, const void *keys, const uint32_t *off, const uint32_t *scan_sizes, uint64_t n,
                    uint32_t out_stride, uint32_t read_ts, int apply_visibility, uint32_t *row_counts, void *key_out,
                    uint32_t key_stride, uint32_t *key_lens, uint32_t *ts_out, void *payload_out) {
    const char *k = static_cast<const char *>(keys);
    const uint32_t P = t->payload_size;
    for (uint64_t i = 0; i < n; ++i) {
        const uint64_t base = i * (uint64_t) out_stride;
        auto iter = t->tree->RangeScanBySize(k + off[i], (uint16_t) (off[i + 1] - off[i]), scan_sizes[i]);
        uint32_t cnt = 0;
        for (uint32_t scanned = 0; scanned < scan_sizes[i]; ++scanned) {
            std::unique_ptr<Record> rec = iter->GetNext();
            if (rec == nullptr) continue;
            const uint32_t cid = rec->meta.meta_data.GetTxnCommitId();
            if (!apply_visibility || read_ts >= cid) {       // executor.h:469; the range branch never walks a chain
                const uint32_t len = rec->meta.meta_data.GetKeyLength();
                key_lens[base + cnt] = len;
                if (ts_out) ts_out[base + cnt] = cid;
                if (key_out) memcpy(static_cast<char *>(key_out) + (base + cnt) * key_stride, rec->GetKey(), len < key_stride ? len : key_stride);
                if (payload_out) memcpy(static_cast<char *>(payload_out) + (base + cnt) * P, rec->GetPayload(), P);
                ++cnt;
            }
            delete[] reinterpret_cast<char *>(rec.release());
        }
        row_counts[i] = cnt;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------------
// timing legs (CPU baseline): T threads, each over its slice of the same key stream, results copied
// into per-thread scratch (the reference allocates a Record per row; that cost is part of its path).
// Returns elapsed seconds; *rows_out = total rows returned.
// ---------------------------------------------------------------------------------------------
static void pin_to_core(int c) {
    cpu_set_t set;
    CPU_ZERO(&set);
    CPU_SET(c, &set);
    pthread_setaffinity_np(pthread_self(), sizeof(set), &set);   // like PinToCore, ycsb_workload.cpp:41-46
}

double ref_time_probe(ref_table *t, const void *keys, uint32_t key_size, uint64_t n, uint32_t read_ts,
                      int threads, int pin, uint64_t *hits_out) {
    std::vector<std::thread> th;
    std::vector<uint64_t> hits(threads, 0);
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    const char *k = static_cast<const char *>(keys);
    for (int w = 0; w < threads; ++w) {
        th.emplace_back([&, w] {
            if (pin) pin_to_core(w);
            std::vector<char> pay(t->payload_size), key(key_size);
            uint64_t b = n * w / threads, e = n * (w + 1) / threads, h = 0;
            ready++;
            while (!go.load()) {}
            for (uint64_t i = b; i < e; ++i) {
                std::unique_ptr<Record> rec = t->tree->Read(k + i * key_size, (uint16_t) key_size, read_ts, false);
                if (rec != nullptr) {
                    uint32_t ts;
                    h += visible_version(rec.get(), key_size, t->payload_size, read_ts, &ts, key.data(), pay.data());
                    delete[] reinterpret_cast<char *>(rec.release());
                }
            }
            hits[w] = h;
        });
    }
    while (ready.load() < threads) {}
    auto t0 = std::chrono::steady_clock::now();
    go.store(true);
    for (auto &x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    uint64_t tot = 0;
    for (auto h : hits) tot += h;
    if (hits_out) *hits_out = tot;
    return std::chrono::duration<double>(t1 - t0).count();
}

double ref_time_scan(ref_table *t, const void *start_keys, uint32_t key_size, uint32_t scan_size, uint64_t n,
                     uint32_t read_ts, int apply_visibility, int threads, int pin, uint64_t *rows_out) {
    std::vector<std::thread> th;
    std::vector<uint64_t> rows(threads, 0);
    std::atomic<int> ready{0};
    std::atomic<bool> go{false};
    const char *k = static_cast<const char *>(start_keys);
    for (int w = 0; w < threads; ++w) {
        th.emplace_back([&, w] {
            if (pin) pin_to_core(w);
            std::vector<char> pay((uint64_t) t->payload_size * scan_size), key((uint64_t) key_size * scan_size);
            std::vector<uint32_t> ts(scan_size);
            uint64_t b = n * w / threads, e = n * (w + 1) / threads, r = 0;
            ready++;
            while (!go.load()) {}
            // MVCC scans run the reference's own IndexScanExecutor (range branch), rows copied out in its predicate
            // callback — the same code the parity tests compare against; unfiltered scans run the restated
            // TableScanExecutor loop over the unmodified BTree::RangeScanBySize / Iterator::GetNext.
            for (uint64_t i = b; i < e; ++i)
                r += apply_visibility && scan_size
                         ? scan_one_executor(t, k + i * key_size, key_size, scan_size, read_ts, ts.data(), key.data(), pay.data())
                         : scan_one(t, k + i * key_size, key_size, scan_size, read_ts, apply_visibility, ts.data(), key.data(),
                                    pay.data());
            rows[w] = r;
        });
    }
    while (ready.load() < threads) {}
    auto t0 = std::chrono::steady_clock::now();
    go.store(true);
    for (auto &x : th) x.join();
    auto t1 = std::chrono::steady_clock::now();
    uint64_t tot = 0;
    for (auto r : rows) tot += r;
    if (rows_out) *rows_out = tot;
    return std::chrono::duration<double>(t1 - t0).count();
}

}  // extern "C"
