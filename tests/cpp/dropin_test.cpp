// tests/cpp/dropin_test.cpp — drop-in test of the C++ adapters (include/stage_gpu_executors.h) against the
// UNMODIFIED reference classes, in one process, on one tree.
//
// The table is built with the reference's own BTree exactly as its loaders do (Insert + FinalizeInsert,
// benchmark/ycsb/ycsb_loader.cpp:93-173). Then, with a real SSNTransactionManager / TransactionContext:
//   * mvstore::IndexScanExecutor<const char*, T>  (point_lookup)   vs  stage_gpu::GpuIndexScanExecutor
//   * mvstore::TableScanExecutor<const char*, T>  (range, by value) vs  stage_gpu::GpuTableScanExecutor
//   * mvstore::BTree::Read / RangeScanBySize / Iterator::GetNext    vs  stage_gpu::GpuTreeView / GpuIterator
// are run on the same key streams (the drivers' FastRandom / ZipfDistribution generators,
// benchmark/benchmark_common.h:12-105) and every returned struct is compared byte for byte. The
// reference's *range* IndexScanExecutor keeps pointers into Records it has already freed
// (include/execute/executor.h:476-484), so its by-value sibling TableScanExecutor is the range checker.
//
// Built only where /root/reference exists (oracle/Makefile.ref -> oracle/_ref/dropin_test, which travels to
// the GPU box); run by tests/test_gpu_dropin_cpp.py. `--ref-only` runs the reference half alone (no GPU).
// TEST CODE: the reference here is the checker, never the product.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <thread>
#include <atomic>
#include <chrono>

#include "execute/executor.h"
#include "benchmark_common.h"
#include "../../include/stage_gpu_executors.h"

using namespace mvstore;

// the tuple shapes the YCSB driver instantiates the executors with (benchmark/ycsb/ycsb_configuration.h:22-58)
struct TupleInt {
    uint32_t key;
    char cols[10][100];
};
struct TupleStr {
    char key[16];
    char cols[10][100];
};
// TPC-C stock shape: 16-byte key, 400-byte payload region (benchmark/tpcc/tpcc_record.h:417-510)
struct TupleStock {
    char key[16];
    char body[400];
};

static int g_fail = 0;
static EphemeralPool *g_undo = nullptr;      // pool of the table made last (make_table)
#define EXPECT(cond, ...)                                   \
    do {                                                    \
        if (!(cond)) {                                      \
            if (g_fail < 20) { printf("FAIL %s:%d: ", __FILE__, __LINE__); printf(__VA_ARGS__); printf("\n"); } \
            ++g_fail;                                       \
        }                                                   \
    } while (0)

struct Table {
    BTree *tree;
    Catalog *catalog;
    uint32_t K, P;
};

static Table make_table(uint32_t K, uint32_t P, uint32_t split, uint64_t blocks) {
    Table t{};
    t.K = K;
    t.P = P;
    t.catalog = new Catalog();
    uint8_t *bitmap = new uint8_t[2]{1, 0};
    t.catalog->init("table0", 2, bitmap, K, 1, false);
    t.catalog->add_col("Key", K, "VARCHAR");
    t.catalog->add_col("payload", P, "VARCHAR");
    ParameterSet param(split, 32 * 1024, 64 * 1024, P);
    auto *leaf_pool = new DramBlockPool(blocks, blocks);
    auto *inner = new InnerNodeBuffer(new RecordBufferPool(5000000, 5000000));
    auto *undo = new EphemeralPool(new UndoBuffer(new RecordBufferPool(5000000, 5000000)));
    g_undo = undo;
    SSNTransactionManager::GetInstance()->Init(undo);
    t.tree = new BTree(param, *t.catalog, leaf_pool, inner, undo);
    return t;
}

static void insert_row(Table &t, const char *key, const char *payload) {
    RecordMeta meta;
    ReturnCode rc = t.tree->Insert(key, (uint16_t) t.K, payload, &meta, INVALID_CID);
    if (!rc.IsOk()) return;
    auto *mp = reinterpret_cast<RecordMetadata *>(meta.GetMetaData().GetLocationPtr()->record_meta_ptr);
    t.tree->FinalizeInsert(mp, INVALID_CID);
    t.tree->IncreaseTupleCount(1);
}

template <class T>
static int run_checks(Table &t, const std::vector<std::string> &probe_keys, const std::vector<std::string> &scan_keys,
                      uint32_t scan_size, bool ref_only, const char *what, stage_gpu::GpuTreeView *reuse = nullptr) {
    auto *mgr = SSNTransactionManager::GetInstance();
    std::unique_ptr<stage_gpu::GpuTreeView> owned;
    if (!ref_only && !reuse) owned.reset(new stage_gpu::GpuTreeView(t.tree->GetRootNodeSafe(), t.K, t.P));
    struct ViewRef {                       // the checks below use view-> / view.get() for either kind of view
        stage_gpu::GpuTreeView *p;
        stage_gpu::GpuTreeView *operator->() const { return p; }
        stage_gpu::GpuTreeView *get() const { return p; }
    } view{reuse ? reuse : owned.get()};
    uint64_t hits = 0, rows = 0, crc = 0;
    const int fail0 = g_fail;

    // --- point lookups through the executors ------------------------------------------------------------
    for (const std::string &k : probe_keys) {
        TransactionContext *txn = mgr->BeginTransaction(0, IsolationLevelType::SERIALIZABLE);
        txn->SetReadOnly();
        IndexScanExecutor<const char *, T> ref(t.tree, (uint16_t) t.K, k.data(), 1, true, nullptr, false, txn, nullptr);
        EXPECT(ref.Execute(), "reference Execute");
        const T *want = ref.GetRecord();
        if (want) { ++hits; for (size_t i = 0; i < sizeof(T); i += 97) crc = crc * 131 + ((const uint8_t *) want)[i]; }
        if (!ref_only) {
            stage_gpu::GpuIndexScanExecutor<const char *, T, TransactionContext> gpu(view.get(), (uint16_t) t.K, k.data(), 1, true,
                                                                                    nullptr, false, txn, nullptr);
            EXPECT(gpu.Execute(), "gpu Execute");
            const T *got = gpu.GetRecord();
            EXPECT((got == nullptr) == (want == nullptr), "%s point: found mismatch", what);
            if (got && want) EXPECT(memcmp(got, want, sizeof(T)) == 0, "%s point: tuple bytes differ", what);
            // tree-level shape: BTree::Read vs GpuTreeView::Read (latest in-leaf version, cstamp = reader id)
            std::unique_ptr<Record> r = t.tree->Read(k.data(), (uint16_t) t.K, txn->GetReadId(), false);
            std::unique_ptr<stage_gpu::Record> g = view->Read(k.data(), (uint16_t) t.K, txn->GetReadId(), false);
            EXPECT((r == nullptr) == (g == nullptr), "%s Read: found mismatch", what);
            if (r && g) {
                EXPECT(memcmp(r->GetData(), g->GetData(), r->GetKeyPaddedLength() + t.P) == 0, "%s Read: data differs", what);
                EXPECT(r->GetCstamp() == g->GetCstamp(), "%s Read: cstamp", what);
                EXPECT(r->meta.meta_data.GetTxnCommitId() == g->commit_id, "%s Read: commit id", what);
                EXPECT(r->GetKeyPaddedLength() == g->GetKeyPaddedLength(), "%s Read: padded key length", what);
            }
            if (r) delete[] reinterpret_cast<char *>(r.release());
        }
        delete want;
        mgr->CommitTransaction(txn);
    }

    // --- range scans through the executors and the iterator ------------------------------------------------
    for (const std::string &k : scan_keys) {
        TransactionContext *txn = mgr->BeginTransaction(0, IsolationLevelType::SERIALIZABLE);
        txn->SetReadOnly();
        TableScanExecutor<const char *, T> ref(t.tree, k.data(), t.K, scan_size, txn, nullptr);
        EXPECT(ref.Execute(), "reference scan Execute");
        std::vector<T> &want = ref.GetResults();
        rows += want.size();
        if (!ref_only) {
            stage_gpu::GpuTableScanExecutor<const char *, T, TransactionContext> gpu(view.get(), k.data(), t.K, scan_size, txn,
                                                                                    nullptr);
            EXPECT(gpu.Execute(), "gpu scan Execute");
            std::vector<T> &got = gpu.GetResults();
            EXPECT(got.size() == want.size(), "%s scan: %zu rows, reference %zu", what, got.size(), want.size());
            EXPECT(gpu.GetFailNum() == ref.GetFailNum(), "%s scan: fail_num %u vs %u", what, gpu.GetFailNum(), ref.GetFailNum());
            if (got.size() == want.size())
                for (size_t i = 0; i < got.size(); ++i)
                    EXPECT(memcmp(&got[i], &want[i], sizeof(T)) == 0, "%s scan: row %zu differs", what, i);
            // IndexScanExecutor range shape with a predicate (called once per visible row, result ignored)
            uint32_t pred_calls = 0;
            stage_gpu::GpuIndexScanExecutor<const char *, T, TransactionContext> gi(
                view.get(), (uint16_t) t.K, k.data(), scan_size, false,
                [&](const T *, bool &) { ++pred_calls; return true; }, false, txn, nullptr);
            EXPECT(gi.Execute(), "gpu index range Execute");
            EXPECT(gi.GetResults().size() == want.size() && pred_calls == want.size(), "%s index range: row count", what);
            for (size_t i = 0; i < gi.GetResults().size() && i < want.size(); ++i)
                EXPECT(memcmp(gi.GetResults()[i], &want[i], sizeof(T)) == 0, "%s index range: row %zu differs", what, i);
            // iterator shape
            auto it_ref = t.tree->RangeScanBySize(k.data(), (uint16_t) t.K, scan_size);
            auto it_gpu = view->RangeScanBySize(k.data(), (uint16_t) t.K, scan_size);
            for (uint32_t s = 0; s < scan_size; ++s) {
                std::unique_ptr<Record> r = it_ref->GetNext();
                if (!r) continue;                                      // executor.h:629-632: a null GetNext is skipped
                std::unique_ptr<stage_gpu::Record> g = it_gpu->GetNext();
                EXPECT(g != nullptr, "%s iterator: gpu ran out at step %u", what, s);
                if (g) EXPECT(memcmp(r->GetData(), g->GetData(), r->GetKeyPaddedLength() + t.P) == 0, "%s iterator: row differs", what);
                delete[] reinterpret_cast<char *>(r.release());
            }
            EXPECT(it_gpu->GetNext() == nullptr, "%s iterator: gpu has extra rows", what);
        }
        mgr->CommitTransaction(txn);
    }
    // --- physical whole-tree scan: TableScanExecutor with scan_sz == -1 (executor.h:581-619) ------------------------
    if (!scan_keys.empty()) {
        TransactionContext *txn = mgr->BeginTransaction(0, IsolationLevelType::SERIALIZABLE);
        txn->SetReadOnly();
        const std::string &k = scan_keys[0];
        TableScanExecutor<const char *, T> ref(t.tree, k.data(), t.K, (uint32_t) -1, txn, nullptr);
        EXPECT(ref.Execute(), "reference physical scan");
        rows += ref.GetResults().size();
        if (!ref_only) {
            stage_gpu::GpuTableScanExecutor<const char *, T, TransactionContext> gpu(view.get(), k.data(), t.K, (uint32_t) -1, txn,
                                                                                    nullptr);
            EXPECT(gpu.Execute(), "gpu physical scan");
            EXPECT(gpu.GetResults().size() == ref.GetResults().size(), "%s physical: %zu rows, reference %zu", what,
                   gpu.GetResults().size(), ref.GetResults().size());
            if (gpu.GetResults().size() == ref.GetResults().size())
                for (size_t i = 0; i < gpu.GetResults().size(); ++i)
                    EXPECT(memcmp(&gpu.GetResults()[i], &ref.GetResults()[i], sizeof(T)) == 0, "%s physical: row %zu differs", what, i);
        }
        mgr->CommitTransaction(txn);
    }
    printf("%-12s probes %zu (hits %lu) scans %zu x %u (rows %lu) crc %016lx : %s\n", what, probe_keys.size(), hits,
           scan_keys.size(), scan_size, rows, crc, g_fail == fail0 ? "OK" : "FAILED");
    return g_fail - fail0;
}


// ---- MVCC: committed version chains and in-flight updates -------------------------------------------------------
// A committed update as SSNTransactionManager::CommitTransaction leaves it (transaction_manager.cpp:610-676): the
// old version retired into a TupleHeader chain, the record overwritten in place, meta = new commit id.
static void commit_update(Table &t, uint32_t key, int fill, uint32_t cid, std::vector<char *> &allocs) {
    LeafNode *leaf = t.tree->TraverseToLeaf(nullptr, reinterpret_cast<const char *>(&key), (uint16_t) t.K);
    RecordMetadata *mp = nullptr;
    if (!leaf || !leaf->Read(reinterpret_cast<const char *>(&key), (uint16_t) t.K, t.P, &mp, false).IsOk()) return;
    char *rk = nullptr, *rp = nullptr;
    leaf->GetRawRecord(*mp, &rk, &rp);
    char *mem = new char[sizeof(TupleHeader) + t.K + t.P];
    allocs.push_back(mem);
    TupleHeader *hd = new (mem) TupleHeader();
    char *slot = mem + sizeof(TupleHeader);
    hd->init(hd, reinterpret_cast<uint64_t>(slot), (oid_t) allocs.size());
    memcpy(slot, rk, t.K);
    memcpy(slot + t.K, rp, t.P);
    hd->SetBeginId(mp->GetTxnCommitId());
    hd->SetCommitId(cid);
    hd->SetNextHeaderPointer(mp->GetNextPointer());
    memset(rp + 100, fill, 100);
    mp->FinalizeForUpdate(cid);
    mp->SetNextPointer(reinterpret_cast<uint64_t>(hd));
}

static int resolve_undo(void *ctx, uint64_t copy_location, uint64_t *next_tuple_header, const void **old_record) {
    EphemeralPool *pool = static_cast<EphemeralPool *>(ctx);
    if (pool->GetOversionHeader(copy_location) == nullptr) return 0;
    *next_tuple_header = pool->GetNext(copy_location);
    *old_record = reinterpret_cast<const void *>(copy_location);
    return 1;
}

static void run_mvcc_checks(bool ref_only, uint32_t n_rows) {
    const int fail0 = g_fail;
    Table t = make_table(4, 1000, 16 * 1024, n_rows / 16 + 4096);
    std::vector<char> payload(1000);
    for (uint32_t rowid = 0; rowid < n_rows; ++rowid) {
        memset(payload.data(), (int) rowid, payload.size());
        insert_row(t, reinterpret_cast<const char *>(&rowid), payload.data());
    }
    benchmark::FastRandom rng(11);
    std::vector<char *> allocs;
    uint32_t cid = 10;
    for (uint32_t i = 0; i < n_rows / 4; ++i) commit_update(t, (uint32_t) (rng.next() % n_rows), (int) cid, cid, allocs), ++cid;   // many rows twice or more
    const uint32_t last_cid = cid - 1;
    std::unique_ptr<stage_gpu::GpuTreeView> view;
    if (!ref_only) view.reset(new stage_gpu::GpuTreeView(t.tree->GetRootNodeSafe(), t.K, t.P));

    // (1) IndexScanExecutor RANGE branch over version chains, at read ids straddling the commit ids. The rows are
    // captured in the predicate (the reference's result vector points into freed Records). This pins the rule that a
    // range scan never walks a chain: a row whose in-leaf version is newer than the reader is dropped.
    uint64_t rows = 0, dropped = 0;
    for (uint32_t read_id : {0u, 9u, 10u, 10u + n_rows / 16, 10u + n_rows / 8, last_cid - 1, last_cid, 0x7FFFFFFFu}) {
        for (uint32_t s = 0; s < 24; ++s) {
            const uint32_t start = (uint32_t) (rng.next() % n_rows), scan_size = s % 3 ? 100 : 400;
            TransactionContext txn(0, IsolationLevelType::SERIALIZABLE, read_id, read_id);
            txn.SetReadOnly();
            std::vector<TupleInt> want, got;
            IndexScanExecutor<const char *, TupleInt> ref(
                t.tree, (uint16_t) t.K, reinterpret_cast<const char *>(&start), scan_size, false,
                [&](const TupleInt *r, bool &) { want.push_back(*r); return true; }, false, &txn, nullptr);
            EXPECT(ref.Execute(), "reference range Execute");
            rows += want.size();
            dropped += scan_size - want.size();
            if (ref_only) continue;
            stage_gpu::GpuIndexScanExecutor<const char *, TupleInt, TransactionContext> gpu(
                view.get(), (uint16_t) t.K, reinterpret_cast<const char *>(&start), scan_size, false,
                [&](const TupleInt *r, bool &) { got.push_back(*r); return true; }, false, &txn, nullptr);
            EXPECT(gpu.Execute(), "gpu range Execute");
            EXPECT(got.size() == want.size(), "range over chains (read id %u): %zu rows, reference %zu", read_id, got.size(), want.size());
            if (got.size() == want.size())
                for (size_t i = 0; i < got.size(); ++i)
                    EXPECT(memcmp(&got[i], &want[i], sizeof(TupleInt)) == 0, "range over chains (read id %u): row %zu differs", read_id, i);
        }
    }
    printf("%-12s range scans over version chains: %lu rows returned, %lu dropped as too new : %s\n", "mvcc", rows, dropped,
           g_fail == fail0 ? "OK" : "FAILED");

    // (2) updates IN FLIGHT (LeafNode::Update without the finalize, b_tree.cpp:1061-1163): BTree::Read serves the undo
    // copy; GpuTreeView built with the undo resolver must return the same Record bytes, commit id and cstamp.
    std::vector<uint32_t> flying;
    std::vector<oid_t> columns{1};
    std::vector<char> delta(1000, (char) 0x5A);
    for (uint32_t i = 0; i < 200; ++i) {
        const uint32_t k = (uint32_t) (rng.next() % n_rows);
        if ((k & 0xFF) == 0x5A) continue;                      // the row's committed bytes must differ from the delta
        RecordMeta meta;
        if (t.tree->Update(reinterpret_cast<const char *>(&k), (uint16_t) t.K, delta.data(), columns, &meta, false, 0x70000000u + i).IsOk())
            flying.push_back(k);
    }
    EXPECT(flying.size() > 100, "in-flight updates were not accepted");
    std::unique_ptr<stage_gpu::GpuTreeView> live;
    if (!ref_only) {
        stage_snapshot_t *refused = nullptr;
        stage_schema_t sc{t.K, t.P, 65536, STAGE_KEY_AUTO};
        EXPECT(stage_gpu_snapshot_from_tree(t.tree->GetRootNodeSafe(), &sc, &refused) == STAGE_ERR_NOT_QUIESCED,
               "plain from_tree must refuse a tree with writers in flight");
        live.reset(new stage_gpu::GpuTreeView(t.tree->GetRootNodeSafe(), t.K, t.P, resolve_undo, g_undo));
    }
    uint32_t checked = 0;
    for (uint32_t i = 0; i < flying.size() + 300; ++i) {
        const uint32_t k = i < flying.size() ? flying[i] : (uint32_t) (rng.next() % (n_rows + 10));
        const cid_t reader = 0x7FFFFFFFu;
        std::unique_ptr<Record> r = t.tree->Read(reinterpret_cast<const char *>(&k), (uint16_t) t.K, reader, false);
        if (i < flying.size()) EXPECT(r != nullptr && (unsigned char) r->GetPayload()[500] != 0x5A, "reference serves uncommitted bytes for key %u", k);
        if (live) {
            std::unique_ptr<stage_gpu::Record> g = live->Read(reinterpret_cast<const char *>(&k), (uint16_t) t.K, reader, false);
            EXPECT((r == nullptr) == (g == nullptr), "in-flight Read: found mismatch for key %u", k);
            if (r && g) {
                // Record::Neww (b_tree.h:430-448) leaves the key's padding bytes uninitialised: compare key and payload
                EXPECT(memcmp(r->GetKey(), g->GetData(), t.K) == 0, "in-flight Read: key %u key bytes differ", k);
                EXPECT(memcmp(r->GetPayload(), g->GetPayload(), t.P) == 0, "in-flight Read: key %u payload differs", k);
                EXPECT(r->meta.meta_data.GetTxnCommitId() == g->commit_id, "in-flight Read: key %u commit id", k);
                ++checked;
            }
        }
        if (r) delete[] reinterpret_cast<char *>(r.release());
    }
    printf("%-12s %zu updates in flight, %u reads compared : %s\n", "mvcc", flying.size(), checked, g_fail == fail0 ? "OK" : "FAILED");
}

// The reference's driver loop shape (benchmark/ycsb/ycsb_mixed.cpp:18-110 RunMixed with -u 0: `ops` point reads per
// transaction, one IndexScanExecutor per read, N worker threads as in ycsb_workload.cpp:522-529) through the GPU adapters,
// first as literal batches of one, then with the batching front-end coalescing the threads' operations. Every tuple is
// compared with the reference executor's; the rates are printed for profiles/README.md.
static void run_driver_shape(Table &t, uint32_t n_rows, int threads, int ops_per_txn, int txns_per_thread) {
    stage_gpu::GpuTreeView view(t.tree->GetRootNodeSafe(), t.K, t.P);
    const int fail0 = g_fail;
    for (int pass = 0; pass < 3; ++pass) {                // 0: reference executors, 1: GPU adapters, 2: GPU adapters + batching
        if (pass == 2) view.EnableBatching(256, 50);
        std::atomic<uint64_t> hits{0}, mism{0};
        std::vector<std::thread> th;
        const auto t0 = std::chrono::steady_clock::now();
        for (int w = 0; w < threads; ++w)
            th.emplace_back([&, w] {
                benchmark::FastRandom rng(1000 + w);
                uint64_t h = 0, bad = 0;
                for (int x = 0; x < txns_per_thread; ++x) {
                    // a read-only context built directly (the manager's transaction table is a stub container here and
                    // not meant for concurrent BeginTransaction calls; reads only need the read id)
                    TransactionContext txn_obj((size_t) w, IsolationLevelType::SERIALIZABLE, 100 + x, 100 + x);
                    TransactionContext *txn = &txn_obj;
                    txn->SetReadOnly();
                    for (int o = 0; o < ops_per_txn; ++o) {
                        uint32_t k = (uint32_t) (1 + rng.next() % n_rows);              // keys 1..N like the driver: key N misses
                        const char *kp = reinterpret_cast<const char *>(&k);
                        if (pass == 0) {
                            IndexScanExecutor<const char *, TupleInt> ref(t.tree, (uint16_t) t.K, kp, 1, true, nullptr, false, txn, nullptr);
                            ref.Execute();
                            const TupleInt *r = ref.GetRecord();
                            h += r != nullptr;
                            delete r;
                        } else {
                            stage_gpu::GpuIndexScanExecutor<const char *, TupleInt, TransactionContext> gpu(&view, (uint16_t) t.K, kp, 1, true,
                                                                                                           nullptr, false, txn, nullptr);
                            gpu.Execute();
                            const TupleInt *g = gpu.GetRecord();
                            h += g != nullptr;
                            if (x % 8 == 0) {                                                // spot-check against the reference
                                IndexScanExecutor<const char *, TupleInt> ref(t.tree, (uint16_t) t.K, kp, 1, true, nullptr, false, txn, nullptr);
                                ref.Execute();
                                const TupleInt *r = ref.GetRecord();
                                bad += (r == nullptr) != (g == nullptr) || (r && g && memcmp(r, g, sizeof(TupleInt)) != 0);
                                delete r;
                            }
                        }
                    }
                }
                hits += h;
                mism += bad;
            });
        for (auto &x : th) x.join();
        const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        const double ops = (double) threads * txns_per_thread * ops_per_txn;
        uint64_t nb = 0, no = 0;
        view.BatchingStats(&nb, &no);
        printf("driver-shape  %-34s %2d threads x %d reads/txn: %10.0f reads/s (hits %lu%s)\n",
               pass == 0 ? "reference IndexScanExecutor" : pass == 1 ? "GPU adapter, batches of one" : "GPU adapter + batching front-end",
               threads, ops_per_txn, ops / sec, (unsigned long) hits.load(),
               pass == 2 ? (", " + std::to_string(no) + " reads in " + std::to_string(nb) + " launches").c_str() : "");
        EXPECT(mism.load() == 0, "driver shape pass %d: tuples differ from the reference", pass);
    }
    printf("%-12s %s\n", "driver-shape", g_fail == fail0 ? "OK" : "FAILED");
}

// Keys of different lengths in one tree — the fixture of test/testing_btree.cpp (std::to_string keys, :393-440):
// BTree::Read / RangeScanBySize / Iterator::GetNext against GpuTreeView built with STAGE_KEY_VARLEN (slot of 8 bytes).
static void run_varlen(bool ref_only, uint32_t n_keys) {
    Table t = make_table(8, 8, 16 * 1024, 4096);
    auto insert = [&](const std::string &k, uint64_t v) {
        RecordMeta meta;
        ReturnCode rc = t.tree->Insert(k.data(), (uint16_t) k.size(), reinterpret_cast<const char *>(&v), &meta, INVALID_CID);
        if (!rc.IsOk()) return;
        auto *mp = reinterpret_cast<RecordMetadata *>(meta.GetMetaData().GetLocationPtr()->record_meta_ptr);
        t.tree->FinalizeInsert(mp, INVALID_CID);
    };
    for (uint32_t i = 0; i < n_keys; ++i) insert(std::to_string((i * 7919u) % n_keys), (i * 7919u) % n_keys);   // shuffled order
    if (ref_only) return;
    stage_gpu::GpuTreeView view(t.tree->GetRootNodeSafe(), 8, 8, 65536, STAGE_KEY_VARLEN);
    const int fail0 = g_fail;
    std::vector<std::string> probes;
    for (uint32_t i = 0; i < n_keys; i += 3) probes.push_back(std::to_string(i));
    for (const char *k : {"", "abc", "00", "1x", "99999999", "-1"}) probes.push_back(k);     // misses ("99999999": too long for the slot)
    probes.push_back(std::to_string(n_keys));
    uint64_t hits = 0;
    for (const std::string &k : probes) {
        std::unique_ptr<Record> r = t.tree->Read(k.data(), (uint16_t) k.size(), 5, false);
        std::unique_ptr<stage_gpu::Record> g = view.Read(k.data(), (uint16_t) k.size(), 5, false);
        EXPECT((r == nullptr) == (g == nullptr), "varlen Read '%s': found mismatch", k.c_str());
        if (r && g) {
            ++hits;
            EXPECT(r->GetKeyLength() == g->GetKeyLength() && r->GetKeyPaddedLength() == g->GetKeyPaddedLength(), "varlen Read: key length");
            EXPECT(memcmp(r->GetKey(), g->GetKey(), k.size()) == 0, "varlen Read: key bytes");
            EXPECT(memcmp(r->GetPayload(), g->GetPayload(), 8) == 0, "varlen Read '%s': payload", k.c_str());
            EXPECT(r->GetCstamp() == g->GetCstamp(), "varlen Read: cstamp");
        }
        if (r) delete[] reinterpret_cast<char *>(r.release());
    }
    uint64_t rows = 0;
    for (const char *start : {"9000", "", "1", "10", "5555", "99", "999999", ":"}) {
        for (uint32_t size : {1u, 100u, 1000u}) {
            const std::string k(start);
            auto it_ref = t.tree->RangeScanBySize(k.data(), (uint16_t) k.size(), size);
            auto it_gpu = view.RangeScanBySize(k.data(), (uint16_t) k.size(), size);
            EXPECT(it_gpu != nullptr, "varlen scan: no iterator");
            if (!it_gpu) continue;
            for (uint32_t s = 0; s < size; ++s) {
                std::unique_ptr<Record> r = it_ref->GetNext();
                if (!r) continue;
                std::unique_ptr<stage_gpu::Record> g = it_gpu->GetNext();
                EXPECT(g != nullptr, "varlen iterator from '%s': gpu ran out at step %u", start, s);
                if (g) {
                    ++rows;
                    EXPECT(r->GetKeyLength() == g->GetKeyLength(), "varlen iterator: key length %u vs %u", r->GetKeyLength(), g->GetKeyLength());
                    EXPECT(memcmp(r->GetKey(), g->GetKey(), r->GetKeyLength()) == 0 && memcmp(r->GetPayload(), g->GetPayload(), 8) == 0,
                           "varlen iterator from '%s': row %u differs", start, s);
                }
                delete[] reinterpret_cast<char *>(r.release());
            }
            EXPECT(it_gpu->GetNext() == nullptr, "varlen iterator from '%s': gpu has extra rows", start);
        }
    }
    printf("variable-length keys: %zu reads (%lu hits), %lu iterator rows %s\n", probes.size(), (unsigned long) hits,
           (unsigned long) rows, g_fail == fail0 ? "OK" : "MISMATCH");
}

int main(int argc, char **argv) {
    setvbuf(stdout, nullptr, _IOLBF, 0);      // a crash must not swallow the lines before it
    bool ref_only = false;
    uint32_t n_rows = 20000, n_probe = 2000, n_scan = 60;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--ref-only")) ref_only = true;
        else if (!strcmp(argv[i], "--rows") && i + 1 < argc) n_rows = (uint32_t) atoi(argv[++i]);
        else if (!strcmp(argv[i], "--probes") && i + 1 < argc) n_probe = (uint32_t) atoi(argv[++i]);
    }
    if (!ref_only) stage_gpu::check(stage_gpu_init(0), "stage_gpu_init");

    // ---- YCSB int table (ycsb_loader.cpp:144-172) and the driver's key streams ---------------------------------
    {
        Table t = make_table(4, 1000, 16 * 1024, n_rows / 16 + 4096);
        std::vector<char> payload(1000);
        for (uint32_t rowid = 0; rowid < n_rows; ++rowid) {
            memset(payload.data(), (int) rowid, payload.size());
            insert_row(t, reinterpret_cast<const char *>(&rowid), payload.data());
        }
        benchmark::FastRandom rng(1);
        benchmark::ZipfDistribution zipf(n_rows - 1, 0.2);
        std::vector<std::string> probes, scans;
        for (uint32_t i = 0; i < n_probe; ++i) {
            uint32_t k = i % 2 ? (uint32_t) (1 + rng.next() % n_rows) : (uint32_t) zipf.GetNextNumber();   // 1..N: key N misses
            probes.emplace_back(reinterpret_cast<const char *>(&k), 4);
        }
        for (uint32_t extra : {0u, n_rows, n_rows + 7, 0xFFFFFFFFu, 0x80808080u})
            probes.emplace_back(reinterpret_cast<const char *>(&extra), 4);
        for (uint32_t i = 0; i < n_scan; ++i) {
            uint32_t k = (uint32_t) (rng.next() % n_rows);
            scans.emplace_back(reinterpret_cast<const char *>(&k), 4);
        }
        run_checks<TupleInt>(t, probes, scans, 1000, ref_only, "ycsb-int");
        run_checks<TupleInt>(t, {}, scans, 10, ref_only, "ycsb-int/10");
        // structural refresh through the adapter: rows inserted after the first snapshot (leaf splits), then
        // GpuTreeView::Refresh stages only the changed leaves; all read shapes again on the refreshed view
        std::unique_ptr<stage_gpu::GpuTreeView> live;
        if (!ref_only) live.reset(new stage_gpu::GpuTreeView(t.tree->GetRootNodeSafe(), t.K, t.P));
        const uint32_t n_more = n_rows / 10;
        for (uint32_t i = 0; i < n_more; ++i) {
            const uint32_t rowid = 0x20000005u + 256u * i;           // one region of the reference key order
            memset(payload.data(), (int) i, payload.size());
            insert_row(t, reinterpret_cast<const char *>(&rowid), payload.data());
            if (i % 16 == 0) {
                probes.emplace_back(reinterpret_cast<const char *>(&rowid), 4);
                scans.emplace_back(reinterpret_cast<const char *>(&rowid), 4);
            }
        }
        if (!ref_only) {
            const stage_refresh_stats_t st = live->Refresh(t.tree->GetRootNodeSafe());
            printf("refresh: %lu leaves, %lu carried over, %lu re-transformed, %lu rows\n", (unsigned long) st.leaves_total,
                   (unsigned long) st.leaves_carried, (unsigned long) st.leaves_transformed, (unsigned long) st.rows);
            EXPECT(st.rows == (uint64_t) n_rows + n_more, "refresh: row count");
            EXPECT(st.leaves_carried > st.leaves_total / 2, "refresh: most leaves must be carried over");
        }
        run_checks<TupleInt>(t, probes, scans, 100, ref_only, "int/refresh", live.get());
        if (!ref_only) run_driver_shape(t, n_rows, 16, 10, 150);
    }
    // ---- YCSB string table (ycsb_loader.cpp:107-143): 16 random bytes per key, payload 1000 x 'a' ---------------
    {
        const uint32_t n = n_rows / 2;
        Table t = make_table(16, 1000, 16 * 1024, n / 16 + 4096);
        benchmark::FastRandom krng(1804289383);
        std::vector<std::string> keys;
        std::vector<char> payload(1000, 'a');
        for (uint32_t i = 0; i < n; ++i) {
            keys.push_back(krng.next_string(16));
            insert_row(t, keys.back().data(), payload.data());
        }
        benchmark::FastRandom rng(1);
        benchmark::ZipfDistribution zipf(n - 1, 0.9);
        std::vector<std::string> probes, scans;
        for (uint32_t i = 0; i < n_probe; ++i) probes.push_back(keys[zipf.GetNextNumber() % n]);
        for (uint32_t i = 0; i < 50; ++i) probes.push_back(rng.next_string(16));                // misses
        for (uint32_t i = 0; i < n_scan; ++i) scans.push_back(i % 4 ? keys[rng.next() % n] : rng.next_string(16));
        run_checks<TupleStr>(t, probes, scans, 1000, ref_only, "ycsb-string");
    }
    // ---- stock-shaped table: 16-byte memcmp keys, 400-byte rows, TPC-C ParameterSet (tpcc.cpp:78) ----------------
    {
        const uint32_t n = n_rows / 2;
        Table t = make_table(16, 400, 64 * 1024, n / 64 + 4096);
        std::vector<std::string> keys;
        std::vector<char> payload(400);
        for (uint32_t i = 0; i < n; ++i) {
            // {s_w_id, s_i_id} as two 8-byte integers, the shape of Stock::StockKey (tpcc_record.h:417-430)
            uint64_t kk[2] = {i % 30, i / 30};
            std::string k(reinterpret_cast<const char *>(kk), 16);
            for (uint32_t b = 0; b < 400; ++b) payload[b] = (char) ((i * 7 + b) & 0xFF);
            memcpy(payload.data(), k.data(), 16);
            keys.push_back(k);
            insert_row(t, k.data(), payload.data());
        }
        benchmark::FastRandom rng(7);
        std::vector<std::string> probes, scans;
        for (uint32_t i = 0; i < n_probe; ++i) probes.push_back(keys[rng.next() % n]);
        for (uint32_t i = 0; i < n_scan; ++i) scans.push_back(keys[rng.next() % n]);
        run_checks<TupleStock>(t, probes, scans, 10, ref_only, "stock");       // stock-level scans 10 rows
    }
    run_mvcc_checks(ref_only, n_rows);
    run_varlen(ref_only, 100000);
    printf("%s: %d failure(s)\n", ref_only ? "reference half only" : "drop-in parity", g_fail);
    return g_fail ? 1 : 0;
}
