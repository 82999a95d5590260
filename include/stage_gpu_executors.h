// stage_gpu_executors.h — header-only C++ adapters that put the reference's own class shapes back on
// top of the C ABI (include/stage_gpu.h), so that the YCSB / TPC-C drivers can switch the read operators
// by changing only the executor type.
//
// Shapes mirrored (paths relative to the reference repository):
//   mvstore::Record               include/vstore/b_tree.h:400-505   ([RecordMeta 48 B][cstamp u32][pad8(K) key][payload])
//   BTree::Read                   include/vstore/b_tree.h:812-813
//   BTree::RangeScanBySize        include/vstore/b_tree.h:830-835, Iterator::GetNext :899-941
//   IndexScanExecutor<Key,T>      include/execute/executor.h:342-570
//   TableScanExecutor<Key,T>      include/execute/executor.h:572-661
//
// Differences, all forced by the boundary: the BTree* argument becomes a GpuTreeView* (a GPU snapshot of
// that tree), the transaction is anything with GetReadId() (the reference's TransactionContext qualifies),
// results are owned copies (the reference's range results point into freed Records, executor.h:476-484),
// and every call is a batch of one — correct but slow; throughput comes from ReadBatch/ScanBatch.
#ifndef STAGE_GPU_EXECUTORS_H_
#define STAGE_GPU_EXECUTORS_H_

#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "stage_gpu.h"

namespace stage_gpu {

typedef uint32_t cid_t;

inline void check(int rc, const char *what) {
    if (rc != STAGE_OK)
        throw std::runtime_error(std::string(what) + ": " + stage_gpu_strerror(rc) + " [" + stage_gpu_last_error() + "]");
}

// Byte-compatible stand-in for mvstore::Record's accessors (b_tree.h:463-504): GetData() points at the
// padded key, the payload follows it, GetCstamp() is the reader's id as BTree::Read sets it (b_tree.cpp:2120).
class Record {
public:
    Record(uint32_t key_size, uint32_t payload_size)
        : key_size_(key_size), pad_(padded(key_size)), buf_(sizeof(cid_t) + padded(key_size) + payload_size, 0) {}
    static uint32_t padded(uint32_t k) { return (k + 7u) / 8u * 8u; }
    char *GetData() { return buf_.data() + sizeof(cid_t); }
    const char *GetKey() const { return buf_.data() + sizeof(cid_t); }
    char *GetPayload() { return buf_.data() + sizeof(cid_t) + pad_; }
    uint32_t GetCstamp() const { uint32_t v; std::memcpy(&v, buf_.data(), 4); return v; }
    void SetCstamp(cid_t c) { std::memcpy(buf_.data(), &c, 4); }
    uint16_t GetKeyLength() const { return (uint16_t) key_size_; }
    uint16_t GetKeyPaddedLength() const { return (uint16_t) pad_; }
    cid_t commit_id = 0;      // meta.meta_data.GetTxnCommitId() of the version returned
private:
    uint32_t key_size_, pad_;
    std::vector<char> buf_;
};

class GpuIterator;

// GPU-resident view of one BTree. Construct from BTree::GetRootNodeSafe() of a quiesced tree.
class GpuTreeView {
public:
    GpuTreeView(const void *root, uint32_t key_size, uint32_t payload_size, uint32_t leaf_node_size = 65536,
                uint32_t key_kind = STAGE_KEY_AUTO) {
        schema_ = stage_schema_t{key_size, payload_size, leaf_node_size, key_kind};
        stage_snapshot_t *h = nullptr;
        check(stage_gpu_snapshot_from_tree(root, &schema_, &h), "stage_gpu_snapshot_from_tree");
        snap_ = wrap(h, 0, 0);
    }
    // Tree with writers in flight: `resolve` looks an undo copy up in the tree's EphemeralPool (stage_undo_resolver_t,
    // include/stage_gpu.h), as BTree::Read does for a record whose update has not committed (b_tree.cpp:2087-2103).
    GpuTreeView(const void *root, uint32_t key_size, uint32_t payload_size, stage_undo_resolver_t resolve, void *resolve_ctx,
                uint32_t leaf_node_size = 65536, uint32_t key_kind = STAGE_KEY_AUTO) {
        schema_ = stage_schema_t{key_size, payload_size, leaf_node_size, key_kind};
        stage_snapshot_t *h = nullptr;
        check(stage_gpu_snapshot_from_tree_live(root, &schema_, resolve, resolve_ctx, &h), "stage_gpu_snapshot_from_tree_live");
        snap_ = wrap(h, 0, 0);
    }
    ~GpuTreeView() = default;          // the snapshot goes when its last user (this view or a reader in flight) lets go
    GpuTreeView(const GpuTreeView &) = delete;
    GpuTreeView &operator=(const GpuTreeView &) = delete;

    // BTree::Read shape. Returns the LATEST in-leaf version like the reference (visibility is the
    // executor's job); nullptr on a miss.
    std::unique_ptr<Record> Read(const char *key, uint16_t key_size, cid_t commit_id, bool /*is_for_update*/) {
        if (varlen()) {          // keys of different lengths travel as slots (stage_gpu_key_pack)
            char slot[32], slot_out[32];
            if (stage_gpu_key_pack(&schema_, key, key_size, slot) != STAGE_OK) return nullptr;   // no such key in this table
            auto rec = std::make_unique<Record>(key_size, schema_.payload_size);
            uint8_t found = 0;
            uint32_t ts = 0;
            ReadBatch(slot, 1, 0xFFFFFFFFu, &found, &ts, slot_out, rec->GetPayload());
            if (!found) return nullptr;
            std::memcpy(rec->GetData(), key, key_size);
            rec->commit_id = ts;
            rec->SetCstamp(commit_id);
            return rec;
        }
        if (key_size != schema_.key_size) return nullptr;
        auto rec = std::make_unique<Record>(schema_.key_size, schema_.payload_size);
        uint8_t found = 0;
        uint32_t ts = 0;
        // read_ts = MAX_CID: the in-leaf version is always "visible" here
        ReadBatch(key, 1, 0xFFFFFFFFu, &found, &ts, rec->GetData(), rec->GetPayload());
        if (!found) return nullptr;
        rec->commit_id = ts;
        rec->SetCstamp(commit_id);
        return rec;
    }

    std::unique_ptr<GpuIterator> RangeScanBySize(const char *key1, uint16_t size1, uint32_t scan_size);
    bool varlen() const { return (schema_.key_kind & STAGE_KEY_VARLEN) != 0; }

    // Batched forms: the ones that reach HBM speed. Outputs are caller-allocated host buffers.
    void ReadBatch(const void *keys, uint64_t n, cid_t read_ts, uint8_t *found, uint32_t *ts, void *key_out,
                   void *payload_out) {
        std::shared_ptr<Handle> h = pin();
        if (n == 1 && h->batcher)      // a single operation: coalesced with the other threads' operations in flight
            check(stage_gpu_batcher_probe(h->batcher, keys, read_ts, found, ts, key_out, payload_out), "stage_gpu_batcher_probe");
        else
            check(stage_gpu_probe_batch(h->snap, keys, n, read_ts, found, ts, key_out, payload_out, STAGE_IO_HOST),
                  "stage_gpu_probe_batch");
    }
    void ScanBatch(const void *start_keys, const uint32_t *scan_sizes, uint64_t n, uint32_t out_stride, cid_t read_ts,
                   bool apply_visibility, uint32_t *row_counts, void *key_out, uint32_t *ts_out, void *payload_out) {
        std::shared_ptr<Handle> h = pin();
        if (n == 1 && h->batcher && scan_sizes[0] <= out_stride)
            check(stage_gpu_batcher_scan(h->batcher, start_keys, scan_sizes[0], read_ts, apply_visibility ? 1 : 0, row_counts, key_out,
                                         ts_out, payload_out), "stage_gpu_batcher_scan");
        else
            check(stage_gpu_scan_batch(h->snap, start_keys, scan_sizes, n, out_stride, read_ts, apply_visibility ? 1 : 0,
                                       row_counts, key_out, ts_out, payload_out, STAGE_IO_HOST), "stage_gpu_scan_batch");
    }

    // After the host tree took committed inserts / updates / deletes (quiesced again): swap in the next snapshot.
    // Only the leaves that changed are staged again (stage_gpu_snapshot_refresh_from_tree).
    // With `resolve` (see the second constructor) update transactions may stay in flight during the call.
    // Readers are never blocked: calls already running finish on the snapshot they pinned (it is destroyed when the
    // last of them returns), calls that start after the swap see the new one. One Refresh at a time.
    stage_refresh_stats_t Refresh(const void *root, stage_undo_resolver_t resolve = nullptr, void *resolve_ctx = nullptr) {
        std::lock_guard<std::mutex> one(refresh_mu_);
        std::shared_ptr<Handle> cur = pin();
        stage_snapshot_t *next = nullptr;
        stage_refresh_stats_t st{};
        if (resolve)
            check(stage_gpu_snapshot_refresh_from_tree_live(cur->snap, root, resolve, resolve_ctx, &next, &st),
                  "stage_gpu_snapshot_refresh_from_tree_live");
        else
            check(stage_gpu_snapshot_refresh_from_tree(cur->snap, root, &next, &st), "stage_gpu_snapshot_refresh_from_tree");
        std::shared_ptr<Handle> fresh = wrap(next, batch_max_, batch_wait_us_);
        {
            std::lock_guard<std::mutex> g(mu_);
            snap_.swap(fresh);
        }
        return st;
    }

    // The reference's drivers run one executor per operation on N worker threads (ycsb_workload.cpp:522-529). With
    // batching enabled, single operations (Read, ReadBatch / ScanBatch of one, i.e. what the executor adapters issue) of
    // concurrent threads are coalesced into batched launches (stage_gpu_batcher_*). Call before the workers start.
    void EnableBatching(uint32_t max_batch = 256, uint32_t max_wait_us = 30) {
        std::lock_guard<std::mutex> one(refresh_mu_);
        batch_max_ = max_batch;
        batch_wait_us_ = max_wait_us;
        std::shared_ptr<Handle> cur = pin();
        if (!cur->batcher) check(stage_gpu_batcher_create(cur->snap, max_batch, max_wait_us, &cur->batcher), "stage_gpu_batcher_create");
    }
    void BatchingStats(uint64_t *batches, uint64_t *operations) {
        std::shared_ptr<Handle> cur = pin();
        *batches = *operations = 0;
        if (cur->batcher) stage_gpu_batcher_stats(cur->batcher, batches, operations);
    }

    const stage_schema_t &schema() const { return schema_; }
    // snapshot (+ its batcher), destroyed together when the last user lets go
    struct Handle {
        stage_snapshot_t *snap = nullptr;
        stage_batcher_t *batcher = nullptr;
        ~Handle() {
            if (batcher) stage_gpu_batcher_destroy(batcher);
            if (snap) stage_gpu_snapshot_destroy(snap);
        }
    };
    // The current snapshot, kept alive for as long as the returned pointer lives.
    std::shared_ptr<Handle> pin() const {
        std::lock_guard<std::mutex> g(mu_);
        return snap_;
    }
    // Raw handle of the current snapshot; only valid until the next Refresh (use pin() from concurrent readers).
    stage_snapshot_t *handle() { return pin()->snap; }

private:
    static std::shared_ptr<Handle> wrap(stage_snapshot_t *h, uint32_t batch_max, uint32_t wait_us) {
        auto p = std::make_shared<Handle>();
        p->snap = h;
        if (batch_max) check(stage_gpu_batcher_create(h, batch_max, wait_us, &p->batcher), "stage_gpu_batcher_create");
        return p;
    }
    stage_schema_t schema_{};
    mutable std::mutex mu_;
    std::mutex refresh_mu_;
    std::shared_ptr<Handle> snap_;
    uint32_t batch_max_ = 0, batch_wait_us_ = 0;
};

// Iterator::GetNext shape over one scan's result slab (the scan runs when the iterator is created).
class GpuIterator {
public:
    GpuIterator(GpuTreeView *tree, const char *key1, uint32_t scan_size, cid_t read_ts, bool apply_visibility)
        : schema_(tree->schema()), K_(tree->schema().key_size), P_(tree->schema().payload_size), scan_size_(scan_size) {
        keys_.resize((size_t) scan_size * K_);
        ts_.resize(scan_size);
        pay_.resize((size_t) scan_size * P_);
        uint32_t sz = scan_size;
        tree->ScanBatch(key1, &sz, 1, scan_size, read_ts, apply_visibility, &count_, keys_.data(), ts_.data(),
                        pay_.data());
    }
    // nullptr once the scan has no more rows (the reference returns nullptr for the rest of the
    // executor loop too, b_tree.h:900-902).
    std::unique_ptr<Record> GetNext() {
        if (next_ >= count_) return nullptr;
        uint32_t len = K_;
        if (schema_.key_kind & STAGE_KEY_VARLEN) stage_gpu_key_unpack(&schema_, keys_.data() + (size_t) next_ * K_, nullptr, &len);
        auto rec = std::make_unique<Record>(len, P_);
        std::memcpy(rec->GetData(), keys_.data() + (size_t) next_ * K_, len);
        std::memcpy(rec->GetPayload(), pay_.data() + (size_t) next_ * P_, P_);
        rec->commit_id = ts_[next_];
        ++next_;
        return rec;
    }
    uint32_t count() const { return count_; }
private:
    stage_schema_t schema_;
    uint32_t K_, P_, scan_size_, count_ = 0, next_ = 0;
    std::vector<char> keys_, pay_;
    std::vector<uint32_t> ts_;
};

inline std::unique_ptr<GpuIterator> GpuTreeView::RangeScanBySize(const char *key1, uint16_t size1, uint32_t scan_size) {
    if (varlen()) {
        char slot[32];
        if (stage_gpu_key_pack(&schema_, key1, size1, slot) != STAGE_OK) return nullptr;
        return std::make_unique<GpuIterator>(this, slot, scan_size, 0xFFFFFFFFu, false);
    }
    if (size1 != schema_.key_size) return nullptr;
    return std::make_unique<GpuIterator>(this, key1, scan_size, 0xFFFFFFFFu, false);
}

// IndexScanExecutor shape (executor.h:353-364, 549-555). T is copied from [padded key][payload] exactly
// as the reference copies sizeof(T) bytes from Record::GetData() (executor.h:397, 476).
template <class Key, class T, class Txn>
class GpuIndexScanExecutor {
public:
    GpuIndexScanExecutor(GpuTreeView *data_table, uint16_t key_sz, const Key &start_key, uint32_t scan_sz,
                         bool point_lookup, std::function<bool(const T *, bool &end_scan)> predicate,
                         bool is_for_update, Txn *txn, void * /*VersionStore* buf_mgr*/)
        : table_(data_table), key_size_(key_sz), start_key_(start_key), scan_size_(scan_sz),
          point_lookup_(point_lookup), predicate_(predicate), is_for_update_(is_for_update), txn_(txn) {}

    bool Execute() {
        const cid_t read_id = txn_->GetReadId();
        const uint32_t K = table_->schema().key_size, P = table_->schema().payload_size;
        const uint32_t pad = Record::padded(K);
        if (point_lookup_) {
            std::vector<char> key(K), pay(P);
            uint8_t found = 0;
            uint32_t ts = 0;
            table_->ReadBatch(start_key_, 1, read_id, &found, &ts, key.data(), pay.data());
            record_.reset();
            if (found) {
                record_.reset(new T);
                fill(record_.get(), key.data(), pay.data(), K, pad, P);
            }
            return true;
        }
        std::vector<char> keys((size_t) scan_size_ * K), pays((size_t) scan_size_ * P);
        std::vector<uint32_t> ts(scan_size_);
        uint32_t count = 0, sz = scan_size_;
        table_->ScanBatch(start_key_, &sz, 1, scan_size_, read_id, true, &count, keys.data(), ts.data(), pays.data());
        owned_.resize(count);
        results_.clear();
        bool end_scan = false;
        for (uint32_t i = 0; i < count; ++i) {
            fill(&owned_[i], keys.data() + (size_t) i * K, pays.data() + (size_t) i * P, K, pad, P);
            if (predicate_) predicate_(&owned_[i], end_scan);     // result ignored, as in executor.h:477-484
            results_.push_back(&owned_[i]);
        }
        return true;
    }
    const std::vector<T *> &GetResults() { return results_; }
    const T *GetRecord() { return record_.get(); }

private:
    static void fill(T *dst, const char *key, const char *pay, uint32_t K, uint32_t pad, uint32_t P) {
        std::vector<char> img(pad + P, 0);
        std::memcpy(img.data(), key, K);
        std::memcpy(img.data() + pad, pay, P);
        std::memset(dst, 0, sizeof(T));
        std::memcpy(dst, img.data(), sizeof(T) < img.size() ? sizeof(T) : img.size());
    }
    GpuTreeView *table_;
    uint16_t key_size_;
    Key start_key_;
    uint32_t scan_size_;
    bool point_lookup_;
    std::function<bool(const T *, bool &)> predicate_;
    bool is_for_update_;
    Txn *txn_;
    std::vector<T> owned_;
    std::vector<T *> results_;
    std::unique_ptr<T> record_;
};

// TableScanExecutor shape (executor.h:576-579, 644-649): no visibility test, T copied by value.
template <class Key, class T, class Txn>
class GpuTableScanExecutor {
public:
    GpuTableScanExecutor(GpuTreeView *data_table, const Key &start_key, uint32_t key_size, uint32_t scan_size,
                         Txn *txn, void * /*buf_mgr*/)
        : table_(data_table), start_key_(start_key), key_sz_(key_size), scan_sz_(scan_size), txn_(txn) {}

    bool Execute() {
        const uint32_t K = table_->schema().key_size, P = table_->schema().payload_size, pad = Record::padded(K);
        if (scan_sz_ == (uint32_t) -1) {          // physical whole-tree walk (executor.h:581-619)
            stage_snapshot_info_t info{};
            check(stage_gpu_snapshot_info(table_->handle(), &info), "stage_gpu_snapshot_info");
            const uint64_t n = info.n_rows;
            std::vector<char> keys((size_t) n * K), pays((size_t) n * P);
            check(stage_gpu_scan_physical(table_->handle(), 0, n, keys.data(), nullptr, pays.data(), STAGE_IO_HOST),
                  "stage_gpu_scan_physical");
            results_.resize(n);
            std::vector<char> img(pad + P);
            for (uint64_t i = 0; i < n; ++i) {
                std::fill(img.begin(), img.end(), 0);
                std::memcpy(img.data(), keys.data() + (size_t) i * K, K);
                std::memcpy(img.data() + pad, pays.data() + (size_t) i * P, P);
                std::memset(&results_[i], 0, sizeof(T));
                std::memcpy(&results_[i], img.data(), sizeof(T) < img.size() ? sizeof(T) : img.size());
            }
            return true;
        }
        std::vector<char> keys((size_t) scan_sz_ * K), pays((size_t) scan_sz_ * P);
        std::vector<uint32_t> ts(scan_sz_);
        uint32_t count = 0, sz = scan_sz_;
        table_->ScanBatch(start_key_, &sz, 1, scan_sz_, 0, false, &count, keys.data(), ts.data(), pays.data());
        results_.resize(count);
        std::vector<char> img(pad + P);
        for (uint32_t i = 0; i < count; ++i) {
            std::fill(img.begin(), img.end(), 0);
            std::memcpy(img.data(), keys.data() + (size_t) i * K, K);
            std::memcpy(img.data() + pad, pays.data() + (size_t) i * P, P);
            std::memset(&results_[i], 0, sizeof(T));
            std::memcpy(&results_[i], img.data(), sizeof(T) < img.size() ? sizeof(T) : img.size());
        }
        fail_num_ = scan_sz_ - count;        // GetNext() == nullptr iterations (executor.h:629-632)
        return true;
    }
    std::vector<T> &GetResults() { return results_; }
    uint32_t GetFailNum() const { return fail_num_; }

private:
    GpuTreeView *table_;
    Key start_key_;
    uint32_t key_sz_, scan_sz_;
    Txn *txn_;
    std::vector<T> results_;
    uint32_t fail_num_ = 0;
};

}  // namespace stage_gpu
#endif  // STAGE_GPU_EXECUTORS_H_
