// Offline stand-in for <tbb/concurrent_unordered_map.h> (oracle/_ref build only).
// Used by include/common/manager.h:68-72, include/execute/txn_context.h:24 (per-transaction
// read/write set: single-threaded) and include/common/ephemeral_pool.h. NOT thread-safe for
// concurrent inserts; the harness only mutates these from one thread at a time.
#pragma once
#include <unordered_map>
#include <functional>
namespace tbb {
template <typename K, typename V, typename H = std::hash<K>, typename E = std::equal_to<K>>
class concurrent_unordered_map : public std::unordered_map<K, V, H, E> {
    using base = std::unordered_map<K, V, H, E>;
public:
    using base::base;
    template <typename It> auto unsafe_erase(It it) { return base::erase(it); }
    size_t unsafe_erase(const K& k) { return base::erase(k); }
};
}
