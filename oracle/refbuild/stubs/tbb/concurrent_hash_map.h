// Offline stand-in for <tbb/concurrent_hash_map.h> (included by the reference, unused).
#pragma once
#include <unordered_map>
namespace tbb { template <typename K, typename V> using concurrent_hash_map = std::unordered_map<K, V>; }
