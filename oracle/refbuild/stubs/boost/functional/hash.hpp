// Offline stand-in for <boost/functional/hash.hpp> (oracle/_ref build only):
// hash_combine as used by include/common/constants.h:212-216, ephemeral_pool.h:158-160.
#pragma once
#include <functional>
#include <type_traits>
#include <cstddef>
#include <cstdint>
namespace boost {
template <typename T>
inline void hash_combine(std::size_t& seed, const T& v) {
    std::size_t h;
    if constexpr (std::is_enum<T>::value) h = std::hash<long long>()((long long)v);
    else h = std::hash<T>()(v);
    seed ^= h + 0x9e3779b9 + (seed << 6) + (seed >> 2);
}
}
