// Offline stand-in for <tbb/spin_mutex.h>, used ONLY to compile the unmodified reference
// sources into oracle/_ref/ (test infrastructure; never linked into the product).
// Members offered = what the reference touches (include/common/spin_latch.h:33-45).
#pragma once
#include <atomic>
namespace tbb {
class spin_mutex {
    std::atomic_flag f_ = ATOMIC_FLAG_INIT;
public:
    spin_mutex() = default;
    spin_mutex(const spin_mutex&) {}
    spin_mutex& operator=(const spin_mutex&) { return *this; }
    void lock() { while (f_.test_and_set(std::memory_order_acquire)) { __builtin_ia32_pause(); } }
    bool try_lock() { return !f_.test_and_set(std::memory_order_acquire); }
    void unlock() { f_.clear(std::memory_order_release); }
};
}
