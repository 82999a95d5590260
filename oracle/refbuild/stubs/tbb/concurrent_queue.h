// Offline stand-in for <tbb/concurrent_queue.h> (oracle/_ref build only).
// Used by include/common/concurrent_queue.h:54, concurrent_blocking_queue.h:55,
// include/vstore/version_store.h:692, src/vstore/version_store.cpp:23,472 (copied by value).
#pragma once
#include <deque>
#include <mutex>
#include <condition_variable>
#include <memory>
#include <cstddef>
namespace tbb {
template <typename T> using cache_aligned_allocator = std::allocator<T>;

template <typename T, typename Alloc = std::allocator<T>>
class concurrent_queue {
protected:
    mutable std::mutex m_;
    std::deque<T> q_;
public:
    concurrent_queue() = default;
    concurrent_queue(const concurrent_queue& o) { std::lock_guard<std::mutex> g(o.m_); q_ = o.q_; }
    concurrent_queue& operator=(const concurrent_queue& o) {
        if (this != &o) { std::lock_guard<std::mutex> g(o.m_); std::lock_guard<std::mutex> h(m_); q_ = o.q_; }
        return *this;
    }
    bool empty() const { std::lock_guard<std::mutex> g(m_); return q_.empty(); }
    void clear() { std::lock_guard<std::mutex> g(m_); q_.clear(); }
    void push(const T& v) { std::lock_guard<std::mutex> g(m_); q_.push_back(v); }
    void push(T&& v) { std::lock_guard<std::mutex> g(m_); q_.push_back(std::move(v)); }
    bool try_pop(T& out) {
        std::lock_guard<std::mutex> g(m_);
        if (q_.empty()) return false;
        out = std::move(q_.front()); q_.pop_front(); return true;
    }
    size_t unsafe_size() const { std::lock_guard<std::mutex> g(m_); return q_.size(); }
};

template <typename T, typename Alloc = std::allocator<T>>
class concurrent_bounded_queue {
    mutable std::mutex m_;
    std::condition_variable cv_;
    std::deque<T> q_;
    bool aborted_ = false;
public:
    bool empty() const { std::lock_guard<std::mutex> g(m_); return q_.empty(); }
    void clear() { std::lock_guard<std::mutex> g(m_); q_.clear(); }
    void push(const T& v) { { std::lock_guard<std::mutex> g(m_); q_.push_back(v); } cv_.notify_one(); }
    bool try_pop(T& out) {
        std::lock_guard<std::mutex> g(m_);
        if (q_.empty()) return false;
        out = std::move(q_.front()); q_.pop_front(); return true;
    }
    void pop(T& out) {
        std::unique_lock<std::mutex> g(m_);
        cv_.wait(g, [&] { return !q_.empty() || aborted_; });
        if (q_.empty()) return;
        out = std::move(q_.front()); q_.pop_front();
    }
    std::ptrdiff_t size() const { std::lock_guard<std::mutex> g(m_); return (std::ptrdiff_t)q_.size(); }
    void set_capacity(std::ptrdiff_t) {}
    void abort() { { std::lock_guard<std::mutex> g(m_); aborted_ = true; } cv_.notify_all(); }
};
}
