// Host thread pool with a chunk-claiming parallel_for.
//
// Role parity: the reference's process-global pool and `parallel_for`
// (/root/reference/pytorch_impl/libs/native/so_threadpool/threadpool.cpp:80-149,
//  include/threadpool.hpp:202-222). Design here: persistent workers sleep on one condition
// variable; a parallel region publishes a [begin, end) range plus a grain, and every
// participant (the caller included) claims grains with one atomic fetch_add until the range is
// exhausted — dynamic load balance instead of the reference's static equal split.
#pragma once

#include <atomic>
#include <condition_variable>
#include <cstddef>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace gb {

class ThreadPool {
public:
    explicit ThreadPool(unsigned workers = 0);
    ~ThreadPool();
    ThreadPool(ThreadPool const&) = delete;
    ThreadPool& operator=(ThreadPool const&) = delete;

    unsigned size() const noexcept { return static_cast<unsigned>(threads_.size()) + 1u; }

    // Calls body(lo, hi) over disjoint sub-ranges covering [begin, end). Blocks until done.
    // Re-entrant calls from inside a body run serially on the calling thread.
    void parallel_for(std::size_t begin, std::size_t end, std::function<void(std::size_t, std::size_t)> const& body,
                      std::size_t grain = 0);

    static ThreadPool& global();

private:
    struct Region {
        std::function<void(std::size_t, std::size_t)> const* body = nullptr;
        std::size_t end = 0, grain = 1;
        std::atomic<std::size_t> next{0};
        std::atomic<unsigned> inside{0};
    };

    void worker_loop();
    void drain(Region& region);

    std::vector<std::thread> threads_;
    std::mutex mutex_;
    std::condition_variable wake_, done_;
    Region* current_ = nullptr;
    unsigned long generation_ = 0;
    bool stopping_ = false;
    std::mutex region_mutex_;  // one parallel region at a time
};

template <class F>
inline void parallel_for(std::size_t begin, std::size_t end, F&& body, std::size_t grain = 0) {
    std::function<void(std::size_t, std::size_t)> fn = std::forward<F>(body);
    ThreadPool::global().parallel_for(begin, end, fn, grain);
}

}  // namespace gb
