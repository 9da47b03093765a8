#include "threadpool.hpp"

#include <algorithm>
#include <cstdlib>

#include <unistd.h>

namespace gb {

namespace {
thread_local bool tl_inside_region = false;

unsigned default_workers() {
    if (char const* env = std::getenv("GARFIELD_NUM_THREADS")) {
        long v = std::strtol(env, nullptr, 10);
        if (v > 0) return static_cast<unsigned>(v);
    }
    long online = ::sysconf(_SC_NPROCESSORS_ONLN);
    return online > 0 ? static_cast<unsigned>(online) : 1u;
}
}  // namespace

ThreadPool::ThreadPool(unsigned workers) {
    if (workers == 0) workers = default_workers();
    threads_.reserve(workers > 0 ? workers - 1 : 0);
    for (unsigned i = 1; i < workers; ++i) threads_.emplace_back([this] { worker_loop(); });
}

ThreadPool::~ThreadPool() {
    {
        std::lock_guard<std::mutex> lock(mutex_);
        stopping_ = true;
    }
    wake_.notify_all();
    for (auto& t : threads_) t.join();
}

ThreadPool& ThreadPool::global() {
    static ThreadPool pool;
    return pool;
}

void ThreadPool::drain(Region& region) {
    bool const was_inside = tl_inside_region;
    tl_inside_region = true;
    for (;;) {
        std::size_t lo = region.next.fetch_add(region.grain, std::memory_order_relaxed);
        if (lo >= region.end) break;
        std::size_t hi = std::min(region.end, lo + region.grain);
        (*region.body)(lo, hi);
    }
    tl_inside_region = was_inside;
}

void ThreadPool::worker_loop() {
    unsigned long seen = 0;
    for (;;) {
        Region* region;
        {
            std::unique_lock<std::mutex> lock(mutex_);
            wake_.wait(lock, [&] { return stopping_ || (current_ != nullptr && generation_ != seen); });
            if (stopping_) return;
            seen = generation_;
            region = current_;
            region->inside.fetch_add(1, std::memory_order_acq_rel);
        }
        drain(*region);
        {
            std::lock_guard<std::mutex> lock(mutex_);
            region->inside.fetch_sub(1, std::memory_order_acq_rel);
        }
        done_.notify_all();
    }
}

void ThreadPool::parallel_for(std::size_t begin, std::size_t end,
                              std::function<void(std::size_t, std::size_t)> const& body, std::size_t grain) {
    if (end <= begin) return;
    std::size_t const total = end - begin;
    if (tl_inside_region || threads_.empty() || total == 1) {
        body(begin, end);
        return;
    }
    if (grain == 0) grain = std::max<std::size_t>(1, total / (4 * size()));
    std::lock_guard<std::mutex> only_one(region_mutex_);
    Region region;
    region.body = &body;
    region.end = end;
    region.grain = grain;
    region.next.store(begin, std::memory_order_relaxed);
    {
        std::lock_guard<std::mutex> lock(mutex_);
        current_ = &region;
        ++generation_;
    }
    wake_.notify_all();
    drain(region);
    {
        std::unique_lock<std::mutex> lock(mutex_);
        current_ = nullptr;  // late wakers see no region and go back to sleep
        done_.wait(lock, [&] { return region.inside.load(std::memory_order_acquire) == 0; });
    }
}

}  // namespace gb
