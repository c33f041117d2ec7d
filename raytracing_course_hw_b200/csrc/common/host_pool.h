// Persistent host worker pool shared by the two libraries (each .so carries its own instance): the device library
// uses it for record building and the BVH work that stays on the host, the host library for triangle assembly and
// image decoding.
#pragma once
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdlib>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace b200rt {

// Host threads one context may use for record building and the BVH rebuild: min(16, cores), or
// B200RT_HOST_THREADS (one process per GPU on a shared host: cores / GPUs avoids oversubscription).
inline unsigned host_threads() {
    unsigned n = std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
    if (const char *e = getenv("B200RT_HOST_THREADS")) n = (unsigned)std::max(1, std::min(64, atoi(e)));
    return n;
}

// Persistent host workers for the upload path (BVH rebuild, record building).  A scene upload runs ~20 short
// parallel sections; spawning 16 std::threads for each cost more than the work inside (measured on the GPU box's
// virtualised cores: ~1 ms per round).  run(n, fn) executes fn(0) .. fn(n-1) on the pool plus the calling thread
// and returns when all are done; one job at a time (contexts driven from different host threads take turns).
class HostPool {
public:
    static HostPool &get() { static HostPool p; return p; }
    void run(unsigned n, const std::function<void(unsigned)> &fn) {
        if (n <= 1) { if (n) fn(0); return; }
        std::lock_guard<std::mutex> serial(run_mu_);
        {
            std::lock_guard<std::mutex> lk(mu_);
            while (workers_.size() + 1 < n) workers_.emplace_back([this] { loop(); });
            job_ = &fn; n_tasks_ = n; next_ = 0; done_ = 0; gen_++;
        }
        cv_work_.notify_all();
        work();
        std::unique_lock<std::mutex> lk(mu_);
        cv_done_.wait(lk, [this] { return done_ == n_tasks_; });
        job_ = nullptr;
    }
    ~HostPool() {
        { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
        cv_work_.notify_all();
        for (auto &t : workers_) t.join();
    }
private:
    void work() {
        for (;;) {
            unsigned t;
            const std::function<void(unsigned)> *job;
            {
                std::lock_guard<std::mutex> lk(mu_);
                if (!job_ || next_ >= n_tasks_) return;
                t = next_++; job = job_;
            }
            (*job)(t);
            std::lock_guard<std::mutex> lk(mu_);
            if (++done_ == n_tasks_) cv_done_.notify_all();
        }
    }
    void loop() {
        uint64_t seen = 0;
        for (;;) {
            // an upload issues its parallel sections back to back: spin briefly before going to sleep, a futex wake
            // of a parked (virtual) core costs more than most sections
            for (int spin = 0; spin < 20000 && gen_.load(std::memory_order_acquire) == seen && !stop_; spin++) {
#if defined(__x86_64__) || defined(__i386__)
                __builtin_ia32_pause();
#endif
            }
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_work_.wait(lk, [&] { return stop_ || gen_.load() != seen; });
                if (stop_) return;
                seen = gen_.load();
            }
            work();
        }
    }
    std::vector<std::thread> workers_;
    std::mutex mu_, run_mu_;
    std::condition_variable cv_work_, cv_done_;
    const std::function<void(unsigned)> *job_ = nullptr;
    unsigned n_tasks_ = 0, next_ = 0, done_ = 0;
    std::atomic<uint64_t> gen_{0};
    std::atomic<bool> stop_{false};
};

// Runs f(begin, end) over [0, n) in chunks of at least `grain` items on the pool.
template <class F>
inline void parallel_for(size_t n, const F &f, size_t grain = 16384) {
    const unsigned nt = (unsigned)std::min<size_t>((size_t)host_threads(), (n + grain - 1) / grain);
    if (nt <= 1) { if (n) f((size_t)0, n); return; }
    const size_t chunk = (n + nt - 1) / nt;
    HostPool::get().run(nt, [&](unsigned t) {
        const size_t a = t * chunk, b = std::min(n, a + chunk);
        if (a < b) f(a, b);
    });
}

}  // namespace b200rt
