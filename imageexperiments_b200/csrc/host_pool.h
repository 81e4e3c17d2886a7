// Small persistent host thread pool for the per-sequence code-table builds (193 independent jobs per image,
// entropy_host.cpp). One pool per encoder context; parallelFor blocks until every index has run. The calling
// thread takes part in the work, so a pool of size 1 degenerates to a plain loop.
#pragma once

#include <atomic>
#include <condition_variable>
#include <exception>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace mnenc {

class HostPool {
public:
    explicit HostPool(int threads) {
        for (int i = 1; i < threads; ++i) workers_.emplace_back([this] { workerLoop(); });
    }
    ~HostPool() {
        {
            std::lock_guard<std::mutex> g(m_);
            stop_ = true;
        }
        wake_.notify_all();
        for (auto& t : workers_) t.join();
    }
    HostPool(const HostPool&) = delete;
    HostPool& operator=(const HostPool&) = delete;

    int size() const { return static_cast<int>(workers_.size()) + 1; }

    // fn(i) for every i in [0, n), dynamically distributed in index order (put the heavy jobs first).
    void parallelFor(int n, const std::function<void(int)>& fn) {
        if (n <= 0) return;
        if (workers_.empty() || n == 1) {
            for (int i = 0; i < n; ++i) fn(i);
            return;
        }
        auto job = std::make_shared<Job>();
        job->fn = &fn;
        job->n = n;
        job->pending = n;
        {
            std::lock_guard<std::mutex> g(m_);
            job_ = job;
            ++epoch_;
        }
        wake_.notify_all();
        drain(*job);
        std::unique_lock<std::mutex> g(m_);
        done_.wait(g, [&] { return job->pending == 0; });
        job_.reset();
        if (job->error) std::rethrow_exception(job->error);  // first exception of any index, on the calling thread
    }

    // body() on the calling thread while every other pool thread runs helper() (which must return once body is done —
    // the caller arranges that, e.g. through a flag body sets last). Used by the decoder's spinning reader team.
    void withHelpers(const std::function<void()>& helper, const std::function<void()>& body) {
        if (workers_.empty()) {
            body();
            return;
        }
        const std::function<void(int)> fn = [&](int) { helper(); };
        auto job = std::make_shared<Job>();
        job->fn = &fn;
        job->n = static_cast<int>(workers_.size());
        job->pending = job->n;
        {
            std::lock_guard<std::mutex> g(m_);
            job_ = job;
            ++epoch_;
        }
        wake_.notify_all();
        std::exception_ptr bodyError;
        try {
            body();
        } catch (...) {  // the helpers still hold `fn`: wait for them first (the caller's body must release them on every path)
            bodyError = std::current_exception();
        }
        std::unique_lock<std::mutex> g(m_);
        done_.wait(g, [&] { return job->pending == 0; });
        job_.reset();
        if (bodyError) std::rethrow_exception(bodyError);
        if (job->error) std::rethrow_exception(job->error);
    }

private:
    struct Job {  // late wakers hold their own reference, so a finished job can never be confused with the next one
        const std::function<void(int)>* fn = nullptr;
        int n = 0;
        std::atomic<int> next{0};
        int pending = 0;  // guarded by m_
        std::exception_ptr error;  // guarded by m_: first exception thrown by fn on any thread
    };
    void drain(Job& job) {
        int ran = 0;
        while (true) {
            const int i = job.next.fetch_add(1, std::memory_order_relaxed);
            if (i >= job.n) break;
            try {
                (*job.fn)(i);
            } catch (...) {  // never out of a worker thread (std::terminate) and never past a job the others still run
                std::lock_guard<std::mutex> g(m_);
                if (!job.error) job.error = std::current_exception();
            }
            ++ran;
        }
        if (ran) {
            std::lock_guard<std::mutex> g(m_);
            job.pending -= ran;
            if (job.pending == 0) done_.notify_all();
        }
    }
    void workerLoop() {
        unsigned long long seen = 0;
        while (true) {
            std::shared_ptr<Job> job;
            {
                std::unique_lock<std::mutex> g(m_);
                wake_.wait(g, [&] { return stop_ || epoch_ != seen; });
                if (stop_) return;
                seen = epoch_;
                job = job_;
            }
            if (job) drain(*job);
        }
    }

    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable wake_, done_;
    std::shared_ptr<Job> job_;
    unsigned long long epoch_ = 0;
    bool stop_ = false;
};

}  // namespace mnenc
