// Spinning helper team of the decoder's parallel Huffman reader (decode_host.cpp). Header-only so that the native stress test
// (tests/native/spin_team_test.cpp) builds it without the rest of the library.
#pragma once

#include <sched.h>

#include <atomic>
#include <cstdint>
#include <exception>
#include <functional>
#include <mutex>

namespace mnenc {

// The threads of a HostPool held inside ONE parallelFor for the duration of an image. A sequence's chunks are tiny (tens of
// microseconds each), far below the wake-up latency of a condition variable, so the helpers spin on an atomic ticket
// instead: ticket = generation << 32 | next index; a job is published by writing its slot (generation parity) and then
// storing the new ticket. A helper that draws a stale ticket sees the slot's generation differ and backs off. The slot's
// generation and size are ONE atomic word (genN): a helper preempted between drawing an over-the-end ticket of job g - 2 and
// checking it must never pair the old generation with the size of job g, which reuses the slot (that ran an index twice).
class SpinTeam {
public:
    explicit SpinTeam(int threads) : threads_(threads) {}
    int size() const { return threads_; }
    void forEach(int n, const std::function<void(int)>& fn) {
        if (n <= 0) return;
        if (threads_ <= 1) {
            for (int i = 0; i < n; ++i) fn(i);
            return;
        }
        const uint64_t g = ++gen_;
        Job& j = jobs_[g & 1];
        j.fn.store(&fn, std::memory_order_relaxed);
        j.pending.store(n, std::memory_order_relaxed);
        j.genN.store(g << 32 | static_cast<uint32_t>(n), std::memory_order_release);
        ticket_.store(g << 32, std::memory_order_release);
        drain();
        while (j.pending.load(std::memory_order_acquire) > 0) pause();
        std::exception_ptr e;
        {
            std::lock_guard<std::mutex> lock(errMutex_);
            std::swap(e, err_);
        }
        if (e) std::rethrow_exception(e);
    }
    void helper() {  // body of every pool thread but the caller's
        for (unsigned idle = 0; !quit_.load(std::memory_order_acquire); ++idle) {
            drain();
            if ((idle & 63) == 63)
                sched_yield();  // stay polite when there are more threads than cores
            else
                pause();
        }
    }
    void finish() { quit_.store(true, std::memory_order_release); }

private:
    struct Job {
        std::atomic<const std::function<void(int)>*> fn{nullptr};
        std::atomic<int> pending{0};
        std::atomic<uint64_t> genN{0};  // generation << 32 | number of indices, read and written as one word
    };
    // true when ticket generation g and index i belong to the job the slot holds right now
    static bool live(const Job& j, uint64_t g, uint64_t i) {
        const uint64_t m = j.genN.load(std::memory_order_acquire);
        return (m >> 32) == g && i < (m & 0xFFFFFFFFu);
    }
    static void pause() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#elif defined(__aarch64__) || defined(__arm__)
        asm volatile("yield" ::: "memory");
#else
        std::atomic_signal_fence(std::memory_order_seq_cst);
#endif
    }
    void drain() {
        for (;;) {
            const uint64_t peek = ticket_.load(std::memory_order_acquire);
            const uint64_t g = peek >> 32;
            if (g == 0) return;
            if (!live(jobs_[g & 1], g, peek & 0xFFFFFFFFu)) return;
            // The drawn ticket, not the peeked one, says which job the index belongs to: between the peek and the fetch_add the
            // job may have completed and the next one been published. An index below the job's size is never thrown away - its
            // job cannot complete (pending > 0) before it has run, so the slot is stable.
            const uint64_t t = ticket_.fetch_add(1, std::memory_order_acq_rel);
            const uint64_t gt = t >> 32;
            Job& j = jobs_[gt & 1];
            const int i = static_cast<int>(t & 0xFFFFFFFFu);
            if (!live(j, gt, t & 0xFFFFFFFFu)) continue;
            try {
                (*j.fn.load(std::memory_order_relaxed))(i);
            } catch (...) {  // first exception of the job reaches the caller of forEach after the job has drained
                std::lock_guard<std::mutex> lock(errMutex_);
                if (!err_) err_ = std::current_exception();
            }
            j.pending.fetch_sub(1, std::memory_order_release);
        }
    }
    int threads_;
    std::atomic<uint64_t> ticket_{0};
    std::atomic<bool> quit_{false};
    uint64_t gen_ = 0;
    Job jobs_[2];
    std::mutex errMutex_;
    std::exception_ptr err_;
};

}  // namespace mnenc
