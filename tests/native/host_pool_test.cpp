// csrc/host_pool.h: an exception thrown by a job — on a worker or on the calling thread — reaches the caller of
// parallelFor / withHelpers after every index has run, never std::terminate, and the pool stays usable.
#include <atomic>
#include <cstdio>
#include <stdexcept>

#include "../../imageexperiments_b200/csrc/host_pool.h"

int main() {
    mnenc::HostPool pool(4);
    for (int round = 0; round < 200; ++round) {
        std::atomic<int> ran{0};
        bool caught = false;
        try {
            pool.parallelFor(64, [&](int i) {
                ++ran;
                if (i % 16 == round % 16) throw std::runtime_error("job failed");
            });
        } catch (const std::runtime_error&) {
            caught = true;
        }
        if (!caught || ran != 64) {
            std::printf("round %d: caught %d, ran %d of 64\n", round, int(caught), int(ran));
            return 1;
        }
        std::atomic<int> sum{0};
        pool.parallelFor(100, [&](int i) { sum += i; });
        if (sum != 4950) {
            std::printf("round %d: pool unusable after an exception (sum %d)\n", round, int(sum));
            return 1;
        }
    }
    for (int round = 0; round < 50; ++round) {
        std::atomic<bool> release{false};
        std::atomic<int> helpers{0};
        bool caught = false;
        try {
            pool.withHelpers(
                [&] {
                    ++helpers;
                    while (!release.load()) std::this_thread::yield();
                },
                [&] {
                    struct Release {
                        std::atomic<bool>& r;
                        ~Release() { r = true; }
                    } rel{release};
                    throw std::runtime_error("body failed");
                });
        } catch (const std::runtime_error&) {
            caught = true;
        }
        if (!caught || helpers != 3) {
            std::printf("withHelpers round %d: caught %d, helpers %d of 3\n", round, int(caught), int(helpers));
            return 1;
        }
    }
    std::puts("host_pool OK");
    return 0;
}
