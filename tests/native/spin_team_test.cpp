// csrc/spin_team.h under oversubscription: more spinning helpers than cores, thousands of tiny jobs published back to back.
// Every index of every job runs exactly once (a helper that is preempted between looking at the ticket and drawing one must
// not throw the drawn index away - that hung forEach), and an exception thrown by a job on any thread reaches the caller of
// forEach after the job has drained, leaving the team usable.
#include <atomic>
#include <cstdio>
#include <stdexcept>
#include <thread>
#include <vector>

#include "../../imageexperiments_b200/csrc/spin_team.h"

int main() {
    const int hw = static_cast<int>(std::thread::hardware_concurrency());
    const int threads = 3 * (hw > 0 ? hw : 4) + 1;  // deliberately more than the cores
    mnenc::SpinTeam team(threads);
    std::vector<std::thread> helpers;
    for (int i = 1; i < threads; ++i) helpers.emplace_back([&] { team.helper(); });
    bool ok = true;
    for (int round = 0; round < 20000 && ok; ++round) {
        const int n = 1 + round % 7;
        std::atomic<int> hits[8] = {};
        team.forEach(n, [&](int i) { ++hits[i]; });
        for (int i = 0; i < n; ++i)
            if (hits[i] != 1) {
                std::printf("round %d: index %d ran %d times\n", round, i, int(hits[i]));
                ok = false;
            }
    }
    for (int round = 0; round < 500 && ok; ++round) {
        std::atomic<int> ran{0};
        bool caught = false;
        try {
            team.forEach(16, [&](int i) {
                ++ran;
                if (i == round % 16) throw std::runtime_error("job failed");
            });
        } catch (const std::runtime_error&) {
            caught = true;
        }
        if (!caught || ran != 16) {
            std::printf("throwing round %d: caught %d, ran %d of 16\n", round, int(caught), int(ran));
            ok = false;
        }
    }
    team.finish();
    for (auto& t : helpers) t.join();
    if (ok) std::printf("spin_team OK (%d threads)\n", threads);
    return ok ? 0 : 1;
}
