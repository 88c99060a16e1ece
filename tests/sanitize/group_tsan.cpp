// group_tsan.cpp -- TEST INFRASTRUCTURE: the meeting point of a single-process multi-GPU group (csrc/group.hpp:
// LocalGroup -- barrier, all-to-all exchange of pointers, abort of the waiting ranks when one rank fails, reset) under
// ThreadSanitizer.  The GPU side of a group cannot run under a race detector here (compute-sanitizer is closed on the
// pool); the host threads that drive it can.  Built with g++ -fsanitize=thread by tests/test_sanitizers.py.
#include <atomic>
#include <cstdio>
#include <thread>
#include <vector>

#include "../../pcsc_eigenvalue_solver_project_b200/csrc/group.hpp"

using namespace eigsol_b200;

int main() {
    for (int P : {2, 3, 8}) {
        LocalGroup g(P);
        std::vector<long long> sums(size_t(P), 0);
        std::atomic<int> aborted{0};
        // 1. many rounds of "all-gather one value": every rank must see every rank's value of the SAME round
        {
            std::vector<std::thread> th;
            for (int r = 0; r < P; ++r)
                th.emplace_back([&, r] {
                    for (int round = 0; round < 2000; ++round) {
                        const long long mine = 1000LL * round + r;
                        g.exchange(r, &mine, [&](const std::vector<const void*>& sl) {
                            for (int q = 0; q < P; ++q) {
                                const long long v = *static_cast<const long long*>(sl[size_t(q)]);
                                if (v != 1000LL * round + q) std::abort();
                                sums[size_t(r)] += v;
                            }
                        });
                    }
                });
            for (auto& t : th) t.join();
            for (int r = 1; r < P; ++r)
                if (sums[size_t(r)] != sums[0]) return 2;
        }
        // 2. one rank fails between two barriers: the others leave their barrier with an exception instead of hanging
        {
            std::vector<std::thread> th;
            for (int r = 0; r < P; ++r)
                th.emplace_back([&, r] {
                    try {
                        g.barrier();
                        if (r == P - 1) {
                            g.fail();
                            return;
                        }
                        g.barrier();  // never completes: rank P-1 is gone
                        g.barrier();
                    } catch (const DeviceError&) {
                        aborted.fetch_add(1);
                    }
                });
            for (auto& t : th) t.join();
            if (aborted.load() != P - 1) return 3;
        }
        // 3. after reset() the group works again
        g.reset();
        {
            std::vector<std::thread> th;
            std::atomic<int> through{0};
            for (int r = 0; r < P; ++r)
                th.emplace_back([&] {
                    for (int i = 0; i < 100; ++i) g.barrier();
                    through.fetch_add(1);
                });
            for (auto& t : th) t.join();
            if (through.load() != P) return 4;
        }
    }
    std::printf("group_tsan: no sanitizer report\n");
    return 0;
}
