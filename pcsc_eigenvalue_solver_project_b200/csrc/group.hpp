// group.hpp -- meeting point of the host threads of a single-process multi-GPU group (no CUDA in this file, so that it
// can be unit-tested under ThreadSanitizer on a CPU-only box: tests/sanitize/group_tsan.cpp).
#pragma once
#include <condition_variable>
#include <cstdint>
#include <mutex>
#include <vector>

#include "runtime.hpp"

namespace eigsol_b200 {

// Meeting point of the P ranks of a SINGLE-PROCESS group (one host thread per GPU, EIGSOL_B200_NGPU=P): the
// setup-time collectives of a process-per-GPU group (NCCL below) become exchanges through this object.
struct LocalGroup {
    explicit LocalGroup(int n) : nranks(n), slot(size_t(n), nullptr) {}
    const int nranks;
    std::mutex mu;
    std::condition_variable cv;
    int arrived = 0;
    uint64_t generation = 0;
    bool failed = false;             // a rank left with an exception: every barrier now throws instead of waiting
    std::vector<const void*> slot;   // slot[r]: what rank r shows its peers in the current exchange

    void barrier() {
        std::unique_lock<std::mutex> lk(mu);
        if (failed) throw DeviceError("powerMethod: device error: another GPU of the group failed");
        const uint64_t gen = generation;
        if (++arrived == nranks) {
            arrived = 0;
            ++generation;
            cv.notify_all();
            return;
        }
        cv.wait(lk, [&] { return generation != gen || failed; });
        if (generation == gen) throw DeviceError("powerMethod: device error: another GPU of the group failed");
    }
    void fail() {
        std::lock_guard<std::mutex> lk(mu);
        failed = true;
        cv.notify_all();
    }
    void reset() {
        std::lock_guard<std::mutex> lk(mu);
        failed = false;
        arrived = 0;
    }
    // every rank shows `mine`, runs `body(slots)` while all are visible, and leaves together
    template <class F> void exchange(int rank, const void* mine, F&& body) {
        slot[size_t(rank)] = mine;
        barrier();
        body(slot);
        barrier();
    }
};

}  // namespace eigsol_b200
