// runtime.hpp -- small host-runtime pieces shared by engine.cu: error plumbing, RAII device memory,
// and a lazily dlopen()ed NCCL (so a single-GPU caller never needs libnccl to be loadable, and a
// process that already carries torch's bundled NCCL reuses that copy instead of loading a second one).
#pragma once
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdint>
#include <cstdio>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

namespace eigsol_b200 {

// CUDA / NCCL failures surface as "powerMethod: device error: ..." (SURVEY.md 8b error conventions;
// the reference itself only throws std::runtime_error, src/power_method/power_method.hpp:54,57,138).
struct DeviceError : std::runtime_error {
    using std::runtime_error::runtime_error;
};
struct StateError : std::runtime_error {
    using std::runtime_error::runtime_error;
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line) {
    if (e != cudaSuccess) {
        char buf[512];
        std::snprintf(buf, sizeof buf, "powerMethod: device error: %s failed: %s (%s:%d)", what,
                      cudaGetErrorString(e), file, line);
        throw DeviceError(buf);
    }
}
#define EIGSOL_CUDA(call) ::eigsol_b200::cuda_check((call), #call, __FILE__, __LINE__)

// Stream the small zero-filled device buffers of the calling thread come from (DevBuf::alloc): the handle's own
// non-blocking stream while one of its operations runs on this thread, else none (plain cudaMalloc).  With peer access
// enabled between 8 GPUs every cudaMalloc / cudaFree maps or unmaps the block on all peers under a process-wide lock
// (~0.6 ms each, serialised over the rank threads: 10 + 9 + 17 ms of an 8-GPU group's set-up and tear-down were ~30 small
// allocations); the stream-ordered pool hands the same blocks out again for nothing.
inline thread_local cudaStream_t g_alloc_stream = nullptr;
struct AllocScope {
    cudaStream_t prev;
    explicit AllocScope(cudaStream_t s) : prev(g_alloc_stream) { g_alloc_stream = s; }
    ~AllocScope() { g_alloc_stream = prev; }
};

// ---- RAII device / pinned memory ---------------------------------------------------------------
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    bool pooled = false;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes), pooled(o.pooled) { o.p = nullptr; o.bytes = 0; o.pooled = false; }
    DevBuf& operator=(DevBuf&& o) noexcept {
        if (this != &o) {
            release();
            p = o.p;
            bytes = o.bytes;
            pooled = o.pooled;
            o.p = nullptr;
            o.bytes = 0;
            o.pooled = false;
        }
        return *this;
    }
    ~DevBuf() { release(); }
    void release() {
        if (p) {
            // pooled memory goes back to the device's stream-ordered pool (no device synchronisation, no driver
            // round trip; the next ingest reuses it); callers only release buffers whose work has completed
            if (pooled) cudaFreeAsync(p, cudaStreamPerThread);
            else cudaFree(p);
        }
        p = nullptr;
        bytes = 0;
        pooled = false;
    }
    // Large, private (never shared with a peer GPU or another process) buffers: from the device's default
    // stream-ordered memory pool.  No zero fill unless asked.  The memory may be used on OTHER streams only after
    // `st` has been synchronised (or an event dependency) -- callers synchronise once after a batch of allocations.
    void alloc_pooled(size_t n, cudaStream_t st, bool zero = false, size_t pad = 64) {
        release();
        bytes = n;
        const cudaError_t e = cudaMallocAsync(&p, n + pad, st);
        if (e != cudaSuccess) {  // no pool support (old driver / MPS quirk): plain allocation
            cudaGetLastError();
            p = nullptr;
            if (zero) alloc(n, pad);
            else alloc_raw(n, pad);
            return;
        }
        pooled = true;
        // the pad is ALWAYS zero: kernels over-read index arrays by a vector width and use what they read as a gather
        // index (recycled pool memory holds garbage)
        if (zero) EIGSOL_CUDA(cudaMemsetAsync(p, 0, n + pad, st));
        else if (pad) EIGSOL_CUDA(cudaMemsetAsync(static_cast<char*>(p) + n, 0, pad, st));
    }
    // pad: extra zeroed bytes behind the payload so vector loads may over-read safely
    // Plain cudaMalloc, zero-filled and synchronised: for memory that is exported to other processes (CUDA IPC needs
    // cudaMalloc memory) or mapped by peers without a shared pool.
    void alloc_plain(size_t n, size_t pad = 64) {
        release();
        bytes = n;
        EIGSOL_CUDA(cudaMalloc(&p, n + pad));
        // cudaMemset runs on the legacy stream and is asynchronous for device memory, while uploads go
        // to non-blocking streams: without this synchronisation the zero fill may land AFTER an upload.
        EIGSOL_CUDA(cudaMemset(p, 0, n + pad));
        EIGSOL_CUDA(cudaDeviceSynchronize());
    }
    // The default for small buffers: zero-filled and ready for use on any stream when the call returns.  From the
    // device's stream-ordered pool on the handle's own stream when one is current on this thread (g_alloc_stream), else
    // plain.  (Routing these through cudaStreamPerThread instead was measured: the peer-map phase of a 2-GPU group went
    // from 0.4 ms to 32 ms -- per-thread streams synchronise with the legacy stream the plain copies use.)
    void alloc(size_t n, size_t pad = 64) {
        if (!g_alloc_stream) {
            alloc_plain(n, pad);
            return;
        }
        release();
        bytes = n;
        if (cudaMallocAsync(&p, n + pad, g_alloc_stream) != cudaSuccess) {
            cudaGetLastError();
            p = nullptr;
            alloc_plain(n, pad);
            return;
        }
        pooled = true;
        EIGSOL_CUDA(cudaMemsetAsync(p, 0, n + pad, g_alloc_stream));
        EIGSOL_CUDA(cudaStreamSynchronize(g_alloc_stream));
    }
    // no zero fill, no synchronisation: for buffers every byte of which is written before it is read
    void alloc_raw(size_t n, size_t pad = 64) {
        release();
        bytes = n;
        EIGSOL_CUDA(cudaMalloc(&p, n + pad));
        if (pad) EIGSOL_CUDA(cudaMemset(static_cast<char*>(p) + n, 0, pad));  // over-read area: always zero
    }
    void upload(const void* src, size_t n, cudaStream_t s) {
        if (n) EIGSOL_CUDA(cudaMemcpyAsync(p, src, n, cudaMemcpyHostToDevice, s));
    }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

// Small pinned host blocks (the 200-byte iteration state a handle polls) are recycled for the life of the process:
// cudaMallocHost takes a process-wide driver lock and milliseconds -- per handle, and serialised over the rank
// threads of a group, it was most of a group handle's 9-11 ms of "streams + state" set-up.
struct SmallPinnedPool {
    static constexpr size_t kBytes = 4096;
    std::mutex mu;
    std::vector<void*> idle;
    static SmallPinnedPool& get() {
        static SmallPinnedPool* p = new SmallPinnedPool;  // never destroyed: blocks may be returned during static destruction
        return *p;
    }
    void* acquire() {
        {
            std::lock_guard<std::mutex> lk(mu);
            if (!idle.empty()) {
                void* q = idle.back();
                idle.pop_back();
                return q;
            }
        }
        void* q = nullptr;
        EIGSOL_CUDA(cudaHostAlloc(&q, kBytes, cudaHostAllocPortable));
        return q;
    }
    void release(void* q) {
        std::lock_guard<std::mutex> lk(mu);
        idle.push_back(q);
    }
};

struct PinnedBuf {
    void* p = nullptr;
    bool small = false;
    PinnedBuf() = default;
    PinnedBuf(const PinnedBuf&) = delete;
    PinnedBuf& operator=(const PinnedBuf&) = delete;
    ~PinnedBuf() { release(); }
    void release() {
        if (p) {
            if (small) SmallPinnedPool::get().release(p);
            else cudaFreeHost(p);
        }
        p = nullptr;
    }
    void alloc(size_t n) {
        release();
        small = n <= SmallPinnedPool::kBytes;
        if (small) p = SmallPinnedPool::get().acquire();
        else EIGSOL_CUDA(cudaMallocHost(&p, n));
    }
    template <class T> T* as() const { return static_cast<T*>(p); }
};

// ---- NCCL, loaded on first use -----------------------------------------------------------------
// Only the handful of entry points this path needs; signatures as in nccl.h 2.27 (ABI-stable).
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclInt8 = 0, ncclChar = 0, ncclUint8 = 1, ncclInt32 = 2, ncclInt = 2, ncclUint32 = 3, ncclInt64 = 4,
       ncclUint64 = 5, ncclFloat16 = 6, ncclFloat32 = 7, ncclFloat = 7, ncclFloat64 = 8, ncclDouble = 8 };
enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 };

struct NcclApi {
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    void* lib = nullptr;
};

inline NcclApi& nccl() {
    static NcclApi api = [] {
        NcclApi a;
        // prefer a copy that is already in the process (torch's bundled NCCL), else the system one
        a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
        if (!a.lib) a.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
        if (!a.lib) a.lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
        if (!a.lib) return a;
        auto sym = [&](const char* n) { return dlsym(a.lib, n); };
        a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
        a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
        a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
        a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
        a.AllReduce = reinterpret_cast<decltype(a.AllReduce)>(sym("ncclAllReduce"));
        a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
        a.Broadcast = reinterpret_cast<decltype(a.Broadcast)>(sym("ncclBroadcast"));
        a.Send = reinterpret_cast<decltype(a.Send)>(sym("ncclSend"));
        a.Recv = reinterpret_cast<decltype(a.Recv)>(sym("ncclRecv"));
        a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
        a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
        return a;
    }();
    if (!api.lib || !api.GetUniqueId || !api.CommInitRank || !api.AllReduce || !api.Send || !api.Recv ||
        !api.GroupStart || !api.GroupEnd || !api.Broadcast)
        throw DeviceError("powerMethod: device error: NCCL (libnccl.so.2) could not be loaded");
    return api;
}

inline void nccl_check(int r, const char* what, const char* file, int line) {
    if (r != ncclSuccess) {
        char buf[512];
        const char* msg = nccl().GetErrorString ? nccl().GetErrorString(r) : "?";
        std::snprintf(buf, sizeof buf, "powerMethod: device error: %s failed: %s (%s:%d)", what, msg, file, line);
        throw DeviceError(buf);
    }
}
#define EIGSOL_NCCL(call) ::eigsol_b200::nccl_check((call), #call, __FILE__, __LINE__)

}  // namespace eigsol_b200
