// ingest.cuh -- device-side ingest of the reference's sparse matrix (single GPU).
//
// EigSol::Matrix::Sparse<Scalar> is Eigen::SparseMatrix<Scalar, ColMajor, int> (reference:
// src/matrix/matrix.hpp:43-44), i.e. CSC.  The matvec kernels want CSR with ascending columns inside a row
// (the reference's accumulation order, eigen/Eigen/src/SparseCore/SparseDenseProduct.h:101-110).  The
// transpose used to be a counting sort on one host core (1.1 s for the 117M entries of the 256^3 Laplacian,
// 20x the 200 power iterations that follow).  Here the three CSC arrays are pushed to the device through
// pinned staging buffers by a few host threads, and the transpose runs there:
//     col[p]    = column of entry p                       (binary search in outer, one thread per entry)
//     perm      = stable radix sort of p by row = inner[p] (cub::DeviceRadixSort -- library code, setup only)
//     rowptr    = exclusive scan of the row histogram
//     colidx[k] = col[perm[k]],  vals[k] = values[perm[k]]
// Stable sort + CSC order (column-major, rows ascending inside a column) = columns ascending inside a row.
// This is setup, not the hot path: it runs once per Matrix.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#if defined(__SSE2__)
#include <emmintrin.h>
#endif
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <thread>
#include <vector>

#include "runtime.hpp"

namespace eigsol_b200 {

// ---- host -> device copy of a large pageable array through pinned staging, several threads ----------
// Pinned staging buffers live for the life of the process (cudaHostAlloc costs milliseconds).  Uploads of different
// handles / ranks may run concurrently, so buffers are handed out and taken back under a lock.
struct PinnedPool {
    static constexpr size_t kBytes = size_t(16) << 20;
    static constexpr size_t kKeep = 64;  // at most 1 GB of pinned staging is kept
    std::mutex mu;
    std::vector<void*> idle;
    static PinnedPool& get() {
        static PinnedPool p;
        return p;
    }
    cudaError_t acquire(void** out) {
        {
            std::lock_guard<std::mutex> lk(mu);
            if (!idle.empty()) {
                *out = idle.back();
                idle.pop_back();
                return cudaSuccess;
            }
        }
        return cudaHostAlloc(out, kBytes, cudaHostAllocPortable);
    }
    void release(void* p) {
        if (!p) return;
        {
            std::lock_guard<std::mutex> lk(mu);
            if (idle.size() < kKeep) {
                idle.push_back(p);
                return;
            }
        }
        cudaFreeHost(p);
    }
};

// Copies between large PAGEABLE host arrays and the device through pinned staging buffers, several host threads, each
// with its own stream and two buffers (the memcpy of one chunk overlaps the DMA of the previous one).  Jobs run in the
// order they were added.  start() returns at once; wait_prefix(k) blocks until the first k jobs have landed (so device
// work that needs only those can start while the rest is still in flight); finish() joins and returns the first CUDA
// error seen.  download = true reverses the direction (device -> pinned -> host array).
// Host copy of one chunk into a pinned staging buffer.  EIGSOL_B200_STAGE_NT=1: non-temporal 16-byte stores (the
// destination is only read by the DMA engine, so the lines need not be fetched nor kept in the caches of the core).
inline void stage_copy(void* dst, const void* src, size_t bytes) {
#if defined(__SSE2__)
    static const bool nt = [] {
        const char* e = std::getenv("EIGSOL_B200_STAGE_NT");
        return e && *e && std::atoi(e) != 0;
    }();
    if (nt && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
        const size_t blocks = bytes / 64;
        const __m128i* s = static_cast<const __m128i*>(src);
        __m128i* d = static_cast<__m128i*>(dst);
        for (size_t i = 0; i < blocks; ++i) {
            const __m128i a = _mm_loadu_si128(s + 4 * i), b = _mm_loadu_si128(s + 4 * i + 1);
            const __m128i c = _mm_loadu_si128(s + 4 * i + 2), e = _mm_loadu_si128(s + 4 * i + 3);
            _mm_stream_si128(d + 4 * i, a);
            _mm_stream_si128(d + 4 * i + 1, b);
            _mm_stream_si128(d + 4 * i + 2, c);
            _mm_stream_si128(d + 4 * i + 3, e);
        }
        _mm_sfence();
        const size_t done = blocks * 64;
        if (done < bytes) std::memcpy(static_cast<char*>(dst) + done, static_cast<const char*>(src) + done, bytes - done);
        return;
    }
#endif
    std::memcpy(dst, src, bytes);
}

struct StagedUploader {
    // bytes per job: small enough that the first DMA starts early and the tail is short, large enough that the per-job
    // cost (one cudaMemcpyAsync + one event) vanishes; EIGSOL_B200_UPLOAD_CHUNK_MB in [1, 16], default 4
    static size_t chunk_bytes() {
        static const size_t c = [] {
            const char* e = std::getenv("EIGSOL_B200_UPLOAD_CHUNK_MB");
            long mb = e && *e ? std::atol(e) : 4;
            mb = mb < 1 ? 1 : (mb > 16 ? 16 : mb);
            return size_t(mb) << 20;
        }();
        return c;
    }
    struct Job {
        char* dst;
        const char* src;
        size_t bytes;
    };
    std::vector<Job> jobs;
    bool download = false;
    std::vector<std::thread> pool;
    std::atomic<size_t> next{0};
    std::atomic<int> err{int(cudaSuccess)};
    std::mutex mu;
    std::condition_variable cv;
    size_t prefix = 0;         // jobs [0, prefix) are the ones wait_prefix() waits for
    int past_prefix = 0;       // threads that have completed everything they took from the prefix
    int nthreads_run = 0;

    ~StagedUploader() {
        for (auto& t : pool)
            if (t.joinable()) t.join();
    }
    size_t add(void* dst, const void* src, size_t bytes) {
        const size_t chunk = chunk_bytes();
        for (size_t off = 0; off < bytes; off += chunk)
            jobs.push_back(Job{static_cast<char*>(dst) + off, static_cast<const char*>(src) + off,
                               std::min(chunk, bytes - off)});
        return jobs.size();
    }
    void worker(int device) {
        cudaSetDevice(device);
        void* pin[2] = {nullptr, nullptr};
        cudaStream_t st = nullptr;
        cudaEvent_t ev[2] = {nullptr, nullptr};
        cudaError_t e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
        for (int b = 0; b < 2 && e == cudaSuccess; ++b) {
            e = PinnedPool::get().acquire(&pin[b]);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming);
        }
        int b = 0;
        bool used[2] = {false, false};
        const Job* pending[2] = {nullptr, nullptr};  // download: chunk sitting in pin[b] that still has to go to the host array
        bool passed = false;
        auto drain = [&](int k) {  // the DMA of buffer k has finished; a download now copies it out
            if (!used[k]) return cudaSuccess;
            const cudaError_t r = cudaEventSynchronize(ev[k]);
            if (r == cudaSuccess && pending[k]) std::memcpy(pending[k]->dst, pin[k], pending[k]->bytes);
            pending[k] = nullptr;
            used[k] = false;
            return r;
        };
        auto pass_prefix = [&] {
            if (passed) return;
            passed = true;
            std::lock_guard<std::mutex> lk(mu);
            ++past_prefix;
            cv.notify_all();
        };
        while (e == cudaSuccess) {
            const size_t i = next.fetch_add(1);
            if (i >= prefix && !passed) {
                // everything this thread took from the prefix must have landed before it reports
                for (int k = 0; k < 2 && e == cudaSuccess; ++k) e = drain(k);
                pass_prefix();
            }
            if (i >= jobs.size() || e != cudaSuccess) break;
            e = drain(b);
            if (e != cudaSuccess) break;
            const Job& j = jobs[i];
            if (download) {
                e = cudaMemcpyAsync(pin[b], j.src, j.bytes, cudaMemcpyDeviceToHost, st);
                pending[b] = &j;
            } else {
                stage_copy(pin[b], j.src, j.bytes);
                e = cudaMemcpyAsync(j.dst, pin[b], j.bytes, cudaMemcpyHostToDevice, st);
            }
            if (e == cudaSuccess) e = cudaEventRecord(ev[b], st);
            used[b] = true;
            b ^= 1;
        }
        for (int k = 0; k < 2; ++k) {
            const cudaError_t e2 = drain(k);
            if (e == cudaSuccess) e = e2;
        }
        pass_prefix();
        if (st) {
            const cudaError_t e2 = cudaStreamSynchronize(st);
            if (e == cudaSuccess) e = e2;
        }
        for (int k = 0; k < 2; ++k) {
            if (ev[k]) cudaEventDestroy(ev[k]);
            PinnedPool::get().release(pin[k]);
        }
        if (st) cudaStreamDestroy(st);
        if (e != cudaSuccess) err.store(int(e));
    }
    void start(int device, int nthreads, size_t prefix_jobs = 0) {
        prefix = std::min(prefix_jobs, jobs.size());
        nthreads_run = int(std::max<size_t>(1, std::min<size_t>(size_t(std::min(nthreads, 16)), std::max<size_t>(jobs.size(), 1))));
        for (int t = 0; t < nthreads_run; ++t) pool.emplace_back([this, device] { worker(device); });
    }
    void wait_prefix() {
        std::unique_lock<std::mutex> lk(mu);
        cv.wait(lk, [&] { return past_prefix >= nthreads_run; });
    }
    cudaError_t finish() {
        for (auto& t : pool) t.join();
        pool.clear();
        jobs.clear();
        return cudaError_t(err.load());
    }
    // Runs all jobs and waits; returns the first CUDA error seen (cudaSuccess if none).
    cudaError_t run(int device, int nthreads) {
        if (jobs.empty()) return cudaSuccess;
        start(device, nthreads);
        return finish();
    }
};

// ---- kernels ----------------------------------------------------------------------------------------
// flags: bit 1 = row index out of range (outer is validated on the host: it sizes the allocations)
// One thread per entry: its column (upper_bound in outer), iota, range check and the row histogram.
__global__ void csc_expand_kernel(const int32_t* __restrict__ outer, const int32_t* __restrict__ inner, int64_t cols,
                                  int64_t rows, int64_t nnz, int32_t* __restrict__ col, int32_t* __restrict__ iota,
                                  int32_t* __restrict__ rowcount, int* flags) {
    const int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= nnz) return;
    int64_t lo = 0, hi = cols;  // largest j with outer[j] <= p
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(outer + mid) <= p) lo = mid;
        else hi = mid;
    }
    col[p] = int32_t(lo);
    iota[p] = int32_t(p);
    const int32_t r = inner[p];
    if (r < 0 || r >= rows) atomicOr(flags, 2);
    else atomicAdd(rowcount + r, 1);
}

// flags bit 0: outer not monotone / out of [0, nnz] (host only checked outer[0] == 0 and read outer[cols])
__global__ void outer_check_kernel(const int32_t* __restrict__ outer, int64_t cols, int64_t nnz, int* flags) {
    const int64_t j = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (j >= cols) return;
    const int32_t a = outer[j], b = outer[j + 1];
    if (b < a || a < 0 || b > nnz) atomicOr(flags, 1);
}

template <class S>
__global__ void csc_gather_kernel(const int32_t* __restrict__ perm, const int32_t* __restrict__ col,
                                  const S* __restrict__ cvals, int64_t nnz, int32_t* __restrict__ colidx,
                                  S* __restrict__ vals) {
    const int64_t k = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    const int32_t p = perm[k];
    colidx[k] = col[p];
    vals[k] = cvals[p];
}

// Row-length statistics on the device (what plan.hpp: row_stats computes on the host).
// out[0] = max len, out[1] = min len, out[2..3] = padded entries of 64-row slices (u64), dsum[0] = sum (len-mean)^2
__global__ void row_stats_kernel(const int32_t* __restrict__ rowptr, int64_t rows, double mean, int* imax, int* imin,
                                 unsigned long long* padded, double* var) {
    // one thread per 64-row slice
    const int64_t sl = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t r0 = sl * 64;
    if (r0 >= rows) return;
    const int64_t r1 = r0 + 64 < rows ? r0 + 64 : rows;
    int mx = 0, mn = INT32_MAX;
    double v = 0.0;
    int32_t prev = rowptr[r0];
    for (int64_t r = r0; r < r1; ++r) {
        const int32_t nx = rowptr[r + 1];
        const int len = nx - prev;
        prev = nx;
        mx = len > mx ? len : mx;
        mn = len < mn ? len : mn;
        const double d = len - mean;
        v += d * d;
    }
    atomicMax(imax, mx);
    atomicMin(imin, mn);
    atomicAdd(padded, (unsigned long long)mx * 64ull);
    atomicAdd(var, v);
}

// ---- single-process group, CSC input: every GPU uploads one column slab; each rank then pulls the entries of
// ---- ITS rows out of every slab over NVLink (peer loads), in slab order = ascending columns
// Global column of every entry of a column slab (outer rebased to the slab: souter[0] == 0) + range check of rows.
__global__ void slab_expand_kernel(const int32_t* __restrict__ souter, const int32_t* __restrict__ sinner, int64_t scols,
                                   int64_t col0, int64_t rows, int64_t snnz, int32_t* __restrict__ scol, int* flags) {
    const int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= snnz) return;
    int64_t lo = 0, hi = scols;  // largest j with souter[j] <= p
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if (__ldg(souter + mid) <= p) lo = mid;
        else hi = mid;
    }
    scol[p] = int32_t(col0 + lo);
    const int32_t r = sinner[p];
    if (r < 0 || r >= rows) atomicOr(flags, 2);
}

// number of entries of a (peer's) slab whose row lies in [lo, hi)
__global__ void slab_count_kernel(const int32_t* __restrict__ sinner, int64_t snnz, int32_t lo, int32_t hi,
                                  unsigned long long* count) {
    unsigned long long c = 0;
    for (int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; p < snnz; p += int64_t(gridDim.x) * blockDim.x) {
        const int32_t r = sinner[p];
        c += (r >= lo && r < hi) ? 1ull : 0ull;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

struct RowInRange {
    const int32_t* sinner;
    int32_t lo, hi;
    __device__ __forceinline__ bool operator()(int32_t p) const {
        const int32_t r = sinner[p];
        return r >= lo && r < hi;
    }
};

// copies the selected entries of a (peer's) slab into this rank's COO arrays: key = row - lo, global column, value
template <class S>
__global__ void slab_pull_kernel(const int32_t* __restrict__ sel, int64_t nsel, const int32_t* __restrict__ sinner,
                                 const int32_t* __restrict__ scol, const S* __restrict__ svals, int32_t lo,
                                 int32_t* __restrict__ key, int32_t* __restrict__ col, S* __restrict__ val,
                                 int32_t* __restrict__ iota, int32_t iota0, int32_t* __restrict__ rowcount) {
    const int64_t k = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (k >= nsel) return;
    const int32_t p = sel[k];
    const int32_t r = sinner[p] - lo;
    key[k] = r;
    col[k] = scol[p];
    val[k] = svals[p];
    iota[k] = iota0 + int32_t(k);
    atomicAdd(rowcount + r, 1);
}

// rowptr of a row block cut out of a larger CSR array: make it start at 0
__global__ void rebase_kernel(int32_t* __restrict__ rowptr, int64_t count, int32_t base) {
    const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < count) rowptr[i] -= base;
}

// entries with |column - row| <= 1 (GLOBAL column indices; row0 = global index of the block's first row): the
// "tridiagonal core" statistic of the kernel choice (plan.hpp).  One thread per row.
__global__ void near_diag_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, int64_t rows,
                                 int64_t row0, unsigned long long* count) {
    unsigned long long c = 0;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < rows; i += int64_t(gridDim.x) * blockDim.x) {
        const int64_t g = row0 + i;
        for (int32_t p = rowptr[i]; p < rowptr[i + 1]; ++p) {
            const int64_t d = int64_t(colidx[p]) - g;
            c += (d >= -1 && d <= 1) ? 1ull : 0ull;
        }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_down_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(count, c);
}

// ---- row-block ingest (multi-GPU: each rank holds rows [lo, hi) in CSR with GLOBAL columns) ----------
// mark[c] = 1 for every referenced column outside [lo, hi); bit 1 of flags = column out of range
__global__ void rows_mark_kernel(const int32_t* __restrict__ colidx, int64_t nnz, int64_t n, int64_t lo, int64_t hi,
                                 int32_t* __restrict__ mark, int* flags) {
    const int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= nnz) return;
    const int32_t c = colidx[p];
    if (c < 0 || c >= n) {
        atomicOr(flags, 2);
        return;
    }
    if ((c < lo || c >= hi) && mark[c] == 0) mark[c] = 1;
}

// pos = exclusive scan of mark: the sorted distinct remote columns, compacted
__global__ void halo_compact_kernel(const int32_t* __restrict__ mark, const int32_t* __restrict__ pos, int64_t n,
                                    int32_t* __restrict__ halo_cols) {
    const int64_t c = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (c < n && mark[c]) halo_cols[pos[c]] = int32_t(c);
}

// global column -> extended local numbering [own rows | pad | halo]
__global__ void rows_localize_kernel(int32_t* __restrict__ colidx, int64_t nnz, int64_t lo, int64_t hi, int64_t nloc_pad,
                                     const int32_t* __restrict__ pos) {
    const int64_t p = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (p >= nnz) return;
    const int32_t c = colidx[p];
    colidx[p] = (c >= lo && c < hi) ? int32_t(c - lo) : int32_t(nloc_pad + pos[c]);
}

// def[u] = 1 if any row of work unit u has an entry in the halo (column >= nloc_pad).  Units are either
// `rows_per_unit` consecutive rows, or the row blocks [rowblk[u], rowblk[u+1]) when rowblk != nullptr.
// One warp per unit.
__global__ void deferred_units_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ colidx, int64_t rows,
                                      int64_t nunits, int rows_per_unit, const int32_t* __restrict__ rowblk,
                                      int32_t nloc_pad, uint8_t* __restrict__ def) {
    const int64_t u = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (u >= nunits) return;
    int64_t r0, r1;
    if (rowblk) {
        r0 = rowblk[u];
        r1 = rowblk[u + 1];
    } else {
        r0 = u * rows_per_unit;
        r1 = r0 + rows_per_unit;
    }
    r1 = r1 < rows ? r1 : rows;
    int any = 0;
    if (r0 < r1) {
        const int32_t p0 = rowptr[r0], p1 = rowptr[r1];
        for (int32_t p = p0 + lane; p < p1; p += 32)
            if (colidx[p] >= nloc_pad) any = 1;
    }
    any = __any_sync(0xffffffffu, any);
    if (lane == 0) def[u] = any ? 1 : 0;
}

}  // namespace eigsol_b200
