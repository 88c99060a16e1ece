// engine.cu -- host runtime + C ABI of libeigsol_b200.so (see include/eigsol_b200.h).
//
// Replaces the loop of EigSol::powerMethodImpl (reference: src/power_method/power_method.hpp:61-98)
// by launches of the fused kernels in kernels.cuh.  The host side here only plans (plan.hpp), uploads,
// issues launches and polls the device-resident PowerState; no arithmetic of the path runs on the CPU
// and there is no CPU fallback: without a CUDA device every compute entry point fails loudly.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <map>
#include <cstdio>
#include <climits>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/eigsol_b200.h"
#include "fileio.hpp"
#include "group.hpp"
#include "ingest.cuh"
#include "kernels.cuh"
#include "plan.hpp"
#include "probe.cuh"
#include "runtime.hpp"

namespace eigsol_b200 {

static thread_local std::string g_last_error;
static thread_local const void* g_preload_x0 = nullptr;  // set by the one-shot CSC call around its matrix creation
static std::atomic<int64_t> g_matrices_created{0};  // device uploads of whole matrices so far (tests of the Matrix cache)

// -------------------------------------------------------------------------------------------------
// One rank of a multi-GPU group: a process-per-GPU rank (NCCL communicator, peers reached through CUDA IPC)
// or one thread of a single-process group (`local`, peers reached through plain peer access).
// -------------------------------------------------------------------------------------------------
struct Comm {
    int rank = 0, nranks = 1, device = 0;
    ncclComm_t nccl_comm = nullptr;
    LocalGroup* local = nullptr;
    std::vector<int> local_devices;  // single-process group: CUDA device of every rank
    bool pool_shared = false;        // ... and this device's memory pool is readable / writable by every peer of the group
    cudaStream_t stream = nullptr;
    DevBuf scratch;  // small device scratch for setup-time collectives
    ~Comm() {
        if (nccl_comm) nccl().CommDestroy(nccl_comm);
        if (stream) cudaStreamDestroy(stream);
    }
};

// EIGSOL_B200_TRACE=1: wall-clock phase marks on stderr (setup paths only)
struct Trace {
    bool on;
    std::chrono::steady_clock::time_point t0;
    Trace() : on(std::getenv("EIGSOL_B200_TRACE") != nullptr), t0(std::chrono::steady_clock::now()) {}
    void mark_async(const char* what) {  // host-side mark: does not wait for the device
        if (!on) return;
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[eigsol_b200] %-36s %8.2f ms (host)\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
    void mark(const char* what) {
        if (!on) return;
        cudaDeviceSynchronize();
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[eigsol_b200] %-36s %8.2f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

static int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v && *v ? std::atoi(v) : dflt;
}

// Host threads feeding the pinned staging buffers of an upload.  One GPU: half the cores (16-core box, one-shot C5 call:
// 2 threads 940 it/s, 4: 1100, 8: 1200, 16: slower).  A single-process group: ALL cores divided by the ranks -- every
// rank feeds its own PCIe link and the calling thread only waits (2 GPUs, 16 cores, 20-trip call: 4 per rank 67.8 ms,
// 6: 61.7, 8: 59.3).  Process-per-GPU ranks share the host with their launcher's other work: half the cores / ranks.
static int upload_threads(int nranks, bool local_group = false) {
    const unsigned hw = std::max(2u, std::thread::hardware_concurrency());
    const unsigned share = local_group ? hw / unsigned(std::max(nranks, 1)) : hw / 2 / unsigned(std::max(nranks, 1));
    const int dflt = int(std::min(8u, std::max(1u, share)));
    return std::max(1, env_int("EIGSOL_B200_UPLOAD_THREADS", std::max(dflt, nranks == 1 ? 2 : 1)));
}

static void require_device() {
    int n = 0;
    const cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        cudaGetLastError();
        throw DeviceError(std::string("powerMethod: device error: no CUDA device available (") +
                          (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                          "); this path has no CPU fallback");
    }
}

// all-gather of one int64 per rank (setup only)
static std::vector<int64_t> comm_allgather_i64(Comm* c, int64_t v) {
    std::vector<int64_t> out(size_t(c->nranks), 0);
    if (c->nranks == 1) {
        out[0] = v;
        return out;
    }
    if (c->local) {
        c->local->exchange(c->rank, &v, [&](const std::vector<const void*>& sl) {
            for (int r = 0; r < c->nranks; ++r) out[size_t(r)] = *static_cast<const int64_t*>(sl[size_t(r)]);
        });
        return out;
    }
    int64_t* d = c->scratch.as<int64_t>();
    EIGSOL_CUDA(cudaMemcpyAsync(d + c->nranks + c->rank, &v, sizeof v, cudaMemcpyHostToDevice, c->stream));
    // AllGather may be absent from a very old libnccl; grouped broadcasts are equivalent
    EIGSOL_NCCL(nccl().GroupStart());
    for (int r = 0; r < c->nranks; ++r)
        EIGSOL_NCCL(nccl().Broadcast(d + c->nranks + c->rank, d + r, 1, ncclInt64, r, c->nccl_comm, c->stream));
    EIGSOL_NCCL(nccl().GroupEnd());
    EIGSOL_CUDA(cudaMemcpyAsync(out.data(), d, sizeof(int64_t) * c->nranks, cudaMemcpyDeviceToHost, c->stream));
    EIGSOL_CUDA(cudaStreamSynchronize(c->stream));
    return out;
}

// all-gather of a fixed-size blob per rank (setup only): out[r * bytes ...] = rank r's blob
static std::vector<char> comm_allgather_bytes(Comm* c, const void* mine, size_t bytes) {
    std::vector<char> out(bytes * size_t(c->nranks));
    if (c->local) {
        c->local->exchange(c->rank, mine, [&](const std::vector<const void*>& sl) {
            for (int r = 0; r < c->nranks; ++r) std::memcpy(out.data() + bytes * size_t(r), sl[size_t(r)], bytes);
        });
        return out;
    }
    DevBuf d;
    d.alloc_plain(out.size());  // NCCL buffer
    EIGSOL_CUDA(cudaMemcpyAsync(d.as<char>() + bytes * c->rank, mine, bytes, cudaMemcpyHostToDevice, c->stream));
    EIGSOL_NCCL(nccl().GroupStart());
    for (int r = 0; r < c->nranks; ++r)
        EIGSOL_NCCL(nccl().Broadcast(d.as<char>() + bytes * r, d.as<char>() + bytes * r, bytes, ncclChar, r, c->nccl_comm,
                                     c->stream));
    EIGSOL_NCCL(nccl().GroupEnd());
    EIGSOL_CUDA(cudaMemcpyAsync(out.data(), d.p, out.size(), cudaMemcpyDeviceToHost, c->stream));
    EIGSOL_CUDA(cudaStreamSynchronize(c->stream));
    return out;
}

// min over ranks of an int (setup only): lets every rank take the same decision
static int comm_min_int(Comm* c, int v) {
    if (c->nranks == 1) return v;
    if (c->local) {
        int mn = v;
        c->local->exchange(c->rank, &v, [&](const std::vector<const void*>& sl) {
            for (int r = 0; r < c->nranks; ++r) mn = std::min(mn, *static_cast<const int*>(sl[size_t(r)]));
        });
        return mn;
    }
    int* d = reinterpret_cast<int*>(c->scratch.as<int64_t>() + 3 * c->nranks + 8);
    EIGSOL_CUDA(cudaMemcpyAsync(d, &v, sizeof v, cudaMemcpyHostToDevice, c->stream));
    EIGSOL_NCCL(nccl().AllReduce(d, d, 1, ncclInt32, ncclMin, c->nccl_comm, c->stream));
    EIGSOL_CUDA(cudaMemcpyAsync(&v, d, sizeof v, cudaMemcpyDeviceToHost, c->stream));
    EIGSOL_CUDA(cudaStreamSynchronize(c->stream));
    return v;
}

// -------------------------------------------------------------------------------------------------
// The device-resident matrix + iteration state behind an eigsol_b200_matrix handle.
// -------------------------------------------------------------------------------------------------
struct MatrixImpl {
    int dtype = 0;
    bool dense = false;
    int64_t n = 0, nnz_global = 0;
    Comm* comm = nullptr;
    int rank = 0, nranks = 1, device = 0, sm_count = 148;

    // partition and halo
    std::vector<int64_t> row_begin;
    int64_t lo = 0, hi = 0, nloc = 0, nloc_pad = 0, nhalo = 0, ext_len = 0, local_nnz = 0;
    HaloPlan halo;
    std::vector<int64_t> send_count, send_off, send_start;  // per peer
    std::vector<char> send_contig;
    bool need_pack = false;
    DevBuf d_send_idx, d_sendbuf;
    // x exchange + scalar all-reduce: MODE_FUSED (one GPU), MODE_SPLIT (NCCL), MODE_PEER (peer memory)
    int xmode = MODE_FUSED;
    PeerArgs peer_h{};
    DevBuf d_peer, d_scal_in, d_flag_in, d_stats;
    unsigned long long gate_epoch = 0;  // MODE_PEER: gates passed so far (identical on every rank)
    std::vector<void*> ipc_opened;

    // sparse layouts
    CsrHost csr;  // local rows, localized columns (kept on the host to rebuild layouts)
    RowStats stats;
    DevBuf d_rowptr, d_colidx, d_vals;
    DevBuf d_sell_off, d_sell_col, d_sell_val;
    int sell_R = 0;
    int64_t sell_nslices = 0, sell_padded = 0;
    DevBuf d_rowblk;
    std::vector<int32_t> rowblk_h;
    int nblk = 0;
    DevBuf d_window;    // CSR_WINDOW: per 256-row unit {first column, count} staged by the TMA bulk copy
    int window_smem = 0;  // dynamic shared memory of the window kernel (bytes)
    DevBuf d_deferred;  // MODE_PEER: per work unit of the chosen kernel, 1 = reads halo entries
    bool have_deferred = false;
    int unit_start = 0;  // rotation of the unit walk: middle of the longest halo-free run of units
    int serpentine = 0;
    bool pdl = true;  // programmatic dependent launch of the matvec kernels
    int scalar_block = 256;  // EIGSOL_B200_SCALAR_BLOCK: threads per CTA (= rows per work unit) of csr_scalar
    int dense_sym = 0;  // one GPU, dense: 1 = A == A^T, 2 = A == A^H (bitwise, checked at upload) -> gemv_sym_kernel
    int64_t sym_per = 0;
    int sym_nctas = 0, sym_smem = 0;
    int gemv_map = 0;  // EIGSOL_B200_GEMV_MAP: 1 = row tiles of one column segment are neighbouring CTAs
    int gemv_var = 0;  // EIGSOL_B200_GEMV_VAR: streaming loop variant of the dense kernel (kernels.cuh)

    // dense layout
    DevBuf d_dense, d_dense_partial, d_tile_ticket;
    int64_t lda = 0;
    int nseg = 1, ntiles = 1;

    // vectors and state
    DevBuf buf[kRing], d_state, d_partials, d_scal, d_xloc, d_xglob;
    PinnedBuf h_state;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    // launch configuration
    int kernel = K_CSR_SCALAR, lanes = 0, grid = 1, block = 256, nslots = 1;
    int64_t stored_bytes = 0;
    double sigma0 = 1.0;   // power of two near 1/max|a_ij|: scale of the first two launches of a solve
    int64_t launches = 0;  // kernel launches issued over the handle's life
    int issued = 0;        // matvec launches issued since begin (ring phase on the host)
    unsigned long long flag_epoch = 0;  // MODE_PEER: flag base of the current solve (identical on every rank)
    bool begun = false;
    int max_trips = 0;
    const void* preload_x0 = nullptr;  // one-shot call: host start vector to upload together with the matrix
    DevBuf d_x0pre;                    // group rank: this rank's rows of the preloaded start vector (copied into buf[0] by begin)
    bool x0_preloaded = false;         // ... and it is in buf[0] already (consumed by the next begin() with that pointer)
    cudaGraphExec_t graph = nullptr, graph_short = nullptr;  // graph_len launches / kRing launches
    std::map<std::pair<int, int>, cudaGraphExec_t> batch_graphs;  // (launches, first ring phase) -> one graph for a whole eigsol_b200_power_step batch
    int graph_len = 0;

    int ss() const { return dtype ? 16 : 8; }
    ~MatrixImpl() {
        if (stream) cudaStreamSynchronize(stream);  // pooled buffers are handed back after their last use
        for (void* q : ipc_opened) cudaIpcCloseMemHandle(q);
        if (graph) cudaGraphExecDestroy(graph);
        if (graph_short) cudaGraphExecDestroy(graph_short);
        for (auto& kv : batch_graphs) cudaGraphExecDestroy(kv.second);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (stream) cudaStreamDestroy(stream);
    }

    void init_common(int dtype_, Comm* comm_);
    void init_sparse_local(CsrHost&& local_global_cols);
    void init_sparse_device_csc(int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner, const void* values);
    void ensure_host_csr();
    void compute_prescale();
    void init_sparse_device_rows(const int32_t* rowptr, const int32_t* colidx, const void* values);
    void finish_rows_on_device();
    double near_diag_on_device(int64_t row0);
    void init_sparse_group_csc(int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner, const void* values);
    void init_dense(int64_t rows, const void* a, int64_t lda_host);
    void check_dense_symmetry();
    void plan_exchange();
    void setup_peer();
    template <class S> IterArgs<S> iter_args() {
        IterArgs<S> a;
        for (int b = 0; b < kRing; ++b) a.buf[b] = buf[b].as<S>();
        a.st = d_state.as<PowerState>();
        a.partials = d_partials.as<double>();
        a.nslots = nslots;
        a.mode = xmode;
        a.peer = xmode == MODE_PEER ? d_peer.as<PeerArgs>() : nullptr;
        a.scal_glob = d_scal.as<double>();
        a.deferred = (xmode == MODE_PEER && have_deferred) ? d_deferred.as<uint8_t>() : nullptr;
        a.serpentine = serpentine;
        a.unit_start = (xmode == MODE_PEER && have_deferred) ? unit_start : 0;
        a.halo_base = xmode == MODE_PEER ? int(nloc_pad) : INT32_MAX;
        return a;
    }
    void alloc_vectors();
    void set_kernel(int kernel_, int lanes_);
    void configure_launch();
    void drop_graph();

    void begin(const void* x0, int max_it, double tol, bool unit_scale = false);
    void exchange_halo(int in);
    void launch_matvec(int phase);
    void build_deferred();
    void issue_trips(int k);
    void ensure_graphs();
    cudaGraphExec_t batch_graph(int len, int phase);
    void rank_barrier();
    void step(int trips, float* ms_out);
    void end(void* x_out, double lambda_out[2], int* iters, int* conv);
    const PowerState& poll_state();
    void gather_download(const void* d_local, void* host_out);
    void run(const void* x0, int max_it, double tol, void* x_out, double lambda_out[2], int* iters, int* conv);
    void matvec(const void* x, void* y);
    void rayleigh(const void* x, double out[4]);
    void fill_info(eigsol_b200_matrix_info* info) const;
};

void MatrixImpl::init_common(int dtype_, Comm* comm_) {
    require_device();
    dtype = dtype_;
    comm = comm_;
    if (comm) {
        rank = comm->rank;
        nranks = comm->nranks;
        device = comm->device;
        EIGSOL_CUDA(cudaSetDevice(device));
    } else {
        EIGSOL_CUDA(cudaGetDevice(&device));
    }
    EIGSOL_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device));  // (cudaGetDeviceProperties costs ms)
    {
        // the device's stream-ordered pool keeps what an ingest frees (up to EIGSOL_B200_POOL_KEEP_MB, default 16 GB of
        // the 180) so that the next upload of a matrix does not pay the driver for its buffers again
        static std::atomic<uint64_t> done_mask{0};  // once per device and process
        const uint64_t bit = uint64_t(1) << (device & 63);
        if (!(done_mask.fetch_or(bit) & bit)) {
            cudaMemPool_t pool = nullptr;
            if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess && pool) {
                uint64_t keep = uint64_t(std::max(0, env_int("EIGSOL_B200_POOL_KEEP_MB", 16384))) << 20;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            cudaGetLastError();
        }
    }
    EIGSOL_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    g_alloc_stream = stream;  // small buffers of this handle: from the pool, on its stream (the API call's AllocScope restores)
    EIGSOL_CUDA(cudaEventCreate(&ev0));
    EIGSOL_CUDA(cudaEventCreate(&ev1));
    d_state.alloc(sizeof(PowerState));
    if (comm && comm->nccl_comm) d_scal.alloc_plain(kRing * 4 * sizeof(double));  // NCCL buffer (split mode, begin barrier)
    else d_scal.alloc(kRing * 4 * sizeof(double));
    serpentine = env_int("EIGSOL_B200_SERPENTINE", 1) ? 1 : 0;
    pdl = env_int("EIGSOL_B200_PDL", 1) != 0;
    scalar_block = env_int("EIGSOL_B200_SCALAR_BLOCK", 256) == 128 ? 128 : 256;
    gemv_var = std::max(0, std::min(3, env_int("EIGSOL_B200_GEMV_VAR", 3)));
    gemv_map = env_int("EIGSOL_B200_GEMV_MAP", 0) ? 1 : 0;
    h_state.alloc(sizeof(PowerState));
}

void MatrixImpl::alloc_vectors() {
    nloc_pad = (nloc + 15) / 16 * 16;
    ext_len = nloc_pad + nhalo;
    // one GPU: the vectors are private -> pool memory (no driver round trip on reuse); a group shares them with its peers
    // (CUDA IPC / peer access), which needs plain allocations
    for (auto& b : buf) {
        if (b.p && b.bytes == size_t(ext_len) * ss()) continue;  // buf[0] may be in place already (preloaded x0)
        if (nranks == 1 || (comm && comm->pool_shared)) b.alloc_pooled(size_t(ext_len) * ss(), stream, true, 256);
        else b.alloc_plain(size_t(ext_len) * ss(), 256);  // exported with CUDA IPC
    }
    if (nranks == 1 || (comm && comm->pool_shared)) d_xloc.alloc_pooled(size_t(std::max<int64_t>(nloc, 1)) * ss(), stream);
    else d_xloc.alloc(size_t(std::max<int64_t>(nloc, 1)) * ss());
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    if (nranks > 1 && !(comm && comm->local)) d_xglob.alloc_plain(size_t(std::max<int64_t>(n, 1)) * ss());  // NCCL gather target
}

// -------------------------------------------------------------------------------------------------
// Exchange plan: tell every peer which of its entries this rank reads each iteration (setup only).
// After this, send_count[q]/d_send_idx describe what to send to q, halo.peer_off what arrives from q.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::plan_exchange() {
    send_count.assign(size_t(nranks), 0);
    send_off.assign(size_t(nranks) + 1, 0);
    send_start.assign(size_t(nranks), 0);
    send_contig.assign(size_t(nranks), 1);
    need_pack = false;
    if (nranks == 1) return;
    Comm* c = comm;
    std::vector<int32_t> idx;  // what each peer reads from me (global columns), grouped by peer
    int64_t total_send = 0;
    if (c->local) {
        // single process: every rank reads its peers' halo plans directly
        c->local->exchange(rank, &halo, [&](const std::vector<const void*>& sl) {
            for (int q = 0; q < nranks; ++q)
                send_count[size_t(q)] = q == rank ? 0 : static_cast<const HaloPlan*>(sl[size_t(q)])->peer_count[size_t(rank)];
            for (int q = 0; q < nranks; ++q) send_off[size_t(q) + 1] = send_off[size_t(q)] + send_count[size_t(q)];
            total_send = send_off[size_t(nranks)];
            idx.resize(size_t(total_send));
            for (int q = 0; q < nranks; ++q) {
                if (send_count[size_t(q)] == 0) continue;
                const HaloPlan* hq = static_cast<const HaloPlan*>(sl[size_t(q)]);
                std::copy_n(hq->cols.begin() + hq->peer_off[size_t(rank)], send_count[size_t(q)], idx.begin() + send_off[size_t(q)]);
            }
        });
        d_send_idx.alloc(sizeof(int32_t) * size_t(std::max<int64_t>(total_send, 1)));
    } else {
        int64_t* d = c->scratch.as<int64_t>();
        // counts: d[0..P) = what I need from each peer, d[P..2P) = what each peer needs from me
        EIGSOL_CUDA(cudaMemcpyAsync(d, halo.peer_count.data(), sizeof(int64_t) * nranks, cudaMemcpyHostToDevice, c->stream));
        EIGSOL_NCCL(nccl().GroupStart());
        for (int q = 0; q < nranks; ++q) {
            if (q == rank) continue;
            EIGSOL_NCCL(nccl().Send(d + q, 1, ncclInt64, q, c->nccl_comm, c->stream));
            EIGSOL_NCCL(nccl().Recv(d + nranks + q, 1, ncclInt64, q, c->nccl_comm, c->stream));
        }
        EIGSOL_NCCL(nccl().GroupEnd());
        EIGSOL_CUDA(cudaMemcpyAsync(send_count.data(), d + nranks, sizeof(int64_t) * nranks, cudaMemcpyDeviceToHost, c->stream));
        EIGSOL_CUDA(cudaStreamSynchronize(c->stream));
        send_count[rank] = 0;
        for (int q = 0; q < nranks; ++q) send_off[q + 1] = send_off[q] + send_count[q];
        total_send = send_off[nranks];

        DevBuf d_need;  // my halo columns (global), grouped by owner
        d_need.alloc_plain(sizeof(int32_t) * size_t(std::max<int64_t>(nhalo, 1)));  // NCCL buffer
        d_need.upload(halo.cols.data(), sizeof(int32_t) * size_t(nhalo), c->stream);
        d_send_idx.alloc_plain(sizeof(int32_t) * size_t(std::max<int64_t>(total_send, 1)));  // NCCL buffer
        EIGSOL_NCCL(nccl().GroupStart());
        for (int q = 0; q < nranks; ++q) {
            if (q == rank) continue;
            if (halo.peer_count[q] > 0)
                EIGSOL_NCCL(nccl().Send(d_need.as<int32_t>() + halo.peer_off[q], size_t(halo.peer_count[q]), ncclInt32, q,
                                        c->nccl_comm, c->stream));
            if (send_count[q] > 0)
                EIGSOL_NCCL(nccl().Recv(d_send_idx.as<int32_t>() + send_off[q], size_t(send_count[q]), ncclInt32, q,
                                        c->nccl_comm, c->stream));
        }
        EIGSOL_NCCL(nccl().GroupEnd());
        idx.assign(size_t(total_send), 0);
        if (total_send)
            EIGSOL_CUDA(cudaMemcpyAsync(idx.data(), d_send_idx.p, sizeof(int32_t) * size_t(total_send), cudaMemcpyDeviceToHost,
                                        c->stream));
        EIGSOL_CUDA(cudaStreamSynchronize(c->stream));
    }
    // global -> local numbering; a contiguous ascending run is sent straight from the vector
    for (int q = 0; q < nranks; ++q) {
        bool contig = true;
        for (int64_t k = send_off[q]; k < send_off[q + 1]; ++k) {
            if (idx[k] < lo || idx[k] >= hi) throw PlanError("peer requested a column this rank does not own");
            idx[k] = int32_t(idx[k] - lo);
            if (k > send_off[q] && idx[k] != idx[k - 1] + 1) contig = false;
        }
        send_contig[q] = contig ? 1 : 0;
        send_start[q] = send_count[q] ? idx[send_off[q]] : 0;
        if (!contig) need_pack = true;
    }
    if (total_send) {
        EIGSOL_CUDA(cudaMemcpy(d_send_idx.p, idx.data(), sizeof(int32_t) * size_t(total_send), cudaMemcpyHostToDevice));
        d_sendbuf.alloc_plain(size_t(total_send) * ss());  // NCCL buffer (split mode)
    }
}

// -------------------------------------------------------------------------------------------------
// MODE_PEER wiring: every rank exports its two vector buffers, its scal_in and its flag_in through
// CUDA IPC; peers map them and from then on the matvec kernel itself stores boundary rows, partial sums
// and completion flags straight into peer memory over NVLink (kernels.cuh, PeerArgs).  Falls back to
// MODE_SPLIT (NCCL send/recv + all-reduce) when IPC mapping is not possible on some rank.
// -------------------------------------------------------------------------------------------------
struct PeerBlob {
    cudaIpcMemHandle_t buf[kRing], scal, flag;  // process-per-GPU group: CUDA IPC handles
    void *raw_buf[kRing], *raw_scal, *raw_flag; // single-process group: plain device pointers (peer access)
    int device;
    int64_t halo_base;                 // nloc_pad: where the halo segment starts in the vector buffers
    int64_t peer_off[kMaxPeers + 1];   // offset of each owner's entries inside the halo segment
};

void MatrixImpl::setup_peer() {
    xmode = nranks == 1 ? MODE_FUSED : MODE_SPLIT;
    if (nranks == 1) return;
    const char* want = std::getenv("EIGSOL_B200_EXCHANGE");
    const bool local = comm->local != nullptr;
    const bool force_nccl = !local && want && std::string(want) == "nccl";
    // mailboxes and flags are written by the peers: plain memory (CUDA IPC export / peer access), unless the group
    // shares this device's pool
    if (local && comm->pool_shared) {
        d_scal_in.alloc(sizeof(double) * kRing * kMaxPeers * 4);
        d_flag_in.alloc(sizeof(unsigned long long) * 2 * kMaxPeers);
    } else {
        d_scal_in.alloc_plain(sizeof(double) * kRing * kMaxPeers * 4);
        d_flag_in.alloc_plain(sizeof(unsigned long long) * 2 * kMaxPeers);  // [0,8): trip flags, [8,16): gate flags
    }
    PeerBlob mine;
    std::memset(&mine, 0, sizeof mine);
    int ok = (nranks <= kMaxPeers && !force_nccl) ? 1 : 0;
    if (ok && !local) {
        for (int b = 0; b < kRing; ++b) ok &= cudaIpcGetMemHandle(&mine.buf[b], buf[b].p) == cudaSuccess;
        ok &= cudaIpcGetMemHandle(&mine.scal, d_scal_in.p) == cudaSuccess;
        ok &= cudaIpcGetMemHandle(&mine.flag, d_flag_in.p) == cudaSuccess;
        cudaGetLastError();
    }
    for (int b = 0; b < kRing; ++b) mine.raw_buf[b] = buf[b].p;
    mine.raw_scal = d_scal_in.p;
    mine.raw_flag = d_flag_in.p;
    mine.device = device;
    mine.halo_base = nloc_pad;
    for (int q = 0; q <= nranks && q <= kMaxPeers; ++q) mine.peer_off[q] = halo.peer_off[q];
    ok = comm_min_int(comm, ok);
    if (!ok) {
        if (local) throw DeviceError("powerMethod: device error: a single-process GPU group holds at most 8 devices");
        return;
    }
    const std::vector<char> all = comm_allgather_bytes(comm, &mine, sizeof mine);
    PeerArgs pa;
    std::memset(&pa, 0, sizeof pa);
    pa.nranks = nranks;
    pa.rank = rank;
    pa.scal_in = d_scal_in.as<double>();
    pa.flag_in = d_flag_in.as<unsigned long long>();
    pa.gate_in = d_flag_in.as<unsigned long long>() + kMaxPeers;
    pa.send_idx = d_send_idx.as<int32_t>();
    pa.spin_limit = (long long)env_int("EIGSOL_B200_PEER_TIMEOUT_MS", 4000) * 2000000LL;  // ~2 GHz ticks
    const size_t es = size_t(ss());
    int mapped = 1;
    for (int q = 0; q < nranks && mapped; ++q) {
        if (q == rank) continue;
        PeerBlob pb;
        std::memcpy(&pb, all.data() + sizeof(PeerBlob) * q, sizeof pb);
        void* rb[kRing] = {nullptr, nullptr, nullptr};
        void *rs = nullptr, *rf = nullptr;
        if (local) {
            // same process: enable peer access once per device pair and use the peer's pointers as they are
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, device, pb.device) != cudaSuccess || !can) {
                cudaGetLastError();
                mapped = 0;
                break;
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(pb.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                mapped = 0;
                break;
            }
            cudaGetLastError();
            for (int b = 0; b < kRing; ++b) rb[b] = pb.raw_buf[b];
            rs = pb.raw_scal;
            rf = pb.raw_flag;
        } else {
            auto open = [&](void** dst, const cudaIpcMemHandle_t& h) {
                if (cudaIpcOpenMemHandle(dst, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    mapped = 0;
                    return;
                }
                ipc_opened.push_back(*dst);
            };
            for (int b = 0; b < kRing && mapped; ++b) open(&rb[b], pb.buf[b]);
            if (mapped) open(&rs, pb.scal);
            if (mapped) open(&rf, pb.flag);
            if (!mapped) break;
        }
        pa.peer_scal[q] = static_cast<double*>(rs);
        pa.peer_flag[q] = static_cast<unsigned long long*>(rf) + rank;
        pa.peer_gate[q] = static_cast<unsigned long long*>(rf) + kMaxPeers + rank;
        if (send_count[q] == 0) continue;
        const size_t dst_off = size_t(pb.halo_base + pb.peer_off[rank]) * es;
        if (send_contig[q]) {
            const int k = pa.nsend++;
            pa.send_start[k] = int(send_start[q]);
            pa.send_count[k] = int(send_count[q]);
            for (int b = 0; b < kRing; ++b) pa.send_dst[k][b] = static_cast<char*>(rb[b]) + dst_off;
        } else {
            const int k = pa.nlist++;
            pa.list_off[k] = int(send_off[q]);
            pa.list_count[k] = int(send_count[q]);
            for (int b = 0; b < kRing; ++b) pa.list_dst[k][b] = static_cast<char*>(rb[b]) + dst_off;
        }
    }
    mapped = comm_min_int(comm, mapped);
    if (!mapped) {
        for (void* q : ipc_opened) cudaIpcCloseMemHandle(q);
        ipc_opened.clear();
        if (local)
            throw DeviceError("powerMethod: device error: the GPUs of the group cannot access each other's memory "
                              "(peer access); run with EIGSOL_B200_NGPU=1");
        return;
    }
    pa.debug_nowait = env_int("EIGSOL_B200_DEBUG_NOWAIT", 0);
    pa.halo_src_mask = 0;
    for (int q = 0; q < nranks; ++q)
        if (q != rank && halo.peer_count[size_t(q)] > 0) pa.halo_src_mask |= 1u << q;
    if (env_int("EIGSOL_B200_DEBUG_NOSEND", 0)) pa.nsend = pa.nlist = 0;  // timing experiments only: results are wrong
    if (env_int("EIGSOL_B200_PEER_STATS", 0)) {
        d_stats.alloc(sizeof(unsigned long long) * kPeerStats);
        pa.stats = d_stats.as<unsigned long long>();
    }
    peer_h = pa;
    d_peer.alloc(sizeof(PeerArgs));
    EIGSOL_CUDA(cudaMemcpy(d_peer.p, &pa, sizeof pa, cudaMemcpyHostToDevice));
    xmode = MODE_PEER;
}

// -------------------------------------------------------------------------------------------------
// Sparse: local rows in CSR with GLOBAL columns -> halo plan, localized columns, layouts, vectors.
// row_begin must be set.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::init_sparse_local(CsrHost&& a) {
    dense = false;
    lo = row_begin[rank];
    hi = row_begin[rank + 1];
    nloc = hi - lo;
    if (a.rows != nloc) throw PlanError("row block size does not match the partition");
    Trace tr;
    csr = std::move(a);
    local_nnz = csr.nnz();
    const double near_diag = row_stats(nloc, csr.rowptr.data(), csr.colidx.data(), lo).near_diag;  // global columns still
    halo = halo_plan(nloc, csr.rowptr.data(), csr.colidx.data(), row_begin, rank);
    nhalo = int64_t(halo.cols.size());
    tr.mark("rows: halo plan");
    alloc_vectors();
    if (ext_len > INT32_MAX) throw PlanError("local vector length exceeds int32 indices");
    localize_columns(local_nnz, csr.colidx.data(), lo, hi, nloc_pad, halo.cols);
    csr.cols = ext_len;
    tr.mark("rows: vectors + localize");
    plan_exchange();
    setup_peer();
    tr.mark("rows: exchange plan + peer map");
    stats = row_stats(nloc, csr.rowptr.data());
    stats.near_diag = near_diag;
    tr.mark("rows: row stats");
    if (nranks > 1) {
        const std::vector<int64_t> all = comm_allgather_i64(comm, local_nnz);
        nnz_global = 0;
        for (int64_t v : all) nnz_global += v;
    } else {
        nnz_global = local_nnz;
    }
    // CSR arrays are always resident (scalar / vector / stream variants share them)
    d_rowptr.alloc(sizeof(int32_t) * size_t(nloc + 1));
    d_rowptr.upload(csr.rowptr.data(), sizeof(int32_t) * size_t(nloc + 1), stream);
    d_colidx.alloc(sizeof(int32_t) * size_t(std::max<int64_t>(local_nnz, 1)));
    d_colidx.upload(csr.colidx.data(), sizeof(int32_t) * size_t(local_nnz), stream);
    d_vals.alloc(size_t(std::max<int64_t>(local_nnz, 1)) * ss());
    d_vals.upload(csr.vals.data(), size_t(local_nnz) * ss(), stream);
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    tr.mark("rows: upload");
    const int forced = env_int("EIGSOL_B200_KERNEL", 0);
    const KernelChoice ch = choose_kernel(dtype, stats);
    if ((forced > 0 && forced <= K_SELL) || forced == K_CSR_WINDOW) set_kernel(forced, env_int("EIGSOL_B200_LANES", 0));
    else set_kernel(ch.kernel, ch.lanes);
    compute_prescale();
}

// -------------------------------------------------------------------------------------------------
// Single GPU, CSC input: staged multi-threaded upload + transpose on the device (ingest.cuh).
// The host keeps no CSR copy; ensure_host_csr() fetches one if a derived layout has to be built.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::init_sparse_device_csc(int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                                        const void* values) {
    Trace tr;
    dense = false;
    if (rows > INT32_MAX || cols > INT32_MAX) throw PlanError("matrix dimension exceeds int32 indices");
    if (!outer) throw PlanError("null outer index array");
    if (outer[0] != 0) throw PlanError("CSC outer index must start at 0 (compressed storage expected)");
    // outer[cols] sizes the allocations; the monotonicity of outer and the range of inner are checked on the device
    const int64_t nnz = outer[cols];
    if (nnz < 0) throw PlanError("CSC outer index not monotone");
    if (nnz > 0 && (!inner || !values)) throw PlanError("null inner index array");
    row_begin.assign({0, rows});
    lo = 0;
    hi = nloc = rows;
    local_nnz = nnz_global = nnz;
    halo = HaloPlan();
    halo.peer_count.assign(1, 0);
    halo.peer_off.assign(2, 0);
    nhalo = 0;
    const size_t s = size_t(ss());
    const size_t nz = size_t(std::max<int64_t>(nnz, 1));
    stats = RowStats();
    stats.rows = rows;
    stats.nnz = nnz;
    csr = CsrHost();
    csr.dtype = dtype;
    csr.rows = rows;
    if (nnz > 0) {
        // The PCIe copy of the three arrays is what bounds this call, so it starts first and everything else hides
        // behind it: the uploader threads push outer, inner, values (in that order) while this thread allocates the
        // rest; as soon as outer + inner have landed the expansion, the sort by row and the row-pointer scan run on
        // the device UNDER the values upload; only the final gather waits for the values.
        DevBuf d_outer, d_inner, d_cvals, d_col, d_iota, d_perm, d_keys, d_flags, d_tmp, d_tmp2;
        d_outer.alloc_pooled(sizeof(int32_t) * size_t(cols + 1), stream);
        d_inner.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_cvals.alloc_pooled(nz * s, stream);
        EIGSOL_CUDA(cudaStreamSynchronize(stream));  // the uploader threads write from their own streams
        StagedUploader up;
        up.add(d_outer.p, outer, sizeof(int32_t) * size_t(cols + 1));
        const size_t index_jobs = up.add(d_inner.p, inner, sizeof(int32_t) * size_t(nnz));
        up.add(d_cvals.p, values, size_t(nnz) * s);
        if (preload_x0) {
            // one-shot call: the start vector rides on the same upload (begin() then finds it in place)
            buf[0].alloc_pooled(size_t((rows + 15) / 16 * 16) * s, stream, true, 256);
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
            up.add(buf[0].p, preload_x0, size_t(rows) * s);
        }
        up.start(device, upload_threads(nranks), index_jobs);
        tr.mark_async("ingest: allocs + upload started");
        alloc_vectors();
        if (ext_len > INT32_MAX) throw PlanError("local vector length exceeds int32 indices");
        plan_exchange();
        setup_peer();
        csr.cols = ext_len;
        d_rowptr.alloc_pooled(sizeof(int32_t) * size_t(rows + 1), stream, true);  // zero-filled: the row histogram
        d_colidx.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_vals.alloc_pooled(nz * s, stream);
        d_col.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_iota.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_perm.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_keys.alloc_pooled(sizeof(int32_t) * nz, stream);
        d_flags.alloc(64);
        int end_bit = 1;
        while (end_bit < 31 && (int64_t(1) << end_bit) < rows) ++end_bit;
        size_t tb = 0, tb2 = 0;
        EIGSOL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tb, d_inner.as<int32_t>(), d_keys.as<int32_t>(),
                                                    d_iota.as<int32_t>(), d_perm.as<int32_t>(), int(nnz), 0, end_bit, stream));
        d_tmp.alloc_pooled(tb, stream);
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb2, d_rowptr.as<int32_t>(), d_rowptr.as<int32_t>(), int(rows + 1), stream));
        d_tmp2.alloc_pooled(tb2, stream);
        tr.mark_async("ingest: remaining allocs");
        up.wait_prefix();
        tr.mark_async("ingest: outer + inner on the device");
        const int B = 256;
        const unsigned gnz = unsigned((nnz + B - 1) / B);
        outer_check_kernel<<<unsigned((cols + B - 1) / B), B, 0, stream>>>(d_outer.as<int32_t>(), cols, nnz, d_flags.as<int>());
        csc_expand_kernel<<<gnz, B, 0, stream>>>(d_outer.as<int32_t>(), d_inner.as<int32_t>(), cols, rows, nnz,
                                                  d_col.as<int32_t>(), d_iota.as<int32_t>(), d_rowptr.as<int32_t>(),
                                                  d_flags.as<int>());
        EIGSOL_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp.p, tb, d_inner.as<int32_t>(), d_keys.as<int32_t>(),
                                                    d_iota.as<int32_t>(), d_perm.as<int32_t>(), int(nnz), 0, end_bit, stream));
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp2.p, tb2, d_rowptr.as<int32_t>(), d_rowptr.as<int32_t>(), int(rows + 1), stream));
        // row-length statistics for the kernel choice (needs only rowptr)
        int* f = d_flags.as<int>();  // [0] error flags, [1] max, [2] min, [4..5] padded (u64), [6..7] var (double)
        const int init_min = INT32_MAX;
        EIGSOL_CUDA(cudaMemcpyAsync(f + 2, &init_min, sizeof(int), cudaMemcpyHostToDevice, stream));
        stats.mean = double(nnz) / double(rows);
        const unsigned gsl = unsigned(((rows + 63) / 64 + B - 1) / B);
        row_stats_kernel<<<gsl, B, 0, stream>>>(d_rowptr.as<int32_t>(), rows, stats.mean, f + 1, f + 2,
                                                reinterpret_cast<unsigned long long*>(f + 4), reinterpret_cast<double*>(f + 6));
        EIGSOL_CUDA(cudaGetLastError());
        EIGSOL_CUDA(up.finish());
        x0_preloaded = preload_x0 != nullptr;
        tr.mark_async("ingest: values on the device");
        if (dtype)
            csc_gather_kernel<double2><<<gnz, B, 0, stream>>>(d_perm.as<int32_t>(), d_col.as<int32_t>(), d_cvals.as<double2>(), nnz,
                                                              d_colidx.as<int32_t>(), d_vals.as<double2>());
        else
            csc_gather_kernel<double><<<gnz, B, 0, stream>>>(d_perm.as<int32_t>(), d_col.as<int32_t>(), d_cvals.as<double>(), nnz,
                                                             d_colidx.as<int32_t>(), d_vals.as<double>());
        launches += 6;
        EIGSOL_CUDA(cudaGetLastError());
        int hf[8];
        EIGSOL_CUDA(cudaMemcpyAsync(hf, f, sizeof hf, cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        tr.mark("ingest: gather + stats");
        if (hf[0] & 1) throw PlanError("CSC outer index not monotone");
        if (hf[0] & 2) throw PlanError("CSC inner index out of range");
        unsigned long long padded;
        double var;
        std::memcpy(&padded, hf + 4, sizeof padded);
        std::memcpy(&var, hf + 6, sizeof var);
        stats.max = hf[1];
        stats.min = hf[2];
        stats.cv = stats.mean > 0 ? std::sqrt(var / double(rows)) / stats.mean : 0.0;
        stats.sell_fill = double(padded) / double(nnz);
        stats.near_diag = near_diag_on_device(0);
    } else {
        alloc_vectors();
        plan_exchange();
        setup_peer();
        csr.cols = ext_len;
        d_rowptr.alloc(sizeof(int32_t) * size_t(rows + 1));
        d_colidx.alloc(sizeof(int32_t) * nz);
        d_vals.alloc(nz * s);
    }
    const int forced = env_int("EIGSOL_B200_KERNEL", 0);
    if ((forced > 0 && forced <= K_SELL) || forced == K_CSR_WINDOW) set_kernel(forced, env_int("EIGSOL_B200_LANES", 0));
    else set_kernel(K_AUTO, 0);
    compute_prescale();
    tr.mark("ingest: kernel choice + prescale");
}

// -------------------------------------------------------------------------------------------------
// Single-process group, CSC input (what EigSol::Matrix holds): the P GPUs split the UPLOAD by column slabs
// -- contiguous pieces of the caller's three arrays, 1/P of the bytes over each GPU's own PCIe link -- and the
// ROWS by the same boundaries (row_begin == slab boundaries: nnz-balanced for a structurally symmetric matrix,
// correct for any).  Every rank then pulls the entries of its rows out of every slab with peer loads over NVLink
// (slab order = ascending columns), and transposes its COO piece to CSR with global columns on its own GPU;
// finish_rows_on_device() does the rest.  row_begin must be set.
// -------------------------------------------------------------------------------------------------
struct SlabShow {
    const int32_t *sinner, *scol;
    const void* svals;
    int64_t snnz;
};

void MatrixImpl::init_sparse_group_csc(int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                                       const void* values) {
    Trace tr;
    dense = false;
    lo = row_begin[size_t(rank)];
    hi = row_begin[size_t(rank) + 1];
    nloc = hi - lo;
    const int64_t c0 = lo, c1 = hi;  // my column slab
    const int64_t e0 = outer[c0], e1 = outer[c1];
    if (e0 < 0 || e1 < e0 || e1 > outer[cols]) throw PlanError("CSC outer index not monotone");
    const int64_t snnz = e1 - e0, scols = c1 - c0;
    const size_t s = size_t(ss());
    const int B = 256;
    DevBuf d_souter, d_sinner, d_svals, d_scol, d_flags;
    const bool pooled = comm->pool_shared;  // peers read the slabs: pool memory only if the pool is shared with them
    auto get = [&](DevBuf& b, size_t bytes, bool zero) {
        if (pooled) b.alloc_pooled(bytes, stream, zero);
        else if (zero) b.alloc(bytes);
        else b.alloc_raw(bytes);
    };
    d_flags.alloc(64);
    get(d_souter, sizeof(int32_t) * size_t(scols + 1), false);
    get(d_sinner, sizeof(int32_t) * size_t(std::max<int64_t>(snnz, 1)), false);
    get(d_scol, sizeof(int32_t) * size_t(std::max<int64_t>(snnz, 1)), false);
    get(d_svals, size_t(std::max<int64_t>(snnz, 1)) * s, false);
    EIGSOL_CUDA(cudaStreamSynchronize(stream));  // the uploader threads write from their own streams
    {
        StagedUploader up;
        up.add(d_souter.p, outer + c0, sizeof(int32_t) * size_t(scols + 1));
        up.add(d_sinner.p, inner + e0, sizeof(int32_t) * size_t(snnz));
        up.add(d_svals.p, static_cast<const char*>(values) + size_t(e0) * s, size_t(snnz) * s);
        if (preload_x0 && nloc > 0) {
            // one-shot call: this rank's rows of the start vector ride on the slab upload
            get(d_x0pre, size_t(nloc) * s, false);
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
            up.add(d_x0pre.p, static_cast<const char*>(preload_x0) + size_t(lo) * s, size_t(nloc) * s);
        }
        tr.mark_async("group csc: slab buffers");
        EIGSOL_CUDA(up.run(device, upload_threads(nranks, true)));
        x0_preloaded = preload_x0 != nullptr;
    }
    tr.mark("group csc: slab upload");
    if (e0 != 0) rebase_kernel<<<unsigned((scols + 1 + B - 1) / B), B, 0, stream>>>(d_souter.as<int32_t>(), scols + 1, int32_t(e0));
    if (scols > 0) outer_check_kernel<<<unsigned((scols + B - 1) / B), B, 0, stream>>>(d_souter.as<int32_t>(), scols, snnz, d_flags.as<int>());
    if (snnz > 0)
        slab_expand_kernel<<<unsigned((snnz + B - 1) / B), B, 0, stream>>>(d_souter.as<int32_t>(), d_sinner.as<int32_t>(), scols, c0,
                                                                           rows, snnz, d_scol.as<int32_t>(), d_flags.as<int>());
    launches += 2;
    int hflags = 0;
    EIGSOL_CUDA(cudaMemcpyAsync(&hflags, d_flags.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    int rows_ok = (hflags & 2) ? 0 : ((hflags & 1) ? 1 : 2);  // 0: bad row index, 1: outer not monotone, 2: fine
    rows_ok = comm_min_int(comm, rows_ok);  // every rank takes the same exit
    if (rows_ok == 0) throw PlanError("CSC inner index out of range");
    if (rows_ok == 1) throw PlanError("CSC outer index not monotone");
    // peer access to every slab
    for (int q = 0; q < nranks; ++q) {
        if (q == rank) continue;
        int can = 0;
        const int qdev = comm->local_devices[size_t(q)];
        if (cudaDeviceCanAccessPeer(&can, device, qdev) != cudaSuccess || !can)
            throw DeviceError("powerMethod: device error: the GPUs of the group cannot access each other's memory (peer access)");
        const cudaError_t e = cudaDeviceEnablePeerAccess(qdev, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) EIGSOL_CUDA(e);
        cudaGetLastError();
    }
    SlabShow mine{d_sinner.as<int32_t>(), d_scol.as<int32_t>(), d_svals.p, snnz};
    const std::vector<char> all = comm_allgather_bytes(comm, &mine, sizeof mine);  // also the barrier: all slabs are resident
    std::vector<SlabShow> slab{size_t(nranks), SlabShow{}};
    std::memcpy(slab.data(), all.data(), all.size());
    // 1. how many entries of each slab belong to my rows
    DevBuf d_counts;
    d_counts.alloc(sizeof(unsigned long long) * size_t(nranks));
    int64_t max_snnz = 1;
    for (int q = 0; q < nranks; ++q) {
        max_snnz = std::max(max_snnz, slab[size_t(q)].snnz);
        if (slab[size_t(q)].snnz == 0) continue;
        const int g = int(std::min<int64_t>((slab[size_t(q)].snnz + B - 1) / B, int64_t(sm_count) * 8));
        slab_count_kernel<<<g, B, 0, stream>>>(slab[size_t(q)].sinner, slab[size_t(q)].snnz, int32_t(lo), int32_t(hi),
                                               d_counts.as<unsigned long long>() + q);
        ++launches;
    }
    std::vector<unsigned long long> cnt(size_t(nranks), 0);
    EIGSOL_CUDA(cudaMemcpyAsync(cnt.data(), d_counts.p, sizeof(unsigned long long) * size_t(nranks), cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    int64_t nnz = 0;
    for (auto v : cnt) nnz += int64_t(v);
    if (nnz > INT32_MAX) throw PlanError("a rank's row block holds more than 2^31-1 entries");
    local_nnz = nnz;
    tr.mark("group csc: expand + count");
    // 2. pull them, slab by slab, into COO arrays (key = local row, global column, value)
    const size_t nz = size_t(std::max<int64_t>(nnz, 1));
    DevBuf d_key, d_col, d_val, d_iota, d_sel, d_nsel, d_tmp;
    get(d_key, sizeof(int32_t) * nz, false);
    get(d_col, sizeof(int32_t) * nz, false);
    get(d_val, nz * s, false);
    get(d_iota, sizeof(int32_t) * nz, false);
    get(d_sel, sizeof(int32_t) * size_t(max_snnz), false);
    d_nsel.alloc(sizeof(int));
    get(d_rowptr, sizeof(int32_t) * size_t(nloc + 1), true);  // zero-filled: the row histogram
    size_t tb = 0;
    {
        cub::CountingInputIterator<int32_t> it0(0);
        EIGSOL_CUDA(cub::DeviceSelect::If(nullptr, tb, it0, d_sel.as<int32_t>(), d_nsel.as<int>(), int(max_snnz),
                                          RowInRange{nullptr, 0, 0}, stream));
    }
    get(d_tmp, tb, false);
    int64_t off = 0;
    for (int q = 0; q < nranks; ++q) {
        const int64_t nq = int64_t(cnt[size_t(q)]);
        if (nq == 0) continue;
        const SlabShow& sq = slab[size_t(q)];
        cub::CountingInputIterator<int32_t> it0(0);
        size_t tbq = tb;
        EIGSOL_CUDA(cub::DeviceSelect::If(d_tmp.p, tbq, it0, d_sel.as<int32_t>(), d_nsel.as<int>(), int(sq.snnz),
                                          RowInRange{sq.sinner, int32_t(lo), int32_t(hi)}, stream));
        const unsigned g = unsigned((nq + B - 1) / B);
        if (dtype)
            slab_pull_kernel<double2><<<g, B, 0, stream>>>(d_sel.as<int32_t>(), nq, sq.sinner, sq.scol,
                                                           static_cast<const double2*>(sq.svals), int32_t(lo),
                                                           d_key.as<int32_t>() + off, d_col.as<int32_t>() + off,
                                                           d_val.as<double2>() + off, d_iota.as<int32_t>() + off, int32_t(off),
                                                           d_rowptr.as<int32_t>());
        else
            slab_pull_kernel<double><<<g, B, 0, stream>>>(d_sel.as<int32_t>(), nq, sq.sinner, sq.scol,
                                                          static_cast<const double*>(sq.svals), int32_t(lo),
                                                          d_key.as<int32_t>() + off, d_col.as<int32_t>() + off,
                                                          d_val.as<double>() + off, d_iota.as<int32_t>() + off, int32_t(off),
                                                          d_rowptr.as<int32_t>());
        launches += 2;
        off += nq;
    }
    EIGSOL_CUDA(cudaGetLastError());
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    comm->local->barrier();  // every rank has pulled: the slabs may go
    d_souter.release();
    d_sinner.release();
    d_scol.release();
    d_svals.release();
    d_sel.release();
    tr.mark("group csc: pull");
    // 3. COO (column-major order) -> CSR: stable sort by local row keeps the columns ascending inside a row
    get(d_colidx, sizeof(int32_t) * nz, false);
    get(d_vals, nz * s, false);
    if (nnz > 0) {
        DevBuf d_perm, d_keys2, d_t1, d_t2;
        get(d_perm, sizeof(int32_t) * nz, false);
        get(d_keys2, sizeof(int32_t) * nz, false);
        int end_bit = 1;
        while (end_bit < 31 && (int64_t(1) << end_bit) < std::max<int64_t>(nloc, 2)) ++end_bit;
        size_t t1 = 0, t2 = 0;
        EIGSOL_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, t1, d_key.as<int32_t>(), d_keys2.as<int32_t>(), d_iota.as<int32_t>(),
                                                    d_perm.as<int32_t>(), int(nnz), 0, end_bit, stream));
        get(d_t1, t1, false);
        EIGSOL_CUDA(cub::DeviceRadixSort::SortPairs(d_t1.p, t1, d_key.as<int32_t>(), d_keys2.as<int32_t>(), d_iota.as<int32_t>(),
                                                    d_perm.as<int32_t>(), int(nnz), 0, end_bit, stream));
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, t2, d_rowptr.as<int32_t>(), d_rowptr.as<int32_t>(), int(nloc + 1), stream));
        get(d_t2, t2, false);
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(d_t2.p, t2, d_rowptr.as<int32_t>(), d_rowptr.as<int32_t>(), int(nloc + 1), stream));
        const unsigned gnz = unsigned((nnz + B - 1) / B);
        if (dtype)
            csc_gather_kernel<double2><<<gnz, B, 0, stream>>>(d_perm.as<int32_t>(), d_col.as<int32_t>(), d_val.as<double2>(), nnz,
                                                              d_colidx.as<int32_t>(), d_vals.as<double2>());
        else
            csc_gather_kernel<double><<<gnz, B, 0, stream>>>(d_perm.as<int32_t>(), d_col.as<int32_t>(), d_val.as<double>(), nnz,
                                                             d_colidx.as<int32_t>(), d_vals.as<double>());
        launches += 3;
        EIGSOL_CUDA(cudaGetLastError());
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
    }
    tr.mark("group csc: transpose");
    finish_rows_on_device();
    tr.mark_async("group csc: finish returned");
}

// -------------------------------------------------------------------------------------------------
// Row block [lo, hi) in CSR with GLOBAL columns (each rank of a multi-GPU run, or the whole matrix on one
// GPU): staged upload straight from the caller's arrays, then halo discovery, column renumbering and row
// statistics on the device (ingest.cuh).  row_begin must be set.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::init_sparse_device_rows(const int32_t* rowptr, const int32_t* colidx, const void* values) {
    Trace tr;
    dense = false;
    lo = row_begin[rank];
    hi = row_begin[rank + 1];
    nloc = hi - lo;
    if (n > INT32_MAX) throw PlanError("matrix dimension exceeds int32 indices");
    if (!rowptr) throw PlanError("null rowptr");
    // rowptr points at this rank's first row; a block cut out of a larger CSR array starts at base != 0
    const int32_t base = rowptr[0];
    if (base < 0) throw PlanError("CSR rowptr must start at 0");
    for (int64_t i = 0; i < nloc; ++i)
        if (rowptr[i + 1] < rowptr[i]) throw PlanError("CSR rowptr not monotone");
    const int64_t nnz = int64_t(rowptr[nloc]) - base;
    if (nnz > 0 && (!colidx || !values)) throw PlanError("null colidx");
    local_nnz = nnz;
    const size_t s = size_t(ss());
    const size_t nz = size_t(std::max<int64_t>(nnz, 1));
    d_rowptr.alloc_pooled(sizeof(int32_t) * size_t(nloc + 1), stream);
    d_colidx.alloc_pooled(sizeof(int32_t) * nz, stream, nnz == 0);
    d_vals.alloc_pooled(nz * s, stream, nnz == 0);
    EIGSOL_CUDA(cudaStreamSynchronize(stream));  // the uploader threads write from their own streams
    StagedUploader up;
    up.add(d_rowptr.p, rowptr, sizeof(int32_t) * size_t(nloc + 1));
    up.add(d_colidx.p, colidx + base, sizeof(int32_t) * size_t(nnz));
    up.add(d_vals.p, static_cast<const char*>(values) + size_t(base) * s, size_t(nnz) * s);
    EIGSOL_CUDA(up.run(device, upload_threads(nranks, comm && comm->local)));
    if (base != 0) {
        const unsigned g = unsigned((nloc + 1 + 255) / 256);
        rebase_kernel<<<g, 256, 0, stream>>>(d_rowptr.as<int32_t>(), nloc + 1, base);
        ++launches;
    }
    tr.mark("rows(dev): staged upload");
    finish_rows_on_device();
}

// Fraction of the resident CSR entries (GLOBAL columns) on the tridiagonal: the stencil statistic of the kernel choice.
double MatrixImpl::near_diag_on_device(int64_t row0) {
    if (local_nnz == 0 || nloc == 0) return 1.0;
    DevBuf d_cnt;
    d_cnt.alloc_pooled(sizeof(unsigned long long), stream, true, 0);
    const int g = int(std::max<int64_t>(1, std::min<int64_t>((nloc + 255) / 256, int64_t(sm_count) * 8)));
    near_diag_kernel<<<g, 256, 0, stream>>>(d_rowptr.as<int32_t>(), d_colidx.as<int32_t>(), nloc, row0, d_cnt.as<unsigned long long>());
    ++launches;
    unsigned long long c = 0;
    EIGSOL_CUDA(cudaMemcpyAsync(&c, d_cnt.p, sizeof c, cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    return double(c) / double(local_nnz);
}

// The rank's rows are resident as CSR with GLOBAL columns (d_rowptr / d_colidx / d_vals, local_nnz set): halo
// discovery, column renumbering, row statistics, exchange plan, kernel choice -- all on the device.
void MatrixImpl::finish_rows_on_device() {
    Trace tr;
    const int64_t nnz = local_nnz;
    const int B = 256;
    const unsigned gnz = unsigned((std::max<int64_t>(nnz, 1) + B - 1) / B);
    const unsigned gn = unsigned((std::max<int64_t>(n, 1) + B - 1) / B);
    DevBuf d_mark, d_pos, d_flags, d_tmp, d_hcols;
    d_flags.alloc(64);
    halo = HaloPlan();
    halo.peer_count.assign(size_t(nranks), 0);
    halo.peer_off.assign(size_t(nranks) + 1, 0);
    if (nnz > 0) {
        d_mark.alloc_pooled(sizeof(int32_t) * size_t(n), stream, true);
        d_pos.alloc_pooled(sizeof(int32_t) * size_t(n), stream);
        rows_mark_kernel<<<gnz, B, 0, stream>>>(d_colidx.as<int32_t>(), nnz, n, lo, hi, d_mark.as<int32_t>(), d_flags.as<int>());
        size_t tb = 0;
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, d_mark.as<int32_t>(), d_pos.as<int32_t>(), int(n), stream));
        d_tmp.alloc(tb);
        EIGSOL_CUDA(cub::DeviceScan::ExclusiveSum(d_tmp.p, tb, d_mark.as<int32_t>(), d_pos.as<int32_t>(), int(n), stream));
        int32_t last[2] = {0, 0};
        int hflags = 0;
        EIGSOL_CUDA(cudaMemcpyAsync(&last[0], d_mark.as<int32_t>() + (n - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaMemcpyAsync(&last[1], d_pos.as<int32_t>() + (n - 1), sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaMemcpyAsync(&hflags, d_flags.p, sizeof(int), cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        if (hflags & 2) throw PlanError("CSR column index out of range");
        nhalo = int64_t(last[0]) + int64_t(last[1]);
        halo.cols.resize(size_t(nhalo));
        if (nhalo > 0) {
            d_hcols.alloc(sizeof(int32_t) * size_t(nhalo));
            halo_compact_kernel<<<gn, B, 0, stream>>>(d_mark.as<int32_t>(), d_pos.as<int32_t>(), n, d_hcols.as<int32_t>());
            EIGSOL_CUDA(cudaMemcpyAsync(halo.cols.data(), d_hcols.p, sizeof(int32_t) * size_t(nhalo), cudaMemcpyDeviceToHost, stream));
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
            // owners of the (sorted) halo columns
            size_t k = 0;
            for (int q = 0; q < nranks; ++q) {
                const size_t k0 = k;
                while (k < halo.cols.size() && halo.cols[k] < row_begin[size_t(q) + 1]) ++k;
                halo.peer_count[size_t(q)] = int64_t(k - k0);
            }
            for (int q = 0; q < nranks; ++q) halo.peer_off[size_t(q) + 1] = halo.peer_off[size_t(q)] + halo.peer_count[size_t(q)];
        }
        launches += 2;
    } else {
        nhalo = 0;
    }
    tr.mark("rows(dev): halo discovery");
    const double near_diag = near_diag_on_device(lo);  // needs the GLOBAL columns: before the renumbering below
    alloc_vectors();
    if (ext_len > INT32_MAX) throw PlanError("local vector length exceeds int32 indices");
    if (nnz > 0) {
        rows_localize_kernel<<<gnz, B, 0, stream>>>(d_colidx.as<int32_t>(), nnz, lo, hi, nloc_pad, d_pos.as<int32_t>());
        ++launches;
    }
    csr = CsrHost();
    csr.dtype = dtype;
    csr.rows = nloc;
    csr.cols = ext_len;
    stats = RowStats();
    stats.rows = nloc;
    stats.nnz = nnz;
    if (nnz > 0 && nloc > 0) {
        int* f = d_flags.as<int>();  // [1] max, [2] min, [4..5] padded (u64), [6..7] var (double)
        const int init_min = INT32_MAX;
        EIGSOL_CUDA(cudaMemcpyAsync(f + 2, &init_min, sizeof(int), cudaMemcpyHostToDevice, stream));
        stats.mean = double(nnz) / double(nloc);
        const unsigned gsl = unsigned(((nloc + 63) / 64 + B - 1) / B);
        row_stats_kernel<<<gsl, B, 0, stream>>>(d_rowptr.as<int32_t>(), nloc, stats.mean, f + 1, f + 2,
                                                reinterpret_cast<unsigned long long*>(f + 4), reinterpret_cast<double*>(f + 6));
        ++launches;
        int hf[8];
        EIGSOL_CUDA(cudaMemcpyAsync(hf, f, sizeof hf, cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        unsigned long long padded;
        double var;
        std::memcpy(&padded, hf + 4, sizeof padded);
        std::memcpy(&var, hf + 6, sizeof var);
        stats.max = hf[1];
        stats.min = hf[2];
        stats.cv = stats.mean > 0 ? std::sqrt(var / double(nloc)) / stats.mean : 0.0;
        stats.sell_fill = double(padded) / double(nnz);
        stats.near_diag = near_diag;
    }
    EIGSOL_CUDA(cudaGetLastError());
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    tr.mark("rows(dev): localize + stats");
    plan_exchange();
    setup_peer();
    tr.mark("rows(dev): exchange plan + peer map");
    if (nranks > 1) {
        const std::vector<int64_t> all = comm_allgather_i64(comm, local_nnz);
        nnz_global = 0;
        for (int64_t v : all) nnz_global += v;
    } else {
        nnz_global = local_nnz;
    }
    const int forced = env_int("EIGSOL_B200_KERNEL", 0);
    if ((forced > 0 && forced <= K_SELL) || forced == K_CSR_WINDOW) set_kernel(forced, env_int("EIGSOL_B200_LANES", 0));
    else set_kernel(K_AUTO, 0);
    compute_prescale();
    tr.mark("rows(dev): kernel choice + unit map");
}

// Derived layouts (SELL, row blocks, deferred maps) are built on the host from the CSR; after a device-side
// ingest the host has none yet.
void MatrixImpl::ensure_host_csr() {
    if (dense || !csr.rowptr.empty()) return;
    EIGSOL_CUDA(cudaSetDevice(device));
    csr.rowptr.resize(size_t(nloc) + 1);
    csr.colidx.resize(size_t(local_nnz));
    csr.vals.resize(size_t(local_nnz) * size_t(csr.spd()));
    EIGSOL_CUDA(cudaMemcpyAsync(csr.rowptr.data(), d_rowptr.p, sizeof(int32_t) * size_t(nloc + 1), cudaMemcpyDeviceToHost, stream));
    if (local_nnz) {
        EIGSOL_CUDA(cudaMemcpyAsync(csr.colidx.data(), d_colidx.p, sizeof(int32_t) * size_t(local_nnz), cudaMemcpyDeviceToHost, stream));
        EIGSOL_CUDA(cudaMemcpyAsync(csr.vals.data(), d_vals.p, size_t(local_nnz) * size_t(ss()), cudaMemcpyDeviceToHost, stream));
    }
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
}

// sigma0 from max |a_ij| over the resident matrix values (all ranks agree on the global maximum).
void MatrixImpl::compute_prescale() {
    sigma0 = 1.0;
    const double* vals = dense ? d_dense.as<double>() : d_vals.as<double>();
    const int64_t count = (dense ? lda * ext_len : local_nnz) * (dtype ? 2 : 1);
    DevBuf d_max;
    d_max.alloc(sizeof(unsigned long long));
    if (count > 0) {
        const int g = int(std::max<int64_t>(1, std::min<int64_t>((count + 255) / 256, int64_t(sm_count) * 8)));
        absmax_kernel<<<g, 256, 0, stream>>>(vals, count, d_max.as<unsigned long long>());
        ++launches;
        EIGSOL_CUDA(cudaGetLastError());
    }
    unsigned long long bits = 0;
    EIGSOL_CUDA(cudaMemcpyAsync(&bits, d_max.p, sizeof bits, cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    if (nranks > 1) {
        const std::vector<int64_t> all = comm_allgather_i64(comm, int64_t(bits));
        for (int64_t v : all) bits = std::max<unsigned long long>(bits, (unsigned long long)v);
    }
    double amax;
    std::memcpy(&amax, &bits, sizeof amax);
    sigma0 = prescale_from_amax(amax);
}

void MatrixImpl::drop_graph() {
    if (graph) cudaGraphExecDestroy(graph);
    if (graph_short) cudaGraphExecDestroy(graph_short);
    graph = graph_short = nullptr;
    graph_len = 0;
    for (auto& kv : batch_graphs) cudaGraphExecDestroy(kv.second);
    batch_graphs.clear();
}

void MatrixImpl::set_kernel(int kernel_, int lanes_) {
    drop_graph();
    if (dense) {
        kernel = K_GEMV;
        lanes = 0;
        configure_launch();
        return;
    }
    if (kernel_ == K_AUTO) {
        const KernelChoice ch = choose_kernel(dtype, stats);
        kernel_ = ch.kernel;
        lanes_ = ch.lanes;
    }
    if ((kernel_ < K_CSR_SCALAR || kernel_ > K_SELL) && kernel_ != K_CSR_WINDOW) throw PlanError("unknown sparse kernel id");
    if (kernel_ == K_CSR_VECTOR) {
        if (lanes_ == 0) lanes_ = 8;
        if (lanes_ != 2 && lanes_ != 4 && lanes_ != 8 && lanes_ != 16 && lanes_ != 32)
            throw PlanError("CSR_VECTOR lanes per row must be 2, 4, 8, 16 or 32");
    } else if (kernel_ == K_SELL) {
        if (lanes_ == 0) lanes_ = dtype ? 1 : 2;
        if (lanes_ != 1 && lanes_ != 2) throw PlanError("SELL rows per lane must be 1 or 2");
        ensure_host_csr();
        if (sell_R != lanes_) {
            SellHost s = build_sell(csr, lanes_);
            sell_R = lanes_;
            sell_nslices = s.nslices;
            sell_padded = s.padded();
            d_sell_off.alloc(sizeof(int64_t) * size_t(s.nslices + 1));
            d_sell_off.upload(s.slice_off.data(), sizeof(int64_t) * size_t(s.nslices + 1), stream);
            d_sell_col.alloc(sizeof(int32_t) * size_t(std::max<int64_t>(sell_padded, 1)));
            d_sell_col.upload(s.colidx.data(), sizeof(int32_t) * size_t(sell_padded), stream);
            d_sell_val.alloc(size_t(std::max<int64_t>(sell_padded, 1)) * ss());
            d_sell_val.upload(s.vals.data(), size_t(sell_padded) * ss(), stream);
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
        }
    } else if (kernel_ == K_CSR_STREAM) {
        lanes_ = 0;
        if (nblk == 0 && nloc > 0) {
            ensure_host_csr();
            const int tile = dtype ? StreamCfg<double2>::TILE : StreamCfg<double>::TILE;
            rowblk_h = build_row_blocks(nloc, csr.rowptr.data(), tile, kStreamMaxRows);
            const std::vector<int32_t>& rb = rowblk_h;
            nblk = int(rb.size()) - 1;
            d_rowblk.alloc(sizeof(int32_t) * rb.size());
            d_rowblk.upload(rb.data(), sizeof(int32_t) * rb.size(), stream);
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
        }
    } else if (kernel_ == K_CSR_WINDOW) {
        // per unit of 256 rows: the contiguous range of LOCAL columns its rows reference, 16-byte aligned and
        // truncated to the shared-memory budget; halo columns and outliers are gathered from global memory
        lanes_ = 0;
        ensure_host_csr();
        const int B = 256;
        const int64_t per16 = 16 / ss();  // elements per 16 bytes
        const int64_t budget = (int64_t(env_int("EIGSOL_B200_WINDOW_KB", 64)) << 10) / ss();
        const int64_t nunits = (nloc + B - 1) / B;
        std::vector<int32_t> w(size_t(std::max<int64_t>(nunits, 1)) * 2, 0);
        int64_t maxcount = 0;
        for (int64_t u = 0; u < nunits; ++u) {
            const int64_t r0 = u * B, r1 = std::min<int64_t>(r0 + B, nloc);
            int64_t cmin = INT64_MAX, cmax = -1;
            for (int32_t p = csr.rowptr[size_t(r0)]; p < csr.rowptr[size_t(r1)]; ++p) {
                const int64_t c = csr.colidx[size_t(p)];
                if (c >= nloc_pad) continue;  // halo entry
                cmin = std::min(cmin, c);
                cmax = std::max(cmax, c);
            }
            if (cmax < 0) continue;
            // centre a too-wide window on the unit's own rows (the diagonal band)
            int64_t first = cmin / per16 * per16;
            int64_t count = (cmax + 1 - first + per16 - 1) / per16 * per16;
            if (count > budget) {
                const int64_t mid = (r0 + r1) / 2;
                first = std::max<int64_t>(cmin, mid - budget / 2) / per16 * per16;
                count = budget / per16 * per16;
            }
            count = std::min<int64_t>(count, (nloc_pad - first) / per16 * per16);
            if (count <= 0) continue;
            w[size_t(u) * 2] = int32_t(first);
            w[size_t(u) * 2 + 1] = int32_t(count);
            maxcount = std::max(maxcount, count);
        }
        window_smem = int(std::max<int64_t>(maxcount * ss(), 16));
        d_window.alloc(sizeof(int32_t) * w.size());
        d_window.upload(w.data(), sizeof(int32_t) * w.size(), stream);
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
        if (dtype) {
            EIGSOL_CUDA(cudaFuncSetAttribute(spmv_csr_window_kernel<double2, 256, false>, attr, window_smem));
            EIGSOL_CUDA(cudaFuncSetAttribute(spmv_csr_window_kernel<double2, 256, true>, attr, window_smem));
        } else {
            EIGSOL_CUDA(cudaFuncSetAttribute(spmv_csr_window_kernel<double, 256, false>, attr, window_smem));
            EIGSOL_CUDA(cudaFuncSetAttribute(spmv_csr_window_kernel<double, 256, true>, attr, window_smem));
        }
    } else {
        lanes_ = 0;
    }
    kernel = kernel_;
    lanes = lanes_;
    configure_launch();
}

// -------------------------------------------------------------------------------------------------
// Dense: rows [lo, hi) of the column-major host matrix, columns re-ordered into the rank's extended
// numbering [own columns | gap | remote columns ascending] so the kernel sees a plain GEMV against
// the extended vector.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::init_dense(int64_t rows, const void* a, int64_t lda_host) {
    dense = true;
    n = rows;
    nnz_global = n * n;
    row_begin = partition_even(n, nranks);
    lo = row_begin[rank];
    hi = row_begin[rank + 1];
    nloc = hi - lo;
    local_nnz = nloc * n;
    halo = HaloPlan();
    halo.peer_count.assign(size_t(nranks), 0);
    halo.peer_off.assign(size_t(nranks) + 1, 0);
    if (nranks > 1) {
        for (int64_t c = 0; c < n; ++c)
            if (c < lo || c >= hi) halo.cols.push_back(int32_t(c));
        for (int q = 0; q < nranks; ++q) halo.peer_count[q] = q == rank ? 0 : row_begin[q + 1] - row_begin[q];
        for (int q = 0; q < nranks; ++q) halo.peer_off[q + 1] = halo.peer_off[q] + halo.peer_count[q];
    }
    nhalo = int64_t(halo.cols.size());
    alloc_vectors();
    plan_exchange();
    setup_peer();
    stats = RowStats();
    stats.rows = nloc;
    stats.nnz = local_nnz;
    stats.mean = stats.max = stats.min = double(n);
    lda = (nloc + 15) / 16 * 16;
    if (lda == 0) lda = 16;
    const char* src = static_cast<const char*>(a);
    const size_t s = size_t(ss());
    if (nranks == 1 && lda == n && lda_host == n && size_t(n) * size_t(n) * s >= (size_t(32) << 20)) {
        // the whole matrix is one contiguous column-major block with the device's leading dimension: several host
        // threads through pinned staging (a pageable cudaMemcpy2D moves ~12 GB/s, this ~40)
        d_dense.alloc_pooled(size_t(lda) * size_t(ext_len) * s, stream);
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        StagedUploader up;
        up.add(d_dense.p, src, size_t(n) * size_t(n) * s);
        EIGSOL_CUDA(up.run(device, upload_threads(nranks)));
        check_dense_symmetry();
        set_kernel(K_GEMV, 0);
        compute_prescale();
        return;
    }
    d_dense.alloc(size_t(lda) * size_t(std::max<int64_t>(ext_len, 1)) * ss());
    auto copy_cols = [&](int64_t gcol0, int64_t ncols, int64_t ecol0) {
        if (ncols <= 0 || nloc <= 0) return;
        EIGSOL_CUDA(cudaMemcpy2DAsync(d_dense.as<char>() + size_t(ecol0) * size_t(lda) * s, size_t(lda) * s,
                                      src + (size_t(gcol0) * size_t(lda_host) + size_t(lo)) * s, size_t(lda_host) * s,
                                      size_t(nloc) * s, size_t(ncols), cudaMemcpyHostToDevice, stream));
    };
    copy_cols(lo, nloc, 0);
    copy_cols(0, lo, nloc_pad);
    copy_cols(hi, n - hi, nloc_pad + lo);
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    check_dense_symmetry();
    set_kernel(K_GEMV, 0);
    compute_prescale();
}

// One GPU: is the resident dense matrix bitwise equal to its transpose / conjugate transpose?  Then the matvec can read
// the stored columns as rows (gemv_sym_kernel).  EIGSOL_B200_GEMV_SYM=0 switches the test (and that kernel) off.
void MatrixImpl::check_dense_symmetry() {
    dense_sym = 0;
    // Off by default: measured on C3 (8192^2, symmetric), the column-dot kernel runs 92.3 us per trip against 89.1 us of
    // the general kernel -- the launch boundary, not the streaming pattern, is what both lose (profiles/r02_notes.md).
    if (nranks != 1 || n < 1 || !env_int("EIGSOL_B200_GEMV_SYM", 0)) return;
    DevBuf d_flags;
    d_flags.alloc(64);
    int three = 3;
    EIGSOL_CUDA(cudaMemcpyAsync(d_flags.p, &three, sizeof three, cudaMemcpyHostToDevice, stream));
    const unsigned t = unsigned((n + 31) / 32);
    if (dtype) sym_check_kernel<double2><<<dim3(t, t), 256, 0, stream>>>(d_dense.as<double2>(), int(n), lda, d_flags.as<int>());
    else sym_check_kernel<double><<<dim3(t, t), 256, 0, stream>>>(d_dense.as<double>(), int(n), lda, d_flags.as<int>());
    ++launches;
    EIGSOL_CUDA(cudaGetLastError());
    int f = 0;
    EIGSOL_CUDA(cudaMemcpyAsync(&f, d_flags.p, sizeof f, cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    if (f & 2) dense_sym = 2;       // Hermitian (a real symmetric matrix is both: the conjugate is a no-op)
    else if (f & 1) dense_sym = 1;  // complex symmetric
}

// -------------------------------------------------------------------------------------------------
// Launch configuration: persistent grids sized from the SM count and the kernel's occupancy.
// -------------------------------------------------------------------------------------------------
template <class K> static int occupancy(K kern, int block, size_t smem = 0) {
    int nb = 0;
    EIGSOL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, block, smem));
    return std::max(nb, 1);
}

template <class S, bool PEER> static int sparse_occupancy_t(int kernel, int lanes, int window_smem, int block) {
    constexpr int B = 256;
    switch (kernel) {
        case K_CSR_WINDOW: return occupancy(spmv_csr_window_kernel<S, B, PEER>, B, size_t(window_smem));
        case K_CSR_SCALAR:
            return block == 128 ? occupancy(spmv_csr_scalar_kernel<S, 128, PEER>, 128) : occupancy(spmv_csr_scalar_kernel<S, B, PEER>, B);
        case K_CSR_STREAM: return occupancy(spmv_csr_stream_kernel<S, B, PEER>, B);
        case K_SELL:
            return lanes == 2 ? occupancy(spmv_sell_kernel<S, 2, B, PEER>, B) : occupancy(spmv_sell_kernel<S, 1, B, PEER>, B);
        case K_CSR_VECTOR:
            switch (lanes) {
                case 2: return occupancy(spmv_csr_vector_kernel<S, 2, B, PEER>, B);
                case 4: return occupancy(spmv_csr_vector_kernel<S, 4, B, PEER>, B);
                case 8: return occupancy(spmv_csr_vector_kernel<S, 8, B, PEER>, B);
                case 16: return occupancy(spmv_csr_vector_kernel<S, 16, B, PEER>, B);
                default: return occupancy(spmv_csr_vector_kernel<S, 32, B, PEER>, B);
            }
    }
    return 1;
}
template <class S> static int sparse_occupancy(int kernel, int lanes, int window_smem, bool peer, int block) {
    return peer ? sparse_occupancy_t<S, true>(kernel, lanes, window_smem, block) : sparse_occupancy_t<S, false>(kernel, lanes, window_smem, block);
}

void MatrixImpl::configure_launch() {
    // csr_scalar may run with 128-thread CTAs (EIGSOL_B200_SCALAR_BLOCK=128): work units of 128 rows halve the
    // quantisation of a small per-rank problem over the persistent grid (8 GPUs on C5: 9.2 units of 256 rows per CTA)
    block = (!dense && kernel == K_CSR_SCALAR && scalar_block == 128) ? 128 : 256;
    const int64_t s = ss();
    if (dense && dense_sym && nranks == 1) {
        // flat, perfectly balanced partition of the stored matrix over the CTAs of a persistent grid (gemv_sym_kernel)
        const int64_t total = lda * n;
        auto plan = [&](int64_t nc) {
            int64_t per = ((total + nc - 1) / nc + 127) / 128 * 128;
            return std::max<int64_t>(per, (lda + 127) / 128 * 128);  // at least one column: a column meets <= 2 CTAs
        };
        // the shared memory (8 warp sums per column of a range) depends on the range, the occupancy on the shared memory
        int64_t nc = std::min<int64_t>(int64_t(sm_count) * 4, n);
        for (int pass = 0; pass < 2; ++pass) {
            const int64_t per = plan(nc);
            sym_smem = int((per / lda + 3) * (block / 32) * s);
            if (sym_smem > 48 * 1024) {
                if (dtype) EIGSOL_CUDA(cudaFuncSetAttribute(gemv_sym_kernel<double2, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, sym_smem));
                else EIGSOL_CUDA(cudaFuncSetAttribute(gemv_sym_kernel<double, 256>, cudaFuncAttributeMaxDynamicSharedMemorySize, sym_smem));
            }
            const int occ = dtype ? occupancy(gemv_sym_kernel<double2, 256>, 256, size_t(sym_smem)) : occupancy(gemv_sym_kernel<double, 256>, 256, size_t(sym_smem));
            nc = std::min<int64_t>(int64_t(sm_count) * occ, n);
        }
        sym_per = plan(nc);
        sym_smem = int((sym_per / lda + 3) * (block / 32) * s);
        sym_nctas = int((total + sym_per - 1) / sym_per);
        nslots = sym_nctas;
        grid = nslots + 1;
        ntiles = int(n);  // tickets: one per column
        d_dense_partial.alloc(size_t(2) * size_t(n) * size_t(s));
        d_tile_ticket.alloc(sizeof(unsigned) * size_t(n));
        stored_bytes = s * lda * n + 2 * s * n;
    } else if (dense) {
        const int RT = block * (dtype ? 1 : 2) * (gemv_var == 2 ? 2 : 1);  // rows per CTA (kernels.cuh: GemvRows)
        ntiles = int(std::max<int64_t>((nloc + RT - 1) / RT, 1));
        const int target = sm_count * env_int("EIGSOL_B200_GEMV_WAVES", 5);
        nseg = std::max(1, target / ntiles);
        nseg = int(std::min<int64_t>(nseg, std::max<int64_t>(1, (ext_len + 31) / 32)));
        grid = ntiles * nseg + 1;  // + the control CTA
        nslots = ntiles;
        d_dense_partial.alloc(size_t(nseg) * size_t(lda) * size_t(s));
        d_tile_ticket.alloc(sizeof(unsigned) * size_t(ntiles));
        stored_bytes = s * lda * ext_len + s * (ext_len + nloc) + 2 * s * int64_t(nseg) * lda;
    } else {
        const bool peer = xmode == MODE_PEER;
        const int occ = dtype ? sparse_occupancy<double2>(kernel, lanes, window_smem, peer, block) : sparse_occupancy<double>(kernel, lanes, window_smem, peer, block);
        const int resident = sm_count * occ;
        int64_t need = 1;
        switch (kernel) {
            case K_CSR_WINDOW:
            case K_CSR_SCALAR: need = (nloc + block - 1) / block; break;
            case K_CSR_VECTOR: need = (nloc * lanes + block - 1) / block; break;
            case K_CSR_STREAM: need = nblk; break;
            case K_SELL: need = (sell_nslices + (block / 32) - 1) / (block / 32); break;
        }
        // persistent grid: every CTA resident at once (the control CTA and the halo waits rely on it)
        nslots = int(std::max<int64_t>(1, std::min<int64_t>(need, resident - 1)));
        grid = nslots + 1;  // + the control CTA
        const int64_t xz = s * (nloc + nhalo) + s * nloc;
        if (kernel == K_SELL) stored_bytes = sell_padded * (s + 4) + 8 * (sell_nslices + 1) + xz;
        else stored_bytes = local_nnz * (s + 4) + 4 * (nloc + 1) + xz + (kernel == K_CSR_STREAM ? 4 * int64_t(nblk + 1) : 0);
    }
    d_partials.alloc(sizeof(double) * 3 * size_t(nslots) * kRing);
    build_deferred();
}

// MODE_PEER: which work units of the chosen kernel contain a row that reads a halo entry.  Those are
// processed after the halo wait, everything else before it (kernels.cuh: walk_units).
void MatrixImpl::build_deferred() {
    have_deferred = false;
    unit_start = 0;
    if (dense || xmode != MODE_PEER || nhalo == 0 || nloc == 0) return;
    if (!env_int("EIGSOL_B200_DEFER", 1)) return;  // 0: every CTA waits for the halo before its first row
    int64_t nunits = 0;
    int rows_per_unit = 0;
    const int32_t* blk = nullptr;
    switch (kernel) {
        case K_CSR_WINDOW:
        case K_CSR_SCALAR: rows_per_unit = block; break;
        case K_CSR_VECTOR: rows_per_unit = block / lanes; break;
        case K_CSR_STREAM: blk = d_rowblk.as<int32_t>(); nunits = nblk; break;
        case K_SELL: rows_per_unit = 32 * sell_R; break;
        default: return;
    }
    if (!blk) nunits = (nloc + rows_per_unit - 1) / rows_per_unit;
    if (nunits <= 0) return;
    // computed on the device from the resident (localized) CSR arrays, one warp per unit
    d_deferred.alloc(size_t(nunits));
    {
        const int B = 256;
        const unsigned g = unsigned((nunits * 32 + B - 1) / B);
        deferred_units_kernel<<<g, B, 0, stream>>>(d_rowptr.as<int32_t>(), d_colidx.as<int32_t>(), nloc, nunits, rows_per_unit, blk,
                                                  int32_t(nloc_pad), d_deferred.as<uint8_t>());
        ++launches;
        EIGSOL_CUDA(cudaGetLastError());
    }
    std::vector<uint8_t> def(size_t(nunits), 0);
    EIGSOL_CUDA(cudaMemcpyAsync(def.data(), d_deferred.p, size_t(nunits), cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    // longest circular run of units that read no halo entry; the walk starts in its middle
    {
        int64_t best_len = 0, best_begin = 0, run = 0;
        for (int64_t k = 0; k < 2 * nunits; ++k) {
            const int64_t u = k % nunits;
            if (def[size_t(u)]) {
                run = 0;
            } else {
                ++run;
                if (run > best_len && run <= nunits) {
                    best_len = run;
                    best_begin = k - run + 1;
                }
            }
        }
        unit_start = nunits > 0 ? int((best_begin + best_len / 2) % nunits) : 0;
    }
    have_deferred = true;
}

// One matvec launch on `stream`.  All launches of a handle are identical (the kernels pick the
// ping-pong buffer from the device state), which is what lets a CUDA graph of them be replayed.
// Kernel launch with the programmatic-stream-serialization attribute (PDL) when enabled.
template <class... KArgs, class... Args>
static void launch_pdl(bool pdl, cudaStream_t st, void (*kern)(KArgs...), dim3 g, dim3 b, size_t smem, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = g;
    cfg.blockDim = b;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl ? 1 : 0;
    EIGSOL_CUDA(cudaLaunchKernelEx(&cfg, kern, KArgs(args)...));
}

template <class S, bool PEER> static void launch_sparse_t(MatrixImpl& m, int ph) {
    constexpr int B = 256;
    IterArgs<S> it = m.iter_args<S>();
    CsrView<S> A{int32_t(m.nloc), m.d_rowptr.as<int32_t>(), m.d_colidx.as<int32_t>(), m.d_vals.as<S>()};
    const dim3 g(m.grid), b(B);
    cudaStream_t st = m.stream;
    switch (m.kernel) {
        case K_CSR_SCALAR:
            if (m.block == 128) launch_pdl(m.pdl, st, spmv_csr_scalar_kernel<S, 128, PEER>, g, dim3(128), 0, A, it, ph);
            else launch_pdl(m.pdl, st, spmv_csr_scalar_kernel<S, B, PEER>, g, b, 0, A, it, ph);
            break;
        case K_CSR_WINDOW:
            launch_pdl(m.pdl, st, spmv_csr_window_kernel<S, B, PEER>, g, b, size_t(m.window_smem), A, m.d_window.as<int2>(), it, ph);
            break;
        case K_CSR_STREAM: launch_pdl(m.pdl, st, spmv_csr_stream_kernel<S, B, PEER>, g, b, 0, A, m.d_rowblk.as<int32_t>(), m.nblk, it, ph); break;
        case K_SELL: {
            SellView<S> V{int32_t(m.nloc), int32_t(m.sell_nslices), m.d_sell_off.as<int64_t>(), m.d_sell_col.as<int32_t>(),
                          m.d_sell_val.as<S>()};
            if (m.lanes == 2) launch_pdl(m.pdl, st, spmv_sell_kernel<S, 2, B, PEER>, g, b, 0, V, it, ph);
            else launch_pdl(m.pdl, st, spmv_sell_kernel<S, 1, B, PEER>, g, b, 0, V, it, ph);
            break;
        }
        case K_CSR_VECTOR:
            switch (m.lanes) {
                case 2: launch_pdl(m.pdl, st, spmv_csr_vector_kernel<S, 2, B, PEER>, g, b, 0, A, it, ph); break;
                case 4: launch_pdl(m.pdl, st, spmv_csr_vector_kernel<S, 4, B, PEER>, g, b, 0, A, it, ph); break;
                case 8: launch_pdl(m.pdl, st, spmv_csr_vector_kernel<S, 8, B, PEER>, g, b, 0, A, it, ph); break;
                case 16: launch_pdl(m.pdl, st, spmv_csr_vector_kernel<S, 16, B, PEER>, g, b, 0, A, it, ph); break;
                default: launch_pdl(m.pdl, st, spmv_csr_vector_kernel<S, 32, B, PEER>, g, b, 0, A, it, ph); break;
            }
            break;
        default: throw StateError("no sparse kernel selected");
    }
}
template <class S> static void launch_sparse(MatrixImpl& m, int ph) {
    if (m.xmode == MODE_PEER) launch_sparse_t<S, true>(m, ph);
    else launch_sparse_t<S, false>(m, ph);
}

template <class S> static void launch_dense(MatrixImpl& m, int ph) {
    constexpr int B = 256;
    IterArgs<S> it = m.iter_args<S>();
    DenseView<S> A{int32_t(m.nloc), int32_t(m.ext_len), m.lda, m.d_dense.as<S>(), m.nseg, m.ntiles,
                   m.d_dense_partial.as<S>(), m.d_tile_ticket.as<unsigned>(), m.gemv_map};
    const dim3 g(m.grid), b(B);
    if (m.dense_sym && m.nranks == 1) {
        SymDenseView<S> V{int32_t(m.n), m.lda, m.d_dense.as<S>(), m.sym_per, m.sym_nctas, m.d_dense_partial.as<S>(),
                          m.d_tile_ticket.as<unsigned>(), m.dense_sym == 2 ? 1 : 0};
        launch_pdl(m.pdl, m.stream, gemv_sym_kernel<S, B>, g, b, size_t(m.sym_smem), V, it, ph);
        return;
    }
    auto go = [&](auto kern) { launch_pdl(m.pdl, m.stream, kern, g, b, 0, A, it, ph); };
    if (m.xmode == MODE_PEER) {
        if (m.gemv_var == 3) go(gemv_colmajor_kernel<S, B, true, 3>);
        else if (m.gemv_var == 2) go(gemv_colmajor_kernel<S, B, true, 2>);
        else if (m.gemv_var == 1) go(gemv_colmajor_kernel<S, B, true, 1>);
        else go(gemv_colmajor_kernel<S, B, true, 0>);
    } else {
        if (m.gemv_var == 3) go(gemv_colmajor_kernel<S, B, false, 3>);
        else if (m.gemv_var == 2) go(gemv_colmajor_kernel<S, B, false, 2>);
        else if (m.gemv_var == 1) go(gemv_colmajor_kernel<S, B, false, 1>);
        else go(gemv_colmajor_kernel<S, B, false, 0>);
    }
}

void MatrixImpl::launch_matvec(int phase) {
    if (dense) {
        if (dtype) launch_dense<double2>(*this, phase);
        else launch_dense<double>(*this, phase);
    } else {
        if (dtype) launch_sparse<double2>(*this, phase);
        else launch_sparse<double>(*this, phase);
    }
    ++launches;
}

// -------------------------------------------------------------------------------------------------
// The loop.
// -------------------------------------------------------------------------------------------------
void MatrixImpl::begin(const void* x0, int max_it, double tol, bool unit_scale) {
    if (!x0) throw PlanError("null start vector");
    EIGSOL_CUDA(cudaSetDevice(device));
    PowerState st;
    std::memset(&st, 0, sizeof st);
    st.q[0] = 1.0;  // x0 arrives normalised (power_method.hpp:62); launch 0 uses it as is
    st.sigma[0] = unit_scale ? 1.0 : sigma0;
    st.tol = tol;
    st.final_q = 1.0;
    st.max_trips = max_it;
    st.stop = max_it <= 0 ? STOP_MAX_TRIPS : STOP_NONE;  // the loop of :68 never runs
    st.ctrl[0].sigma = st.sigma[0];
    st.ctrl[0].launch_no = 0;
    st.ctrl[0].active = max_it > 0 ? 1 : 0;
    max_trips = max_it;
    flag_epoch += (unsigned long long)issued + 1;  // flags are monotonic across solves: stale late stores stay below
    st.flag_base = flag_epoch;
    issued = 0;
    const size_t s = size_t(ss());
    EIGSOL_CUDA(cudaMemcpyAsync(d_state.p, &st, sizeof st, cudaMemcpyHostToDevice, stream));
    bool in_place = x0_preloaded && x0 == preload_x0;  // one-shot call: x0 came up with the matrix
    x0_preloaded = false;
    preload_x0 = nullptr;
    const bool staged_slice = in_place && d_x0pre.p != nullptr;  // group rank: the rows sit in d_x0pre, not in buf[0] yet
    if (staged_slice) in_place = false;
    for (int b = in_place ? 1 : 0; b < kRing; ++b) EIGSOL_CUDA(cudaMemsetAsync(buf[b].p, 0, size_t(ext_len) * s, stream));
    if (staged_slice) {
        if (nloc) EIGSOL_CUDA(cudaMemcpyAsync(buf[0].p, d_x0pre.p, size_t(nloc) * s, cudaMemcpyDeviceToDevice, stream));
        EIGSOL_CUDA(cudaStreamSynchronize(stream));
        d_x0pre.release();
    } else if (nloc && !in_place) {
        const char* src = static_cast<const char*>(x0) + size_t(lo) * s;
        if (size_t(nloc) * s >= (size_t(32) << 20)) {
            // a large pageable vector: several host threads through pinned staging instead of one synchronous copy
            EIGSOL_CUDA(cudaStreamSynchronize(stream));  // the zero fill above must not land after the upload
            StagedUploader up;
            up.add(buf[0].p, src, size_t(nloc) * s);
            EIGSOL_CUDA(up.run(device, upload_threads(nranks, comm && comm->local)));
        } else {
            EIGSOL_CUDA(cudaMemcpyAsync(buf[0].p, src, size_t(nloc) * s, cudaMemcpyHostToDevice, stream));
        }
    }
    if (dense) EIGSOL_CUDA(cudaMemsetAsync(d_tile_ticket.p, 0, sizeof(unsigned) * size_t(ntiles), stream));
    if (nranks > 1) {
        // x0 is global on every rank: the halo of launch 0 comes straight from it (no exchange needed).  The halo
        // columns are sorted, so a stencil's halo is a handful of contiguous runs: one copy per run; a scattered
        // halo is gathered on the host first.
        if (nhalo) {
            const char* src = static_cast<const char*>(x0);
            char* dst = buf[0].as<char>() + size_t(nloc_pad) * s;
            std::vector<std::pair<int64_t, int64_t>> runs;  // {first halo slot, length}
            for (int64_t k = 0; k < nhalo;) {
                int64_t e = k + 1;
                while (e < nhalo && halo.cols[size_t(e)] == halo.cols[size_t(e - 1)] + 1) ++e;
                runs.emplace_back(k, e - k);
                k = e;
                if (runs.size() > 64) break;
            }
            if (runs.size() <= 64) {
                for (const auto& r : runs)
                    EIGSOL_CUDA(cudaMemcpyAsync(dst + size_t(r.first) * s, src + size_t(halo.cols[size_t(r.first)]) * s,
                                                size_t(r.second) * s, cudaMemcpyHostToDevice, stream));
            } else {
                std::vector<char> hx(size_t(nhalo) * s);
                for (int64_t k = 0; k < nhalo; ++k) std::memcpy(hx.data() + size_t(k) * s, src + size_t(halo.cols[size_t(k)]) * s, s);
                EIGSOL_CUDA(cudaMemcpyAsync(dst, hx.data(), hx.size(), cudaMemcpyHostToDevice, stream));
                EIGSOL_CUDA(cudaStreamSynchronize(stream));
            }
        }
        // barrier: no rank may start launching (and storing into peers) before every rank has reset
        rank_barrier();
    }
    ensure_graphs();  // captured and instantiated here, never inside a timed batch of trips
    EIGSOL_CUDA(cudaStreamSynchronize(stream));  // x0 is a borrowed host pointer
    begun = true;
}

// Barrier of the ranks on the device timeline of `stream`: what is queued behind it starts on every rank within
// microseconds of the last rank's arrival.  MODE_PEER: the gate kernel (flags in peer memory); NCCL: an all-reduce.
void MatrixImpl::rank_barrier() {
    if (nranks == 1) return;
    if (xmode == MODE_PEER) {
        ++gate_epoch;
        peer_gate_kernel<<<1, 32, 0, stream>>>(d_peer.as<PeerArgs>(), gate_epoch, d_state.as<PowerState>());
        ++launches;
        EIGSOL_CUDA(cudaGetLastError());
    } else {
        EIGSOL_NCCL(nccl().AllReduce(d_scal.p, d_scal.p, 1, ncclDouble, ncclSum, comm->nccl_comm, stream));
    }
}

// x exchange of the multi-GPU path: every rank sends the entries its peers' rows reference and
// receives its own halo (semantically the all-gather of x; only the referenced entries move).
template <class S> static void exchange_impl(MatrixImpl& m, int in) {
    S* u = m.buf[in].as<S>();  // in = launch number % kRing
    const int ncclT = ncclDouble;
    const size_t mul = sizeof(S) / sizeof(double);
    if (m.need_pack) {
        for (int q = 0; q < m.nranks; ++q) {
            if (m.send_count[q] == 0 || m.send_contig[q]) continue;
            const int cnt = int(m.send_count[q]);
            const int g = std::min((cnt + 255) / 256, m.sm_count * 4);
            pack_kernel<S><<<g, 256, 0, m.stream>>>(u, m.d_send_idx.as<int32_t>() + m.send_off[q],
                                                   m.d_sendbuf.as<S>() + m.send_off[q], cnt);
            ++m.launches;
        }
    }
    EIGSOL_NCCL(nccl().GroupStart());
    for (int q = 0; q < m.nranks; ++q) {
        if (q == m.rank) continue;
        if (m.send_count[q] > 0) {
            const S* src = m.send_contig[q] ? u + m.send_start[q] : m.d_sendbuf.as<S>() + m.send_off[q];
            EIGSOL_NCCL(nccl().Send(src, size_t(m.send_count[q]) * mul, ncclT, q, m.comm->nccl_comm, m.stream));
        }
        if (m.halo.peer_count[q] > 0)
            EIGSOL_NCCL(nccl().Recv(u + m.nloc_pad + m.halo.peer_off[q], size_t(m.halo.peer_count[q]) * mul, ncclT, q,
                                    m.comm->nccl_comm, m.stream));
    }
    EIGSOL_NCCL(nccl().GroupEnd());
}

void MatrixImpl::exchange_halo(int in) {
    if (nranks == 1) return;
    if (dtype) exchange_impl<double2>(*this, in);
    else exchange_impl<double>(*this, in);
}

// CUDA graphs of kGraphLong and kRing matvec launches (multiples of the ring length, so every node's ring phase is
// a constant), with the programmatic-dependent-launch edges between the nodes.  Built once per launch
// configuration, at begin() -- outside any timed region.
void MatrixImpl::ensure_graphs() {
    if (xmode == MODE_SPLIT) return;
    int glen = env_int("EIGSOL_B200_GRAPH", 15);
    glen = glen / kRing * kRing;
    auto capture = [&](int len) {
        cudaGraph_t g = nullptr;
        cudaGraphExec_t ge = nullptr;
        EIGSOL_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
        const int64_t keep = launches;
        try {
            for (int i = 0; i < len; ++i) launch_matvec(i % kRing);
        } catch (...) {
            cudaStreamEndCapture(stream, &g);
            if (g) cudaGraphDestroy(g);
            throw;
        }
        launches = keep;
        EIGSOL_CUDA(cudaStreamEndCapture(stream, &g));
        const cudaError_t e = cudaGraphInstantiate(&ge, g, 0);
        cudaGraphDestroy(g);
        EIGSOL_CUDA(e);
        EIGSOL_CUDA(cudaGraphUpload(ge, stream));
        return ge;
    };
    if (glen >= kRing && !graph_short) graph_short = capture(kRing);
    if (glen > kRing && (!graph || graph_len != glen)) {
        if (graph) cudaGraphExecDestroy(graph);
        graph = capture(glen);
        graph_len = glen;
    }
}

// One graph of `len` launches starting at ring phase `phase`; a handful are kept per launch configuration.
cudaGraphExec_t MatrixImpl::batch_graph(int len, int phase) {
    const std::pair<int, int> key(len, phase);
    auto it = batch_graphs.find(key);
    if (it != batch_graphs.end()) return it->second;
    if (batch_graphs.size() >= 6) {
        for (auto& kv : batch_graphs) cudaGraphExecDestroy(kv.second);
        batch_graphs.clear();
    }
    cudaGraph_t g = nullptr;
    cudaGraphExec_t ge = nullptr;
    EIGSOL_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
    const int64_t keep = launches;
    try {
        for (int i = 0; i < len; ++i) launch_matvec((phase + i) % kRing);
    } catch (...) {
        cudaStreamEndCapture(stream, &g);
        if (g) cudaGraphDestroy(g);
        throw;
    }
    launches = keep;
    EIGSOL_CUDA(cudaStreamEndCapture(stream, &g));
    const cudaError_t e = cudaGraphInstantiate(&ge, g, 0);
    cudaGraphDestroy(g);
    EIGSOL_CUDA(e);
    EIGSOL_CUDA(cudaGraphUpload(ge, stream));
    batch_graphs[key] = ge;
    return ge;
}

void MatrixImpl::issue_trips(int k) {
    if (k <= 0) return;
    EIGSOL_CUDA(cudaSetDevice(device));
    if (xmode != MODE_SPLIT) {
        // single launches up to a ring boundary, then the long graph, then the 3-launch graph, then the rest
        while (k > 0 && issued % kRing != 0) {
            launch_matvec(issued % kRing);
            ++issued;
            --k;
        }
        while (graph && k >= graph_len) {
            EIGSOL_CUDA(cudaGraphLaunch(graph, stream));
            launches += graph_len;
            issued += graph_len;
            k -= graph_len;
        }
        while (graph_short && k >= kRing) {
            EIGSOL_CUDA(cudaGraphLaunch(graph_short, stream));
            launches += kRing;
            issued += kRing;
            k -= kRing;
        }
        for (; k > 0; --k) {
            launch_matvec(issued % kRing);
            ++issued;
        }
        EIGSOL_CUDA(cudaGetLastError());
        return;
    }
    // MODE_SPLIT (NCCL): halo send/recv -> matvec (its control CTA folds the all-reduced sums of the previous
    // launch) -> fixed-order sum of the CTA partials -> all-reduce of 3 doubles
    for (int i = 0; i < k; ++i) {
        const int ph = issued % kRing;
        if (issued > 0) exchange_halo(ph);  // the halo of launch 0 came from the host's x0
        launch_matvec(ph);
        double* sc = d_scal.as<double>() + 4 * ph;
        reduce_partials_kernel<256><<<1, 256, 0, stream>>>(d_partials.as<double>() + size_t(ph) * size_t(nslots) * 3, nslots, sc);
        EIGSOL_NCCL(nccl().AllReduce(sc, sc, 3, ncclDouble, ncclSum, comm->nccl_comm, stream));
        launches += 1;
        ++issued;
    }
    EIGSOL_CUDA(cudaGetLastError());
}

const PowerState& MatrixImpl::poll_state() {
    // fold the sums of the last executed launch into the state (MODE_PEER: exchanges them with the peers)
    if (dtype) control_kernel<double2><<<1, 256, 0, stream>>>(iter_args<double2>());
    else control_kernel<double><<<1, 256, 0, stream>>>(iter_args<double>());
    ++launches;
    EIGSOL_CUDA(cudaMemcpyAsync(h_state.p, d_state.p, sizeof(PowerState), cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    if (h_state.as<PowerState>()->stop == STOP_ERROR)
        throw DeviceError("powerMethod: device error: a peer GPU did not signal within the timeout "
                          "(EIGSOL_B200_PEER_TIMEOUT_MS); the ranks are out of step");
    return *h_state.as<PowerState>();
}

void MatrixImpl::step(int trips, float* ms_out) {
    if (!begun) throw StateError("eigsol_b200_power_step called before eigsol_b200_power_begin");
    if (trips < 0) throw PlanError("negative trip count");
    EIGSOL_CUDA(cudaSetDevice(device));
    // A short batch runs as ONE graph of exactly `trips` launches (every launch chained to its predecessor by a
    // programmatic-dependent-launch edge; the 15 + 3 + single scheme of issue_trips has a full dependency at every graph
    // boundary).  The graph is captured and instantiated HERE, before the gate and the first event: host time, not in
    // the timed window.
    cudaGraphExec_t whole = nullptr;
    if (xmode != MODE_SPLIT && trips >= 4 && trips <= 64 && env_int("EIGSOL_B200_GRAPH", 15) >= kRing) whole = batch_graph(trips, issued % kRing);
    rank_barrier();  // the ranks' timers start together, whatever the skew of their host threads
    EIGSOL_CUDA(cudaEventRecord(ev0, stream));
    if (whole) {
        EIGSOL_CUDA(cudaGraphLaunch(whole, stream));
        launches += trips;
        issued += trips;
    } else {
        issue_trips(trips);
    }
    EIGSOL_CUDA(cudaEventRecord(ev1, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    if (ms_out) EIGSOL_CUDA(cudaEventElapsedTime(ms_out, ev0, ev1));
}

// Local slice on the device -> global vector on the host (every rank gets the full vector).
void MatrixImpl::gather_download(const void* d_local, void* host_out) {
    const size_t s = size_t(ss());
    auto download = [&](void* dst, const void* src, size_t bytes) {
        EIGSOL_CUDA(cudaStreamSynchronize(stream));  // src is complete
        if (bytes >= (size_t(32) << 20)) {
            StagedUploader down;  // device -> pinned -> pageable host array, several threads
            down.download = true;
            down.add(dst, src, bytes);
            EIGSOL_CUDA(down.run(device, upload_threads(nranks, comm && comm->local)));
        } else if (bytes) {
            EIGSOL_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, stream));
            EIGSOL_CUDA(cudaStreamSynchronize(stream));
        }
    };
    if (nranks == 1) {
        download(host_out, d_local, size_t(n) * s);
        return;
    }
    if (comm->local) {
        // single process: host_out is the SAME buffer on every rank; each writes its own rows
        download(static_cast<char*>(host_out) + size_t(lo) * s, d_local, size_t(nloc) * s);
        comm->local->barrier();  // the whole vector is in place when any rank returns
        return;
    }
    const size_t mul = s / sizeof(double);
    EIGSOL_NCCL(nccl().GroupStart());
    for (int r = 0; r < nranks; ++r) {
        const int64_t cnt = row_begin[r + 1] - row_begin[r];
        if (cnt == 0) continue;
        EIGSOL_NCCL(nccl().Broadcast(d_local, d_xglob.as<char>() + size_t(row_begin[r]) * s, size_t(cnt) * mul, ncclDouble,
                                     r, comm->nccl_comm, stream));
    }
    EIGSOL_NCCL(nccl().GroupEnd());
    EIGSOL_CUDA(cudaMemcpyAsync(host_out, d_xglob.p, size_t(n) * s, cudaMemcpyDeviceToHost, stream));
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
}

void MatrixImpl::end(void* x_out, double lambda_out[2], int* iters, int* conv) {
    if (!begun) throw StateError("eigsol_b200_power_end called before eigsol_b200_power_begin");
    EIGSOL_CUDA(cudaSetDevice(device));
    const PowerState st = poll_state();
    if (x_out) {
        const int g = int(std::max<int64_t>(1, std::min<int64_t>((nloc + 255) / 256, sm_count * 8)));
        if (dtype)
            normalize_result_kernel<double2><<<g, 256, 0, stream>>>(buf[0].as<double2>(), buf[1].as<double2>(), buf[2].as<double2>(),
                                                                   d_state.as<PowerState>(), d_xloc.as<double2>(), int(nloc));
        else
            normalize_result_kernel<double><<<g, 256, 0, stream>>>(buf[0].as<double>(), buf[1].as<double>(), buf[2].as<double>(),
                                                                  d_state.as<PowerState>(), d_xloc.as<double>(), int(nloc));
        ++launches;
        EIGSOL_CUDA(cudaGetLastError());
        gather_download(d_xloc.p, x_out);
    }
    if (lambda_out) {
        lambda_out[0] = st.lambda[0];
        lambda_out[1] = st.lambda[1];
    }
    if (iters) *iters = st.trips;
    if (conv) *conv = st.converged;
}

// Whole solve: power_method.hpp:61-98.  maxIterations trips need maxIterations+1 launches (launch 0
// is y = A x0 of trip 0; the sums of launch l >= 1 close trip l-1; the poll folds the last ones).  The
// host polls the device state between batches; launches issued past the stopping point are no-ops.
void MatrixImpl::run(const void* x0, int max_it, double tol, void* x_out, double lambda_out[2], int* iters, int* conv) {
    Trace tr;
    begin(x0, max_it, tol);
    tr.mark("run: begin (x0 upload)");
    if (max_it > 0) {
        int64_t remaining = int64_t(max_it) + 1;
        int batch = env_int("EIGSOL_B200_POLL", 64);
        if (batch < 1) batch = 1;
        while (remaining > 0) {
            const int k = int(std::min<int64_t>(remaining, batch));
            issue_trips(k);
            remaining -= k;
            if (poll_state().stop != STOP_NONE) break;
        }
    }
    tr.mark("run: iterate");
    end(x_out, lambda_out, iters, conv);
    tr.mark("run: end (normalise+download)");
}

// y = A x alone: launch 0 of the loop with ||x|| taken as 1 leaves A x un-normalised in buffer 1.
void MatrixImpl::matvec(const void* x, void* y) {
    begin(x, 1, -1.0, true);
    issue_trips(1);
    EIGSOL_CUDA(cudaStreamSynchronize(stream));
    gather_download(buf[1].p, y);
    begun = false;
}

// x^H (A x) and ||A x||^2 in one fused launch (the Rayleigh-quotient step of the reference's shifted inverse
// method, src/power_method/shifted_inverse_power_solver.hpp:62, `x.dot(A * x)`): launch 0 of the loop with
// sigma = 1 leaves exactly these sums; the control-only launch folds them into the state.
void MatrixImpl::rayleigh(const void* x, double out[4]) {
    begin(x, 1, -1.0, true);
    issue_trips(1);
    const PowerState st = poll_state();
    out[0] = st.last_a[0];
    out[1] = st.last_a[1];
    out[2] = st.q[1];
    out[3] = 0.0;
    begun = false;
}

void MatrixImpl::fill_info(eigsol_b200_matrix_info* info) const {
    std::memset(info, 0, sizeof *info);
    info->dtype = dtype;
    info->is_dense = dense ? 1 : 0;
    info->n = n;
    info->nnz = nnz_global;
    info->local_rows = nloc;
    info->local_nnz = local_nnz;
    info->halo = nhalo;
    info->kernel = kernel;
    info->lanes_per_row = lanes;
    info->grid = grid;
    info->block = block;
    info->row_len_mean = stats.mean;
    info->row_len_max = stats.max;
    info->row_len_cv = stats.cv;
    const int64_t s = ss();
    info->alg_bytes = dense ? s * nloc * n + s * (n + nloc)
                            : local_nnz * (s + 4) + 4 * (nloc + 1) + s * (nloc + nhalo) + s * nloc;
    info->stored_bytes = stored_bytes;
    info->nranks = nranks;
    info->rank = rank;
    int per_trip = 1;
    if (xmode == MODE_SPLIT) {
        per_trip += 1;
        for (int q = 0; q < nranks; ++q)
            if (send_count[q] > 0 && !send_contig[q]) ++per_trip;
    }
    info->exchange = xmode;
    info->kernels_per_trip = per_trip;
}

// -------------------------------------------------------------------------------------------------
// Single-process multi-GPU group behind ONE matrix handle (SURVEY.md 8e "Process model"; EIGSOL_B200_NGPU=P):
// P persistent host threads, one per device, each driving one rank (MatrixImpl) exactly as a process-per-GPU
// rank would -- the same partition, the same fused peer-memory kernel -- so that the reference-facing call
// EigSol::powerMethod(M, opts) shards over the GPUs of the box without the caller knowing.
// -------------------------------------------------------------------------------------------------
// The host threads, streams and meeting point of a group over devices [first, first + P) live for the life of the
// process (thread start, stream creation and granting the peers access to the memory pools cost milliseconds -- more
// than the 20 trips of a short solve); every group handle over the same devices borrows them, one job at a time.
struct GroupWorkers {
    const int P;
    LocalGroup group;
    std::vector<int> devices;
    std::vector<std::unique_ptr<Comm>> comms;
    std::vector<std::thread> threads;
    std::mutex job_mu;  // one job (of any handle) at a time: the ranks of two handles never interleave in the barriers
    std::mutex mu;
    std::condition_variable cv_work, cv_done;
    std::function<void(int)> job;
    uint64_t job_id = 0;
    int pending = 0;
    std::vector<std::exception_ptr> errors;
    std::vector<uint64_t> error_seq;
    uint64_t seq = 0;

    GroupWorkers(int nranks, int first_device) : P(nranks), group(nranks) {
        for (int r = 0; r < P; ++r) devices.push_back(first_device + r);
        comms.resize(size_t(P));
        errors.resize(size_t(P));
        error_seq.assign(size_t(P), 0);
        for (int r = 0; r < P; ++r) threads.emplace_back([this, r] { worker(r); });
        for (auto& t : threads) t.detach();  // they serve until the process exits
        run_all([this](int r) {
            std::unique_ptr<Comm> c(new Comm);
            c->rank = r;
            c->nranks = P;
            c->device = devices[size_t(r)];
            c->local = &group;
            c->local_devices = devices;
            EIGSOL_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
            {
                // let the peers of the group read and write this device's stream-ordered pool: slabs, ring buffers and
                // temporaries of a group can then come from the pool like those of a single GPU
                cudaMemPool_t pool = nullptr;
                bool ok = cudaDeviceGetDefaultMemPool(&pool, c->device) == cudaSuccess && pool;
                for (int q = 0; q < P && ok; ++q) {
                    if (q == r) continue;
                    int can = 0;
                    ok = cudaDeviceCanAccessPeer(&can, devices[size_t(q)], c->device) == cudaSuccess && can;
                    if (!ok) break;
                    cudaMemAccessDesc d;
                    std::memset(&d, 0, sizeof d);
                    d.location.type = cudaMemLocationTypeDevice;
                    d.location.id = devices[size_t(q)];
                    d.flags = cudaMemAccessFlagsProtReadWrite;
                    ok = cudaMemPoolSetAccess(pool, &d, 1) == cudaSuccess;
                }
                cudaGetLastError();
                c->pool_shared = ok && env_int("EIGSOL_B200_GROUP_POOL", 1) != 0;
            }
            comms[size_t(r)] = std::move(c);
        });
    }
    // process-wide, never destroyed (worker threads must not be joined during static destruction)
    static GroupWorkers& get(int nranks, int first_device) {
        static std::mutex mu;
        static std::map<std::pair<int, int>, GroupWorkers*>* pools = new std::map<std::pair<int, int>, GroupWorkers*>();
        std::lock_guard<std::mutex> lk(mu);
        GroupWorkers*& w = (*pools)[{nranks, first_device}];
        if (!w) w = new GroupWorkers(nranks, first_device);
        return *w;
    }
    void worker(int r) {
        cudaSetDevice(devices[size_t(r)]);
        uint64_t seen = 0;
        for (;;) {
            std::function<void(int)> fn;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv_work.wait(lk, [&] { return job_id != seen; });
                seen = job_id;
                fn = job;
            }
            std::exception_ptr err;
            try {
                fn(r);
            } catch (...) {
                err = std::current_exception();
                group.fail();  // peers blocked in a barrier leave it with an exception instead of waiting forever
            }
            {
                std::lock_guard<std::mutex> lk(mu);
                if (err) {
                    errors[size_t(r)] = err;
                    error_seq[size_t(r)] = ++seq;
                }
                if (--pending == 0) cv_done.notify_all();
            }
        }
    }
    // runs fn(rank) on every rank's thread; rethrows the FIRST failure (the later ones are its echoes)
    void run_all(std::function<void(int)> fn) {
        std::lock_guard<std::mutex> one_job(job_mu);
        std::unique_lock<std::mutex> lk(mu);
        for (auto& e : errors) e = nullptr;
        seq = 0;
        job = std::move(fn);
        pending = P;
        ++job_id;
        cv_work.notify_all();
        cv_done.wait(lk, [&] { return pending == 0; });
        int first = -1;
        for (int r = 0; r < P; ++r)
            if (errors[size_t(r)] && (first < 0 || error_seq[size_t(r)] < error_seq[size_t(first)])) first = r;
        if (first >= 0) {
            group.reset();
            std::rethrow_exception(errors[size_t(first)]);
        }
    }
};

struct LocalSolver {
    const int P;
    GroupWorkers& workers;
    std::vector<std::unique_ptr<Comm>>& comms;  // the workers' (shared by every handle of the group)
    std::vector<std::unique_ptr<MatrixImpl>> ranks;

    LocalSolver(int nranks, int first_device)
        : P(nranks), workers(GroupWorkers::get(nranks, first_device)), comms(workers.comms) {
        ranks.resize(size_t(P));
        run_all([this](int r) { ranks[size_t(r)].reset(new MatrixImpl); });
    }
    ~LocalSolver() {
        try {
            run_all([this](int r) { ranks[size_t(r)].reset(); });  // every rank frees its own device memory
        } catch (...) {
        }
    }
    void run_all(std::function<void(int)> fn) { workers.run_all(std::move(fn)); }
    MatrixImpl& r0() { return *ranks[0]; }
};

// GPUs a comm == NULL matrix is sharded over: EIGSOL_B200_NGPU (default 1), at most the visible devices, at most 8,
// and no more than one GPU per EIGSOL_B200_MIN_ROWS_PER_GPU rows (default 1024): tiny matrices stay on one GPU.
static int group_size_for(int64_t n) {
    int want = env_int("EIGSOL_B200_NGPU", 1);
    if (want <= 1) return 1;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess) {
        cudaGetLastError();
        return 1;
    }
    want = std::min({want, ndev, kMaxPeers});
    const int64_t per = std::max(1, env_int("EIGSOL_B200_MIN_ROWS_PER_GPU", 1024));
    want = int(std::min<int64_t>(want, std::max<int64_t>(1, n / per)));
    return std::max(want, 1);
}

}  // namespace eigsol_b200

#include "block_host.cuh"

// =================================================================================================
// C ABI
// =================================================================================================
using namespace eigsol_b200;

// One handle = one rank (a single GPU, or one rank of a process-per-GPU group) or a whole single-process group.
struct eigsol_b200_matrix {
    std::unique_ptr<MatrixImpl> one;
    std::unique_ptr<LocalSolver> group;
    std::mutex mu;  // calls on ONE handle are serialised; different handles (e.g. the ranks of a thread-per-GPU
                    // group) run concurrently
    MatrixImpl& lead() { return group ? group->r0() : *one; }
    const MatrixImpl& lead() const { return group ? *group->ranks[0] : *one; }
    // runs f(rank impl) on the handle's rank(s)
    template <class F> void each(F&& f) {
        if (group) {
            group->run_all([&](int r) {
                AllocScope sc(group->ranks[size_t(r)]->stream);
                f(*group->ranks[size_t(r)], r);
            });
        } else {
            AllocScope sc(one->stream);
            f(*one, 0);
        }
    }
};
struct eigsol_b200_comm {
    Comm impl;
};

namespace {
template <class F> int guarded(F&& f) {
    try {
        g_last_error.clear();
        return f();
    } catch (const DeviceError& e) {
        g_last_error = e.what();
        return EIGSOL_B200_ERR_DEVICE;
    } catch (const StateError& e) {
        g_last_error = e.what();
        return EIGSOL_B200_ERR_STATE;
    } catch (const IoError& e) {
        g_last_error = e.what();
        return EIGSOL_B200_ERR_IO;
    } catch (const PlanError& e) {
        g_last_error = std::string("eigsol_b200: invalid argument: ") + e.what();
        return EIGSOL_B200_ERR_INVALID;
    } catch (const std::bad_alloc&) {
        g_last_error = "eigsol_b200: out of host memory";
        return EIGSOL_B200_ERR_INVALID;
    } catch (const std::exception& e) {
        g_last_error = std::string("powerMethod: device error: ") + e.what();
        return EIGSOL_B200_ERR_DEVICE;
    }
}

int fail(int code, const char* msg) {
    g_last_error = msg;
    return code;
}

// power_method.hpp:52-58, same order, same messages
int check_shape(int64_t rows, int64_t cols) {
    if (rows < 0 || cols < 0) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: negative dimension");
    if (rows != cols) return fail(EIGSOL_B200_ERR_NOT_SQUARE, "powerMethod: matrix must be square");
    if (rows == 0) return fail(EIGSOL_B200_ERR_ZERO_SIZE, "powerMethod: matrix has zero size");
    return EIGSOL_B200_OK;
}
int check_dtype(int dtype) {
    if (dtype != EIGSOL_B200_F64 && dtype != EIGSOL_B200_C128)
        return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: dtype must be 0 (f64) or 1 (c128)");
    return EIGSOL_B200_OK;
}

// comm == NULL and EIGSOL_B200_NGPU > 1: the handle becomes a single-process group over the first P devices
// (counted from the calling thread's current device)
std::unique_ptr<LocalSolver> make_group(int64_t n) {
    require_device();
    const int P = group_size_for(n);
    if (P <= 1) return nullptr;
    int dev0 = 0, ndev = 0;
    EIGSOL_CUDA(cudaGetDevice(&dev0));
    EIGSOL_CUDA(cudaGetDeviceCount(&ndev));
    if (dev0 + P > ndev) dev0 = 0;
    return std::unique_ptr<LocalSolver>(new LocalSolver(P, dev0));
}
}  // namespace

extern "C" {

const char* eigsol_b200_last_error(void) { return g_last_error.c_str(); }

int eigsol_b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

const char* eigsol_b200_version(void) { return "eigsol_b200 0.2 (sm_100a)"; }

int eigsol_b200_matrix_create_csc(int dtype, int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                                  const void* values, eigsol_b200_comm* comm, eigsol_b200_matrix** out) {
    if (!out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null output handle");
    *out = nullptr;
    if (int e = check_dtype(dtype)) return e;
    if (int e = check_shape(rows, cols)) return e;
    return guarded([&] {
        if (!outer || (outer[cols] > 0 && (!inner || !values))) throw PlanError("null CSC array");
        std::unique_ptr<eigsol_b200_matrix> h(new eigsol_b200_matrix);
        Trace tg;
        if (!comm) h->group = make_group(rows);
        if (h->group) {
            tg.mark_async("group: workers ready");
            // column-slab boundaries from the column prefix `outer` (nnz-balanced), reused as the row partition
            // (the monotonicity of outer is checked on the devices, slab by slab: a host loop over 16.7M columns costs
            // 10 ms of a 55 ms call)
            if (outer[0] != 0) throw PlanError("CSC outer index must start at 0 (compressed storage expected)");
            if (outer[cols] < 0) throw PlanError("CSC outer index not monotone");
            const std::vector<int64_t> rb = partition_rows(cols, outer, h->group->P);
            const void* pre = g_preload_x0;  // (thread-local of the calling thread: hand it to the rank threads)
            h->each([&](MatrixImpl& m, int r) {
                Trace tr;
                m.init_common(dtype, h->group->comms[size_t(r)].get());
                tr.mark_async("group rank: streams + state");
                m.n = rows;
                m.row_begin = rb;
                m.preload_x0 = pre;
                m.init_sparse_group_csc(rows, cols, outer, inner, values);
            });
            tg.mark_async("group: ranks initialised");
            *out = h.release();
            ++g_matrices_created;
            return int(EIGSOL_B200_OK);
        }
        Trace tr0;
        AllocScope alloc_scope(nullptr);  // init_common makes the handle's stream current for small allocations
        h->one.reset(new MatrixImpl);
        MatrixImpl& m = *h->one;
        m.init_common(dtype, comm ? &comm->impl : nullptr);
        m.n = rows;
        tr0.mark_async("create: handle + streams");
        if (m.nranks == 1 && env_int("EIGSOL_B200_DEVICE_INGEST", 1)) {
            m.preload_x0 = g_preload_x0;
            m.init_sparse_device_csc(rows, cols, outer, inner, values);
            *out = h.release();
            ++g_matrices_created;
            return int(EIGSOL_B200_OK);
        }
        CsrHost g = csc_to_csr(dtype, rows, cols, outer, inner, values);
        m.row_begin = partition_rows(rows, g.rowptr.data(), m.nranks);
        if (m.nranks == 1) {
            m.init_sparse_local(std::move(g));
        } else {
            const int64_t lo = m.row_begin[m.rank], hi = m.row_begin[m.rank + 1];
            CsrHost loc;
            loc.dtype = dtype;
            loc.rows = hi - lo;
            loc.cols = cols;
            loc.rowptr.resize(size_t(hi - lo) + 1);
            const int32_t p0 = g.rowptr[lo];
            for (int64_t i = lo; i <= hi; ++i) loc.rowptr[i - lo] = g.rowptr[i] - p0;
            loc.colidx.assign(g.colidx.begin() + p0, g.colidx.begin() + g.rowptr[hi]);
            loc.vals.assign(g.vals.begin() + int64_t(p0) * g.spd(), g.vals.begin() + int64_t(g.rowptr[hi]) * g.spd());
            g = CsrHost();
            m.init_sparse_local(std::move(loc));
        }
        *out = h.release();
        ++g_matrices_created;
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_matrix_create_csr_rows(int dtype, int64_t n, int64_t row_begin, int64_t row_count, const int32_t* rowptr,
                                       const int32_t* colidx, const void* values, eigsol_b200_comm* comm,
                                       eigsol_b200_matrix** out) {
    if (!out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null output handle");
    *out = nullptr;
    if (int e = check_dtype(dtype)) return e;
    if (int e = check_shape(n, n)) return e;
    return guarded([&] {
        validate_csr_rows(n, row_count, rowptr, colidx);
        if (rowptr[row_count] > 0 && !values) throw PlanError("null values array");
        std::unique_ptr<eigsol_b200_matrix> h(new eigsol_b200_matrix);
        if (!comm) {
            if (row_begin != 0 || row_count != n) throw PlanError("without a comm the row block must be the whole matrix");
            h->group = make_group(n);
        }
        if (h->group) {
            // nnz-balanced row blocks; every rank uploads its own block straight from the caller's arrays
            const std::vector<int64_t> rb = partition_rows(n, rowptr, h->group->P);
            h->each([&](MatrixImpl& m, int r) {
                m.init_common(dtype, h->group->comms[size_t(r)].get());
                m.n = n;
                m.row_begin = rb;
                m.init_sparse_device_rows(rowptr + rb[size_t(r)], colidx, values);
            });
            *out = h.release();
            ++g_matrices_created;
            return int(EIGSOL_B200_OK);
        }
        AllocScope alloc_scope(nullptr);  // init_common makes the handle's stream current for small allocations
        h->one.reset(new MatrixImpl);
        MatrixImpl& m = *h->one;
        m.init_common(dtype, comm ? &comm->impl : nullptr);
        m.n = n;
        if (m.nranks == 1) {
            m.row_begin = {0, n};
        } else {
            const std::vector<int64_t> begins = comm_allgather_i64(m.comm, row_begin);
            m.row_begin = begins;
            m.row_begin.push_back(n);
            for (int r = 0; r < m.nranks; ++r)
                if (m.row_begin[r] > m.row_begin[r + 1]) throw PlanError("row blocks of the ranks are not ascending");
            if (m.row_begin[0] != 0 || m.row_begin[m.rank + 1] - m.row_begin[m.rank] != row_count)
                throw PlanError("row blocks of the ranks do not tile [0, n)");
        }
        if (env_int("EIGSOL_B200_DEVICE_INGEST", 1)) {
            m.init_sparse_device_rows(rowptr, colidx, values);
            *out = h.release();
            ++g_matrices_created;
            return int(EIGSOL_B200_OK);
        }
        CsrHost loc;
        loc.dtype = dtype;
        loc.rows = row_count;
        loc.cols = n;
        loc.rowptr.assign(rowptr, rowptr + row_count + 1);
        const int64_t nnz = rowptr[row_count];
        loc.colidx.assign(colidx, colidx + nnz);
        loc.vals.assign(static_cast<const double*>(values), static_cast<const double*>(values) + nnz * loc.spd());
        m.init_sparse_local(std::move(loc));
        *out = h.release();
        ++g_matrices_created;
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_matrix_create_dense(int dtype, int64_t rows, int64_t cols, const void* a, int64_t lda,
                                    eigsol_b200_comm* comm, eigsol_b200_matrix** out) {
    if (!out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null output handle");
    *out = nullptr;
    if (int e = check_dtype(dtype)) return e;
    if (int e = check_shape(rows, cols)) return e;
    if (!a || lda < rows) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: bad dense matrix pointer / lda");
    if (rows > INT32_MAX / 2) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: dense order too large");
    return guarded([&] {
        std::unique_ptr<eigsol_b200_matrix> h(new eigsol_b200_matrix);
        if (!comm) h->group = make_group(rows);
        if (!h->group) h->one.reset(new MatrixImpl);
        h->each([&](MatrixImpl& m, int r) {
            m.init_common(dtype, h->group ? h->group->comms[size_t(r)].get() : (comm ? &comm->impl : nullptr));
            m.init_dense(rows, a, lda);
        });
        *out = h.release();
        ++g_matrices_created;
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_matrix_destroy(eigsol_b200_matrix* m) {
    if (!m) return EIGSOL_B200_OK;
    return guarded([&] {
        Trace tr;
        if (m->one) cudaSetDevice(m->one->device);
        delete m;
        tr.mark_async("destroy");
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_matrix_get_info(const eigsol_b200_matrix* m, eigsol_b200_matrix_info* info) {
    if (!m || !info) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle / info");
    m->lead().fill_info(info);
    if (m->group) {
        // a single-process group reports rank 0's launch geometry and the LARGEST per-rank byte counts (the rank that
        // bounds a trip)
        for (const auto& r : m->group->ranks) {
            eigsol_b200_matrix_info ri;
            r->fill_info(&ri);
            info->alg_bytes = std::max(info->alg_bytes, ri.alg_bytes);
            info->stored_bytes = std::max(info->stored_bytes, ri.stored_bytes);
            info->halo = std::max(info->halo, ri.halo);
        }
    }
    return EIGSOL_B200_OK;
}

int eigsol_b200_matrix_set_kernel(eigsol_b200_matrix* m, int kernel, int lanes) {
    if (!m) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        m->each([&](MatrixImpl& r, int) {
            EIGSOL_CUDA(cudaSetDevice(r.device));
            r.begun = false;
            r.set_kernel(kernel, lanes);
        });
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_power_run(eigsol_b200_matrix* m, const void* x0, int max_iterations, double tolerance, void* x_out,
                          double lambda_out[2], int* iterations_out, int* converged_out) {
    if (!m) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        m->each([&](MatrixImpl& r, int k) {
            // every rank of a group writes its own rows of x_out; the scalars come from rank 0 (identical everywhere)
            r.run(x0, max_iterations, tolerance, x_out, k == 0 ? lambda_out : nullptr, k == 0 ? iterations_out : nullptr,
                  k == 0 ? converged_out : nullptr);
        });
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_power_begin(eigsol_b200_matrix* m, const void* x0, int max_iterations, double tolerance) {
    if (!m) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        m->each([&](MatrixImpl& r, int) { r.begin(x0, max_iterations, tolerance); });
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_power_step(eigsol_b200_matrix* m, int trips, float* ms_out) {
    if (!m) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        float ms[kMaxPeers] = {};
        m->each([&](MatrixImpl& r, int k) { r.step(trips, &ms[k]); });
        if (ms_out) *ms_out = *std::max_element(ms, ms + kMaxPeers);  // a group: the slowest rank
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_power_end(eigsol_b200_matrix* m, void* x_out, double lambda_out[2], int* iterations_out,
                          int* converged_out) {
    if (!m) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null handle");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        m->each([&](MatrixImpl& r, int k) {
            r.end(x_out, k == 0 ? lambda_out : nullptr, k == 0 ? iterations_out : nullptr, k == 0 ? converged_out : nullptr);
        });
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_matvec(eigsol_b200_matrix* m, const void* x, void* y) {
    if (!m || !x || !y) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        m->each([&](MatrixImpl& r, int) { r.matvec(x, y); });
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_rayleigh(eigsol_b200_matrix* m, const void* x, double dot_out[2], double* norm2_out) {
    if (!m || !x) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        double o[kMaxPeers][4] = {};
        m->each([&](MatrixImpl& r, int k) { r.rayleigh(x, o[k]); });
        if (dot_out) {
            dot_out[0] = o[0][0];
            dot_out[1] = o[0][1];
        }
        if (norm2_out) *norm2_out = o[0][2];
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_block_power_run(eigsol_b200_matrix* m, int block, int nev, const void* x0, int max_iterations,
                                double tolerance, void* x_out, double* lambda_out, int* iterations_out,
                                int* converged_out, int* breakdown_out, float* ms_out) {
    if (!m || !x0 || !lambda_out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        BlockResult rr[kMaxPeers];
        double lam[kMaxPeers][2 * kBlockMax];
        m->each([&](MatrixImpl& impl, int k) {
            // every rank of a group writes its own rows of x_out; the scalars come from rank 0 (identical everywhere)
            block_power_run(impl, block, nev, x0, max_iterations, tolerance, x_out, lam[k], rr[k]);
        });
        const BlockResult& r = rr[0];
        for (int i = 0; i < 2 * block; ++i) lambda_out[i] = lam[0][i];
        if (iterations_out) *iterations_out = r.iterations;
        if (converged_out) *converged_out = r.converged;
        if (breakdown_out) *breakdown_out = r.breakdown;
        if (ms_out) *ms_out = r.ms;
        return int(EIGSOL_B200_OK);
    });
}

int64_t eigsol_b200_matrices_created(void) { return g_matrices_created.load(); }

int64_t eigsol_b200_matrix_launch_count(const eigsol_b200_matrix* m) {
    if (!m) return 0;
    if (!m->group) return m->one->launches;
    int64_t total = 0;
    for (const auto& r : m->group->ranks) total += r->launches;
    return total;
}

// EIGSOL_B200_PEER_STATS=1: where the multi-GPU trips of this rank waited since the last call (kernels.cuh, PeerArgs::stats)
int eigsol_b200_matrix_peer_stats(eigsol_b200_matrix* m, uint64_t out[8]) {
    if (!m || !out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::lock_guard<std::mutex> lock(m->mu);
    return guarded([&] {
        std::memset(out, 0, sizeof(uint64_t) * 8);
        MatrixImpl& r = m->lead();
        if (!r.d_stats.p) return int(EIGSOL_B200_OK);
        EIGSOL_CUDA(cudaSetDevice(r.device));
        EIGSOL_CUDA(cudaStreamSynchronize(r.stream));
        EIGSOL_CUDA(cudaMemcpy(out, r.d_stats.p, sizeof(uint64_t) * kPeerStats, cudaMemcpyDeviceToHost));
        EIGSOL_CUDA(cudaMemset(r.d_stats.p, 0, sizeof(uint64_t) * kPeerStats));
        return int(EIGSOL_B200_OK);
    });
}

}  // extern "C"

// ---- measurement: HBM ceiling of a read-dominated stream (probe.cuh) -------------------------------------
extern "C" int eigsol_b200_hbm_probe(int64_t plane_bytes, int planes, int reps, double* best_gbps, double* mean_gbps) {
    if (plane_bytes < 4096 || reps < 1 || (planes != 1 && planes != 4 && planes != 8 && planes != 12))
        return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: hbm probe takes 1, 4, 8 or 12 read planes");
    return guarded([&] {
        require_device();
        const int64_t n4 = plane_bytes / 32;
        DevBuf in, out;
        in.alloc(size_t(n4) * 32 * size_t(planes));
        out.alloc(size_t(n4) * 32);
        int dev = 0, sms = 148;
        EIGSOL_CUDA(cudaGetDevice(&dev));
        EIGSOL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        cudaStream_t st;
        EIGSOL_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        cudaEvent_t e0, e1;
        EIGSOL_CUDA(cudaEventCreate(&e0));
        EIGSOL_CUDA(cudaEventCreate(&e1));
        const int grid = sms * std::max(1, env_int("EIGSOL_B200_PROBE_CTAS_PER_SM", 4));
        auto launch = [&] {
            switch (planes) {
                case 1: read_mix_kernel<1><<<grid, 256, 0, st>>>(in.as<double>(), n4, out.as<double>()); break;
                case 4: read_mix_kernel<4><<<grid, 256, 0, st>>>(in.as<double>(), n4, out.as<double>()); break;
                case 8: read_mix_kernel<8><<<grid, 256, 0, st>>>(in.as<double>(), n4, out.as<double>()); break;
                default: read_mix_kernel<12><<<grid, 256, 0, st>>>(in.as<double>(), n4, out.as<double>()); break;
            }
        };
        for (int w = 0; w < 3; ++w) launch();
        double best = 0.0, sum = 0.0;
        const double bytes = double(n4) * 32.0 * double(planes + 1);
        for (int r = 0; r < reps; ++r) {
            EIGSOL_CUDA(cudaEventRecord(e0, st));
            launch();
            EIGSOL_CUDA(cudaEventRecord(e1, st));
            EIGSOL_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            EIGSOL_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            const double g = bytes / (double(ms) * 1e-3) / 1e9;
            best = std::max(best, g);
            sum += g;
        }
        EIGSOL_CUDA(cudaGetLastError());
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaStreamDestroy(st);
        if (best_gbps) *best_gbps = best;
        if (mean_gbps) *mean_gbps = sum / reps;
        return int(EIGSOL_B200_OK);
    });
}

// ---- one-shot entry points ------------------------------------------------------------------------
extern "C" {

// The caller's result vector is usually fresh memory (np.empty, Eigen's Vector(n)): the download's first touch of its
// pages costs more than the PCIe copy (134 MB: 5-9 ms of page faults, worse with several ranks writing at once).  The
// one-shot calls touch the pages -- writing back what is there -- on a few helper threads while the matrix uploads.
struct Prefault {
    std::vector<std::thread> pool;
    void start(void* p, size_t bytes) {
        if (!p || bytes < (size_t(32) << 20) || !env_int("EIGSOL_B200_PREFAULT", 1)) return;
        const int nt = 4;
        char* base = static_cast<char*>(p);
        for (int t = 0; t < nt; ++t)
            pool.emplace_back([=] {
                const size_t lo = bytes * size_t(t) / nt, hi = bytes * size_t(t + 1) / nt;
                for (size_t off = lo; off < hi; off += 4096) {
                    volatile char* q = base + off;
                    *q = *q;
                }
            });
    }
    void join() {
        for (auto& t : pool) t.join();
        pool.clear();
    }
    ~Prefault() { join(); }
};

// Solve and, if keep_out != NULL, hand the device-resident matrix to the caller instead of destroying it: the FIRST
// solve on a matrix and its upload are one call (the start vector rides on the matrix upload), later solves use
// eigsol_b200_power_run on *keep_out.
int eigsol_b200_power_csc_keep(int dtype, int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                               const void* values, const void* x0, int max_iterations, double tolerance, void* x_out,
                               double lambda_out[2], int* iterations_out, int* converged_out, eigsol_b200_matrix** keep_out) {
    if (keep_out) *keep_out = nullptr;
    eigsol_b200_matrix* m = nullptr;
    Prefault pf;
    if (rows > 0 && (dtype == EIGSOL_B200_F64 || dtype == EIGSOL_B200_C128)) pf.start(x_out, size_t(rows) * (dtype ? 16 : 8));
    g_preload_x0 = x0;  // the start vector goes up together with the matrix (single GPU)
    int e = eigsol_b200_matrix_create_csc(dtype, rows, cols, outer, inner, values, nullptr, &m);
    g_preload_x0 = nullptr;
    pf.join();
    if (e) return e;
    e = eigsol_b200_power_run(m, x0, max_iterations, tolerance, x_out, lambda_out, iterations_out, converged_out);
    if (!e && keep_out) {
        *keep_out = m;
        return e;
    }
    const std::string keep = g_last_error;
    eigsol_b200_matrix_destroy(m);
    if (e) g_last_error = keep;
    return e;
}

int eigsol_b200_power_dense_keep(int dtype, int64_t rows, int64_t cols, const void* a, int64_t lda, const void* x0,
                                 int max_iterations, double tolerance, void* x_out, double lambda_out[2],
                                 int* iterations_out, int* converged_out, eigsol_b200_matrix** keep_out) {
    if (keep_out) *keep_out = nullptr;
    eigsol_b200_matrix* m = nullptr;
    int e = eigsol_b200_matrix_create_dense(dtype, rows, cols, a, lda, nullptr, &m);
    if (e) return e;
    e = eigsol_b200_power_run(m, x0, max_iterations, tolerance, x_out, lambda_out, iterations_out, converged_out);
    if (!e && keep_out) {
        *keep_out = m;
        return e;
    }
    const std::string keep = g_last_error;
    eigsol_b200_matrix_destroy(m);
    if (e) g_last_error = keep;
    return e;
}

int eigsol_b200_power_csc(int dtype, int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                          const void* values, const void* x0, int max_iterations, double tolerance, void* x_out,
                          double lambda_out[2], int* iterations_out, int* converged_out) {
    return eigsol_b200_power_csc_keep(dtype, rows, cols, outer, inner, values, x0, max_iterations, tolerance, x_out,
                                      lambda_out, iterations_out, converged_out, nullptr);
}

int eigsol_b200_power_dense(int dtype, int64_t rows, int64_t cols, const void* a, int64_t lda, const void* x0,
                            int max_iterations, double tolerance, void* x_out, double lambda_out[2],
                            int* iterations_out, int* converged_out) {
    return eigsol_b200_power_dense_keep(dtype, rows, cols, a, lda, x0, max_iterations, tolerance, x_out, lambda_out,
                                        iterations_out, converged_out, nullptr);
}

// ---- multi-GPU group ------------------------------------------------------------------------------
int eigsol_b200_comm_unique_id(void* id128) {
    if (!id128) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null id buffer");
    return guarded([&] {
        ncclUniqueId id;
        EIGSOL_NCCL(nccl().GetUniqueId(&id));
        std::memcpy(id128, &id, sizeof id);
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_comm_create(int rank, int nranks, const void* id128, int device, eigsol_b200_comm** out) {
    if (!out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null output handle");
    *out = nullptr;
    if (nranks < 1 || rank < 0 || rank >= nranks || (nranks > 1 && !id128))
        return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: bad rank / nranks / id");
    return guarded([&] {
        require_device();
        std::unique_ptr<eigsol_b200_comm> h(new eigsol_b200_comm);
        Comm& c = h->impl;
        c.rank = rank;
        c.nranks = nranks;
        c.device = device;
        EIGSOL_CUDA(cudaSetDevice(device));
        EIGSOL_CUDA(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
        c.scratch.alloc_plain(sizeof(int64_t) * size_t(4 * nranks + 16));  // NCCL buffer
        if (nranks > 1) {
            ncclUniqueId id;
            std::memcpy(&id, id128, sizeof id);
            EIGSOL_NCCL(nccl().CommInitRank(&c.nccl_comm, nranks, id, rank));
        }
        *out = h.release();
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_comm_destroy(eigsol_b200_comm* comm) {
    if (!comm) return EIGSOL_B200_OK;
    return guarded([&] {
        cudaSetDevice(comm->impl.device);
        delete comm;
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_comm_max_float(eigsol_b200_comm* comm, float* inout) {
    if (!comm || !inout) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    return guarded([&] {
        Comm& c = comm->impl;
        if (c.nranks == 1) return int(EIGSOL_B200_OK);
        EIGSOL_CUDA(cudaSetDevice(c.device));
        float* d = c.scratch.as<float>();
        EIGSOL_CUDA(cudaMemcpyAsync(d, inout, sizeof(float), cudaMemcpyHostToDevice, c.stream));
        EIGSOL_NCCL(nccl().AllReduce(d, d, 1, ncclFloat, ncclMax, c.nccl_comm, c.stream));
        EIGSOL_CUDA(cudaMemcpyAsync(inout, d, sizeof(float), cudaMemcpyDeviceToHost, c.stream));
        EIGSOL_CUDA(cudaStreamSynchronize(c.stream));
        return int(EIGSOL_B200_OK);
    });
}

// ---- matrix files (no GPU) ----------------------------------------------------------------------------
namespace {
struct FileOwner {
    std::unique_ptr<FileMatrixHost> parsed;
    std::unique_ptr<MappedMatrix> mapped;
};
}  // namespace

int eigsol_b200_read_matrix_text(const char* path, int dtype, eigsol_b200_file_matrix* out) {
    if (!path || !out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::memset(out, 0, sizeof *out);
    if (int e = check_dtype(dtype)) return e;
    return guarded([&] {
        std::unique_ptr<FileOwner> o(new FileOwner);
        o->parsed.reset(new FileMatrixHost(read_matrix_text(path, dtype)));
        const FileMatrixHost& m = *o->parsed;
        out->dtype = m.dtype;
        out->is_dense = m.dense ? 1 : 0;
        out->rows = m.rows;
        out->cols = m.cols;
        out->nnz = m.dense ? m.rows * m.cols : m.nnz;
        out->outer = m.dense ? nullptr : m.outer.data();
        out->inner = m.dense ? nullptr : m.inner.data();
        out->values = m.values.data();
        out->owner = o.release();
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_read_matrix_bin(const char* path, eigsol_b200_file_matrix* out) {
    if (!path || !out) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    std::memset(out, 0, sizeof *out);
    return guarded([&] {
        std::unique_ptr<FileOwner> o(new FileOwner);
        o->mapped.reset(new MappedMatrix(path));
        const MappedMatrix& m = *o->mapped;
        out->dtype = m.h.dtype;
        out->is_dense = m.h.dense;
        out->rows = m.h.rows;
        out->cols = m.h.cols;
        out->nnz = m.h.nnz;
        out->outer = m.outer;
        out->inner = m.inner;
        out->values = m.values;
        out->owner = o.release();
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_write_matrix_bin(const char* path, int dtype, int is_dense, int64_t rows, int64_t cols, int64_t nnz,
                                 const int32_t* outer, const int32_t* inner, const void* values) {
    if (!path || !values || (!is_dense && (!outer || !inner)))
        return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: null pointer");
    if (int e = check_dtype(dtype)) return e;
    if (rows <= 0 || cols <= 0 || nnz < 0) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: bad dimensions");
    return guarded([&] {
        write_matrix_bin(path, dtype, is_dense != 0, rows, cols, nnz, outer, inner, values);
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_file_matrix_release(eigsol_b200_file_matrix* m) {
    if (!m) return EIGSOL_B200_OK;
    delete static_cast<FileOwner*>(m->owner);
    std::memset(m, 0, sizeof *m);
    return EIGSOL_B200_OK;
}

// ---- host-side planning (no GPU) ------------------------------------------------------------------
int eigsol_b200_csc_to_csr(int dtype, int64_t rows, int64_t cols, const int32_t* outer, const int32_t* inner,
                           const void* values, int32_t* rowptr, int32_t* colidx, void* values_out) {
    if (int e = check_dtype(dtype)) return e;
    return guarded([&] {
        validate_csc(rows, cols, outer, inner);
        if (!rowptr || (outer[cols] > 0 && (!colidx || !values || !values_out))) throw PlanError("null array");
        csc_to_csr_raw(dtype, rows, cols, outer, inner, static_cast<const double*>(values), rowptr, colidx,
                       static_cast<double*>(values_out));
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_partition_rows(int64_t rows, const int32_t* rowptr, int nranks, int64_t* row_begin) {
    return guarded([&] {
        if (rows < 0 || !rowptr || !row_begin) throw PlanError("null array / negative size");
        const std::vector<int64_t> rb = partition_rows(rows, rowptr, nranks);
        std::copy(rb.begin(), rb.end(), row_begin);
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_halo_plan(int64_t row_count, const int32_t* rowptr, const int32_t* colidx, int nranks,
                          const int64_t* row_begin, int rank, int64_t* halo_count, int32_t* halo_cols,
                          int64_t* peer_count) {
    return guarded([&] {
        if (row_count < 0 || !rowptr || !row_begin || !halo_count || nranks < 1 || rank < 0 || rank >= nranks)
            throw PlanError("bad halo plan arguments");
        const std::vector<int64_t> rb(row_begin, row_begin + nranks + 1);
        const HaloPlan h = halo_plan(row_count, rowptr, colidx, rb, rank);
        *halo_count = int64_t(h.cols.size());
        if (halo_cols) std::copy(h.cols.begin(), h.cols.end(), halo_cols);
        if (peer_count) std::copy(h.peer_count.begin(), h.peer_count.end(), peer_count);
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_choose_kernel_csr(int dtype, int64_t rows, const int32_t* rowptr, const int32_t* colidx, int* kernel_out,
                                  int* lanes_out) {
    if (int e = check_dtype(dtype)) return e;
    return guarded([&] {
        if (rows < 0 || !rowptr || !kernel_out || !lanes_out) throw PlanError("null array / negative size");
        const KernelChoice c = choose_kernel(dtype, row_stats(rows, rowptr, colidx, 0));
        *kernel_out = c.kernel;
        *lanes_out = c.lanes;
        return int(EIGSOL_B200_OK);
    });
}

int eigsol_b200_choose_kernel(int dtype, int64_t rows, const int32_t* rowptr, int* kernel_out, int* lanes_out) {
    if (int e = check_dtype(dtype)) return e;
    return guarded([&] {
        if (rows < 0 || !rowptr || !kernel_out || !lanes_out) throw PlanError("null array / negative size");
        const KernelChoice c = choose_kernel(dtype, row_stats(rows, rowptr));
        *kernel_out = c.kernel;
        *lanes_out = c.lanes;
        return int(EIGSOL_B200_OK);
    });
}

}  // extern "C"

// ---- x0 for callers without Eigen -----------------------------------------------------------------
namespace {
inline double eigen_style_random_real() {
    // RandomImpl.h:66-69: fill 52 mantissa bits from 31-bit rand() words, low word first;
    // :89-99: OR into the bit pattern of 2.0 (-> [2,4)) and subtract 3 (-> [-1,1)).
    uint64_t bits = 0;
    for (int shift = 0; shift < 52; shift += 31) bits |= uint64_t(std::rand()) << shift;
    bits &= (~uint64_t(0)) >> 12;
    const double two = 2.0;
    uint64_t tb;
    std::memcpy(&tb, &two, sizeof tb);
    bits |= tb;
    double r;
    std::memcpy(&r, &bits, sizeof r);
    return r - 3.0;
}
}  // namespace

extern "C" int eigsol_b200_random_unit(int dtype, int64_t n, int seed, void* x_out) {
    if (int e = check_dtype(dtype)) return e;
    if (n < 0 || (n > 0 && !x_out)) return fail(EIGSOL_B200_ERR_INVALID, "eigsol_b200: invalid argument: bad vector");
    if (seed >= 0) std::srand(unsigned(seed));
    double* x = static_cast<double*>(x_out);
    if (dtype == EIGSOL_B200_F64) {
        for (int64_t i = 0; i < n; ++i) x[i] = eigen_style_random_real();
    } else {
        for (int64_t i = 0; i < n; ++i) {
            const double im = eigen_style_random_real();
            const double re = eigen_style_random_real();
            x[2 * i] = re;
            x[2 * i + 1] = im;
        }
    }
    const int64_t nd = dtype ? 2 * n : n;
    double z = 0.0;
    for (int64_t i = 0; i < nd; ++i) z += x[i] * x[i];
    if (z > 0.0) {
        const double nrm = std::sqrt(z);
        for (int64_t i = 0; i < nd; ++i) x[i] /= nrm;
    }
    return EIGSOL_B200_OK;
}
