// kernels.cuh -- hand-written sm_100a kernels of the power-iteration path.
//
// One launch == one power iteration's matvec with everything else fused into it:
//     x  = u / ||u||                 (reference: src/power_method/power_method.hpp:72,78)
//     z  = A x                       (:69 of the NEXT trip == the A*x temporary of :81, computed once)
//     d  = x^H z  (conjugated dot)   (:81, eigen/Eigen/src/Core/InnerProduct.h:255)
//     q' = ||z||^2                   (:72 of the next trip, eigen/Eigen/src/Core/Dot.h:61-77)
// and, on one GPU, the convergence test of :83-95 (src/core/tolerance.hpp:28-33) in the last CTA to
// finish.  u is the un-normalised output of the previous launch; x is never written to HBM, so an
// iteration reads A once, reads u once (gathers served from L1/L2), writes z once: exactly the
// algorithmic bytes of SURVEY.md 8(d).
//
// Bandwidth-bound integer/fp64 streaming work: no tensor cores.  What matters here is coalescing
// (128-bit loads of values / column indices), enough bytes in flight per SM, a persistent grid sized
// from the SM count, and keeping the x gathers in L1/L2 (matrix streams bypass L1 allocation).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eigsol_b200 {

// -------------------------------------------------------------------------------------------------
// Device-resident iteration state (one per matrix handle).  Mirrors the locals of powerMethodImpl
// (power_method.hpp:61-66): lambda, usedIters, converged, initialized.
// -------------------------------------------------------------------------------------------------
struct PowerState {
    double q[2];        // ||u||^2 of ping-pong buffer 0 / 1
    double lambda[2];   // current eigenvalue estimate (re, im)
    double tol;         // opts.tolerance
    double final_q;     // ||u||^2 of the buffer holding the result vector
    int max_trips;      // opts.maxIterations
    int launches;       // effective launches done; buffer (launches & 1) is the next input
    int trips;          // usedIters
    int initialized;    // power_method.hpp:66
    int stop;           // 0 running, 1 converged, 2 zero norm (:73-76), 3 maxIterations reached
    int converged;
    int final_buf;      // buffer whose normalisation is EigenResult::eigenvector
    unsigned ticket;    // CTA completion counter of the running launch
    int pending;        // MODE_PEER: launch `launches` has run but its global sums are not folded in yet
    int reserved;
};

enum { STOP_NONE = 0, STOP_CONVERGED = 1, STOP_ZERO_NORM = 2, STOP_MAX_TRIPS = 3,
       STOP_ERROR = 4 };  // a peer never signalled (MODE_PEER wait timed out)

// -------------------------------------------------------------------------------------------------
// Scalar traits: double and complex<double> (as double2 = interleaved re, im).
// -------------------------------------------------------------------------------------------------
template <class S> struct Sc;

template <> struct Sc<double> {
    static constexpr bool is_complex = false;
    static __device__ __forceinline__ double zero() { return 0.0; }
    static __device__ __forceinline__ void mac(double& acc, double a, double b) { acc = fma(a, b, acc); }
    static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
    static __device__ __forceinline__ void add(double& acc, double a) { acc += a; }
    static __device__ __forceinline__ double scale(double a, double s) { return a * s; }
    static __device__ __forceinline__ double div(double a, double s) { return a / s; }
    // re/im of conj(x)*z accumulated
    static __device__ __forceinline__ void dotc(double x, double z, double& re, double&) { re = fma(x, z, re); }
    static __device__ __forceinline__ double abs2(double z) { return z * z; }
    static __device__ __forceinline__ double ldg(const double* p) { return __ldg(p); }
    static __device__ __forceinline__ double ldcg(const double* p) { return __ldcg(p); }
    static __device__ __forceinline__ double shfl_down(double v, int d, int w) { return __shfl_down_sync(0xffffffffu, v, d, w); }
};

template <> struct Sc<double2> {
    static constexpr bool is_complex = true;
    static __device__ __forceinline__ double2 zero() { return make_double2(0.0, 0.0); }
    static __device__ __forceinline__ void mac(double2& acc, double2 a, double2 b) {
        acc.x = fma(a.x, b.x, acc.x);
        acc.x = fma(-a.y, b.y, acc.x);
        acc.y = fma(a.x, b.y, acc.y);
        acc.y = fma(a.y, b.x, acc.y);
    }
    static __device__ __forceinline__ double2 mul(double2 a, double2 b) {
        return make_double2(fma(a.x, b.x, -a.y * b.y), fma(a.x, b.y, a.y * b.x));
    }
    static __device__ __forceinline__ void add(double2& acc, double2 a) { acc.x += a.x; acc.y += a.y; }
    static __device__ __forceinline__ double2 scale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }
    static __device__ __forceinline__ double2 div(double2 a, double s) { return make_double2(a.x / s, a.y / s); }
    static __device__ __forceinline__ void dotc(double2 x, double2 z, double& re, double& im) {
        re = fma(x.x, z.x, re);
        re = fma(x.y, z.y, re);
        im = fma(x.x, z.y, im);
        im = fma(-x.y, z.x, im);
    }
    static __device__ __forceinline__ double abs2(double2 z) { return fma(z.x, z.x, z.y * z.y); }
    static __device__ __forceinline__ double2 ldg(const double2* p) { return __ldg(p); }
    static __device__ __forceinline__ double2 ldcg(const double2* p) { return __ldcg(p); }
    static __device__ __forceinline__ double2 shfl_down(double2 v, int d, int w) {
        return make_double2(__shfl_down_sync(0xffffffffu, v.x, d, w), __shfl_down_sync(0xffffffffu, v.y, d, w));
    }
};

// -------------------------------------------------------------------------------------------------
// Streaming loads for the matrix arrays: read-only path, no L1 allocation (L1 is kept for x gathers).
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ double2 ld_stream_f64x2(const double* p) {
    double2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f64 {%0, %1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ double ld_stream_f64(const double* p) {
    double r;
    asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ int2 ld_stream_s32x2(const int32_t* p) {
    int2 r;
    asm volatile("ld.global.nc.L1::no_allocate.v2.s32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
__device__ __forceinline__ int4 ld_stream_s32x4(const int32_t* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ int32_t ld_stream_s32(const int32_t* p) {
    int32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
template <class S> __device__ __forceinline__ S ld_stream_val(const S* p);
template <> __device__ __forceinline__ double ld_stream_val<double>(const double* p) { return ld_stream_f64(p); }
template <> __device__ __forceinline__ double2 ld_stream_val<double2>(const double2* p) {
    return ld_stream_f64x2(reinterpret_cast<const double*>(p));
}

// -------------------------------------------------------------------------------------------------
// Views passed by value to the kernels.
// -------------------------------------------------------------------------------------------------
template <class S> struct CsrView {
    int32_t rows;            // local rows
    const int32_t* rowptr;   // rows + 1
    const int32_t* colidx;   // nnz (+ padding), local extended numbering
    const S* vals;           // nnz (+ padding)
};

template <class S> struct SellView {
    int32_t rows;
    int32_t nslices;
    const int64_t* slice_off;  // nslices + 1 (entries)
    const int32_t* colidx;     // padded entries, -1 = padding
    const S* vals;
};

template <class S> struct DenseView {
    int32_t rows;       // local rows
    int32_t cols;       // all columns (extended numbering)
    int64_t lda;        // device leading dimension (even, >= rows)
    const S* a;         // column-major
    int32_t nseg;       // column segments per row tile
    int32_t ntiles;     // row tiles
    S* partial;         // [nseg][lda] partial sums
    unsigned* tile_ticket;  // [ntiles]
};

// How the grid-level finish of a launch is done.
enum { MODE_FUSED = 0,   // one GPU: the last CTA sums the partials and runs the convergence test
       MODE_SPLIT = 1,   // N GPUs over NCCL: partials only; reduce + all-reduce + finalize_kernel follow
       MODE_PEER = 2 };  // N GPUs over peer memory: ONE kernel per iteration does matvec + x exchange +
                         // scalar all-reduce (P2P stores over NVLink, flags with system-scope release/acquire)

constexpr int kMaxPeers = 8;

// Multi-GPU peer-memory wiring of one rank (MODE_PEER).  Every rank runs the same launch sequence in
// lock step: launch l reads buffer l&1 (own rows + halo) and writes buffer (l+1)&1.  Rows of z that a
// peer's rows reference are ALSO stored straight into that peer's halo region of ITS buffer (l+1)&1,
// the rank's three partial sums go to every peer's scal_in[l&1][rank], and finally flag_in[rank] on every
// peer is set to l+1 (release).  Launch l+1 starts by waiting for all its flag_in[q] >= l+1 (acquire).
struct PeerArgs {
    int nranks, rank;
    int nsend;                        // contiguous send ranges (rows stored remotely as they are produced)
    int send_start[kMaxPeers];        // first local row of the range
    int send_count[kMaxPeers];
    void* send_dst[kMaxPeers][2];     // remote address of the peer's halo segment, per buffer parity
    int nlist;                        // scattered send lists (copied by the last CTA)
    int list_off[kMaxPeers], list_count[kMaxPeers];
    void* list_dst[kMaxPeers][2];
    const int32_t* send_idx;          // local row indices of the scattered lists
    double* peer_scal[kMaxPeers];     // remote scal_in base of rank q (nullptr for self)
    unsigned long long* peer_flag[kMaxPeers];  // remote &flag_in[rank] of rank q
    double* scal_in;                  // local [2][kMaxPeers][4]
    unsigned long long* flag_in;      // local [kMaxPeers]
    long long spin_limit;             // clock64 ticks before a wait gives up (STOP_ERROR)
    int debug;                        // TIMING EXPERIMENTS ONLY (EIGSOL_B200_PEER_DEBUG): 1 no sys fence per CTA, 2 no halo stores, 4 no wait, 8 no publish
};

// Everything a matvec launch needs besides the matrix.
template <class S> struct IterArgs {
    S* buf0;             // ping-pong vector buffers, extended layout [local | halo]
    S* buf1;
    PowerState* st;
    double* partials;    // 3 doubles per CTA slot: d_re, d_im, q'
    int mode;            // MODE_*
    const PeerArgs* peer;  // device copy (MODE_PEER), else nullptr
};

// -------------------------------------------------------------------------------------------------
// State transitions.
// -------------------------------------------------------------------------------------------------
// tolerance.hpp:28-33 with a = lambdaNew, b = lambda: |a-b| <= tol*(1+|a|), |.| the complex modulus.
__device__ __forceinline__ bool close_relative(double are, double aim, double bre, double bim, double tol) {
    const double diff = hypot(are - bre, aim - bim);
    const double scale = 1.0 + hypot(are, aim);
    return diff <= tol * scale;
}

// Runs once per effective launch l, after the global sums (d_re, d_im, q') of that launch are known.
// Launch l consumed u_l (buffer in = l & 1, squared norm qin) and produced u_{l+1}.  For l >= 1 the
// sums are trip k = l-1 of power_method.hpp:68-96: lambdaNew = d, and x_{k+1} = u_l / sqrt(qin).
__device__ __forceinline__ void finalize_launch(PowerState* st, double dre, double dim, double qn, int l, int in,
                                                double qin) {
    st->q[in ^ 1] = qn;
    if (l >= 1) {
        const bool close = st->initialized && close_relative(dre, dim, st->lambda[0], st->lambda[1], st->tol);
        st->lambda[0] = dre;
        st->lambda[1] = dim;
        st->trips = l;
        st->final_buf = in;
        st->final_q = qin;
        st->initialized = 1;
        if (close) {
            st->converged = 1;
            st->stop = STOP_CONVERGED;
        } else if (l >= st->max_trips) {
            st->stop = STOP_MAX_TRIPS;
        }
    }
    st->launches = l + 1;
    st->ticket = 0;
}

// trip k = l-1 found normY == 0 (power_method.hpp:73-76): x and lambda stay those of the previous
// trip, usedIters = k+1.
__device__ __forceinline__ void zero_norm_stop(PowerState* st, int l, int in) {
    st->trips = l;
    st->final_buf = in ^ 1;
    st->final_q = st->q[in ^ 1];
    st->stop = STOP_ZERO_NORM;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ double ld_volatile_f64(const double* p) {
    double v;
    asm volatile("ld.volatile.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}

// Per-launch context (kept small: it lives in registers across the matvec loop).  The state AFTER
// everything a launch learns in its prologue (the deferred finalize of the previous launch in MODE_PEER,
// a zero-norm stop) is a pure function of the state in memory and the peers' sums, so every thread of
// every CTA derives the same decision, and the last CTA to finish re-derives the full state and is the
// only one that writes it back: no CTA ever reads a half-updated state.
struct LaunchCtx {
    int l, in;
    double qin, s;
    int dirty;    // the derived state differs from the state in memory
    int wait_ok;  // MODE_PEER: every peer signalled in time
};

// -------------------------------------------------------------------------------------------------
// CTA reduction of the three running sums.  Deterministic: fixed shuffle tree inside the CTA.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
    return v;
}

template <int BLOCK>
__device__ __forceinline__ void cta_sum3(double& a, double& b, double& c, double (*sred)[BLOCK / 32]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    a = warp_sum(a);
    b = warp_sum(b);
    c = warp_sum(c);
    if (lane == 0) {
        sred[0][warp] = a;
        sred[1][warp] = b;
        sred[2][warp] = c;
    }
    __syncthreads();
    if (warp == 0) {
        a = lane < BLOCK / 32 ? sred[0][lane] : 0.0;
        b = lane < BLOCK / 32 ? sred[1][lane] : 0.0;
        c = lane < BLOCK / 32 ? sred[2][lane] : 0.0;
        a = warp_sum(a);
        b = warp_sum(b);
        c = warp_sum(c);
    }
    __syncthreads();
}

// -------------------------------------------------------------------------------------------------
// Per-launch prologue, executed by every thread: returns false when the launch must be a no-op
// (already stopped, past maxIterations, ||y|| == 0, or a peer wait timed out).
// MODE_PEER: first wait until every peer has finished the previous launch (its halo rows and partial
// sums have landed here), add the partial sums in rank order and run the deferred finalize.
// -------------------------------------------------------------------------------------------------
// Derives (ns, c) from the state in memory; returns true when the launch is active.  No barriers.
template <class S>
__device__ __noinline__ bool derive_state(const IterArgs<S>& it, PowerState& ns, LaunchCtx& c) {
    ns = *it.st;
    c.dirty = 0;
    if (ns.stop != STOP_NONE) return false;
    if (it.mode == MODE_PEER && ns.pending) {
        const PeerArgs& pa = *it.peer;
        const int lp = ns.launches;  // executed, not yet finalized
        c.dirty = 1;
        if (!c.wait_ok) {
            ns.stop = STOP_ERROR;
            return false;
        }
        double a = 0.0, b = 0.0, d = 0.0;
        const double* sc = pa.scal_in + (lp & 1) * kMaxPeers * 4;
        for (int q = 0; q < pa.nranks; ++q) {
            a += ld_volatile_f64(sc + 4 * q + 0);
            b += ld_volatile_f64(sc + 4 * q + 1);
            d += ld_volatile_f64(sc + 4 * q + 2);
        }
        finalize_launch(&ns, a, b, d, lp, lp & 1, ns.q[lp & 1]);
        ns.pending = 0;
        if (ns.stop != STOP_NONE) return false;
    }
    c.l = ns.launches;
    if (c.l > ns.max_trips) return false;
    c.in = c.l & 1;
    c.qin = ns.q[c.in];
    if (c.qin == 0.0) {
        if (it.mode != MODE_SPLIT) {  // MODE_SPLIT: finalize_kernel takes the decision
            zero_norm_stop(&ns, c.l, c.in);
            c.dirty = 1;
        }
        return false;
    }
    c.s = 1.0 / sqrt(c.qin);
    return true;
}

template <class S, int BLOCK>
__device__ __forceinline__ bool launch_prologue(const IterArgs<S>& it, LaunchCtx& c) {
    __shared__ int s_wait_ok;
    c.wait_ok = 1;
    if (it.mode == MODE_PEER) {
        if (threadIdx.x == 0) {
            int ok = 1;
            if (it.st->stop == STOP_NONE && it.st->pending && !(it.peer->debug & 4)) {
                const PeerArgs& pa = *it.peer;
                const unsigned long long want = (unsigned long long)(it.st->launches + 1);
                const long long t0 = clock64();
                for (int q = 0; q < pa.nranks && ok; ++q) {
                    if (q == pa.rank) continue;
                    while (ld_acquire_sys(pa.flag_in + q) < want) {
                        if (clock64() - t0 > pa.spin_limit) {
                            ok = 0;
                            break;
                        }
                    }
                }
            }
            s_wait_ok = ok;
        }
        __syncthreads();
        c.wait_ok = s_wait_ok;
    }
    PowerState ns;
    return derive_state<S>(it, ns, c);
}

// No-op launch: if the prologue changed the state, the last CTA to get here writes it back.
template <class S>
__device__ __forceinline__ void finish_inactive(const IterArgs<S>& it, LaunchCtx& c) {
    if (!c.dirty || threadIdx.x != 0) return;
    const unsigned t = atomicAdd(&it.st->ticket, 1u);
    if (t == gridDim.x - 1) {
        PowerState ns;
        derive_state<S>(it, ns, c);
        ns.ticket = 0;
        *it.st = ns;
    }
}

// Active launch: CTA sums -> partial slot; the last CTA adds the slots in fixed order and either runs
// the convergence test (MODE_FUSED) or publishes this rank's sums and its completion flag to every peer
// (MODE_PEER).  slot / nslots: which partial slot this CTA owns and how many CTAs will report.
template <class S, int BLOCK>
__device__ __forceinline__ void grid_finish(const IterArgs<S>& it, LaunchCtx& c, double dre, double dim, double qn,
                                            int slot, int nslots) {
    __shared__ double sred[3][BLOCK / 32];
    __shared__ int s_last;
    cta_sum3<BLOCK>(dre, dim, qn, sred);
    if (threadIdx.x == 0) {
        it.partials[3 * slot + 0] = dre;
        it.partials[3 * slot + 1] = dim;
        it.partials[3 * slot + 2] = qn;
    }
    if (it.mode == MODE_SPLIT) return;
    if (threadIdx.x == 0) {
        if (it.mode == MODE_PEER && !(it.peer->debug & 1)) __threadfence_system();  // this CTA's remote halo stores
        else __threadfence();
        const unsigned t = atomicAdd(&it.st->ticket, 1u);
        s_last = (t == unsigned(nslots - 1));
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double a = 0.0, b = 0.0, d = 0.0;
    for (int i = threadIdx.x; i < nslots; i += BLOCK) {
        a += __ldcg(it.partials + 3 * i + 0);
        b += __ldcg(it.partials + 3 * i + 1);
        d += __ldcg(it.partials + 3 * i + 2);
    }
    cta_sum3<BLOCK>(a, b, d, sred);
    if (it.mode == MODE_FUSED) {
        if (threadIdx.x == 0) {
            PowerState ns;
            derive_state<S>(it, ns, c);
            finalize_launch(&ns, a, b, d, c.l, c.in, c.qin);
            ns.ticket = 0;
            *it.st = ns;
        }
        return;
    }
    // MODE_PEER
    const PeerArgs& pa = *it.peer;
    const int par = c.l & 1;
    // scattered send lists: rows other CTAs wrote, copied to the peers' halo segments
    const S* outv = par ? it.buf0 : it.buf1;
    for (int k = 0; k < pa.nlist; ++k) {
        S* dst = static_cast<S*>(pa.list_dst[k][par ^ 1]);
        const int32_t* idx = pa.send_idx + pa.list_off[k];
        for (int i = threadIdx.x; i < pa.list_count[k]; i += BLOCK) dst[i] = Sc<S>::ldcg(outv + idx[i]);
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        double* mine = pa.scal_in + par * kMaxPeers * 4 + pa.rank * 4;
        mine[0] = a;
        mine[1] = b;
        mine[2] = d;
        for (int q = 0; q < pa.nranks && !(pa.debug & 8); ++q) {
            if (q == pa.rank) continue;
            double* dst = pa.peer_scal[q] + par * kMaxPeers * 4 + pa.rank * 4;
            dst[0] = a;
            dst[1] = b;
            dst[2] = d;
        }
        PowerState ns;
        derive_state<S>(it, ns, c);
        ns.pending = 1;
        ns.ticket = 0;
        *it.st = ns;
        __threadfence_system();
        for (int q = 0; q < pa.nranks && !(pa.debug & 8); ++q)
            if (q != pa.rank) st_release_sys(pa.peer_flag[q], (unsigned long long)(c.l + 1));
    }
}

// Row epilogue shared by every matvec kernel: z is (A u)_row * s already.  In MODE_PEER a row that a
// peer's rows reference is also stored into that peer's halo segment (fire-and-forget NVLink store).
template <class S>
__device__ __forceinline__ void row_epilogue(const IterArgs<S>& it, const LaunchCtx& c, S z, S u_own, S* out, int row,
                                             double& dre, double& dim, double& qn) {
    const S x = Sc<S>::scale(u_own, c.s);
    Sc<S>::dotc(x, z, dre, dim);
    qn += Sc<S>::abs2(z);
    out[row] = z;
    if (it.mode == MODE_PEER) {
        const PeerArgs& pa = *it.peer;
        for (int k = 0; k < pa.nsend && !(pa.debug & 2); ++k) {
            const unsigned off = unsigned(row - pa.send_start[k]);
            if (off < unsigned(pa.send_count[k])) static_cast<S*>(pa.send_dst[k][(c.in ^ 1)])[off] = z;
        }
    }
}

// =================================================================================================
// CSR, one thread per row.  Right for very short rows; loads of vals/colidx are strided across the
// warp but every fetched sector is consumed by the following loop trips out of L1.
// =================================================================================================
template <class S, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmv_csr_scalar_kernel(CsrView<S> A, IterArgs<S> it) {
    LaunchCtx c;
    if (!launch_prologue<S, BLOCK>(it, c)) {
        finish_inactive<S>(it, c);
        return;
    }
    const S* __restrict__ u = c.in ? it.buf1 : it.buf0;
    S* __restrict__ out = c.in ? it.buf0 : it.buf1;
    double dre = 0.0, dim = 0.0, qn = 0.0;
    for (int row = blockIdx.x * BLOCK + threadIdx.x; row < A.rows; row += gridDim.x * BLOCK) {
        const int p0 = __ldg(A.rowptr + row), p1 = __ldg(A.rowptr + row + 1);
        S acc = Sc<S>::zero();
        for (int p = p0; p < p1; ++p) Sc<S>::mac(acc, __ldg(A.vals + p), Sc<S>::ldg(u + __ldg(A.colidx + p)));
        row_epilogue<S>(it, c, Sc<S>::scale(acc, c.s), Sc<S>::ldg(u + row), out, row, dre, dim, qn);
    }
    grid_finish<S, BLOCK>(it, c, dre, dim, qn, blockIdx.x, gridDim.x);
}

// =================================================================================================
// CSR, LANES lanes per row (LANES = 32 is warp-per-row).  Lanes stride over the row so each warp
// instruction reads LANES consecutive values / indices; partial sums meet in a shuffle tree.
// =================================================================================================
template <class S, int LANES, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmv_csr_vector_kernel(CsrView<S> A, IterArgs<S> it) {
    LaunchCtx c;
    if (!launch_prologue<S, BLOCK>(it, c)) {
        finish_inactive<S>(it, c);
        return;
    }
    const S* __restrict__ u = c.in ? it.buf1 : it.buf0;
    S* __restrict__ out = c.in ? it.buf0 : it.buf1;
    double dre = 0.0, dim = 0.0, qn = 0.0;
    const int sub = threadIdx.x % LANES;
    const int group = (blockIdx.x * BLOCK + threadIdx.x) / LANES;
    const int ngroups = gridDim.x * (BLOCK / LANES);
    // all lanes of a warp run the same number of trips so the full-mask shuffles stay converged
    const int rows_rounded = (A.rows + (32 / LANES) - 1) / (32 / LANES) * (32 / LANES);
    for (int row = group; row < rows_rounded; row += ngroups) {
        S acc = Sc<S>::zero();
        if (row < A.rows) {
            const int p0 = __ldg(A.rowptr + row), p1 = __ldg(A.rowptr + row + 1);
            for (int p = p0 + sub; p < p1; p += LANES)
                Sc<S>::mac(acc, ld_stream_val<S>(A.vals + p), Sc<S>::ldg(u + ld_stream_s32(A.colidx + p)));
        }
#pragma unroll
        for (int d = LANES / 2; d > 0; d >>= 1) Sc<S>::add(acc, Sc<S>::shfl_down(acc, d, LANES));
        if (sub == 0 && row < A.rows)
            row_epilogue<S>(it, c, Sc<S>::scale(acc, c.s), Sc<S>::ldg(u + row), out, row, dre, dim, qn);
    }
    grid_finish<S, BLOCK>(it, c, dre, dim, qn, blockIdx.x, gridDim.x);
}

// =================================================================================================
// CSR streamed through shared memory (CSR-stream / CSR-adaptive shape).  A CTA owns a block of
// consecutive rows whose entries fit one tile: the entries are read as one contiguous stream
// (16-byte value loads, 8-byte index loads, fully coalesced), multiplied with the gathered x and
// parked in shared memory; then one thread per row sums its segment in ascending column order.
// A row longer than the tile is its own block and is reduced by the whole CTA.
// =================================================================================================
template <class S> struct StreamCfg;
template <> struct StreamCfg<double> { static constexpr int TILE = 4096; };   // 32 KB of products
template <> struct StreamCfg<double2> { static constexpr int TILE = 2048; };  // 32 KB of products
constexpr int kStreamMaxRows = 512;

template <class S, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmv_csr_stream_kernel(CsrView<S> A, const int32_t* __restrict__ rowblk,
                                                                int nblk, IterArgs<S> it) {
    constexpr int TILE = StreamCfg<S>::TILE;
    __shared__ __align__(16) S prod[TILE];
    __shared__ int32_t srow[kStreamMaxRows + 1];
    __shared__ double slong[3][BLOCK / 32];
    LaunchCtx c;
    if (!launch_prologue<S, BLOCK>(it, c)) {
        finish_inactive<S>(it, c);
        return;
    }
    const S* __restrict__ u = c.in ? it.buf1 : it.buf0;
    S* __restrict__ out = c.in ? it.buf0 : it.buf1;
    double dre = 0.0, dim = 0.0, qn = 0.0;
    const double* vals_d = reinterpret_cast<const double*>(A.vals);
    for (int b = blockIdx.x; b < nblk; b += gridDim.x) {
        const int r0 = __ldg(rowblk + b), r1 = __ldg(rowblk + b + 1);
        const int nr = r1 - r0;
        for (int i = threadIdx.x; i <= nr; i += BLOCK) srow[i] = __ldg(A.rowptr + r0 + i);
        __syncthreads();
        const int p0 = srow[0], p1 = srow[nr];
        const int base = p0 & ~3;
        if (p1 - base > TILE) {
            // one long row: CTA-strided partial sums, then a CTA reduction (real and imaginary parts).
            S acc = Sc<S>::zero();
            for (int p = p0 + threadIdx.x; p < p1; p += BLOCK)
                Sc<S>::mac(acc, ld_stream_val<S>(A.vals + p), Sc<S>::ldg(u + ld_stream_s32(A.colidx + p)));
            double ar, ai = 0.0, dummy = 0.0;
            if constexpr (Sc<S>::is_complex) {
                ar = acc.x;
                ai = acc.y;
            } else {
                ar = acc;
            }
            cta_sum3<BLOCK>(ar, ai, dummy, slong);
            if (threadIdx.x == 0) {
                S z;
                if constexpr (Sc<S>::is_complex) z = make_double2(ar * c.s, ai * c.s);
                else z = ar * c.s;
                row_epilogue<S>(it, c, z, Sc<S>::ldg(u + r0), out, r0, dre, dim, qn);
            }
        } else {
            for (int e = base + 2 * threadIdx.x; e < p1; e += 2 * BLOCK) {
                const int2 cc = ld_stream_s32x2(A.colidx + e);
                S v0, v1;
                if constexpr (Sc<S>::is_complex) {
                    v0 = ld_stream_f64x2(vals_d + 2 * size_t(e));
                    v1 = ld_stream_f64x2(vals_d + 2 * size_t(e) + 2);
                } else {
                    const double2 vv = ld_stream_f64x2(vals_d + e);
                    v0 = vv.x;
                    v1 = vv.y;
                }
                prod[e - base] = Sc<S>::mul(v0, Sc<S>::ldg(u + cc.x));
                prod[e - base + 1] = Sc<S>::mul(v1, Sc<S>::ldg(u + cc.y));
            }
            __syncthreads();
            for (int i = threadIdx.x; i < nr; i += BLOCK) {
                S acc = Sc<S>::zero();
                const int q1 = srow[i + 1] - base;
                for (int q = srow[i] - base; q < q1; ++q) Sc<S>::add(acc, prod[q]);
                row_epilogue<S>(it, c, Sc<S>::scale(acc, c.s), Sc<S>::ldg(u + r0 + i), out, r0 + i, dre, dim, qn);
            }
        }
        __syncthreads();
    }
    grid_finish<S, BLOCK>(it, c, dre, dim, qn, blockIdx.x, gridDim.x);
}

// =================================================================================================
// SELL: a warp owns a slice of 32*R consecutive rows, lane l the R rows  slice*32R + l*R + {0..R-1}.
// Entry j of those rows is stored at  slice_off + (j*32 + l)*R : one warp instruction reads 32*R
// consecutive values (R=2, f64: 16-byte loads; c128: 16 bytes per value) and 32*R consecutive column
// indices, and for stencil-like rows the x gathers of neighbouring lanes fall into the same lines.
// No shared memory, no block barrier in the main loop.
// =================================================================================================
template <class S, int R, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmv_sell_kernel(SellView<S> A, IterArgs<S> it) {
    LaunchCtx c;
    if (!launch_prologue<S, BLOCK>(it, c)) {
        finish_inactive<S>(it, c);
        return;
    }
    const S* __restrict__ u = c.in ? it.buf1 : it.buf0;
    S* __restrict__ out = c.in ? it.buf0 : it.buf1;
    double dre = 0.0, dim = 0.0, qn = 0.0;
    const int lane = threadIdx.x & 31;
    const int gwarp = (blockIdx.x * BLOCK + threadIdx.x) >> 5;
    const int nwarps = gridDim.x * (BLOCK / 32);
    for (int sl = gwarp; sl < A.nslices; sl += nwarps) {
        const int64_t off = __ldg(A.slice_off + sl);
        const int width = int((__ldg(A.slice_off + sl + 1) - off) / (32 * R));
        const S* vp = A.vals + off + lane * R;
        const int32_t* cp = A.colidx + off + lane * R;
        S acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = Sc<S>::zero();
#pragma unroll 4
        for (int j = 0; j < width; ++j) {
            S v[R];
            int32_t col[R];
            if constexpr (R == 2) {
                const int2 cc = ld_stream_s32x2(cp);
                col[0] = cc.x;
                col[1] = cc.y;
                if constexpr (Sc<S>::is_complex) {
                    v[0] = ld_stream_val<S>(vp);
                    v[1] = ld_stream_val<S>(vp + 1);
                } else {
                    const double2 vv = ld_stream_f64x2(reinterpret_cast<const double*>(vp));
                    v[0] = vv.x;
                    v[1] = vv.y;
                }
            } else {
                col[0] = ld_stream_s32(cp);
                v[0] = ld_stream_val<S>(vp);
            }
#pragma unroll
            for (int r = 0; r < R; ++r)
                if (col[r] >= 0) Sc<S>::mac(acc[r], v[r], Sc<S>::ldg(u + col[r]));
            vp += 32 * R;
            cp += 32 * R;
        }
        const int row = (sl * 32 + lane) * R;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (row + r < A.rows)
                row_epilogue<S>(it, c, Sc<S>::scale(acc[r], c.s), Sc<S>::ldg(u + row + r), out, row + r, dre, dim, qn);
    }
    grid_finish<S, BLOCK>(it, c, dre, dim, qn, blockIdx.x, gridDim.x);
}

// =================================================================================================
// Dense column-major GEMV.  CTA (tile, seg) owns RT = BLOCK*R rows (R = 2 for f64 -> 16-byte loads
// down a column, 1 for c128) and the columns of segment seg.  x = u*s of the segment is staged in
// shared memory once per CTA.  Segment sums go to partial[seg][row]; the last CTA of a row tile adds
// them in segment order (deterministic), runs the row epilogue and reports the tile's sums.
// =================================================================================================
constexpr int kGemvChunk = 512;  // columns of x staged per pass

template <class S, int BLOCK>
__global__ void __launch_bounds__(BLOCK) gemv_colmajor_kernel(DenseView<S> A, IterArgs<S> it) {
    constexpr int R = Sc<S>::is_complex ? 1 : 2;
    constexpr int RT = BLOCK * R;
    __shared__ __align__(16) S xs[kGemvChunk];
    __shared__ int s_tile_last;
    LaunchCtx c;
    if (!launch_prologue<S, BLOCK>(it, c)) {
        finish_inactive<S>(it, c);
        return;
    }
    const S* __restrict__ u = c.in ? it.buf1 : it.buf0;
    S* __restrict__ out = c.in ? it.buf0 : it.buf1;
    const int tile = blockIdx.x / A.nseg, seg = blockIdx.x % A.nseg;
    const int c0 = int(int64_t(A.cols) * seg / A.nseg), c1 = int(int64_t(A.cols) * (seg + 1) / A.nseg);
    const int r = tile * RT + threadIdx.x * R;
    const bool live = r < A.lda;  // lda is even and rows beyond A.rows are zero-filled on the device
    S acc[R];
#pragma unroll
    for (int k = 0; k < R; ++k) acc[k] = Sc<S>::zero();
    for (int cb = c0; cb < c1; cb += kGemvChunk) {
        const int nc = min(kGemvChunk, c1 - cb);
        __syncthreads();
        for (int j = threadIdx.x; j < nc; j += BLOCK) xs[j] = Sc<S>::scale(Sc<S>::ldg(u + cb + j), c.s);
        __syncthreads();
        if (live) {
            const S* col = A.a + int64_t(cb) * A.lda + r;
#pragma unroll 8
            for (int j = 0; j < nc; ++j) {
                if constexpr (R == 2) {
                    const double2 a = ld_stream_f64x2(reinterpret_cast<const double*>(col));
                    acc[0] = fma(a.x, xs[j], acc[0]);
                    acc[1] = fma(a.y, xs[j], acc[1]);
                } else {
                    Sc<S>::mac(acc[0], ld_stream_val<S>(col), xs[j]);
                }
                col += A.lda;
            }
        }
    }
    if (live) {
#pragma unroll
        for (int k = 0; k < R; ++k) A.partial[int64_t(seg) * A.lda + r + k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned t = atomicAdd(A.tile_ticket + tile, 1u);
        s_tile_last = (t == unsigned(A.nseg - 1));
        if (s_tile_last) A.tile_ticket[tile] = 0;
    }
    __syncthreads();
    if (!s_tile_last) return;
    __threadfence();
    double dre = 0.0, dim = 0.0, qn = 0.0;
#pragma unroll
    for (int k = 0; k < R; ++k) {
        const int row = r + k;
        if (row < A.rows) {
            S z = Sc<S>::zero();
            for (int sgi = 0; sgi < A.nseg; ++sgi) Sc<S>::add(z, Sc<S>::ldcg(A.partial + int64_t(sgi) * A.lda + row));
            row_epilogue<S>(it, c, z, Sc<S>::ldg(u + row), out, row, dre, dim, qn);
        }
    }
    grid_finish<S, BLOCK>(it, c, dre, dim, qn, tile, A.ntiles);
}

// =================================================================================================
// Small helper kernels.
// =================================================================================================
// Multi-GPU: fixed-order sum of the CTA partials into scal[3] (then NCCL all-reduces scal).
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) reduce_partials_kernel(const double* partials, int nslots, PowerState* st,
                                                                double* scal) {
    __shared__ double sred[3][BLOCK / 32];
    double a = 0.0, b = 0.0, d = 0.0;
    const int stop = st->stop;
    const bool active = stop == STOP_NONE && st->launches <= st->max_trips && st->q[st->launches & 1] != 0.0;
    if (active)
        for (int i = threadIdx.x; i < nslots; i += BLOCK) {
            a += partials[3 * i + 0];
            b += partials[3 * i + 1];
            d += partials[3 * i + 2];
        }
    cta_sum3<BLOCK>(a, b, d, sred);
    if (threadIdx.x == 0) {
        scal[0] = a;
        scal[1] = b;
        scal[2] = d;
    }
}

// Multi-GPU: the state transition after the all-reduce; identical inputs on every rank, so every
// rank takes the same decision without a broadcast.
__global__ void finalize_kernel(PowerState* st, const double* scal) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const int l = st->launches;
    if (st->stop != STOP_NONE || l > st->max_trips) return;
    const int in = l & 1;
    const double qin = st->q[in];
    if (qin == 0.0) {
        zero_norm_stop(st, l, in);
        return;
    }
    finalize_launch(st, scal[0], scal[1], scal[2], l, in, qin);
}

// MODE_PEER: folds the sums of the last executed launch into the state (run before the host reads it,
// so that it sees the same state a single-GPU run would have after that launch).
template <class S>
__global__ void peer_flush_kernel(IterArgs<S> it) {
    LaunchCtx c;
    launch_prologue<S, 32>(it, c);
    if (c.dirty && threadIdx.x == 0) {
        PowerState ns;
        derive_state<S>(it, ns, c);
        ns.ticket = 0;
        *it.st = ns;
    }
}

// x = y / normY (power_method.hpp:78, true division) of the result buffer -> contiguous output.
template <class S>
__global__ void normalize_result_kernel(const S* buf0, const S* buf1, const PowerState* st, S* x, int n) {
    const S* src = st->final_buf ? buf1 : buf0;
    const double nrm = sqrt(st->final_q);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        x[i] = Sc<S>::div(src[i], nrm);
}

// Multi-GPU: gather the entries a peer needs into a contiguous send buffer.
template <class S>
__global__ void pack_kernel(const S* __restrict__ u, const int32_t* __restrict__ idx, S* __restrict__ sendbuf, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        sendbuf[i] = u[idx[i]];
}

}  // namespace eigsol_b200
