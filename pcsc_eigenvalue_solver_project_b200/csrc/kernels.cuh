// kernels.cuh -- hand-written sm_100a kernels of the power-iteration path.
//
// One launch == one power iteration's matvec with everything else fused into it or overlapped with it:
//     v' = sigma * A v               (reference: src/power_method/power_method.hpp:69; the A*x temporary of
//                                     :81 is the next trip's y, computed once; sigma = power-of-two scale)
//     a  = v^H v'  (conjugated dot)  (:81, eigen/Eigen/src/Core/InnerProduct.h:255)
//     q' = ||v'||^2                  (:72 of the next trip, eigen/Eigen/src/Core/Dot.h:61-77)
// and, one launch later and in parallel with that launch's matvec CTAs, the Rayleigh quotient,
// the convergence test of :83-95 (src/core/tolerance.hpp:28-33) and -- on N GPUs -- the all-reduce of
// (a, q') in the launch's control CTA.  The normalised x is never written to HBM, so an iteration reads
// A once, reads v once (gathers served from L1/L2), writes v' once: exactly the algorithmic bytes of
// SURVEY.md 8(d).
//
// Bandwidth-bound integer/fp64 streaming work: no tensor cores.  What matters here is coalescing
// (128-bit loads of values / column indices), enough bytes in flight per SM, a persistent grid sized
// from the SM count, and keeping the x gathers in L1/L2 (matrix streams bypass L1 allocation).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eigsol_b200 {

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
__device__ __forceinline__ void ld_stream_f64x4(const double* p, double (&o)[4]) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p));
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
    int32_t tile_major;     // launch order of the CTAs: 1 = row tiles of one column segment are neighbours
};

// -------------------------------------------------------------------------------------------------
// Device-resident iteration state (one per matrix handle).  Mirrors the locals of powerMethodImpl
// (power_method.hpp:61-66): lambda, usedIters, converged, initialized.
//
// Pipeline.  Launch m reads the vector v_m (ring buffer m % 3) and writes v_{m+1} = sigma_m * A v_m into
// buffer (m+1) % 3; every matvec CTA leaves three partial sums
//     a_m = v_m^H (sigma_m A v_m)      (re, im)          q_{m+1} = ||v_{m+1}||^2
// in its slot and exits -- no atomics, no fences, no "last CTA".  CTA 0 of every launch is the CONTROL
// CTA: while the other CTAs of launch m stream the matrix, it folds the sums of launch m-1 into the state
// (on N GPUs: publishes them to every peer and adds the peers' sums in rank order), closes trip m-2 of
// power_method.hpp:68-96
//     lambda = a_{m-1} / (sigma_{m-1} q_{m-1})   ( = x^H A x with x = v_{m-1}/||v_{m-1}|| , :81 )
//     convergence test (:83-91), maxIterations, ||y|| == 0 exit (:73-76)
// and writes the control record of launch m+1 (scale, active flag).  So the scalar reduction, the
// cross-GPU all-reduce and the state transition are OFF the critical path: they overlap a whole matvec.
// The scale sigma is a power of two chosen from the norms seen two launches ago, so the stored vector
// stays O(1) and the scaling itself is exact; the Rayleigh quotient is scale-invariant and the returned
// eigenvector is normalised by a true division at the end (:78).  Three buffers keep the vector of the
// stopping trip intact while the (at most one) launch already in flight completes.
// -------------------------------------------------------------------------------------------------
constexpr int kRing = 3;

struct LaunchCtrl {
    double sigma;   // v_{m+1} = sigma * A v_m
    int launch_no;  // m
    int active;     // 0: the launch is a no-op (stopped, or past maxIterations)
};

struct PowerState {
    double lambda[2];      // current eigenvalue estimate (re, im)
    double tol;            // opts.tolerance
    double q[kRing];       // ||v_m||^2 at slot m % 3
    double sigma[kRing];   // sigma_m at slot m % 3
    double final_q;        // ||v||^2 of the buffer holding the result vector
    double last_a[2];      // v^H (sigma A v) of the launch folded last (eigsol_b200_rayleigh reads it)
    unsigned long long flag_base;  // peer flags of this solve are flag_base + launch number (never reset: monotonic)
    int max_trips;         // opts.maxIterations
    int launch_no;         // launches executed since begin (the next launch is number launch_no)
    int folded;            // launches whose sums have been folded into the state
    int trips;             // usedIters
    int initialized;       // power_method.hpp:66
    int stop;              // STOP_*
    int converged;
    int final_buf;         // ring buffer whose normalisation is EigenResult::eigenvector
    int error;             // a peer never signalled (wait timed out)
    int reserved;
    LaunchCtrl ctrl[kRing];  // control record of launch m at slot m % 3
};

enum { STOP_NONE = 0, STOP_CONVERGED = 1, STOP_ZERO_NORM = 2, STOP_MAX_TRIPS = 3, STOP_ERROR = 4 };

// tolerance.hpp:28-33 with a = lambdaNew, b = lambda: |a-b| <= tol*(1+|a|), |.| the complex modulus.
__host__ __device__ __forceinline__ bool close_relative(double are, double aim, double bre, double bim, double tol) {
    const double diff = hypot(are - bre, aim - bim);
    const double scale = 1.0 + hypot(are, aim);
    return diff <= tol * scale;
}

// Folds the global sums of launch j (a = v_j^H z, qn = ||v_{j+1}||^2) into the state.  For j >= 1 this
// closes trip k = j-1 of power_method.hpp:68-96: x_{k+1} = v_j / ||v_j|| and lambdaNew = x^H A x.
__host__ __device__ __forceinline__ void fold_launch(PowerState* st, int j, double are, double aim, double qn) {
    const int r = j % kRing;
    st->q[(j + 1) % kRing] = qn;
    st->last_a[0] = are;
    st->last_a[1] = aim;
    if (j >= 1) {
        const double denom = st->sigma[r] * st->q[r];
        const double lre = are / denom, lim = aim / denom;
        const bool close = st->initialized && close_relative(lre, lim, st->lambda[0], st->lambda[1], st->tol);
        st->lambda[0] = lre;
        st->lambda[1] = lim;
        st->trips = j;
        st->final_buf = r;
        st->final_q = st->q[r];
        st->initialized = 1;
        if (close) {
            st->converged = 1;
            st->stop = STOP_CONVERGED;
        } else if (j >= st->max_trips) {
            st->stop = STOP_MAX_TRIPS;
        }
    }
    if (st->stop == STOP_NONE && qn == 0.0) {
        // trip j found normY == 0 (power_method.hpp:73-76): x and lambda stay those of the previous
        // trip, usedIters = j+1.
        st->trips = j + 1;
        st->final_buf = r;
        st->final_q = st->q[r];
        st->stop = STOP_ZERO_NORM;
    }
    st->folded = j + 1;
}

// Scale of launch m+1, from what is known after folding launch m-1: growth^2 g2 = ||A v||^2/||v||^2 seen
// in launch m-1 and the predicted ||v_{m+1}||^2 = sigma_m^2 g2 q_m.  Rounded to a power of two.
// Scale of the first two launches, before any norm has been seen: a power of two near 1/max|a_ij| (found once
// at upload), so that matrices scaled by 1e+-150 neither overflow nor underflow (which would read as the
// ||y|| == 0 exit) during the two launches the adaptive scale needs to catch up.
__host__ __device__ __forceinline__ double prescale_from_amax(double amax) {
    if (!(amax > 0.0) || !isfinite(amax)) return 1.0;
    const int k = int(rint(log2(amax)));
    if (k > 1000 || k < -1000) return 1.0;
    return ldexp(1.0, -k);
}

__host__ __device__ __forceinline__ double next_sigma(const PowerState* st, int m) {
    if (m == 0) return st->sigma[0];
    const int r = m % kRing, p = (m - 1) % kRing;
    const double qm = st->q[r], qp = st->q[p], sm = st->sigma[r], sp = st->sigma[p];
    const double g2 = qm / (sp * sp * qp);
    const double pred2 = sm * sm * g2 * qm;
    const double x = g2 * pred2;
    if (!(x > 0.0) || !isfinite(x)) return 1.0;
    const int k = int(rint(0.5 * log2(x)));
    if (k > 900 || k < -900) return 1.0;
    return ldexp(1.0, -k);
}

// The control CTA's decisions, separated from how the sums reach it so that tests can drive the same state
// machine on the host (tests/statemachine/sim.cu).  `control_wants_fold`: launch launch_no-1 ran actively and
// its sums are not folded yet.  `control_apply`: fold them (or note a no-op launch, or an error), and --
// in a matvec launch (advance) -- write the control record of launch m+1 and count launch m.
struct ControlPlan {
    int m, j;
    bool have, fold;
};
__host__ __device__ __forceinline__ ControlPlan control_wants_fold(const PowerState* st) {
    ControlPlan p;
    p.m = st->launch_no;
    p.j = p.m - 1;
    p.have = p.j >= 0 && st->folded == p.j;
    p.fold = p.have && st->ctrl[(p.j + kRing) % kRing].active && st->stop == STOP_NONE && !st->error;
    return p;
}
__host__ __device__ __forceinline__ void control_apply(PowerState* st, const ControlPlan& p, bool ok, double a, double b,
                                                       double d, bool advance) {
    if (!ok || *(volatile int*)&st->error) {
        st->error = 1;
        st->stop = STOP_ERROR;
    } else if (p.fold) {
        fold_launch(st, p.j, a, b, d);
    } else if (p.have) {
        st->folded = p.j + 1;  // launch j was a no-op
    }
    if (advance) {
        LaunchCtrl nc;
        nc.sigma = next_sigma(st, p.m);
        nc.launch_no = p.m + 1;
        nc.active = (st->stop == STOP_NONE && p.m + 1 <= st->max_trips) ? 1 : 0;
        st->sigma[(p.m + 1) % kRing] = nc.sigma;
        st->ctrl[(p.m + 1) % kRing] = nc;
        st->launch_no = p.m + 1;
    }
}

// How the sums of a launch become global and how the halo of x travels.
enum { MODE_FUSED = 0,   // one GPU
       MODE_SPLIT = 1,   // N GPUs over NCCL: send/recv of the halo before, all-reduce of the sums after each launch
       MODE_PEER = 2 };  // N GPUs over peer memory: the matvec kernel itself stores boundary rows into the peers'
                         // halo segments and the control CTA exchanges the sums (P2P stores over NVLink, flags with
                         // system-scope release/acquire)

constexpr int kMaxPeers = 8;

// Multi-GPU peer-memory wiring of one rank (MODE_PEER).  Launch m reads ring buffer m%3 (own rows + halo)
// and writes (m+1)%3.  Rows of v_{m+1} that a peer's rows reference are ALSO stored straight into that
// peer's halo segment of ITS buffer (m+1)%3 as they are produced.  The control CTA of launch m+1 then
// writes this rank's sums of launch m into every peer's scal_in[m%3][rank] and sets flag_in[rank] = m+1
// on every peer (release): "my launches <= m are complete -- your halo of v_{m+1} has landed -- and my
// sums of launch m are in your mailbox".  A peer's launch m+1 processes the rows that read halo entries
// LAST and waits for the flag (acquire) only before those, so the exchange overlaps its interior rows.
struct PeerArgs {
    int nranks, rank;
    int nsend;                        // contiguous send ranges (rows stored remotely as they are produced)
    int send_start[kMaxPeers];        // first local row of the range
    int send_count[kMaxPeers];
    void* send_dst[kMaxPeers][kRing]; // remote address of the peer's halo segment, per ring buffer
    int nlist;                        // scattered send lists (copied by the next launch's control CTA)
    int list_off[kMaxPeers], list_count[kMaxPeers];
    void* list_dst[kMaxPeers][kRing];
    const int32_t* send_idx;          // local row indices of the scattered lists
    double* peer_scal[kMaxPeers];     // remote scal_in base of rank q (nullptr for self)
    unsigned long long* peer_flag[kMaxPeers];  // remote &flag_in[rank] of rank q
    double* scal_in;                  // local [kRing][kMaxPeers][4]
    unsigned long long* flag_in;      // local [kMaxPeers]
    long long spin_limit;             // clock64 ticks before a wait gives up (error -> STOP_ERROR)
    int debug_nowait;                 // timing experiments only: waits return at once (results are wrong)
    unsigned halo_src_mask;           // ranks that own halo entries of this rank (the ones a matvec warp waits for)
    // device-side barrier of the ranks (peer_gate_kernel): gate_in[q] = last gate epoch rank q has entered
    unsigned long long* gate_in;               // local [kMaxPeers]
    unsigned long long* peer_gate[kMaxPeers];  // remote &gate_in[rank] of rank q
    // EIGSOL_B200_PEER_STATS=1: where a multi-GPU trip waits (nanoseconds of %globaltimer), else nullptr
    //   [0] control CTA: ns between publishing the flag and having every peer's flag, summed   [1] launches counted
    //   [2] matvec warps that found the halo flag missing: ns spun, summed                     [3] such warps
    //   [4] the longest single halo wait                [5] matvec warps that reached a halo unit (waited or not)
    unsigned long long* stats;
};

constexpr int kPeerStats = 8;

// Everything a matvec launch needs besides the matrix.
template <class S> struct IterArgs {
    S* buf[kRing];           // ring of vector buffers, extended layout [local | halo]
    PowerState* st;
    double* partials;        // [kRing][nslots][3]: a_re, a_im, q' per matvec CTA (or GEMV row tile)
    int nslots;
    int mode;                // MODE_*
    const PeerArgs* peer;    // device copy (MODE_PEER), else nullptr
    const double* scal_glob; // MODE_SPLIT: [kRing][4] all-reduced sums of launch j at slot j%3
    const uint8_t* deferred; // MODE_PEER: per work unit, 1 = reads halo entries (needs the halo wait first); may be null
    int serpentine;          // odd launches walk the work units backwards (what was touched last is re-read first: L2)
    int unit_start;          // rotation of the unit walk (see walk_units)
    int halo_base;           // MODE_PEER: first halo index of the vector buffers; INT32_MAX otherwise (diagnostics)
};

// Gather of one x entry, per exchange mode (the matvec kernels are compiled twice, PEER = false / true):
//   PEER = false (one GPU, NCCL mode): nothing writes the input vector while the kernel runs -> the non-coherent
//       read-only path (ld.global.nc via __ldg), which also lets the compiler batch the gathers of an unrolled row.
//   PEER = true (MODE_PEER): the halo segment of the input vector is stored by peer GPUs WHILE this kernel is already
//       running and is read only after the acquire of the owner's flag.  PTX defines ld.global.nc only for data nobody
//       writes during the kernel's lifetime, so this variant uses plain (coherent-at-acquire, L1-cached) loads for every
//       gather and carries no __restrict__ on the vector.  No halo line can sit in L1 before the flag: the halo starts
//       on a 128-byte boundary (nloc_pad) and nothing reads it before the wait.
template <class S, bool PEER> struct XPtr { typedef const S* __restrict__ type; };
template <class S> struct XPtr<S, true> { typedef const S* type; };

template <class S, bool PEER> __device__ __forceinline__ S ld_x(const S* u, int idx) {
    if constexpr (PEER) return u[idx];
    else return Sc<S>::ldg(u + idx);
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

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// Spins (one thread, acquire loads: measured cheaper than relaxed loads + a system fence) until the flags of
// the peers in `mask` have reached `want`; false on timeout.
__device__ __forceinline__ bool wait_peer_flags(const PeerArgs& pa, unsigned mask, unsigned long long want) {
    if (pa.debug_nowait) return true;
    const long long t0 = clock64();
    for (int q = 0; q < pa.nranks; ++q) {
        if (q == pa.rank || !((mask >> q) & 1u)) continue;
        while (ld_acquire_sys(pa.flag_in + q) < want)
            if (clock64() - t0 > pa.spin_limit) return false;
    }
    return true;
}

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

// Programmatic dependent launch: every matvec kernel lets its successor's CTAs become resident at once
// (they fill the SM slots this launch's tail frees) and itself waits here for its predecessor to have
// completed and flushed before it reads anything that launch wrote.  No-ops without the launch attribute.
__device__ __forceinline__ void pdl_prologue() {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
// The two halves, for kernels that use the time between becoming resident and the predecessor's completion: the
// matrix is read-only across launches, so its first bytes can be pulled into L2 (prefetch_l2) while the previous
// launch's tail -- which leaves HBM idle -- is still running; only the vector depends on the predecessor.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// one bulk L2 prefetch (TMA engine, no register or shared-memory destination); bytes a multiple of 16, 16-byte aligned
__device__ __forceinline__ void prefetch_l2(const void* p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// -------------------------------------------------------------------------------------------------
// The control CTA: blockIdx.x == 0 of every matvec launch (advance = true), and the whole of
// control_kernel (advance = false: folds what is pending so the host reads a settled state).
// Control CTAs are the only writers of the state and are ordered by the stream.
// -------------------------------------------------------------------------------------------------
template <class S, int BLOCK>
__device__ void control_cta(const IterArgs<S>& it, bool advance) {
    __shared__ double sred[3][BLOCK / 32];
    PowerState* st = it.st;
    const ControlPlan plan = control_wants_fold(st);
    const int m = plan.m;
    const int j = plan.j;  // the launch whose sums are folded now
    const unsigned long long flagv = st->flag_base + (unsigned long long)m;
    const bool fold = plan.fold;
    double a = 0.0, b = 0.0, d = 0.0;
    if (fold) {
        const double* part = it.partials + size_t(j % kRing) * size_t(it.nslots) * 3;
        for (int i = threadIdx.x; i < it.nslots; i += BLOCK) {
            a += __ldcg(part + 3 * i + 0);
            b += __ldcg(part + 3 * i + 1);
            d += __ldcg(part + 3 * i + 2);
        }
    }
    cta_sum3<BLOCK>(a, b, d, sred);
    if (it.mode == MODE_PEER && fold) {
        // scattered send lists: rows of v_m the peers' rows reference, copied into their halo segments
        const PeerArgs& pa = *it.peer;
        const S* v = it.buf[m % kRing];
        for (int k = 0; k < pa.nlist; ++k) {
            S* dst = static_cast<S*>(pa.list_dst[k][m % kRing]);
            const int32_t* idx = pa.send_idx + pa.list_off[k];
            for (int i = threadIdx.x; i < pa.list_count[k]; i += BLOCK) dst[i] = Sc<S>::ldcg(v + idx[i]);
        }
        if (pa.nlist) __threadfence_system();
    }
    __shared__ double s_tot[3];
    __shared__ int s_okv[kMaxPeers];
    if (threadIdx.x == 0) {
        s_tot[0] = a;
        s_tot[1] = b;
        s_tot[2] = d;
    }
    __syncthreads();
    if (it.mode == MODE_PEER) {
        // thread q talks to rank q, all peers in parallel: sums of launch j into q's mailbox, then (release) the
        // flag "my launches < m are complete (your halo of v_m has landed) and my sums of launch m-1
        // are with you"; then wait for q's flag in return.
        const PeerArgs& pa = *it.peer;
        const int q = threadIdx.x;
        const unsigned long long t_pub = (pa.stats && threadIdx.x == 0) ? global_ns() : 0ull;
        if (q < pa.nranks) {
            int okq = 1;
            const int slot = ((j + kRing) % kRing) * kMaxPeers * 4 + pa.rank * 4;
            if (q == pa.rank) {
                if (fold) {
                    pa.scal_in[slot + 0] = s_tot[0];
                    pa.scal_in[slot + 1] = s_tot[1];
                    pa.scal_in[slot + 2] = s_tot[2];
                }
            } else {
                if (fold) {
                    double* dst = pa.peer_scal[q] + slot;
                    dst[0] = s_tot[0];
                    dst[1] = s_tot[1];
                    dst[2] = s_tot[2];
                }
                st_release_sys(pa.peer_flag[q], flagv);
                if (fold) okq = wait_peer_flags(pa, 1u << q, flagv) ? 1 : 0;
            }
            s_okv[q] = okq;
        }
        __syncthreads();
        if (pa.stats && threadIdx.x == 0 && fold) {
            pa.stats[0] += global_ns() - t_pub;  // control CTAs are ordered by the stream: plain updates
            pa.stats[1] += 1ull;
        }
    }
    if (threadIdx.x != 0) return;
    int ok = 1;
    if (it.mode == MODE_PEER) {
        const PeerArgs& pa = *it.peer;
        for (int q = 0; q < pa.nranks; ++q) ok &= s_okv[q];
        if (fold && ok) {
            a = b = d = 0.0;
            const double* sc = pa.scal_in + (j % kRing) * kMaxPeers * 4;
            for (int q = 0; q < pa.nranks; ++q) {  // rank order: the same sum, bit for bit, on every rank
                a += ld_volatile_f64(sc + 4 * q + 0);
                b += ld_volatile_f64(sc + 4 * q + 1);
                d += ld_volatile_f64(sc + 4 * q + 2);
            }
        }
    } else if (it.mode == MODE_SPLIT && fold) {
        const double* sc = it.scal_glob + (j % kRing) * 4;
        a = __ldcg(sc + 0);
        b = __ldcg(sc + 1);
        d = __ldcg(sc + 2);
    }
    control_apply(st, plan, ok != 0, a, b, d, advance);
}

// Control-only launch (host poll / end): one CTA.
template <class S>
__global__ void __launch_bounds__(256) control_kernel(IterArgs<S> it) {
    control_cta<S, 256>(it, false);
}

// Device-side barrier of the ranks of a MODE_PEER group: thread q announces this rank's arrival at gate
// `epoch` in rank q's gate_in[rank] and spins until rank q has announced the same.  The kernel (and with it
// everything queued behind it on the stream) completes within a couple of microseconds of the LAST rank's
// arrival, on every rank.  Used where the host would otherwise need a collective: before the first launch of a
// solve (no rank may store into a peer's ring buffers before that peer has reset them) and in front of a timed
// batch of trips, so that the ranks' event timers start together whatever the skew of the host threads.
__global__ void __launch_bounds__(32) peer_gate_kernel(const PeerArgs* pa_ptr, unsigned long long epoch, PowerState* st) {
    const PeerArgs& pa = *pa_ptr;
    const int q = threadIdx.x;
    if (q >= pa.nranks || q == pa.rank) return;
    st_release_sys(pa.peer_gate[q], epoch);
    if (pa.debug_nowait) return;
    const long long t0 = clock64();
    while (ld_acquire_sys(pa.gate_in + q) < epoch)
        if (clock64() - t0 > pa.spin_limit) {
            *(volatile int*)&st->error = 1;
            return;
        }
}

// -------------------------------------------------------------------------------------------------
// Matvec CTAs: per-launch context, halo wait, peer sends, CTA finish.
// -------------------------------------------------------------------------------------------------
struct LaunchCtx {
    int m, in, out, reverse;
    double s;
    unsigned long long flagv;  // peers' flag value that says the halo of v_m has landed
};

template <class S>
__device__ __forceinline__ bool launch_begin(const IterArgs<S>& it, int phase, LaunchCtx& c) {
    const LaunchCtrl ct = it.st->ctrl[phase];
    if (!ct.active) return false;
    c.m = ct.launch_no;
    c.s = ct.sigma;
    c.in = phase;
    c.out = phase == kRing - 1 ? 0 : phase + 1;
    c.reverse = it.serpentine & ct.launch_no & 1;
    c.flagv = it.st->flag_base + (unsigned long long)ct.launch_no;
    return true;
}

template <class S> __device__ __forceinline__ S* ring_buf(const IterArgs<S>& it, int i) {
    return i == 0 ? it.buf[0] : (i == 1 ? it.buf[1] : it.buf[2]);
}

// Waits until the halo of v_m has landed.  On a timeout the error flag is raised (the control CTA turns it
// into STOP_ERROR) and the caller carries on.  Warp form: lane 0 spins, the warp follows (used mid-walk,
// where the flags have normally arrived long ago).  CTA form: thread 0 spins, the CTA follows.
template <class S>
__device__ __forceinline__ void halo_wait_warp(const IterArgs<S>& it, const LaunchCtx& c) {
    if ((threadIdx.x & 31) == 0) {
        const PeerArgs& pa = *it.peer;
        if (pa.stats) {
            // instrumented form: was the flag already there, and if not, how long did this warp spin
            bool there = true;
            for (int q = 0; q < pa.nranks; ++q)
                if (q != pa.rank && ((pa.halo_src_mask >> q) & 1u) && ld_acquire_sys(pa.flag_in + q) < c.flagv) there = false;
            atomicAdd(pa.stats + 5, 1ull);
            if (!there) {
                const unsigned long long t0 = global_ns();
                if (!wait_peer_flags(pa, pa.halo_src_mask, c.flagv)) *(volatile int*)&it.st->error = 1;
                const unsigned long long dt = global_ns() - t0;
                atomicAdd(pa.stats + 2, dt);
                atomicAdd(pa.stats + 3, 1ull);
                atomicMax(pa.stats + 4, dt);
            }
        } else if (!wait_peer_flags(pa, pa.halo_src_mask, c.flagv)) {
            *(volatile int*)&it.st->error = 1;
        }
    }
    __syncwarp();
}
template <class S>
__device__ __forceinline__ void halo_wait_cta(const IterArgs<S>& it, const LaunchCtx& c) {
    if (threadIdx.x == 0) {
        if (!wait_peer_flags(*it.peer, it.peer->halo_src_mask, c.flagv)) *(volatile int*)&it.st->error = 1;
    }
    __syncthreads();
}

// Contiguous send ranges of this launch, staged in shared memory once per CTA.
struct SendSm {
    int n;
    int start[kMaxPeers];
    int count[kMaxPeers];
    void* dst[kMaxPeers];
};

template <class S>
__device__ __forceinline__ void load_sends(const IterArgs<S>& it, const LaunchCtx& c, SendSm& sm) {
    if (threadIdx.x == 0) sm.n = it.mode == MODE_PEER ? it.peer->nsend : 0;
    if (it.mode == MODE_PEER && threadIdx.x < kMaxPeers) {
        sm.start[threadIdx.x] = it.peer->send_start[threadIdx.x];
        sm.count[threadIdx.x] = it.peer->send_count[threadIdx.x];
        sm.dst[threadIdx.x] = it.peer->send_dst[threadIdx.x][c.out];
    }
    __syncthreads();
}

// Row epilogue shared by every matvec kernel: z = sigma * (A v)_row.  A row that a peer's rows reference
// is also stored into that peer's halo segment (fire-and-forget NVLink store).
template <class S>
__device__ __forceinline__ void row_epilogue(const SendSm& sm, S z, S v_own, S* out, int row, double& are, double& aim,
                                             double& qn) {
    Sc<S>::dotc(v_own, z, are, aim);
    qn += Sc<S>::abs2(z);
    out[row] = z;
    for (int k = 0; k < sm.n; ++k) {
        const unsigned off = unsigned(row - sm.start[k]);
        if (off < unsigned(sm.count[k])) static_cast<S*>(sm.dst[k])[off] = z;
    }
}

// CTA sums -> this CTA's partial slot of the launch (read by the next launch's control CTA).
template <class S, int BLOCK>
__device__ __forceinline__ void cta_finish(const IterArgs<S>& it, const LaunchCtx& c, double are, double aim, double qn,
                                           int slot) {
    __shared__ double sred[3][BLOCK / 32];
    cta_sum3<BLOCK>(are, aim, qn, sred);
    if (threadIdx.x == 0) {
        double* part = it.partials + (size_t(c.in) * size_t(it.nslots) + size_t(slot)) * 3;
        part[0] = are;
        part[1] = aim;
        part[2] = qn;
    }
}

// Work-unit walk shared by the sparse kernels.  Units are visited in a rotated order that starts in the
// middle of the longest run of units that read no halo entry (it.unit_start, planned on the host), so the
// units that do read halo entries come up mid-kernel -- long after the peers' flags have arrived -- in
// both walking directions; a warp waits for the flags (once) right before its first such unit.  Without a
// deferred map a multi-GPU peer launch waits first.  Odd launches walk backwards (serpentine).
template <class S, class F>
__device__ __forceinline__ void walk_units(const IterArgs<S>& it, const LaunchCtx& c, int first, int stride, int nunits,
                                           F&& body) {
    const uint8_t* def = it.mode == MODE_PEER ? it.deferred : nullptr;
    bool waited = it.mode != MODE_PEER;
    if (!waited && !def) {
        halo_wait_cta<S>(it, c);
        waited = true;
    }
    for (int i = first; i < nunits; i += stride) {
        int unit = (c.reverse ? nunits - 1 - i : i) + it.unit_start;
        if (unit >= nunits) unit -= nunits;
        if (!waited && def[unit]) {
            halo_wait_warp<S>(it, c);
            waited = true;
        }
        body(unit);
    }
}

// =================================================================================================
// CSR, one thread per row.  Right for very short rows; loads of vals/colidx are strided across the
// warp but every fetched sector is consumed by the following loop trips out of L1.
// Work unit = BLOCK consecutive rows.
// =================================================================================================
template <class S, int BLOCK, bool PEER>
__global__ void __launch_bounds__(BLOCK) spmv_csr_scalar_kernel(CsrView<S> A, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    double are = 0.0, aim = 0.0, qn = 0.0;
    const int nunits = (A.rows + BLOCK - 1) / BLOCK;
    walk_units<S>(it, c, blockIdx.x - 1, gridDim.x - 1, nunits, [&](int unit) {
        const int row = unit * BLOCK + threadIdx.x;
        if (row < A.rows) {
            const int p0 = __ldg(A.rowptr + row), p1 = __ldg(A.rowptr + row + 1);
            S acc = Sc<S>::zero();
            for (int p = p0; p < p1; ++p) Sc<S>::mac(acc, __ldg(A.vals + p), ld_x<S, PEER>(u, __ldg(A.colidx + p)));
            row_epilogue<S>(sends, Sc<S>::scale(acc, c.s), ld_x<S, PEER>(u, row), out, row, are, aim, qn);
        }
    });
    cta_finish<S, BLOCK>(it, c, are, aim, qn, blockIdx.x - 1);
}

// =================================================================================================
// CSR, LANES lanes per row (LANES = 32 is warp-per-row).  Lanes stride over the row so each warp
// instruction reads LANES consecutive values / indices; partial sums meet in a shuffle tree.
// Work unit = BLOCK / LANES consecutive rows (one CTA-wide step).
// =================================================================================================
template <class S, int LANES, int BLOCK, bool PEER>
__global__ void __launch_bounds__(BLOCK) spmv_csr_vector_kernel(CsrView<S> A, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    double are = 0.0, aim = 0.0, qn = 0.0;
    constexpr int RPU = BLOCK / LANES;  // rows per unit
    const int sub = threadIdx.x % LANES;
    const int nunits = (A.rows + RPU - 1) / RPU;
    walk_units<S>(it, c, blockIdx.x - 1, gridDim.x - 1, nunits, [&](int unit) {
        const int row = unit * RPU + threadIdx.x / LANES;
        S acc = Sc<S>::zero();
        if (row < A.rows) {
            const int p0 = __ldg(A.rowptr + row), p1 = __ldg(A.rowptr + row + 1);
            for (int p = p0 + sub; p < p1; p += LANES)
                Sc<S>::mac(acc, ld_stream_val<S>(A.vals + p), ld_x<S, PEER>(u, ld_stream_s32(A.colidx + p)));
        }
#pragma unroll
        for (int d = LANES / 2; d > 0; d >>= 1) Sc<S>::add(acc, Sc<S>::shfl_down(acc, d, LANES));
        if (sub == 0 && row < A.rows)
            row_epilogue<S>(sends, Sc<S>::scale(acc, c.s), ld_x<S, PEER>(u, row), out, row, are, aim, qn);
    });
    cta_finish<S, BLOCK>(it, c, are, aim, qn, blockIdx.x - 1);
}

// =================================================================================================
// CSR streamed through shared memory (CSR-stream / CSR-adaptive shape).  A CTA owns a block of
// consecutive rows whose entries fit one tile: the entries are read as one contiguous stream
// (16-byte value loads, 8-byte index loads, fully coalesced), multiplied with the gathered x and
// parked in shared memory; then one thread per row sums its segment in ascending column order.
// A row longer than the tile is its own block and is reduced by the whole CTA.
// Work unit = one row block.
// =================================================================================================
template <class S> struct StreamCfg;
template <> struct StreamCfg<double> { static constexpr int TILE = 4096; };   // 32 KB of products
template <> struct StreamCfg<double2> { static constexpr int TILE = 2048; };  // 32 KB of products
constexpr int kStreamMaxRows = 512;

template <class S, int BLOCK, bool PEER>
__global__ void __launch_bounds__(BLOCK) spmv_csr_stream_kernel(CsrView<S> A, const int32_t* __restrict__ rowblk,
                                                                int nblk, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    constexpr int TILE = StreamCfg<S>::TILE;
    __shared__ __align__(16) S prod[TILE];
    __shared__ int32_t srow[kStreamMaxRows + 1];
    __shared__ double slong[3][BLOCK / 32];
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    double are = 0.0, aim = 0.0, qn = 0.0;
    const double* vals_d = reinterpret_cast<const double*>(A.vals);
    walk_units<S>(it, c, blockIdx.x - 1, gridDim.x - 1, nblk, [&](int b) {
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
                Sc<S>::mac(acc, ld_stream_val<S>(A.vals + p), ld_x<S, PEER>(u, ld_stream_s32(A.colidx + p)));
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
                row_epilogue<S>(sends, z, ld_x<S, PEER>(u, r0), out, r0, are, aim, qn);
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
                prod[e - base] = Sc<S>::mul(v0, ld_x<S, PEER>(u, cc.x));
                prod[e - base + 1] = Sc<S>::mul(v1, ld_x<S, PEER>(u, cc.y));
            }
            __syncthreads();
            for (int i = threadIdx.x; i < nr; i += BLOCK) {
                S acc = Sc<S>::zero();
                const int q1 = srow[i + 1] - base;
                for (int q = srow[i] - base; q < q1; ++q) Sc<S>::add(acc, prod[q]);
                row_epilogue<S>(sends, Sc<S>::scale(acc, c.s), ld_x<S, PEER>(u, r0 + i), out, r0 + i, are, aim, qn);
            }
        }
        __syncthreads();
    });
    cta_finish<S, BLOCK>(it, c, are, aim, qn, blockIdx.x - 1);
}

// =================================================================================================
// CSR with the x window of a row block staged in shared memory by the TMA engine ("where locality
// allows": banded / stencil matrices).  The host records for every unit of BLOCK rows the contiguous range
// [first, first+count) of local columns its rows reference (truncated to the shared-memory budget); one
// elected thread issues a single bulk copy (cp.async.bulk, completion on an mbarrier) of that range of v
// while all threads fetch their row pointers; products then read x from shared memory, and any column
// outside the window (halo entries, outliers) falls back to a global gather.  One thread per row.
// =================================================================================================
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    }
}

template <class S, int BLOCK, bool PEER>
__global__ void __launch_bounds__(BLOCK) spmv_csr_window_kernel(CsrView<S> A, const int2* __restrict__ win, IterArgs<S> it,
                                                                int phase) {
    extern __shared__ __align__(128) unsigned char window_smem[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ SendSm sends;
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    if (threadIdx.x == 0) mbar_init(&bar, 1);
    load_sends<S>(it, c, sends);  // ends with a CTA barrier: the mbarrier is initialised for everyone
    S* xs = reinterpret_cast<S*>(window_smem);
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    double are = 0.0, aim = 0.0, qn = 0.0;
    unsigned parity = 0;
    const int nunits = (A.rows + BLOCK - 1) / BLOCK;
    walk_units<S>(it, c, blockIdx.x - 1, gridDim.x - 1, nunits, [&](int unit) {
        const int2 w = __ldg(win + unit);  // {first column, count}; 16-byte aligned range, count may be 0
        if (w.y > 0 && threadIdx.x == 0) tma_load_1d(xs, u + w.x, unsigned(w.y) * unsigned(sizeof(S)), &bar);
        const int row = unit * BLOCK + threadIdx.x;
        int p0 = 0, p1 = 0;
        if (row < A.rows) {
            p0 = __ldg(A.rowptr + row);
            p1 = __ldg(A.rowptr + row + 1);
        }
        if (w.y > 0) {
            mbar_wait(&bar, parity);
            parity ^= 1u;
        }
        if (row < A.rows) {
            S acc = Sc<S>::zero();
            for (int p = p0; p < p1; ++p) {
                const int col = __ldg(A.colidx + p);
                const unsigned k = unsigned(col - w.x);
                const S x = k < unsigned(w.y) ? xs[k] : ld_x<S, PEER>(u, col);
                Sc<S>::mac(acc, __ldg(A.vals + p), x);
            }
            const unsigned kd = unsigned(row - w.x);
            const S own = kd < unsigned(w.y) ? xs[kd] : ld_x<S, PEER>(u, row);
            row_epilogue<S>(sends, Sc<S>::scale(acc, c.s), own, out, row, are, aim, qn);
        }
        __syncthreads();  // everyone is done reading the window before the next bulk copy lands in it
    });
    cta_finish<S, BLOCK>(it, c, are, aim, qn, blockIdx.x - 1);
}

// =================================================================================================
// SELL: a warp owns a slice of 32*R consecutive rows, lane l the R rows  slice*32R + l*R + {0..R-1}.
// Entry j of those rows is stored at  slice_off + (j*32 + l)*R : one warp instruction reads 32*R
// consecutive values (R=2, f64: 16-byte loads; c128: 16 bytes per value) and 32*R consecutive column
// indices, and for stencil-like rows the x gathers of neighbouring lanes fall into the same lines.
// No shared memory, no block barrier in the main loop.  Work unit = one slice (per warp).
// =================================================================================================
template <class S, int R, int BLOCK, bool PEER>
__global__ void __launch_bounds__(BLOCK) spmv_sell_kernel(SellView<S> A, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    double are = 0.0, aim = 0.0, qn = 0.0;
    const int lane = threadIdx.x & 31;
    const int gwarp = ((blockIdx.x - 1) * BLOCK + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x - 1) * (BLOCK / 32);
    walk_units<S>(it, c, gwarp, nwarps, A.nslices, [&](int sl) {
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
                if (col[r] >= 0) Sc<S>::mac(acc[r], v[r], ld_x<S, PEER>(u, col[r]));
            vp += 32 * R;
            cp += 32 * R;
        }
        const int row = (sl * 32 + lane) * R;
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (row + r < A.rows)
                row_epilogue<S>(sends, Sc<S>::scale(acc[r], c.s), ld_x<S, PEER>(u, row + r), out, row + r, are, aim, qn);
    });
    cta_finish<S, BLOCK>(it, c, are, aim, qn, blockIdx.x - 1);
}

// =================================================================================================
// Dense column-major GEMV.  CTA (tile, seg) owns RT = BLOCK*R rows (R = 2 for f64 -> 16-byte loads
// down a column, 1 for c128) and the columns of segment seg.  v of the segment is staged in shared
// memory once per CTA.  Segment sums go to partial[seg][row]; the last CTA of a row tile adds them in
// segment order (deterministic), runs the row epilogue and leaves the tile's sums in the tile's slot.
// =================================================================================================
constexpr int kGemvChunk = 512;  // columns of x staged per pass
template <class S, int VAR> struct GemvRows { static constexpr int value = (Sc<S>::is_complex ? 1 : 2) * (VAR == 2 ? 2 : 1); };  // VAR 3 = VAR 0 rows

// Sum of the nseg segment partials of R consecutive rows, eight segments in flight, added in segment order.  Kept out
// of line on purpose: inlined into the kernel, its registers ended up costing loads in flight in the streaming loop
// (107 vs 91 us on C3).
template <class S, int R>
__device__ __noinline__ void gemv_combine_rows(const S* __restrict__ pp, int64_t lda, int nseg, S (&z)[R]) {
#pragma unroll
    for (int k = 0; k < R; ++k) z[k] = Sc<S>::zero();
    int s0 = 0;
    for (; s0 + 8 <= nseg; s0 += 8) {
        S t[8][R];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int k = 0; k < R; ++k) t[i][k] = Sc<S>::ldcg(pp + int64_t(s0 + i) * lda + k);
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int k = 0; k < R; ++k) Sc<S>::add(z[k], t[i][k]);
    }
    for (; s0 < nseg; ++s0)
#pragma unroll
        for (int k = 0; k < R; ++k) Sc<S>::add(z[k], Sc<S>::ldcg(pp + int64_t(s0) * lda + k));
}

// VAR 0: x staged 512 columns at a time, 8 column loads in flight per thread.  VAR 3: the same loop with the combine of
// the last CTA done by gemv_combine_rows (eight partial loads in flight, out of line).  VAR 1: 64 columns, 4 loads in flight
// (fewer registers).  A.tile_major selects which neighbours a CTA has in the launch order: 0 = the column segments of
// one row tile, 1 = the row tiles of one column segment (adjacent CTAs then read adjacent 4 KB pieces of the SAME
// columns, i.e. whole contiguous columns between them).
template <class S, int BLOCK, bool PEER, int VAR>
__global__ void __launch_bounds__(BLOCK) gemv_colmajor_kernel(DenseView<S> A, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    constexpr int R = GemvRows<S, VAR>::value;  // rows per thread: 16 bytes (VAR 0, 1) or 32 bytes (VAR 2) per column
    constexpr int RT = BLOCK * R;
    constexpr int CHUNK = (VAR == 0 || VAR == 3) ? kGemvChunk : (VAR == 1 ? 64 : 128);
    constexpr int UNROLL = (VAR == 0 || VAR == 3) ? 8 : 4;
    __shared__ __align__(16) S xs[CHUNK];
    __shared__ int s_tile_last;
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);
    if (it.mode == MODE_PEER) halo_wait_cta<S>(it, c);  // every row reads the whole halo
    typename XPtr<S, PEER>::type u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    const int cta = blockIdx.x - 1;
    const int tile = A.tile_major ? cta % A.ntiles : cta / A.nseg;
    const int seg = A.tile_major ? cta / A.ntiles : cta % A.nseg;
    const int c0 = int(int64_t(A.cols) * seg / A.nseg), c1 = int(int64_t(A.cols) * (seg + 1) / A.nseg);
    const int r = tile * RT + threadIdx.x * R;
    const bool live = r < A.lda;  // lda is even and rows beyond A.rows are zero-filled on the device
    S acc[R];
#pragma unroll
    for (int k = 0; k < R; ++k) acc[k] = Sc<S>::zero();
    for (int cb = c0; cb < c1; cb += CHUNK) {
        const int nc = min(CHUNK, c1 - cb);
        __syncthreads();
        for (int j = threadIdx.x; j < nc; j += BLOCK) xs[j] = Sc<S>::scale(ld_x<S, PEER>(u, cb + j), c.s);
        __syncthreads();
        if (live) {
            const S* col = A.a + int64_t(cb) * A.lda + r;
#pragma unroll UNROLL
            for (int j = 0; j < nc; ++j) {
                if constexpr (R * sizeof(S) == 32) {
                    // one 256-bit load per column (SASS LDG.E.ENL2.256): 4 real or 2 complex rows
                    double a[4];
                    ld_stream_f64x4(reinterpret_cast<const double*>(col), a);
                    if constexpr (Sc<S>::is_complex) {
                        Sc<S>::mac(acc[0], make_double2(a[0], a[1]), xs[j]);
                        Sc<S>::mac(acc[1], make_double2(a[2], a[3]), xs[j]);
                    } else {
#pragma unroll
                        for (int k = 0; k < 4; ++k) acc[k] = fma(a[k], xs[j], acc[k]);
                    }
                } else if constexpr (R == 2) {
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
    // The tail of the launch: the last CTA of a row tile adds the segment partials (L2-resident) in segment order.
    double are = 0.0, aim = 0.0, qn = 0.0;
    if constexpr (VAR == 3) {
        if (live) {
            S z[R];
            gemv_combine_rows<S, R>(A.partial + r, A.lda, A.nseg, z);
#pragma unroll
            for (int k = 0; k < R; ++k)
                if (r + k < A.rows) row_epilogue<S>(sends, z[k], ld_x<S, PEER>(u, r + k), out, r + k, are, aim, qn);
        }
    } else {
#pragma unroll
        for (int k = 0; k < R; ++k) {
            const int row = r + k;
            if (row < A.rows) {
                S z = Sc<S>::zero();
                for (int sgi = 0; sgi < A.nseg; ++sgi) Sc<S>::add(z, Sc<S>::ldcg(A.partial + int64_t(sgi) * A.lda + row));
                row_epilogue<S>(sends, z, ld_x<S, PEER>(u, row), out, row, are, aim, qn);
            }
        }
    }
    cta_finish<S, BLOCK>(it, c, are, aim, qn, tile);
}

// =================================================================================================
// Dense GEMV for a matrix that equals its own (conjugate) transpose: y_j = sum_i op(A[i,j]) x_i, op = conj for a
// Hermitian matrix, identity for a symmetric one -- the dot product of the STORED column j with x.  The column-major
// array is then read as one flat stream: CTA c owns the elements [c * per, (c + 1) * per) (a multiple of 128, at least
// one column long) and walks them column by column in steps of 8 KB of consecutive memory (256 threads x 32-byte loads,
// SASS LDG.E.ENL2.256), x coming from L1; a column's sum is a shuffle tree per warp plus eight warp sums parked in shared
// memory -- one barrier per launch, no column-segment partials of the whole vector, no last-CTA combine.  All CTAs carry
// the same number of bytes, so they finish together.  Only the column a range starts or ends in the middle of is shared
// by two CTAs: its two partial sums are added, in row order, by whichever CTA arrives second (one ticket per column).
// (A first version gave every WARP its own range: 4736 concurrent 4 KB streams ran 3 % slower than the general kernel.)  The engine selects this kernel when a bitwise symmetry check at upload succeeds
// (sym_check_kernel); everything else of the launch (control CTA, state machine, epilogue sums) is the common frame.
// One GPU only (the row blocks of a sharded dense matrix are not symmetric).
// =================================================================================================
template <class S> struct SymDenseView {
    int32_t n;            // order of the matrix
    int64_t lda;          // device leading dimension (multiple of 16, rows >= n are zero)
    const S* a;           // column-major
    int64_t per;          // elements per CTA (multiple of 128, >= lda)
    int32_t nctas;        // CTAs that own a range
    S* partial;           // [n][2]: partial sums of the columns shared by two warps (head part, tail part)
    unsigned* ticket;     // [n]
    int32_t conj;         // 1: Hermitian (use conj of the stored entries)
};

template <class S> __device__ __forceinline__ void ld_stream_row4(const S* p, S (&o)[32 / sizeof(S)]);
template <> __device__ __forceinline__ void ld_stream_row4<double>(const double* p, double (&o)[4]) { ld_stream_f64x4(p, o); }
template <> __device__ __forceinline__ void ld_stream_row4<double2>(const double2* p, double2 (&o)[2]) {
    double t[4];
    ld_stream_f64x4(reinterpret_cast<const double*>(p), t);
    o[0] = make_double2(t[0], t[1]);
    o[1] = make_double2(t[2], t[3]);
}

template <class S, int BLOCK>
__global__ void __launch_bounds__(BLOCK) gemv_sym_kernel(SymDenseView<S> A, IterArgs<S> it, int phase) {
    pdl_prologue();
    if (blockIdx.x == 0) {
        control_cta<S, BLOCK>(it, true);
        return;
    }
    constexpr int V = 32 / int(sizeof(S));  // elements per 32-byte load: 4 real, 2 complex
    constexpr int STEP = BLOCK * V;         // elements per CTA-wide step: 8 KB of consecutive matrix
    constexpr int NW = BLOCK / 32;
    extern __shared__ __align__(16) unsigned char sym_smem[];
    S* wsum = reinterpret_cast<S*>(sym_smem);  // [columns of this CTA's range][NW]: per-warp sums, no barrier per column
    __shared__ SendSm sends;
    LaunchCtx c;
    if (!launch_begin<S>(it, phase, c)) return;
    load_sends<S>(it, c, sends);  // (no sends on one GPU; keeps the common epilogue)
    const S* __restrict__ u = ring_buf<S>(it, c.in);
    S* __restrict__ out = ring_buf<S>(it, c.out);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cta = int(blockIdx.x) - 1;
    const int64_t total = A.lda * int64_t(A.n);
    const int64_t e_begin = int64_t(cta) * A.per, e_end = min(total, e_begin + A.per);
    const int col_first = int(e_begin / A.lda);
    int ncols = 0;
    for (int64_t e = e_begin; e < e_end; ++ncols) {
        const int col = col_first + ncols;
        const int64_t r0 = e - int64_t(col) * A.lda;
        const int64_t r1 = min(A.lda, r0 + (e_end - e));  // rows [r0, r1) of column `col`
        const S* ap = A.a + int64_t(col) * A.lda;
        S acc[V];
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = Sc<S>::zero();
#pragma unroll 4
        for (int64_t i = r0 + int64_t(threadIdx.x) * V; i + V <= r1; i += STEP) {
            S a[V], x[V];
            ld_stream_row4<S>(ap + i, a);
            {
                // the matching 32 bytes of x, through L1 (x is read by every CTA over and over)
                const double2* xp = reinterpret_cast<const double2*>(u + i);
                const double2 xa = __ldg(xp), xb = __ldg(xp + 1);
                if constexpr (Sc<S>::is_complex) {
                    x[0] = xa;
                    x[1] = xb;
                } else {
                    x[0] = xa.x;
                    x[1] = xa.y;
                    x[2] = xb.x;
                    x[3] = xb.y;
                }
            }
#pragma unroll
            for (int k = 0; k < V; ++k) {
                if constexpr (Sc<S>::is_complex) {
                    if (A.conj) a[k].y = -a[k].y;
                }
                Sc<S>::mac(acc[k], a[k], x[k]);
            }
        }
        S sum = acc[0];
#pragma unroll
        for (int k = 1; k < V; ++k) Sc<S>::add(sum, acc[k]);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) Sc<S>::add(sum, Sc<S>::shfl_down(sum, d, 32));
        if (lane == 0) wsum[ncols * NW + warp] = sum;
        e += r1 - r0;
    }
    __syncthreads();
    // thread t finishes column col_first + t of the range: the NW warp sums in warp order, then -- for the (at most two)
    // columns shared with a neighbouring CTA -- head + tail in row order by whichever CTA arrives second
    double are = 0.0, aim = 0.0, qn = 0.0;
    for (int t = threadIdx.x; t < ncols; t += BLOCK) {
        S z = Sc<S>::zero();
#pragma unroll
        for (int w = 0; w < NW; ++w) Sc<S>::add(z, wsum[t * NW + w]);
        const int col = col_first + t;
        const int64_t cs = int64_t(col) * A.lda;
        const bool head = cs >= e_begin, tail = cs + A.lda <= e_end;  // the range holds the column's first / last row
        bool finalize = head && tail;
        if (!finalize) {
            A.partial[2 * int64_t(col) + (head ? 0 : 1)] = z;
            __threadfence();
            const unsigned tk = atomicAdd(A.ticket + col, 1u);
            if (tk == 1u) {
                A.ticket[col] = 0;
                __threadfence();
                z = Sc<S>::ldcg(A.partial + 2 * int64_t(col));
                Sc<S>::add(z, Sc<S>::ldcg(A.partial + 2 * int64_t(col) + 1));
                finalize = true;
            }
        }
        if (finalize) row_epilogue<S>(sends, Sc<S>::scale(z, c.s), Sc<S>::ldg(u + col), out, col, are, aim, qn);
    }
    cta_finish<S, BLOCK>(it, c, are, aim, qn, cta);
}

// Bitwise test A == A^T (flags bit 0 cleared if not) and A == A^H (bit 1 cleared if not) of a column-major n x n matrix:
// 32 x 32 tiles of the upper triangle against their mirror tiles, both read coalesced through shared memory.
template <class S>
__global__ void __launch_bounds__(256) sym_check_kernel(const S* __restrict__ a, int n, int64_t lda, int* flags) {
    __shared__ S tile[32][33];
    const int tj = blockIdx.x, ti = blockIdx.y;
    if (ti > tj) return;  // upper triangle of tiles (ti <= tj)
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
    // mirror tile (rows of tile tj, columns of tile ti), element (r, c) = a[tj*32 + r + (ti*32 + c) * lda]
    for (int cc = ty; cc < 32; cc += 8) {
        const int r = tj * 32 + tx, cidx = ti * 32 + cc;
        tile[cc][tx] = (r < n && cidx < n) ? a[r + int64_t(cidx) * lda] : Sc<S>::zero();
    }
    __syncthreads();
    int bad_sym = 0, bad_herm = 0;
    for (int cc = ty; cc < 32; cc += 8) {
        const int r = ti * 32 + tx, cidx = tj * 32 + cc;  // element (r, cidx) of the upper tile; its mirror is tile[tx][cc]
        if (r < n && cidx < n) {
            const S v = a[r + int64_t(cidx) * lda];
            const S m = tile[tx][cc];
            if constexpr (Sc<S>::is_complex) {
                if (!(v.x == m.x && v.y == m.y)) bad_sym = 1;
                if (!(v.x == m.x && v.y == -m.y)) bad_herm = 1;
            } else {
                if (!(v == m)) bad_sym = bad_herm = 1;
            }
        }
    }
    if (bad_sym) atomicAnd(flags, ~1);
    if (bad_herm) atomicAnd(flags, ~2);
}

// =================================================================================================
// Small helper kernels.
// =================================================================================================
// MODE_SPLIT: fixed-order sum of the CTA partials of launch j (ring slot) into scal[0..2]; NCCL
// all-reduces scal in place and the next launch's control CTA folds it.
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) reduce_partials_kernel(const double* partials, int nslots, double* scal) {
    __shared__ double sred[3][BLOCK / 32];
    double a = 0.0, b = 0.0, d = 0.0;
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

// max |a| over an array of doubles (complex: over re and im parts), as the bit pattern of a non-negative double
__global__ void absmax_kernel(const double* __restrict__ a, int64_t count, unsigned long long* out) {
    double m = 0.0;
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < count; i += int64_t(gridDim.x) * blockDim.x) {
        const double v = fabs(a[i]);
        m = v > m ? v : m;  // NaNs are skipped
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const double o = __shfl_down_sync(0xffffffffu, m, d);
        m = o > m ? o : m;
    }
    if ((threadIdx.x & 31) == 0) atomicMax(out, (unsigned long long)__double_as_longlong(m));
}

// x = y / normY (power_method.hpp:78, true division) of the result buffer -> contiguous output.
template <class S>
__global__ void normalize_result_kernel(const S* buf0, const S* buf1, const S* buf2, const PowerState* st, S* x, int n) {
    const int fb = st->final_buf;
    const S* src = fb == 0 ? buf0 : (fb == 1 ? buf1 : buf2);
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
