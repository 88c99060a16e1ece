// block.cuh -- device code of the BLOCK power iteration: k vectors per pass over A.
//
// The single-vector path streams A (12 bytes per entry for fp64 CSR) to advance ONE vector; here the same
// stream advances K of them, so the bytes per vector-trip fall from  12 nnz + 16 n  to  12 nnz / K + 16 n
// (C5, K = 4: 1.74 GB -> 0.64 GB per vector).  The block U is stored ROW-major, n x K: the K operands a
// matrix entry needs are one contiguous 8K-byte piece (one 32-byte sector for K = 4), and a thread's K
// results are one contiguous store.
//
// One iteration = one SpMM launch + one single-CTA control launch (blockmath.hpp):
//     launch L:   U_{L+1}[i,:] = sigma * (sum_j a_ij U_L[j,:]) T_L          (= sigma A Q_L,  Q_L = U_L T_L)
//                 per CTA partial sums of  C = Q_L^H U_{L+1}  (k x k)  and  W = U_{L+1}^H U_{L+1}  (upper)
//     control:    folds the partials in slot order (deterministic), H = C / sigma, Ritz pairs, convergence
//                 test of power_method.hpp:83-91 on each wanted value, W = R^H R, T_{L+1} = R^{-1}
// Nothing n-sized is read twice and the host only polls the stop flag every few trips.
#pragma once
#include "blockmath.hpp"
#include "kernels.cuh"

namespace eigsol_b200 {

template <class S> struct BlockArgs {
    S* u[2];            // the two n x K row-major blocks; launch L reads u[L & 1], writes u[(L + 1) & 1]
    BlockState* st;
    double* partials;   // [entry][slot]
    int nslots;
    int chunk;          // SpMM kernels: units are dealt to the CTAs in runs of `chunk` consecutive units
};

// Multi-GPU wiring of the block iteration (one rank; nullptr on one GPU).  The block is row-sharded like the single
// vector: every rank holds [own rows | halo rows] x K, the SpMM writes the own rows of the next block, a small push
// kernel copies the rows the peers reference into THEIR halo segments (NVLink stores), and the control launch --
// which is a synchronisation point of the iteration anyway -- all-reduces the k x k Gram sums through mailboxes in
// peer memory (rank-ordered sum: the same bits, hence the same decisions, on every rank) and doubles as the barrier
// that says "every peer's halo rows of this trip have landed".
constexpr int kBlockMailStride = 2 * (kBlockElems + kBlockMax * (kBlockMax + 1) / 2);  // doubles per rank and parity

struct BlockPeer {
    int nranks, rank;
    double* mail_in;                              // local [2][kMaxPeers][kBlockMailStride]
    double* peer_mail[kMaxPeers];                 // remote mail_in base of rank q (nullptr for self)
    unsigned long long* flag_in;                  // local [kMaxPeers]
    unsigned long long* peer_flag[kMaxPeers];     // remote &flag_in[rank] of rank q
    long long spin_limit;
    // halo pushes: rows of the freshly written block that peer ranks read
    int nsend;
    int send_start[kMaxPeers], send_count[kMaxPeers];
    void* send_dst[kMaxPeers][2];                 // peer's halo segment, per block buffer (bytes address)
    int nlist;
    int list_off[kMaxPeers], list_count[kMaxPeers];
    void* list_dst[kMaxPeers][2];
    const int32_t* send_idx;
};

// Rows of block buffer `which` (the one the SpMM of this trip just wrote) that peers reference -> their halo segments.
template <class S, int K>
__global__ void __launch_bounds__(256) block_halo_push_kernel(const BlockPeer* pp, const S* __restrict__ u, int which) {
    pdl_prologue();
    const BlockPeer& p = *pp;
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (int k = 0; k < p.nsend; ++k) {
        S* dst = static_cast<S*>(p.send_dst[k][which]);
        const S* src = u + int64_t(p.send_start[k]) * K;
        const int64_t cnt = int64_t(p.send_count[k]) * K;
        for (int64_t i = tid; i < cnt; i += nth) dst[i] = src[i];
    }
    for (int k = 0; k < p.nlist; ++k) {
        S* dst = static_cast<S*>(p.list_dst[k][which]);
        const int32_t* idx = p.send_idx + p.list_off[k];
        const int64_t cnt = int64_t(p.list_count[k]) * K;
        for (int64_t i = tid; i < cnt; i += nth) dst[i] = u[int64_t(idx[i / K]) * K + (i % K)];
    }
}

// Block-cyclic unit order: CTA b works on units (g * grid + b) * chunk + j, j = 0..chunk-1, g = 0, 1, ...
// Consecutive units of a run share their near-diagonal operand rows through L1 (a stencil's +-1 line
// neighbours), while the CTAs running at the same time stay within grid * chunk units of each other, so the
// far neighbours (+-1 plane) are still shared through L2.  -1 past the end.
__device__ __forceinline__ int block_unit_at(int it, int chunk, int nunits) {
    const int g = it / chunk, j = it - g * chunk;
    const long long u = (long long)(g * int(gridDim.x) + int(blockIdx.x)) * chunk + j;
    return u < nunits ? int(u) : -1;
}

// number of accumulated sums: C (K x K) and the upper triangle of W
template <int K> struct BlockSums {
    static constexpr int NC = K * K;
    static constexpr int NW = K * (K + 1) / 2;
    static constexpr int NE = NC + NW;
};
template <class S, int K> struct BlockStride { static constexpr int value = (Sc<S>::is_complex ? 2 : 1) * BlockSums<K>::NE; };

template <class S, int K> struct BlockAcc {
    S c[BlockSums<K>::NC];
    S w[BlockSums<K>::NW];
    __device__ __forceinline__ void clear() {
#pragma unroll
        for (int i = 0; i < BlockSums<K>::NC; ++i) c[i] = Sc<S>::zero();
#pragma unroll
        for (int i = 0; i < BlockSums<K>::NW; ++i) w[i] = Sc<S>::zero();
    }
};

// acc += conj(x) * z
__device__ __forceinline__ void cmac(double& acc, double x, double z) { acc = fma(x, z, acc); }
__device__ __forceinline__ void cmac(double2& acc, double2 x, double2 z) {
    acc.x = fma(x.x, z.x, acc.x);
    acc.x = fma(x.y, z.y, acc.x);
    acc.y = fma(x.x, z.y, acc.y);
    acc.y = fma(-x.y, z.x, acc.y);
}

// ---- one row of a row-major n x K block ----------------------------------------------------------------
template <int K> __device__ __forceinline__ void load_block_row(const double* __restrict__ U, int64_t r, double (&out)[K]) {
    if constexpr (K == 1) {
        out[0] = __ldg(U + r);
    } else {
        const double2* p = reinterpret_cast<const double2*>(U + r * K);
#pragma unroll
        for (int b = 0; b < K / 2; ++b) {
            const double2 t = __ldg(p + b);
            out[2 * b] = t.x;
            out[2 * b + 1] = t.y;
        }
    }
}
template <int K> __device__ __forceinline__ void load_block_row(const double2* __restrict__ U, int64_t r, double2 (&out)[K]) {
#pragma unroll
    for (int b = 0; b < K; ++b) out[b] = __ldg(U + r * K + b);
}
template <int K> __device__ __forceinline__ void store_block_row(double* __restrict__ U, int64_t r, const double (&v)[K]) {
    if constexpr (K == 1) {
        U[r] = v[0];
    } else {
        double2* p = reinterpret_cast<double2*>(U + r * K);
#pragma unroll
        for (int b = 0; b < K / 2; ++b) p[b] = make_double2(v[2 * b], v[2 * b + 1]);
    }
}
template <int K> __device__ __forceinline__ void store_block_row(double2* __restrict__ U, int64_t r, const double2 (&v)[K]) {
#pragma unroll
    for (int b = 0; b < K; ++b) U[r * K + b] = v[b];
}

__device__ __forceinline__ double cx_to(const Cx& t, double*) { return t.re; }
__device__ __forceinline__ double2 cx_to(const Cx& t, double2*) { return make_double2(t.re, t.im); }

// T_L (upper triangular) of the state into shared memory, K x K compact
template <class S, int K> __device__ __forceinline__ void load_T(const BlockState* st, int L, S* sT) {
    if (threadIdx.x < K * K) {
        const int a = threadIdx.x / K, b = threadIdx.x % K;
        sT[threadIdx.x] = cx_to(st->T[L & 1][a * kBlockLd + b], static_cast<S*>(nullptr));
    }
}

// y = (A U_L)[row,:] and own = U_L[row,:]  ->  v = sigma * y T (stored by the caller), q = own T, and the sums
template <class S, int K>
__device__ __forceinline__ void block_row_epilogue(const S (&y)[K], const S (&own)[K], const S* __restrict__ sT, double sigma,
                                                   S (&v)[K], BlockAcc<S, K>& acc) {
    S q[K];
#pragma unroll
    for (int b = 0; b < K; ++b) {
        S qb = Sc<S>::zero(), vb = Sc<S>::zero();
#pragma unroll
        for (int a = 0; a <= b; ++a) {
            const S t = sT[a * K + b];
            Sc<S>::mac(qb, own[a], t);
            Sc<S>::mac(vb, y[a], t);
        }
        q[b] = qb;
        v[b] = Sc<S>::scale(vb, sigma);
    }
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
        for (int b = 0; b < K; ++b) cmac(acc.c[a * K + b], q[a], v[b]);
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
        for (int b = a; b < K; ++b) cmac(acc.w[a * K - a * (a - 1) / 2 + (b - a)], v[a], v[b]);
}

__device__ __forceinline__ double comp_of(double v, int) { return v; }
__device__ __forceinline__ double comp_of(double2 v, int c) { return c ? v.y : v.x; }

// CTA reduction of the accumulators into partials[entry][slot]; fixed shuffle tree, fixed warp order.
// Entry order: re of C (row-major), re of W (upper, row-major) [, then the imaginary parts].
template <class S, int K, int BLOCK>
__device__ __forceinline__ void block_cta_reduce(const BlockAcc<S, K>& acc, double* __restrict__ partials, int nslots, int slot) {
    constexpr int NE = BlockSums<K>::NE, NC = BlockSums<K>::NC, STRIDE = BlockStride<S, K>::value;
    __shared__ double sred[BLOCK / 32][STRIDE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int comp = 0; comp < STRIDE / NE; ++comp) {
#pragma unroll
        for (int e = 0; e < NE; ++e) {
            double v = e < NC ? comp_of(acc.c[e < NC ? e : 0], comp) : comp_of(acc.w[e >= NC ? e - NC : 0], comp);
            v = warp_sum(v);
            if (lane == 0) sred[warp][comp * NE + e] = v;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < STRIDE; i += BLOCK) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < BLOCK / 32; ++w) s += sred[w][i];
        partials[int64_t(i) * nslots + slot] = s;
    }
}

// =================================================================================================
// SpMM, CSR, one thread per row, K right-hand sides.  Persistent grid; unit = BLOCK consecutive rows.
// =================================================================================================
template <class S, int K, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmm_csr_block_kernel(CsrView<S> A, BlockArgs<S> ba) {
    pdl_prologue();
    __shared__ S sT[K * K];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const int L = st->launch_no;
    const double sigma = st->sigma;
    const S* __restrict__ U = ba.u[L & 1];
    S* __restrict__ out = ba.u[(L + 1) & 1];
    load_T<S, K>(st, L, sT);
    __syncthreads();
    BlockAcc<S, K> acc;
    acc.clear();
    const int nunits = (A.rows + BLOCK - 1) / BLOCK;
    for (int it = 0, unit; (unit = block_unit_at(it, ba.chunk, nunits)) >= 0; ++it) {
        const int row = unit * BLOCK + threadIdx.x;
        if (row < A.rows) {
            const int p0 = __ldg(A.rowptr + row), p1 = __ldg(A.rowptr + row + 1);
            S y[K];
#pragma unroll
            for (int b = 0; b < K; ++b) y[b] = Sc<S>::zero();
            for (int p = p0; p < p1; ++p) {
                const S a = __ldg(A.vals + p);
                S u[K];
                load_block_row<K>(U, __ldg(A.colidx + p), u);
#pragma unroll
                for (int b = 0; b < K; ++b) Sc<S>::mac(y[b], a, u[b]);
            }
            S own[K], v[K];
            load_block_row<K>(U, row, own);
            block_row_epilogue<S, K>(y, own, sT, sigma, v, acc);
            store_block_row<K>(out, row, v);
        }
    }
    block_cta_reduce<S, K, BLOCK>(acc, ba.partials, ba.nslots, blockIdx.x);
}

// =================================================================================================
// SpMM, CSR, K lanes per row: lane b of a row's lane group owns column b of the block.  A matrix entry is
// loaded once per group (same address in K lanes = one transaction) and its K operands U[col, 0..K) are one
// coalesced 8K-byte piece across the group.  A lane keeps ONE running product and 2K Gram sums, so the
// register count stays near 64 for every K (the thread-per-row kernel above needs 100 registers at K = 4 and
// spills at K = 8: too few warps in flight to cover the col -> U[col] load chain), and the entry loop is
// unrolled UNROLL-wide with all gathers of a chunk issued before the first use.
// Work unit = BLOCK / K consecutive rows.  The epilogue moves y and own across the group with shuffles.
// =================================================================================================
template <class S> __device__ __forceinline__ S shfl_grp(S v, int src, int width);
template <> __device__ __forceinline__ double shfl_grp<double>(double v, int src, int width) {
    return __shfl_sync(0xffffffffu, v, src, width);
}
template <> __device__ __forceinline__ double2 shfl_grp<double2>(double2 v, int src, int width) {
    return make_double2(__shfl_sync(0xffffffffu, v.x, src, width), __shfl_sync(0xffffffffu, v.y, src, width));
}
__device__ __forceinline__ double shfl_xor_d(double v, int d) { return __shfl_xor_sync(0xffffffffu, v, d); }

// Sums held as: lane (g, b) has c[a] = sum over its rows of conj(q_a) v_b and w[a] = sum of conj(v_a) v_b.
template <class S, int K, int BLOCK>
__device__ __forceinline__ void block_cta_reduce_lanes(const S (&c)[K], const S (&w)[K], double* __restrict__ partials,
                                                       int nslots, int slot) {
    constexpr int NE = BlockSums<K>::NE, NC = BlockSums<K>::NC, STRIDE = BlockStride<S, K>::value;
    __shared__ double sred[BLOCK / 32][STRIDE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, b = lane % K;
#pragma unroll
    for (int comp = 0; comp < STRIDE / NE; ++comp) {
#pragma unroll
        for (int a = 0; a < K; ++a) {
            double vc = comp_of(c[a], comp), vw = comp_of(w[a], comp);
#pragma unroll
            for (int d = 16; d >= K; d >>= 1) {  // lanes with the same b: b, b + K, b + 2K, ...
                vc += shfl_xor_d(vc, d);
                vw += shfl_xor_d(vw, d);
            }
            if (lane < K) {
                sred[warp][comp * NE + a * K + b] = vc;
                if (a <= b) sred[warp][comp * NE + NC + a * K - a * (a - 1) / 2 + (b - a)] = vw;
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < STRIDE; i += BLOCK) {
        double s = 0.0;
#pragma unroll
        for (int wp = 0; wp < BLOCK / 32; ++wp) s += sred[wp][i];
        partials[int64_t(i) * nslots + slot] = s;
    }
}

template <class S, int K, int BLOCK, int UNROLL>
__global__ void __launch_bounds__(BLOCK) spmm_csr_lanes_kernel(CsrView<S> A, BlockArgs<S> ba) {
    pdl_prologue();
    constexpr int RPU = BLOCK / K;  // rows per unit
    __shared__ S sT[K * K];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const int L = st->launch_no;
    const double sigma = st->sigma;
    const S* __restrict__ U = ba.u[L & 1];
    S* __restrict__ out = ba.u[(L + 1) & 1];
    load_T<S, K>(st, L, sT);
    __syncthreads();
    const int grp = threadIdx.x / K, b = threadIdx.x % K;
    S tcol[K];  // column b of T (zero below the diagonal)
#pragma unroll
    for (int a = 0; a < K; ++a) tcol[a] = sT[a * K + b];
    S cacc[K], wacc[K];
#pragma unroll
    for (int a = 0; a < K; ++a) cacc[a] = wacc[a] = Sc<S>::zero();
    const int nunits = (A.rows + RPU - 1) / RPU;
    for (int it = 0, unit; (unit = block_unit_at(it, ba.chunk, nunits)) >= 0; ++it) {
        const int row = unit * RPU + grp;
        const bool valid = row < A.rows;
        const int safe = valid ? row : 0;
        const int p0 = valid ? __ldg(A.rowptr + row) : 0, p1 = valid ? __ldg(A.rowptr + row + 1) : 0;
        S y = Sc<S>::zero();
        for (int p = p0; p < p1; p += UNROLL) {
            int col[UNROLL];
            S a[UNROLL], u[UNROLL];
#pragma unroll
            for (int i = 0; i < UNROLL; ++i) {
                const bool ok = p + i < p1;
                col[i] = ok ? __ldg(A.colidx + p + i) : safe;
                a[i] = ok ? __ldg(A.vals + p + i) : Sc<S>::zero();
            }
#pragma unroll
            for (int i = 0; i < UNROLL; ++i) u[i] = __ldg(U + int64_t(col[i]) * K + b);
#pragma unroll
            for (int i = 0; i < UNROLL; ++i) Sc<S>::mac(y, a[i], u[i]);
        }
        const S own = valid ? __ldg(U + int64_t(row) * K + b) : Sc<S>::zero();
        S q = Sc<S>::zero(), v = Sc<S>::zero();
#pragma unroll
        for (int a = 0; a < K; ++a) {
            Sc<S>::mac(v, shfl_grp<S>(y, a, K), tcol[a]);
            Sc<S>::mac(q, shfl_grp<S>(own, a, K), tcol[a]);
        }
        v = Sc<S>::scale(v, sigma);
        if (valid) out[int64_t(row) * K + b] = v;
#pragma unroll
        for (int a = 0; a < K; ++a) {
            cmac(cacc[a], shfl_grp<S>(q, a, K), v);
            cmac(wacc[a], shfl_grp<S>(v, a, K), v);
        }
    }
    block_cta_reduce_lanes<S, K, BLOCK>(cacc, wacc, ba.partials, ba.nslots, blockIdx.x);
}

// =================================================================================================
// SpMM, CSR, real K = 4 / 8: the kernel the dispatcher uses for wide real blocks.
// What limits the two kernels above at K >= 4 is not HBM but the L1 data stage (distinct 128-byte lines per
// warp instruction) and, for the lane kernel, the 4K shuffles per row.  Here
//   * the unit's matrix entries (256 rows) are staged in shared memory by fully coalesced loads -- the
//     thread-per-row walk of vals / colidx costs ~4.6 line accesses per row, the staged copy 0.7;
//   * a lane owns FOUR columns of a row (K / 4 lanes per row) and fetches them with one 256-bit load, so a
//     warp's gather of a stencil neighbour touches each line once;
//   * the Gram sums C = Q^H V and W = V^H V: K = 4 keeps the 26 sums of a thread's rows in registers;
//     K = 8 (100 sums: too many registers) runs them on the fp64 tensor cores -- q and v go through a
//     per-warp shared tile and every 4 rows are two mma.m8n8k4.f64, no shuffles.
// =================================================================================================
__device__ __forceinline__ void ld_nc_f64x4(const double* p, double (&o)[4]) {
    asm volatile("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(o[0]), "=d"(o[1]), "=d"(o[2]), "=d"(o[3]) : "l"(p));
}
__device__ __forceinline__ void st_f64x4(double* p, const double (&v)[4]) {
    asm volatile("st.global.v4.f64 [%0], {%1, %2, %3, %4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
}
// D (8 x 8) += A (8 x 4, row) * B (4 x 8, col); lane l holds A[l/4][l%4], B[l%4][l/4], D[l/4][2(l%4) + {0,1}]
__device__ __forceinline__ void dmma_8x8x4(double& d0, double& d1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

__device__ __forceinline__ void mbar_expect(unsigned long long* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_copy_1d(void* dst_smem, const void* src_gmem, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_addr(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_addr(bar))
                 : "memory");
}

// mbar_wait with a watchdog: a byte-count mismatch would otherwise spin forever; after ~2 s the kernel traps and
// the host sees a launch failure instead of a hung GPU.
__device__ __forceinline__ void mbar_wait_bounded(unsigned long long* bar, unsigned parity) {
    const long long t0 = clock64();
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
        if (!done && clock64() - t0 > 4000000000ll) __trap();
    }
}

// Shared-memory stage of one unit: the unit's row pointers, values and column indices, copied by the TMA
// engine (cp.async.bulk, 16-byte aligned source ranges, completion on an mbarrier) while the CTA computes the
// previous unit.  Two stages.
template <int UR> struct TileStage {
    static constexpr int RP = UR + 4;  // row pointers (+ alignment slack)
    static __host__ __device__ constexpr size_t bytes(int cap) { return size_t(cap + 2) * 8 + size_t(cap + 4) * 4 + size_t(RP) * 4; }
};

template <int K, int BLOCK, int UR, bool REGGRAM>
__global__ void __launch_bounds__(BLOCK, REGGRAM ? 2 : 3) spmm_csr_tile_kernel(CsrView<double> A, BlockArgs<double> ba, int cap) {
    static_assert(K == 4 || K == 8, "tile kernel: K = 4 or 8");
    static_assert(!REGGRAM || K == 4, "register Gram sums: K = 4 only");
    pdl_prologue();  // programmatic dependent launch: resident early, reads nothing before the predecessor is complete
    constexpr int LPR = K / 4;          // lanes per row
    constexpr int RPP = BLOCK / LPR;    // rows per CTA pass
    constexpr int PASSES = UR / RPP;
    static_assert(PASSES >= 1 && UR % RPP == 0, "unit = whole passes");
    constexpr int RW = 32 / LPR;        // rows per warp pass
    constexpr int TW = 8;               // K = 8 tile row (one tile for q, one for v)
    constexpr int NT = K / 4;           // tiles per warp
    // K = 4: four planes of 32 x double2 -- (q0,q1) (q2,q3) (v0,v1) (v2,v3) -- 576 bytes apart: a lane's 16-byte
    // stores are consecutive across the warp and the fragment loads of a half warp fall in disjoint banks
    constexpr int PL = 72;
    constexpr int NE = BlockSums<K>::NE, NC = BlockSums<K>::NC;
    extern __shared__ __align__(128) unsigned char stage_smem[];
    __shared__ __align__(8) unsigned long long bar[2];
    __shared__ int meta[2][4];  // e0a (first staged value), e0c (first staged column), staged?, rows
    __shared__ double sT[K * K];
    __shared__ __align__(16) double tile[REGGRAM ? 1 : BLOCK / 32][NT][K == 4 ? 4 * PL : RW * TW];
    __shared__ double sred[BLOCK / 32][NE];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const int L = st->launch_no;
    const double sigma = st->sigma;
    const double* __restrict__ U = ba.u[L & 1];
    double* __restrict__ out = ba.u[(L + 1) & 1];
    load_T<double, K>(st, L, sT);
    const size_t sbytes = TileStage<UR>::bytes(cap);
    auto stage_val = [&](int s) { return reinterpret_cast<double*>(stage_smem + s * sbytes); };
    auto stage_col = [&](int s) { return reinterpret_cast<int*>(stage_smem + s * sbytes + size_t(cap + 2) * 8); };
    auto stage_rp = [&](int s) { return reinterpret_cast<int*>(stage_smem + s * sbytes + size_t(cap + 2) * 8 + size_t(cap + 4) * 4); };
    const int nunits = (A.rows + UR - 1) / UR;
    // one thread: start the copies of `unit` into stage s
    auto issue = [&](int unit, int s) {
        const int r0 = unit * UR, nr = min(UR, A.rows - r0);
        const int e0 = __ldg(A.rowptr + r0), e1 = __ldg(A.rowptr + r0 + nr);
        const int e0a = e0 & ~1, e0c = e0 & ~3;
        const bool fits = e1 - e0 <= cap;
        meta[s][0] = e0a;
        meta[s][1] = e0c;
        meta[s][2] = fits ? 1 : 0;
        meta[s][3] = nr;
        const unsigned brp = unsigned((nr + 1) * 4 + 15) & ~15u;
        const unsigned bv = fits ? unsigned((e1 - e0a) * 8 + 15) & ~15u : 0u;
        const unsigned bc = fits ? unsigned((e1 - e0c) * 4 + 15) & ~15u : 0u;
        mbar_expect(&bar[s], brp + bv + bc);
        tma_copy_1d(stage_rp(s), A.rowptr + r0, brp, &bar[s]);
        if (bv) tma_copy_1d(stage_val(s), A.vals + e0a, bv, &bar[s]);
        if (bc) tma_copy_1d(stage_col(s), A.colidx + e0c, bc, &bar[s]);
    };
    if (threadIdx.x == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], 1);
        const int first = block_unit_at(0, ba.chunk, nunits);
        if (first >= 0) issue(first, 0);
    }
    __syncthreads();  // barriers initialised, sT loaded
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = threadIdx.x % LPR;        // which 4 columns of the row this lane owns
    const int fr = lane & 3, fc = lane >> 2;  // mma fragment coordinates
    // (c0, c1): K = 8: C fragment; K = 4: the stacked tile [Q^T V ; V^T V | ...].  (w0, w1): K = 8: W fragment.
    double c0 = 0.0, c1 = 0.0, w0 = 0.0, w1 = 0.0;
    BlockAcc<double, 4> acc4;  // REGGRAM: the 26 sums of a thread's rows in registers
    acc4.clear();
    double treg[10];  // K = 4: the upper triangle of T in registers
    if constexpr (K == 4) {
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = a; b < 4; ++b) treg[a * 4 - a * (a - 1) / 2 + (b - a)] = sT[a * 4 + b];
    }
    for (int it = 0, unit; (unit = block_unit_at(it, ba.chunk, nunits)) >= 0; ++it) {
        const int s = it & 1;
        if (threadIdx.x == 0) {
            const int next = block_unit_at(it + 1, ba.chunk, nunits);
            if (next >= 0) issue(next, s ^ 1);
        }
        mbar_wait_bounded(&bar[s], unsigned(it >> 1) & 1u);
        const int e0a = meta[s][0], e0c = meta[s][1], nr = meta[s][3];
        const bool staged = meta[s][2] != 0;
        const double* __restrict__ sval = stage_val(s);
        const int* __restrict__ scol = stage_col(s);
        const int* __restrict__ srp = stage_rp(s);
        const int r0 = unit * UR;
#pragma unroll 1
        for (int pass = 0; pass < PASSES; ++pass) {
            const int lr = pass * RPP + threadIdx.x / LPR, row = r0 + lr;
            const bool valid = lr < nr;
            double y[4] = {0.0, 0.0, 0.0, 0.0}, own[4] = {0.0, 0.0, 0.0, 0.0};
            if (valid) {
                const int p0 = srp[lr], p1 = srp[lr + 1];
                ld_nc_f64x4(U + int64_t(row) * K + 4 * sub, own);
                for (int p = p0; p < p1; p += 4) {  // four gathers in flight
                    int col[4];
                    double a[4], u[4][4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const bool ok = p + i < p1;
                        const int pp = ok ? p + i : p;
                        col[i] = staged ? scol[pp - e0c] : __ldg(A.colidx + pp);
                        a[i] = staged ? sval[pp - e0a] : __ldg(A.vals + pp);
                        if (!ok) a[i] = 0.0;
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) ld_nc_f64x4(U + int64_t(col[i]) * K + 4 * sub, u[i]);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) y[j] = fma(a[i], u[i][j], y[j]);
                }
            }
            double v[4], q[4];
            if constexpr (K == 4) {
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    double vv = 0.0, qq = 0.0;
#pragma unroll
                    for (int a = 0; a <= b; ++a) {
                        const double t = treg[a * 4 - a * (a - 1) / 2 + (b - a)];
                        vv = fma(y[a], t, vv);
                        qq = fma(own[a], t, qq);
                    }
                    v[b] = vv * sigma;
                    q[b] = qq;
                }
            } else {
                // the whole row of y and own in both lanes of the row
                double yf[K], of[K];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double yo = __shfl_xor_sync(0xffffffffu, y[j], 1), oo = __shfl_xor_sync(0xffffffffu, own[j], 1);
                    yf[j] = sub == 0 ? y[j] : yo;
                    yf[4 + j] = sub == 0 ? yo : y[j];
                    of[j] = sub == 0 ? own[j] : oo;
                    of[4 + j] = sub == 0 ? oo : own[j];
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    double vv = 0.0, qq = 0.0;
#pragma unroll
                    for (int a = 0; a < K; ++a) {
                        const double t = sT[a * K + 4 * sub + j];  // zero below the diagonal
                        vv = fma(yf[a], t, vv);
                        qq = fma(of[a], t, qq);
                    }
                    v[j] = vv * sigma;
                    q[j] = qq;
                }
            }
            if (valid) st_f64x4(out + int64_t(row) * K + 4 * sub, v);
            // q and v of the warp's rows into its tile(s) (zeros for rows past the end), then the Gram updates
            // on the fp64 tensor cores: every 4 rows are one (K = 4) or two (K = 8) mma.m8n8k4
            if constexpr (REGGRAM) {
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b) acc4.c[a * 4 + b] = fma(q[a], v[b], acc4.c[a * 4 + b]);
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = a; b < 4; ++b) {
                        const int e = a * 4 - a * (a - 1) / 2 + (b - a);
                        acc4.w[e] = fma(v[a], v[b], acc4.w[e]);
                    }
            } else if constexpr (K == 4) {
                double2* tp = reinterpret_cast<double2*>(tile[warp][0]);
                tp[lane] = make_double2(q[0], q[1]);
                tp[PL / 2 + lane] = make_double2(q[2], q[3]);
                tp[2 * (PL / 2) + lane] = make_double2(v[0], v[1]);
                tp[3 * (PL / 2) + lane] = make_double2(v[2], v[3]);
                __syncwarp();
                const double* tw = tile[warp][0];
                const int ca = fc, cb = (fc + 4) & 7;  // A rows: q_0..q_3, v_0..v_3; B columns: v_0..v_3, q_0..q_3
                const double* pa = tw + (ca >> 1) * PL + (ca & 1) + 2 * fr;
                const double* pb = tw + (cb >> 1) * PL + (cb & 1) + 2 * fr;
#pragma unroll
                for (int j = 0; j < 8; ++j) dmma_8x8x4(c0, c1, pa[8 * j], pb[8 * j]);  // rows 4j + fr
                __syncwarp();
            } else {
                double* tq = tile[warp][0] + (lane / LPR) * TW + 4 * sub;
                double* tv = tile[warp][1] + (lane / LPR) * TW + 4 * sub;
                *reinterpret_cast<double2*>(tq) = make_double2(q[0], q[1]);
                *reinterpret_cast<double2*>(tq + 2) = make_double2(q[2], q[3]);
                *reinterpret_cast<double2*>(tv) = make_double2(v[0], v[1]);
                *reinterpret_cast<double2*>(tv + 2) = make_double2(v[2], v[3]);
                __syncwarp();
#pragma unroll
                for (int j = 0; j < RW / 4; ++j) {
                    const double qa = tile[warp][0][(4 * j + fr) * TW + fc], vb = tile[warp][1][(4 * j + fr) * TW + fc];
                    dmma_8x8x4(c0, c1, qa, vb);  // C += Q^T V
                    dmma_8x8x4(w0, w1, vb, vb);  // W += V^T V
                }
                __syncwarp();
            }
        }
        __syncthreads();  // everyone is done with stage s before the copies of unit + 2 * grid land in it
    }
    if constexpr (REGGRAM) {
        block_cta_reduce<double, 4, BLOCK>(acc4, ba.partials, ba.nslots, blockIdx.x);
        return;
    }
    // fragments -> entries.  Lane holds D[fc][2 fr], D[fc][2 fr + 1].
    for (int i = lane; i < NE; i += 32) sred[warp][i] = 0.0;
    __syncwarp();
    if constexpr (K == 8) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int a = fc, b = 2 * fr + h;
            sred[warp][a * K + b] = h ? c1 : c0;
            if (a <= b) sred[warp][NC + a * K - a * (a - 1) / 2 + (b - a)] = h ? w1 : w0;
        }
    } else if (fr < 2) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int b = 2 * fr + h;
            const double d = h ? c1 : c0;
            if (fc < 4) {
                sred[warp][fc * K + b] = d;
            } else if (fc - 4 <= b) {
                const int a = fc - 4;
                sred[warp][NC + a * K - a * (a - 1) / 2 + (b - a)] = d;
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < NE; i += BLOCK) {
        double s = 0.0;
#pragma unroll
        for (int wp = 0; wp < BLOCK / 32; ++wp) s += sred[wp][i];
        ba.partials[int64_t(i) * ba.nslots + blockIdx.x] = s;
    }
}

// W = U^H U of the start block (C is left zero): same reduction, no matrix.
template <class S, int K, int BLOCK>
__global__ void __launch_bounds__(BLOCK) block_gram_kernel(const S* __restrict__ U, int64_t rows, double* partials, int nslots) {
    BlockAcc<S, K> acc;
    acc.clear();
    for (int64_t row = int64_t(blockIdx.x) * BLOCK + threadIdx.x; row < rows; row += int64_t(gridDim.x) * BLOCK) {
        S v[K];
        load_block_row<K>(U, row, v);
#pragma unroll
        for (int a = 0; a < K; ++a)
#pragma unroll
            for (int b = a; b < K; ++b) cmac(acc.w[a * K - a * (a - 1) / 2 + (b - a)], v[a], v[b]);
    }
    block_cta_reduce<S, K, BLOCK>(acc, partials, nslots, blockIdx.x);
}

// =================================================================================================
// Dense A (column-major, lda even): partial products over column segments, then a combine pass.
//   partial kernel: CTA = (row tile, column segment); a thread owns RPT consecutive rows (one 16-byte load per
//   column), the segment's rows of U are staged in shared memory 64 columns at a time and broadcast.
//   combine kernel: thread per row adds the segment partials and runs the common epilogue.
// A is read exactly once; the partials are nseg * n * K values (a few MB next to the 512 MB matrix of C3).
// =================================================================================================
template <class S> struct DenseRpt { static constexpr int value = Sc<S>::is_complex ? 1 : 2; };

template <class S, int K, int BLOCK>
__global__ void __launch_bounds__(BLOCK) gemm_block_partial_kernel(const S* __restrict__ a, int64_t lda, int rows, int cols,
                                                                   int ntiles, int nseg, BlockArgs<S> ba,
                                                                   S* __restrict__ partial) {
    pdl_prologue();
    constexpr int RPT = DenseRpt<S>::value, CHUNK = 64;
    __shared__ S su[CHUNK * K];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const S* __restrict__ U = ba.u[st->launch_no & 1];
    const int tile = blockIdx.x % ntiles, seg = blockIdx.x / ntiles;
    const int c0 = int(int64_t(cols) * seg / nseg), c1 = int(int64_t(cols) * (seg + 1) / nseg);
    const int row0 = (tile * BLOCK + threadIdx.x) * RPT;
    S acc[RPT][K];
#pragma unroll
    for (int r = 0; r < RPT; ++r)
#pragma unroll
        for (int b = 0; b < K; ++b) acc[r][b] = Sc<S>::zero();
    for (int cb = c0; cb < c1; cb += CHUNK) {
        const int nc = min(CHUNK, c1 - cb);
        __syncthreads();
        for (int i = threadIdx.x; i < nc * K; i += BLOCK) su[i] = __ldg(U + int64_t(cb) * K + i);
        __syncthreads();
        if (row0 < rows) {
#pragma unroll 4
            for (int j = 0; j < nc; ++j) {
                const double2 t = ld_stream_f64x2(reinterpret_cast<const double*>(a + (int64_t(cb + j) * lda + row0)));
                if constexpr (RPT == 2) {
#pragma unroll
                    for (int b = 0; b < K; ++b) {
                        const S ub = su[j * K + b];
                        Sc<S>::mac(acc[0][b], *reinterpret_cast<const S*>(&t.x), ub);
                        Sc<S>::mac(acc[1 % RPT][b], *reinterpret_cast<const S*>(&t.y), ub);
                    }
                } else {
#pragma unroll
                    for (int b = 0; b < K; ++b) Sc<S>::mac(acc[0][b], *reinterpret_cast<const S*>(&t), su[j * K + b]);
                }
            }
        }
    }
#pragma unroll
    for (int r = 0; r < RPT; ++r)
        if (row0 + r < rows) store_block_row<K>(partial + int64_t(seg) * rows * K, row0 + r, acc[r]);
}

template <class S, int K, int BLOCK>
__global__ void __launch_bounds__(BLOCK) block_combine_kernel(const S* __restrict__ partial, int rows, int nseg, BlockArgs<S> ba) {
    pdl_prologue();
    __shared__ S sT[K * K];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const int L = st->launch_no;
    const double sigma = st->sigma;
    const S* __restrict__ U = ba.u[L & 1];
    S* __restrict__ out = ba.u[(L + 1) & 1];
    load_T<S, K>(st, L, sT);
    __syncthreads();
    BlockAcc<S, K> acc;
    acc.clear();
    for (int64_t row = int64_t(blockIdx.x) * BLOCK + threadIdx.x; row < rows; row += int64_t(gridDim.x) * BLOCK) {
        S y[K];
#pragma unroll
        for (int b = 0; b < K; ++b) y[b] = Sc<S>::zero();
        for (int s = 0; s < nseg; ++s) {
            S t[K];
            load_block_row<K>(partial + int64_t(s) * rows * K, row, t);
#pragma unroll
            for (int b = 0; b < K; ++b) Sc<S>::add(y[b], t[b]);
        }
        S own[K], v[K];
        load_block_row<K>(U, row, own);
        block_row_epilogue<S, K>(y, own, sT, sigma, v, acc);
        store_block_row<K>(out, row, v);
    }
    block_cta_reduce<S, K, BLOCK>(acc, ba.partials, ba.nslots, blockIdx.x);
}

// =================================================================================================
// Control: one CTA.  Warps fold the per-CTA partials (lanes stride over the slots, fixed order), thread 0
// runs the k x k step of blockmath.hpp.
// =================================================================================================
template <int BLOCK>
__global__ void __launch_bounds__(BLOCK) block_control_kernel(BlockState* st, const double* __restrict__ partials, int nslots,
                                                              int k, int is_complex, int begin, const BlockPeer* peer,
                                                              unsigned long long epoch) {
    pdl_prologue();
    __shared__ double tot[2 * (kBlockElems + kBlockMax * (kBlockMax + 1) / 2)];
    __shared__ Cx sC[kBlockElems], sW[kBlockElems], sH[kBlockElems], swork[kBlockElems];
    __shared__ Cx stheta[kBlockMax];
    __shared__ int flags[2];  // eigenvalues ok, Cholesky ok
    if (!begin && st->stop != BSTOP_NONE) return;
    const int nc = k * k, ne = nc + k * (k + 1) / 2, stride = (is_complex ? 2 : 1) * ne;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = warp; i < stride; i += BLOCK / 32) {
        double s = 0.0;
        for (int slot = lane; slot < nslots; slot += 32) s += partials[int64_t(i) * nslots + slot];
        s = warp_sum(s);
        if (lane == 0) tot[i] = s;
    }
    __syncthreads();
    if (peer) {
        // all-reduce of this rank's sums over the ranks + barrier (see BlockPeer): thread q talks to rank q
        const BlockPeer& p = *peer;
        __shared__ int s_ok;
        if (threadIdx.x == 0) s_ok = 1;
        __syncthreads();
        const int par = int(epoch & 1ull);
        const int q = threadIdx.x;
        if (q < p.nranks) {
            double* box = (q == p.rank ? p.mail_in : p.peer_mail[q]) + (size_t(par) * kMaxPeers + p.rank) * kBlockMailStride;
            for (int i = 0; i < stride; ++i) box[i] = tot[i];
            if (q != p.rank) {
                st_release_sys(p.peer_flag[q], epoch);
                const long long t0 = clock64();
                while (ld_acquire_sys(p.flag_in + q) < epoch)
                    if (clock64() - t0 > p.spin_limit) {
                        s_ok = 0;
                        break;
                    }
            }
        }
        __syncthreads();
        if (!s_ok) {
            if (threadIdx.x == 0) st->stop = BSTOP_ERROR;
            return;
        }
        for (int i = threadIdx.x; i < stride; i += BLOCK) {
            double s = 0.0;
            for (int r = 0; r < p.nranks; ++r)  // rank order: the same sum, bit for bit, on every rank
                s += ld_volatile_f64(p.mail_in + (size_t(par) * kMaxPeers + r) * kBlockMailStride + i);
            tot[i] = s;
        }
        __syncthreads();
    }
    for (int e = threadIdx.x; e < nc; e += BLOCK) {
        const int a = e / k, b = e % k;
        sC[a * kBlockLd + b] = cx(tot[e], is_complex ? tot[ne + e] : 0.0);
        if (b >= a) {
            const int ew = nc + a * k - a * (a - 1) / 2 + (b - a);
            sW[a * kBlockLd + b] = cx(tot[ew], is_complex ? tot[ne + ew] : 0.0);
        } else {
            sW[a * kBlockLd + b] = cx(0.0);
        }
    }
    __syncthreads();
    if (begin) {
        if (threadIdx.x == 0) block_control_begin(st, sW);
        return;
    }
    // the three pieces of the step (blockmath.hpp): warp 0 the Ritz values, warp 1 the Cholesky factor, then thread 0
    if (warp == 0) {
        block_step_H(st, sC, sH, lane, 32);
        __syncwarp();
        Cx theta[kBlockMax];
        const bool ok = eig_values_coop(sH, k, theta, swork, lane, 32);
        if (lane == 0) {
            flags[0] = ok ? 1 : 0;
            for (int i = 0; i < k; ++i) stheta[i] = theta[i];
        }
    } else if (threadIdx.x == 32) {
        flags[1] = block_step_T(st, sW) ? 1 : 0;
    }
    __syncthreads();
    if (threadIdx.x == 0) block_step_decide(st, sH, stheta, flags[0] != 0, flags[1] != 0);
}

// ---- layout changes at the API boundary (Eigen holds n x k column-major) ---------------------------------
template <class S, int K> __global__ void block_pack_kernel(const S* __restrict__ xcol, int64_t n, S* __restrict__ u) {
    const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= n) return;
    S v[K];
#pragma unroll
    for (int b = 0; b < K; ++b) v[b] = xcol[int64_t(b) * n + row];
    store_block_row<K>(u, row, v);
}

template <class S, int K> struct BlockMat {
    S m[K * K];
};

// X (column-major) = U M
template <class S, int K>
__global__ void block_finish_kernel(const S* __restrict__ u, BlockMat<S, K> M, int64_t n, S* __restrict__ xcol) {
    const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    if (row >= n) return;
    S own[K];
    load_block_row<K>(u, row, own);
#pragma unroll
    for (int b = 0; b < K; ++b) {
        S s = Sc<S>::zero();
#pragma unroll
        for (int a = 0; a < K; ++a) Sc<S>::mac(s, own[a], M.m[a * K + b]);
        xcol[int64_t(b) * n + row] = s;
    }
}

}  // namespace eigsol_b200
