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
};

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
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
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
                                                              int k, int is_complex, int begin) {
    __shared__ double tot[2 * (kBlockElems + kBlockMax * (kBlockMax + 1) / 2)];
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
    if (threadIdx.x != 0) return;
    Cx C[kBlockElems], W[kBlockElems];
    for (int a = 0; a < k; ++a)
        for (int b = 0; b < k; ++b) {
            const int e = a * k + b;
            C[a * kBlockLd + b] = cx(tot[e], is_complex ? tot[ne + e] : 0.0);
            if (b >= a) {
                const int ew = nc + a * k - a * (a - 1) / 2 + (b - a);
                W[a * kBlockLd + b] = cx(tot[ew], is_complex ? tot[ne + ew] : 0.0);
            } else {
                W[a * kBlockLd + b] = cx(0.0);
            }
        }
    if (begin) block_control_begin(st, W);
    else block_control_step(st, C, W);
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
