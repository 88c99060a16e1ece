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
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
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
//   * the Gram sums C = Q^H V and W = V^H V run on the fp64 tensor cores: q and v go through a per-warp
//     shared tile and every 4 rows are one (K = 4: C and W stacked in one 8 x 8 tile) or two (K = 8)
//     mma.m8n8k4.f64 -- no shuffles, and ~60 accumulator registers fewer than the thread-per-row kernel.
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

template <int K, int BLOCK>
__global__ void __launch_bounds__(BLOCK) spmm_csr_tile_kernel(CsrView<double> A, BlockArgs<double> ba) {
    static_assert(K == 4 || K == 8, "tile kernel: K = 4 or 8");
    constexpr int LPR = K / 4;          // lanes per row
    constexpr int RPP = BLOCK / LPR;    // rows per CTA pass
    constexpr int UR = 256;             // rows per unit (staged together)
    constexpr int PASSES = UR / RPP;
    constexpr int CAP = 2048;           // staged entries per unit; a unit with more reads the matrix from global
    constexpr int RW = 32 / LPR;        // rows per warp pass
    constexpr int TW = 2 * K;           // tile row: q[0..K) then v[0..K)
    constexpr int NE = BlockSums<K>::NE, NC = BlockSums<K>::NC;
    __shared__ double sT[K * K];
    __shared__ double sval[CAP];
    __shared__ int scol[CAP];
    __shared__ int srp[UR + 1];
    __shared__ __align__(16) double tile[BLOCK / 32][RW * TW];
    __shared__ double sred[BLOCK / 32][NE];
    const BlockState* st = ba.st;
    if (st->stop != BSTOP_NONE) return;
    const int L = st->launch_no;
    const double sigma = st->sigma;
    const double* __restrict__ U = ba.u[L & 1];
    double* __restrict__ out = ba.u[(L + 1) & 1];
    load_T<double, K>(st, L, sT);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = threadIdx.x % LPR;        // which 4 columns of the row this lane owns
    const int fr = lane & 3, fc = lane >> 2;  // mma fragment coordinates
    double* wt = tile[warp];
    // accumulators: K = 8: (c0, c1) = C fragment, (w0, w1) = W fragment; K = 4: (c0, c1) = the stacked 8 x 8 tile
    double c0 = 0.0, c1 = 0.0, w0 = 0.0, w1 = 0.0;
    const int nunits = (A.rows + UR - 1) / UR;
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
        const int r0 = unit * UR, nr = min(UR, A.rows - r0);
        __syncthreads();  // the previous unit's readers are done with the staged arrays
        for (int i = threadIdx.x; i <= nr; i += BLOCK) srp[i] = __ldg(A.rowptr + r0 + i);
        __syncthreads();
        const int e0 = srp[0], ne = srp[nr] - e0;
        const bool staged = ne <= CAP;
        if (staged)
            for (int i = threadIdx.x; i < ne; i += BLOCK) {
                sval[i] = ld_stream_f64(A.vals + e0 + i);
                scol[i] = ld_stream_s32(A.colidx + e0 + i);
            }
        __syncthreads();
#pragma unroll 1
        for (int pass = 0; pass < PASSES; ++pass) {
            const int lr = pass * RPP + threadIdx.x / LPR, row = r0 + lr;
            const bool valid = lr < nr;
            double y[4] = {0.0, 0.0, 0.0, 0.0}, own[4] = {0.0, 0.0, 0.0, 0.0};
            if (valid) {
                const int p0 = srp[lr] - e0, p1 = srp[lr + 1] - e0;
                int p = p0;
                for (; p + 1 < p1; p += 2) {  // two gathers in flight
                    const int ca = staged ? scol[p] : __ldg(A.colidx + e0 + p);
                    const int cb = staged ? scol[p + 1] : __ldg(A.colidx + e0 + p + 1);
                    const double va = staged ? sval[p] : __ldg(A.vals + e0 + p);
                    const double vb = staged ? sval[p + 1] : __ldg(A.vals + e0 + p + 1);
                    double ua[4], ub[4];
                    ld_nc_f64x4(U + int64_t(ca) * K + 4 * sub, ua);
                    ld_nc_f64x4(U + int64_t(cb) * K + 4 * sub, ub);
#pragma unroll
                    for (int j = 0; j < 4; ++j) y[j] = fma(va, ua[j], y[j]);
#pragma unroll
                    for (int j = 0; j < 4; ++j) y[j] = fma(vb, ub[j], y[j]);
                }
                if (p < p1) {
                    const int ca = staged ? scol[p] : __ldg(A.colidx + e0 + p);
                    const double va = staged ? sval[p] : __ldg(A.vals + e0 + p);
                    double ua[4];
                    ld_nc_f64x4(U + int64_t(ca) * K + 4 * sub, ua);
#pragma unroll
                    for (int j = 0; j < 4; ++j) y[j] = fma(va, ua[j], y[j]);
                }
                ld_nc_f64x4(U + int64_t(row) * K + 4 * sub, own);
            }
            // the whole row of y and own in every lane of the row's lane group
            double yf[K], of[K];
            if constexpr (LPR == 1) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    yf[j] = y[j];
                    of[j] = own[j];
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double yo = __shfl_xor_sync(0xffffffffu, y[j], 1), oo = __shfl_xor_sync(0xffffffffu, own[j], 1);
                    yf[j] = sub == 0 ? y[j] : yo;
                    yf[4 + j] = sub == 0 ? yo : y[j];
                    of[j] = sub == 0 ? own[j] : oo;
                    of[4 + j] = sub == 0 ? oo : own[j];
                }
            }
            double v[4], q[4];
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
            if (valid) st_f64x4(out + int64_t(row) * K + 4 * sub, v);
            // q and v of the warp's RW rows into its tile (zeros for rows past the end), then the Gram updates
            double* tr = wt + (lane / LPR) * TW;
            *reinterpret_cast<double4*>(tr + 4 * sub) = make_double4(q[0], q[1], q[2], q[3]);
            *reinterpret_cast<double4*>(tr + K + 4 * sub) = make_double4(v[0], v[1], v[2], v[3]);
            __syncwarp();
#pragma unroll
            for (int j = 0; j < RW / 4; ++j) {
                const double* r4 = wt + (4 * j + fr) * TW;
                if constexpr (K == 8) {
                    const double qa = r4[fc], vb = r4[K + fc];
                    dmma_8x8x4(c0, c1, qa, vb);  // C += Q^T V
                    dmma_8x8x4(w0, w1, vb, vb);  // W += V^T V
                } else {
                    const double a8 = r4[fc];            // rows of A: q_0..q_3, v_0..v_3
                    const double b8 = r4[(fc + 4) & 7];  // columns of B: v_0..v_3, q_0..q_3
                    dmma_8x8x4(c0, c1, a8, b8);          // top-left 4 x 4 = Q^T V, bottom-left = V^T V
                }
            }
            __syncwarp();
        }
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
    } else {
        if (fr < 2) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int b = 2 * fr + h;
                const double d = h ? c1 : c0;
                if (fc < 4) {
                    sred[warp][fc * K + b] = d;
                } else {
                    const int a = fc - 4;
                    if (a <= b) sred[warp][NC + a * K - a * (a - 1) / 2 + (b - a)] = d;
                }
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
