// block_host.cuh -- host side of the block power iteration (included by engine.cu after MatrixImpl).
// One GPU; uses the matrix copy already resident behind the handle (CSR arrays or the column-major dense
// matrix) and its own two n x K blocks.  Per trip: one SpMM launch (dense: partial + combine) and the
// single-CTA control launch; the host reads the 2 KB state every `poll` trips.
#pragma once
#include "block.cuh"

namespace eigsol_b200 {

struct BlockResult {
    int iterations = 0, converged = 0, breakdown = 0, launches = 0;
    float ms = 0.f;  // device time of the iteration loop (without upload, Gram pass and download)
};

template <class S, int K>
static void block_run_t(MatrixImpl& m, int nev, const void* x0, int max_it, double tol, void* x_out, double* lambda_out,
                        BlockResult& res) {
    constexpr int B = 256;
    const int64_t n = m.n;
    const size_t s = sizeof(S);
    cudaStream_t st = m.stream;
    DevBuf u[2], d_xcol, d_state, d_partials, d_dpart;
    u[0].alloc(size_t(n) * K * s);
    u[1].alloc(size_t(n) * K * s);
    d_xcol.alloc(size_t(n) * K * s);
    d_state.alloc(sizeof(BlockState));
    PinnedBuf h_state;
    h_state.alloc(sizeof(BlockState));

    // launch geometry
    int grid, ntiles = 1, nseg = 1, grid_partial = 0;
    void (*spmm)(CsrView<S>, BlockArgs<S>) = nullptr;
    void (*spmm_tile)(CsrView<double>, BlockArgs<double>, int) = nullptr;  // real K >= 4: TMA-staged tile kernel
    int tile_cap = 0;
    size_t tile_smem = 0;
    if (m.dense) {
        grid = int(std::min<int64_t>((n + B - 1) / B, int64_t(m.sm_count) * occupancy(block_combine_kernel<S, K, B>, B)));
        const int rt = B * DenseRpt<S>::value;
        ntiles = int((n + rt - 1) / rt);
        const int waves = env_int("EIGSOL_B200_BLOCK_GEMM_WAVES", 4);
        nseg = int(std::max<int64_t>(1, std::min<int64_t>(n / 64, (int64_t(m.sm_count) * waves + ntiles - 1) / ntiles)));
        grid_partial = ntiles * nseg;
        d_dpart.alloc(size_t(nseg) * size_t(n) * K * s);
    } else {
        // real K >= 4: staged tile kernel; complex K >= 4: K lanes per row; K <= 2: one thread per row (block.cuh).
        // EIGSOL_B200_BLOCK_KERNEL=row|lanes|tile overrides.
        enum { V_ROW, V_LANES, V_TILE };
        constexpr bool can_tile = !Sc<S>::is_complex && K >= 4, can_row = K <= 4;
        int variant = can_tile ? V_TILE : (K >= 4 ? V_LANES : V_ROW);
        if (const char* kv = std::getenv("EIGSOL_B200_BLOCK_KERNEL")) {
            if (!std::strcmp(kv, "row")) variant = V_ROW;
            else if (!std::strcmp(kv, "lanes")) variant = V_LANES;
            else if (!std::strcmp(kv, "tile")) variant = V_TILE;
        }
        if (variant == V_TILE && !can_tile) variant = K >= 4 ? V_LANES : V_ROW;
        if (variant == V_ROW && !can_row) variant = V_LANES;
        const int unroll = env_int("EIGSOL_B200_BLOCK_UNROLL", 4);
        int rows_per_unit = B;
        int occ = 1;
        if (variant == V_TILE) {
            if constexpr (can_tile) {
                constexpr int UR = K == 4 ? 256 : 128;  // rows staged together (one CTA pass)
                if constexpr (K == 4) {
                    spmm_tile = env_int("EIGSOL_B200_BLOCK_REGGRAM", 0) ? spmm_csr_tile_kernel<K, B, UR, true>
                                                                      : spmm_csr_tile_kernel<K, B, UR, false>;
                } else {
                    spmm_tile = spmm_csr_tile_kernel<K, B, UR, false>;
                }
                rows_per_unit = UR;
                // staged entries per unit: the mean unit plus 12.5 %; a unit with more reads the matrix from global
                const double mean_row = n > 0 ? double(m.local_nnz) / double(n) : 0.0;
                int64_t cap = int64_t(double(UR) * mean_row * 1.125) + 32;
                cap = std::min<int64_t>(std::max<int64_t>(cap, 256), 6144);
                tile_cap = env_int("EIGSOL_B200_BLOCK_CAP", int((cap + 3) / 4 * 4)) / 4 * 4;
                tile_smem = 2 * TileStage<UR>::bytes(tile_cap);
                EIGSOL_CUDA(cudaFuncSetAttribute(spmm_tile, cudaFuncAttributeMaxDynamicSharedMemorySize, int(tile_smem)));
                occ = occupancy(spmm_tile, B, tile_smem);
            }
        } else if (variant == V_LANES) {
            spmm = unroll >= 8 ? spmm_csr_lanes_kernel<S, K, B, 8>
                               : (unroll <= 2 ? spmm_csr_lanes_kernel<S, K, B, 2> : spmm_csr_lanes_kernel<S, K, B, 4>);
            rows_per_unit = B / K;
            occ = occupancy(spmm, B);
        } else {
            if constexpr (can_row) spmm = spmm_csr_block_kernel<S, K, B>;
            occ = occupancy(spmm, B);
        }
        grid = int(std::min<int64_t>((n + rows_per_unit - 1) / rows_per_unit, int64_t(m.sm_count) * occ));
    }
    grid = std::max(grid, 1);
    constexpr int stride = BlockStride<S, K>::value;
    d_partials.alloc(sizeof(double) * size_t(stride) * size_t(grid));

    BlockArgs<S> ba;
    ba.u[0] = u[0].as<S>();
    ba.u[1] = u[1].as<S>();
    ba.st = d_state.as<BlockState>();
    ba.partials = d_partials.as<double>();
    ba.nslots = grid;
    ba.chunk = std::max(1, env_int("EIGSOL_B200_BLOCK_CHUNK", 1));

    // start block: column-major n x K from the caller -> row-major U_0, W_0 = U_0^H U_0, T_0
    EIGSOL_CUDA(cudaMemcpyAsync(d_xcol.p, x0, size_t(n) * K * s, cudaMemcpyHostToDevice, st));
    const int gn = int((n + B - 1) / B);
    block_pack_kernel<S, K><<<gn, B, 0, st>>>(d_xcol.as<S>(), n, ba.u[0]);
    BlockState* hs = h_state.as<BlockState>();
    std::memset(hs, 0, sizeof(BlockState));
    hs->k = K;
    hs->nev = nev;
    hs->max_trips = max_it;
    hs->tol = tol;
    hs->sigma = m.sigma0;
    EIGSOL_CUDA(cudaMemcpyAsync(d_state.p, hs, sizeof(BlockState), cudaMemcpyHostToDevice, st));
    block_gram_kernel<S, K, B><<<grid, B, 0, st>>>(ba.u[0], n, ba.partials, grid);
    block_control_kernel<1024><<<1, 1024, 0, st>>>(ba.st, ba.partials, grid, K, Sc<S>::is_complex ? 1 : 0, 1);
    EIGSOL_CUDA(cudaGetLastError());
    res.launches += 3;

    auto one_trip = [&] {
        if (m.dense) {
            launch_pdl(m.pdl, st, gemm_block_partial_kernel<S, K, B>, dim3(grid_partial), dim3(B), 0, m.d_dense.as<S>(), m.lda,
                       int(n), int(n), ntiles, nseg, ba, d_dpart.as<S>());
            launch_pdl(m.pdl, st, block_combine_kernel<S, K, B>, dim3(grid), dim3(B), 0, d_dpart.as<S>(), int(n), nseg, ba);
            res.launches += 3;
        } else {
            CsrView<S> A{int32_t(n), m.d_rowptr.as<int32_t>(), m.d_colidx.as<int32_t>(), m.d_vals.as<S>()};
            if constexpr (!Sc<S>::is_complex) {
                if (spmm_tile) launch_pdl(m.pdl, st, spmm_tile, dim3(grid), dim3(B), tile_smem, A, ba, tile_cap);
                else launch_pdl(m.pdl, st, spmm, dim3(grid), dim3(B), 0, A, ba);
            } else {
                launch_pdl(m.pdl, st, spmm, dim3(grid), dim3(B), 0, A, ba);
            }
            res.launches += 2;
        }
        launch_pdl(m.pdl, st, block_control_kernel<1024>, dim3(1), dim3(1024), 0, ba.st, ba.partials, grid, K,
                   Sc<S>::is_complex ? 1 : 0, 0);
    };

    const int poll = std::max(1, env_int("EIGSOL_B200_BLOCK_POLL", 16));
    EIGSOL_CUDA(cudaEventRecord(m.ev0, st));
    int issued = 0;
    for (;;) {
        EIGSOL_CUDA(cudaMemcpyAsync(hs, d_state.p, sizeof(BlockState), cudaMemcpyDeviceToHost, st));
        EIGSOL_CUDA(cudaStreamSynchronize(st));
        if (hs->stop != BSTOP_NONE) break;
        // launch 0 is the set-up product, so max_it trips need max_it + 1 launches; never issue more than that
        const int left = max_it + 1 - issued;
        if (left <= 0) throw DeviceError("powerMethod: device error: block iteration did not stop at maxIterations");
        const int batch = std::min(poll, left);
        for (int i = 0; i < batch; ++i) one_trip();
        EIGSOL_CUDA(cudaGetLastError());
        issued += batch;
    }
    EIGSOL_CUDA(cudaEventRecord(m.ev1, st));
    EIGSOL_CUDA(cudaEventSynchronize(m.ev1));
    EIGSOL_CUDA(cudaEventElapsedTime(&res.ms, m.ev0, m.ev1));
    m.launches += res.launches;

    res.iterations = hs->iterations;
    res.converged = hs->converged;
    res.breakdown = hs->stop == BSTOP_BREAKDOWN ? 1 : 0;
    for (int i = 0; i < K; ++i) {
        lambda_out[2 * i] = hs->theta[i].re;
        lambda_out[2 * i + 1] = hs->theta[i].im;
    }
    if (!x_out) return;
    // X = U_f T_f S: the Ritz vectors of the last folded launch (power_method.hpp:98 returns x of the last trip)
    Cx M[kBlockElems];
    block_result_vectors(hs);  // S: the loop itself only needs the Ritz values
    block_result_transform(hs, M);
    BlockMat<S, K> bm;
    if constexpr (Sc<S>::is_complex) {
        for (int a = 0; a < K; ++a)
            for (int b = 0; b < K; ++b) bm.m[a * K + b] = make_double2(M[a * kBlockLd + b].re, M[a * kBlockLd + b].im);
    } else {
        // real matrix: a conjugate pair (theta_j, theta_{j+1}) returns Re and Im of its vector in columns j, j+1
        // (the LAPACK dgeev convention); real Ritz values return their (real) vector
        for (int b = 0; b < K; ++b) {
            const Cx t = hs->theta[b];
            const bool cpx = std::fabs(t.im) > 1e-12 * std::hypot(t.re, t.im);
            const bool second = cpx && t.im < 0.0 && b > 0;
            for (int a = 0; a < K; ++a) bm.m[a * K + b] = second ? M[a * kBlockLd + b - 1].im : M[a * kBlockLd + b].re;
        }
    }
    const int f = std::max(hs->final_launch, 0);
    block_finish_kernel<S, K><<<gn, B, 0, st>>>(ba.u[f & 1], bm, n, d_xcol.as<S>());
    EIGSOL_CUDA(cudaGetLastError());
    ++m.launches;
    EIGSOL_CUDA(cudaMemcpyAsync(x_out, d_xcol.p, size_t(n) * K * s, cudaMemcpyDeviceToHost, st));
    EIGSOL_CUDA(cudaStreamSynchronize(st));
}

template <class S>
static void block_run_k(MatrixImpl& m, int k, int nev, const void* x0, int max_it, double tol, void* x_out,
                        double* lambda_out, BlockResult& res) {
    switch (k) {
        case 1: return block_run_t<S, 1>(m, nev, x0, max_it, tol, x_out, lambda_out, res);
        case 2: return block_run_t<S, 2>(m, nev, x0, max_it, tol, x_out, lambda_out, res);
        case 4: return block_run_t<S, 4>(m, nev, x0, max_it, tol, x_out, lambda_out, res);
        case 8: return block_run_t<S, 8>(m, nev, x0, max_it, tol, x_out, lambda_out, res);
    }
    throw PlanError("block width must be 1, 2, 4 or 8");
}

static void block_power_run(MatrixImpl& m, int k, int nev, const void* x0, int max_it, double tol, void* x_out,
                            double* lambda_out, BlockResult& res) {
    if (m.nranks != 1) throw PlanError("the block power iteration runs on one GPU in this version");
    if (nev < 1 || nev > k) throw PlanError("nev must satisfy 1 <= nev <= block width");
    if (!x0 || !lambda_out) throw PlanError("null pointer");
    EIGSOL_CUDA(cudaSetDevice(m.device));
    if (m.dtype) block_run_k<double2>(m, k, nev, x0, max_it, tol, x_out, lambda_out, res);
    else block_run_k<double>(m, k, nev, x0, max_it, tol, x_out, lambda_out, res);
}

}  // namespace eigsol_b200
