// block_host.cuh -- host side of the block power iteration (included by engine.cu after MatrixImpl).
// One GPU or a row-sharded group (sparse); uses the matrix copy already resident behind the handle (CSR arrays or the column-major dense
// matrix) and its own two n x K blocks.  Per trip: one SpMM launch (dense: partial + combine) and the
// single-CTA control launch; the host reads the 2 KB state every `poll` trips.
#pragma once
#include "block.cuh"

namespace eigsol_b200 {

struct BlockResult {
    int iterations = 0, converged = 0, breakdown = 0, launches = 0;
    float ms = 0.f;  // device time of the iteration loop (without upload, Gram pass and download)
};

// What a rank shows its peers for one block solve (single-process group: raw pointers; process-per-GPU: CUDA IPC).
struct BlockPeerBlob {
    cudaIpcMemHandle_t u[2], mail, flag;
    void *raw_u[2], *raw_mail, *raw_flag;
    int64_t halo_base;
    int64_t peer_off[kMaxPeers + 1];
};

struct BlockPeerHost {
    DevBuf d_mail, d_flag, d_peer;
    std::vector<void*> opened;
    ~BlockPeerHost() {
        for (void* q : opened) cudaIpcCloseMemHandle(q);
    }
};

// Fills and uploads the BlockPeer of this rank for the two block buffers u0 / u1 (ext x K row-major each).
static const BlockPeer* block_setup_peer(MatrixImpl& m, void* u0, void* u1, size_t row_bytes, BlockPeerHost& host) {
    if (m.nranks == 1) return nullptr;
    if (m.xmode != MODE_PEER) throw PlanError("the multi-GPU block power iteration needs the peer-memory exchange (EIGSOL_B200_EXCHANGE=peer)");
    const bool local = m.comm->local != nullptr;
    const bool shared_pool = local && m.comm->pool_shared;
    if (shared_pool) {
        host.d_mail.alloc(sizeof(double) * 2 * kMaxPeers * kBlockMailStride);
        host.d_flag.alloc(sizeof(unsigned long long) * kMaxPeers);
    } else {
        host.d_mail.alloc_plain(sizeof(double) * 2 * kMaxPeers * kBlockMailStride);
        host.d_flag.alloc_plain(sizeof(unsigned long long) * kMaxPeers);
    }
    BlockPeerBlob mine;
    std::memset(&mine, 0, sizeof mine);
    if (!local) {
        int ok = 1;
        ok &= cudaIpcGetMemHandle(&mine.u[0], u0) == cudaSuccess;
        ok &= cudaIpcGetMemHandle(&mine.u[1], u1) == cudaSuccess;
        ok &= cudaIpcGetMemHandle(&mine.mail, host.d_mail.p) == cudaSuccess;
        ok &= cudaIpcGetMemHandle(&mine.flag, host.d_flag.p) == cudaSuccess;
        cudaGetLastError();
        if (!comm_min_int(m.comm, ok)) throw DeviceError("powerMethod: device error: cannot export the block buffers to the peer GPUs (CUDA IPC)");
    }
    mine.raw_u[0] = u0;
    mine.raw_u[1] = u1;
    mine.raw_mail = host.d_mail.p;
    mine.raw_flag = host.d_flag.p;
    mine.halo_base = m.nloc_pad;
    for (int q = 0; q <= m.nranks && q <= kMaxPeers; ++q) mine.peer_off[q] = m.halo.peer_off[size_t(q)];
    const std::vector<char> all = comm_allgather_bytes(m.comm, &mine, sizeof mine);
    BlockPeer bp;
    std::memset(&bp, 0, sizeof bp);
    bp.nranks = m.nranks;
    bp.rank = m.rank;
    bp.mail_in = host.d_mail.as<double>();
    bp.flag_in = host.d_flag.as<unsigned long long>();
    bp.spin_limit = m.peer_h.spin_limit;
    bp.send_idx = m.d_send_idx.as<int32_t>();
    int mapped = 1;
    for (int q = 0; q < m.nranks && mapped; ++q) {
        if (q == m.rank) continue;
        BlockPeerBlob pb;
        std::memcpy(&pb, all.data() + sizeof(BlockPeerBlob) * size_t(q), sizeof pb);
        void *ru[2] = {pb.raw_u[0], pb.raw_u[1]}, *rm = pb.raw_mail, *rf = pb.raw_flag;
        if (!local) {
            auto open = [&](void** dst, const cudaIpcMemHandle_t& h) {
                if (cudaIpcOpenMemHandle(dst, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
                    cudaGetLastError();
                    mapped = 0;
                    return;
                }
                host.opened.push_back(*dst);
            };
            open(&ru[0], pb.u[0]);
            if (mapped) open(&ru[1], pb.u[1]);
            if (mapped) open(&rm, pb.mail);
            if (mapped) open(&rf, pb.flag);
            if (!mapped) break;
        }
        bp.peer_mail[q] = static_cast<double*>(rm);
        bp.peer_flag[q] = static_cast<unsigned long long*>(rf) + m.rank;
        if (m.send_count[size_t(q)] == 0) continue;
        const size_t dst_off = size_t(pb.halo_base + pb.peer_off[m.rank]) * row_bytes;
        if (m.send_contig[size_t(q)]) {
            const int k = bp.nsend++;
            bp.send_start[k] = int(m.send_start[size_t(q)]);
            bp.send_count[k] = int(m.send_count[size_t(q)]);
            for (int b = 0; b < 2; ++b) bp.send_dst[k][b] = static_cast<char*>(ru[b]) + dst_off;
        } else {
            const int k = bp.nlist++;
            bp.list_off[k] = int(m.send_off[size_t(q)]);
            bp.list_count[k] = int(m.send_count[size_t(q)]);
            for (int b = 0; b < 2; ++b) bp.list_dst[k][b] = static_cast<char*>(ru[b]) + dst_off;
        }
    }
    if (!comm_min_int(m.comm, mapped)) throw DeviceError("powerMethod: device error: cannot map the peers' block buffers");
    host.d_peer.alloc(sizeof(BlockPeer));
    EIGSOL_CUDA(cudaMemcpy(host.d_peer.p, &bp, sizeof bp, cudaMemcpyHostToDevice));
    return host.d_peer.as<BlockPeer>();
}

template <class S, int K>
static void block_run_t(MatrixImpl& m, int nev, const void* x0, int max_it, double tol, void* x_out, double* lambda_out,
                        BlockResult& res) {
    constexpr int B = 256;
    const int64_t nglob = m.n;
    // a rank of a group holds [own rows | pad | halo rows]: n = own rows (what the kernels loop over), ext = rows of
    // the block buffers (what the gathers may address)
    const bool multi = m.nranks > 1;
    const int64_t n = multi ? m.nloc : m.n;
    const int64_t ext = multi ? m.ext_len : m.n;
    const size_t s = sizeof(S);
    cudaStream_t st = m.stream;
    DevBuf u[2], d_xcol, d_state, d_partials, d_dpart;
    const bool plain = multi && !(m.comm->local && m.comm->pool_shared);  // exported with CUDA IPC / mapped by peers
    for (auto& b : u) {
        if (plain) b.alloc_plain(size_t(ext) * K * s);
        else b.alloc(size_t(ext) * K * s);
    }
    d_xcol.alloc(size_t(ext) * K * s);
    d_state.alloc(sizeof(BlockState));
    BlockPeerHost peer_host;
    const BlockPeer* peer = block_setup_peer(m, u[0].p, u[1].p, size_t(K) * s, peer_host);
    unsigned long long epoch = 0;
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

    // start block: column-major nglob x K from the caller -> row-major U_0 (own rows + halo rows), W_0 = U_0^H U_0, T_0
    if (!multi) {
        EIGSOL_CUDA(cudaMemcpyAsync(d_xcol.p, x0, size_t(n) * K * s, cudaMemcpyHostToDevice, st));
    } else {
        const char* src = static_cast<const char*>(x0);
        std::vector<char> hx(size_t(m.nhalo) * s);
        for (int b = 0; b < K; ++b) {
            const char* col = src + size_t(b) * size_t(nglob) * s;
            char* dcol = d_xcol.as<char>() + size_t(b) * size_t(ext) * s;
            if (n) EIGSOL_CUDA(cudaMemcpyAsync(dcol, col + size_t(m.lo) * s, size_t(n) * s, cudaMemcpyHostToDevice, st));
            if (m.nhalo) {
                for (int64_t k = 0; k < m.nhalo; ++k) std::memcpy(hx.data() + size_t(k) * s, col + size_t(m.halo.cols[size_t(k)]) * s, s);
                EIGSOL_CUDA(cudaMemcpyAsync(dcol + size_t(m.nloc_pad) * s, hx.data(), hx.size(), cudaMemcpyHostToDevice, st));
                EIGSOL_CUDA(cudaStreamSynchronize(st));  // hx is reused for the next column
            }
        }
    }
    const int gn = int((ext + B - 1) / B);
    block_pack_kernel<S, K><<<gn, B, 0, st>>>(d_xcol.as<S>(), ext, ba.u[0]);
    BlockState* hs = h_state.as<BlockState>();
    std::memset(hs, 0, sizeof(BlockState));
    hs->k = K;
    hs->nev = nev;
    hs->max_trips = max_it;
    hs->tol = tol;
    hs->sigma = m.sigma0;
    EIGSOL_CUDA(cudaMemcpyAsync(d_state.p, hs, sizeof(BlockState), cudaMemcpyHostToDevice, st));
    block_gram_kernel<S, K, B><<<grid, B, 0, st>>>(ba.u[0], n, ba.partials, grid);
    block_control_kernel<1024><<<1, 1024, 0, st>>>(ba.st, ba.partials, grid, K, Sc<S>::is_complex ? 1 : 0, 1, peer, ++epoch);
    EIGSOL_CUDA(cudaGetLastError());
    res.launches += 3;
    int trip_no = 0;  // launch number of the next SpMM on the host side (the device counts in BlockState::launch_no)

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
        if (peer) {
            // the rows of the block just written that peers read -> their halo segments (buffer (L + 1) & 1)
            const int gp = std::max(1, std::min(m.sm_count * 2, int((int64_t(std::max<int64_t>(m.nhalo, 1)) * K + B - 1) / B)));
            launch_pdl(m.pdl, st, block_halo_push_kernel<S, K>, dim3(gp), dim3(B), 0, peer, static_cast<const S*>(ba.u[(trip_no + 1) & 1]),
                       (trip_no + 1) & 1);
            ++res.launches;
        }
        ++trip_no;
        launch_pdl(m.pdl, st, block_control_kernel<1024>, dim3(1), dim3(1024), 0, ba.st, ba.partials, grid, K,
                   Sc<S>::is_complex ? 1 : 0, 0, peer, ++epoch);
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

    if (hs->stop == BSTOP_ERROR)
        throw DeviceError("powerMethod: device error: a peer GPU did not signal within the timeout "
                          "(EIGSOL_B200_PEER_TIMEOUT_MS); the ranks are out of step");
    if (multi) comm_min_int(m.comm, 1);  // every rank is past its last launch before any buffer of this solve is freed
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
    block_finish_kernel<S, K><<<gn, B, 0, st>>>(ba.u[f & 1], bm, ext, d_xcol.as<S>());
    EIGSOL_CUDA(cudaGetLastError());
    ++m.launches;
    if (!multi) {
        EIGSOL_CUDA(cudaMemcpyAsync(x_out, d_xcol.p, size_t(n) * K * s, cudaMemcpyDeviceToHost, st));
        EIGSOL_CUDA(cudaStreamSynchronize(st));
    } else {
        // column by column: own rows of every rank -> the global column of the caller's n x K block
        for (int b = 0; b < K; ++b)
            m.gather_download(d_xcol.as<char>() + size_t(b) * size_t(ext) * s, static_cast<char*>(x_out) + size_t(b) * size_t(nglob) * s);
    }
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
    if (m.nranks != 1 && m.dense) throw PlanError("the multi-GPU block power iteration handles sparse matrices (dense: one GPU)");
    if (nev < 1 || nev > k) throw PlanError("nev must satisfy 1 <= nev <= block width");
    if (!x0 || !lambda_out) throw PlanError("null pointer");
    EIGSOL_CUDA(cudaSetDevice(m.device));
    if (m.dtype) block_run_k<double2>(m, k, nev, x0, max_it, tol, x_out, lambda_out, res);
    else block_run_k<double>(m, k, nev, x0, max_it, tol, x_out, lambda_out, res);
}

}  // namespace eigsol_b200
