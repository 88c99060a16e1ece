// Test harness (CPU): the k x k algebra and the control step of the block power iteration
// (csrc/blockmath.hpp) compiled by g++ and driven from tests/test_block_math.py.  Matrices cross as compact
// row-major k x k arrays of (re, im) doubles.
#include <cstring>

#include "../../pcsc_eigenvalue_solver_project_b200/csrc/blockmath.hpp"

using namespace eigsol_b200;

static void widen(const double* in, int k, Cx* out) {
    for (int i = 0; i < kBlockElems; ++i) out[i] = cx(0.0);
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) out[i * kBlockLd + j] = cx(in[2 * (i * k + j)], in[2 * (i * k + j) + 1]);
}
static void narrow(const Cx* in, int k, double* out) {
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) {
            out[2 * (i * k + j)] = in[i * kBlockLd + j].re;
            out[2 * (i * k + j) + 1] = in[i * kBlockLd + j].im;
        }
}

static BlockState g_state;

extern "C" {
int bm_chol(const double* w, int k, double* r_out) {
    Cx W[kBlockElems], R[kBlockElems];
    widen(w, k, W);
    std::memset(R, 0, sizeof R);
    const bool ok = chol_upper(W, k, R);
    narrow(R, k, r_out);
    return ok ? 1 : 0;
}
void bm_triinv(const double* r, int k, double* t_out) {
    Cx R[kBlockElems], T[kBlockElems];
    widen(r, k, R);
    tri_inverse_upper(R, k, T);
    narrow(T, k, t_out);
}
int bm_eig(const double* h, int k, double* theta_out, double* s_out) {
    Cx H[kBlockElems], S[kBlockElems], th[kBlockMax];
    widen(h, k, H);
    const bool ok = eig_general(H, k, th, S);
    for (int i = 0; i < k; ++i) {
        theta_out[2 * i] = th[i].re;
        theta_out[2 * i + 1] = th[i].im;
    }
    narrow(S, k, s_out);
    return ok ? 1 : 0;
}
int bm_schur(const double* h, int k, double* t_out, double* z_out) {
    Cx M[kBlockElems], Z[kBlockElems];
    widen(h, k, M);
    const bool ok = schur_complex(M, k, Z);
    narrow(M, k, t_out);
    narrow(Z, k, z_out);
    return ok ? 1 : 0;
}
// the iteration: begin(W0) then step(C, W) after every product; T of the NEXT launch and the result transform
void bm_begin(int k, int nev, int max_trips, double tol, double sigma, const double* w0) {
    std::memset(&g_state, 0, sizeof g_state);
    g_state.k = k;
    g_state.nev = nev;
    g_state.max_trips = max_trips;
    g_state.tol = tol;
    g_state.sigma = sigma;
    Cx W[kBlockElems];
    widen(w0, k, W);
    block_control_begin(&g_state, W);
}
void bm_step(const double* c, const double* w) {
    Cx C[kBlockElems], W[kBlockElems];
    widen(c, g_state.k, C);
    widen(w, g_state.k, W);
    block_control_step(&g_state, C, W);
}
// ints: launch_no, iterations, converged, stop, final_launch
void bm_state(int* ints, double* theta, double* t_next, double* result_transform) {
    ints[0] = g_state.launch_no;
    ints[1] = g_state.iterations;
    ints[2] = g_state.converged;
    ints[3] = g_state.stop;
    ints[4] = g_state.final_launch;
    for (int i = 0; i < g_state.k; ++i) {
        theta[2 * i] = g_state.theta[i].re;
        theta[2 * i + 1] = g_state.theta[i].im;
    }
    narrow(g_state.T[g_state.launch_no & 1], g_state.k, t_next);
    BlockState fin = g_state;  // the host's end-of-loop step, on a copy
    block_result_vectors(&fin);
    Cx M[kBlockElems];
    block_result_transform(&fin, M);
    narrow(M, g_state.k, result_transform);
}
}
