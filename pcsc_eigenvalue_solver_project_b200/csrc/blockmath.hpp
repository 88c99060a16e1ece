// blockmath.hpp -- the k x k dense algebra and the control step of the BLOCK power iteration.
//
// Block (simultaneous) power iteration keeps k vectors instead of the one of power_method.hpp:68-96:
//     Q_m = orth(U_m)            U_{m+1} = A Q_m            H_m = Q_m^H A Q_m  (k x k)
// and reads the k dominant eigenvalues off H_m (Rayleigh-Ritz).  With k = 1 this IS the reference's loop:
// Q = x / ||x||, H = x^H A x (power_method.hpp:78,81), same convergence test (:83-91, tolerance.hpp:28-33) on
// every wanted Ritz value, same `initialized` / usedIters bookkeeping.  The n-sized work (A Q, the Gram
// sums) is one pass of the SpMM kernel in block.cuh; everything here is k x k, k <= 8, and runs in ONE
// thread of the control kernel between two launches -- and, compiled by g++, in tests/blockmath on the CPU.
//
//   orth by Cholesky QR:   W = U^H U = R^H R  ->  Q = U T,  T = R^{-1}      (chol_upper, tri_inverse_upper)
//   Ritz pairs:            H = Z Tm Z^H (complex Schur: Givens Hessenberg reduction + shifted QR), then the
//                          eigenvectors of the triangular factor by back substitution   (eig_general)
// All arithmetic is complex double; a real matrix simply has conjugate-pair Ritz values.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define EB_HD __host__ __device__ __forceinline__
#define EB_HDN __host__ __device__ __noinline__
#else
#define EB_HD inline
#define EB_HDN inline
#endif

namespace eigsol_b200 {

constexpr int kBlockMax = 8;               // largest block width
constexpr int kBlockLd = kBlockMax;        // row stride of every k x k array below
constexpr int kBlockElems = kBlockMax * kBlockMax;

struct Cx {
    double re, im;
};
EB_HD Cx cx(double re, double im = 0.0) {
    Cx r;
    r.re = re;
    r.im = im;
    return r;
}
EB_HD Cx operator+(Cx a, Cx b) { return cx(a.re + b.re, a.im + b.im); }
EB_HD Cx operator-(Cx a, Cx b) { return cx(a.re - b.re, a.im - b.im); }
EB_HD Cx operator-(Cx a) { return cx(-a.re, -a.im); }
EB_HD Cx operator*(Cx a, Cx b) { return cx(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
EB_HD Cx operator*(double s, Cx a) { return cx(s * a.re, s * a.im); }
EB_HD Cx conj(Cx a) { return cx(a.re, -a.im); }
EB_HD double cabs(Cx a) { return hypot(a.re, a.im); }
EB_HD double cabs2(Cx a) { return a.re * a.re + a.im * a.im; }
EB_HD Cx operator/(Cx a, Cx b) {  // Smith's algorithm: no overflow in |b|^2
    if (fabs(b.re) >= fabs(b.im)) {
        const double r = b.im / b.re, d = b.re + b.im * r;
        return cx((a.re + a.im * r) / d, (a.im - a.re * r) / d);
    }
    const double r = b.re / b.im, d = b.re * r + b.im;
    return cx((a.re * r + a.im) / d, (a.im * r - a.re) / d);
}
// Cheap forms for data already scaled to O(1) (eig_general divides H by its largest entry first): no hypot,
// one real division per complex division.  On the device these run in ONE thread between two launches.
EB_HD double cabs_fast(Cx a) { return sqrt(a.re * a.re + a.im * a.im); }
EB_HD double cabs1(Cx a) { return fabs(a.re) + fabs(a.im); }
EB_HD Cx cdiv_fast(Cx a, Cx b) {
    const double inv = 1.0 / (b.re * b.re + b.im * b.im);
    return cx((a.re * b.re + a.im * b.im) * inv, (a.im * b.re - a.re * b.im) * inv);
}
EB_HD Cx csqrt_(Cx a) {  // principal square root
    const double m = cabs_fast(a);
    if (m == 0.0) return cx(0.0, 0.0);
    const double s = sqrt(0.5 * (m + fabs(a.re)));
    const double t = 0.5 * a.im / s;
    if (a.re >= 0.0) return cx(s, t);
    return cx(fabs(t), a.im >= 0.0 ? s : -s);
}

// ---- Cholesky QR ------------------------------------------------------------------------------------
// W (k x k Hermitian, only the upper triangle is read) = R^H R with R upper triangular, real positive
// diagonal.  Returns false when a pivot falls below 1e-13 of its diagonal entry: the block has (numerically)
// lost rank and R^{-1} would not orthonormalise it.
EB_HDN bool chol_upper(const Cx* W, int k, Cx* R) {
    for (int j = 0; j < k; ++j) {
        double d = W[j * kBlockLd + j].re;
        for (int p = 0; p < j; ++p) d -= cabs2(R[p * kBlockLd + j]);
        const double wjj = W[j * kBlockLd + j].re;
        if (!(d > 1e-13 * wjj) || !(wjj > 0.0) || !(wjj < HUGE_VAL)) return false;  // also NaN / inf / zero block
        const double r = sqrt(d);
        for (int c = 0; c < j; ++c) R[j * kBlockLd + c] = cx(0.0);
        R[j * kBlockLd + j] = cx(r);
        for (int c = j + 1; c < k; ++c) {
            Cx s = W[j * kBlockLd + c];
            for (int p = 0; p < j; ++p) s = s - conj(R[p * kBlockLd + j]) * R[p * kBlockLd + c];
            R[j * kBlockLd + c] = cx(s.re / r, s.im / r);
        }
    }
    return true;
}

// T = R^{-1}, R upper triangular with non-zero diagonal.
EB_HDN void tri_inverse_upper(const Cx* R, int k, Cx* T) {
    for (int c = 0; c < k; ++c) {
        for (int r = c + 1; r < k; ++r) T[r * kBlockLd + c] = cx(0.0);
        T[c * kBlockLd + c] = cx(1.0) / R[c * kBlockLd + c];
        for (int r = c - 1; r >= 0; --r) {
            Cx s = cx(0.0);
            for (int p = r + 1; p <= c; ++p) s = s + R[r * kBlockLd + p] * T[p * kBlockLd + c];
            T[r * kBlockLd + c] = -(s / R[r * kBlockLd + r]);
        }
    }
}

// ---- complex Schur decomposition of a small general matrix -------------------------------------------
// Rotation G = [c s; -conj(s) c] (c real) with G [a; b] = [r; 0].
EB_HD void givens(Cx a, Cx b, double& c, Cx& s) {
    const double nb2 = cabs2(b);
    if (nb2 == 0.0) {
        c = 1.0;
        s = cx(0.0);
        return;
    }
    const double na2 = cabs2(a);
    if (na2 == 0.0) {
        const double inv = 1.0 / sqrt(nb2);
        c = 0.0;
        s = cx(b.re * inv, -b.im * inv);
        return;
    }
    const double na = sqrt(na2), inv_nrm = 1.0 / sqrt(na2 + nb2);
    c = na * inv_nrm;
    const double f = inv_nrm / na;  // s = (a / |a|) conj(b) / nrm
    const Cx ab = a * conj(b);
    s = cx(ab.re * f, ab.im * f);
}
// rows (i, j) of M <- G * rows, columns q0..k-1
EB_HD void rot_rows(Cx* M, int k, int i, int j, double c, Cx s, int q0) {
    for (int q = q0; q < k; ++q) {
        const Cx x = M[i * kBlockLd + q], y = M[j * kBlockLd + q];
        M[i * kBlockLd + q] = c * x + s * y;
        M[j * kBlockLd + q] = c * y - conj(s) * x;
    }
}
// columns (i, j) of M <- columns * G^H, rows 0..p1-1
EB_HD void rot_cols(Cx* M, int i, int j, double c, Cx s, int p1) {
    for (int p = 0; p < p1; ++p) {
        const Cx x = M[p * kBlockLd + i], y = M[p * kBlockLd + j];
        M[p * kBlockLd + i] = c * x + conj(s) * y;
        M[p * kBlockLd + j] = c * y - s * x;
    }
}

// In place: M (general k x k) -> upper triangular Tm, Z unitary with M_in = Z Tm Z^H.  Returns false if
// the QR iteration did not converge (60 sweeps on one eigenvalue).
//
// Written for `nl` cooperating lanes (lane = 0..nl-1) that share M and Z and step together: every lane
// evaluates the (identical) scalar decisions and rotation coefficients, and applies a rotation to its own
// columns / rows; EB_SYNC() separates the phases.  On the device the lanes are the 32 threads of one warp
// and M, Z live in shared memory; the host runs the same code with nl = 1.  Element by element the arithmetic
// is the same for every nl, so the host run is a bit-exact model of the device run.
// Z == nullptr: eigenvalues only -- rotations are then confined to the active window [l, hi].
#if defined(__CUDA_ARCH__)
#define EB_SYNC() __syncwarp()
#else
#define EB_SYNC() ((void)0)
#endif

EB_HDN bool schur_coop(Cx* M, int k, Cx* Z, int lane, int nl) {
    if (Z)
        for (int e = lane; e < k * k; e += nl) Z[(e / k) * kBlockLd + e % k] = cx(e / k == e % k ? 1.0 : 0.0);
    EB_SYNC();
    // Hessenberg form by Givens similarity transforms
    for (int j = 0; j + 2 < k; ++j)
        for (int i = k - 1; i >= j + 2; --i) {
            const Cx a = M[(i - 1) * kBlockLd + j], b = M[i * kBlockLd + j];
            EB_SYNC();  // everyone has read the pivot pair
            if (cabs1(b) == 0.0) continue;
            double c;
            Cx s;
            givens(a, b, c, s);
            for (int q = j + lane; q < k; q += nl) {  // rows (i-1, i) <- G rows
                const Cx x = M[(i - 1) * kBlockLd + q], y = M[i * kBlockLd + q];
                M[(i - 1) * kBlockLd + q] = c * x + s * y;
                M[i * kBlockLd + q] = q == j ? cx(0.0) : c * y - conj(s) * x;
            }
            EB_SYNC();
            for (int p = lane; p < k; p += nl) {  // columns (i-1, i) <- columns G^H
                const Cx x = M[p * kBlockLd + i - 1], y = M[p * kBlockLd + i];
                M[p * kBlockLd + i - 1] = c * x + conj(s) * y;
                M[p * kBlockLd + i] = c * y - s * x;
                if (Z) {
                    const Cx zx = Z[p * kBlockLd + i - 1], zy = Z[p * kBlockLd + i];
                    Z[p * kBlockLd + i - 1] = c * zx + conj(s) * zy;
                    Z[p * kBlockLd + i] = c * zy - s * zx;
                }
            }
            EB_SYNC();
        }
    const double eps = 2.220446049250313e-16;
    double norm = 0.0;
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) norm = fmax(norm, cabs1(M[i * kBlockLd + j]));
    int hi = k - 1, sweeps = 0;
    double cs[kBlockMax];
    Cx sn[kBlockMax];
    while (hi > 0) {
        int l = hi;
        for (; l > 0; --l) {
            const double sub = cabs1(M[l * kBlockLd + l - 1]);  // |re| + |im|, as LAPACK's zlahqr deflation test
            double diag = cabs1(M[(l - 1) * kBlockLd + l - 1]) + cabs1(M[l * kBlockLd + l]);
            if (diag == 0.0) diag = norm;
            if (sub <= eps * diag) break;
        }
        EB_SYNC();
        if (l > 0 && lane == 0) M[l * kBlockLd + l - 1] = cx(0.0);
        EB_SYNC();
        if (l == hi) {  // M[hi][hi] is an eigenvalue
            --hi;
            sweeps = 0;
            continue;
        }
        if (++sweeps > 60) return false;
        // Wilkinson shift: the eigenvalue of the trailing 2 x 2 nearer to its last diagonal entry
        const Cx a = M[(hi - 1) * kBlockLd + hi - 1], b = M[(hi - 1) * kBlockLd + hi];
        const Cx c2 = M[hi * kBlockLd + hi - 1], d = M[hi * kBlockLd + hi];
        const double below = hi >= 2 ? cabs1(M[(hi - 1) * kBlockLd + hi - 2]) : 0.0;
        EB_SYNC();
        Cx mu;
        if (sweeps == 10 || sweeps == 20 || sweeps == 30) {
            mu = cx(cabs1(c2) + below) + d;  // exceptional shift
        } else {
            const Cx half = 0.5 * (a - d);
            const Cx disc = csqrt_(half * half + b * c2);
            const Cx m1 = 0.5 * (a + d) + disc, m2 = 0.5 * (a + d) - disc;
            mu = cabs2(m1 - d) <= cabs2(m2 - d) ? m1 : m2;
        }
        for (int i = l + lane; i <= hi; i += nl) M[i * kBlockLd + i] = M[i * kBlockLd + i] - mu;
        EB_SYNC();
        const int qend = Z ? k : hi + 1;  // eigenvalues only: stay inside the window
        for (int i = l; i < hi; ++i) {    // M - mu = Q R
            const Cx pa = M[i * kBlockLd + i], pb = M[(i + 1) * kBlockLd + i];
            EB_SYNC();
            givens(pa, pb, cs[i], sn[i]);
            for (int q = i + lane; q < qend; q += nl) {
                const Cx x = M[i * kBlockLd + q], y = M[(i + 1) * kBlockLd + q];
                M[i * kBlockLd + q] = cs[i] * x + sn[i] * y;
                M[(i + 1) * kBlockLd + q] = q == i ? cx(0.0) : cs[i] * y - conj(sn[i]) * x;
            }
            EB_SYNC();
        }
        const int pbeg = Z ? 0 : l;
        for (int i = l; i < hi; ++i) {  // R Q (and, with Z, the rows above the window: the full Schur form)
            for (int p = pbeg + lane; p < (Z ? k : i + 2); p += nl) {
                if (p < i + 2) {
                    const Cx x = M[p * kBlockLd + i], y = M[p * kBlockLd + i + 1];
                    M[p * kBlockLd + i] = cs[i] * x + conj(sn[i]) * y;
                    M[p * kBlockLd + i + 1] = cs[i] * y - sn[i] * x;
                }
                if (Z) {
                    const Cx zx = Z[p * kBlockLd + i], zy = Z[p * kBlockLd + i + 1];
                    Z[p * kBlockLd + i] = cs[i] * zx + conj(sn[i]) * zy;
                    Z[p * kBlockLd + i + 1] = cs[i] * zy - sn[i] * zx;
                }
            }
            EB_SYNC();
        }
        for (int i = l + lane; i <= hi; i += nl) M[i * kBlockLd + i] = M[i * kBlockLd + i] + mu;
        EB_SYNC();
    }
    for (int e = lane; e < k * k; e += nl)
        if (e / k > e % k) M[(e / k) * kBlockLd + e % k] = cx(0.0);
    EB_SYNC();
    return true;
}
EB_HDN bool schur_complex(Cx* M, int k, Cx* Z) { return schur_coop(M, k, Z, 0, 1); }

// Order of the Ritz values: larger modulus first; moduli equal to 1e-9 relative (conjugate pairs, multiple
// eigenvalues) are ordered by imaginary part, then real part, so the order is stable from trip to trip.
EB_HD bool ritz_before(Cx a, Cx b) {
    const double ma = cabs_fast(a), mb = cabs_fast(b), sc = fmax(ma, mb);  // called on the scaled matrix
    const double tie = 1e-9 * sc;
    if (fabs(ma - mb) > tie) return ma > mb;
    if (fabs(a.im - b.im) > tie) return a.im > b.im;
    if (fabs(a.re - b.re) > tie) return a.re > b.re;
    return false;
}

// 2^e ~ max|h_ij| (e clamped): the Schur iteration runs on H * 2^-e, exactly representable.
EB_HD bool eig_scale(const Cx* H, int k, double& hs, double& hback) {
    double hmax = 0.0;
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) hmax = fmax(hmax, fmax(fabs(H[i * kBlockLd + j].re), fabs(H[i * kBlockLd + j].im)));
    if (!(hmax < HUGE_VAL)) return false;  // inf / NaN
    int he = 0;
    if (hmax > 0.0) frexp(hmax, &he);
    he = he < -1000 ? -1000 : (he > 1000 ? 1000 : he);
    hs = ldexp(1.0, -he);
    hback = ldexp(1.0, he);
    return true;
}
// ord[c] = index of the c-th diagonal entry of the triangular M in ritz_before order (insertion sort, stable)
EB_HD void ritz_order(const Cx* M, int k, int* ord) {
    for (int i = 0; i < k; ++i) ord[i] = i;
    for (int i = 1; i < k; ++i) {
        const int o = ord[i];
        int j = i - 1;
        while (j >= 0 && ritz_before(M[o * kBlockLd + o], M[ord[j] * kBlockLd + ord[j]])) {
            ord[j + 1] = ord[j];
            --j;
        }
        ord[j + 1] = o;
    }
}

// The ordered eigenvalues only, by `nl` cooperating lanes (see schur_coop); `work` is a k x k scratch shared by
// the lanes, theta is each lane's own (identical) copy.  This is what runs between two launches: the convergence
// test needs no vectors.  The values are bit-identical to eig_general's.
EB_HDN bool eig_values_coop(const Cx* H, int k, Cx* theta, Cx* work, int lane, int nl) {
    double hs, hback;
    if (!eig_scale(H, k, hs, hback)) return false;
    for (int e = lane; e < k * k; e += nl) {
        const int i = (e / k) * kBlockLd + e % k;
        work[i] = cx(H[i].re * hs, H[i].im * hs);
    }
    EB_SYNC();
    if (!schur_coop(work, k, nullptr, lane, nl)) return false;
    int ord[kBlockMax];
    ritz_order(work, k, ord);
    for (int c = 0; c < k; ++c) {
        const Cx d = work[ord[c] * kBlockLd + ord[c]];
        theta[c] = cx(d.re * hback, d.im * hback);
    }
    EB_SYNC();
    return true;
}

// Eigenvalues theta[0..k) (ordered by ritz_before) and unit-norm eigenvectors (columns of S, largest component
// real and positive) of the general matrix H.  H is not modified.  false: the QR iteration failed.
EB_HDN bool eig_general(const Cx* H, int k, Cx* theta, Cx* S) {
    Cx M[kBlockElems], Z[kBlockElems], V[kBlockElems];
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j)  // fmax() in eig_scale skips NaN: look at every entry here
            if (!(fabs(H[i * kBlockLd + j].re) < HUGE_VAL) || !(fabs(H[i * kBlockLd + j].im) < HUGE_VAL)) return false;
    // work on H * 2^-e with 2^e ~ max|h_ij| (exact; the Ritz values are multiplied back at the end), so the
    // plain sqrt / reciprocal forms above can neither overflow nor underflow
    double hs, hback;
    if (!eig_scale(H, k, hs, hback)) return false;
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) M[i * kBlockLd + j] = cx(H[i * kBlockLd + j].re * hs, H[i * kBlockLd + j].im * hs);
    if (!schur_complex(M, k, Z)) return false;
    double norm = 0.0;
    for (int i = 0; i < k; ++i)
        for (int j = i; j < k; ++j) norm = fmax(norm, cabs1(M[i * kBlockLd + j]));
    const double small = fmax(norm * 2.220446049250313e-16, 1e-300);
    // eigenvectors of the triangular factor: column i of V
    for (int i = 0; i < k; ++i) {
        const Cx lam = M[i * kBlockLd + i];
        for (int r = i + 1; r < k; ++r) V[r * kBlockLd + i] = cx(0.0);
        V[i * kBlockLd + i] = cx(1.0);
        for (int r = i - 1; r >= 0; --r) {
            Cx s = cx(0.0);
            for (int p = r + 1; p <= i; ++p) s = s + M[r * kBlockLd + p] * V[p * kBlockLd + i];
            Cx den = M[r * kBlockLd + r] - lam;
            if (cabs1(den) < small) den = cx(small);
            V[r * kBlockLd + i] = -(s / den);  // Smith's division: den may be as small as 1e-300 (H = 0)
        }
    }
    int ord[kBlockMax];
    ritz_order(M, k, ord);
    for (int c = 0; c < k; ++c) {
        const int i = ord[c];
        theta[c] = cx(M[i * kBlockLd + i].re * hback, M[i * kBlockLd + i].im * hback);
        double n2 = 0.0, best = -1.0;
        Cx big = cx(1.0);
        for (int r = 0; r < k; ++r) {
            Cx s = cx(0.0);
            for (int p = 0; p <= i; ++p) s = s + Z[r * kBlockLd + p] * V[p * kBlockLd + i];
            S[r * kBlockLd + c] = s;
            const double a2 = cabs2(s);
            n2 += a2;
            if (a2 > best * (1.0 + 1e-12)) {  // first of (nearly) equal maxima: a stable choice
                best = a2;
                big = s;
            }
        }
        const double nb = cabs_fast(big);
        const Cx ph = nb > 0.0 ? cx(big.re / nb, -big.im / nb) : cx(1.0);  // conj(big)/|big|
        const double inv = n2 > 0.0 ? 1.0 / sqrt(n2) : 1.0;
        for (int r = 0; r < k; ++r) S[r * kBlockLd + c] = inv * (ph * S[r * kBlockLd + c]);
    }
    return true;
}

// ---- the iteration state and its control step ---------------------------------------------------------
enum { BSTOP_NONE = 0, BSTOP_CONVERGED = 1, BSTOP_BREAKDOWN = 2, BSTOP_MAX_TRIPS = 3, BSTOP_ERROR = 4 /* a peer GPU never arrived */ };

struct BlockState {
    int k;            // block width
    int nev;          // Ritz values that must pass the convergence test (1 <= nev <= k)
    int max_trips;    // opts.maxIterations
    int launch_no;    // SpMM launches folded so far = number of the next launch; launch L reads U buffer L & 1
    int iterations;   // usedIters (power_method.hpp:63)
    int initialized;  // power_method.hpp:66
    int converged;
    int stop;         // BSTOP_*
    int final_launch; // launch whose input block, T and S give the returned Ritz vectors: X = U T S
    int reserved;
    double tol;       // opts.tolerance
    double sigma;     // power-of-two prescale of A (U_{m+1} = sigma A Q_m), H = C / sigma
    Cx theta[kBlockMax];        // current Ritz values, ordered
    Cx T[2][kBlockElems];       // T[L & 1] = R^{-1} used by launch L
    Cx S[kBlockElems];          // Ritz vectors of H_{final_launch} in the Q basis (block_result_vectors, after the loop)
    Cx H[kBlockElems];          // H_{final_launch}
};

EB_HD bool block_close_relative(Cx a, Cx b, double tol) {  // tolerance.hpp:28-33, a = new value
    return cabs(a - b) <= tol * (1.0 + cabs(a));
}

// After the Gram pass over the start block: W = U_0^H U_0.
EB_HDN void block_control_begin(BlockState* st, const Cx* W) {
    Cx R[kBlockElems];
    st->launch_no = 0;
    st->iterations = 0;
    st->initialized = 0;
    st->converged = 0;
    st->stop = BSTOP_NONE;
    st->final_launch = -1;
    for (int i = 0; i < st->k; ++i) st->theta[i] = cx(0.0);  // Scalar lambda = Scalar(0), power_method.hpp:61
    for (int i = 0; i < kBlockElems; ++i) st->S[i] = cx((i / kBlockLd) == (i % kBlockLd) ? 1.0 : 0.0);
    if (!chol_upper(W, st->k, R)) {
        st->stop = BSTOP_BREAKDOWN;
        return;
    }
    tri_inverse_upper(R, st->k, st->T[0]);
    if (st->max_trips <= 0) st->stop = BSTOP_MAX_TRIPS;
}

// After SpMM launch L = st->launch_no:  C = Q_L^H (sigma A Q_L),  W = U_{L+1}^H U_{L+1}.
// Launch 0 is the set-up product A Q_0 (the reference never forms x0^H A x0); launch L >= 1 closes trip
// L - 1 of power_method.hpp:68-96 for every wanted Ritz value.  The step is three independent pieces so the
// control kernel can run them side by side:
//     H = C / sigma and its ordered eigenvalues                       (one warp, cooperatively)
//     W = R^H R,  T_{L+1} = R^{-1} written into st->T[(L + 1) & 1]     (another warp; harmless if the loop stops)
//     the decision below                                              (one thread, after both)
EB_HD void block_step_H(const BlockState* st, const Cx* C, Cx* Hn, int lane, int nl) {
    const double inv_sigma = 1.0 / st->sigma;
    const int k = st->k;
    for (int e = lane; e < k * k; e += nl) {
        const int i = (e / k) * kBlockLd + e % k;
        Hn[i] = inv_sigma * C[i];
    }
}
EB_HDN bool block_step_T(BlockState* st, const Cx* W) {
    Cx R[kBlockElems];
    if (!chol_upper(W, st->k, R)) return false;  // A Q lost rank: the block counterpart of normY == 0 (power_method.hpp:73-76)
    tri_inverse_upper(R, st->k, st->T[(st->launch_no + 1) & 1]);
    return true;
}
EB_HDN void block_step_decide(BlockState* st, const Cx* Hn, const Cx* theta, bool eig_ok, bool chol_ok) {
    const int k = st->k, L = st->launch_no;
    st->launch_no = L + 1;
    if (!eig_ok) {
        st->stop = BSTOP_BREAKDOWN;
        return;
    }
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) st->H[i * kBlockLd + j] = Hn[i * kBlockLd + j];
    st->final_launch = L;
    if (L >= 1) {
        bool all_close = st->initialized != 0;
        for (int i = 0; i < st->nev && all_close; ++i) all_close = block_close_relative(theta[i], st->theta[i], st->tol);
        for (int i = 0; i < k; ++i) st->theta[i] = theta[i];
        st->iterations = L;  // usedIters = k + 1 with trip k = L - 1
        if (all_close) {
            st->converged = 1;
            st->stop = BSTOP_CONVERGED;
            return;
        }
        st->initialized = 1;
        if (st->iterations >= st->max_trips) {
            st->stop = BSTOP_MAX_TRIPS;
            return;
        }
    }
    if (!chol_ok) st->stop = BSTOP_BREAKDOWN;
}
// The same step on one thread (the host restatement in tests/blockmath).
EB_HDN void block_control_step(BlockState* st, const Cx* C, const Cx* W) {
    Cx Hn[kBlockElems], work[kBlockElems], theta[kBlockMax];
    block_step_H(st, C, Hn, 0, 1);
    const bool eig_ok = eig_values_coop(Hn, st->k, theta, work, 0, 1);
    const bool chol_ok = block_step_T(st, W);
    block_step_decide(st, Hn, theta, eig_ok, chol_ok);
}

// After the loop (host): the Ritz vectors of H_{final_launch} in the Q basis.
EB_HDN void block_result_vectors(BlockState* st) {
    Cx theta[kBlockMax];
    if (st->final_launch >= 0 && eig_general(st->H, st->k, theta, st->S)) return;
    for (int i = 0; i < kBlockElems; ++i) st->S[i] = cx((i / kBlockLd) == (i % kBlockLd) ? 1.0 : 0.0);
}

// M = T S (k x k): the returned Ritz vectors are X = U M.
EB_HD void block_result_transform(const BlockState* st, Cx* M) {
    const int k = st->k;
    const Cx* T = st->T[(st->final_launch < 0 ? 0 : st->final_launch) & 1];
    for (int i = 0; i < k; ++i)
        for (int j = 0; j < k; ++j) {
            Cx s = cx(0.0);
            for (int p = i; p < k; ++p) s = s + T[i * kBlockLd + p] * st->S[p * kBlockLd + j];
            M[i * kBlockLd + j] = s;
        }
}

}  // namespace eigsol_b200
