/*
 * eigsol_oracle.c -- CPU restatement of the reference's power-iteration hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker (or as the timed CPU baseline), never as the thing shipped.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (1) the reference itself compiled from /root/reference (oracle/_ref/libeigsol_ref.so, built by
 *       oracle/Makefile from oracle/ref_shim.cpp) when that library is present, and
 *   (2) golden vectors produced by that compiled reference and committed under tests/golden/
 *       (generator: tests/golden/make_golden.py),
 * plus the five known-answer cases of the reference's own test/power_method_test.cpp:38-134.
 *
 * All file:line citations are relative to the reference checkout (/root/reference).
 * Complex numbers cross this C ABI as interleaved (re, im) doubles, i.e. the memory layout of
 * std::complex<double>.
 *
 * Plain C, single-threaded on purpose: the reference's path is single-threaded (the col-major
 * sparse x dense kernel has no OpenMP branch, eigen/Eigen/src/SparseCore/SparseDenseProduct.h:94-111).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORACLE_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * matvec: sparse (CSC, int32) -- eigen/Eigen/src/SparseCore/SparseDenseProduct.h:94-111
 *   res = 0; for each column j: rhs_j = alpha * x[j] (alpha = 1); for (i, v) in column j:
 *   res[i] += v * rhs_j.  The zeroing of res is the setZero() the product evaluator performs
 *   before scaleAndAddTo (eigen/Eigen/src/SparseCore/SparseDenseProduct.h:  evalTo -> dst.setZero()).
 * ------------------------------------------------------------------------------------------ */
ORACLE_API void oracle_spmv_csc_f64(int64_t rows, int64_t cols, const int32_t *outer,
                                    const int32_t *inner, const double *vals, const double *x,
                                    double *y)
{
    for (int64_t i = 0; i < rows; ++i) y[i] = 0.0;
    for (int64_t j = 0; j < cols; ++j) {
        const double xj = 1.0 * x[j];
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p) y[inner[p]] += vals[p] * xj;
    }
}

ORACLE_API void oracle_spmv_csc_c128(int64_t rows, int64_t cols, const int32_t *outer,
                                     const int32_t *inner, const double *vals, const double *x,
                                     double *y)
{
    for (int64_t i = 0; i < 2 * rows; ++i) y[i] = 0.0;
    for (int64_t j = 0; j < cols; ++j) {
        const double xr = x[2 * j], xi = x[2 * j + 1];
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p) {
            const double ar = vals[2 * (int64_t)p], ai = vals[2 * (int64_t)p + 1];
            const int64_t i = inner[p];
            y[2 * i] += ar * xr - ai * xi;
            y[2 * i + 1] += ar * xi + ai * xr;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * matvec: dense, column-major -- eigen/Eigen/src/Core/GeneralProduct.h:272 ->
 * eigen/Eigen/src/Core/products/GeneralMatrixVector.h:80-108 (column-panel axpy:
 * res += A[:, j] * x[j], columns visited in ascending order).
 * ------------------------------------------------------------------------------------------ */
ORACLE_API void oracle_gemv_colmajor_f64(int64_t rows, int64_t cols, const double *A, int64_t lda,
                                         const double *x, double *y)
{
    for (int64_t i = 0; i < rows; ++i) y[i] = 0.0;
    for (int64_t j = 0; j < cols; ++j) {
        const double xj = x[j];
        const double *col = A + j * lda;
        for (int64_t i = 0; i < rows; ++i) y[i] += col[i] * xj;
    }
}

ORACLE_API void oracle_gemv_colmajor_c128(int64_t rows, int64_t cols, const double *A, int64_t lda,
                                          const double *x, double *y)
{
    for (int64_t i = 0; i < 2 * rows; ++i) y[i] = 0.0;
    for (int64_t j = 0; j < cols; ++j) {
        const double xr = x[2 * j], xi = x[2 * j + 1];
        const double *col = A + 2 * j * lda;
        for (int64_t i = 0; i < rows; ++i) {
            const double ar = col[2 * i], ai = col[2 * i + 1];
            y[2 * i] += ar * xr - ai * xi;
            y[2 * i + 1] += ar * xi + ai * xr;
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * norm: sqrt(sum |y_i|^2), no overflow rescaling -- eigen/Eigen/src/Core/Dot.h:61-77.
 * (Eigen reduces with a vectorised tree; this sums left to right.  The two differ by a few
 * ulp, which is why every comparison against the reference is a tolerance test.)
 * ------------------------------------------------------------------------------------------ */
ORACLE_API double oracle_squared_norm(int64_t n_doubles, const double *v)
{
    double s = 0.0;
    for (int64_t i = 0; i < n_doubles; ++i) s += v[i] * v[i];
    return s;
}

/* conjugated inner product x^H z -- eigen/Eigen/src/Core/Dot.h:49-51,
 * eigen/Eigen/src/Core/InnerProduct.h:255 (first argument conjugated). */
ORACLE_API void oracle_dotc(int is_complex, int64_t n, const double *x, const double *z,
                            double out[2])
{
    double re = 0.0, im = 0.0;
    if (!is_complex) {
        for (int64_t i = 0; i < n; ++i) re += x[i] * z[i];
    } else {
        for (int64_t i = 0; i < n; ++i) {
            const double xr = x[2 * i], xi = x[2 * i + 1], zr = z[2 * i], zi = z[2 * i + 1];
            re += xr * zr + xi * zi;
            im += xr * zi - xi * zr;
        }
    }
    out[0] = re;
    out[1] = im;
}

/* relative closeness -- src/core/tolerance.hpp:28-33: |a-b| <= tol*(1+|a|); |.| is the complex
 * modulus (std::abs(std::complex) == hypot). */
ORACLE_API int oracle_is_close_relative(const double a[2], const double b[2], double tol)
{
    const double diff = hypot(a[0] - b[0], a[1] - b[1]);
    const double scale = 1.0 + hypot(a[0], a[1]);
    return diff <= tol * scale;
}

/* ------------------------------------------------------------------------------------------
 * x0 -- src/power_method/power_method.hpp:62: Vector<Scalar>::Random(n); x.normalize();
 *   Random: eigen/Eigen/src/Core/Random.h:20-23 -> RandomImpl.h:46-51 (std::rand()),
 *   :66-69 (bit assembly, kEntropy = 31 bits per rand() with glibc RAND_MAX = 2^31-1, so a
 *   52-bit mantissa takes two calls: first call -> bits 0..30, second call << 31, masked to 52),
 *   :89-99 (bits | bit_cast(2.0), minus 3.0 -> [-1, 1)), :256 (complex = two reals; g++ evaluates
 *   the two constructor arguments right to left, so the IMAGINARY part is drawn first -- verified
 *   against the compiled reference in tests/test_oracle.py).
 *   normalize: eigen/Eigen/src/Core/Dot.h:110-114 (divide by sqrt(squaredNorm) if > 0).
 * The caller seeds libc with srand() exactly as a caller of the reference would.
 * ------------------------------------------------------------------------------------------ */
static double eigen_random_real(void)
{
    uint64_t bits = 0;
    for (int shift = 0; shift < 52; shift += 31) bits |= ((uint64_t)rand()) << shift;
    bits &= (~(uint64_t)0) >> 12; /* keep 52 bits */
    const double two = 2.0;
    uint64_t tb;
    memcpy(&tb, &two, sizeof tb);
    bits |= tb;
    double r;
    memcpy(&r, &bits, sizeof r);
    return r - 3.0;
}

ORACLE_API void oracle_random_unit(int is_complex, int64_t n, double *x)
{
    if (!is_complex) {
        for (int64_t i = 0; i < n; ++i) x[i] = eigen_random_real();
    } else {
        for (int64_t i = 0; i < n; ++i) {
            const double im = eigen_random_real(); /* right-to-left argument evaluation */
            const double re = eigen_random_real();
            x[2 * i] = re;
            x[2 * i + 1] = im;
        }
    }
    const int64_t nd = is_complex ? 2 * n : n;
    const double z = oracle_squared_norm(nd, x);
    if (z > 0.0) {
        const double nrm = sqrt(z);
        for (int64_t i = 0; i < nd; ++i) x[i] /= nrm;
    }
}

/* ------------------------------------------------------------------------------------------
 * The power iteration -- src/power_method/power_method.hpp:47-99, restated line by line,
 * including the second matvec at :81.  x0 is explicit (the reference draws it at :62).
 *   kind: 0 = CSC sparse, 1 = dense column-major.   is_complex: 0 = double, 1 = complex<double>.
 * Returns 0, or 1/2 for the two runtime_error cases (:52-58) so tests can pin the order of checks.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    int kind, is_complex;
    int64_t rows, cols, lda;
    const int32_t *outer, *inner;
    const double *vals;
} oracle_mat;

static void oracle_matvec(const oracle_mat *m, const double *x, double *y)
{
    if (m->kind == 0) {
        if (m->is_complex)
            oracle_spmv_csc_c128(m->rows, m->cols, m->outer, m->inner, m->vals, x, y);
        else
            oracle_spmv_csc_f64(m->rows, m->cols, m->outer, m->inner, m->vals, x, y);
    } else {
        if (m->is_complex)
            oracle_gemv_colmajor_c128(m->rows, m->cols, m->vals, m->lda, x, y);
        else
            oracle_gemv_colmajor_f64(m->rows, m->cols, m->vals, m->lda, x, y);
    }
}

static int oracle_power_impl(const oracle_mat *m, const double *x0, int max_iterations,
                             double tolerance, double *x_out, double lambda_out[2],
                             int *iterations_out, int *converged_out)
{
    if (m->rows != m->cols) return 1; /* :52-55 "powerMethod: matrix must be square" */
    if (m->rows == 0) return 2;       /* :56-58 "powerMethod: matrix has zero size"  */
    const int64_t n = m->rows;
    const int64_t nd = m->is_complex ? 2 * n : n;
    double *x = (double *)malloc(sizeof(double) * (size_t)nd);
    double *y = (double *)malloc(sizeof(double) * (size_t)nd);
    double *ax = (double *)malloc(sizeof(double) * (size_t)nd);
    memcpy(x, x0, sizeof(double) * (size_t)nd);

    double lambda[2] = {0.0, 0.0}; /* :61 */
    int used = 0, converged = 0, initialized = 0;
    for (int k = 0; k < max_iterations; ++k) { /* :68 */
        oracle_matvec(m, x, y);                 /* :69 */
        const double norm_y = sqrt(oracle_squared_norm(nd, y)); /* :72 */
        if (norm_y == 0.0) {                    /* :73-76 */
            used = k + 1;
            break;
        }
        for (int64_t i = 0; i < nd; ++i) x[i] = y[i] / norm_y; /* :78, true division */
        oracle_matvec(m, x, ax);                /* :81, second matvec */
        double lambda_new[2];
        oracle_dotc(m->is_complex, n, x, ax, lambda_new); /* :81 */
        if (initialized && oracle_is_close_relative(lambda_new, lambda, tolerance)) { /* :83-91 */
            lambda[0] = lambda_new[0];
            lambda[1] = lambda_new[1];
            used = k + 1;
            converged = 1;
            break;
        }
        lambda[0] = lambda_new[0]; /* :93-95 */
        lambda[1] = lambda_new[1];
        initialized = 1;
        used = k + 1;
    }
    memcpy(x_out, x, sizeof(double) * (size_t)nd); /* :98 */
    lambda_out[0] = lambda[0];
    lambda_out[1] = lambda[1];
    *iterations_out = used;
    *converged_out = converged;
    free(x);
    free(y);
    free(ax);
    return 0;
}

ORACLE_API int oracle_power_csc(int is_complex, int64_t rows, int64_t cols, const int32_t *outer,
                                const int32_t *inner, const double *vals, const double *x0,
                                int max_iterations, double tolerance, double *x_out,
                                double lambda_out[2], int *iterations_out, int *converged_out)
{
    oracle_mat m = {0, is_complex, rows, cols, 0, outer, inner, vals};
    return oracle_power_impl(&m, x0, max_iterations, tolerance, x_out, lambda_out, iterations_out,
                             converged_out);
}

ORACLE_API int oracle_power_dense(int is_complex, int64_t rows, int64_t cols, const double *A,
                                  int64_t lda, const double *x0, int max_iterations,
                                  double tolerance, double *x_out, double lambda_out[2],
                                  int *iterations_out, int *converged_out)
{
    oracle_mat m = {1, is_complex, rows, cols, lda, NULL, NULL, A};
    return oracle_power_impl(&m, x0, max_iterations, tolerance, x_out, lambda_out, iterations_out,
                             converged_out);
}

/* Seeds libc the way a caller of the reference would (std::srand) -- lets a ctypes caller seed the
 * same libc instance this library's rand() uses. */
ORACLE_API void oracle_srand(unsigned seed) { srand(seed); }
