// ref_shim.cpp -- C-ABI doorway onto the UNMODIFIED reference, compiled from the sources where they
// lie under /root/reference (see oracle/Makefile; output goes to oracle/_ref/libeigsol_ref.so, which
// is git-ignored).  TEST INFRASTRUCTURE ONLY: used to pin oracle/eigsol_oracle.c, to generate
// tests/golden/, and as the timed CPU baseline (bench.py --impl reference, cpu_baseline.kind =
// "reference").  Nothing here is on the product path.
//
// The file contains no reference code: it includes the reference's public headers
// (src/power_method/power_method.hpp:135-148 `EigSol::powerMethod`, src/matrix/matrix.hpp:36-246
// `EigSol::Matrix`, src/option/solver_option.hpp:14-20, src/core/tolerance.hpp:28-33) and calls them
// through their public API on caller-supplied raw arrays.
#include <chrono>
#include <complex>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>

#include <Eigen/Dense>
#include <Eigen/Sparse>

#include "src/core/tolerance.hpp"
#include "src/matrix/matrix.hpp"
#include "src/option/solver_option.hpp"
#include "src/power_method/power_method.hpp"
#include "src/reader/file_matrix_reader.hpp"  // EigSol::readMatrixFromFile (:170-200), included once: it has no include guard

#define REF_API extern "C" __attribute__((visibility("default")))

namespace {
thread_local std::string g_err;
using cd = std::complex<double>;

template <typename S>
std::unique_ptr<EigSol::Matrix> wrap_csc(int64_t rows, int64_t cols, const int32_t* outer,
                                         const int32_t* inner, const void* vals) {
    using SpMat = Eigen::SparseMatrix<S>;
    Eigen::Map<const SpMat> view(rows, cols, outer[cols], outer, inner, static_cast<const S*>(vals));
    SpMat owned(view);
    return std::make_unique<EigSol::Matrix>(owned);
}

template <typename S>
std::unique_ptr<EigSol::Matrix> wrap_dense(int64_t rows, int64_t cols, const void* a, int64_t lda) {
    using Mat = Eigen::Matrix<S, Eigen::Dynamic, Eigen::Dynamic>;
    Eigen::Map<const Mat, 0, Eigen::OuterStride<>> view(static_cast<const S*>(a), rows, cols,
                                                        Eigen::OuterStride<>(lda));
    Mat owned(view);
    return std::make_unique<EigSol::Matrix>(owned);
}

template <typename S>
int run_power(const EigSol::Matrix& m, int seed, int max_it, double tol, void* x_out, double* lam,
              int* iters, int* conv, double* seconds) {
    EigSol::SolverOptions opts;
    opts.maxIterations = max_it;
    opts.tolerance = tol;
    if (seed >= 0) std::srand(static_cast<unsigned>(seed));
    const auto t0 = std::chrono::steady_clock::now();
    EigSol::EigenResult<S> r = EigSol::powerMethod<S>(m, opts);
    const auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (x_out) std::memcpy(x_out, r.eigenvector.data(), sizeof(S) * size_t(r.eigenvector.size()));
    const cd l(r.eigenvalue);
    lam[0] = l.real();
    lam[1] = l.imag();
    *iters = r.iterations;
    *conv = r.converged ? 1 : 0;
    return 0;
}

struct RefHandle {
    std::unique_ptr<EigSol::Matrix> m;
    int dtype;
};
}  // namespace

REF_API const char* ref_last_error() { return g_err.c_str(); }

// dtype: 0 = double, 1 = std::complex<double> (interleaved re, im).
REF_API void* ref_matrix_create_csc(int dtype, int64_t rows, int64_t cols, const int32_t* outer,
                                    const int32_t* inner, const void* vals) {
    try {
        auto h = new RefHandle;
        h->dtype = dtype;
        h->m = dtype ? wrap_csc<cd>(rows, cols, outer, inner, vals)
                     : wrap_csc<double>(rows, cols, outer, inner, vals);
        return h;
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}

REF_API void* ref_matrix_create_dense(int dtype, int64_t rows, int64_t cols, const void* a,
                                      int64_t lda) {
    try {
        auto h = new RefHandle;
        h->dtype = dtype;
        h->m = dtype ? wrap_dense<cd>(rows, cols, a, lda) : wrap_dense<double>(rows, cols, a, lda);
        return h;
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}

REF_API void ref_matrix_destroy(void* h) { delete static_cast<RefHandle*>(h); }

// Runs EigSol::powerMethod<Scalar>(M, {max_it, tol}).  seed >= 0: std::srand(seed) first (x0 is drawn
// inside the call, power_method.hpp:62).  `scalar_dtype` is the template argument, so a mismatch
// with the stored type exercises the "scalar type mismatch" throw (:137-139).
// Returns 0, or 1 when the reference threw std::runtime_error, 2 for std::bad_cast, 3 otherwise.
REF_API int ref_power_run(void* hv, int scalar_dtype, int seed, int max_it, double tol, void* x_out,
                          double lambda_out[2], int* iters, int* conv, double* seconds) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        return scalar_dtype ? run_power<cd>(*h->m, seed, max_it, tol, x_out, lambda_out, iters, conv, seconds)
                            : run_power<double>(*h->m, seed, max_it, tol, x_out, lambda_out, iters, conv, seconds);
    } catch (const std::runtime_error& e) {
        g_err = e.what();
        return 1;
    } catch (const std::bad_cast& e) {
        g_err = e.what();
        return 2;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 3;
    }
}

// The start vector the reference would draw after std::srand(seed): Vector<S>::Random(n), normalize()
// (power_method.hpp:62).
REF_API void ref_random_unit(int dtype, int64_t n, int seed, void* out) {
    if (seed >= 0) std::srand(static_cast<unsigned>(seed));
    if (dtype) {
        EigSol::Vector<cd> x = EigSol::Vector<cd>::Random(n);
        x.normalize();
        std::memcpy(out, x.data(), sizeof(cd) * size_t(n));
    } else {
        EigSol::Vector<double> x = EigSol::Vector<double>::Random(n);
        x.normalize();
        std::memcpy(out, x.data(), sizeof(double) * size_t(n));
    }
}

// One product y = A*x through Eigen, exactly the expression at power_method.hpp:69.
REF_API int ref_matvec(void* hv, const void* x, void* y) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        auto go = [&](auto tag) {
            using S = decltype(tag);
            if (h->m->isDense()) {
                const auto& a = h->m->cast<EigSol::Matrix::Dense<S>>();
                Eigen::Map<const EigSol::Vector<S>> xv(static_cast<const S*>(x), a.cols());
                EigSol::Vector<S> yv = a * xv;
                std::memcpy(y, yv.data(), sizeof(S) * size_t(yv.size()));
            } else {
                const auto& a = h->m->cast<EigSol::Matrix::Sparse<S>>();
                Eigen::Map<const EigSol::Vector<S>> xv(static_cast<const S*>(x), a.cols());
                EigSol::Vector<S> yv = a * xv;
                std::memcpy(y, yv.data(), sizeof(S) * size_t(yv.size()));
            }
        };
        if (h->dtype) go(cd{}); else go(double{});
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

REF_API int ref_is_close_relative(int dtype, const double a[2], const double b[2], double tol) {
    if (dtype) return EigSol::is_close_relative(cd(a[0], a[1]), cd(b[0], b[1]), tol) ? 1 : 0;
    return EigSol::is_close_relative(a[0], b[0], tol) ? 1 : 0;
}

REF_API void ref_srand(unsigned seed) { std::srand(seed); }

// ---- the reference's file reader (src/reader/file_matrix_reader.hpp:170-200), for pinning the fast reader ----
// Reads `path` as Scalar = double (dtype 0) or complex<double> (1) and copies out what the returned Matrix owns.
// Call once with null arrays to get the sizes.  Returns 0, or 1 with the reference's exception text in
// ref_last_error().  seconds_out = time spent inside readMatrixFromFile.
REF_API int ref_read_matrix_file(const char* path, int dtype, int* is_dense, int64_t* rows, int64_t* cols, int64_t* nnz,
                                 int32_t* outer, int32_t* inner, void* values, double* seconds_out) {
    try {
        const auto t0 = std::chrono::steady_clock::now();
        auto copy_out = [&](auto tag) {
            using S = decltype(tag);
            // Matrix is neither copyable nor movable: read in place and copy the arrays out
            const EigSol::Matrix mat = EigSol::readMatrixFromFile<S>(std::string(path));
            const auto t1 = std::chrono::steady_clock::now();
            if (seconds_out) *seconds_out = std::chrono::duration<double>(t1 - t0).count();
            *is_dense = mat.isDense() ? 1 : 0;
            if (mat.isDense()) {
                const auto& d = mat.cast<EigSol::Matrix::Dense<S>>();
                *rows = d.rows();
                *cols = d.cols();
                *nnz = int64_t(d.rows()) * d.cols();
                if (values) std::memcpy(values, d.data(), sizeof(S) * size_t(*nnz));
            } else {
                const auto& sp = mat.cast<EigSol::Matrix::Sparse<S>>();
                *rows = sp.rows();
                *cols = sp.cols();
                *nnz = sp.nonZeros();
                if (outer) std::memcpy(outer, sp.outerIndexPtr(), sizeof(int32_t) * size_t(sp.cols() + 1));
                if (inner) std::memcpy(inner, sp.innerIndexPtr(), sizeof(int32_t) * size_t(*nnz));
                if (values) std::memcpy(values, sp.valuePtr(), sizeof(S) * size_t(*nnz));
            }
        };
        if (dtype) copy_out(cd{});
        else copy_out(double{});
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}
