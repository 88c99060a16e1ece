#pragma once
// Drop-in for the reference's src/power_method/power_method.hpp -- same entry points, same checks in
// the same order, same exceptions and messages, same start vector; the LOOP (reference :61-98) is
// replaced by one call through the C ABI of libeigsol_b200.so (include/eigsol_b200.h) into the
// hand-written sm_100a kernels.  There is no host fallback: if the library cannot reach a B200 the
// call throws std::runtime_error("powerMethod: device error: ...").
//
//   powerMethodImpl<Mat>(A, opts)      reference :47-99    (Mat = an Eigen dense or sparse matrix type)
//   powerMethod<Scalar>(M, opts = {})  reference :135-148  (M = EigSol::Matrix)
//
// What stays on the host, exactly as in the reference:
//   * n != cols  -> "powerMethod: matrix must be square"  (:52-55), then
//     n == 0     -> "powerMethod: matrix has zero size"   (:56-58), both BEFORE any rand() is consumed;
//   * x0 = Vector<Scalar>::Random(n); x0.normalize();     (:62) -- the same Eigen expression, hence the
//     same std::rand() stream and bit-identical start vector;
//   * powerMethod: "powerMethod: scalar type mismatch"    (:137-139), then the dense/sparse dispatch.
// Device scalars are double and std::complex<double>.  float / std::complex<float> inputs (ScalarConcept admits
// them, reference src/core/types.hpp:28-30) draw x0 in their own precision exactly as the reference does (same
// rand() stream), are promoted to double on the host, run the same device loop and are cast back -- the result
// is at least as accurate as the reference's float loop, iteration counts can differ by rounding.  long double has
// no device type: "powerMethod: scalar type not supported on the B200 path" (after the shape checks).
//
// powerMethod(M, ...) keeps M's device-resident copy alive in M (a Matrix cannot move), so repeated
// solves on the same Matrix upload and transpose only once -- unless a mutable reference to the stored matrix
// was ever handed out (matrix.hpp), in which case every solve uploads afresh.
#include <Eigen/Dense>
#include <Eigen/Sparse>

#include <cmath>
#include <complex>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <type_traits>

#include "../core/tolerance.hpp"
#include "../core/types.hpp"
#include "../matrix/matrix.hpp"
#include "../option/solver_option.hpp"
#include "../result/eigen_result.hpp"
#include "eigsol_b200.h"

namespace EigSol {
namespace b200_detail {

template <typename S>
inline constexpr int dtype_code = -1;
template <>
inline constexpr int dtype_code<double> = EIGSOL_B200_F64;
template <>
inline constexpr int dtype_code<std::complex<double>> = EIGSOL_B200_C128;

// the double-precision scalar a single-precision input is promoted to (void: no device path)
template <typename S>
struct promoted {
    using type = void;
};
template <>
struct promoted<float> {
    using type = double;
};
template <>
struct promoted<std::complex<float>> {
    using type = std::complex<double>;
};

template <typename Mat>
inline constexpr bool is_sparse_v = std::is_base_of_v<Eigen::SparseMatrixBase<Mat>, Mat>;

[[noreturn]] inline void raise_last_error() { throw std::runtime_error(eigsol_b200_last_error()); }

inline void check(int status) {
    if (status != EIGSOL_B200_OK) raise_last_error();
}

// The device loop on A.  First solve on a matrix (`slot` empty): upload and solve in ONE library call -- the start
// vector rides on the matrix upload -- and the device copy is kept in `slot`; later solves run on the resident copy.
template <typename Mat>
void run_on(const Mat& A, std::shared_ptr<void>& slot, const typename Mat::Scalar* x0, const SolverOptions& opts,
            typename Mat::Scalar* x, double lambda[2], int* iterations, int* converged) {
    using Scalar = typename Mat::Scalar;
    if (slot) {
        check(eigsol_b200_power_run(static_cast<eigsol_b200_matrix*>(slot.get()), x0, opts.maxIterations, opts.tolerance, x,
                                    lambda, iterations, converged));
        return;
    }
    eigsol_b200_matrix* h = nullptr;
    if constexpr (is_sparse_v<Mat>) {
        using Canon = Eigen::SparseMatrix<Scalar, Eigen::ColMajor, int>;
        Canon tmp;
        const Canon* csc = nullptr;
        if constexpr (std::is_same_v<Mat, Canon>) {
            if (A.isCompressed()) csc = &A;
        }
        if (!csc) {
            tmp = A;
            tmp.makeCompressed();
            csc = &tmp;
        }
        check(eigsol_b200_power_csc_keep(dtype_code<Scalar>, csc->rows(), csc->cols(), csc->outerIndexPtr(),
                                         csc->innerIndexPtr(), csc->valuePtr(), x0, opts.maxIterations, opts.tolerance, x,
                                         lambda, iterations, converged, &h));
    } else {
        using Canon = Eigen::Matrix<Scalar, Eigen::Dynamic, Eigen::Dynamic>;
        if constexpr (std::is_same_v<Mat, Canon>) {
            check(eigsol_b200_power_dense_keep(dtype_code<Scalar>, A.rows(), A.cols(), A.data(), A.outerStride(), x0,
                                               opts.maxIterations, opts.tolerance, x, lambda, iterations, converged, &h));
        } else {
            const Canon tmp(A);
            check(eigsol_b200_power_dense_keep(dtype_code<Scalar>, tmp.rows(), tmp.cols(), tmp.data(), tmp.outerStride(), x0,
                                               opts.maxIterations, opts.tolerance, x, lambda, iterations, converged, &h));
        }
    }
    slot = std::shared_ptr<void>(h, [](void* p) { eigsol_b200_matrix_destroy(static_cast<eigsol_b200_matrix*>(p)); });
}

// Shape checks + x0 exactly as the reference (:52-62), then the device loop on `slot`'s matrix
// (uploaded on demand).
template <typename Mat>
EigenResult<typename Mat::Scalar> solve(const Mat& A, const SolverOptions& opts, std::shared_ptr<void>& slot) {
    using Scalar = typename Mat::Scalar;
    const auto n = A.rows();
    if (n != A.cols()) throw std::runtime_error("powerMethod: matrix must be square");
    if (n == 0) throw std::runtime_error("powerMethod: matrix has zero size");
    if constexpr (dtype_code<Scalar> < 0 && !std::is_void_v<typename promoted<Scalar>::type>) {
        // single precision: x0 exactly as the reference draws it, then everything in double on the device
        using Wide = typename promoted<Scalar>::type;
        Vector<Scalar> x0s = Vector<Scalar>::Random(n);
        x0s.normalize();
        Vector<Wide> x0 = x0s.template cast<Wide>();
        x0.normalize();
        Vector<Wide> x(n);
        double lambda[2] = {0.0, 0.0};
        int iterations = 0, converged = 0;
        if (slot) {
            run_on(Eigen::Matrix<Wide, Eigen::Dynamic, Eigen::Dynamic>(), slot, x0.data(), opts, x.data(), lambda, &iterations,
                   &converged);  // resident already: the matrix argument is not looked at
        } else if constexpr (is_sparse_v<Mat>) {
            const Eigen::SparseMatrix<Wide, Eigen::ColMajor, int> wide = A.template cast<Wide>();
            run_on(wide, slot, x0.data(), opts, x.data(), lambda, &iterations, &converged);
        } else {
            const Eigen::Matrix<Wide, Eigen::Dynamic, Eigen::Dynamic> wide = A.template cast<Wide>();
            run_on(wide, slot, x0.data(), opts, x.data(), lambda, &iterations, &converged);
        }
        Scalar value;
        if constexpr (std::is_same_v<Scalar, float>) value = static_cast<float>(lambda[0]);
        else value = Scalar(static_cast<float>(lambda[0]), static_cast<float>(lambda[1]));
        return EigenResult<Scalar>(value, x.template cast<Scalar>(), iterations, converged != 0);
    } else if constexpr (dtype_code<Scalar> < 0) {
        throw std::runtime_error("powerMethod: scalar type not supported on the B200 path");
    } else {
        Vector<Scalar> x0 = Vector<Scalar>::Random(n);
        x0.normalize();
        Vector<Scalar> x(n);
        double lambda[2] = {0.0, 0.0};
        int iterations = 0, converged = 0;
        run_on(A, slot, x0.data(), opts, x.data(), lambda, &iterations, &converged);
        Scalar value;
        if constexpr (std::is_same_v<Scalar, double>) value = lambda[0];
        else value = Scalar(lambda[0], lambda[1]);
        return EigenResult<Scalar>(value, x, iterations, converged != 0);
    }
}

}  // namespace b200_detail

template <typename Mat>
    requires ScalarConcept<typename Mat::Scalar>
EigenResult<typename Mat::Scalar> powerMethodImpl(const Mat& A, const SolverOptions& opts) {
    std::shared_ptr<void> one_shot;  // device copy lives for this call only
    return b200_detail::solve(A, opts, one_shot);
}

template <ScalarConcept Scalar>
EigenResult<Scalar> powerMethod(const Matrix& M, const SolverOptions& opts = SolverOptions{}) {
    if (M.scalar_type() != typeid(Scalar)) throw std::runtime_error("powerMethod: scalar type mismatch");
    std::lock_guard<std::mutex> lock(M.device_mutex());
    std::shared_ptr<void> one_shot;  // used instead of M's slot once mutable access to M was handed out
    std::shared_ptr<void>& slot = M.device_cacheable() ? M.device_slot() : one_shot;
    if (M.isDense()) return b200_detail::solve(M.cast<Matrix::Dense<Scalar>>(), opts, slot);
    return b200_detail::solve(M.cast<Matrix::Sparse<Scalar>>(), opts, slot);
}

}  // namespace EigSol
