// Device-resident Matrix cache of the drop-in headers: repeated solves on one EigSol::Matrix upload
// once, and a device copy can never go stale -- not even through a mutable reference the caller keeps, which is how the
// reference's own tests use Matrix::cast (test/matrix_wrapper_test.cpp:44,82).  Built by eigsol/Makefile against the
// drop-in headers + libeigsol_b200.so + the mini gtest; run on the GPU by tests/test_dropin_cpp.py.
#include <gtest/gtest.h>

#include <cmath>
#include <complex>
#include <cstdlib>

#include "eigsol_b200.h"
#include "src/matrix/matrix.hpp"
#include "src/power_method/power_method.hpp"

using EigSol::Matrix;
using EigSol::powerMethod;
using EigSol::SolverOptions;

namespace {
Eigen::MatrixXd diag3(double a, double b, double c) {
    Eigen::MatrixXd m = Eigen::MatrixXd::Zero(3, 3);
    m(0, 0) = a;
    m(1, 1) = b;
    m(2, 2) = c;
    return m;
}
}  // namespace

TEST(DeviceCache, TwoSolvesOnOneConstMatrixUploadOnce) {
    const Matrix M(diag3(5.0, 2.0, 1.0));
    const int64_t before = eigsol_b200_matrices_created();
    std::srand(3);
    const auto r1 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    std::srand(3);
    const auto r2 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    EXPECT_EQ(eigsol_b200_matrices_created() - before, 1);
    EXPECT_TRUE(r1.converged);
    EXPECT_NEAR(r1.eigenvalue, 5.0, 1e-9);
    EXPECT_EQ(r1.iterations, r2.iterations);
    EXPECT_DOUBLE_EQ(r1.eigenvalue, r2.eigenvalue);
}

TEST(DeviceCache, SparseMatrixIsCachedToo) {
    Eigen::SparseMatrix<std::complex<double>> s(3, 3);
    s.insert(0, 0) = {1.0, 3.0};
    s.insert(1, 1) = {2.0, 4.0};
    s.insert(2, 2) = {5.0, -1.0};
    s.insert(0, 2) = {1.0, 4.0};
    s.makeCompressed();
    const Matrix M(s);
    const int64_t before = eigsol_b200_matrices_created();
    for (int k = 0; k < 3; ++k) {
        const auto r = powerMethod<std::complex<double>>(M, SolverOptions{2000, 1e-12});
        EXPECT_TRUE(r.converged);
        EXPECT_NEAR(r.eigenvalue.real(), 5.0, 1e-8);
        EXPECT_NEAR(r.eigenvalue.imag(), -1.0, 1e-8);
    }
    EXPECT_EQ(eigsol_b200_matrices_created() - before, 1);
}

TEST(DeviceCache, EditThroughAHeldMutableReferenceIsSeenByTheNextSolve) {
    Matrix M(diag3(5.0, 2.0, 1.0));
    const auto r1 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    EXPECT_NEAR(r1.eigenvalue, 5.0, 1e-9);
    auto& A = M.cast<Matrix::Dense<double>>();  // the reference's idiom: the reference outlives the statement
    const auto r2 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    EXPECT_NEAR(r2.eigenvalue, 5.0, 1e-9);
    A(1, 1) = 9.0;  // edit AFTER a solve that followed the cast: a cache refilled by that solve would now be stale
    const int64_t before = eigsol_b200_matrices_created();
    const auto r3 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    EXPECT_NEAR(r3.eigenvalue, 9.0, 1e-8);
    EXPECT_EQ(eigsol_b200_matrices_created() - before, 1);  // uploaded afresh
    A(2, 2) = -20.0;
    const auto r4 = powerMethod<double>(M, SolverOptions{500, 1e-12});
    EXPECT_NEAR(r4.eigenvalue, -20.0, 1e-8);
}

TEST(DeviceCache, PowerMethodImplOnARawEigenMatrixIsOneShot) {
    const Eigen::MatrixXd a = diag3(3.0, 1.0, 0.5);
    const int64_t before = eigsol_b200_matrices_created();
    const auto r1 = EigSol::powerMethodImpl(a, SolverOptions{500, 1e-12});
    const auto r2 = EigSol::powerMethodImpl(a, SolverOptions{500, 1e-12});
    EXPECT_EQ(eigsol_b200_matrices_created() - before, 2);
    EXPECT_NEAR(r1.eigenvalue, 3.0, 1e-9);
    EXPECT_NEAR(r2.eigenvalue, 3.0, 1e-9);
}

TEST(DeviceCache, SinglePrecisionInputsRunPromotedToDouble) {
    Eigen::MatrixXf a = Eigen::MatrixXf::Zero(2, 2);
    a(0, 0) = 2.0f;
    a(1, 1) = 1.0f;
    const Matrix M(a);
    const auto r = powerMethod<float>(M, SolverOptions{1000, 1e-7});
    EXPECT_TRUE(r.converged);
    EXPECT_NEAR(r.eigenvalue, 2.0f, 1e-4f);
    EXPECT_EQ(r.eigenvector.size(), 2);
    Eigen::SparseMatrix<std::complex<float>> s(2, 2);
    s.insert(0, 0) = {0.0f, 3.0f};
    s.insert(1, 1) = {1.0f, 0.0f};
    s.makeCompressed();
    const Matrix Ms(s);
    const auto rc = powerMethod<std::complex<float>>(Ms, SolverOptions{1000, 1e-7});
    EXPECT_TRUE(rc.converged);
    EXPECT_NEAR(rc.eigenvalue.imag(), 3.0f, 1e-4f);
    // long double still has no device type
    Eigen::Matrix<long double, Eigen::Dynamic, Eigen::Dynamic> l(2, 2);
    l << 2.0L, 0.0L, 0.0L, 1.0L;
    const Matrix Ml(l);
    EXPECT_THROW(powerMethod<long double>(Ml), std::runtime_error);
}
