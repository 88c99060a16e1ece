#pragma once
// Drop-in for the reference's src/result/eigen_result.hpp:22-52: what powerMethod returns.
// Filled from the out-parameters of eigsol_b200_power_* (x_out, lambda_out, iterations_out, converged_out).
#include <Eigen/Dense>

#include "../core/types.hpp"

namespace EigSol {

template <ScalarConcept Scalar>
struct EigenResult {
    Scalar eigenvalue;            // Rayleigh-quotient estimate of the dominant eigenvalue
    Vector<Scalar> eigenvector;   // unit 2-norm iterate (no phase normalisation)
    int iterations = 0;           // loop trips performed
    bool converged = false;       // relative-change test passed

    EigenResult() = default;
    EigenResult(const Scalar& lambda, const Vector<Scalar>& vec, int iters, bool conv)
        : eigenvalue(lambda), eigenvector(vec), iterations(iters), converged(conv) {}
};

}  // namespace EigSol
