#pragma once
// Drop-in for the reference's src/option/solver_option.hpp:14-20 (same layout: int, double; 16 bytes).
// The two fields cross the C ABI as the max_iterations / tolerance arguments of eigsol_b200_power_*.
namespace EigSol {

struct SolverOptions {
    int maxIterations = 1000;   // upper bound on loop trips
    double tolerance = 1e-10;   // relative change of the eigenvalue estimate that stops the loop
};

}  // namespace EigSol
