#pragma once
// Drop-in for the reference's src/core/tolerance.hpp:28-33.
//   is_close_relative(a, b, tol)  <=>  |a - b| <= tol * (1 + |a|), evaluated in double; |.| is the
//   modulus for complex scalars.  `a` is the reference value (the scale uses |a| only).
// The device applies the identical formula in kernels.cuh (close_relative).
#include <cmath>
#include <complex>

#include "types.hpp"

namespace EigSol {

template <ScalarConcept S>
inline bool is_close_relative(S a, S b, double tol) {
    const double gap = std::abs(a - b);
    const double bound = tol * (1.0 + std::abs(a));
    return gap <= bound;
}

}  // namespace EigSol
