#pragma once
// Drop-in for the reference's src/core/types.hpp:28-34 -- same names, same meaning:
//   ScalarConcept : floating point, or std::complex of floating point
//   Vector<S>     : Eigen column vector
#include <Eigen/Dense>

#include <complex>
#include <concepts>
#include <type_traits>

namespace EigSol {

template <typename T>
struct is_complex_of_floating : std::false_type {};
template <typename R>
struct is_complex_of_floating<std::complex<R>> : std::is_floating_point<R> {};

template <typename S>
concept ScalarConcept = std::floating_point<S> || is_complex_of_floating<S>::value;

template <ScalarConcept S>
using Vector = Eigen::Matrix<S, Eigen::Dynamic, 1>;

}  // namespace EigSol
