#pragma once
// Drop-in for the reference's src/matrix/matrix.hpp:36-246 -- EigSol::Matrix, the type-erased owner
// of exactly one dense or sparse Eigen matrix.  Same public surface and behaviour:
//   Dense<S>  = Eigen::Matrix<S, Dynamic, Dynamic>   (column-major)          reference :39-40
//   Sparse<S> = Eigen::SparseMatrix<S>               (CSC, int indices)      reference :43-44
//   constructors from any dense expression (:70-76), any Eigen::SparseMatrix (:89-94) and a flat
//   ROW-major std::vector (:109-116, throws "Matrix: size mismatch in vector constructor");
//   isDense() :123, scalar_type() :133, type() :145, cast<T>() :160-183 (std::bad_cast on mismatch);
//   not default-constructible, not copyable, not movable (:47-59).
// This is the input side of the drop-in boundary: power_method.hpp reads the raw CSC / column-major
// arrays of the stored value and hands them to the C ABI (include/eigsol_b200.h).
//
// One addition that is invisible to callers: because a Matrix can never move, it can carry a handle
// to its device-resident copy.  The slot is opaque here (power_method.hpp owns its
// meaning).  The non-const cast<T>() returns a long-lived T& (the reference's own tests keep one:
// test/matrix_wrapper_test.cpp:44,82), through which the host matrix can change at any later time without
// this class noticing -- so the first mutable cast drops the device copy AND switches caching off for this
// Matrix for good: from then on every solve uploads what the host holds at that moment, exactly what the
// reference would compute on.  A Matrix only ever read through the const interface (what powerMethod
// itself uses) uploads once.  The slot is guarded by a mutex: concurrent powerMethod calls on one const
// Matrix take turns instead of racing on it.
#include <Eigen/Dense>
#include <Eigen/Sparse>

#include <cstddef>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <typeinfo>
#include <vector>

#include "../box/box.hpp"
#include "../box/box_typed.hpp"
#include "../core/types.hpp"

namespace EigSol {

class Matrix {
public:
    template <ScalarConcept S>
    using Dense = Eigen::Matrix<S, Eigen::Dynamic, Eigen::Dynamic>;
    template <ScalarConcept S>
    using Sparse = Eigen::SparseMatrix<S>;

    // A Matrix always holds a value and has a stable address for its whole life.
    Matrix() = delete;
    Matrix(const Matrix&) = delete;
    Matrix(Matrix&&) = delete;
    Matrix& operator=(const Matrix&) = delete;
    Matrix& operator=(Matrix&&) = delete;

    // From any dense Eigen expression: evaluated into Dense<Scalar> and stored by value.
    template <typename Derived>
        requires ScalarConcept<typename Derived::Scalar>
    explicit Matrix(const Eigen::MatrixBase<Derived>& expr)
        : dense_(true),
          scalar_(&typeid(typename Derived::Scalar)),
          box_(hold(Dense<typename Derived::Scalar>(expr))) {}

    // From any Eigen::SparseMatrix: storage order and index type are normalised to Sparse<Scalar>.
    template <ScalarConcept S, int Options, typename Index>
    explicit Matrix(const Eigen::SparseMatrix<S, Options, Index>& m)
        : dense_(false), scalar_(&typeid(S)), box_(hold(Sparse<S>(m))) {}

    // From flat row-major data.
    template <ScalarConcept S>
    Matrix(const std::vector<S>& data, std::size_t rows, std::size_t cols)
        : dense_(true), scalar_(&typeid(S)), box_(hold(from_row_major(data, rows, cols))) {}

    bool isDense() const { return dense_; }
    const std::type_info& scalar_type() const { return *scalar_; }
    const std::type_info& type() const { return box_->type(); }

    template <typename T>
    T& cast() {
        require<T>();
        {
            // the caller may modify the matrix through this reference now or at any later time
            std::lock_guard<std::mutex> lock(device_mutex_);
            device_slot_.reset();
            device_cacheable_ = false;
        }
        return static_cast<BoxTyped<T>*>(box_.get())->get();
    }
    template <typename T>
    const T& cast() const {
        require<T>();
        return static_cast<const BoxTyped<T>*>(box_.get())->get();
    }

    // Device-resident copy (used by power_method.hpp only; hold device_mutex() around both).
    std::shared_ptr<void>& device_slot() const { return device_slot_; }
    bool device_cacheable() const { return device_cacheable_; }
    std::mutex& device_mutex() const { return device_mutex_; }

private:
    template <typename T>
    static std::unique_ptr<Box> hold(const T& value) {
        return std::make_unique<BoxTyped<T>>(value);
    }

    template <ScalarConcept S>
    static Dense<S> from_row_major(const std::vector<S>& data, std::size_t rows, std::size_t cols) {
        if (data.size() != rows * cols) throw std::runtime_error("Matrix: size mismatch in vector constructor");
        using RowMajor = Eigen::Matrix<S, Eigen::Dynamic, Eigen::Dynamic, Eigen::RowMajor>;
        if (data.empty()) return Dense<S>(static_cast<Eigen::Index>(rows), static_cast<Eigen::Index>(cols));
        return Dense<S>(Eigen::Map<const RowMajor>(data.data(), static_cast<Eigen::Index>(rows),
                                                   static_cast<Eigen::Index>(cols)));
    }

    template <typename T>
    void require() const {
        if (box_->type() != typeid(T)) throw std::bad_cast{};
    }

    bool dense_;
    const std::type_info* scalar_;
    std::unique_ptr<Box> box_;
    mutable std::shared_ptr<void> device_slot_;
    mutable std::mutex device_mutex_;
    bool device_cacheable_ = true;
};

}  // namespace EigSol
