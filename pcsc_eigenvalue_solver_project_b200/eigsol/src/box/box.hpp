#pragma once
// Drop-in for the reference's src/box/box.hpp:32-81: the abstract base of the type-erasure used by
// EigSol::Matrix.  Non-copyable; reports the dynamic type of the held value; can clone itself.
#include <memory>
#include <typeinfo>

namespace EigSol {

class Box {
public:
    Box() = default;
    virtual ~Box() = default;
    Box(const Box&) = delete;
    Box& operator=(const Box&) = delete;

    virtual const std::type_info& type() const noexcept = 0;
    virtual std::unique_ptr<Box> clone() const = 0;
};

}  // namespace EigSol
