#pragma once
// Drop-in for the reference's src/box/box_typed.hpp:27-79: holds one T by value behind a Box.
#include <memory>
#include <typeinfo>

#include "box.hpp"

namespace EigSol {

template <typename T>
class BoxTyped : public Box {
public:
    BoxTyped() = default;
    explicit BoxTyped(const T& v) : held_(v) {}
    ~BoxTyped() override = default;
    BoxTyped(const BoxTyped&) = delete;
    BoxTyped& operator=(const BoxTyped&) = delete;

    const std::type_info& type() const noexcept override { return typeid(T); }
    std::unique_ptr<Box> clone() const override { return std::make_unique<BoxTyped<T>>(held_); }

    T& get() noexcept { return held_; }
    const T& get() const noexcept { return held_; }
    T& operator*() noexcept { return held_; }
    const T& operator*() const noexcept { return held_; }
    T* operator->() noexcept { return &held_; }
    const T* operator->() const noexcept { return &held_; }

private:
    T held_;
};

}  // namespace EigSol
