// N-dimensional strided view over a flat buffer with a random-access iterator along any axis, so that
// <algorithm> (std::nth_element, std::sort, ...) can run IN PLACE along an axis that is not contiguous.
// Role of the reference's include/array.hpp (`Array<Array<T>>` + `Iterator`, used by the CPU Bulyan's
// coordinate-wise phase, native/py_bulyan/bulyan.cpp:167-171). Own design: a fixed-rank
// (≤ 4) descriptor of shape + element strides; `at(axis, i)` drops an axis, `line(axis)` is the 1-D range
// through the view's origin along `axis`.
#pragma once

#include <cstddef>
#include <initializer_list>
#include <iterator>
#include <stdexcept>

namespace gb {

template <class T>
class StrideIterator {
public:
    using iterator_category = std::random_access_iterator_tag;
    using value_type = T;
    using difference_type = std::ptrdiff_t;
    using pointer = T*;
    using reference = T&;

    StrideIterator() = default;
    StrideIterator(T* at, std::ptrdiff_t stride) : at_(at), stride_(stride) {}

    reference operator*() const { return *at_; }
    pointer operator->() const { return at_; }
    reference operator[](difference_type n) const { return at_[n * stride_]; }
    StrideIterator& operator++() { at_ += stride_; return *this; }
    StrideIterator operator++(int) { StrideIterator old = *this; at_ += stride_; return old; }
    StrideIterator& operator--() { at_ -= stride_; return *this; }
    StrideIterator operator--(int) { StrideIterator old = *this; at_ -= stride_; return old; }
    StrideIterator& operator+=(difference_type n) { at_ += n * stride_; return *this; }
    StrideIterator& operator-=(difference_type n) { at_ -= n * stride_; return *this; }
    friend StrideIterator operator+(StrideIterator it, difference_type n) { return it += n; }
    friend StrideIterator operator+(difference_type n, StrideIterator it) { return it += n; }
    friend StrideIterator operator-(StrideIterator it, difference_type n) { return it -= n; }
    friend difference_type operator-(StrideIterator const& a, StrideIterator const& b) { return (a.at_ - b.at_) / a.stride_; }
    friend bool operator==(StrideIterator const& a, StrideIterator const& b) { return a.at_ == b.at_; }
    friend bool operator!=(StrideIterator const& a, StrideIterator const& b) { return a.at_ != b.at_; }
    friend bool operator<(StrideIterator const& a, StrideIterator const& b) { return (b - a) > 0; }
    friend bool operator>(StrideIterator const& a, StrideIterator const& b) { return (a - b) > 0; }
    friend bool operator<=(StrideIterator const& a, StrideIterator const& b) { return !(a > b); }
    friend bool operator>=(StrideIterator const& a, StrideIterator const& b) { return !(a < b); }

private:
    T* at_ = nullptr;
    std::ptrdiff_t stride_ = 1;
};

template <class T>
struct StrideRange {
    StrideIterator<T> first, last;
    StrideIterator<T> begin() const { return first; }
    StrideIterator<T> end() const { return last; }
    std::size_t size() const { return static_cast<std::size_t>(last - first); }
};

template <class T>
class StridedView {
public:
    static constexpr int kMaxRank = 4;

    // Row-major view of `data` with the given shape.
    StridedView(T* data, std::initializer_list<std::size_t> shape) : base_(data), rank_(static_cast<int>(shape.size())) {
        if (rank_ < 1 || rank_ > kMaxRank) throw std::invalid_argument("StridedView: rank must be 1..4");
        int k = 0;
        for (std::size_t s : shape) shape_[k++] = s;
        std::ptrdiff_t run = 1;
        for (int a = rank_ - 1; a >= 0; --a) { stride_[a] = run; run *= static_cast<std::ptrdiff_t>(shape_[a]); }
    }
    StridedView(T* data, int rank, std::size_t const* shape) : base_(data), rank_(rank) {
        if (rank_ < 1 || rank_ > kMaxRank) throw std::invalid_argument("StridedView: rank must be 1..4");
        std::ptrdiff_t run = 1;
        for (int a = rank_ - 1; a >= 0; --a) { shape_[a] = shape[a]; stride_[a] = run; run *= static_cast<std::ptrdiff_t>(shape_[a]); }
    }

    int rank() const { return rank_; }
    std::size_t extent(int axis) const { return shape_[axis]; }
    std::ptrdiff_t stride(int axis) const { return stride_[axis]; }
    T* data() const { return base_; }

    // Fix `axis` at `index`: a view of rank − 1 (rank-1 views collapse to a single-element rank-1 view).
    StridedView at(int axis, std::size_t index) const {
        StridedView out = *this;
        out.base_ = base_ + static_cast<std::ptrdiff_t>(index) * stride_[axis];
        if (rank_ == 1) { out.shape_[0] = 1; return out; }
        for (int a = axis; a + 1 < rank_; ++a) { out.shape_[a] = shape_[a + 1]; out.stride_[a] = stride_[a + 1]; }
        out.rank_ = rank_ - 1;
        return out;
    }
    StridedView swapped(int a, int b) const {
        StridedView out = *this;
        std::swap(out.shape_[a], out.shape_[b]);
        std::swap(out.stride_[a], out.stride_[b]);
        return out;
    }
    // The elements (0, …, i, …, 0), i = 0 … extent(axis) − 1: usable with <algorithm>.
    StrideRange<T> line(int axis = 0) const {
        StrideIterator<T> first(base_, stride_[axis]);
        return {first, first + static_cast<std::ptrdiff_t>(shape_[axis])};
    }
    T& operator()(std::size_t i) const { return base_[static_cast<std::ptrdiff_t>(i) * stride_[0]]; }
    T& operator()(std::size_t i, std::size_t j) const { return base_[static_cast<std::ptrdiff_t>(i) * stride_[0] + static_cast<std::ptrdiff_t>(j) * stride_[1]]; }

private:
    T* base_;
    int rank_;
    std::size_t shape_[kMaxRank] = {1, 1, 1, 1};
    std::ptrdiff_t stride_[kMaxRank] = {1, 1, 1, 1};
};

}  // namespace gb
