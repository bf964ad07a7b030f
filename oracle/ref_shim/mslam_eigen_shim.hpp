// Stand-in for the handful of Eigen 3.4 fixed-size types the reference's hot-path headers name
// (modular_slam/types/basic_types.hpp:14-21).  TEST INFRASTRUCTURE: only oracle/Makefile's `_ref` target uses this tree,
// to compile reference sources UNCHANGED where they lie under /root/reference (Eigen itself is not in the image).
// Layout follows Eigen's rule for fixed-size vectorizable types: a 16-byte object (Vector2d) is 16-byte aligned.
#pragma once
// the standard headers Eigen/Core itself pulls in (the reference's sources rely on them transitively)
#include <algorithm>
#include <array>
#include <cassert>
#include <climits>
#include <cmath>
#include <complex>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <initializer_list>
#include <iosfwd>
#include <limits>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>

namespace Eigen
{
template <typename T, int N>
struct ArrayShim;

template <typename T, int R, int C>
struct alignas((sizeof(T) * R * C) % 16 == 0 ? 16 : alignof(T)) Matrix
{
    T v[R * C];
    Matrix() = default;
    Matrix(const T& a, const T& b) : v{a, b} { static_assert(R * C == 2, "two coefficients"); }
    Matrix(const T& a, const T& b, const T& c) : v{a, b, c} { static_assert(R * C == 3, "three coefficients"); }
    T& x() { return v[0]; }
    T& y() { return v[1]; }
    T& z() { return v[2]; }
    const T& x() const { return v[0]; }
    const T& y() const { return v[1]; }
    const T& z() const { return v[2]; }
    T& operator[](std::size_t i) { return v[i]; }
    const T& operator[](std::size_t i) const { return v[i]; }
    T& operator()(std::size_t i) { return v[i]; }
    const T& operator()(std::size_t i) const { return v[i]; }
    // coefficient-wise static_cast, as MatrixBase::cast<U>() does
    template <typename U>
    Matrix<U, R, C> cast() const
    {
        Matrix<U, R, C> out;
        for(int i = 0; i < R * C; ++i)
            out.v[i] = static_cast<U>(v[i]);
        return out;
    }
    ArrayShim<T, R * C> array() const;
    Matrix(const ArrayShim<T, R * C>& a);
};

template <typename T, int N>
struct ArrayShim
{
    T v[N];
};

template <typename T, int R, int C>
ArrayShim<T, R * C> Matrix<T, R, C>::array() const
{
    ArrayShim<T, R * C> a;
    for(int i = 0; i < R * C; ++i)
        a.v[i] = v[i];
    return a;
}

template <typename T, int R, int C>
Matrix<T, R, C>::Matrix(const ArrayShim<T, R * C>& a)
{
    for(int i = 0; i < R * C; ++i)
        v[i] = a.v[i];
}

// scalar / array: coefficient-wise quotient (Eigen: operator/(Scalar, ArrayBase) -> scalar_inverse/quotient op)
template <typename T, int N>
ArrayShim<T, N> operator/(const T& s, const ArrayShim<T, N>& a)
{
    ArrayShim<T, N> out;
    for(int i = 0; i < N; ++i)
        out.v[i] = s / a.v[i];
    return out;
}

using Vector2d = Matrix<double, 2, 1>;
using Vector2i = Matrix<int, 2, 1>;
using Vector3d = Matrix<double, 3, 1>;
using Matrix3d = Matrix<double, 3, 3>;
using Matrix4d = Matrix<double, 4, 4>;
struct Quaterniond
{
    double x, y, z, w;
};
struct AngleAxisd
{
    double angle;
    Vector3d axis;
};
} // namespace Eigen
