// HEADLESS SHIM of lms::math::vertex2f (un-vendored `lms_math`).  Float semantics adopted
// by this project (PARITY UNPINNED — the real library is absent): every operation in float, each rounded on its own.
#pragma once
#include <cmath>
namespace lms {
namespace math {
template <typename T>
struct vertex2 {
    T x, y;
    vertex2() : x(0), y(0) {}
    vertex2(T x_, T y_) : x(x_), y(y_) {}
    vertex2 operator+(const vertex2& o) const { return vertex2(x + o.x, y + o.y); }
    vertex2 operator-(const vertex2& o) const { return vertex2(x - o.x, y - o.y); }
    vertex2 operator*(T s) const { return vertex2(x * s, y * s); }
    vertex2 operator/(T s) const { return vertex2(x / s, y / s); }
    T lengthSquared() const { return x * x + y * y; }
    T length() const { return std::sqrt(x * x + y * y); }
    T distance(const vertex2& o) const { return (*this - o).length(); }
    vertex2 normalize() const { return *this / length(); }
    T angle() const { return std::atan2(y, x); }
};
typedef vertex2<float> vertex2f;
template <typename T> int sgn(T v) { return (T(0) < v) - (v < T(0)); }
}  // namespace math
}  // namespace lms
