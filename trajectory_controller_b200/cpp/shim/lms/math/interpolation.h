// HEADLESS SHIM: the module includes this header but calls nothing from it on the hot path.
#pragma once
namespace lms { namespace math {
template <typename T> T linearInterpolation(T x0, T y0, T x1, T y1, T x) { return y0 + (y1 - y0) * (x - x0) / (x1 - x0); }
} }
