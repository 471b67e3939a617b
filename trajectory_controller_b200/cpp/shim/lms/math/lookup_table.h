// HEADLESS SHIM of lms::math::LookupTable (un-vendored `lms_math`): public sample vectors vx
// (ascending) / vy and linearSearch(x) = piecewise-linear interpolation clamped at both ends.  PARITY UNPINNED.
#pragma once
#include <cstddef>
#include <vector>
namespace lms {
namespace math {
enum class LookupTableOrder { ASC, DESC };
template <typename T, LookupTableOrder ORDER>
class LookupTable {
public:
    std::vector<T> vx, vy;
    T linearSearch(T x) const
    {
        const size_t n = vx.size() < vy.size() ? vx.size() : vy.size();
        if (n == 0) return x;                       // an unconfigured table is the identity
        if (x <= vx[0]) return vy[0];
        for (size_t i = 1; i < n; ++i)
            if (x <= vx[i]) return vy[i - 1] + (vy[i] - vy[i - 1]) * (x - vx[i - 1]) / (vx[i] - vx[i - 1]);
        return vy[n - 1];
    }
};
}  // namespace math
}  // namespace lms
