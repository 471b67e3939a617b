// Structural patterns of the models this build specialises for (see solve_kernel.cuh "STRUCTURAL SPARSITY").
// Bit e = row-major entry e.  A model qualifies for a pattern iff every non-zero entry has its NZ bit set and every
// entry with a ONE bit equals exactly 1.0 (checked on the host for every instance, bmpc_core.cu).
#pragma once
#include "solve_kernel.cuh"

namespace bmpc {
namespace pat {
template <int S, int I> struct dense { using type = DensePattern<S, I>; };

// Synthetic lateral bicycle model, SURVEY.md §8(d): A = [1 a 0 0; 0 1 a 0; 0 0 1 a; 0 0 a a], B = (0 0 0 b)',
// C = 0, Q = (q q 0 0).
template <int S, int I> struct bicycle4 {
    static_assert(S == 4 && I == 1, "bicycle4 is a 4-state / 1-input pattern");
    using type = Pattern<0xCC63u, 0x0421u, 0x8u, 0x0u, 0x0u, 0x3u>;
};
// The reference module's model, src/trajectory_point_follower.cpp:326-333: A = [1 a; 0 1], B = [0 b; b b], C = 0.
template <int S, int I> struct tobi {
    static_assert(S == 2 && I == 2, "tobi is a 2-state / 2-input pattern");
    using type = Pattern<0xBu, 0x9u, 0xEu, 0x0u, 0x0u, 0x3u>;
};
// dlib's known-answer test, dlib_files/dlib/test/mpc.cpp:269-281: A = [1 1; 0 1], B = (0 1)', C dense, Q = (q 0).
template <int S, int I> struct kat {
    static_assert(S == 2 && I == 1, "kat is a 2-state / 1-input pattern");
    using type = Pattern<0xBu, 0xBu, 0x2u, 0x2u, 0x3u, 0x1u>;
};
}  // namespace pat
}  // namespace bmpc
