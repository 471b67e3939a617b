// Arithmetic policies of the batched MPC kernels.
//
// EXACT mode reproduces the floating-point evaluation order of the reference solver (dlib::mpc on dlib::matrix
// expression templates): every product and sum is rounded separately, exactly as g++ emits
// them without -ffp-contract (the reference's CMakeLists.txt sets no -march / contraction flags).  The *_rn
// intrinsics are never contracted into FMA by nvcc, whatever -fmad says.
// FAST mode keeps the same recursion but lets each multiply-add round once (FMA).
#pragma once
#include <cuda_runtime.h>

namespace bmpc {

template <typename T, bool EXACT> struct Ar;

template <> struct Ar<double, true> {
    static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
    static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
    static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
    static __device__ __forceinline__ double madd(double acc, double a, double b) { return __dadd_rn(acc, __dmul_rn(a, b)); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
};
template <> struct Ar<double, false> {
    static __device__ __forceinline__ double mul(double a, double b) { return a * b; }
    static __device__ __forceinline__ double add(double a, double b) { return a + b; }
    static __device__ __forceinline__ double sub(double a, double b) { return a - b; }
    static __device__ __forceinline__ double madd(double acc, double a, double b) { return fma(a, b, acc); }
    static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
    static __device__ __forceinline__ double sqrt(double a) { return __dsqrt_rn(a); }
};
template <> struct Ar<float, true> {
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float madd(float acc, float a, float b) { return __fadd_rn(acc, __fmul_rn(a, b)); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
};
template <> struct Ar<float, false> {
    static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
    static __device__ __forceinline__ float add(float a, float b) { return a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
    static __device__ __forceinline__ float madd(float acc, float a, float b) { return fmaf(a, b, acc); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
};

// dlib::clamp(m, lo, hi) element function, matrix_utilities.h:2829-2846 (a NaN maps to hi).
template <typename T> __device__ __forceinline__ T clamp_m(T val, T lo, T hi) { return val <= hi ? (lo <= val ? val : lo) : hi; }
// dlib::put_in_range(a, b, val), algs.h:716-748 (a NaN passes through; reversed bounds handled like the reference).
template <typename T> __device__ __forceinline__ T put_in_range(T a, T b, T val)
{
    if (a < b) { if (val < a) return a; else if (val > b) return b; }
    else       { if (val < b) return b; else if (val > a) return a; }
    return val;
}
template <typename T> __device__ __forceinline__ T absval(T x);
template <> __device__ __forceinline__ double absval<double>(double x) { return fabs(x); }
template <> __device__ __forceinline__ float absval<float>(float x) { return fabsf(x); }

}  // namespace bmpc
