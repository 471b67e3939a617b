// Constructor math of dlib::mpc (mpc.h:114-123) for a batch of models, one thread per model:
//   lambda = sum(R)*N + sum_c trace(B'T_cB);  Q_diag[N-1-c] = diag(B'T_cB);  T_{c+1} = A'T_cA + diagm(Q),  T_0 = diagm(Q)
// Evaluation order as dlib::matrix instantiates it: (trans(B)*T)*B and (trans(A)*T)*A are two dense products each
// (matrix.h:43-60), "+ diagm(Q)" adds the diagonal entry (a zero elsewhere), sum()/trace() accumulate from 0
// (matrix_utilities.h:1420-1437, matrix_la.h:1662-1680).  Models are SoA [elem][stride] on the device.
#pragma once
#include "arith.cuh"

namespace bmpc {

template <typename T, int S, int I, int N, bool EXACT>
__global__ void ctor_kernel(long long count, long long stride, const T* __restrict__ mA, const T* __restrict__ mB,
                            const T* __restrict__ mQ, const T* __restrict__ mR, T* __restrict__ mlambda,
                            T* __restrict__ mqdiag)
{
    using A_ = Ar<T, EXACT>;
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    T A[S * S], B[S * I], Q[S], Tm[S * S];
#pragma unroll
    for (int e = 0; e < S * S; ++e) A[e] = mA[(long long)e * stride + k];
#pragma unroll
    for (int e = 0; e < S * I; ++e) B[e] = mB[(long long)e * stride + k];
#pragma unroll
    for (int e = 0; e < S; ++e) Q[e] = mQ[(long long)e * stride + k];
    T sumR = T(0);
#pragma unroll
    for (int j = 0; j < I; ++j) sumR = A_::add(sumR, mR[(long long)j * stride + k]);
    T lambda = A_::mul(sumR, T(N));                                     // mpc.h:116
#pragma unroll
    for (int r = 0; r < S; ++r)
#pragma unroll
        for (int c = 0; c < S; ++c) Tm[r * S + c] = (r == c) ? Q[r] : T(0);   // mpc.h:117
#pragma unroll 1
    for (int c = 0; c < N; ++c) {                                       // mpc.h:118-123
        T BtT[I * S];
#pragma unroll
        for (int i = 0; i < I; ++i)
#pragma unroll
            for (int q = 0; q < S; ++q) {
                T t = A_::mul(B[i], Tm[q]);
#pragma unroll
                for (int l = 1; l < S; ++l) t = A_::madd(t, B[l * I + i], Tm[l * S + q]);
                BtT[i * S + q] = t;
            }
        T tr = T(0);
#pragma unroll
        for (int j = 0; j < I; ++j) {
            T t = A_::mul(BtT[j * S], B[j]);
#pragma unroll
            for (int q = 1; q < S; ++q) t = A_::madd(t, BtT[j * S + q], B[q * I + j]);
            tr = A_::add(tr, t);
            mqdiag[(long long)((N - 1 - c) * I + j) * stride + k] = t;  // mpc.h:121
        }
        lambda = A_::add(lambda, tr);                                   // mpc.h:120
        T AtT[S * S], Tn[S * S];
#pragma unroll
        for (int r = 0; r < S; ++r)
#pragma unroll
            for (int q = 0; q < S; ++q) {
                T t = A_::mul(A[r], Tm[q]);
#pragma unroll
                for (int l = 1; l < S; ++l) t = A_::madd(t, A[l * S + r], Tm[l * S + q]);
                AtT[r * S + q] = t;
            }
#pragma unroll
        for (int r = 0; r < S; ++r)
#pragma unroll
            for (int q = 0; q < S; ++q) {
                T t = A_::mul(AtT[r * S], A[q]);
#pragma unroll
                for (int l = 1; l < S; ++l) t = A_::madd(t, AtT[r * S + l], A[l * S + q]);
                Tn[r * S + q] = A_::add(t, (r == q) ? Q[r] : T(0));     // mpc.h:122
            }
#pragma unroll
        for (int e = 0; e < S * S; ++e) Tm[e] = Tn[e];
    }
    mlambda[k] = lambda;
}

}  // namespace bmpc
