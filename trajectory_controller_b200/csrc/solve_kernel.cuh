// Batched box-constrained linear-MPC solver for sm_100a: one MPC instance per LANE, persistent warps that
// pull instances from a device-wide queue, everything an iteration touches held in registers and shared memory.
//
// What it computes is dlib::mpc<S,I,N>::operator() — /root/reference/dlib_files/dlib/control/mpc.h:216-240 and
// solve_linear_mpc, mpc.h:253-347 — for `count` independent controllers; SURVEY.md Appendix A is the arithmetic
// contract (dot products accumulate k = 0..n-1 with the first term initialising, sums associate left, no FMA
// in EXACT mode).
//
// Why this shape:
//  * the gradient (mpc.h:275-283) is two length-N dependency chains of SxS mat-vecs; its rounding order forbids
//    splitting one instance over lanes, but S rows x S products give each lane enough ILP to keep the FP64 pipe
//    (16 lanes/SMSP/clk on B200) busy from ONE warp per SM sub-partition.  So occupancy is spent on storage, not
//    on latency hiding: u[] and the linear term MM[] live in registers (fully unrolled horizon), the stashed
//    Q*M[] rows, the gradient and the Nesterov v[] in shared memory as [element][lane] (conflict-free 8-byte
//    accesses, no HBM traffic inside the iteration loop).
//  * iteration counts are data dependent (p99/p50 ~ 2-14x).  A lane therefore never waits for its warp: it is a
//    small state machine (SETUP -> ITERATE ... -> finished -> fetch next) and every trip of the warp-wide loop
//    runs exactly one gradient-shaped pass for all 32 lanes.  The per-solve linear term (mpc.h:258-266) has the
//    same recursion structure as the gradient, so SETUP lanes ride the same pass with predicated operands
//    (MIXED pass); there is no divergent code of any size.
#pragma once
#include <stdint.h>

#include "arith.cuh"

namespace bmpc {

constexpr int kSmemMax = 232448;  // 227 KB opt-in dynamic shared memory per CTA on sm_100

template <typename T, int S, int I, int N>
struct SolveParams {
    // shared model (unused by the per-instance-model kernels); row-major like dlib::matrix
    T A[S * S], B[S * I], C[S], Q[S], R[I], lo[I], hi[I];
    T lambda;            // mpc.h:116-120
    T qdiag[N * I];      // mpc.h:121 (B'TB diagonal WITHOUT R)
    T eps;               // mpc.h:104
    uint32_t max_iter;   // mpc.h:103
    int warm;            // BMPC_WARM
    int keep_controls;   // write controls[] back (always when warm)
    int target_full;     // 0: per-instance constant target [S][stride]; 1: ring [N][S][stride]
    int head;            // ring: physical slot of horizon step 0
    long long count, stride;
    const T* x0;         // [S][stride]
    T* controls;         // [N*I][stride]
    T* vstate;           // [N*I][stride]  (dlib's member v, mpc.h:249 — see "stale v" below)
    T* target;
    T* u0;               // [I][stride]
    uint32_t* iters;     // [stride]
    uint8_t* status;     // [stride]
    // per-instance models, SoA [elem][stride]
    const T *mA, *mB, *mC, *mQ, *mR, *mlo, *mhi, *mlambda, *mqdiag;
    unsigned long long* queue;   // next unclaimed instance
};

constexpr int pow2_floor(int x) { return x >= 16 ? 16 : x >= 8 ? 8 : x >= 4 ? 4 : x >= 2 ? 2 : x >= 1 ? 1 : 0; }

template <typename T, int S, int I, int N, bool PIM>
struct Layout {
    static constexpr int NI = N * I;
    // u[] in registers (fully unrolled horizon) when it costs at most 64 32-bit registers
    static constexpr bool UREG = (NI * int(sizeof(T) / 4) <= 64);
    static constexpr int O_QM = 0;                       // Q*M rows of horizon steps 0..N-2
    static constexpr int O_DF = O_QM + (N - 1) * S;      // gradient
    static constexpr int O_V = O_DF + NI;                // Nesterov v
    static constexpr int O_U = O_V + NI;
    static constexpr int O_MM = O_U + (UREG ? 0 : NI);   // linear term MM
    static constexpr int O_MODEL = O_MM + NI;
    static constexpr int M_A = 0, M_B = M_A + S * S, M_C = M_B + S * I, M_Q = M_C + S, M_R = M_Q + S,
                         M_LO = M_R + I, M_HI = M_LO + I, M_LEN = M_HI + I;
    static constexpr int ELEMS = O_MODEL + (PIM ? M_LEN : 0);
    static constexpr int WARP_BYTES = ELEMS * 32 * int(sizeof(T));
    static constexpr int FIT = kSmemMax / WARP_BYTES;
    static constexpr int WARPS = pow2_floor(FIT < (UREG ? 8 : 16) ? FIT : (UREG ? 8 : 16));
    static constexpr int THREADS = WARPS * 32;
    static constexpr int SMEM_BYTES = WARPS * WARP_BYTES;
    static_assert(WARPS >= 1, "horizon too long for the shared-memory resident layout");
};

#define BMPC_LOOP_N _Pragma("unroll")

template <typename T, int S, int I, int N, bool EXACT, bool PIM>
struct Solver {
    using L = Layout<T, S, I, N, PIM>;
    using P = SolveParams<T, S, I, N>;
    using A_ = Ar<T, EXACT>;
    static constexpr int NI = L::NI;
    static constexpr bool UREG = L::UREG;
    static constexpr int UNROLL = UREG ? N : 2;        // horizon loops: full unroll keeps u[]/mm[] in registers
    static constexpr int UNROLL_K = UREG ? NI : 2;     // loops over the N*I control entries

    const P& p;
    T* sm;            // this lane's column of the warp's shared-memory block: element e at sm[e*32]
    T u[UREG ? NI : 1];

    __device__ __forceinline__ Solver(const P& p_, T* sm_) : p(p_), sm(sm_) {}

    // ---- model access: kernel-parameter constant bank (shared model) or this lane's shared-memory copy ----------
    __device__ __forceinline__ T mA(int r, int k) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_A + r * S + k) * 32]; else return p.A[r * S + k]; }
    __device__ __forceinline__ T mB(int r, int j) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_B + r * I + j) * 32]; else return p.B[r * I + j]; }
    __device__ __forceinline__ T mC(int r) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_C + r) * 32]; else return p.C[r]; }
    __device__ __forceinline__ T mQ(int r) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_Q + r) * 32]; else return p.Q[r]; }
    __device__ __forceinline__ T mR(int j) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_R + j) * 32]; else return p.R[j]; }
    __device__ __forceinline__ T mLo(int j) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_LO + j) * 32]; else return p.lo[j]; }
    __device__ __forceinline__ T mHi(int j) const { if constexpr (PIM) return sm[(L::O_MODEL + L::M_HI + j) * 32]; else return p.hi[j]; }

    __device__ __forceinline__ T getU(int k) const { if constexpr (UREG) return u[k]; else return sm[(L::O_U + k) * 32]; }
    __device__ __forceinline__ void setU(int k, T v) { if constexpr (UREG) u[k] = v; else sm[(L::O_U + k) * 32] = v; }
    __device__ __forceinline__ T getMM(int k) const { return sm[(L::O_MM + k) * 32]; }
    __device__ __forceinline__ void setMM(int k, T v) { sm[(L::O_MM + k) * 32] = v; }

    // The Q*M stash is written in the forward sweep and read back in the backward sweep of the SAME unrolled pass.
    // Through plain pointers nvcc forwards every stored value to its load, i.e. keeps N*S live values in registers
    // and spills them to local memory; volatile accesses make the round trip through shared memory real.
    __device__ __forceinline__ void stash_st(int off, T v) { *reinterpret_cast<volatile T*>(sm + off) = v; }
    __device__ __forceinline__ T stash_ld(int off) const { return *reinterpret_cast<const volatile T*>(sm + off); }

    // matrix.h:43-60 dot products, first term initialises
    __device__ __forceinline__ T dotA(int r, const T* z) const
    {
        T t = A_::mul(mA(r, 0), z[0]);
#pragma unroll
        for (int k = 1; k < S; ++k) t = A_::madd(t, mA(r, k), z[k]);
        return t;
    }
    __device__ __forceinline__ T dotAT(int r, const T* z) const   // trans(A)*z, matrix_utilities.h:375-410
    {
        T t = A_::mul(mA(0, r), z[0]);
#pragma unroll
        for (int k = 1; k < S; ++k) t = A_::madd(t, mA(k, r), z[k]);
        return t;
    }
    __device__ __forceinline__ T dotB(int r, const T* uu) const
    {
        T t = A_::mul(mB(r, 0), uu[0]);
#pragma unroll
        for (int k = 1; k < I; ++k) t = A_::madd(t, mB(r, k), uu[k]);
        return t;
    }
    __device__ __forceinline__ T dotBT(int j, const T* z) const
    {
        T t = A_::mul(mB(0, j), z[0]);
#pragma unroll
        for (int k = 1; k < S; ++k) t = A_::madd(t, mB(k, j), z[k]);
        return t;
    }

    // One gradient-shaped pass.
    //   ITERATE lanes: df = H u + MM via the forward/backward recursion (mpc.h:275-283) fused with the stopping
    //                  test's arg-max over the free coordinates (mpc.h:289-309).
    //   SETUP lanes (MIXED only): MM = trans(K) Q (free response - target) (mpc.h:258-266) through the same chain.
    // The arg-max is taken while walking the horizon backwards with ">=" — that selects the same element as the
    // reference's forward scan with strict ">" (the lowest row-major index among equal maxima); "a > 0" keeps
    // the reference's (0,0) default when every free gradient entry is zero.
    template <bool MIXED>
    __device__ __forceinline__ void pass(bool setup, const T* xs, const T* tgc, long long inst, T& max_df, int& max_idx,
                                         T& max_val)
    {
        T m[S];
        // ---- horizon step 0 -----------------------------------------------------------------------------------
        {
            T u0v[I];
#pragma unroll
            for (int j = 0; j < I; ++j) u0v[j] = getU(j);
#pragma unroll
            for (int r = 0; r < S; ++r) {
                T tb = dotB(r, u0v);                                   // M[0] = B*controls[0]          mpc.h:275
                if constexpr (MIXED) {
                    T ta = A_::add(dotA(r, xs), mC(r));               // M[0] = A*initial_state + C   mpc.h:258
                    m[r] = setup ? ta : tb;
                } else m[r] = tb;
            }
        }
        T w[S];
        // ---- forward recursion, Q scaling, stash ----------------------------------------------------------------
#pragma unroll(UNROLL)
        for (int i = 0; i < N; ++i) {
            if (i > 0) {
                T ui[I], mn[S];
#pragma unroll
                for (int j = 0; j < I; ++j) ui[j] = getU(i * I + j);
#pragma unroll
                for (int r = 0; r < S; ++r) {
                    T t1 = dotA(r, m);                                  // A*M[i-1]
                    T ad = dotB(r, ui);                                 // + B*controls[i]               mpc.h:277
                    if constexpr (MIXED) ad = setup ? mC(r) : ad;      // + C                           mpc.h:260
                    mn[r] = A_::add(t1, ad);
                }
#pragma unroll
                for (int r = 0; r < S; ++r) m[r] = mn[r];
            }
#pragma unroll
            for (int r = 0; r < S; ++r) {
                T d = m[r];
                if constexpr (MIXED) {                                  // M[i] - target[i]              mpc.h:262
                    T tg = tgc[r];
                    if (p.target_full && setup) {
                        int ph = p.head + i; if (ph >= N) ph -= N;
                        tg = p.target[((long long)ph * S + r) * p.stride + inst];
                    }
                    d = setup ? A_::sub(m[r], tg) : d;
                }
                T q = A_::mul(mQ(r), d);                                // diagm(Q)*M[i]                 mpc.h:262,279
                if (i < N - 1) stash_st((L::O_QM + i * S + r) * 32, q); else w[r] = q;
            }
        }
        // ---- backward recursion, gradient, arg-max ---------------------------------------------------------------
        max_df = T(0); max_idx = 0; max_val = T(0);
#pragma unroll(UNROLL)
        for (int i = N - 1; i >= 0; --i) {
#pragma unroll
            for (int j = I - 1; j >= 0; --j) {
                const int k = i * I + j;
                T bt = dotBT(j, w);                                     // trans(B)*M[i]                 mpc.h:266,283
                T uij = getU(k);
                T mmv = getMM(k);
                if constexpr (MIXED) { mmv = setup ? bt : mmv; setMM(k, mmv); }
                T dfv = A_::add(A_::add(mmv, bt), A_::mul(mR(j), uij)); // MM + B'M + diagm(R)*u         mpc.h:283
                sm[(L::O_DF + k) * 32] = dfv;
                const bool active = (uij <= mLo(j) && dfv > T(0)) || (uij >= mHi(j) && dfv < T(0));   // mpc.h:296-299
                const T a = absval(dfv);
                if (!active && a >= max_df && a > T(0)) { max_df = a; max_idx = k; max_val = dfv; }  // mpc.h:301-306
            }
            if (i > 0) {
                T wn[S];
#pragma unroll
                for (int r = 0; r < S; ++r)                              // M[i-1] += trans(A)*M[i]       mpc.h:264,281
                    wn[r] = A_::add(stash_ld((L::O_QM + (i - 1) * S + r) * 32), dotAT(r, w));
#pragma unroll
                for (int r = 0; r < S; ++r) w[r] = wn[r];
            }
        }
    }
};

// The kernel.  grid = #SMs (persistent), block = WARPS*32, dynamic smem = Layout::SMEM_BYTES.
template <typename T, int S, int I, int N, bool EXACT, bool PIM>
__global__ void __launch_bounds__(Layout<T, S, I, N, PIM>::THREADS, 1)
solve_kernel(const __grid_constant__ SolveParams<T, S, I, N> p)
{
    using L = Layout<T, S, I, N, PIM>;
    using A_ = Ar<T, EXACT>;
    using SV = Solver<T, S, I, N, EXACT, PIM>;
    constexpr int NI = L::NI;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T* sm = reinterpret_cast<T*>(smem_raw) + (size_t)warp * L::ELEMS * 32 + lane;
    SV sv(p, sm);

    long long inst = -1;
    bool need = true, active = false, setup = false;
    uint32_t iter = 0;
    bool v_dirty = false;
    T xs[S], tgc[S];
#pragma unroll
    for (int r = 0; r < S; ++r) { xs[r] = T(0); tgc[r] = T(0); }
    T lambda = p.lambda, inv_lambda = T(0), beta = T(0);
    if constexpr (!PIM) {
        inv_lambda = A_::div(T(1), lambda);                                                   // 1.0/lambda   mpc.h:342
        const T sq = A_::sqrt(lambda);
        beta = A_::div(A_::sub(sq, T(1)), A_::add(sq, T(1)));                                // mpc.h:343
    }

    while (true) {
        // ---------------- fetch: idle lanes claim the next instances (one aggregated atomic per warp) ----------
        const unsigned needmask = __ballot_sync(FULL, need);
        if (needmask) {
            const int leader = __ffs(needmask) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(p.queue, (unsigned long long)__popc(needmask));
            base = __shfl_sync(FULL, base, leader);
            if (need) {
                need = false;
                const long long my = (long long)base + __popc(needmask & ((1u << lane) - 1u));
                if (my < p.count) {
                    inst = my; active = true; setup = true; iter = 0; v_dirty = false;
#pragma unroll
                    for (int r = 0; r < S; ++r) xs[r] = p.x0[(long long)r * p.stride + inst];
                    if (!p.target_full) {
#pragma unroll
                        for (int r = 0; r < S; ++r) tgc[r] = p.target[(long long)r * p.stride + inst];
                    }
                    // controls[i-1] = controls[i] (mpc.h:231-232) folded into the load; fresh controller: zeros
#pragma unroll(SV::UNROLL)
                    for (int i = 0; i < N; ++i) {
                        const int src = (i + 1 < N) ? i + 1 : N - 1;
#pragma unroll
                        for (int j = 0; j < I; ++j)
                            sv.setU(i * I + j, p.warm ? p.controls[(long long)(src * I + j) * p.stride + inst] : T(0));
                    }
                    if constexpr (PIM) {
#pragma unroll
                        for (int e = 0; e < S * S; ++e) sm[(L::O_MODEL + L::M_A + e) * 32] = p.mA[(long long)e * p.stride + inst];
#pragma unroll
                        for (int e = 0; e < S * I; ++e) sm[(L::O_MODEL + L::M_B + e) * 32] = p.mB[(long long)e * p.stride + inst];
#pragma unroll
                        for (int e = 0; e < S; ++e) {
                            sm[(L::O_MODEL + L::M_C + e) * 32] = p.mC[(long long)e * p.stride + inst];
                            sm[(L::O_MODEL + L::M_Q + e) * 32] = p.mQ[(long long)e * p.stride + inst];
                        }
#pragma unroll
                        for (int e = 0; e < I; ++e) {
                            sm[(L::O_MODEL + L::M_R + e) * 32] = p.mR[(long long)e * p.stride + inst];
                            sm[(L::O_MODEL + L::M_LO + e) * 32] = p.mlo[(long long)e * p.stride + inst];
                            sm[(L::O_MODEL + L::M_HI + e) * 32] = p.mhi[(long long)e * p.stride + inst];
                        }
                        lambda = p.mlambda[inst];
                        inv_lambda = A_::div(T(1), lambda);
                        const T sq = A_::sqrt(lambda);
                        beta = A_::div(A_::sub(sq, T(1)), A_::add(sq, T(1)));
                    }
                } else { inst = -1; active = false; }
            }
        }
        if (!__any_sync(FULL, active)) break;

        // ---------------- one pass for the whole warp ------------------------------------------------------------
        T max_df, max_val; int max_idx;
        if (__any_sync(FULL, active && setup)) sv.template pass<true>(setup && active, xs, tgc, inst < 0 ? 0 : inst, max_df, max_idx, max_val);
        else sv.template pass<false>(false, xs, tgc, 0, max_df, max_idx, max_val);

        // ---------------- per-lane state machine ---------------------------------------------------------------------
        if (active) {
            bool finished = false, hit_max = false;
            if (setup) {
                setup = false; iter = 0;                                   // mpc.h:270
                if (p.max_iter == 0) { finished = true; hit_max = true; }  // loop body never runs (mpc.h:271)
            } else if (max_df < p.eps) {                                   // mpc.h:310-311
                finished = true;
            } else {
                if (iter < 50) {                                           // mpc.h:319-335 greedy coordinate step
                    T qd;
                    if constexpr (PIM) qd = p.mqdiag[(long long)max_idx * p.stride + inst]; else qd = p.qdiag[max_idx];
                    const int j = max_idx % I;
                    if (qd != T(0)) {                                      // else `continue` (mpc.h:321-322)
                        T uo;
                        if constexpr (SV::UREG) {
                            uo = T(0);
#pragma unroll
                            for (int k = 0; k < NI; ++k) uo = (k == max_idx) ? sv.u[k] : uo;
                        } else uo = sv.getU(max_idx);
                        T un = A_::div(-A_::sub(max_val, A_::mul(qd, uo)), qd);                   // mpc.h:327
                        un = put_in_range(sv.mLo(j), sv.mHi(j), un);                               // mpc.h:328
                        if constexpr (SV::UREG) {
#pragma unroll
                            for (int k = 0; k < NI; ++k) sv.u[k] = (k == max_idx) ? un : sv.u[k];
                        } else sv.setU(max_idx, un);
                        if (iter + 1 == 50) {                                                      // mpc.h:331-334 v = controls
#pragma unroll(SV::UNROLL_K)
                            for (int k = 0; k < NI; ++k) sm[(L::O_V + k) * 32] = sv.getU(k);
                            v_dirty = true;
                        }
                    } else if (iter + 1 == 50) {
                        // "stale v": the reference enters the projected-gradient phase with whatever its member v
                        // holds (mpc.h:322 skipped :331-334).  Persistent controller: v of its previous solve;
                        // fresh controller: zero storage (the single de-undefinition).
#pragma unroll(SV::UNROLL_K)
                        for (int k = 0; k < NI; ++k)
                            sm[(L::O_V + k) * 32] = p.warm ? p.vstate[(long long)k * p.stride + inst] : T(0);
                    }
                } else {                                                   // mpc.h:336-345 accelerated projected gradient
#pragma unroll(SV::UNROLL_K)
                    for (int k = 0; k < NI; ++k) {
                        const int j = k % I;
                        const T lo = sv.mLo(j), hi = sv.mHi(j);
                        const T v_old = sm[(L::O_V + k) * 32];
                        const T vn = clamp_m(A_::sub(sv.getU(k), A_::mul(inv_lambda, sm[(L::O_DF + k) * 32])), lo, hi);
                        sm[(L::O_V + k) * 32] = vn;
                        sv.setU(k, clamp_m(A_::madd(vn, beta, A_::sub(vn, v_old)), lo, hi));
                    }
                    v_dirty = true;
                }
                ++iter;
                if (iter >= p.max_iter) { finished = true; hit_max = true; }   // mpc.h:271
            }
            if (finished) {
#pragma unroll
                for (int j = 0; j < I; ++j) p.u0[(long long)j * p.stride + inst] = sv.getU(j);   // mpc.h:239
                p.iters[inst] = iter;
                p.status[inst] = hit_max ? 1 : 0;
                if (p.warm || p.keep_controls) {
#pragma unroll(SV::UNROLL_K)
                    for (int k = 0; k < NI; ++k) p.controls[(long long)k * p.stride + inst] = sv.getU(k);
                }
                if (p.warm) {
                    if (v_dirty) {
#pragma unroll(SV::UNROLL_K)
                        for (int k = 0; k < NI; ++k) p.vstate[(long long)k * p.stride + inst] = sm[(L::O_V + k) * 32];
                    }
                    if (p.target_full) {
                        // target[i-1] = target[i] (mpc.h:236-237) on the ring: the host advances `head`; the slot
                        // that becomes the last horizon step must repeat the old last step.
                        int last = p.head + N - 1; if (last >= N) last -= N;
#pragma unroll
                        for (int r = 0; r < S; ++r)
                            p.target[((long long)p.head * S + r) * p.stride + inst] =
                                p.target[((long long)last * S + r) * p.stride + inst];
                    }
                }
                active = false; need = true;
            }
        }
    }
}

}  // namespace bmpc
