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
//    (16 lanes/SMSP/clk on B200) busy from two warps per SM sub-partition.  So occupancy is spent on storage, not
//    on latency hiding: u[] and the linear term MM[] live in registers (fully unrolled horizon), the stashed
//    Q*M[] rows, the gradient and the Nesterov v[] in shared memory as [element][lane] (conflict-free 8-byte
//    accesses, no HBM traffic inside the iteration loop).
//  * iteration counts are data dependent (p99/p50 ~ 2-14x).  A lane therefore never waits for its warp: it is a
//    small state machine (SETUP -> ITERATE ... -> finished -> fetch next) and every trip of the warp-wide loop
//    runs exactly one gradient-shaped pass for all 32 lanes.  The per-solve linear term (mpc.h:258-266) has the
//    same recursion structure as the gradient, so SETUP lanes ride the same pass with predicated operands
//    (MIXED pass); there is no divergent code of any size.
//  * STRUCTURAL SPARSITY (Pattern): vehicle models are mostly zeros and ones (the (4,1) bicycle model has 7 non-zero
//    entries in A, two of them exactly 1, and one in B; Q weights two of four states).  dlib multiplies them out
//    densely.  A kernel instantiated for a pattern drops the terms whose coefficient is structurally zero and the
//    multiplications by exactly 1 — both are exact in IEEE arithmetic (x + 0*y == x and 1*x == x for finite y; only
//    the sign of a zero result can differ), so EXACT-mode parity is unaffected while the instruction count per
//    horizon step falls ~3x and the fully unrolled pass fits the instruction cache.  The host picks, per model,
//    the sparsest compiled pattern the model's entries satisfy (dispatch.h); the dense pattern always qualifies.
#pragma once
#include <stdint.h>

#include "arith.cuh"

namespace bmpc {

constexpr int kSmemMax = 232448;  // 227 KB opt-in dynamic shared memory per CTA on sm_100

// Bit e of *NZ: entry e (row-major) may be non-zero.  Bit e of *ONE: entry e is exactly 1.0 (implies NZ).
template <unsigned ANZ_, unsigned AONE_, unsigned BNZ_, unsigned BONE_, unsigned CNZ_, unsigned QNZ_>
struct Pattern {
    static constexpr unsigned ANZ = ANZ_, AONE = AONE_, BNZ = BNZ_, BONE = BONE_, CNZ = CNZ_, QNZ = QNZ_;
};
template <int S, int I>
using DensePattern = Pattern<(1u << (S * S)) - 1u, 0u, (1u << (S * I)) - 1u, 0u, (1u << S) - 1u, (1u << S) - 1u>;

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
    unsigned long long* queue;   // [0] next unclaimed instance; [1] ring tail; [2] ring head; [3] instances stage B has
                                 // not finished; [4] error word (ring overrun); [6] stage-C list length; [7] stage-C tickets
    // split solve (pg_kernel.cuh): stage A parks an instance that reaches iter == 50 in a slot of the hand-over ring
    int split;
    uint32_t* pg_list;   // ring of instance ids, capacity pg_mask + 1 (a power of two), 0xffffffff = empty
    unsigned pg_mask;
    T* park;             // [3*N*I][pg_mask + 1]  state of parked instances BY RING SLOT: controls, v, linear term MM (mpc.h:266)
    uint32_t* park_iter; // [pg_mask + 1]         their iteration counters
    int warp_limit;      // warps of a CTA that take work (small batches are spread one warp per SM sub-partition first)
    int pg_slice;        // stage B may time-slice small batches (pg_kernel.cuh)
    int pg_reuse;        // stage B recycles ring slots (parking enabled)
    int sym;             // stage B runs the symmetric-bound kernel: instances whose linear term is not finite stay in stage A
    // stage C (pair_kernel.cuh): one instance per lane pair.  c_direct: stage B is skipped, stage C consumes stage A's
    // hand-over ring [0, queue[1]); else stage B parks its last c_threshold instances in c_list / c_park / c_iter
    // (by-slot rows of stride c_stride, tail in queue[6]) and stage C consumes those.  queue[7] = stage C's tickets.
    int c_direct;
    unsigned c_threshold;   // stage B hands over once this many instances or fewer are unfinished (0 = never)
    uint32_t* c_list;
    T* c_park;
    uint32_t* c_iter;
    unsigned c_stride;
};

__host__ __device__ constexpr int popc_c(unsigned x) { int n = 0; while (x) { n += int(x & 1u); x >>= 1; } return n; }
__host__ __device__ constexpr int imax_c(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int imin_c(int a, int b) { return a < b ? a : b; }

#ifndef BMPC_UREG_MAX
#define BMPC_UREG_MAX 64
#endif
template <typename T, int S, int I, int N, bool PIM, typename PAT>
struct Layout {
    static constexpr int NI = N * I;
    static constexpr int WORDS = int(sizeof(T) / 4);
    // u[] in registers (fully unrolled horizon) when it costs at most 64 32-bit registers; the linear term MM[] too
    // when both together stay within 128
    static constexpr bool UREG = (NI * WORDS <= BMPC_UREG_MAX);
    static constexpr bool MMREG = UREG && (2 * NI * WORDS <= 128);
    // Per horizon step: one slot per Q-weighted state row (the Q*M stash of the forward sweep, read back by the
    // backward sweep) — and, once the backward sweep has consumed a step's stash, the same slots hold that step's
    // gradient entries df[i][0..I).
    static constexpr int SQ = popc_c(PAT::QNZ & ((1u << S) - 1u));
    static constexpr int W = imax_c(SQ, I);
    static constexpr int O_QM = 0;
    static constexpr int O_V = O_QM + N * W;             // Nesterov v
    static constexpr int O_U = O_V + NI;
    static constexpr int O_MM = O_U + (UREG ? 0 : NI);   // linear term MM
    static constexpr int O_MODEL = O_MM + (MMREG ? 0 : NI);
    static constexpr int M_A = 0, M_B = M_A + S * S, M_C = M_B + S * I, M_Q = M_C + S, M_R = M_Q + S,
                         M_LO = M_R + I, M_HI = M_LO + I, M_LEN = M_HI + I;
    static constexpr int ELEMS = O_MODEL + (PIM ? M_LEN : 0);
    static constexpr int WARP_BYTES = ELEMS * 32 * int(sizeof(T));
    static constexpr int FIT = kSmemMax / WARP_BYTES;
#ifndef BMPC_CAP_WARPS
#define BMPC_CAP_WARPS 8
#endif
    static constexpr int CAP = UREG ? BMPC_CAP_WARPS : 16;   // 255 registers/thread allow 8 warps per SM
    static constexpr int WARPS = imin_c(FIT, CAP);
    static constexpr int THREADS = WARPS * 32;
    static constexpr int SMEM_BYTES = WARPS * WARP_BYTES;
    static_assert(WARPS >= 1, "horizon too long for the shared-memory resident layout");
    // stash slot of state row r (rank of r among the Q-weighted rows)
    __host__ __device__ static constexpr int qslot(int r) { return popc_c(PAT::QNZ & ((1u << r) - 1u)); }
};

template <typename T> struct BitsOf;
template <> struct BitsOf<double> {
    using type = unsigned long long;
    static constexpr type abs_mask = 0x7fffffffffffffffull, inf = 0x7ff0000000000000ull;
    __device__ __forceinline__ static type to(double x) { return (type)__double_as_longlong(x); }
    __device__ __forceinline__ static double from(type b) { return __longlong_as_double((long long)b); }
};
template <> struct BitsOf<float> {
    using type = unsigned;
    static constexpr type abs_mask = 0x7fffffffu, inf = 0x7f800000u;
    __device__ __forceinline__ static type to(float x) { return __float_as_uint(x); }
    __device__ __forceinline__ static float from(type b) { return __uint_as_float(b); }
};

template <typename T, int S, int I, int N, bool EXACT, bool PIM, typename PAT>
struct Solver {
    using L = Layout<T, S, I, N, PIM, PAT>;
    using P = SolveParams<T, S, I, N>;
    using A_ = Ar<T, EXACT>;
    static constexpr int NI = L::NI;
    static constexpr bool UREG = L::UREG, MMREG = L::MMREG;
    static constexpr int UNROLL = UREG ? N : 2;        // horizon loops: full unroll keeps u[]/mm[] in registers
    static constexpr int UNROLL_K = UREG ? NI : 2;     // loops over the N*I control entries

    using Bits = typename BitsOf<T>::type;
    static constexpr Bits kAbsMask = BitsOf<T>::abs_mask, kInfBits = BitsOf<T>::inf;
    __device__ __forceinline__ static Bits to_bits(T x) { return BitsOf<T>::to(x); }
    __device__ __forceinline__ static T from_bits(Bits b) { return BitsOf<T>::from(b); }

    const P& p;
    T* sm;            // this lane's column of the warp's shared-memory block: element e at sm[e*32]
    T u[UREG ? NI : 1];
    T mm[MMREG ? NI : 1];

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
    __device__ __forceinline__ T getMM(int k) const { if constexpr (MMREG) return mm[k]; else return sm[(L::O_MM + k) * 32]; }
    __device__ __forceinline__ void setMM(int k, T v) { if constexpr (MMREG) mm[k] = v; else sm[(L::O_MM + k) * 32] = v; }

    // The Q*M stash is written in the forward sweep and read back in the backward sweep of the SAME unrolled pass.
    // Through plain pointers nvcc forwards every stored value to its load, i.e. keeps N*S live values in registers
    // and spills them to local memory; volatile accesses make the round trip through shared memory real.
    __device__ __forceinline__ void stash_st(int off, T v) { *reinterpret_cast<volatile T*>(sm + off) = v; }
    __device__ __forceinline__ T stash_ld(int off) const { return *reinterpret_cast<const volatile T*>(sm + off); }
    __device__ __forceinline__ T getDF(int k) const { return stash_ld((L::O_QM + (k / I) * L::W + (k % I)) * 32); }

    // One term of a matrix.h:43-60 dot product under the structural pattern: skipped when the coefficient is
    // structurally zero, a plain copy/add when it is exactly one; the first surviving term initialises.
    // (`nz`, `one`, `have` are compile-time constants after unrolling.)
    __device__ __forceinline__ static void term(T& t, bool& have, bool nz, bool one, T coef, T z)
    {
        if (!nz) return;
        if (one) t = have ? A_::add(t, z) : z;
        else t = have ? A_::madd(t, coef, z) : A_::mul(coef, z);
        have = true;
    }
    __host__ __device__ static constexpr bool bit(unsigned m, int e) { return ((m >> e) & 1u) != 0u; }
    __device__ __forceinline__ bool dotA(int r, const T* z, T& t) const        // A*z, row r
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::ANZ, r * S + k), bit(PAT::AONE, r * S + k), mA(r, k), z[k]);
        return have;
    }
    __device__ __forceinline__ bool dotAT(int r, const T* z, T& t) const       // trans(A)*z, matrix_utilities.h:375-410
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::ANZ, k * S + r), bit(PAT::AONE, k * S + r), mA(k, r), z[k]);
        return have;
    }
    __device__ __forceinline__ bool dotB(int r, const T* uu, T& t) const       // B*u, row r
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < I; ++k) term(t, have, bit(PAT::BNZ, r * I + k), bit(PAT::BONE, r * I + k), mB(r, k), uu[k]);
        return have;
    }
    __device__ __forceinline__ bool dotBT(int j, const T* z, T& t) const       // trans(B)*z, row j
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::BNZ, k * I + j), bit(PAT::BONE, k * I + j), mB(k, j), z[k]);
        return have;
    }
    // a + b where either side may be structurally absent (x + 0 == x exactly)
    __device__ __forceinline__ static T add_opt(bool ha, T a, bool hb, T b)
    {
        return ha ? (hb ? A_::add(a, b) : a) : (hb ? b : T(0));
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
                T tb; const bool hb = dotB(r, u0v, tb);               // M[0] = B*controls[0]          mpc.h:275
                if constexpr (MIXED) {
                    T ta; const bool ha = dotA(r, xs, ta);            // M[0] = A*initial_state + C   mpc.h:258
                    const T tas = add_opt(ha, ta, bit(PAT::CNZ, r), mC(r));
                    m[r] = setup ? tas : (hb ? tb : T(0));
                } else m[r] = hb ? tb : T(0);
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
                    T t1; const bool h1 = dotA(r, m, t1);              // A*M[i-1]
                    T ad; bool hb = dotB(r, ui, ad);                   // + B*controls[i]               mpc.h:277
                    if constexpr (MIXED) {                             // + C                           mpc.h:260
                        const bool hc = bit(PAT::CNZ, r);
                        if (hb || hc) { ad = setup ? (hc ? mC(r) : T(0)) : (hb ? ad : T(0)); hb = true; }
                    }
                    mn[r] = add_opt(h1, t1, hb, ad);
                }
#pragma unroll
                for (int r = 0; r < S; ++r) m[r] = mn[r];
            }
#pragma unroll
            for (int r = 0; r < S; ++r) {
                if (bit(PAT::QNZ, r)) {
                    T d = m[r];
                    if constexpr (MIXED) {                              // M[i] - target[i]              mpc.h:262
                        T tg = tgc[r];
                        if (p.target_full && setup) {
                            int ph = p.head + i; if (ph >= N) ph -= N;
                            tg = p.target[((long long)ph * S + r) * p.stride + inst];
                        }
                        d = setup ? A_::sub(m[r], tg) : d;
                    }
                    T q = A_::mul(mQ(r), d);                            // diagm(Q)*M[i]                 mpc.h:262,279
                    if (i < N - 1) stash_st((L::O_QM + i * L::W + L::qslot(r)) * 32, q); else w[r] = q;
                } else if (i == N - 1) w[r] = T(0);                     // Q(r) == 0: the scaled row is a zero
            }
        }
        // ---- backward recursion, gradient, arg-max ---------------------------------------------------------------
        max_df = T(0); max_idx = 0; max_val = T(0);
        Bits max_key = 0; bool max_neg = false;
        T lo_[I], hi_[I];
#pragma unroll
        for (int j = 0; j < I; ++j) { lo_[j] = mLo(j); hi_[j] = mHi(j); }
#pragma unroll(UNROLL)
        for (int i = N - 1; i >= 0; --i) {
#pragma unroll
            for (int j = I - 1; j >= 0; --j) {
                const int k = i * I + j;
                T bt; const bool hbt = dotBT(j, w, bt);                 // trans(B)*M[i]                 mpc.h:266,283
                T uij = getU(k);
                T mmv = getMM(k);
                if constexpr (MIXED) { mmv = setup ? (hbt ? bt : T(0)) : mmv; setMM(k, mmv); }
                T dfv = A_::add(add_opt(true, mmv, hbt, bt), A_::mul(mR(j), uij)); // MM + B'M + diagm(R)*u  mpc.h:283
                stash_st((L::O_QM + i * L::W + j) * 32, dfv);           // this step's stash slots are dead by now
                // all four comparisons evaluated (no short-circuit: a lazily fetched bound serialises the sweep)
                // The bound tests stay floating-point compares; sign, magnitude and the running maximum are done on
                // the bit patterns (integer pipe): for |df| in (0, +inf] the IEEE order equals the unsigned order of
                // the bits, zeros and NaNs are never selected (as in mpc.h:301: abs(NaN) > max_df is false).
                const Bits bits = to_bits(dfv);
                const Bits a = bits & kAbsMask;
                const bool neg = (bits & ~kAbsMask) != 0;
                const bool sel = (Bits)(a - 1) < kInfBits;                       // 0 < |df| <= inf, not NaN
                const bool active = neg ? (uij >= hi_[j]) : (uij <= lo_[j]);     // mpc.h:296-299 given df != 0
                if (sel & !active & (a >= max_key)) { max_key = a; max_idx = k; max_neg = neg; }   // mpc.h:301-306
            }
            if (i > 0) {
                T wn[S];
#pragma unroll
                for (int r = 0; r < S; ++r) {                            // M[i-1] += trans(A)*M[i]       mpc.h:264,281
                    T t; const bool h = dotAT(r, w, t);
                    const bool hq = bit(PAT::QNZ, r);
                    T st = T(0);
                    if (hq) st = stash_ld((L::O_QM + (i - 1) * L::W + L::qslot(r)) * 32);
                    wn[r] = add_opt(hq, st, h, t);
                }
#pragma unroll
                for (int r = 0; r < S; ++r) w[r] = wn[r];
            }
        }
        max_df = from_bits(max_key);
        max_val = from_bits(max_key | (max_neg ? ~kAbsMask : Bits(0)));
    }
};

// The kernel.  grid = #SMs (persistent), block = WARPS*32, dynamic smem = Layout::SMEM_BYTES.
template <typename T, int S, int I, int N, bool EXACT, bool PIM, typename PAT>
__global__ void __launch_bounds__((Layout<T, S, I, N, PIM, PAT>::THREADS), 1)
solve_kernel(const __grid_constant__ SolveParams<T, S, I, N> p)
{
    using L = Layout<T, S, I, N, PIM, PAT>;
    using A_ = Ar<T, EXACT>;
    using SV = Solver<T, S, I, N, EXACT, PIM, PAT>;
    constexpr int NI = L::NI;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= p.warp_limit) return;          // no block-wide barrier anywhere in this kernel
    T* sm = reinterpret_cast<T*>(smem_raw) + (size_t)warp * L::ELEMS * 32 + lane;
    SV sv(p, sm);

    long long inst = -1;
    bool need = true, active = false, setup = false, nonfinite = false;
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
        bool hand = false;
        if (active) {
            bool finished = false, hit_max = false;
            if (setup) {
                setup = false; iter = 0;                                   // mpc.h:270
                // a NaN / inf in x0, the targets or the model reaches every later number through the linear term
                // (x0 and a constant target are looked at directly: a pattern-specialised kernel never reads the
                // entries that only meet structural zeros)
                bool finite = true;
#pragma unroll(SV::UNROLL_K)
                for (int k = 0; k < NI; ++k) finite &= (SV::to_bits(sv.getMM(k)) & SV::kInfBits) != SV::kInfBits;
#pragma unroll
                for (int r = 0; r < S; ++r)
                    finite &= (SV::to_bits(xs[r]) & SV::kInfBits) != SV::kInfBits && (SV::to_bits(tgc[r]) & SV::kInfBits) != SV::kInfBits;
                nonfinite = !finite;
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
                        const T vn = clamp_m(A_::sub(sv.getU(k), A_::mul(inv_lambda, sv.getDF(k))), lo, hi);
                        sm[(L::O_V + k) * 32] = vn;
                        sv.setU(k, clamp_m(A_::madd(vn, beta, A_::sub(vn, v_old)), lo, hi));
                    }
                    v_dirty = true;
                }
                ++iter;
                if (iter >= p.max_iter) { finished = true; hit_max = true; }   // mpc.h:271
                else if (p.split && iter == 50) {
                    // Hand over to stage B (pg_kernel.cuh).  Its symmetric-bound kernel assumes finite numbers; an
                    // instance whose linear term is not (NaN / inf in x0 or the targets) finishes here instead, with
                    // the literal arithmetic of the one-kernel path.
                    hand = !(p.sym && nonfinite);
                }
            }
            if (finished) {
#pragma unroll
                for (int j = 0; j < I; ++j) p.u0[(long long)j * p.stride + inst] = sv.getU(j);   // mpc.h:239
                p.iters[inst] = iter;
                p.status[inst] = (hit_max ? 1 : 0) | (nonfinite ? 2 : 0);
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
        // ---------------- hand-over to stage B: the projected-gradient phase continues there from exactly this state —
        // controls, v (set by mpc.h:331-334, or stale) and MM.  The lanes of a warp reach iter == 50 together, so they
        // take consecutive ring slots (one aggregated atomic) and the by-slot state rows are written coalesced.
        const unsigned handmask = __ballot_sync(FULL, hand);
        if (handmask) {
            const int leader = __ffs(handmask) - 1;
            unsigned long long pos = 0;
            if (lane == leader) {
                pos = atomicAdd(p.queue + 1, (unsigned long long)__popc(handmask));
                atomicAdd(p.queue + 3, (unsigned long long)__popc(handmask));   // instances stage B still has to finish
            }
            pos = __shfl_sync(FULL, pos, leader);
            if (hand) {
                const unsigned ps = (unsigned)(pos + (unsigned long long)__popc(handmask & ((1u << lane) - 1u))) & p.pg_mask;
                const long long rstride = (long long)p.pg_mask + 1;
#pragma unroll(SV::UNROLL_K)
                for (int k = 0; k < NI; ++k) {
                    p.park[(long long)k * rstride + ps] = sv.getU(k);
                    p.park[(long long)(NI + k) * rstride + ps] = sm[(L::O_V + k) * 32];
                    p.park[(long long)(2 * NI + k) * rstride + ps] = sv.getMM(k);
                }
                p.park_iter[ps] = iter;                                // stage B resumes from here
                p.status[inst] = nonfinite ? 2 : 0;                    // stage B adds BMPC_ST_MAXITER
                p.pg_list[ps] = (uint32_t)inst;
                active = false; need = true;
            }
        }
    }
}

}  // namespace bmpc
