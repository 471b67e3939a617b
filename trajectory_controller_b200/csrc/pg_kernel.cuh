// Stage B of the split solve: the accelerated projected-gradient phase of dlib::mpc's solve_linear_mpc
// (/root/reference/dlib_files/dlib/control/mpc.h:271-345 for iter >= 50) as its own persistent kernel.
//
// Why a second kernel.  A cold (4,1,30) solve of BASELINE config 2 runs 704 iterations on average: 50 greedy
// coordinate steps (mpc.h:319-335) and ~650 projected-gradient steps (mpc.h:336-345).  The one-kernel solver
// (solve_kernel.cuh) serves both phases — and the set-up pass — from one warp-wide loop, so every trip pays for the
// arg-max bookkeeping the coordinate step needs, for the coordinate step's register-indexed selects and for a
// separate sweep that re-reads the gradient to apply the projected-gradient update.  Stage A (solve_kernel with
// p.split) stops at iter == 50, parks u / v / MM in a slot of the hand-over ring and this kernel continues from there
// with a pass that contains nothing but the phase's own arithmetic:
//
//   * the stopping test (mpc.h:289-311) needs no arg-max here, only "is any free |df| >= eps";
//   * the update (mpc.h:339-344) is applied inside the backward sweep, as soon as df[i] exists, while the chain
//     of the next horizon step is in flight (more independent FP64 work per warp).  It is provisional: the reference
//     tests convergence BEFORE updating, so the old (u, v) of each entry are parked in the stash slots the backward
//     sweep has just consumed, and a lane that turns out to have converged reports those;
//   * shared-memory traffic is paired: the two Q-weighted rows of a step, and (u_old, v_old), move as one 128-bit
//     access (layout [pair][lane][2], conflict-free), 5 LSU instructions per horizon step instead of 8;
//   * SYM kernels (bounds lower == -upper, the case of every model this project ships: +-22 deg,
//     src/trajectory_point_follower.cpp:16-18) replace the seven FP64 compares per control entry of the literal
//     formulation by at most three, and those can run on the integer pipe (see PgSolver::pass).
//
// Arithmetic, rounding order and iteration counts are those of solve_kernel.cuh (same Ar<T,EXACT> policies, same
// structural-pattern term dropping): the parity tests compare the two-stage result with the oracle bit for bit.
#pragma once
#include "solve_kernel.cuh"

// Number of entries of the linear term MM kept in shared memory instead of registers when the horizon is
// register-resident (profiles/r1_experiments.md rows 16, 18; 0 = all in registers), and the resident-warp cap.
#ifndef BMPC_PG_MM_SMEM
#define BMPC_PG_MM_SMEM 23
#endif
#ifndef BMPC_PG_CAP_WARPS
#define BMPC_PG_CAP_WARPS 8
#endif
// SYM kernels (symmetric bounds, see PgSolver::pass): which of the remaining floating-point compares run on the
// integer pipe instead (bit-pattern compares; exact for the non-NaN values the SYM preconditions guarantee).
#ifndef BMPC_PG_IBIG
#define BMPC_PG_IBIG 0      // |df| >= eps
#endif
#ifndef BMPC_PG_ICLAMP
#define BMPC_PG_ICLAMP 0    // |t| < h of the two clamps
#endif
#ifndef BMPC_PG_SLICE
#define BMPC_PG_SLICE 128   // passes between two fairness parks while more instances than lanes are unfinished
#endif
#ifndef BMPC_PG_CHECK
#define BMPC_PG_CHECK 32    // passes between two looks at the number of unfinished instances (a power of two)
#endif

namespace bmpc {

template <typename T> struct SmemIO;
template <> struct SmemIO<double> {
    __device__ __forceinline__ static void st2(unsigned a, double x, double y) { asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(x), "d"(y)); }
    __device__ __forceinline__ static void ld2(unsigned a, double& x, double& y) { asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(a)); }
    __device__ __forceinline__ static void st1(unsigned a, double x) { asm volatile("st.shared.f64 [%0], %1;" ::"r"(a), "d"(x)); }
    __device__ __forceinline__ static double ld1(unsigned a) { double x; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(a)); return x; }
};
template <> struct SmemIO<float> {
    __device__ __forceinline__ static void st2(unsigned a, float x, float y) { asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y)); }
    __device__ __forceinline__ static void ld2(unsigned a, float& x, float& y) { asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(x), "=f"(y) : "r"(a)); }
    __device__ __forceinline__ static void st1(unsigned a, float x) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(x)); }
    __device__ __forceinline__ static float ld1(unsigned a) { float x; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(a)); return x; }
};

template <typename T, int S, int I, int N, bool PIM, typename PAT>
struct PgLayout {
    using L0 = Layout<T, S, I, N, PIM, PAT>;
    static constexpr int NI = N * I;
    // Short horizons keep u[] and MM[] in registers (fully unrolled horizon).  Long ones (N*I*sizeof(T) > 256 B) keep
    // them in this lane's shared-memory column and run the same pass as a rolled loop.
    static constexpr bool UREG = L0::UREG && L0::MMREG;
    static constexpr bool OK = true;
    // the SYM pass carries one "entry sits on a bound" bit per control entry in a 64-bit mask
    static constexpr bool SYM_OK = UREG && NI <= 64;
    static constexpr int SQ = L0::SQ;
    // slots per horizon step: the Q*M stash of the forward sweep, later (u_old, v_old) per control entry; even, so
    // that a step's slots are whole pairs
    static constexpr int W = ((imax_c(SQ, 2 * I) + 1) / 2) * 2;
    static constexpr int O_QM = 0;                       // pairs, element index always even
    static constexpr int O_V = O_QM + N * W;             // Nesterov v, single elements
    static constexpr int O_U = O_V + NI;                 // controls, when not in registers
    static constexpr int O_MM = O_U + (UREG ? 0 : NI);   // linear term MM: entries [0, MM_S) live here, the rest in registers
    // Register-resident horizons may still keep the first BMPC_PG_MM_SMEM entries of MM in shared memory — 23 of 30
    // take the (4,1,30) kernel from 255 registers with spills to 252 without (single-warp pass -6 %,
    // profiles/r1_experiments.md row 18) — but only where that costs no resident warp.
    static constexpr int MODEL_LEN = PIM ? L0::M_LEN : 0;
    static constexpr int MM_TRY = imin_c(BMPC_PG_MM_SMEM, NI);
    static constexpr int WARPS_PLAIN = imin_c(kSmemMax / ((O_MM + MODEL_LEN) * 32 * int(sizeof(T))), BMPC_PG_CAP_WARPS);
    static constexpr int WARPS_TRY = imin_c(kSmemMax / ((O_MM + MM_TRY + MODEL_LEN) * 32 * int(sizeof(T))), BMPC_PG_CAP_WARPS);
    static constexpr int MM_S = UREG ? (WARPS_TRY == WARPS_PLAIN ? MM_TRY : 0) : NI;
    static constexpr int O_MODEL = O_MM + MM_S;
    static constexpr int ELEMS = O_MODEL + (PIM ? L0::M_LEN : 0);
    static constexpr int WARP_BYTES = ELEMS * 32 * int(sizeof(T));
    static constexpr int FIT = kSmemMax / WARP_BYTES;
    static constexpr int WARPS = imax_c(1, imin_c(FIT, BMPC_PG_CAP_WARPS));
    static constexpr int THREADS = WARPS * 32;
    // One more 32-bit word per lane, after the warps' blocks, where it still fits: the instance id, which only the
    // finish / park code needs (one register less across the pass).
    static constexpr bool ID_SMEM = WARPS * WARP_BYTES + THREADS * 4 <= kSmemMax;
    static constexpr int SMEM_BYTES = WARPS * WARP_BYTES + (ID_SMEM ? THREADS * 4 : 0);
    static_assert(FIT >= 1, "horizon too long for the shared-memory resident layout");
    static constexpr int UNR = UREG ? N : 2;             // horizon loops: full unroll keeps u[] / MM[] in registers
    static constexpr int UNR_K = UREG ? NI : 2;
    __host__ __device__ static constexpr int qslot(int r) { return popc_c(PAT::QNZ & ((1u << r) - 1u)); }
};

template <typename T, int S, int I, int N, bool EXACT, bool PIM, typename PAT, bool SYM>
struct PgSolver {
    using L = PgLayout<T, S, I, N, PIM, PAT>;
    using L0 = typename L::L0;
    using P = SolveParams<T, S, I, N>;
    using A_ = Ar<T, EXACT>;
    using IO = SmemIO<T>;
    using Bits = typename BitsOf<T>::type;
    static constexpr int NI = L::NI;
    static constexpr int ES = int(sizeof(T));
    static constexpr Bits kAbs = BitsOf<T>::abs_mask;
    static_assert(!SYM || L::SYM_OK, "SYM needs a register-resident horizon of at most 64 control entries");

    const P& p;
    const T* smodel;      // PIM: this lane's column of the model block (element e at smodel[e*32])
    unsigned a1, a2;      // shared-window byte addresses of this lane's column: single elements / pairs
    T u[L::UREG ? NI : 1], mm[NI - L::MM_S > 0 ? NI - L::MM_S : 1];
    unsigned long long sat;   // SYM: bit k set <=> |u[k]| == h (the entry sits on one of its bounds)
    __device__ __forceinline__ T getU(int k) const { if constexpr (L::UREG) return u[k]; else return IO::ld1(elem_addr(L::O_U + k)); }
    __device__ __forceinline__ void setU(int k, T v) { if constexpr (L::UREG) u[k] = v; else IO::st1(elem_addr(L::O_U + k), v); }
    __device__ __forceinline__ T getMM(int k) const { if (k < L::MM_S) return IO::ld1(elem_addr(L::O_MM + k)); else return mm[k - L::MM_S]; }
    __device__ __forceinline__ void setMM(int k, T v) { if (k < L::MM_S) IO::st1(elem_addr(L::O_MM + k), v); else mm[k - L::MM_S] = v; }
    __device__ __forceinline__ T getV(int k) const { return IO::ld1(elem_addr(L::O_V + k)); }
    __device__ __forceinline__ void setV(int k, T v) { IO::st1(elem_addr(L::O_V + k), v); }

    __device__ __forceinline__ PgSolver(const P& p_, const T* smodel_, unsigned a1_, unsigned a2_) : p(p_), smodel(smodel_), a1(a1_), a2(a2_), sat(0) {}

    __device__ __forceinline__ T mA(int r, int k) const { if constexpr (PIM) return smodel[(L0::M_A + r * S + k) * 32]; else return p.A[r * S + k]; }
    __device__ __forceinline__ T mB(int r, int j) const { if constexpr (PIM) return smodel[(L0::M_B + r * I + j) * 32]; else return p.B[r * I + j]; }
    __device__ __forceinline__ T mQ(int r) const { if constexpr (PIM) return smodel[(L0::M_Q + r) * 32]; else return p.Q[r]; }
    __device__ __forceinline__ T mR(int j) const { if constexpr (PIM) return smodel[(L0::M_R + j) * 32]; else return p.R[j]; }
    __device__ __forceinline__ T mLo(int j) const { if constexpr (PIM) return smodel[(L0::M_LO + j) * 32]; else return p.lo[j]; }
    __device__ __forceinline__ T mHi(int j) const { if constexpr (PIM) return smodel[(L0::M_HI + j) * 32]; else return p.hi[j]; }

    // byte addresses: pair slot `e` (even element index) and single element `e`
    __device__ __forceinline__ unsigned pair_addr(int e) const { return a2 + (unsigned)(e * 32 * ES); }
    __device__ __forceinline__ unsigned elem_addr(int e) const { return a1 + (unsigned)(e * 32 * ES); }

    __host__ __device__ static constexpr bool bit(unsigned m, int e) { return ((m >> e) & 1u) != 0u; }
    __device__ __forceinline__ static void term(T& t, bool& have, bool nz, bool one, T coef, T z)
    {
        if (!nz) return;
        if (one) t = have ? A_::add(t, z) : z;
        else t = have ? A_::madd(t, coef, z) : A_::mul(coef, z);
        have = true;
    }
    __device__ __forceinline__ bool dotA(int r, const T* z, T& t) const
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::ANZ, r * S + k), bit(PAT::AONE, r * S + k), mA(r, k), z[k]);
        return have;
    }
    __device__ __forceinline__ bool dotAT(int r, const T* z, T& t) const
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::ANZ, k * S + r), bit(PAT::AONE, k * S + r), mA(k, r), z[k]);
        return have;
    }
    __device__ __forceinline__ bool dotB(int r, const T* uu, T& t) const
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < I; ++k) term(t, have, bit(PAT::BNZ, r * I + k), bit(PAT::BONE, r * I + k), mB(r, k), uu[k]);
        return have;
    }
    __device__ __forceinline__ bool dotBT(int j, const T* z, T& t) const
    {
        bool have = false; t = T(0);
#pragma unroll
        for (int k = 0; k < S; ++k) term(t, have, bit(PAT::BNZ, k * I + j), bit(PAT::BONE, k * I + j), mB(k, j), z[k]);
        return have;
    }
    __device__ __forceinline__ static T add_opt(bool ha, T a, bool hb, T b) { return ha ? (hb ? A_::add(a, b) : a) : (hb ? b : T(0)); }

    // SYM helpers.  h > 0 is the common magnitude of both bounds.
    // |x| < h; ICLAMP: on the bit patterns (non-negative IEEE values order like unsigned integers; x is never a NaN)
    // (the magnitude is taken on the high word by hand: written as `bits & abs_mask` nvcc turns it back into an FP64
    // |x| move, i.e. one more instruction on the pipe these kernels are bound by)
    __device__ __forceinline__ static Bits abs_bits(T x)
    {
        if constexpr (sizeof(T) == 8)
            return ((Bits)((unsigned)__double2hiint(x) & 0x7fffffffu) << 32) | (Bits)(unsigned)__double2loint(x);
        else return BitsOf<T>::to(x) & kAbs;
    }
    __device__ __forceinline__ static bool sign_differs(T a, T b)
    {
        if constexpr (sizeof(T) == 8) return (__double2hiint(a) ^ __double2hiint(b)) < 0;
        else return (int)(BitsOf<T>::to(a) ^ BitsOf<T>::to(b)) < 0;
    }
    __device__ __forceinline__ static bool inside(T x, T h)
    {
#if BMPC_PG_ICLAMP
        return abs_bits(x) < BitsOf<T>::to(h);
#else
        return absval(x) < h;
#endif
    }
    // dlib::clamp(x, -h, h) (matrix_utilities.h:2829-2846) for a non-NaN x: x itself strictly inside, else the bound
    // of x's sign — |x| == h returns the same value either way, so "not strictly inside" doubles as "on a bound".
    __device__ __forceinline__ static T clamp_sym(T x, T h, bool in)
    {
        if constexpr (sizeof(T) == 8) {
            const int xh = __double2hiint(x), hh = __double2hiint(h);
            return __hiloint2double(in ? xh : ((xh & 0x80000000) | hh), in ? __double2loint(x) : __double2loint(h));
        } else return in ? x : BitsOf<T>::from((BitsOf<T>::to(x) & ~kAbs) | BitsOf<T>::to(h));
    }
    // rebuilds the on-a-bound mask from the iterate (after a load)
    __device__ __forceinline__ void refresh_sat()
    {
        if constexpr (SYM) {
            sat = 0;
#pragma unroll
            for (int k = 0; k < NI; ++k) sat |= (unsigned long long)(!inside(u[k], mHi(k % I))) << k;
        }
    }

    // One projected-gradient iteration: gradient (mpc.h:275-283), stopping test (mpc.h:289-311) and the provisional
    // update (mpc.h:339-344).  Returns true when the reference would NOT break at mpc.h:310, i.e. the update stands.
    //
    // SYM (lower == -upper == -h, every value finite — the host and stage A check the preconditions, see
    // bmpc_solve.cu): the literal form spends 7 FP64 compares per control entry (|df| >= eps, u <= lower, u >= upper,
    // and two per clamp).  Here
    //   * u <= lower / u >= upper are not recomputed: u is the result of last pass's clamp, which already knows
    //     whether it returned a bound (mask `sat`); "the bound blocks df" (mpc.h:296-299) is then
    //     sat & (sign(u) != sign(df)), one integer op;
    //   * clamp(x, -h, h) is one test |x| < h and a sign transplant instead of two compares and two selects;
    //   * the remaining tests may run on the integer pipe (BMPC_PG_IBIG / BMPC_PG_ICLAMP).
    // Same values, same iteration counts: every replaced predicate is the same boolean function of the same numbers.
    __device__ __forceinline__ bool pass(T inv_lambda, T beta)
    {
        T m[S];
        {
            T u0v[I];
#pragma unroll
            for (int j = 0; j < I; ++j) u0v[j] = getU(j);
#pragma unroll
            for (int r = 0; r < S; ++r) { T tb; const bool hb = dotB(r, u0v, tb); m[r] = hb ? tb : T(0); }     // mpc.h:275
        }
        T w[S];
#pragma unroll(L::UNR)
        for (int i = 0; i < N; ++i) {
            if (i > 0) {
                T ui[I], mn[S];
#pragma unroll
                for (int j = 0; j < I; ++j) ui[j] = getU(i * I + j);
#pragma unroll
                for (int r = 0; r < S; ++r) {                                                                   // mpc.h:277
                    T t1; const bool h1 = dotA(r, m, t1);
                    T ad; const bool hb = dotB(r, ui, ad);
                    mn[r] = add_opt(h1, t1, hb, ad);
                }
#pragma unroll
                for (int r = 0; r < S; ++r) m[r] = mn[r];
            }
            T qs[L::W];
#pragma unroll
            for (int q = 0; q < L::W; ++q) qs[q] = T(0);
#pragma unroll
            for (int r = 0; r < S; ++r) {                                                                       // mpc.h:279
                if (bit(PAT::QNZ, r)) {
                    const T q = A_::mul(mQ(r), m[r]);
                    if (i < N - 1) qs[L::qslot(r)] = q; else w[r] = q;
                } else if (i == N - 1) w[r] = T(0);
            }
            if (i < N - 1) {
#pragma unroll
                for (int q = 0; q < (L::SQ + 1) / 2; ++q) IO::st2(pair_addr(L::O_QM + i * L::W + 2 * q), qs[2 * q], qs[2 * q + 1]);
            }
        }
        bool notconv = false;
        T lo_[I], hi_[I];
#pragma unroll
        for (int j = 0; j < I; ++j) { lo_[j] = mLo(j); hi_[j] = mHi(j); }
        const Bits eps_bits = BitsOf<T>::to(p.eps);
        unsigned long long nsat = 0;
        // The shared-memory accesses are volatile asm (they must really go through shared memory, solve_kernel.cuh), so
        // they issue in program order, and a warp issues in order: a store waiting for the end of an entry's update
        // chain (df -> v -> u, ~6 dependent operations) would hold back the stash load of the NEXT step's recursion, the
        // pass's critical path.  So every load is issued one horizon step ahead of its use and a step's stores one step
        // late, when their operands have long been ready (profiles/r2_experiments.md: lone-warp pass 3 050 -> ~2 300 cycles).
        T vo[I], mmv[I], pu[I], pvo[I], pvn[I];
#pragma unroll
        for (int j = 0; j < I; ++j) { vo[j] = getV((N - 1) * I + j); mmv[j] = getMM((N - 1) * I + j); pu[j] = pvo[j] = pvn[j] = T(0); }
#pragma unroll(L::UNR)
        for (int i = N - 1; i >= 0; --i) {
            // ---- loads of this step's recursion and of the next step's update
            T st[L::W];
#pragma unroll
            for (int q = 0; q < L::W; ++q) st[q] = T(0);
            if (i > 0) {
#pragma unroll
                for (int q = 0; q < (L::SQ + 1) / 2; ++q) IO::ld2(pair_addr(L::O_QM + (i - 1) * L::W + 2 * q), st[2 * q], st[2 * q + 1]);
            }
            T vof[I], mmf[I];
#pragma unroll
            for (int j = 0; j < I; ++j) { vof[j] = T(0); mmf[j] = T(0); if (i > 0) { vof[j] = getV((i - 1) * I + j); mmf[j] = getMM((i - 1) * I + j); } }
            // ---- stores of the previous step (i + 1): its stash slot was consumed a step ago
            if (i < N - 1) {
#pragma unroll
                for (int j = 0; j < I; ++j) { IO::st2(pair_addr(L::O_QM + (i + 1) * L::W + 2 * j), pu[j], pvo[j]); setV((i + 1) * I + j, pvn[j]); }
            }
#pragma unroll
            for (int j = I - 1; j >= 0; --j) {
                const int k = i * I + j;
                T bt; const bool hbt = dotBT(j, w, bt);
                const T uij = getU(k);
                const T dfv = A_::add(add_opt(true, mmv[j], hbt, bt), A_::mul(mR(j), uij));                     // mpc.h:283
                const T v_old = vo[j];
                T vn, un;
                if constexpr (SYM) {
#if BMPC_PG_IBIG
                    const bool big = abs_bits(dfv) >= eps_bits;                   // |df| >= eps, df is not a NaN here
#else
                    const bool big = absval(dfv) >= p.eps;
#endif
                    // mpc.h:296-299 for df != 0: the bound u sits on blocks df iff their signs differ
                    const bool opp = sign_differs(dfv, uij);
                    const bool blocked = ((sat >> k) & 1ull) != 0 && opp;
                    notconv |= big & !blocked;
                    // mpc.h:339-344 (v_old = v; v = clamp(u - 1/lambda*df); u = clamp(v + beta*(v - v_old)))
                    const T t1 = A_::sub(uij, A_::mul(inv_lambda, dfv));
                    vn = clamp_sym(t1, hi_[j], inside(t1, hi_[j]));
                    const T t2 = A_::madd(vn, beta, A_::sub(vn, v_old));
                    const bool in2 = inside(t2, hi_[j]);
                    un = clamp_sym(t2, hi_[j], in2);
                    nsat |= (unsigned long long)(!in2) << k;
                } else {
                    // mpc.h:289-311 reduced to its outcome: some free entry has |df| >= eps.  |df| >= eps > 0 implies
                    // df != 0 and df is not a NaN, so the sign bit tells which bound could block it (mpc.h:296-299).
                    const bool big = absval(dfv) >= p.eps;
                    const bool neg = (BitsOf<T>::to(dfv) & ~kAbs) != 0;
                    const bool at_hi = uij >= hi_[j], at_lo = uij <= lo_[j];
                    notconv |= big & !(neg ? at_hi : at_lo);
                    vn = clamp_m(A_::sub(uij, A_::mul(inv_lambda, dfv)), lo_[j], hi_[j]);
                    un = clamp_m(A_::madd(vn, beta, A_::sub(vn, v_old)), lo_[j], hi_[j]);
                }
                pu[j] = uij; pvo[j] = v_old; pvn[j] = vn;                 // stored during the next step
                setU(k, un);
            }
#pragma unroll
            for (int j = 0; j < I; ++j) { vo[j] = vof[j]; mmv[j] = mmf[j]; }
            if (i > 0) {
                T wn[S];
#pragma unroll
                for (int r = 0; r < S; ++r) {                                                                   // mpc.h:281
                    T t; const bool h = dotAT(r, w, t);
                    const bool hq = bit(PAT::QNZ, r);
                    wn[r] = add_opt(hq, hq ? st[L::qslot(r)] : T(0), h, t);
                }
#pragma unroll
                for (int r = 0; r < S; ++r) w[r] = wn[r];
            }
        }
#pragma unroll
        for (int j = 0; j < I; ++j) { IO::st2(pair_addr(L::O_QM + 2 * j), pu[j], pvo[j]); setV(j, pvn[j]); }   // step 0's stores
        if constexpr (SYM) sat = nsat;
        return notconv;
    }
};

// grid = #SMs (persistent), block = PgLayout::THREADS, dynamic smem = PgLayout::SMEM_BYTES.
//
// Work distribution.  p.pg_list is a ring of instance ids (capacity p.pg_mask + 1, empty slots hold kEmpty); the state
// of a parked instance (u, v, MM, iter) lives BY RING SLOT in p.park ([3*N*I][ring]) / p.park_iter, so that the 32
// lanes of a warp, which always hold consecutive slots, park and resume with fully coalesced accesses.
// queue[1] = tail (producers: stage A, and warps of this kernel that park), queue[2] = head (consumers take tickets
// with one aggregated atomicAdd per warp; a ticket beyond the tail is a reservation for an instance parked later),
// queue[3] = instances not finished yet.  A producer publishes the state with a fence BEFORE the ring entry; a
// consumer reads the entry, fences, and reads the state with L2 loads (ld.cg).
//
// Scheduling (p.pg_reuse; arithmetic is untouched — an instance simply continues in another lane):
//   * fairness slices (p.pg_slice, host decision for batches of 1-4 instances per lane): while more instances than
//     lanes are unfinished, every warp parks its live lanes every BMPC_PG_SLICE passes and resumes the oldest parked
//     ones, so that all instances advance at the same rate.  Iteration counts are heavy-tailed and unpredictable
//     from the state at iter 50; with run-to-completion a long solve claimed late finishes long after the rest.
//   * retirement: once fewer instances are left than the lanes of `t` warps per CTA, warps with index >= t park their
//     lanes for good (the freed lanes of the lower warps hold reservations and pick them up).  A warp alone on its
//     SM sub-partition runs a pass ~1.4x faster than two sharing it, and the end of every launch is its longest solve.
template <typename T, int S, int I, int N, bool EXACT, bool PIM, typename PAT, bool SYM>
__global__ void __launch_bounds__((PgLayout<T, S, I, N, PIM, PAT>::THREADS), 1)
pg_kernel(const __grid_constant__ SolveParams<T, S, I, N> p)
{
    using L = PgLayout<T, S, I, N, PIM, PAT>;
    using L0 = typename L::L0;
    using A_ = Ar<T, EXACT>;
    using SV = PgSolver<T, S, I, N, EXACT, PIM, PAT, SYM>;
    constexpr int NI = L::NI;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp >= p.warp_limit) return;          // no block-wide barrier anywhere in this kernel
    T* wbase = reinterpret_cast<T*>(smem_raw) + (size_t)warp * L::ELEMS * 32;
    const unsigned wb = (unsigned)__cvta_generic_to_shared(wbase);
    T* smodel = wbase + L::O_MODEL * 32 + lane;
    SV sv(p, smodel, wb + (unsigned)lane * sizeof(T), wb + (unsigned)lane * 2u * sizeof(T));

    constexpr uint32_t kEmpty = 0xffffffffu;
    constexpr int kSpinLimit = 1 << 22;
    volatile unsigned long long* const q = p.queue;
    volatile uint32_t* const ring = p.pg_list;
    const unsigned mask = p.pg_mask;
    // by-slot state rows: row k of u at park + k*rstride, of v at park + (NI+k)*rstride, of MM at park + (2NI+k)*rstride
    // (addresses are re-derived from the parameter block where they are used: no registers held across the pass)
    const unsigned total0 = (unsigned)q[1];                   // entries stage A handed over (instance ids are 32-bit)
    const bool reuse = p.pg_reuse != 0;                       // ring slots are recycled: parking is possible
    const bool retire = (p.pg_reuse & 2) != 0;                // surplus warps may retire near the end of the launch
    // the instance id of the lane: in the spare shared-memory word where there is one, else a register
    uint32_t inst_reg = 0;
    volatile uint32_t* const inst_sm = reinterpret_cast<volatile uint32_t*>(smem_raw + (size_t)L::WARPS * L::WARP_BYTES) + threadIdx.x;
    auto get_inst = [&]() -> uint32_t { if constexpr (L::ID_SMEM) return *inst_sm; else return inst_reg; };
    auto set_inst = [&](uint32_t v) { if constexpr (L::ID_SMEM) *inst_sm = v; else inst_reg = v; };
    // Per-lane state machine: kNeed (no work, no ticket) -> kWait (holds the ticket of ring slot `aux`) -> kActive
    // (runs the instance, `aux` is its iteration counter) -> kNeed ...;  kDone: a lane that can get no more work.
    enum : int { kNeed = 0, kWait = 1, kActive = 2, kDone = 3 };
    int st = kNeed;
    uint32_t aux = 0;
    int since = 0;           // passes since this warp's last fairness park (warp-uniform)
    bool eligible = true;    // this warp still takes new tickets (warp-uniform; false once retired)
    T lambda = p.lambda, inv_lambda = T(0), beta = T(0);
    if constexpr (!PIM) {
        inv_lambda = A_::div(T(1), lambda);                                                   // mpc.h:342
        const T sq = A_::sqrt(lambda);
        beta = A_::div(A_::sub(sq, T(1)), A_::add(sq, T(1)));                                // mpc.h:343
    }
#pragma unroll(L::UNR_K)
    for (int k = 0; k < NI; ++k) { sv.setU(k, T(0)); sv.setMM(k, T(0)); }

    // The loop alternates a SLOW section (tickets, ring poll, resume, results, parking decisions — everything that
    // touches global memory) with a tight inner loop of passes that ends as soon as a lane finishes or after
    // BMPC_PG_CHECK passes: the inner loop carries nothing but the solver state and one counter.
    while (true) {
        // ---------------- tickets: one aggregated atomicAdd per warp, never fails; a retired warp takes none
        const unsigned needmask = __ballot_sync(FULL, st == kNeed && eligible);
        if (needmask) {
            const int leader = __ffs(needmask) - 1;
            unsigned long long h = 0;
            if (lane == leader) h = atomicAdd(p.queue + 2, (unsigned long long)__popc(needmask));
            h = __shfl_sync(FULL, h, leader);
            if (st == kNeed && eligible) {
                const unsigned long long t = h + (unsigned long long)__popc(needmask & ((1u << lane) - 1u));
                aux = (unsigned)t & mask;
                // without slot reuse nothing is ever added behind stage A's entries: a ticket beyond them ends the lane
                st = (reuse || t < (unsigned long long)total0) ? kWait : kDone;
            }
        }
        // ---------------- poll the ticket's ring slot; resume the parked instance found there
        if (st == kWait) {
            const uint32_t e = ring[aux];
            if (e != kEmpty) {
                __threadfence();                               // the entry was published after the state
                const unsigned slot = aux;
                const uint32_t inst = e;
                set_inst(inst); st = kActive;
                aux = __ldcg(p.park_iter + slot);
                const long long rstride = (long long)mask + 1;
                const T* src = p.park + slot;
#pragma unroll(L::UNR_K)
                for (int k = 0; k < NI; ++k) {
                    sv.setU(k, __ldcg(src + (long long)k * rstride));
                    sv.setV(k, __ldcg(src + (long long)(NI + k) * rstride));
                    sv.setMM(k, __ldcg(src + (long long)(2 * NI + k) * rstride));
                }
                if constexpr (PIM) {
#pragma unroll
                    for (int e2 = 0; e2 < S * S; ++e2) smodel[(L0::M_A + e2) * 32] = p.mA[(long long)e2 * p.stride + inst];
#pragma unroll
                    for (int e2 = 0; e2 < S * I; ++e2) smodel[(L0::M_B + e2) * 32] = p.mB[(long long)e2 * p.stride + inst];
#pragma unroll
                    for (int e2 = 0; e2 < S; ++e2) smodel[(L0::M_Q + e2) * 32] = p.mQ[(long long)e2 * p.stride + inst];
#pragma unroll
                    for (int e2 = 0; e2 < I; ++e2) {
                        smodel[(L0::M_R + e2) * 32] = p.mR[(long long)e2 * p.stride + inst];
                        smodel[(L0::M_LO + e2) * 32] = p.mlo[(long long)e2 * p.stride + inst];
                        smodel[(L0::M_HI + e2) * 32] = p.mhi[(long long)e2 * p.stride + inst];
                    }
                    lambda = p.mlambda[inst];
                    inv_lambda = A_::div(T(1), lambda);
                    const T sq = A_::sqrt(lambda);
                    beta = A_::div(A_::sub(sq, T(1)), A_::add(sq, T(1)));
                }
                sv.refresh_sat();
                if (reuse) { __threadfence(); ring[slot] = kEmpty; }   // the slot and its state rows are free again
            }
        }
        // ---------------- nothing to run: finished, or wait for parked instances
        if (!__any_sync(FULL, st == kActive)) {
            if (!__any_sync(FULL, st == kWait)) break;             // no ticket of this warp is still open
            if (q[3] == 0) break;                                  // every instance of the launch has finished
            if (p.c_threshold != 0 && q[3] == q[6]) break;         // ... or waits in stage C's list (nothing is active or parked here)
            if (q[4] != 0) break;                                  // another warp gave up (ring overrun)
            __nanosleep(500); since = 0;
            continue;
        }

        // ---------------- passes, until a lane finishes or BMPC_PG_CHECK of them are done
        bool fin = false, hit_max = false;
        int budget = BMPC_PG_CHECK;
        do {
            const bool notconv = sv.pass(inv_lambda, beta);
            if (st == kActive) {
                if (!notconv) fin = true;                                              // mpc.h:310-311, update discarded
                else if (++aux >= p.max_iter) { fin = true; hit_max = true; }          // mpc.h:271
            }
            --budget;
        } while (budget > 0 && !__any_sync(FULL, fin));
        since += BMPC_PG_CHECK - budget;

        // ---------------- results of the lanes that finished
        if (fin) {
            const uint32_t inst = get_inst();
            p.iters[inst] = aux;
            p.status[inst] = (hit_max ? 1 : 0) | (p.status[inst] & 2);   // stage A's BMPC_ST_NONFINITE stays
            const bool wc = p.warm || p.keep_controls;
            if (hit_max) {
#pragma unroll(L::UNR_K)
                for (int k = 0; k < NI; ++k) {
                    const T uk = sv.getU(k);
                    if (k < I) p.u0[(long long)k * p.stride + inst] = uk;                             // mpc.h:239
                    if (wc) p.controls[(long long)k * p.stride + inst] = uk;
                    if (p.warm) p.vstate[(long long)k * p.stride + inst] = sv.getV(k);
                }
            } else {
                // converged: the iterate and v the test was taken on are parked in the stash slots
#pragma unroll(L::UNR_K)
                for (int k = 0; k < NI; ++k) {
                    T uo, vo;
                    SV::IO::ld2(sv.pair_addr(L::O_QM + (k / I) * L::W + 2 * (k % I)), uo, vo);
                    if (k < I) p.u0[(long long)k * p.stride + inst] = uo;
                    if (wc) p.controls[(long long)k * p.stride + inst] = uo;
                    if (p.warm) p.vstate[(long long)k * p.stride + inst] = vo;
                }
            }
            if (p.warm && p.target_full) {                                          // mpc.h:236-237 on the ring
                int last = p.head + N - 1; if (last >= N) last -= N;
#pragma unroll
                for (int r = 0; r < S; ++r)
                    p.target[((long long)p.head * S + r) * p.stride + inst] =
                        p.target[((long long)last * S + r) * p.stride + inst];
            }
            st = kNeed;
        }
        const unsigned finmask = __ballot_sync(FULL, fin);
        if (finmask && lane == (__ffs(finmask) - 1)) atomicAdd(p.queue + 3, 0ull - (unsigned long long)__popc(finmask));

        // ---------------- hand-over to stage C (pair_kernel.cuh): once p.c_threshold or fewer instances are unfinished, the
        // launch is bound by the latency of its longest solves, which the lane-pair kernel halves.  Every warp parks its
        // live instances in stage C's list (by-slot rows like the ring's, tail = queue[6]); instances that open tickets
        // still bring are forwarded the same way after their next passes.  `unfinished` only falls, so the decision
        // is taken once and for all by every warp on its own.
        if (p.c_threshold != 0 && (unsigned)q[3] <= p.c_threshold) {
            const unsigned actmask = __ballot_sync(FULL, st == kActive);
            if (actmask) {
                const int leader = __ffs(actmask) - 1;
                unsigned long long pos = 0;
                if (lane == leader) pos = atomicAdd(p.queue + 6, (unsigned long long)__popc(actmask));
                pos = __shfl_sync(FULL, pos, leader);
                if (st == kActive) {
                    const unsigned ps = (unsigned)pos + (unsigned)__popc(actmask & ((1u << lane) - 1u));
                    T* dst = p.c_park + ps;
                    const long long cs = (long long)p.c_stride;
#pragma unroll(L::UNR_K)
                    for (int k = 0; k < NI; ++k) {
                        dst[(long long)k * cs] = sv.getU(k);
                        dst[(long long)(NI + k) * cs] = sv.getV(k);
                        dst[(long long)(2 * NI + k) * cs] = sv.getMM(k);
                    }
                    p.c_iter[ps] = aux;
                    p.c_list[ps] = get_inst();
                    st = kNeed;
                }
            }
            continue;
        }

        // ---------------- scheduling decisions (see the kernel comment)
        if (reuse) {
            bool park = !eligible;                              // a retired warp forwards what its open tickets still bring
            if ((retire && eligible) || (p.pg_slice != 0 && since >= BMPC_PG_SLICE)) {
                const unsigned unfinished = (unsigned)q[3];    // (one request per warp: the lanes agree)
                if (retire && eligible) {
                    const unsigned per = gridDim.x * 32u;
                    unsigned want = (unfinished + per - 1u) / per;
                    const unsigned min_warps = p.warp_limit < 4 ? p.warp_limit : 4;   // one warp per SM sub-partition is the floor
                    if (want < min_warps) want = min_warps;
                    if ((unsigned)warp >= want) { eligible = false; park = true; }    // `unfinished` only falls: for good
                }
                if (p.pg_slice != 0 && since >= BMPC_PG_SLICE) {
                    since = 0;
                    if (unfinished > gridDim.x * (unsigned)p.warp_limit * 32u) park = true;
                }
            }
            if (park) {
                const unsigned actmask = __ballot_sync(FULL, st == kActive);
                if (actmask) {
                    const int leader = __ffs(actmask) - 1;
                    unsigned long long pos = 0;
                    if (lane == leader) pos = atomicAdd(p.queue + 1, (unsigned long long)__popc(actmask));
                    pos = __shfl_sync(FULL, pos, leader);
                    bool overrun = false;
                    if (st == kActive) {
                        const unsigned ps = (unsigned)(pos + (unsigned long long)__popc(actmask & ((1u << lane) - 1u))) & mask;
                        int spin = 0;
                        while (ring[ps] != kEmpty) if (++spin > kSpinLimit) { overrun = true; break; }
                        if (!overrun) {
                            const long long rstride = (long long)mask + 1;
                            T* dst = p.park + ps;
#pragma unroll(L::UNR_K)
                            for (int k = 0; k < NI; ++k) {
                                dst[(long long)k * rstride] = sv.getU(k);
                                dst[(long long)(NI + k) * rstride] = sv.getV(k);
                                dst[(long long)(2 * NI + k) * rstride] = sv.getMM(k);
                            }
                            p.park_iter[ps] = aux;
                            __threadfence();
                            ring[ps] = get_inst();
                            st = kNeed;
                        }
                    }
                    if (__any_sync(FULL, overrun)) {                // cannot happen with a ring >= instances + lanes
                        if (overrun) atomicExch(p.queue + 4, 1ull);
                        break;
                    }
                }
            }
        }
    }
}

}  // namespace bmpc
