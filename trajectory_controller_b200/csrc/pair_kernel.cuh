// Stage C of the split solve: the projected-gradient phase (mpc.h:271-345 for iter >= 50) with ONE INSTANCE PER LANE
// PAIR, for the block-triangular (4,1) "bicycle4" pattern (patterns.h).
//
// Why.  The per-lane kernel (pg_kernel.cuh) gives every lane a whole instance: 1 086 FP64-pipe instructions per pass and
// warp, i.e. >= 2 172 issue cycles at one FP64 warp instruction per two cycles, whatever the number of live lanes — a
// lone solve (the metric's "p50 solve latency"), a small batch, and the tail of every launch (its longest solve) all
// run at ~3 050 cycles per iteration against a dependency chain of ~1 550.  Splitting an instance over lanes shortens
// the instruction stream, but dlib's rounding order makes every row of a mat-vec one sequential dot product and the
// two horizon recursions sequential in i, so a split pays only where it adds no cross-lane round trip per horizon step.
// The bicycle4 model has exactly that structure: A is block upper triangular,
//       A = [ 1 a01  .   . ]      B = (. . . b3)'     Q = (Q0 Q1 . .)
//           [ .  1  a12  . ]
//           [ .  .   1  a23]      rows 0-1 = path errors (e_y, e_psi), rows 2-3 = steering actuator (delta, delta')
//           [ .  .  a32 a33]
// so in the FORWARD recursion (mpc.h:275-277) rows 2-3 never read rows 0-1, and in the BACKWARD recursion (mpc.h:281,
// trans(A)) rows 0-1 never read rows 2-3.  Lane L of a pair owns rows 2-3, lane H rows 0-1; per horizon step ONE value
// crosses, one way, and it is consumed one step later (m2 from L to H going forward, w1 from H to L going backward):
// the exchange shuffle is off the critical path.  The same shuffle carries, in the other direction, what the partner
// needs of the control entries (below), so one 64-bit shuffle per horizon step serves both lanes.
//   * the N control entries (u, the linear term MM, Nesterov's v) are split by parity: L owns the even entries, H the
//     odd ones, and the stopping test + provisional update (mpc.h:289-311, 339-344) of the entries (2j, 2j+1) run in
//     the same instructions on both lanes — half the registers per lane (N = 60 becomes register-resident, N = 100
//     keeps only MM in shared memory) and half the update instructions per instance;
//   * both lanes execute ONE instruction stream: the two rows of a lane are written in a common form whose per-lane
//     coefficient registers (set up once per instance) make it the lane's own rows.  Where the common form multiplies by
//     an exact 1 or adds an exact 0 the value is unchanged (IEEE: 1*x == x, x + 0 == x up to the sign of a zero result,
//     the freedom the engine's contract already has); every other product and sum is the one dlib
//     evaluates, in dlib's order, so EXACT parity holds and FAST mode contracts exactly the pairs pg_kernel contracts.
// Per pass and warp: ~710 FP64-pipe instructions for 16 instances (vs 1 086 for 32), a dependency chain of ~24 cycles
// per horizon step in both sweeps.  Per instance-pass that is 30 % more issue slots than the per-lane kernel, so
// stage C is used where latency, not throughput, bounds the launch: small batches, the tail of large ones (stage B
// hands over its last instances), per-instance models and long horizons (where the per-lane layout ran from shared
// memory at 1-3 warps per SM).
//
// Requires symmetric bounds (lower == -upper, finite tame model: the SYM preconditions of pg_kernel.cuh, checked by
// bmpc_set_model / stage A) and an even horizon.
#pragma once
#include "pg_kernel.cuh"

#ifndef BMPC_PAIR_CAP_WARPS
#define BMPC_PAIR_CAP_WARPS 8
#endif
// Horizons from this length on run the pass as a ROLLED loop (two trips per iteration) with u / MM in shared memory:
// the fully unrolled pass of a long horizon no longer fits the instruction cache (profiles/r2_experiments.md).
#ifndef BMPC_PAIR_ROLL_MIN
#define BMPC_PAIR_ROLL_MIN 1000
#endif

namespace bmpc {

template <typename T, int N, bool PIM>
struct PairLayout {
    static_assert(N % 2 == 0, "the lane-pair kernel splits the control entries by parity: even horizons only");
    static constexpr int NH = N / 2;                       // control entries per lane
    static constexpr int WORDS = int(sizeof(T) / 4);
    // u (NH entries) always lives in registers; MM too while both stay within 128 32-bit registers
    static constexpr bool ROLLED = (N >= BMPC_PAIR_ROLL_MIN);
    static constexpr bool MMREG = !ROLLED && (2 * NH * WORDS <= 128);
    static constexpr int UNR = ROLLED ? 2 : N + 2;         // horizon loops: full unroll keeps u[] / MM[] in registers
    // shared memory per lane, in elements: N pair slots (H: the Q*M stash of the forward sweep, slot s = horizon step s;
    // both lanes: (u_old, v_old) of the lane's own entry e in slot e once the backward sweep has consumed it), v, MM
    static constexpr int O_Q = 0;                          // slot s lives at pair index s + 1: one pad slot in front
    static constexpr int O_V = 2 * (N + 1);
    static constexpr int O_MM = O_V + NH;
    static constexpr int O_U = O_MM + (MMREG ? 0 : NH);
    static constexpr int ELEMS = O_U + (ROLLED ? NH : 0);
    static constexpr int WARP_BYTES = ELEMS * 32 * int(sizeof(T));
    static constexpr int FIT = kSmemMax / WARP_BYTES;
    static constexpr int WARPS = imax_c(1, imin_c(FIT, BMPC_PAIR_CAP_WARPS));
    static constexpr int THREADS = WARPS * 32;
    static constexpr int SMEM_BYTES = WARPS * WARP_BYTES;
    static_assert(FIT >= 1, "horizon too long for the lane-pair layout");
};

template <typename T, int N, bool EXACT, bool PIM>
struct PairSolver {
    using L = PairLayout<T, N, PIM>;
    using A_ = Ar<T, EXACT>;
    using IO = SmemIO<T>;
    using Bits = typename BitsOf<T>::type;
    static constexpr int NH = L::NH;
    static constexpr int ES = int(sizeof(T));

    // per-lane coefficients of the common row form (see pass())
    T cP, c1, c2, c3, qP, qQ;        // forward
    T h1, mc, g1, g2;                // backward
    T b3, R, hi, inv_lambda, beta, eps;
    bool isH;
    unsigned a1, a2, a2own, a2lag;   // shared-window byte addresses: single elements / pair slots / slot + role (own entries) / slot + (L ? 1 : 0)
    T u[L::ROLLED ? 1 : NH];
    T mm[L::MMREG ? NH : 1];
    unsigned long long sat;          // bit j set <=> |u[j]| == hi (entry sits on a bound), own entries

    __device__ __forceinline__ unsigned pair_addr(int slot) const { return a2 + (unsigned)(2 * (slot + 1) * 32 * ES); }
    __device__ __forceinline__ unsigned own_addr(int j) const { return a2own + (unsigned)(2 * (2 * j + 1) * 32 * ES); }   // slot 2j + role
    __device__ __forceinline__ unsigned lag_addr(int slot) const { return a2lag + (unsigned)(2 * (slot + 1) * 32 * ES); }  // slot + 1 on L
    __device__ __forceinline__ unsigned elem_addr(int e) const { return a1 + (unsigned)(e * 32 * ES); }
    __device__ __forceinline__ T getV(int j) const { return IO::ld1(elem_addr(L::O_V + j)); }
    __device__ __forceinline__ void setV(int j, T v) { IO::st1(elem_addr(L::O_V + j), v); }
    __device__ __forceinline__ T getU(int j) const { if constexpr (L::ROLLED) return IO::ld1(elem_addr(L::O_U + j)); else return u[j]; }
    __device__ __forceinline__ void setU(int j, T v) { if constexpr (L::ROLLED) IO::st1(elem_addr(L::O_U + j), v); else u[j] = v; }
    __device__ __forceinline__ T getMM(int j) const { if constexpr (L::MMREG) return mm[j]; else return IO::ld1(elem_addr(L::O_MM + j)); }
    __device__ __forceinline__ void setMM(int j, T v) { if constexpr (L::MMREG) mm[j] = v; else IO::st1(elem_addr(L::O_MM + j), v); }

    // Model of this instance -> the lane's coefficient registers.  A row-major [16], B [4], Q [4].
    __device__ __forceinline__ void set_model(bool high, T a01, T a12, T a23, T a32, T a33, T b3_, T q0, T q1, T r, T hi_, T lambda, T eps_)
    {
        isH = high;
        // forward (mpc.h:277), state (p, q) = (m2, m3) on L, (m0, m1) on H:
        //   p' = p + cP*q                               L: m2 + a23*m3            H: m0 + a01*m1
        //   q' = (c1*p + c2*q) + c3*Y                   L: (a32*m2 + a33*m3) + b3*u_i   H: (0*m0 + 1*m1) + a12*m2
        cP = high ? a01 : a23;
        c1 = high ? T(0) : a32; c2 = high ? T(1) : a33; c3 = high ? a12 : b3_;
        qP = high ? q0 : T(0); qQ = high ? q1 : T(0);                       // mpc.h:279, H only (L's products are unused zeros)
        // backward (mpc.h:281), state (p, q) = (w2, w3) on L, (w0, w1) on H, X = the partner's q of the previous step:
        //   p' = ((h1*X + p) + stash0) + mc*q           L: ((a12*w1 + w2) + 0) + a32*w3   H: ((0*w3 + w0) + stash0) + 0*w1
        //   q' = stash1 + (g1*p + g2*q)                 L: 0 + (a23*w2 + a33*w3)          H: stash1 + (a01*w0 + 1*w1)
        // (L's pair slots hold the zeros its forward sweep stored)
        h1 = high ? T(0) : a12; mc = high ? T(0) : a32;
        g1 = high ? a01 : a23; g2 = high ? T(1) : a33;
        b3 = b3_; R = r; hi = hi_; eps = eps_;
        inv_lambda = A_::div(T(1), lambda);                                                      // mpc.h:342
        const T sq = A_::sqrt(lambda);
        beta = A_::div(A_::sub(sq, T(1)), A_::add(sq, T(1)));                                    // mpc.h:343
    }

    __device__ __forceinline__ static T xchg(T v) { return __shfl_xor_sync(0xffffffffu, v, 1); }
    using PS = PgSolver<T, 4, 1, 2, EXACT, false, DensePattern<4, 1>, false>;   // only for its static SYM helpers

    __device__ __forceinline__ void refresh_sat()
    {
        sat = 0;
#pragma unroll
        for (int j = 0; j < NH; ++j) sat |= (unsigned long long)(!PS::inside(getU(j), hi)) << j;
    }

    // One projected-gradient iteration of the pair's instance.  Returns true when the reference would NOT break at
    // mpc.h:310 (the update stands); the value is the same on both lanes.
    //
    // Schedule.  A warp issues in order, so a shuffle whose result is consumed in the next horizon step would stall the
    // whole pair for its latency.  The lane that CONSUMES the crossing value therefore runs one horizon step behind the
    // lane that produces it: every exchange issued at the end of trip j is first read in trip j+2.
    //   forward  trip k = 0..N  : L computes step k (k < N), H computes step k-1 (k >= 1) from L's m2 of step k-2
    //   backward trip k = 0..N+1: H computes step N-1-k (k < N), L computes step N-k (1 <= k <= N) from H's w1 of step
    //                             N-k+1; at odd k >= 3 both lanes test and update one control entry: H its entry N+2-k
    //                             (w3 received two trips ago), L its entry N+1-k (w3 = its own q of the trip before)
    __device__ __forceinline__ bool pass()
    {
        // ---------------- forward sweep: M[i] = A*M[i-1] + B*u[i] (mpc.h:275-277), Q-scaling and stash (mpc.h:279)
        const T u_first = getU(0);
        T r_old = xchg(isH ? u_first : T(0));                         // L: control entry 1;  H: "m2 of step -1" = 0
        T pcur = T(0);
        T qcur = A_::mul(c3, isH ? T(0) : u_first);                      // trip 0, L: M[0] = B*u[0] (m3 = b3*u0);  H: zeros
        T r_new = xchg(pcur);                                         // H: m2 of step 0;  (L: unused)
        T wp = T(0), wq = T(0);
        T ucur = T(0);                                                // own control entry (k+1)/2, fetched at the start of odd trip k
#pragma unroll(L::UNR)
        for (int k = 1; k <= N; ++k) {
            if ((k & 1) && k <= N - 3) ucur = getU((k + 1) / 2);      // used by the exchange of this trip (H) and as Y of the next (L)
            // Y: L reads control entry k (own if k is even, received if odd); H reads L's m2 of step k-2
            T Y;
            if (k == N) Y = isH ? r_old : T(0);                       // L has no step N: its numbers of this trip are unused
            else if (k & 1) Y = r_old;
            else Y = isH ? r_old : ucur;
            const T pn = A_::madd(pcur, cP, qcur);
            const T t12 = A_::madd(A_::mul(c1, pcur), c2, qcur);
            T qn;
            if constexpr (EXACT) qn = A_::add(t12, A_::mul(c3, Y));
            else qn = isH ? A_::madd(qcur, c3, Y) : A_::add(t12, A_::mul(c3, Y));   // H: m1 + a12*m2 is one contracted term in FAST mode
            pcur = pn; qcur = qn;
            // diagm(Q)*M[k-1], rows 0-1, on H; zeros on L (its pair slots must read as zero in the backward sweep)
            const T sp = A_::mul(qP, pcur), sq_ = A_::mul(qQ, qcur);
            IO::st2(pair_addr(k - 1), sp, sq_);
            if (k == N) { wp = sp; wq = sq_; }                        // the backward sweep starts from the scaled last step
            r_old = r_new;
            if (k <= N - 2) {
                // end of trip k: H sends the control entry of L's step k+2 (needed when k is odd), L sends m2 of step k
                if (k & 1) r_new = xchg(isH ? ucur : pcur);
                else r_new = xchg(pcur);
            }
        }
        // ---------------- backward sweep: M[i-1] += trans(A)*M[i] (mpc.h:281), gradient (mpc.h:283), stopping test
        // (mpc.h:289-311) and provisional update (mpc.h:339-344)
        pcur = wp; qcur = wq;                                         // trip 0: H = step N-1 (Q0*m0, Q1*m1);  L = (0, 0)
        r_old = T(0);                                                 // L's trip 1 keeps its zeros
        r_new = xchg(qcur);                                           // end of trip 0: L gets w1 of step N-1, H gets w3 of step N-1 (0)
        bool notconv = false;
        unsigned long long nsat = 0;
        // The shared-memory accesses are volatile (they must really go through shared memory, solve_kernel.cuh), so
        // they issue in program order: a store waiting for the end of an update's dependency chain would hold back
        // every later load.  Hence loads are issued one trip (stash) / one update (v) ahead of their use and an
        // update's stores one update late, when their operands have long been ready.
        T s0, s1;
        IO::ld2(lag_addr(N - 2), s0, s1);                             // stash of trip 1
        T vo = getV(NH - 1);                                          // v of the first update (trip 3)
        T mmv = getMM(NH - 1);
        T uo_ = getU(NH - 1);
        T pend_u = T(0), pend_vold = T(0), pend_vn = T(0), pend_un = T(0);   // the previous update's stores
#pragma unroll(L::UNR)
        for (int k = 1; k <= N + 1; ++k) {
            const T X = r_old;                                        // L: w1 of step N-k+1;  H: w3 of step N-k+2
            T s0f = T(0), s1f = T(0);
            if (k + 1 <= N) IO::ld2(lag_addr(N - 2 - k), s0f, s1f);   // stash of trip k+1
            const bool upd = (k & 1) && k >= 3;
            const int j = (N + 1 - k) / 2;                            // L: entry N+1-k = 2j, H: entry N+2-k = 2j+1
            T vof = T(0), mmf = T(0), uf = T(0);
            if (upd && j >= 1) { vof = getV(j - 1); mmf = getMM(j - 1); uf = getU(j - 1); }           // for the update two trips on
            if (upd && j + 1 <= NH - 1) {                                                             // ... and of two trips ago
                IO::st2(own_addr(j + 1), pend_u, pend_vold); setV(j + 1, pend_vn);
                if constexpr (L::ROLLED) setU(j + 1, pend_un);
            }
            if (upd) {
                const T w3 = isH ? X : qcur;
                const T uij = uo_;
                const T dfv = A_::add(A_::add(mmv, A_::mul(b3, w3)), A_::mul(R, uij));               // mpc.h:283
                const bool big = absval(dfv) >= eps;
                const bool blocked = ((sat >> j) & 1ull) != 0 && PS::sign_differs(dfv, uij);           // mpc.h:296-299
                notconv |= big & !blocked;
                const T t1 = A_::sub(uij, A_::mul(inv_lambda, dfv));                                   // mpc.h:341
                const T vn = PS::clamp_sym(t1, hi, PS::inside(t1, hi));
                const T t2 = A_::madd(vn, beta, A_::sub(vn, vo));                                      // mpc.h:344
                const bool in2 = PS::inside(t2, hi);
                const T un = PS::clamp_sym(t2, hi, in2);
                if constexpr (!L::ROLLED) setU(j, un);                 // registers; the rolled form stores it with the next update
                nsat |= (unsigned long long)(!in2) << j;
                pend_u = uij; pend_vold = vo; pend_vn = vn; pend_un = un;   // the slot of this entry's step has been consumed: stored next update
                vo = vof; mmv = mmf; uo_ = uf;
            }
            if (k <= N) {
                // H: step N-1-k from slot N-1-k;  L: step N-k, whose slot holds the zeros the forward sweep stored
                // (a12*X + p) + s0 + mc*q  and  s1 + (g1*p + g2*q): on H the first term is 0*w3 + w0 and mc = 0,
                // on L both stash words are zero
                const T tt = A_::add(A_::mul(h1, X), pcur);
                const T pn = A_::madd(A_::add(tt, s0), mc, qcur);
                const T cq = A_::madd(A_::mul(g1, pcur), g2, qcur);
                const T qn = A_::add(s1, cq);
                pcur = pn; qcur = qn;
            }
            s0 = s0f; s1 = s1f;
            r_old = r_new;
            if (k <= N - 1) r_new = xchg(qcur);                       // end of trip k: H's w1 of step N-1-k, L's w3 of step N-k
        }
        IO::st2(own_addr(0), pend_u, pend_vold); setV(0, pend_vn);    // the last update's stores
        if constexpr (L::ROLLED) setU(0, pend_un);
        sat = nsat;
        return __shfl_xor_sync(0xffffffffu, (int)notconv, 1) != 0 || notconv;
    }
};

// grid = #SMs (persistent), block = PairLayout::THREADS, dynamic smem = PairLayout::SMEM_BYTES.
// Work: a list of parked instances, complete before the launch (stage A's hand-over ring when stage B is skipped,
// p.c_direct; else the list stage B parks its last instances in, p.c_list).  Pairs take tickets (one aggregated
// atomicAdd per warp on queue[7]) and run every instance to completion.
template <typename T, int N, bool EXACT, bool PIM>
__global__ void __launch_bounds__((PairLayout<T, N, PIM>::THREADS), 1)
pair_kernel(const __grid_constant__ SolveParams<T, 4, 1, N> p)
{
    using L = PairLayout<T, N, PIM>;
    using SV = PairSolver<T, N, EXACT, PIM>;
    constexpr int NH = L::NH;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int role = lane & 1;                                        // 0: L (rows 2-3, even entries), 1: H (rows 0-1, odd entries)
    // the list and the by-slot state rows this launch consumes
    const uint32_t* const list = p.c_direct ? p.pg_list : p.c_list;
    const T* const park = p.c_direct ? p.park : p.c_park;
    const uint32_t* const park_iter = p.c_direct ? p.park_iter : p.c_iter;
    const long long rstride = p.c_direct ? (long long)p.pg_mask + 1 : (long long)p.c_stride;
    const unsigned total = (unsigned)p.queue[p.c_direct ? 1 : 6];
    // spread over the SMs first, one warp per SM sub-partition before a second one: only as many warps take work
    // as the list needs
    {
        const unsigned per = gridDim.x * 16u;
        unsigned want = (total + per - 1u) / per;
        if (want < 1u) want = 1u;
        if ((unsigned)warp >= want) return;
    }
    T* wbase = reinterpret_cast<T*>(smem_raw) + (size_t)warp * L::ELEMS * 32;
    const unsigned wb = (unsigned)__cvta_generic_to_shared(wbase);
    SV sv;
    sv.a1 = wb + (unsigned)lane * sizeof(T);
    sv.a2 = wb + (unsigned)lane * 2u * sizeof(T);
    sv.a2own = sv.a2 + (unsigned)(role * 2 * 32 * (int)sizeof(T));
    sv.a2lag = sv.a2 + (unsigned)((1 - role) * 2 * 32 * (int)sizeof(T));
    sv.sat = 0;
    if constexpr (!PIM)
        sv.set_model(role != 0, p.A[1], p.A[6], p.A[11], p.A[14], p.A[15], p.B[3], p.Q[0], p.Q[1], p.R[0], p.hi[0], p.lambda, p.eps);
#pragma unroll
    for (int j = 0; j < NH; ++j) { sv.setU(j, T(0)); sv.setMM(j, T(0)); }

    bool need = true, active = false;
    uint32_t iter = 0, inst = 0;
    while (true) {
        // ---------------- tickets: one per pair
        const unsigned needmask = __ballot_sync(FULL, need) & 0x55555555u;
        if (needmask) {
            const int leader = __ffs(needmask) - 1;
            unsigned long long h = 0;
            if (lane == leader) h = atomicAdd(p.queue + 7, (unsigned long long)__popc(needmask));
            h = __shfl_sync(FULL, h, leader);
            if (need) {
                need = false;
                const unsigned long long t = h + (unsigned long long)__popc(needmask & ((1u << (lane & ~1)) - 1u));
                if (t < (unsigned long long)total) {
                    const unsigned slot = (unsigned)t;
                    inst = list[slot];
                    iter = park_iter[slot];
                    active = true;
                    // rows u | v | MM of the slot, this lane's parity; the pointers advance opaquely so that the
                    // compiler does not precompute (and keep alive) one 64-bit offset per entry
                    const T* su = park + slot + (long long)role * rstride;
                    const T* sv_ = su + (long long)N * rstride;
                    const T* sm_ = sv_ + (long long)N * rstride;
                    const long long step2 = 2 * rstride;
#pragma unroll
                    for (int j = 0; j < NH; ++j) {
                        sv.setU(j, __ldcg(su));
                        sv.setV(j, __ldcg(sv_));
                        sv.setMM(j, __ldcg(sm_));
                        su += step2; sv_ += step2; sm_ += step2;
                        asm volatile("" : "+l"(su), "+l"(sv_), "+l"(sm_));
                    }
                    if constexpr (PIM) {
                        const long long st = p.stride;
                        sv.set_model(role != 0, p.mA[1 * st + inst], p.mA[6 * st + inst], p.mA[11 * st + inst], p.mA[14 * st + inst],
                                     p.mA[15 * st + inst], p.mB[3 * st + inst], p.mQ[inst], p.mQ[st + inst], p.mR[inst], p.mhi[inst],
                                     p.mlambda[inst], p.eps);
                    }
                    sv.refresh_sat();
                }
            }
        }
        if (!__any_sync(FULL, active)) break;

        // ---------------- passes until a pair finishes
        bool fin = false, hit_max = false;
        do {
            const bool notconv = sv.pass();
            if (active) {
                if (!notconv) fin = true;                                              // mpc.h:310-311, update discarded
                else if (++iter >= p.max_iter) { fin = true; hit_max = true; }          // mpc.h:271
            }
        } while (!__any_sync(FULL, fin));

        // ---------------- results
        if (fin) {
            const bool wc = p.warm || p.keep_controls;
            if (role == 0) {
                p.iters[inst] = iter;
                p.status[inst] = (hit_max ? 1 : 0) | (p.status[inst] & 2);
            }
            T* dc = p.controls + (long long)role * p.stride + inst;
            T* dv = p.vstate + (long long)role * p.stride + inst;
            const long long step2 = 2 * p.stride;
#pragma unroll
            for (int j = 0; j < NH; ++j) {
                T uo, vo;
                if (hit_max) { uo = sv.getU(j); vo = sv.getV(j); }
                else SV::IO::ld2(sv.own_addr(j), uo, vo);              // converged: the iterate the test was taken on
                if (j == 0 && role == 0) p.u0[inst] = uo;                                             // mpc.h:239
                if (wc) *dc = uo;
                if (p.warm) *dv = vo;
                dc += step2; dv += step2;
                asm volatile("" : "+l"(dc), "+l"(dv));
            }
            if (p.warm && p.target_full && role == 0) {                                 // mpc.h:236-237 on the ring
                int last = p.head + N - 1; if (last >= N) last -= N;
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    p.target[((long long)p.head * 4 + r) * p.stride + inst] = p.target[((long long)last * 4 + r) * p.stride + inst];
            }
            active = false; need = true;
        }
    }
}

}  // namespace bmpc
