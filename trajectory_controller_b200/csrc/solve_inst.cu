// One translation unit per (template shape, structural pattern):
//   nvcc ... -DBMPC_S=4 -DBMPC_I=1 -DBMPC_N=30 -DBMPC_PAT=bicycle4 solve_inst.cu
// Instantiates the solve kernel for {fp64,fp32} x {exact,fast} x {shared,per-instance model} and the constructor
// kernel (mpc.h:114-123), and exports them through a shape-erased ShapeOps table (dispatch.h).
#include "ctor_kernel.cuh"
#include "dispatch.h"
#include "patterns.h"
#include "pair_kernel.cuh"
#include "pg_kernel.cuh"
#include "solve_kernel.cuh"

#ifndef BMPC_S
#error "compile with -DBMPC_S= -DBMPC_I= -DBMPC_N= -DBMPC_PAT="
#endif
#ifndef BMPC_PAT
#define BMPC_PAT dense
#endif

namespace bmpc {
namespace {
constexpr int S_ = BMPC_S, I_ = BMPC_I, N_ = BMPC_N;
using PAT_ = typename pat::BMPC_PAT<S_, I_>::type;
constexpr bool kDense = (PAT_::ANZ == DensePattern<S_, I_>::ANZ && PAT_::AONE == 0u && PAT_::BNZ == DensePattern<S_, I_>::BNZ &&
                         PAT_::BONE == 0u && PAT_::CNZ == DensePattern<S_, I_>::CNZ && PAT_::QNZ == DensePattern<S_, I_>::QNZ);

template <typename T, bool PIM>
void fill_params(SolveParams<T, S_, I_, N_>& p, const GenericSolve& g)
{
    for (int e = 0; e < S_ * S_; ++e) p.A[e] = PIM ? T(0) : (T)g.A[e];
    for (int e = 0; e < S_ * I_; ++e) p.B[e] = PIM ? T(0) : (T)g.B[e];
    for (int e = 0; e < S_; ++e) { p.C[e] = PIM ? T(0) : (T)g.C[e]; p.Q[e] = PIM ? T(0) : (T)g.Q[e]; }
    for (int e = 0; e < I_; ++e) {
        p.R[e] = PIM ? T(0) : (T)g.R[e]; p.lo[e] = PIM ? T(0) : (T)g.lo[e]; p.hi[e] = PIM ? T(0) : (T)g.hi[e];
    }
    p.lambda = PIM ? T(1) : (T)g.lambda;
    for (int e = 0; e < N_ * I_; ++e) p.qdiag[e] = PIM ? T(0) : (T)g.qdiag[e];
    p.eps = (T)g.eps; p.max_iter = g.max_iter;
    p.warm = g.warm; p.keep_controls = g.keep_controls; p.target_full = g.target_full; p.head = g.head;
    p.count = g.count; p.stride = g.stride;
    // every per-instance array is [element][stride]: a launch on instances [first, first + count) shifts the bases
    const long long f = g.first;
    p.x0 = (const T*)g.x0 + f; p.controls = (T*)g.controls + f; p.vstate = (T*)g.vstate + f; p.target = (T*)g.target + f;
    p.u0 = (T*)g.u0 + f; p.iters = g.iters + f; p.status = g.status + f;
    p.mA = (const T*)g.mA + f; p.mB = (const T*)g.mB + f; p.mC = (const T*)g.mC + f; p.mQ = (const T*)g.mQ + f; p.mR = (const T*)g.mR + f;
    p.mlo = (const T*)g.mlo + f; p.mhi = (const T*)g.mhi + f; p.mlambda = (const T*)g.mlambda + f; p.mqdiag = (const T*)g.mqdiag + f;
    p.queue = g.queue;
    p.split = g.split; p.pg_list = g.pg_list; p.pg_mask = g.pg_mask; p.pg_slice = g.pg_slice;
    p.park = (T*)g.park; p.park_iter = g.park_iter; p.pg_reuse = g.pg_reuse; p.sym = g.sym;
    p.c_direct = g.c_direct; p.c_threshold = g.c_threshold; p.c_list = g.c_list; p.c_park = (T*)g.c_park; p.c_iter = g.c_iter;
    p.c_stride = g.c_stride;
}

template <typename T, bool EXACT, bool PIM>
cudaError_t launch_solve(const GenericSolve& g, int grid, cudaStream_t st)
{
    using L = Layout<T, S_, I_, N_, PIM, PAT_>;
    SolveParams<T, S_, I_, N_> p;
    fill_params<T, PIM>(p, g);
    p.warp_limit = g.warp_limit_a;
    auto kern = solve_kernel<T, S_, I_, N_, EXACT, PIM, PAT_>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    kern<<<grid, L::THREADS, L::SMEM_BYTES, st>>>(p);
    return cudaGetLastError();
}

template <typename T, bool EXACT, bool PIM>
cudaError_t info_solve(KernelInfo* out)
{
    using L = Layout<T, S_, I_, N_, PIM, PAT_>;
    cudaFuncAttributes a;
    cudaError_t e = cudaFuncGetAttributes(&a, solve_kernel<T, S_, I_, N_, EXACT, PIM, PAT_>);
    if (e != cudaSuccess) return e;
    out->block = L::THREADS; out->smem_bytes = L::SMEM_BYTES; out->regs = a.numRegs; out->warps = L::WARPS;
    out->local_bytes = (int)a.localSizeBytes; out->sym_ok = 0;
    return cudaSuccess;
}

constexpr bool kHasPg = PgLayout<double, S_, I_, N_, false, PAT_>::OK && PgLayout<float, S_, I_, N_, false, PAT_>::OK;

template <typename T, bool EXACT, bool PIM, bool SYM>
cudaError_t launch_pg_t(const GenericSolve& g, int grid, cudaStream_t st)
{
    using L = PgLayout<T, S_, I_, N_, PIM, PAT_>;
    SolveParams<T, S_, I_, N_> p;
    fill_params<T, PIM>(p, g);
    p.warp_limit = g.warp_limit_b;
    auto kern = pg_kernel<T, S_, I_, N_, EXACT, PIM, PAT_, SYM>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES);
    if (e != cudaSuccess) return e;
    kern<<<grid, L::THREADS, L::SMEM_BYTES, st>>>(p);
    return cudaGetLastError();
}

template <typename T, bool EXACT, bool PIM>
cudaError_t launch_pg(const GenericSolve& g, int grid, cudaStream_t st)
{
    if constexpr (!kHasPg) return cudaErrorNotSupported;
    else {
        if constexpr (PgLayout<T, S_, I_, N_, PIM, PAT_>::SYM_OK) {
            if (g.sym) return launch_pg_t<T, EXACT, PIM, true>(g, grid, st);
        }
        return launch_pg_t<T, EXACT, PIM, false>(g, grid, st);
    }
}

template <typename T, bool EXACT, bool PIM>
cudaError_t info_pg(int sym, KernelInfo* out)
{
    if constexpr (!kHasPg) return cudaErrorNotSupported;
    else {
        using L = PgLayout<T, S_, I_, N_, PIM, PAT_>;
        cudaFuncAttributes a;
        cudaError_t e;
        if constexpr (L::SYM_OK) {
            if (sym) e = cudaFuncGetAttributes(&a, pg_kernel<T, S_, I_, N_, EXACT, PIM, PAT_, true>);
            else e = cudaFuncGetAttributes(&a, pg_kernel<T, S_, I_, N_, EXACT, PIM, PAT_, false>);
        } else e = cudaFuncGetAttributes(&a, pg_kernel<T, S_, I_, N_, EXACT, PIM, PAT_, false>);
        if (e != cudaSuccess) return e;
        out->block = L::THREADS; out->smem_bytes = L::SMEM_BYTES; out->regs = a.numRegs; out->warps = L::WARPS;
        out->local_bytes = (int)a.localSizeBytes; out->sym_ok = L::SYM_OK ? 1 : 0;
        return cudaSuccess;
    }
}

// stage C: the lane-pair kernel exists for the block-triangular (4,1) pattern and even horizons
constexpr bool kHasPair = (S_ == 4 && I_ == 1 && N_ % 2 == 0 && PAT_::ANZ == 0xCC63u && PAT_::AONE == 0x0421u && PAT_::BNZ == 0x8u &&
                           PAT_::BONE == 0u && PAT_::CNZ == 0u && PAT_::QNZ == 0x3u);

template <typename T, bool EXACT, bool PIM>
cudaError_t launch_pair(const GenericSolve& g, int grid, cudaStream_t st)
{
    if constexpr (!kHasPair) return cudaErrorNotSupported;
    else {
        using L = PairLayout<T, N_, PIM>;
        SolveParams<T, S_, I_, N_> p;
        fill_params<T, PIM>(p, g);
        auto kern = pair_kernel<T, N_, EXACT, PIM>;
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::SMEM_BYTES);
        if (e != cudaSuccess) return e;
        kern<<<grid, L::THREADS, L::SMEM_BYTES, st>>>(p);
        return cudaGetLastError();
    }
}

template <typename T, bool EXACT, bool PIM>
cudaError_t info_pair(KernelInfo* out)
{
    if constexpr (!kHasPair) return cudaErrorNotSupported;
    else {
        using L = PairLayout<T, N_, PIM>;
        cudaFuncAttributes a;
        cudaError_t e = cudaFuncGetAttributes(&a, pair_kernel<T, N_, EXACT, PIM>);
        if (e != cudaSuccess) return e;
        out->block = L::THREADS; out->smem_bytes = L::SMEM_BYTES; out->regs = a.numRegs; out->warps = L::WARPS;
        out->local_bytes = (int)a.localSizeBytes; out->sym_ok = 1;
        return cudaSuccess;
    }
}

template <typename T, bool EXACT>
cudaError_t launch_ctor(const GenericCtor& g, cudaStream_t st)
{
    const int block = 128;
    const long long grid = (g.count + block - 1) / block;
    ctor_kernel<T, S_, I_, N_, EXACT><<<(unsigned)grid, block, 0, st>>>(
        g.count, g.stride, (const T*)g.mA, (const T*)g.mB, (const T*)g.mQ, (const T*)g.mR, (T*)g.mlambda, (T*)g.mqdiag);
    return cudaGetLastError();
}

#define BMPC_DISPATCH(FN, ...)                                                                                   \
    do {                                                                                                         \
        const bool ex = (mode == 0);                                                                             \
        if (precision == 64) {                                                                                   \
            if (ex) { if (pim) return FN<double, true, true>(__VA_ARGS__); return FN<double, true, false>(__VA_ARGS__); }   \
            if (pim) return FN<double, false, true>(__VA_ARGS__);                                                \
            return FN<double, false, false>(__VA_ARGS__);                                                        \
        }                                                                                                        \
        if (ex) { if (pim) return FN<float, true, true>(__VA_ARGS__); return FN<float, true, false>(__VA_ARGS__); }         \
        if (pim) return FN<float, false, true>(__VA_ARGS__);                                                     \
        return FN<float, false, false>(__VA_ARGS__);                                                             \
    } while (0)

cudaError_t solve_entry(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st)
{ BMPC_DISPATCH(launch_solve, g, grid, st); }
cudaError_t info_entry(int precision, int mode, int pim, KernelInfo* out)
{ BMPC_DISPATCH(info_solve, out); }
cudaError_t pg_entry(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st)
{ BMPC_DISPATCH(launch_pg, g, grid, st); }
cudaError_t pg_info_entry(int precision, int mode, int pim, int sym, KernelInfo* out)
{ BMPC_DISPATCH(info_pg, sym, out); }
cudaError_t pair_entry(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st)
{ BMPC_DISPATCH(launch_pair, g, grid, st); }
cudaError_t pair_info_entry(int precision, int mode, int pim, KernelInfo* out)
{ BMPC_DISPATCH(info_pair, out); }
template <bool DENSE>
cudaError_t ctor_entry_t(int precision, int mode, const GenericCtor& g, cudaStream_t st)
{
    if constexpr (!DENSE) return cudaErrorNotSupported;   // the constructor recursion lives in the dense unit only
    else {
    if (precision == 64) return mode == 0 ? launch_ctor<double, true>(g, st) : launch_ctor<double, false>(g, st);
    return mode == 0 ? launch_ctor<float, true>(g, st) : launch_ctor<float, false>(g, st);
    }
}
cudaError_t ctor_entry(int precision, int mode, const GenericCtor& g, cudaStream_t st)
{ return ctor_entry_t<kDense>(precision, mode, g, st); }
}  // namespace

#define BMPC_CAT_(a, b, c, d) bmpc_ops_##a##_##b##_##c##_##d
#define BMPC_CAT(a, b, c, d) BMPC_CAT_(a, b, c, d)
#define BMPC_STR_(x) #x
#define BMPC_STR(x) BMPC_STR_(x)
extern const ShapeOps BMPC_CAT(BMPC_S, BMPC_I, BMPC_N, BMPC_PAT);
const ShapeOps BMPC_CAT(BMPC_S, BMPC_I, BMPC_N, BMPC_PAT) = {
    S_, I_, N_, BMPC_STR(BMPC_PAT), PAT_::ANZ, PAT_::AONE, PAT_::BNZ, PAT_::BONE, PAT_::CNZ, PAT_::QNZ,
    solve_entry, ctor_entry, info_entry, kHasPg ? 1 : 0, pg_entry, pg_info_entry, kHasPair ? 1 : 0, pair_entry, pair_info_entry};

}  // namespace bmpc
