// Host-side, shape-erased view of one template instantiation family (internal to libbmpc.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace bmpc {

// Everything a solve launch needs, in doubles / untyped device pointers; the shape's translation unit converts it
// into its SolveParams<T,S,I,N> kernel-parameter block.
struct GenericSolve {
    const double *A, *B, *C, *Q, *R, *lo, *hi;   // shared model (host), may be null for per-instance models
    double lambda;
    const double* qdiag;                         // host [N*I]
    double eps;
    uint32_t max_iter;
    int warm, keep_controls, target_full, head;
    long long count, stride;
    long long first;                             // the launch works on instances [first, first + count) of the handle
    const void* x0; void* controls; void* vstate; void* target; void* u0;
    uint32_t* iters; uint8_t* status;
    const void *mA, *mB, *mC, *mQ, *mR, *mlo, *mhi, *mlambda, *mqdiag;
    unsigned long long* queue;                   // 64 bytes: [0] stage-A claims, [1] stage-B list length, [2] stage-B claims
    int split;                                   // run the two-stage solve (pg_kernel.cuh)
    uint32_t* pg_list; unsigned pg_mask; int pg_slice;
    void* park; uint32_t* park_iter;             // by-ring-slot state of parked instances (pg_kernel.cuh)
    int pg_reuse;                                // stage B recycles ring slots (parking enabled)
    int sym;                                     // bounds are symmetric and the model is tame: stage B may run its SYM kernel
    int warp_limit_a, warp_limit_b;              // warps per CTA that take work in stage A / stage B
    // stage C (pair_kernel.cuh)
    int c_direct; unsigned c_threshold; uint32_t* c_list; void* c_park; uint32_t* c_iter; unsigned c_stride;
};

// mpc.h:114-123 for `count` models held SoA [elem][stride] on the device.
struct GenericCtor {
    long long count, stride;
    const void *mA, *mB, *mQ, *mR;
    void* mlambda; void* mqdiag;
};

struct KernelInfo { int block, smem_bytes, regs, warps, local_bytes, sym_ok; };

// One (shape, structural pattern) instantiation family.  `ctor` is only provided by the dense pattern.
struct ShapeOps {
    int S, I, N;
    const char* pattern;                       // "dense", "bicycle4", ...
    unsigned anz, aone, bnz, bone, cnz, qnz;   // solve_kernel.cuh Pattern bits
    cudaError_t (*solve)(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st);
    cudaError_t (*ctor)(int precision, int mode, const GenericCtor& g, cudaStream_t st);
    cudaError_t (*info)(int precision, int mode, int pim, KernelInfo* out);
    // stage B of the split solve; has_pg == 0 when the shape keeps u[]/MM[] out of registers (long horizons)
    int has_pg;
    cudaError_t (*pg)(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st);
    cudaError_t (*pg_info)(int precision, int mode, int pim, int sym, KernelInfo* out);
    // stage C: one instance per lane pair (bicycle4 pattern, even horizons); has_pair == 0 elsewhere
    int has_pair;
    cudaError_t (*pair)(int precision, int mode, int pim, const GenericSolve& g, int grid, cudaStream_t st);
    cudaError_t (*pair_info)(int precision, int mode, int pim, KernelInfo* out);
};

}  // namespace bmpc
