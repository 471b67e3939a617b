// Arithmetic-pipe peak probe: the roofline denominator for the solver is FP64 (or FP32) CUDA-core throughput,
// which MEASURED_PEAKS.json does not contain (it has HBM and bf16 tensor only), so the engine measures it on the
// device it runs on.  Register-resident dependent chains, 8 independent chains per thread, no memory traffic.
#include "bmpc_internal.h"

namespace bmpc {
namespace {

template <typename T, bool FUSED>
__global__ void __launch_bounds__(256) k_peak(T* sink, int iters, T a, T b)
{
    T x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = T(threadIdx.x + k) * T(1e-3);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int rep = 0; rep < 8; ++rep) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                if (FUSED) x[k] = Ar<T, false>::madd(b, x[k], a);                 // 1 instruction, 2 flops
                else x[k] = (rep & 1) ? Ar<T, true>::add(x[k], b) : Ar<T, true>::mul(x[k], a);   // 1 instruction, 1 flop
            }
        }
    }
    T s = T(0);
#pragma unroll
    for (int k = 0; k < 8; ++k) s += x[k];
    if (s == T(123.456)) sink[0] = s;   // never true; keeps the chains alive
}

template <typename T, bool FUSED>
int run_peak(bmpc_handle* h, double* tflops, double* ginstr)
{
    const int iters = 4096, block = 256, grid = h->sm_count * 8;
    BMPC_CUDA(h->small.reserve(64));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        BMPC_CUDA(cudaEventRecord(h->ev0, h->stream));
        k_peak<T, FUSED><<<grid, block, 0, h->stream>>>((T*)h->small.p, iters, T(0.999999), T(1e-9));
        BMPC_CUDA(cudaGetLastError());
        BMPC_CUDA(cudaEventRecord(h->ev1, h->stream));
        BMPC_CUDA(cudaEventSynchronize(h->ev1));
        float ms = 0;
        BMPC_CUDA(cudaEventElapsedTime(&ms, h->ev0, h->ev1));
        if (rep > 0 && ms < best) best = ms;
        h->launches++;
    }
    const double instr = (double)grid * block * iters * 64.0;
    *ginstr = instr / (best * 1e-3) / 1e9;
    *tflops = *ginstr * (FUSED ? 2.0 : 1.0) / 1e3;
    return BMPC_OK;
}
}  // namespace
}  // namespace bmpc

using namespace bmpc;

extern "C" int bmpc_probe_peak(bmpc_t* h, int precision, int fused, double* tflops, double* ginstr_per_s)
{
    if (!h || !tflops || !ginstr_per_s) return fail(BMPC_E_INVALID, "bmpc_probe_peak: null argument");
    BMPC_CUDA(cudaSetDevice(h->device));
    h->timed = false;
    if (precision == BMPC_FP64) return fused ? run_peak<double, true>(h, tflops, ginstr_per_s) : run_peak<double, false>(h, tflops, ginstr_per_s);
    if (precision == BMPC_FP32) return fused ? run_peak<float, true>(h, tflops, ginstr_per_s) : run_peak<float, false>(h, tflops, ginstr_per_s);
    return fail(BMPC_E_UNSUPPORTED, "bmpc_probe_peak: precision must be 64 or 32");
}
