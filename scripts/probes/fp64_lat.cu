// Developer probe: dependent-issue latency and per-SMSP throughput of FP64 ops on the device (not product code).
#include <cstdio>
#include <cuda_runtime.h>
template <int OP, int ILP>
__global__ void chain(double* out, long long* cyc, int n, double a, double b)
{
    double x[ILP];
#pragma unroll
    for (int k = 0; k < ILP; ++k) x[k] = a + k + threadIdx.x;
    long long t0 = clock64();
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r)
#pragma unroll
            for (int k = 0; k < ILP; ++k) {
                if (OP == 0) x[k] = __dadd_rn(x[k], b);
                if (OP == 1) x[k] = __dmul_rn(x[k], b);
                if (OP == 2) x[k] = fma(x[k], b, a);
                if (OP == 3) x[k] = (x[k] <= a) ? __dadd_rn(x[k], b) : b;   // DSETP -> select -> DADD
            }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < ILP; ++k) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int OP, int ILP>
void run(const char* name, int warps)
{
    double* out; long long* cyc; cudaMalloc(&out, 8 * 1024); cudaMalloc(&cyc, 8);
    const int n = 256;
    chain<OP, ILP><<<1, warps * 32>>>(out, cyc, n, 1.0, 1.0000001);
    chain<OP, ILP><<<1, warps * 32>>>(out, cyc, n, 1.0, 1.0000001);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double per = (double)c / (n * 16.0);
    printf("%-6s ILP=%d warps/CTA=%2d : %.2f cycles per dependent step, %.2f cycles per warp-instruction per SMSP\n", name, ILP, warps, per,
           per / ILP / ((warps + 3) / 4));
    cudaFree(out); cudaFree(cyc);
}
int main()
{
    run<0, 1>("DADD", 1); run<1, 1>("DMUL", 1); run<2, 1>("DFMA", 1); run<3, 1>("DSETP+", 1);
    run<0, 2>("DADD", 1); run<0, 4>("DADD", 1); run<0, 8>("DADD", 1);
    run<0, 1>("DADD", 4); run<0, 1>("DADD", 8); run<0, 2>("DADD", 8); run<0, 4>("DADD", 8); run<0, 1>("DADD", 16); run<0,1>("DADD", 32);
    return 0;
}
