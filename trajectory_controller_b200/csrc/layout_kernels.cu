// Generic (shape-independent) device kernels of libbmpc.so:
//  * layout changes between the ABI's instance-major double arrays and the engine's [element][instance] arrays
//  * trajectory-point extraction, src/trajectory_point_follower.cpp:392-443
#include "bmpc_internal.h"

namespace bmpc {
namespace {

// Transposes between instance-major doubles [count][len] and [len][stride] of T, a tile of kTile instances per block
// staged through shared memory (row pitch len | 1: conflict-free both ways), so that BOTH sides move whole 32-byte
// sectors: the flat instance-major side is walked linearly, the element-major side row by row.  Pure HBM traffic:
// (8 + sizeof T) * len bytes per instance.
constexpr int kTile = 128;
template <typename T>
__global__ void k_aos_to_soa(const double* __restrict__ in, T* __restrict__ out, long long count, int len, long long stride)
{
    extern __shared__ double tile[];
    const long long b0 = (long long)blockIdx.x * kTile;
    const int nb = (int)((count - b0) < kTile ? (count - b0) : kTile), pitch = len | 1;
    for (int f = threadIdx.x; f < nb * len; f += blockDim.x) tile[(f / len) * pitch + f % len] = in[b0 * len + f];
    __syncthreads();
    for (int f = threadIdx.x; f < nb * len; f += blockDim.x) {
        const int e = f / nb, b = f % nb;
        out[(long long)e * stride + b0 + b] = (T)tile[b * pitch + e];
    }
}
template <typename T>
__global__ void k_soa_to_aos(const T* __restrict__ in, double* __restrict__ out, long long count, int len, long long stride)
{
    extern __shared__ double tile[];
    const long long b0 = (long long)blockIdx.x * kTile;
    const int nb = (int)((count - b0) < kTile ? (count - b0) : kTile), pitch = len | 1;
    for (int f = threadIdx.x; f < nb * len; f += blockDim.x) {
        const int e = f / nb, b = f % nb;
        tile[b * pitch + e] = (double)in[(long long)e * stride + b0 + b];
    }
    __syncthreads();
    for (int f = threadIdx.x; f < nb * len; f += blockDim.x) out[b0 * len + f] = tile[(f / len) * pitch + f % len];
}
// Un-tiled forms for rows too long for a shared-memory tile (len > kTileMaxLen): one thread per instance.
constexpr int kTileMaxLen = 40;
template <typename T>
__global__ void k_aos_to_soa_long(const double* __restrict__ in, T* __restrict__ out, long long count, int len, long long stride)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    for (int e = 0; e < len; ++e) out[(long long)e * stride + b] = (T)in[b * len + e];
}
template <typename T>
__global__ void k_soa_to_aos_long(const T* __restrict__ in, double* __restrict__ out, long long count, int len, long long stride)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    for (int e = 0; e < len; ++e) out[b * len + e] = (double)in[(long long)e * stride + b];
}
// Undoes, for the instances a ragged warm solve did NOT touch, the advance of the shared target-ring head
// (mpc.h:236-237 runs only inside operator()): their rows move one physical slot forward so that every logical
// horizon step keeps its value under the new head.  One thread per (instance, state row).
template <typename T>
__global__ void k_rotate_targets(T* __restrict__ target, long long first, long long count, int S, int N, long long stride)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= count * S) return;
    const long long b = first + t / S;
    const int r = (int)(t % S);
    T carry = target[((long long)(N - 1) * S + r) * stride + b];
    for (int ph = 0; ph < N; ++ph) {
        T* cell = target + ((long long)ph * S + r) * stride + b;
        const T keep = *cell; *cell = carry; carry = keep;
    }
}
template <typename T>
__global__ void k_broadcast(const double* __restrict__ vec, T* __restrict__ out, long long count, int len, long long stride)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    for (int e = 0; e < len; ++e) out[(long long)e * stride + b] = (T)vec[e];
}

// getTrajectoryPoint's scan + interpolation (src/trajectory_point_follower.cpp:392-443), one thread per polyline.
// The arc length is a SEQUENTIAL float sum (:423-427): a parallel scan would round differently and could flip the
// selected index, so each thread walks its polyline in order; every float operation is rounded separately
// (no FMA), matching the float semantics adopted for the un-vendored lms::math::vertex2f:
// distance = sqrtf(ex*ex + ey*ey), normalize = (ex/len, ey/len), operator*(float), operator+.
// With the point-major layout (stride_point = count, stride_instance = 1) every load is coalesced.
__global__ void k_extract(long long count, int P, const int32_t* __restrict__ npoints, const float* __restrict__ px,
                          const float* __restrict__ py, const float* __restrict__ dx, const float* __restrict__ dy,
                          const float* __restrict__ vel, long long si, long long sp, const float* __restrict__ dist,
                          int32_t* __restrict__ idx, uint8_t* __restrict__ flag, float* __restrict__ opx,
                          float* __restrict__ opy, float* __restrict__ odx, float* __restrict__ ody,
                          float* __restrict__ ovel)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    const float d = dist[b];
    int n = npoints ? npoints[b] : P;
    if (n > P) n = P;
    const long long o = b * si;
    float rx = d, ry = 0.0f, rdx = 1.0f, rdy = 0.0f, rv = 0.0f;     // :394-403 default point
    int ri = -1; uint8_t rf = BMPC_PT_EMPTY;
    if (n > 0) {                                                     // :404-407
        float cur = 0.0f;                                            // :423
        float bx = px[o], by = py[o];
        bool found = false;
        for (int i = 1; i < n; ++i) {                                // :424
            const float tx = px[o + i * sp], ty = py[o + i * sp];
            const float ex = __fsub_rn(bx, tx), ey = __fsub_rn(by, ty);
            const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));   // :427 distance()
            cur = __fadd_rn(cur, len);
            if (cur > d) {                                           // :428 strict
                const float delta = __fsub_rn(cur, d);              // :431
                const float ax = __fmul_rn(__fdiv_rn(ex, len), delta), ay = __fmul_rn(__fdiv_rn(ey, len), delta);   // :432
                rx = __fadd_rn(tx, ax); ry = __fadd_rn(ty, ay);      // :433-434
                rdx = dx[o + i * sp]; rdy = dy[o + i * sp]; rv = vel[o + i * sp];
                ri = i; rf = BMPC_PT_INTERPOLATED; found = true;
                break;                                               // :435-436
            }
            bx = tx; by = ty;
        }
        if (!found) {                                                // :439-442
            const long long l = o + (long long)(n - 1) * sp;
            rx = px[l]; ry = py[l]; rdx = dx[l]; rdy = dy[l]; rv = vel[l];
            ri = n - 1; rf = BMPC_PT_LAST;
        }
    }
    idx[b] = ri; flag[b] = rf; opx[b] = rx; opy[b] = ry; odx[b] = rdx; ody[b] = rdy; ovel[b] = rv;
}

// getTrajectoryPoint for `steps` ascending look-ahead distances d_t = d0 + t*step of the same polyline in ONE sequential
// scan (a target per horizon step).  The arc length `cur` is the same float sum whatever the
// distance asked for, so step t's point is produced by the first segment with cur > d_t — exactly what k_extract
// returns for d_t; several steps may fall into one segment, the remaining ones get the last point (:439-442).
// Outputs are instance-major [count][steps] (the host layout): a thread collects its `steps` results in its row of a
// shared-memory tile (row stride steps|1 words: conflict-free), then the block writes its contiguous slice of each
// output array with coalesced stores.
__global__ void k_extract_horizon(long long count, int P, const int32_t* __restrict__ npoints, const float* __restrict__ px,
                                  const float* __restrict__ py, const float* __restrict__ dx, const float* __restrict__ dy,
                                  const float* __restrict__ vel, long long si, long long sp, const float* __restrict__ d0,
                                  const float* __restrict__ step, int steps, int32_t* __restrict__ idx,
                                  uint8_t* __restrict__ flag, float* __restrict__ opx, float* __restrict__ opy,
                                  float* __restrict__ odx, float* __restrict__ ody, float* __restrict__ ovel)
{
    extern __shared__ __align__(16) unsigned char hz_raw[];
    const int rs = steps | 1;                                        // row stride in words
    const int tile = blockDim.x * rs;
    float* sm = reinterpret_cast<float*>(hz_raw);                     // 5 float planes, then the idx plane, then flags
    int32_t* sm_i = reinterpret_cast<int32_t*>(sm + 5 * tile);
    uint8_t* sm_f = reinterpret_cast<uint8_t*>(sm_i + tile);
    const long long b0 = (long long)blockIdx.x * blockDim.x, b = b0 + threadIdx.x;
    const int row = threadIdx.x * rs;
    auto emit = [&](int t, int i, uint8_t f, float x, float y, float ddx, float ddy, float v) {
        sm[row + t] = x; sm[tile + row + t] = y; sm[2 * tile + row + t] = ddx; sm[3 * tile + row + t] = ddy;
        sm[4 * tile + row + t] = v; sm_i[row + t] = i; sm_f[row + t] = f;
    };
    if (b < count) {
        const float d_0 = d0[b], st = step[b];
        int n = npoints ? npoints[b] : P;
        if (n > P) n = P;
        const long long o = b * si;
        auto dist_of = [&](int t) { return __fadd_rn(d_0, __fmul_rn((float)t, st)); };
        int t = 0;
        if (n <= 0) {                                                // :394-407 for every step
            for (; t < steps; ++t) emit(t, -1, BMPC_PT_EMPTY, dist_of(t), 0.0f, 1.0f, 0.0f, 0.0f);
        } else {
            float cur = 0.0f;                                        // :423
            float bx = px[o], by = py[o];
            float d = dist_of(0);
            for (int i = 1; i < n && t < steps; ++i) {               // :424
                const float tx = px[o + i * sp], ty = py[o + i * sp];
                const float ex = __fsub_rn(bx, tx), ey = __fsub_rn(by, ty);
                const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));   // :427
                cur = __fadd_rn(cur, len);
                if (cur > d) {                                       // :428 strict
                    const float nx = __fdiv_rn(ex, len), ny = __fdiv_rn(ey, len);                // :432
                    const float ddx = dx[o + i * sp], ddy = dy[o + i * sp], v = vel[o + i * sp];
                    do {
                        const float delta = __fsub_rn(cur, d);      // :431
                        emit(t, i, BMPC_PT_INTERPOLATED, __fadd_rn(tx, __fmul_rn(nx, delta)), __fadd_rn(ty, __fmul_rn(ny, delta)), ddx, ddy, v);
                        ++t;
                        d = dist_of(t);
                    } while (t < steps && cur > d);
                }
                bx = tx; by = ty;
            }
            if (t < steps) {                                         // :439-442
                const long long l = o + (long long)(n - 1) * sp;
                const float x = px[l], y = py[l], ddx = dx[l], ddy = dy[l], v = vel[l];
                for (; t < steps; ++t) emit(t, n - 1, BMPC_PT_LAST, x, y, ddx, ddy, v);
            }
        }
    }
    __syncthreads();
    const long long rows = count - b0 < blockDim.x ? count - b0 : blockDim.x;
    const long long total = rows * steps, base = b0 * steps;
    float* const outs[5] = {opx, opy, odx, ody, ovel};
    for (long long e = threadIdx.x; e < total; e += blockDim.x) {
        const int at = (int)(e / steps) * rs + (int)(e % steps);
#pragma unroll
        for (int a = 0; a < 5; ++a) outs[a][base + e] = sm[a * tile + at];
        idx[base + e] = sm_i[at];
        flag[base + e] = sm_f[at];
    }
}

// ---- small-call paths: a handful of numbers travel as KERNEL ARGUMENTS instead of through a staging copy, so that a
// capacity-1 controller (the dlib::mpc facade, one object per cycle() in the reference module) pays no H2D copy and no
// stream synchronisation for its inputs.
// vals[off[a] .. off[a] + len[a]) -> dst[a][e * stride + b] for b in [first, first + count): `count` == capacity
// broadcasts one vector to every instance (set_target(val)), count == 1 writes one instance (shared model, x0).
template <typename T>
__global__ void k_put_vals(SmallVals v, PutPlan plan, long long first, long long count, long long stride, unsigned long long* zero8)
{
    if (zero8 && blockIdx.x == 0 && threadIdx.x < 8) zero8[threadIdx.x] = 0ull;   // the launch counters of the solve that follows
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    for (int a = 0; a < plan.n; ++a) {
        T* dst = reinterpret_cast<T*>(plan.dst[a]);
        for (int e = 0; e < plan.len[a]; ++e) dst[(long long)e * stride + first + b] = (T)v.x[plan.off[a] + e];
    }
}
// Results of a small solve, packed for ONE device-to-host copy: [count][I] doubles | iters u32[count] | status u8[count]
// (padded to 8) | the launch's error word (queue[4]).
template <typename T>
__global__ void k_pack_results(const T* __restrict__ u0, const uint32_t* __restrict__ iters, const uint8_t* __restrict__ status,
                               const unsigned long long* __restrict__ queue, int count, int I, long long stride, unsigned char* out,
                               unsigned long long* flag, unsigned long long seq)
{
    const int t = threadIdx.x;
    double* ou = reinterpret_cast<double*>(out);
    uint32_t* oi = reinterpret_cast<uint32_t*>(out + (size_t)count * I * 8);
    uint8_t* os = out + (size_t)count * I * 8 + (size_t)count * 4;
    unsigned long long* oe = reinterpret_cast<unsigned long long*>(out + small_results_err_offset(count, I));
    for (int f = t; f < count * I; f += blockDim.x) ou[f] = (double)u0[(long long)(f % I) * stride + f / I];
    for (int b = t; b < count; b += blockDim.x) { oi[b] = iters[b]; os[b] = status[b]; }
    if (t == 0) *oe = queue[4];
    if (flag) {   // `out` is pinned host memory the caller polls: everything above must be visible there before the flag
        __threadfence_system();
        __syncthreads();
        if (t == 0) { *reinterpret_cast<volatile unsigned long long*>(flag) = seq; __threadfence_system(); }
    }
}

inline unsigned blocks_for(long long count, int block) { return (unsigned)((count + block - 1) / block); }
}  // namespace

cudaError_t aos_to_soa(const double* d_aos, void* d_soa, long long count, int len, long long stride, int elem_size,
                       cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    if (len > kTileMaxLen) {
        if (elem_size == 8) k_aos_to_soa_long<double><<<blocks_for(count, 256), 256, 0, st>>>(d_aos, (double*)d_soa, count, len, stride);
        else k_aos_to_soa_long<float><<<blocks_for(count, 256), 256, 0, st>>>(d_aos, (float*)d_soa, count, len, stride);
        return cudaGetLastError();
    }
    const size_t sm = (size_t)kTile * (len | 1) * 8;
    if (elem_size == 8) k_aos_to_soa<double><<<blocks_for(count, kTile), 256, sm, st>>>(d_aos, (double*)d_soa, count, len, stride);
    else k_aos_to_soa<float><<<blocks_for(count, kTile), 256, sm, st>>>(d_aos, (float*)d_soa, count, len, stride);
    return cudaGetLastError();
}
cudaError_t soa_to_aos(const void* d_soa, double* d_aos, long long count, int len, long long stride, int elem_size,
                       cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    if (len > kTileMaxLen) {
        if (elem_size == 8) k_soa_to_aos_long<double><<<blocks_for(count, 256), 256, 0, st>>>((const double*)d_soa, d_aos, count, len, stride);
        else k_soa_to_aos_long<float><<<blocks_for(count, 256), 256, 0, st>>>((const float*)d_soa, d_aos, count, len, stride);
        return cudaGetLastError();
    }
    const size_t sm = (size_t)kTile * (len | 1) * 8;
    if (elem_size == 8) k_soa_to_aos<double><<<blocks_for(count, kTile), 256, sm, st>>>((const double*)d_soa, d_aos, count, len, stride);
    else k_soa_to_aos<float><<<blocks_for(count, kTile), 256, sm, st>>>((const float*)d_soa, d_aos, count, len, stride);
    return cudaGetLastError();
}
cudaError_t rotate_targets(void* d_target, long long first, long long count, int S, int N, long long stride, int elem_size,
                           cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    if (elem_size == 8) k_rotate_targets<double><<<blocks_for(count * S, 256), 256, 0, st>>>((double*)d_target, first, count, S, N, stride);
    else k_rotate_targets<float><<<blocks_for(count * S, 256), 256, 0, st>>>((float*)d_target, first, count, S, N, stride);
    return cudaGetLastError();
}
cudaError_t broadcast_rows(const double* d_vec, void* d_soa, long long count, int len, long long stride, int elem_size,
                           cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    if (elem_size == 8) k_broadcast<double><<<blocks_for(count, 256), 256, 0, st>>>(d_vec, (double*)d_soa, count, len, stride);
    else k_broadcast<float><<<blocks_for(count, 256), 256, 0, st>>>(d_vec, (float*)d_soa, count, len, stride);
    return cudaGetLastError();
}
cudaError_t launch_extract(long long count, int P, const int32_t* npoints, const float* px, const float* py,
                           const float* dx, const float* dy, const float* vel, long long stride_instance,
                           long long stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* opx,
                           float* opy, float* odx, float* ody, float* ovel, cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    k_extract<<<blocks_for(count, 128), 128, 0, st>>>(count, P, npoints, px, py, dx, dy, vel, stride_instance,
                                                     stride_point, dist, idx, flag, opx, opy, odx, ody, ovel);
    return cudaGetLastError();
}

cudaError_t put_vals(const SmallVals& v, const PutPlan& plan, long long first, long long count, long long stride, int elem_size,
                     cudaStream_t st, unsigned long long* zero8)
{
    if (count <= 0) return cudaSuccess;
    const int block = count < 128 ? 32 : 128;
    if (elem_size == 8) k_put_vals<double><<<blocks_for(count, block), block, 0, st>>>(v, plan, first, count, stride, zero8);
    else k_put_vals<float><<<blocks_for(count, block), block, 0, st>>>(v, plan, first, count, stride, zero8);
    return cudaGetLastError();
}
cudaError_t pack_results(const void* u0, const uint32_t* iters, const uint8_t* status, const unsigned long long* queue, int count,
                         int I, long long stride, int elem_size, unsigned char* out, cudaStream_t st, unsigned long long* flag,
                         unsigned long long seq)
{
    if (elem_size == 8) k_pack_results<double><<<1, 64, 0, st>>>((const double*)u0, iters, status, queue, count, I, stride, out, flag, seq);
    else k_pack_results<float><<<1, 64, 0, st>>>((const float*)u0, iters, status, queue, count, I, stride, out, flag, seq);
    return cudaGetLastError();
}

cudaError_t launch_extract_horizon(long long count, int P, const int32_t* npoints, const float* px, const float* py,
                                   const float* dx, const float* dy, const float* vel, long long stride_instance,
                                   long long stride_point, const float* d0, const float* step, int steps, int32_t* idx,
                                   uint8_t* flag, float* opx, float* opy, float* odx, float* ody, float* ovel, cudaStream_t st)
{
    if (count <= 0 || steps <= 0) return cudaSuccess;
    // 25 bytes per (thread, step) of shared-memory tile; the block shrinks for long horizons
    int block = 128;
    auto bytes = [&](int blk) { return (size_t)blk * (steps | 1) * 25 + 16; };
    while (block > 32 && bytes(block) > 96 * 1024) block >>= 1;
    if (bytes(block) > 227 * 1024) return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(k_extract_horizon, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    k_extract_horizon<<<blocks_for(count, block), block, bytes(block), st>>>(count, P, npoints, px, py, dx, dy, vel, stride_instance,
                                                                            stride_point, d0, step, steps, idx, flag, opx, opy, odx, ody, ovel);
    return cudaGetLastError();
}

}  // namespace bmpc
