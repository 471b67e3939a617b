// libbmpc.so — trajectory-point extraction WITHOUT any allocation: bmpc_extract_points_static (include/bmpc.h).
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code that runs next to a CUDA graph capture, tight serving loops, harnesses that forbid it).  bmpc_extract_points
// stages its host arrays in growable device buffers of the handle (DevBuf::reserve -> cudaMalloc); this variant stages
// them in ONE file-scope device array of BMPC_STATIC_EXTRACT_BYTES, which every GPU that loads the library holds until
// the process exits, and runs a variant of k_extract (layout_kernels.cu) that addresses everything relative to that
// array.  Nothing here calls cudaMalloc, cudaFree, cudaHostAlloc or a library routine that allocates; there is no handle.
//
// The array's name is a HOST-SIDE STAND-IN in host code, not a device address: its address on a device is asked for once
// per device, with that device current, through cudaGetSymbolAddress (static_base), and that pointer is what the copies
// and the kernel get.  All calls on a device share the array and it keeps the last call's contents: nothing in it has to
// start at zero for the arithmetic (every byte the kernel reads is uploaded by the same call), but the OUTPUT window is
// cleared first so that a stale result of an earlier call can never be handed back.  Calls from several host threads are
// serialised by a mutex and the entry point is synchronous on return, so calls through this ABI never overlap on the
// GPU; anything else that made two of them overlap would race on the array.
#include <mutex>

#include "bmpc_internal.h"

namespace {

static __device__ __align__(256) unsigned char g_static_scratch[BMPC_STATIC_EXTRACT_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_static_mutex;                       // guards g_static_ptr and serialises the users of the array
unsigned char* g_static_ptr[kMaxDevices] = {};   // device address of g_static_scratch on device d (nullptr: not asked yet)

// Byte offsets of the windows inside the array (all multiples of 256).
struct StaticPlan {
    unsigned long long px, py, dx, dy, vel, dist, npoints, out_idx, out_flag, out_px, out_py, out_dx, out_dy, out_vel;
    int has_npoints;
};

constexpr size_t align256(size_t n) { return (n + 255) & ~(size_t)255; }

// k_extract (layout_kernels.cu; src/trajectory_point_follower.cpp:392-443) with every array addressed as base + offset
// of the static scratch array: the same sequential float scan, every operation rounded on its own (no FMA), so the
// results are bit-identical to bmpc_extract_points'.
__global__ void k_extract_static(unsigned char* __restrict__ base, StaticPlan pl, long long count, int P, long long si,
                                 long long sp)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    const float* px = (const float*)(base + pl.px);
    const float* py = (const float*)(base + pl.py);
    const float* dx = (const float*)(base + pl.dx);
    const float* dy = (const float*)(base + pl.dy);
    const float* vel = (const float*)(base + pl.vel);
    const float d = ((const float*)(base + pl.dist))[b];
    int n = pl.has_npoints ? ((const int32_t*)(base + pl.npoints))[b] : P;
    if (n > P) n = P;
    const long long o = b * si;
    float rx = d, ry = 0.0f, rdx = 1.0f, rdy = 0.0f, rv = 0.0f;     // :394-403 default point
    int ri = -1; uint8_t rf = BMPC_PT_EMPTY;
    if (n > 0) {                                                     // :404-407
        float cur = 0.0f;                                            // :423
        float bx = px[o], by = py[o];
        bool found = false;
        for (int i = 1; i < n; ++i) {                                // :424
            const long long k = o + i * sp;
            const float tx = px[k], ty = py[k];
            const float ex = __fsub_rn(bx, tx), ey = __fsub_rn(by, ty);
            const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));   // :427 distance()
            cur = __fadd_rn(cur, len);
            if (cur > d) {                                           // :428 strict
                const float delta = __fsub_rn(cur, d);              // :431
                const float ax = __fmul_rn(__fdiv_rn(ex, len), delta), ay = __fmul_rn(__fdiv_rn(ey, len), delta);   // :432
                rx = __fadd_rn(tx, ax); ry = __fadd_rn(ty, ay);      // :433-434
                rdx = dx[k]; rdy = dy[k]; rv = vel[k];
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
    ((int32_t*)(base + pl.out_idx))[b] = ri;
    (base + pl.out_flag)[b] = rf;
    ((float*)(base + pl.out_px))[b] = rx;
    ((float*)(base + pl.out_py))[b] = ry;
    ((float*)(base + pl.out_dx))[b] = rdx;
    ((float*)(base + pl.out_dy))[b] = rdy;
    ((float*)(base + pl.out_vel))[b] = rv;
}

// Device address of the array on `device` (made current here).  Asked for once per device.
int static_base(int device, unsigned char** out)
{
    BMPC_CUDA(cudaSetDevice(device));
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        BMPC_CUDA(cudaGetSymbolAddress((void**)&p, g_static_scratch));
        if (!p) return bmpc::fail(BMPC_E_CUDA, "bmpc_extract_points_static: cudaGetSymbolAddress returned a null address");
        g_static_ptr[device] = (unsigned char*)p;
    }
    *out = g_static_ptr[device];
    return BMPC_OK;
}

}  // namespace

extern "C" int bmpc_extract_points_static(int device, void* cuda_stream, int64_t count, int P, const int32_t* npoints,
                                          const float* px, const float* py, const float* dx, const float* dy,
                                          const float* vel, int64_t stride_instance, int64_t stride_point,
                                          const float* dist, int32_t* idx, uint8_t* flag, float* out_px, float* out_py,
                                          float* out_dx, float* out_dy, float* out_vel)
{
    using bmpc::fail;
    if (device < 0 || device >= kMaxDevices) return fail(BMPC_E_INVALID, "bmpc_extract_points_static: bad device index");
    if (count < 1 || P < 0 || !dist || !idx || !flag || !out_px || !out_py || !out_dx || !out_dy || !out_vel)
        return fail(BMPC_E_INVALID, "bmpc_extract_points_static: bad argument");
    if (P > 0 && (!px || !py || !dx || !dy || !vel))
        return fail(BMPC_E_INVALID, "bmpc_extract_points_static: null trajectory array");
    if (stride_instance < 0 || stride_point < 0) return fail(BMPC_E_INVALID, "bmpc_extract_points_static: negative stride");

    // lay the windows out; every factor is bounded by the array size BEFORE anything is multiplied or added up
    const unsigned long long cap = BMPC_STATIC_EXTRACT_BYTES;
    const std::string too_big = "bmpc_extract_points_static: problem exceeds the static scratch array of " + std::to_string(cap) +
                                " bytes (BMPC_STATIC_EXTRACT_BYTES); use bmpc_extract_points";
    if ((unsigned long long)count > cap / 4 || (unsigned long long)P > cap / 4 ||
        (count > 1 && (unsigned long long)stride_instance > cap / 4) || (P > 1 && (unsigned long long)stride_point > cap / 4))
        return fail(BMPC_E_INVALID, too_big);
    const unsigned long long span_i = (unsigned long long)(count - 1) * (unsigned long long)stride_instance;
    const unsigned long long span_p = P > 0 ? (unsigned long long)(P - 1) * (unsigned long long)stride_point : 0;
    if (span_i > cap / 4 || span_p > cap / 4) return fail(BMPC_E_INVALID, too_big);
    const size_t nflt = P > 0 ? (size_t)(span_i + span_p + 1) : 0;
    StaticPlan pl{};
    size_t at = 0;
    auto take = [&at](size_t bytes) { const size_t o = at; at += align256(bytes); return (unsigned long long)o; };
    pl.px = take(nflt * 4 + 4); pl.py = take(nflt * 4 + 4); pl.dx = take(nflt * 4 + 4); pl.dy = take(nflt * 4 + 4);
    pl.vel = take(nflt * 4 + 4);
    pl.dist = take((size_t)count * 4);
    pl.has_npoints = npoints ? 1 : 0;
    pl.npoints = npoints ? take((size_t)count * 4) : 0;
    const size_t out_begin = at;
    pl.out_idx = take((size_t)count * 4); pl.out_px = take((size_t)count * 4); pl.out_py = take((size_t)count * 4);
    pl.out_dx = take((size_t)count * 4); pl.out_dy = take((size_t)count * 4); pl.out_vel = take((size_t)count * 4);
    pl.out_flag = take((size_t)count);
    if (at > cap)
        return fail(BMPC_E_INVALID, "bmpc_extract_points_static: needs " + std::to_string(at) +
                                        " bytes of scratch, the static array holds " + std::to_string(cap) +
                                        " (BMPC_STATIC_EXTRACT_BYTES); use bmpc_extract_points");

    std::lock_guard<std::mutex> lock(g_static_mutex);
    unsigned char* base = nullptr;
    if (int rc = static_base(device, &base)) return rc;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    BMPC_CUDA(cudaMemsetAsync(base + out_begin, 0, at - out_begin, st));           // the array keeps the last call's results
    const float* src[5] = {px, py, dx, dy, vel};
    const unsigned long long dst[5] = {pl.px, pl.py, pl.dx, pl.dy, pl.vel};
    for (int a = 0; a < 5; ++a)
        if (nflt) BMPC_CUDA(cudaMemcpyAsync(base + dst[a], src[a], nflt * 4, cudaMemcpyHostToDevice, st));
    BMPC_CUDA(cudaMemcpyAsync(base + pl.dist, dist, (size_t)count * 4, cudaMemcpyHostToDevice, st));
    if (npoints) BMPC_CUDA(cudaMemcpyAsync(base + pl.npoints, npoints, (size_t)count * 4, cudaMemcpyHostToDevice, st));
    const unsigned blocks = (unsigned)((count + 127) / 128);
    k_extract_static<<<blocks, 128, 0, st>>>(base, pl, (long long)count, P, (long long)stride_instance, (long long)stride_point);
    BMPC_CUDA(cudaGetLastError());
    BMPC_CUDA(cudaMemcpyAsync(idx, base + pl.out_idx, (size_t)count * 4, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(flag, base + pl.out_flag, (size_t)count, cudaMemcpyDeviceToHost, st));
    float* outs[5] = {out_px, out_py, out_dx, out_dy, out_vel};
    const unsigned long long from[5] = {pl.out_px, pl.out_py, pl.out_dx, pl.out_dy, pl.out_vel};
    for (int a = 0; a < 5; ++a)
        BMPC_CUDA(cudaMemcpyAsync(outs[a], base + from[a], (size_t)count * 4, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaStreamSynchronize(st));
    return BMPC_OK;
}
