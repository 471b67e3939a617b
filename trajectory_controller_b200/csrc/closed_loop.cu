// libbmpc.so — closed loop on the device (include/bmpc.h "closed loop"): per tick
//   k_cl_prepare : world path -> vehicle-frame polyline -> getTrajectoryPoint (src/trajectory_point_follower.cpp:392-443)
//                  -> target (src/...:84-85,368-371) written straight into the solver's target / x0 arrays
//   solve_kernel : persistent-controller solve (BMPC_WARM)
//   k_cl_plant   : harness plant step with u0, iteration statistics
// Nothing crosses PCIe inside bmpc_cl_run.  The harness arithmetic (path sampling, atan, plant) is specified in
// include/bmpc.h; every operation below is a separately rounded IEEE operation (no FMA) so that the CPU restatement
// the tests check against reproduces it bit for bit.
#include <cmath>
#include <cstring>
#include <vector>

#include "bmpc_internal.h"

namespace bmpc {
namespace {

__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double ddiv(double a, double b) { return __ddiv_rn(a, b); }

// bmpc_harness_atan (include/bmpc.h): two half-angle reductions, then the odd series to z^11 by Horner.
__device__ double harness_atan(double t)
{
    const double t1 = ddiv(t, dadd(1.0, __dsqrt_rn(dadd(1.0, dmul(t, t)))));
    const double t2 = ddiv(t1, dadd(1.0, __dsqrt_rn(dadd(1.0, dmul(t1, t1)))));
    const double z = dmul(t2, t2);
    double p = ddiv(1.0, 23.0);
#pragma unroll
    for (int k = 21; k >= 1; k -= 2) p = dsub(ddiv(1.0, (double)k), dmul(z, p));
    return dmul(4.0, dmul(t2, p));
}

struct ClParams {
    long long count, stride;
    int P, npath;
    float lookahead, vel;
    const double *path_x, *path_y, *path_tx, *path_ty;
    const double* pose;      // [4][stride]
    double* tgt_dbg;         // [2][stride]
    int32_t* idx_dbg;
    // a target per horizon step (bmpc_cl_set_horizon_targets): N steps at d_t = h_d0 + t*h_step, written to the ring
    // [N][4][stride] whose physical slot of step 0 is `head`
    int N, head;
    float h_d0, h_step;
};

// One thread per vehicle.  The arc-length scan is sequential per polyline (float sum, reference :423-427), so a
// thread walks its P points in order; the world path is shared by all vehicles and stays in L1/L2.
template <typename T>
__global__ void k_cl_prepare(ClParams q, T* __restrict__ x0, T* __restrict__ target)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= q.count) return;
    const double X = q.pose[b], Y = q.pose[q.stride + b], c = q.pose[2 * q.stride + b], s = q.pose[3 * q.stride + b];
    // k0 = first index with path_x[k] > X (binary search on the ascending x), clamped so that P points exist
    int lo = 0, hi = q.npath;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (q.path_x[mid] > X) hi = mid; else lo = mid + 1; }
    int k0 = lo;
    if (k0 > q.npath - q.P) k0 = q.npath - q.P;
    if (k0 < 0) k0 = 0;
    const int n = q.P < q.npath - k0 ? q.P : q.npath - k0;

    auto vx = [&](int k) { const double dx = dsub(q.path_x[k], X), dy = dsub(q.path_y[k], Y); return (float)dadd(dmul(c, dx), dmul(s, dy)); };
    auto vy = [&](int k) { const double dx = dsub(q.path_x[k], X), dy = dsub(q.path_y[k], Y); return (float)dsub(dmul(c, dy), dmul(s, dx)); };

    const float d = q.lookahead;
    float rx = d, ry = 0.0f;                       // :394-403 default point (empty polyline)
    int ri = -1, src = -1;
    if (n > 0) {
        float cur = 0.0f, bx = vx(k0), by = vy(k0);
        bool found = false;
        for (int i = 1; i < n; ++i) {                                    // :424
            const float tx = vx(k0 + i), ty = vy(k0 + i);
            const float ex = __fsub_rn(bx, tx), ey = __fsub_rn(by, ty);
            const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));   // :427
            cur = __fadd_rn(cur, len);
            if (cur > d) {                                               // :428
                const float delta = __fsub_rn(cur, d);
                rx = __fadd_rn(tx, __fmul_rn(__fdiv_rn(ex, len), delta));                    // :432-434
                ry = __fadd_rn(ty, __fmul_rn(__fdiv_rn(ey, len), delta));
                ri = i; src = k0 + i; found = true;
                break;
            }
            bx = tx; by = ty;
        }
        if (!found) { rx = vx(k0 + n - 1); ry = vy(k0 + n - 1); ri = n - 1; src = k0 + n - 1; }   // :439-442
    }
    float fdx = 1.0f, fdy = 0.0f;
    if (src >= 0) {
        const double tx = q.path_tx[src], ty = q.path_ty[src];
        fdx = (float)dadd(dmul(c, tx), dmul(s, ty));
        fdy = (float)dsub(dmul(c, ty), dmul(s, tx));
    }
    const double y_soll = (double)ry;                                    // :85
    double phi;                                                          // :84 (harness atan instead of libm atan2)
    if (fdx > 0.0f) phi = harness_atan(ddiv((double)fdy, (double)fdx));
    else phi = fdy >= 0.0f ? 1.5707963267948966 : -1.5707963267948966;
    (void)rx;
    target[b] = (T)y_soll; target[q.stride + b] = (T)phi; target[2 * q.stride + b] = T(0); target[3 * q.stride + b] = T(0);
    x0[b] = T(0); x0[q.stride + b] = T(0);                               // current_state = (0, 0, delta, ddelta)  :377-379
    q.tgt_dbg[b] = y_soll; q.tgt_dbg[q.stride + b] = phi;
    q.idx_dbg[b] = ri;
}

// The same tick preparation with a target per horizon step: one sequential scan of the
// vehicle-frame polyline yields the look-ahead point of every step (see k_extract_horizon, layout_kernels.cu — the point
// of step t is getTrajectoryPoint(d_t)), and target[t] = (point_t.y, atan(dir_t.y/dir_t.x), 0, 0) goes straight into
// the solver's target ring (dlib: set_target(val, t), mpc.h:142-155).
template <typename T>
__global__ void k_cl_prepare_horizon(ClParams q, T* __restrict__ x0, T* __restrict__ target)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= q.count) return;
    const double X = q.pose[b], Y = q.pose[q.stride + b], c = q.pose[2 * q.stride + b], s = q.pose[3 * q.stride + b];
    int lo = 0, hi = q.npath;
    while (lo < hi) { const int mid = (lo + hi) >> 1; if (q.path_x[mid] > X) hi = mid; else lo = mid + 1; }
    int k0 = lo;
    if (k0 > q.npath - q.P) k0 = q.npath - q.P;
    if (k0 < 0) k0 = 0;
    const int n = q.P < q.npath - k0 ? q.P : q.npath - k0;

    auto vx = [&](int k) { const double dx = dsub(q.path_x[k], X), dy = dsub(q.path_y[k], Y); return (float)dadd(dmul(c, dx), dmul(s, dy)); };
    auto vy = [&](int k) { const double dx = dsub(q.path_x[k], X), dy = dsub(q.path_y[k], Y); return (float)dsub(dmul(c, dy), dmul(s, dx)); };
    auto dist_of = [&](int t) { return __fadd_rn(q.h_d0, __fmul_rn((float)t, q.h_step)); };
    // target of step t from the point's y and the world tangent at path index src (src < 0: the default direction (1, 0))
    auto put = [&](int t, float ry, int src, int ri) {
        float fdx = 1.0f, fdy = 0.0f;
        if (src >= 0) {
            const double tx = q.path_tx[src], ty = q.path_ty[src];
            fdx = (float)dadd(dmul(c, tx), dmul(s, ty));
            fdy = (float)dsub(dmul(c, ty), dmul(s, tx));
        }
        const double y_soll = (double)ry;
        double phi;
        if (fdx > 0.0f) phi = harness_atan(ddiv((double)fdy, (double)fdx));
        else phi = fdy >= 0.0f ? 1.5707963267948966 : -1.5707963267948966;
        int ph = q.head + t; if (ph >= q.N) ph -= q.N;
        T* row = target + (long long)ph * 4 * q.stride + b;
        row[0] = (T)y_soll; row[q.stride] = (T)phi; row[2 * q.stride] = T(0); row[3 * q.stride] = T(0);
        if (t == 0) { q.tgt_dbg[b] = y_soll; q.tgt_dbg[q.stride + b] = phi; q.idx_dbg[b] = ri; }
    };
    int t = 0;
    if (n <= 0) {
        for (; t < q.N; ++t) put(t, 0.0f, -1, -1);
    } else {
        float cur = 0.0f, bx = vx(k0), by = vy(k0);
        float d = dist_of(0);
        for (int i = 1; i < n && t < q.N; ++i) {
            const float tx = vx(k0 + i), ty = vy(k0 + i);
            const float ex = __fsub_rn(bx, tx), ey = __fsub_rn(by, ty);
            const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)));
            cur = __fadd_rn(cur, len);
            if (cur > d) {
                const float ny = __fdiv_rn(ey, len);
                do {
                    const float delta = __fsub_rn(cur, d);
                    put(t, __fadd_rn(ty, __fmul_rn(ny, delta)), k0 + i, i);
                    ++t;
                    d = dist_of(t);
                } while (t < q.N && cur > d);
            }
            bx = tx; by = ty;
        }
        if (t < q.N) {
            const float ly = vy(k0 + n - 1);
            for (; t < q.N; ++t) put(t, ly, k0 + n - 1, n - 1);
        }
    }
    x0[b] = T(0); x0[q.stride + b] = T(0);
}

struct PlantParams {
    long long count, stride;
    double a22, a23, a32, a33, b2, b3, Tv, Tv_l;
    double* pose;
    const uint32_t* iters;
    unsigned long long* itot;
    uint32_t* imax;
};

template <typename T>
__global__ void k_cl_plant(PlantParams q, T* __restrict__ x0, const T* __restrict__ u0)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= q.count) return;
    const double u = (double)u0[b];
    const double de = (double)x0[2 * q.stride + b], dd = (double)x0[3 * q.stride + b];
    const double de2 = dadd(dadd(dmul(q.a22, de), dmul(q.a23, dd)), dmul(q.b2, u));
    const double dd2 = dadd(dadd(dmul(q.a32, de), dmul(q.a33, dd)), dmul(q.b3, u));
    x0[2 * q.stride + b] = (T)de2; x0[3 * q.stride + b] = (T)dd2;
    const double X = q.pose[b], Y = q.pose[q.stride + b], c = q.pose[2 * q.stride + b], s = q.pose[3 * q.stride + b];
    const double dpsi = dmul(q.Tv_l, de);
    const double c1 = dsub(c, dmul(s, dpsi)), s1 = dadd(s, dmul(c, dpsi));
    const double nrm = __dsqrt_rn(dadd(dmul(c1, c1), dmul(s1, s1)));
    q.pose[b] = dadd(X, dmul(q.Tv, c));
    q.pose[q.stride + b] = dadd(Y, dmul(q.Tv, s));
    q.pose[2 * q.stride + b] = ddiv(c1, nrm);
    q.pose[3 * q.stride + b] = ddiv(s1, nrm);
    const uint32_t it = q.iters[b];
    q.itot[b] += it;
    if (it > q.imax[b]) q.imax[b] = it;
}

}  // namespace
}  // namespace bmpc

using namespace bmpc;

extern "C" {

int bmpc_cl_init(bmpc_t* h, int64_t count, const bmpc_cl_config* cfg, const double* path_x, const double* path_y,
                 const double* path_tx, const double* path_ty, const double* pose0, const double* act0)
{
    if (!h || !cfg || !path_x || !path_y || !path_tx || !path_ty || !pose0) return fail(BMPC_E_INVALID, "bmpc_cl_init: null argument");
    if (h->S != 4 || h->I != 1) return fail(BMPC_E_UNSUPPORTED, "bmpc_cl_init: the closed-loop harness is defined for S = 4, I = 1");
    if (!h->have_model || h->per_instance) return fail(BMPC_E_STATE, "bmpc_cl_init: needs a shared model (bmpc_set_model, per_instance = 0)");
    if (count < 1 || count > h->cap) return fail(BMPC_E_INVALID, "bmpc_cl_init: count out of capacity");
    if (cfg->P < 1 || cfg->npath < cfg->P) return fail(BMPC_E_INVALID, "bmpc_cl_init: need 1 <= P <= npath");
    BMPC_CUDA(cudaSetDevice(h->device));
    int rc = bmpc_reset(h);                                 // fresh controllers: controls = 0, targets = 0, v = 0
    if (rc) return rc;
    h->full_targets = false;                                // the loop uses set_target(val): one target for all steps
    h->cl_horizon = false;                                  // ... unless bmpc_cl_set_horizon_targets asks for one per step
    h->cl_cfg = *cfg; h->cl_count = count;
    const size_t np = (size_t)cfg->npath, cap = (size_t)h->cap;
    BMPC_CUDA(h->cl_path.reserve(4 * np * 8));
    const double* src[4] = {path_x, path_y, path_tx, path_ty};
    for (int a = 0; a < 4; ++a)
        BMPC_CUDA(cudaMemcpyAsync((char*)h->cl_path.p + a * np * 8, src[a], np * 8, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(h->cl_pose.reserve(4 * cap * 8));
    BMPC_CUDA(h->cl_tgt.alloc_zero(2 * cap * 8, h->stream));
    BMPC_CUDA(h->cl_idx.alloc_zero(cap * 4, h->stream));
    BMPC_CUDA(h->cl_itot.alloc_zero(cap * 8, h->stream));
    BMPC_CUDA(h->cl_imax.alloc_zero(cap * 4, h->stream));
    // pose0 [count][4] -> [4][cap]; act0 [count][2] -> x0 rows 2, 3 (rows 0, 1 are written every tick)
    BMPC_CUDA(h->stage.reserve((size_t)count * 4 * 8));
    BMPC_CUDA(cudaMemcpyAsync(h->stage.p, pose0, (size_t)count * 4 * 8, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(aos_to_soa((const double*)h->stage.p, h->cl_pose.p, count, 4, h->cap, 8, h->stream));
    h->launches++;
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    BMPC_CUDA(cudaMemsetAsync(h->x0.p, 0, h->x0.bytes, h->stream));
    if (act0) {
        BMPC_CUDA(cudaMemcpyAsync(h->stage.p, act0, (size_t)count * 2 * 8, cudaMemcpyHostToDevice, h->stream));
        BMPC_CUDA(aos_to_soa((const double*)h->stage.p, (char*)h->x0.p + 2 * cap * h->elem, count, 2, h->cap, h->elem, h->stream));
        h->launches++;
    }
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    h->cl_ready = true;
    return BMPC_OK;
}

int bmpc_cl_set_horizon_targets(bmpc_t* h, int enable, double d0, double step)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_cl_set_horizon_targets: null handle");
    if (!h->cl_ready) return fail(BMPC_E_STATE, "bmpc_cl_set_horizon_targets: bmpc_cl_init has not been called");
    if (enable && !((float)step >= 0.0f && std::isfinite(d0))) return fail(BMPC_E_INVALID, "bmpc_cl_set_horizon_targets: need a finite d0 and step >= 0");
    BMPC_CUDA(cudaSetDevice(h->device));
    if (enable) {
        // every tick rewrites all N steps of every vehicle, so the ring only has to exist
        const size_t bytes = (size_t)h->cap * h->S * h->elem * h->N;
        BMPC_CUDA(h->target_full.reserve(bytes));
        BMPC_CUDA(cudaMemsetAsync(h->target_full.p, 0, bytes, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
        h->head = 0;
    }
    h->full_targets = enable != 0;
    h->cl_horizon = enable != 0;
    h->cl_h_d0 = (float)d0; h->cl_h_step = (float)step;
    return BMPC_OK;
}

int bmpc_cl_run(bmpc_t* h, int ticks)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_cl_run: null handle");
    if (!h->cl_ready) return fail(BMPC_E_STATE, "bmpc_cl_run: bmpc_cl_init has not been called");
    if (ticks < 0) return fail(BMPC_E_INVALID, "bmpc_cl_run: ticks < 0");
    BMPC_CUDA(cudaSetDevice(h->device));
    const long long count = h->cl_count;
    const size_t np = (size_t)h->cl_cfg.npath;
    const double* path = (const double*)h->cl_path.p;
    ClParams cp{count, h->cap, h->cl_cfg.P, h->cl_cfg.npath, (float)h->cl_cfg.lookahead, (float)h->cl_cfg.v,
                path, path + np, path + 2 * np, path + 3 * np, (const double*)h->cl_pose.p, (double*)h->cl_tgt.p,
                (int32_t*)h->cl_idx.p, h->N, 0, h->cl_h_d0, h->cl_h_step};
    const double Tv = h->cl_cfg.T * h->cl_cfg.v;
    PlantParams pp{count, h->cap, h->hA[10], h->hA[11], h->hA[14], h->hA[15], h->hB[2], h->hB[3], Tv, Tv / h->cl_cfg.l,
                   (double*)h->cl_pose.p, (const uint32_t*)h->iters.p, (unsigned long long*)h->cl_itot.p, (uint32_t*)h->cl_imax.p};
    const unsigned grid = (unsigned)((count + 127) / 128);
    for (int t = 0; t < ticks; ++t) {
        if (h->cl_horizon) {
            cp.head = h->head;                              // the ring's step-0 slot moves by one per warm solve (mpc.h:236-237)
            if (h->elem == 8) k_cl_prepare_horizon<double><<<grid, 128, 0, h->stream>>>(cp, (double*)h->x0.p, (double*)h->target_full.p);
            else k_cl_prepare_horizon<float><<<grid, 128, 0, h->stream>>>(cp, (float*)h->x0.p, (float*)h->target_full.p);
        } else if (h->elem == 8) k_cl_prepare<double><<<grid, 128, 0, h->stream>>>(cp, (double*)h->x0.p, (double*)h->target_const.p);
        else k_cl_prepare<float><<<grid, 128, 0, h->stream>>>(cp, (float*)h->x0.p, (float*)h->target_const.p);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
        int rc = bmpc_solve_resident_async(h, count, BMPC_WARM);
        if (rc) return rc;
        if (h->elem == 8) k_cl_plant<double><<<grid, 128, 0, h->stream>>>(pp, (double*)h->x0.p, (const double*)h->u0.p);
        else k_cl_plant<float><<<grid, 128, 0, h->stream>>>(pp, (float*)h->x0.p, (const float*)h->u0.p);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
    }
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_cl_get(bmpc_t* h, int64_t count, double* pose, double* act, double* u0, double* target, int32_t* idx,
                uint64_t* iters_total, uint32_t* iters_max)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_cl_get: null handle");
    if (!h->cl_ready) return fail(BMPC_E_STATE, "bmpc_cl_get: bmpc_cl_init has not been called");
    if (count < 1 || count > h->cl_count) return fail(BMPC_E_INVALID, "bmpc_cl_get: count out of range");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t cap = (size_t)h->cap;
    BMPC_CUDA(h->stage.reserve((size_t)count * 4 * 8));
    auto fetch = [&](const void* soa, int len, int elem, double* out) -> int {
        if (!out) return BMPC_OK;
        BMPC_CUDA(soa_to_aos(soa, (double*)h->stage.p, count, len, h->cap, elem, h->stream));
        h->launches++;
        BMPC_CUDA(cudaMemcpyAsync(out, h->stage.p, (size_t)count * len * 8, cudaMemcpyDeviceToHost, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
        return BMPC_OK;
    };
    int rc;
    if ((rc = fetch(h->cl_pose.p, 4, 8, pose))) return rc;
    if ((rc = fetch((const char*)h->x0.p + 2 * cap * h->elem, 2, h->elem, act))) return rc;
    if ((rc = fetch(h->u0.p, 1, h->elem, u0))) return rc;
    if ((rc = fetch(h->cl_tgt.p, 2, 8, target))) return rc;
    if (idx) BMPC_CUDA(cudaMemcpyAsync(idx, h->cl_idx.p, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    if (iters_total) BMPC_CUDA(cudaMemcpyAsync(iters_total, h->cl_itot.p, (size_t)count * 8, cudaMemcpyDeviceToHost, h->stream));
    if (iters_max) BMPC_CUDA(cudaMemcpyAsync(iters_max, h->cl_imax.p, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

}  // extern "C"
