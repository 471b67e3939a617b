// libbmpc.so — one tobiMPC control tick of TrajectoryPointController::cycle() for a FLEET of vehicles, resident on the
// device.  Per vehicle, /root/reference/src/trajectory_point_follower.cpp:
//   :66-76    look-ahead distance (evaluated by the caller) -> getTrajectoryPoint (k_extract, :392-443)
//   :77-85    |v| < 0.1 -> v = 0.1;  phi_soll = atan2(dir.y, dir.x);  y_soll = position.y
//   :301-389  mpcControllerTobi: A = [1 Tv; 0 1], B = [0 Tv; Tv/l -Tv/l], C = 0, Q = (w_y, w_phi), R = (w_sf, w_sr),
//             bounds -+22 deg (:16-18), a FRESH dlib::mpc<2,2,4> (constructor math on the device, ctor_kernel.cuh),
//             set_target((y_soll, phi_soll)), operator()((0,0))      -> cold batched solve
//   :114-125  CarCommand fields;  :276-283 targetSpeed < 0.5 -> both steering angles 0
// Kernels: k_extract -> k_tick_prepare -> ctor_kernel -> solve (stage A / stage B) -> k_tick_finish; nothing crosses
// PCIe between them.  Every operation is a separately rounded IEEE operation; the one function that cannot be made
// bit-identical to the reference's host is atan2 (CUDA's double atan2, <= 2 ulp, against glibc's): see bmpc.h.
#include <cmath>

#include "bmpc_internal.h"

namespace bmpc {
namespace {

struct TickParams {
    long long count, stride;
    double T, l, w_y, w_phi, w_sf, w_sr, alpha_max;
    const float* velocity;      // [count] car->velocity()
    const double* v_model;      // [count] result of m_mpcLookupVelocity.linearSearch(v) (:323), or null = v itself
    const float *opx, *opy, *odx, *ody;   // extracted trajectory point
};

template <typename T>
__global__ void k_tick_prepare(TickParams q, T* __restrict__ mA, T* __restrict__ mB, T* __restrict__ mC,
                               T* __restrict__ mQ, T* __restrict__ mR, T* __restrict__ mlo, T* __restrict__ mhi,
                               T* __restrict__ x0, T* __restrict__ target, double* __restrict__ tgt_dbg)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= q.count) return;
    const long long s = q.stride;
    double v = (double)q.velocity[b];                                   // :77
    if (fabs(v) < 0.1) v = 0.1;                                         // :78-82
    const double phi_soll = atan2((double)q.ody[b], (double)q.odx[b]); // :84 (float components widened first)
    const double y_soll = (double)q.opy[b];                             // :85
    if (q.v_model) v = q.v_model[b];                                    // :323
    const double Tv = __dmul_rn(q.T, v);                                // T*v
    const double Tvl = __ddiv_rn(Tv, q.l);                              // T*v/l
    mA[0 * s + b] = T(1); mA[1 * s + b] = T(Tv); mA[2 * s + b] = T(0); mA[3 * s + b] = T(1);       // :326-327
    mB[0 * s + b] = T(0); mB[1 * s + b] = T(Tv); mB[2 * s + b] = T(Tvl); mB[3 * s + b] = T(-Tvl);  // :329-330 (-T*v/l)
    mC[0 * s + b] = T(0); mC[1 * s + b] = T(0);                                                     // :332-333
    mQ[0 * s + b] = T(q.w_y); mQ[1 * s + b] = T(q.w_phi);                                           // :336-337
    mR[0 * s + b] = T(q.w_sf); mR[1 * s + b] = T(q.w_sr);                                           // :339-340
    mlo[0 * s + b] = T(-q.alpha_max); mlo[1 * s + b] = T(-q.alpha_max);                             // :17
    mhi[0 * s + b] = T(q.alpha_max); mhi[1 * s + b] = T(q.alpha_max);                               // :18
    target[0 * s + b] = T(y_soll); target[1 * s + b] = T(phi_soll);                                  // :368-371
    x0[0 * s + b] = T(0); x0[1 * s + b] = T(0);                                                       // :374-375
    tgt_dbg[b] = y_soll; tgt_dbg[s + b] = phi_soll;
}

template <typename T>
__global__ void k_tick_finish(long long count, long long stride, const T* __restrict__ u0, const float* __restrict__ opx,
                              const float* __restrict__ opy, const float* __restrict__ ovel, double* __restrict__ sf,
                              double* __restrict__ sr, float* __restrict__ tspeed, float* __restrict__ tdist,
                              uint8_t* __restrict__ state)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    double f = (double)u0[b], r = (double)u0[stride + b];              // :384-385 action(0,0), action(1,0)
    const float sp = ovel[b];                                           // :116
    const float x = opx[b], y = opy[b];
    tdist[b] = __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));   // :117 position.length()
    state[b] = sp == 0.0f ? BMPC_CAR_IDLE : BMPC_CAR_DRIVING;          // :121-125
    if (sp < 0.5f) { f = 0.0; r = 0.0; }                               // :276-283
    sf[b] = f; sr[b] = r; tspeed[b] = sp;
}

}  // namespace

const ShapeOps* find_pattern(int S, int I, int N, const char* name);   // bmpc_core.cu
}  // namespace bmpc

using namespace bmpc;

extern "C" int bmpc_fleet_tick_tobi(bmpc_t* h, int64_t count, const bmpc_tick_config* cfg, const float* velocity,
                                    const double* v_model, const float* dist, int P, const int32_t* npoints,
                                    const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                                    int64_t stride_instance, int64_t stride_point, double* steering_front,
                                    double* steering_rear, float* target_speed, float* target_distance, uint8_t* state,
                                    int32_t* idx, uint8_t* flag, double* target, uint32_t* iters)
{
    if (!h || !cfg || !velocity || !dist || !steering_front || !steering_rear || !target_speed || !target_distance || !state)
        return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: null argument");
    if (h->S != 2 || h->I != 2) return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: the handle must be a (2,2,N) controller (mpcControllerTobi, src/...:303-366)");
    if (count < 1 || count > h->cap || P < 0) return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: count / P out of range");
    if (P > 0 && (!px || !py || !dx || !dy || !vel)) return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: null trajectory array");
    if (stride_instance < 0 || stride_point < 0) return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: negative stride");
    BMPC_CUDA(cudaSetDevice(h->device));
    cudaStream_t st = h->stream;
    const size_t c = (size_t)count, es = (size_t)h->elem, cap = (size_t)h->cap;

    // ---- trajectory point (same buffers and kernel as bmpc_extract_points, results stay on the device) -------------
    const size_t nflt = P > 0 ? (size_t)((count - 1) * stride_instance + (int64_t)(P - 1) * stride_point + 1) : 0;
    const float* src[5] = {px, py, dx, dy, vel};
    for (int a = 0; a < 5; ++a) {
        BMPC_CUDA(h->ex[a].reserve(nflt * 4 + 4));
        if (nflt) BMPC_CUDA(cudaMemcpyAsync(h->ex[a].p, src[a], nflt * 4, cudaMemcpyHostToDevice, st));
    }
    BMPC_CUDA(h->ex[5].reserve(c * 4));
    BMPC_CUDA(cudaMemcpyAsync(h->ex[5].p, dist, c * 4, cudaMemcpyHostToDevice, st));
    if (npoints) {
        BMPC_CUDA(h->ex[6].reserve(c * 4));
        BMPC_CUDA(cudaMemcpyAsync(h->ex[6].p, npoints, c * 4, cudaMemcpyHostToDevice, st));
    }
    for (int a = 7; a < 14; ++a) BMPC_CUDA(h->ex[a].reserve(c * 4));
    BMPC_CUDA(launch_extract(count, P, npoints ? (const int32_t*)h->ex[6].p : nullptr, (const float*)h->ex[0].p,
                             (const float*)h->ex[1].p, (const float*)h->ex[2].p, (const float*)h->ex[3].p,
                             (const float*)h->ex[4].p, stride_instance, stride_point, (const float*)h->ex[5].p,
                             (int32_t*)h->ex[7].p, (uint8_t*)h->ex[8].p, (float*)h->ex[9].p, (float*)h->ex[10].p,
                             (float*)h->ex[11].p, (float*)h->ex[12].p, (float*)h->ex[13].p, st));
    h->launches++;

    // ---- per-vehicle model, target and zero state ---------------------------------------------------------------------
    const size_t voff = ((c * 4 + 7) / 8) * 8;
    BMPC_CUDA(h->ex[14].reserve(voff + c * 8));                           // velocity (float) | v_model (double)
    BMPC_CUDA(h->ex[15].reserve(cap * 8 * 2 + c * (8 + 8 + 4 + 4 + 1)));   // tgt_dbg [2][cap] | outputs
    BMPC_CUDA(cudaMemcpyAsync(h->ex[14].p, velocity, c * 4, cudaMemcpyHostToDevice, st));
    double* d_vmodel = nullptr;
    if (v_model) {
        d_vmodel = (double*)((char*)h->ex[14].p + voff);
        BMPC_CUDA(cudaMemcpyAsync(d_vmodel, v_model, c * 8, cudaMemcpyHostToDevice, st));
    }
    DevBuf* mb[7] = {&h->mA, &h->mB, &h->mC, &h->mQ, &h->mR, &h->mlo, &h->mhi};
    const int mlen[7] = {4, 4, 2, 2, 2, 2, 2};
    for (int a = 0; a < 7; ++a) BMPC_CUDA(mb[a]->reserve(cap * mlen[a] * es));
    BMPC_CUDA(h->mlambda.reserve(cap * es));
    BMPC_CUDA(h->mqdiag.reserve(cap * h->NI * es));
    TickParams q{};
    q.count = count; q.stride = h->cap; q.T = cfg->T; q.l = cfg->l; q.w_y = cfg->weight_y; q.w_phi = cfg->weight_phi;
    q.w_sf = cfg->weight_steering_front; q.w_sr = cfg->weight_steering_rear; q.alpha_max = cfg->alpha_max;
    q.velocity = (const float*)h->ex[14].p; q.v_model = d_vmodel;
    q.opx = (const float*)h->ex[9].p; q.opy = (const float*)h->ex[10].p; q.odx = (const float*)h->ex[11].p; q.ody = (const float*)h->ex[12].p;
    double* tgt_dbg = (double*)h->ex[15].p;
    const unsigned grid = (unsigned)((count + 127) / 128);
    if (h->elem == 8)
        k_tick_prepare<double><<<grid, 128, 0, st>>>(q, (double*)h->mA.p, (double*)h->mB.p, (double*)h->mC.p, (double*)h->mQ.p,
                                                     (double*)h->mR.p, (double*)h->mlo.p, (double*)h->mhi.p, (double*)h->x0.p,
                                                     (double*)h->target_const.p, tgt_dbg);
    else
        k_tick_prepare<float><<<grid, 128, 0, st>>>(q, (float*)h->mA.p, (float*)h->mB.p, (float*)h->mC.p, (float*)h->mQ.p,
                                                    (float*)h->mR.p, (float*)h->mlo.p, (float*)h->mhi.p, (float*)h->x0.p,
                                                    (float*)h->target_const.p, tgt_dbg);
    BMPC_CUDA(cudaGetLastError());
    h->launches++;

    // ---- fresh controllers: constructor recursion (mpc.h:114-123), then one cold operator() each ------------------------
    GenericCtor g{count, h->cap, h->mA.p, h->mB.p, h->mQ.p, h->mR.p, h->mlambda.p, h->mqdiag.p};
    BMPC_CUDA(h->ops->ctor(h->precision, h->mode, g, st));
    h->launches++;
    h->per_instance = true; h->model_count = count; h->have_model = true;
    h->full_targets = false; h->head = 0;                         // set_target(val): one target for all horizon steps
    const ShapeOps* tobi = h->specialize ? find_pattern(h->S, h->I, h->N, "tobi") : nullptr;   // holds by construction
    h->solve_ops = tobi ? tobi : h->ops;
    int rc = bmpc_solve_resident_async(h, count, 0);
    if (rc) return rc;

    // ---- CarCommand fields ----------------------------------------------------------------------------------------------
    char* ob = (char*)h->ex[15].p + cap * 8 * 2;
    double* d_sf = (double*)ob; double* d_sr = d_sf + c;
    float* d_ts = (float*)(d_sr + c); float* d_td = d_ts + c;
    uint8_t* d_state = (uint8_t*)(d_td + c);
    if (h->elem == 8)
        k_tick_finish<double><<<grid, 128, 0, st>>>(count, h->cap, (const double*)h->u0.p, (const float*)h->ex[9].p,
                                                    (const float*)h->ex[10].p, (const float*)h->ex[13].p, d_sf, d_sr, d_ts, d_td, d_state);
    else
        k_tick_finish<float><<<grid, 128, 0, st>>>(count, h->cap, (const float*)h->u0.p, (const float*)h->ex[9].p,
                                                   (const float*)h->ex[10].p, (const float*)h->ex[13].p, d_sf, d_sr, d_ts, d_td, d_state);
    BMPC_CUDA(cudaGetLastError());
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(steering_front, d_sf, c * 8, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(steering_rear, d_sr, c * 8, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(target_speed, d_ts, c * 4, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(target_distance, d_td, c * 4, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(state, d_state, c, cudaMemcpyDeviceToHost, st));
    if (idx) BMPC_CUDA(cudaMemcpyAsync(idx, h->ex[7].p, c * 4, cudaMemcpyDeviceToHost, st));
    if (flag) BMPC_CUDA(cudaMemcpyAsync(flag, h->ex[8].p, c, cudaMemcpyDeviceToHost, st));
    if (iters) BMPC_CUDA(cudaMemcpyAsync(iters, h->iters.p, c * 4, cudaMemcpyDeviceToHost, st));
    if (target) {                                                  // [count][2] = y_soll, phi_soll
        BMPC_CUDA(h->stage.reserve(c * 2 * 8));
        BMPC_CUDA(soa_to_aos(tgt_dbg, (double*)h->stage.p, count, 2, h->cap, 8, st));
        h->launches++;
        BMPC_CUDA(cudaMemcpyAsync(target, h->stage.p, c * 2 * 8, cudaMemcpyDeviceToHost, st));
    }
    BMPC_CUDA(cudaStreamSynchronize(st));
    return BMPC_OK;
}
