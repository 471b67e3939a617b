// libbmpc.so — one control tick of TrajectoryPointController::cycle() for a FLEET of vehicles, resident on the device
// (SURVEY.md §8(f) ranks 1 and 3).  Per vehicle, /root/reference/src/trajectory_point_follower.cpp:
//   :66-76    look-ahead distance (caller-side, or LookupTable / FOH formula on the device: k_lookahead)
//             -> getTrajectoryPoint (k_extract, :392-443) -> the "stop at crossing" clamp (k_crossing, :445-473)
//   tobiMPC branch (bmpc_fleet_tick_tobi):
//   :77-85    |v| < 0.1 -> v = 0.1;  phi_soll = atan2(dir.y, dir.x);  y_soll = position.y
//   :301-389  mpcControllerTobi: v = lookup(v) (:323), A = [1 Tv; 0 1], B = [0 Tv; Tv/l -Tv/l], C = 0, Q = (w_y, w_phi),
//             R = (w_sf, w_sr), bounds -+22 deg (:16-18), a FRESH dlib::mpc<2,2,4> (constructor math on the device,
//             ctor_kernel.cuh), set_target((y_soll, phi_soll)), operator()((0,0))      -> cold batched solve
//   :114-125  CarCommand fields;  :276-283 targetSpeed < 0.5 -> both steering angles 0
//   PID branch (bmpc_fleet_tick_pid): :214-225 steering = pid(position.y), pid(directory.angle())
// Kernels: [k_lookahead] -> k_extract -> [k_crossing] -> k_tick_prepare -> ctor_kernel -> solve (stage A / stage B) ->
// k_tick_finish; nothing crosses PCIe between them.  Every operation is a separately rounded IEEE operation; the
// functions that cannot be made bit-identical to the reference's host are atan2 / atan2f (CUDA's, <= 2 ulp, against
// glibc's): see bmpc.h.  PID / LookupTable semantics: include/bmpc.h (sensor_utils, lms_math are not vendored).
#include <cmath>
#include <cstdlib>
#include <string>

#include "bmpc_internal.h"

namespace bmpc {
namespace {

// lms::math::LookupTable<float, ASC>::linearSearch (float, every operation rounded on its own)
__device__ __forceinline__ float lut_eval(const float* __restrict__ t, int n, float x)
{
    if (n <= 0) return x;
    const float* vx = t; const float* vy = t + n;
    if (x <= vx[0]) return vy[0];
    for (int i = 1; i < n; ++i)
        if (x <= vx[i])
            return __fadd_rn(vy[i - 1], __fdiv_rn(__fmul_rn(__fsub_rn(vy[i], vy[i - 1]), __fsub_rn(x, vx[i - 1])),
                                                  __fsub_rn(vx[i], vx[i - 1])));
    return vy[n - 1];
}

// :66 / :70-73 look-ahead distance per vehicle; also the plain table evaluation of bmpc_lookup (foh = 0)
__global__ void k_lookahead(long long count, const float* __restrict__ x, const float* __restrict__ table, int n, int foh,
                            float foh_min, float foh_slope, float* __restrict__ y)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    const float v = x[b];
    y[b] = foh ? __fadd_rn(foh_min, __fmul_rn(v, foh_slope)) : lut_eval(table, n, v);
}

struct PidGains { float kp, ki, kd, dt; };
// sensor_utils::PID::pid (float): state in registers, the caller stores it back
__device__ __forceinline__ float pid_step(const PidGains& g, float& integral, float& prev, bool& first, float e)
{
    integral = __fadd_rn(integral, __fmul_rn(e, g.dt));
    const float deriv = first ? 0.0f : __fdiv_rn(__fsub_rn(e, prev), g.dt);
    prev = e; first = false;
    return __fadd_rn(__fadd_rn(__fmul_rn(g.kp, e), __fmul_rn(g.ki, integral)), __fmul_rn(g.kd, deriv));
}

// :445-473, one thread per vehicle: a sequential scan (the PID is stateful along the polyline).  Reads the velocity
// row of every point and the position only of stop points; with the point-major layout every load is coalesced.
__global__ void k_crossing(long long count, int P, const int32_t* __restrict__ npoints, const float* __restrict__ px,
                           const float* __restrict__ py, const float* __restrict__ vel, long long si, long long sp,
                           bmpc_crossing_config cfg, PidGains g, float* __restrict__ integral, float* __restrict__ prev,
                           uint8_t* __restrict__ first, float* __restrict__ point_velocity)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    int n = npoints ? npoints[b] : P;
    if (n > P) n = P;
    if (n <= 0) return;                                               // :404-407 returned before the clamp
    const long long o = b * si;
    float in = integral[b], pr = prev[b]; bool fi = first[b] != 0, touched = false;
    float minVelocity = cfg.max_velocity_crossing;                    // :446
    float pv = 0.0f; bool set = false;
    for (int k = 0; k < n; ++k) {                                     // :447
        if (vel[o + k * sp] != 0.0f) continue;                        // :448
        const float x = px[o + k * sp], y = py[o + k * sp];
        const float len = __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));
        const float sg = (float)((0.0f < x) - (x < 0.0f));            // lms::math::sgn
        const float dts = __fsub_rn(__fmul_rn(sg, len), cfg.stopping_distance);   // :449
        touched = true;
        if (dts < cfg.distance_to_stop) {                             // :450
            float v = pid_step(g, in, pr, fi, dts);                   // :453
            if (isnan(v) || v >= cfg.max_velocity_crossing) v = cfg.max_velocity_crossing;   // :454-456
            if (dts <= cfg.crossing_safety_zone || v < 0.0f) v = 0.0f;                       // :457-459
            if (v < minVelocity) minVelocity = v;                     // :460-462
            pv = minVelocity; set = true;                             // :464
            if (minVelocity == 0.0f) break;                           // :466-468
        } else { in = 0.0f; pr = 0.0f; fi = true; }                   // :469-471 slowDownCar.reset()
    }
    if (touched) { integral[b] = in; prev[b] = pr; first[b] = fi ? 1 : 0; }
    if (set) point_velocity[b] = pv;
}

// :214-225 + :276-283, one thread per vehicle
__global__ void k_pid_branch(long long count, const float* __restrict__ opx, const float* __restrict__ opy,
                             const float* __restrict__ odx, const float* __restrict__ ody, const float* __restrict__ ovel,
                             PidGains gf, PidGains gr, float* __restrict__ fi_, float* __restrict__ fp_, uint8_t* __restrict__ ff_,
                             float* __restrict__ ri_, float* __restrict__ rp_, uint8_t* __restrict__ rf_,
                             double* __restrict__ sf, double* __restrict__ sr, float* __restrict__ tspeed, float* __restrict__ tdist)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    float in = fi_[b], pr = fp_[b]; bool fi = ff_[b] != 0;
    double f = (double)pid_step(gf, in, pr, fi, opy[b]);                          // :219
    fi_[b] = in; fp_[b] = pr; ff_[b] = 0;
    in = ri_[b]; pr = rp_[b]; fi = rf_[b] != 0;
    double r = (double)pid_step(gr, in, pr, fi, atan2f(ody[b], odx[b]));          // :220 directory.angle()
    ri_[b] = in; rp_[b] = pr; rf_[b] = 0;
    const float x = opx[b], y = opy[b], sp = ovel[b];
    tdist[b] = __fsqrt_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)));           // :224 position.length()
    if (sp < 0.5f) { f = 0.0; r = 0.0; }                                          // :276-283
    sf[b] = f; sr[b] = r; tspeed[b] = sp;                                         // :221-223
}

__global__ void k_pid_fill(long long count, float* __restrict__ integral, float* __restrict__ prev, uint8_t* __restrict__ first)
{
    const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= count) return;
    integral[b] = 0.0f; prev[b] = 0.0f; first[b] = 1;
}

struct TickParams {
    long long count, stride;
    double T, l, w_y, w_phi, w_sf, w_sr, alpha_max;
    const float* velocity;      // [count] car->velocity()
    const double* v_model;      // [count] result of m_mpcLookupVelocity.linearSearch(v) (:323) from the caller, or null
    const float* vtable; int vn;   // ... then the BMPC_LOOKUP_VELOCITY table on the device (vn = 0: v itself)
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
    else if (q.vn > 0) v = (double)lut_eval(q.vtable, q.vn, (float)v);  // LookupTable<float>: double -> float -> double
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

namespace {

inline unsigned grid_for(long long count, int block) { return (unsigned)((count + block - 1) / block); }

int pid_state_ready(bmpc_t* h, int which)
{
    if (h->pid_ready[which]) return BMPC_OK;
    const size_t cap = (size_t)h->cap;
    BMPC_CUDA(h->pid_state[which].reserve(cap * 9));
    float* in = (float*)h->pid_state[which].p;
    k_pid_fill<<<grid_for(h->cap, 256), 256, 0, h->stream>>>(h->cap, in, in + cap, (uint8_t*)(in + 2 * cap));
    BMPC_CUDA(cudaGetLastError());
    h->launches++;
    h->pid_ready[which] = true;
    return BMPC_OK;
}
inline PidGains gains_of(const bmpc_t* h, int which) { return PidGains{h->pid_gain[which][0], h->pid_gain[which][1], h->pid_gain[which][2], h->pid_gain[which][3]}; }

// The front end both branches share: uploads, look-ahead, getTrajectoryPoint, crossing clamp.  Device results:
// ex[7] idx, ex[8] flag, ex[9..13] the trajectory point (x, y, dir x, dir y, velocity), ex[14] car velocity (float).
// It works on the vehicles [b0, b0 + n) of a tick of `count`: a large tick is cut into chunks whose uploads (on the
// handle's copy stream) overlap the kernels of the chunk before (fleet_chunks).
struct FrontArgs {
    int64_t count; const float* velocity; const float* dist; int P; const int32_t* npoints;
    const float *px, *py, *dx, *dy, *vel; int64_t si, sp;
};

int fleet_check(bmpc_t* h, const char* who, const FrontArgs& a)
{
    if (a.count < 1 || a.count > h->cap || a.P < 0) return fail(BMPC_E_INVALID, std::string(who) + ": count / P out of range");
    if (a.P > 0 && (!a.px || !a.py || !a.dx || !a.dy || !a.vel)) return fail(BMPC_E_INVALID, std::string(who) + ": null trajectory array");
    if (a.si < 0 || a.sp < 0) return fail(BMPC_E_INVALID, std::string(who) + ": negative stride");
    if (!a.dist && !h->foh && h->lut_n[BMPC_LOOKUP_DISTANCE] <= 0)
        return fail(BMPC_E_STATE, std::string(who) + ": dist == NULL needs the BMPC_LOOKUP_DISTANCE table (bmpc_set_lookup) or FOH drive mode");
    return BMPC_OK;
}

// Device buffers of a tick (sized for all `count` vehicles, same layout as the host arrays).
int fleet_reserve(bmpc_t* h, const FrontArgs& a)
{
    const size_t c = (size_t)a.count;
    const size_t nflt = a.P > 0 ? (size_t)((a.count - 1) * a.si + (int64_t)(a.P - 1) * a.sp + 1) : 0;
    for (int k = 0; k < 5; ++k) BMPC_CUDA(h->ex[k].reserve(nflt * 4 + 4));
    const size_t voff = ((c * 4 + 7) / 8) * 8;
    BMPC_CUDA(h->ex[14].reserve(voff + c * 8));                           // velocity (float) | v_model (double)
    BMPC_CUDA(h->ex[5].reserve(c * 4));
    if (a.npoints) BMPC_CUDA(h->ex[6].reserve(c * 4));
    for (int k = 7; k < 14; ++k) BMPC_CUDA(h->ex[k].reserve(c * 4));
    return BMPC_OK;
}

// How a tick is cut: 4 chunks when it is large and the polylines are point-major ([P][count], a chunk is P rows of a 2-D
// copy) or instance-major ([count][P], a chunk is contiguous); one chunk otherwise.  BMPC_TICK_CHUNK_MIN: developer
// override of the size from which ticks are chunked (tests).
int fleet_chunks(const FrontArgs& a)
{
    const char* e = std::getenv("BMPC_TICK_CHUNK_MIN");
    const long long min_count = e ? std::atoll(e) : 32768ll;
    if (a.count < min_count || a.P <= 0) return 1;
    const bool point_major = a.si == 1 && a.sp >= a.count;
    const bool instance_major = a.sp == 1 && a.si >= a.P;
    return (point_major || instance_major) ? 4 : 1;
}

// Uploads of the vehicles [b0, b0 + n) on stream `st`.
int fleet_upload(bmpc_t* h, const FrontArgs& a, int64_t b0, int64_t n, const double* v_model, cudaStream_t st)
{
    const float* src[5] = {a.px, a.py, a.dx, a.dy, a.vel};
    if (a.P > 0) {
        if (b0 == 0 && n == a.count) {                                    // whole tick: one flat copy per array
            const size_t nflt = (size_t)((a.count - 1) * a.si + (int64_t)(a.P - 1) * a.sp + 1);
            for (int k = 0; k < 5; ++k) BMPC_CUDA(cudaMemcpyAsync(h->ex[k].p, src[k], nflt * 4, cudaMemcpyHostToDevice, st));
        } else if (a.si == 1) {                                           // point-major: P rows of n floats, pitch sp
            for (int k = 0; k < 5; ++k)
                BMPC_CUDA(cudaMemcpy2DAsync((float*)h->ex[k].p + b0, (size_t)a.sp * 4, src[k] + b0, (size_t)a.sp * 4, (size_t)n * 4,
                                            (size_t)a.P, cudaMemcpyHostToDevice, st));
        } else {                                                          // instance-major: n polylines of pitch si
            const size_t nflt = (size_t)((n - 1) * a.si + a.P);
            for (int k = 0; k < 5; ++k)
                BMPC_CUDA(cudaMemcpyAsync((float*)h->ex[k].p + b0 * a.si, src[k] + b0 * a.si, nflt * 4, cudaMemcpyHostToDevice, st));
        }
    }
    BMPC_CUDA(cudaMemcpyAsync((float*)h->ex[14].p + b0, a.velocity + b0, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if (a.dist) BMPC_CUDA(cudaMemcpyAsync((float*)h->ex[5].p + b0, a.dist + b0, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if (a.npoints) BMPC_CUDA(cudaMemcpyAsync((int32_t*)h->ex[6].p + b0, a.npoints + b0, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if (v_model) {
        const size_t voff = (((size_t)a.count * 4 + 7) / 8) * 8;
        BMPC_CUDA(cudaMemcpyAsync((double*)((char*)h->ex[14].p + voff) + b0, v_model + b0, (size_t)n * 8, cudaMemcpyHostToDevice, st));
    }
    return BMPC_OK;
}

// Look-ahead, getTrajectoryPoint and the crossing clamp for the vehicles [b0, b0 + n) on stream `st`.
int fleet_front_kernels(bmpc_t* h, const FrontArgs& a, int64_t b0, int64_t n, cudaStream_t st)
{
    if (!a.dist) {                                                        // :66 / :70-73 on the device
        k_lookahead<<<grid_for(n, 256), 256, 0, st>>>(n, (const float*)h->ex[14].p + b0, (const float*)h->lut[BMPC_LOOKUP_DISTANCE].p,
                                                      h->lut_n[BMPC_LOOKUP_DISTANCE], h->foh ? 1 : 0, h->foh_min, h->foh_slope,
                                                      (float*)h->ex[5].p + b0);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
    }
    const int32_t* d_np = a.npoints ? (const int32_t*)h->ex[6].p + b0 : nullptr;
    const int64_t o = b0 * a.si;
    BMPC_CUDA(launch_extract(n, a.P, d_np, (const float*)h->ex[0].p + o, (const float*)h->ex[1].p + o, (const float*)h->ex[2].p + o,
                             (const float*)h->ex[3].p + o, (const float*)h->ex[4].p + o, a.si, a.sp,
                             (const float*)h->ex[5].p + b0, (int32_t*)h->ex[7].p + b0, (uint8_t*)h->ex[8].p + b0, (float*)h->ex[9].p + b0,
                             (float*)h->ex[10].p + b0, (float*)h->ex[11].p + b0, (float*)h->ex[12].p + b0, (float*)h->ex[13].p + b0, st));
    h->launches++;
    if (h->crossing_on && a.P > 0) {                                      // :445-473
        float* in = (float*)h->pid_state[BMPC_PID_SLOWDOWN].p;
        const size_t cap = (size_t)h->cap;
        k_crossing<<<grid_for(n, 128), 128, 0, st>>>(n, a.P, d_np, (const float*)h->ex[0].p + o, (const float*)h->ex[1].p + o,
                                                     (const float*)h->ex[4].p + o, a.si, a.sp, h->crossing,
                                                     gains_of(h, BMPC_PID_SLOWDOWN), in + b0, in + cap + b0, (uint8_t*)(in + 2 * cap) + b0,
                                                     (float*)h->ex[13].p + b0);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
    }
    return BMPC_OK;
}

// One-chunk front end (the PID branch and small ticks).
int fleet_front(bmpc_t* h, const char* who, int64_t count, const float* velocity, const float* dist, int P,
                const int32_t* npoints, const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                int64_t stride_instance, int64_t stride_point)
{
    const FrontArgs a{count, velocity, dist, P, npoints, px, py, dx, dy, vel, stride_instance, stride_point};
    int rc = fleet_check(h, who, a);
    if (rc) return rc;
    BMPC_CUDA(cudaSetDevice(h->device));
    if ((rc = fleet_reserve(h, a))) return rc;
    if (h->crossing_on && P > 0 && (rc = pid_state_ready(h, BMPC_PID_SLOWDOWN))) return rc;
    if ((rc = fleet_upload(h, a, 0, count, nullptr, h->stream))) return rc;
    return fleet_front_kernels(h, a, 0, count, h->stream);
}

}  // namespace

extern "C" {

int bmpc_set_lookup(bmpc_t* h, int which, int n, const float* vx, const float* vy)
{
    if (!h || which < 0 || which > 1 || n < 0 || n > 4096 || (n > 0 && (!vx || !vy)))
        return fail(BMPC_E_INVALID, "bmpc_set_lookup: bad argument (which in {0,1}, 0 <= n <= 4096)");
    for (int i = 1; i < n; ++i)
        if (!(vx[i] > vx[i - 1])) return fail(BMPC_E_INVALID, "bmpc_set_lookup: vx must be strictly ascending (LookupTableOrder::ASC)");
    BMPC_CUDA(cudaSetDevice(h->device));
    if (n > 0) {
        BMPC_CUDA(h->lut[which].reserve((size_t)n * 8));
        BMPC_CUDA(cudaMemcpyAsync(h->lut[which].p, vx, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
        BMPC_CUDA(cudaMemcpyAsync((float*)h->lut[which].p + n, vy, (size_t)n * 4, cudaMemcpyHostToDevice, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->lut_n[which] = n;
    return BMPC_OK;
}

int bmpc_lookup(bmpc_t* h, int which, int64_t count, const float* x, float* y)
{
    if (!h || which < 0 || which > 1 || !x || !y || count < 1) return fail(BMPC_E_INVALID, "bmpc_lookup: bad argument");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(h->stage.reserve((size_t)count * 8));
    float* d = (float*)h->stage.p;
    BMPC_CUDA(cudaMemcpyAsync(d, x, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    k_lookahead<<<grid_for(count, 256), 256, 0, h->stream>>>(count, d, (const float*)h->lut[which].p, h->lut_n[which], 0, 0.f, 0.f, d + count);
    BMPC_CUDA(cudaGetLastError());
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(y, d + count, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_fleet_set_drive_mode(bmpc_t* h, int foh, float regelpunkt_min, float regelpunkt_slope)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_fleet_set_drive_mode: null handle");
    h->foh = foh != 0; h->foh_min = regelpunkt_min; h->foh_slope = regelpunkt_slope;
    return BMPC_OK;
}

int bmpc_fleet_set_crossing(bmpc_t* h, const bmpc_crossing_config* cfg)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_fleet_set_crossing: null handle");
    h->crossing_on = cfg != nullptr;
    if (cfg) h->crossing = *cfg;
    return BMPC_OK;
}

int bmpc_fleet_set_pid(bmpc_t* h, int which, float kp, float ki, float kd, float dt)
{
    if (!h || which < 0 || which > 2) return fail(BMPC_E_INVALID, "bmpc_fleet_set_pid: bad argument");
    h->pid_gain[which][0] = kp; h->pid_gain[which][1] = ki; h->pid_gain[which][2] = kd; h->pid_gain[which][3] = dt;
    return BMPC_OK;
}

int bmpc_fleet_reset_pid(bmpc_t* h, int which)
{
    if (!h || which < 0 || which > 2) return fail(BMPC_E_INVALID, "bmpc_fleet_reset_pid: bad argument");
    BMPC_CUDA(cudaSetDevice(h->device));
    h->pid_ready[which] = false;
    return pid_state_ready(h, which);
}

int bmpc_fleet_get_pid(bmpc_t* h, int which, int64_t count, float* integral, float* prev, uint8_t* first)
{
    if (!h || which < 0 || which > 2 || count < 1 || count > h->cap) return fail(BMPC_E_INVALID, "bmpc_fleet_get_pid: bad argument");
    BMPC_CUDA(cudaSetDevice(h->device));
    int rc = pid_state_ready(h, which);
    if (rc) return rc;
    const float* in = (const float*)h->pid_state[which].p; const size_t cap = (size_t)h->cap, c = (size_t)count;
    if (integral) BMPC_CUDA(cudaMemcpyAsync(integral, in, c * 4, cudaMemcpyDeviceToHost, h->stream));
    if (prev) BMPC_CUDA(cudaMemcpyAsync(prev, in + cap, c * 4, cudaMemcpyDeviceToHost, h->stream));
    if (first) BMPC_CUDA(cudaMemcpyAsync(first, in + 2 * cap, c, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_fleet_put_pid(bmpc_t* h, int which, int64_t count, const float* integral, const float* prev, const uint8_t* first)
{
    if (!h || which < 0 || which > 2 || count < 1 || count > h->cap) return fail(BMPC_E_INVALID, "bmpc_fleet_put_pid: bad argument");
    BMPC_CUDA(cudaSetDevice(h->device));
    int rc = pid_state_ready(h, which);
    if (rc) return rc;
    float* in = (float*)h->pid_state[which].p; const size_t cap = (size_t)h->cap, c = (size_t)count;
    if (integral) BMPC_CUDA(cudaMemcpyAsync(in, integral, c * 4, cudaMemcpyHostToDevice, h->stream));
    if (prev) BMPC_CUDA(cudaMemcpyAsync(in + cap, prev, c * 4, cudaMemcpyHostToDevice, h->stream));
    if (first) BMPC_CUDA(cudaMemcpyAsync(in + 2 * cap, first, c, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_fleet_tick_pid(bmpc_t* h, int64_t count, const float* velocity, const float* dist, int P, const int32_t* npoints,
                        const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                        int64_t stride_instance, int64_t stride_point, double* steering_front, double* steering_rear,
                        float* target_speed, float* target_distance, int32_t* idx, uint8_t* flag)
{
    if (!h || !velocity || !steering_front || !steering_rear || !target_speed || !target_distance)
        return fail(BMPC_E_INVALID, "bmpc_fleet_tick_pid: null argument");
    int rc = fleet_front(h, "bmpc_fleet_tick_pid", count, velocity, dist, P, npoints, px, py, dx, dy, vel, stride_instance, stride_point);
    if (rc) return rc;
    if ((rc = pid_state_ready(h, BMPC_PID_FRONT)) || (rc = pid_state_ready(h, BMPC_PID_REAR))) return rc;
    cudaStream_t st = h->stream;
    const size_t c = (size_t)count, cap = (size_t)h->cap;
    BMPC_CUDA(h->ex[15].reserve(c * (8 + 8 + 4 + 4)));
    double* d_sf = (double*)h->ex[15].p; double* d_sr = d_sf + c;
    float* d_ts = (float*)(d_sr + c); float* d_td = d_ts + c;
    float* fs = (float*)h->pid_state[BMPC_PID_FRONT].p; float* rs = (float*)h->pid_state[BMPC_PID_REAR].p;
    k_pid_branch<<<grid_for(count, 128), 128, 0, st>>>(count, (const float*)h->ex[9].p, (const float*)h->ex[10].p,
                                                       (const float*)h->ex[11].p, (const float*)h->ex[12].p, (const float*)h->ex[13].p,
                                                       gains_of(h, BMPC_PID_FRONT), gains_of(h, BMPC_PID_REAR), fs, fs + cap,
                                                       (uint8_t*)(fs + 2 * cap), rs, rs + cap, (uint8_t*)(rs + 2 * cap), d_sf, d_sr, d_ts, d_td);
    BMPC_CUDA(cudaGetLastError());
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(steering_front, d_sf, c * 8, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(steering_rear, d_sr, c * 8, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(target_speed, d_ts, c * 4, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaMemcpyAsync(target_distance, d_td, c * 4, cudaMemcpyDeviceToHost, st));
    if (idx) BMPC_CUDA(cudaMemcpyAsync(idx, h->ex[7].p, c * 4, cudaMemcpyDeviceToHost, st));
    if (flag) BMPC_CUDA(cudaMemcpyAsync(flag, h->ex[8].p, c, cudaMemcpyDeviceToHost, st));
    BMPC_CUDA(cudaStreamSynchronize(st));
    return BMPC_OK;
}

int bmpc_fleet_tick_tobi(bmpc_t* h, int64_t count, const bmpc_tick_config* cfg, const float* velocity,
                         const double* v_model, const float* dist, int P, const int32_t* npoints,
                         const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                         int64_t stride_instance, int64_t stride_point, double* steering_front,
                         double* steering_rear, float* target_speed, float* target_distance, uint8_t* state,
                         int32_t* idx, uint8_t* flag, double* target, uint32_t* iters)
{
    if (!h || !cfg || !velocity || !steering_front || !steering_rear || !target_speed || !target_distance || !state)
        return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: null argument");
    if (h->S != 2 || h->I != 2) return fail(BMPC_E_INVALID, "bmpc_fleet_tick_tobi: the handle must be a (2,2,N) controller (mpcControllerTobi, src/...:303-366)");
    const FrontArgs fa{count, velocity, dist, P, npoints, px, py, dx, dy, vel, stride_instance, stride_point};
    int rc = fleet_check(h, "bmpc_fleet_tick_tobi", fa);
    if (rc) return rc;
    BMPC_CUDA(cudaSetDevice(h->device));
    if ((rc = fleet_reserve(h, fa))) return rc;
    if (h->crossing_on && P > 0 && (rc = pid_state_ready(h, BMPC_PID_SLOWDOWN))) return rc;
    cudaStream_t st = h->stream;
    const size_t c = (size_t)count, es = (size_t)h->elem, cap = (size_t)h->cap;

    // ---- buffers: per-vehicle model, target debug rows, outputs --------------------------------------------------------
    const size_t voff = ((c * 4 + 7) / 8) * 8;
    BMPC_CUDA(h->ex[15].reserve(cap * 8 * 2 + c * (8 + 8 + 4 + 4 + 1)));   // tgt_dbg [2][cap] | outputs
    double* d_vmodel = v_model ? (double*)((char*)h->ex[14].p + voff) : nullptr;
    DevBuf* mb[7] = {&h->mA, &h->mB, &h->mC, &h->mQ, &h->mR, &h->mlo, &h->mhi};
    const int mlen[7] = {4, 4, 2, 2, 2, 2, 2};
    for (int a = 0; a < 7; ++a) BMPC_CUDA(mb[a]->reserve(cap * mlen[a] * es));
    BMPC_CUDA(h->mlambda.reserve(cap * es));
    BMPC_CUDA(h->mqdiag.reserve(cap * h->NI * es));
    if (target) BMPC_CUDA(h->stage.reserve(c * 2 * 8));
    double* tgt_dbg = (double*)h->ex[15].p;
    char* ob = (char*)h->ex[15].p + cap * 8 * 2;
    double* d_sf = (double*)ob; double* d_sr = d_sf + c;
    float* d_ts = (float*)(d_sr + c); float* d_td = d_ts + c;
    uint8_t* d_state = (uint8_t*)(d_td + c);
    h->per_instance = true; h->model_count = count; h->have_model = true;
    h->full_targets = false; h->head = 0;                         // set_target(val): one target for all horizon steps
    const ShapeOps* tobi = h->specialize ? find_pattern(h->S, h->I, h->N, "tobi") : nullptr;   // holds by construction
    h->solve_ops = tobi ? tobi : h->ops;

    // ---- chunks: the uploads of chunk k+1 (copy stream) overlap the kernels and the downloads of chunk k ------------------
    const int chunks = fleet_chunks(fa);
    if (chunks > 1 && !h->copy_stream) {
        BMPC_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 8; ++k) BMPC_CUDA(cudaEventCreateWithFlags(&h->ev_chunk[k], cudaEventDisableTiming));
    }
    auto range = [&](int k, int64_t* b0, int64_t* n) { *b0 = count * k / chunks; *n = count * (k + 1) / chunks - *b0; };
    if (chunks > 1) {
        // the copy stream starts after whatever the handle's stream still has queued on these buffers
        BMPC_CUDA(cudaEventRecord(h->ev_chunk[7], st));
        BMPC_CUDA(cudaStreamWaitEvent(h->copy_stream, h->ev_chunk[7], 0));
        for (int k = 0; k < chunks; ++k) {
            int64_t b0, n; range(k, &b0, &n);
            if ((rc = fleet_upload(h, fa, b0, n, v_model, h->copy_stream))) return rc;
            BMPC_CUDA(cudaEventRecord(h->ev_chunk[k], h->copy_stream));
        }
    } else if ((rc = fleet_upload(h, fa, 0, count, v_model, st))) return rc;

    for (int k = 0; k < chunks; ++k) {
        int64_t b0, n; range(k, &b0, &n);
        if (n <= 0) continue;
        if (chunks > 1) BMPC_CUDA(cudaStreamWaitEvent(st, h->ev_chunk[k], 0));
        if ((rc = fleet_front_kernels(h, fa, b0, n, st))) return rc;
        // ---- per-vehicle model, target and zero state
        TickParams q{};
        q.count = n; q.stride = h->cap; q.T = cfg->T; q.l = cfg->l; q.w_y = cfg->weight_y; q.w_phi = cfg->weight_phi;
        q.w_sf = cfg->weight_steering_front; q.w_sr = cfg->weight_steering_rear; q.alpha_max = cfg->alpha_max;
        q.velocity = (const float*)h->ex[14].p + b0; q.v_model = d_vmodel ? d_vmodel + b0 : nullptr;
        q.vtable = (const float*)h->lut[BMPC_LOOKUP_VELOCITY].p; q.vn = h->lut_n[BMPC_LOOKUP_VELOCITY];
        q.opx = (const float*)h->ex[9].p + b0; q.opy = (const float*)h->ex[10].p + b0; q.odx = (const float*)h->ex[11].p + b0;
        q.ody = (const float*)h->ex[12].p + b0;
        const unsigned grid = (unsigned)((n + 127) / 128);
        if (h->elem == 8)
            k_tick_prepare<double><<<grid, 128, 0, st>>>(q, (double*)h->mA.p + b0, (double*)h->mB.p + b0, (double*)h->mC.p + b0,
                                                         (double*)h->mQ.p + b0, (double*)h->mR.p + b0, (double*)h->mlo.p + b0,
                                                         (double*)h->mhi.p + b0, (double*)h->x0.p + b0, (double*)h->target_const.p + b0,
                                                         tgt_dbg + b0);
        else
            k_tick_prepare<float><<<grid, 128, 0, st>>>(q, (float*)h->mA.p + b0, (float*)h->mB.p + b0, (float*)h->mC.p + b0,
                                                        (float*)h->mQ.p + b0, (float*)h->mR.p + b0, (float*)h->mlo.p + b0,
                                                        (float*)h->mhi.p + b0, (float*)h->x0.p + b0, (float*)h->target_const.p + b0,
                                                        tgt_dbg + b0);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
        // ---- fresh controllers: constructor recursion (mpc.h:114-123), then one cold operator() each
        const size_t off = (size_t)b0 * es;
        GenericCtor g{n, h->cap, (char*)h->mA.p + off, (char*)h->mB.p + off, (char*)h->mQ.p + off, (char*)h->mR.p + off,
                      (char*)h->mlambda.p + off, (char*)h->mqdiag.p + off};
        BMPC_CUDA(h->ops->ctor(h->precision, h->mode, g, st));
        h->launches++;
        if ((rc = solve_range_async(h, b0, n, 0))) return rc;
        // ---- CarCommand fields
        if (h->elem == 8)
            k_tick_finish<double><<<grid, 128, 0, st>>>(n, h->cap, (const double*)h->u0.p + b0, (const float*)h->ex[9].p + b0,
                                                        (const float*)h->ex[10].p + b0, (const float*)h->ex[13].p + b0, d_sf + b0,
                                                        d_sr + b0, d_ts + b0, d_td + b0, d_state + b0);
        else
            k_tick_finish<float><<<grid, 128, 0, st>>>(n, h->cap, (const float*)h->u0.p + b0, (const float*)h->ex[9].p + b0,
                                                       (const float*)h->ex[10].p + b0, (const float*)h->ex[13].p + b0, d_sf + b0,
                                                       d_sr + b0, d_ts + b0, d_td + b0, d_state + b0);
        BMPC_CUDA(cudaGetLastError());
        h->launches++;
        const size_t m = (size_t)n;
        BMPC_CUDA(cudaMemcpyAsync(steering_front + b0, d_sf + b0, m * 8, cudaMemcpyDeviceToHost, st));
        BMPC_CUDA(cudaMemcpyAsync(steering_rear + b0, d_sr + b0, m * 8, cudaMemcpyDeviceToHost, st));
        BMPC_CUDA(cudaMemcpyAsync(target_speed + b0, d_ts + b0, m * 4, cudaMemcpyDeviceToHost, st));
        BMPC_CUDA(cudaMemcpyAsync(target_distance + b0, d_td + b0, m * 4, cudaMemcpyDeviceToHost, st));
        BMPC_CUDA(cudaMemcpyAsync(state + b0, d_state + b0, m, cudaMemcpyDeviceToHost, st));
        if (idx) BMPC_CUDA(cudaMemcpyAsync(idx + b0, (int32_t*)h->ex[7].p + b0, m * 4, cudaMemcpyDeviceToHost, st));
        if (flag) BMPC_CUDA(cudaMemcpyAsync(flag + b0, (uint8_t*)h->ex[8].p + b0, m, cudaMemcpyDeviceToHost, st));
        if (iters) BMPC_CUDA(cudaMemcpyAsync(iters + b0, (uint32_t*)h->iters.p + b0, m * 4, cudaMemcpyDeviceToHost, st));
        if (target) {                                              // [count][2] = y_soll, phi_soll
            BMPC_CUDA(soa_to_aos(tgt_dbg + b0, (double*)h->stage.p + 2 * b0, n, 2, h->cap, 8, st));
            h->launches++;
            BMPC_CUDA(cudaMemcpyAsync(target + 2 * b0, (double*)h->stage.p + 2 * b0, m * 2 * 8, cudaMemcpyDeviceToHost, st));
        }
    }
    BMPC_CUDA(cudaStreamSynchronize(st));
    return BMPC_OK;
}

}  // extern "C"
