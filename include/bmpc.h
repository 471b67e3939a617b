/* bmpc.h — thin C ABI of the B200-native batched linear-MPC engine (libbmpc.so).
 *
 * This is the drop-in boundary for ONE hot path of lms-org/trajectory_controller:
 *   getTrajectoryPoint()  (src/trajectory_point_follower.cpp:392-443)   -> bmpc_extract_points
 *   mpcControllerTobi()   (src/trajectory_point_follower.cpp:301-389)   -> bmpc_set_model / bmpc_set_target / bmpc_solve
 *   dlib::mpc<S,I,N>      (dlib_files/dlib/control/mpc.h:13-365)        -> one engine slot per controller object
 * The reference crosses no FFI on this path (everything is in-process C++); these entry points are what a C++
 * facade with dlib::mpc's method set (trajectory_controller_b200/cpp/bmpc_mpc.h) and the module glue bind.
 *
 * Conventions
 *   - every function returns BMPC_OK (0) or a negative bmpc_status; bmpc_last_error() gives the message of the
 *     last failure on the calling thread.  Nothing throws across this boundary.
 *   - the caller owns all host pointers; a handle owns its device memory, stream and events.
 *   - a handle is NOT thread-safe; distinct handles are independent.  One handle = one CUDA device.
 *   - host arrays are instance-major ("AoS"): element e of instance b at [b*len + e]; matrices are row-major
 *     like dlib::matrix (matrix_data_layout.h:131-155).  All host numbers are double (dlib's type) whatever the
 *     compute precision of the handle.
 *   - calls are synchronous on return unless the name ends in _async; bmpc_sync() waits for those.
 *   - there is no CPU fallback: without a CUDA device bmpc_create fails with BMPC_E_CUDA.
 */
#ifndef BMPC_H
#define BMPC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bmpc_handle bmpc_t;

typedef enum bmpc_status {
    BMPC_OK = 0,
    BMPC_E_INVALID = -1,      /* bad argument (null pointer, count > capacity, time >= horizon, ...) */
    BMPC_E_UNSUPPORTED = -2,  /* (S,I,N) / precision / mode combination not instantiated in this build */
    BMPC_E_CUDA = -3,         /* CUDA runtime error, or no usable device */
    BMPC_E_NOMEM = -4,        /* host or device allocation failed */
    BMPC_E_STATE = -5         /* call order violated (e.g. solve before set_model) */
} bmpc_status;

#define BMPC_FP64 64
#define BMPC_FP32 32

/* EXACT: the reference's floating-point evaluation order, no FMA contraction: results equal dlib::mpc's
 *        (IEEE value equality; the sign of a zero is the only freedom).  Iteration counts identical.
 * FAST : same recursion with FMA contraction; tolerance parity only. */
#define BMPC_MODE_EXACT 0
#define BMPC_MODE_FAST 1

/* bmpc_solve_* flags */
#define BMPC_WARM 1u            /* persistent-controller semantics of mpc.h:229-237: shift the stored controls before the
                                   solve, shift the targets after it, keep controls/v for the next call.  Without it
                                   every instance behaves like a freshly constructed dlib::mpc
                                   (src/trajectory_point_follower.cpp:366: controls = 0), and nothing is persisted. */
#define BMPC_KEEP_CONTROLS 2u   /* cold solves only: also write the full control sequence so bmpc_get_controls works */

/* per-instance status byte written by a solve */
#define BMPC_ST_MAXITER 1u      /* the loop ended at max_iterations (mpc.h:271) instead of the eps test (mpc.h:310) */
#define BMPC_ST_NONFINITE 2u    /* a NaN / inf reached the solve through x0, the targets or the model: the result follows
                                   dlib's arithmetic for such inputs but is not a control to act on (the reference logs
                                   isnan() of its outputs, src/trajectory_point_follower.cpp:100-103) */

/* ---- lifetime ------------------------------------------------------------------------------------------- */
/* 1 if the (S,I,N) template shape is instantiated (dlib::mpc is a compile-time template too, mpc.h:13-18). */
int bmpc_shape_supported(int S, int I, int N);
/* mpc.h:51-125 for `capacity` controllers at once.  device < 0 = current device. */
int bmpc_create(bmpc_t** out, int S, int I, int N, int64_t capacity, int precision, int mode, int device);
void bmpc_destroy(bmpc_t* h);
const char* bmpc_last_error(void);
/* Build identity: "bmpc <version> sm_100a ..." */
const char* bmpc_version(void);

/* ---- model / options  (mpc.h:51-125, :172-214) ------------------------------------------------------------- */
/* A[S*S] B[S*I] C[S] Q[S] R[I] lower[I] upper[I]; per_instance = 0: one model for every instance;
 * 1: `count` models, instance-major.  Runs the lambda / Q_diag recursion (mpc.h:114-123) on the device.
 * dlib's requires clause (mpc.h:92-94: min(Q) >= 0, min(R) > 0, upper >= lower) and finiteness are checked for every
 * instance: BMPC_E_INVALID instead of dlib's undefined behaviour. */
int bmpc_set_model(bmpc_t* h, const double* A, const double* B, const double* C, const double* Q, const double* R,
                   const double* lower, const double* upper, int per_instance, int64_t count);
/* Structural specialisation.  bmpc_set_model picks, among the kernels compiled for the shape, the one whose structural
 * pattern (entries that are exactly 0 / exactly 1 in A, B, C, Q of EVERY instance) is the sparsest the model satisfies;
 * dropping those terms is exact in IEEE arithmetic (only the sign of a zero can differ), so parity is unaffected.
 * enable = 0 forces the dense kernel (takes effect at the next bmpc_set_model).  Default: enabled. */
int bmpc_set_specialization(bmpc_t* h, int enable);
/* Staged solve (engine extension; default enabled).  A solve runs as up to three kernels:
 *   stage A  set-up (mpc.h:258-266) plus the 50 coordinate steps (mpc.h:319-335), one instance per lane;
 *   stage B  the projected-gradient phase (mpc.h:336-345), one instance per lane, in a kernel specialised for it.  For
 *            batches of one to four instances per resident lane it time-slices: every 128 iterations a warp parks its
 *            instances at the tail of a work ring and resumes the oldest parked ones, so that heavy-tailed iteration
 *            counts do not leave most lanes idle behind a few late long solves;
 *   stage C  the same phase with one instance per LANE PAIR (block-triangular (4,1) models with symmetric bounds, even
 *            horizons: the synthetic bicycle model of every (4,1,N) configuration): ~40 % less latency per iteration at
 *            ~30 % more issue slots per instance.  It takes the last sm_count*64 unfinished instances of every launch
 *            from stage B, and replaces stage B for batches up to that size, for per-instance models and for horizons
 *            whose controls do not fit a lane's registers (N >= 60 in FP64).
 * Results are bit-identical to the one-kernel solve in every setting.  enable: 0 = one kernel; 1 = staged with fairness
 * slices, retirement of surplus warps near the end of a launch, the symmetric-bound stage-B kernel for models with
 * lower == -upper, and stage C (default); 2 = stages A + B only, every instance runs to completion in the lane that
 * claimed it; 3 = like 2 with fairness slices and retirement, but the literal stage-B kernel whatever the bounds
 * (2 and 3 exist for A/B tests). */
int bmpc_set_split(bmpc_t* h, int enable);
/* Name of the pattern the current model runs with: "dense", "bicycle4", "tobi", "kat". */
const char* bmpc_pattern_name(bmpc_t* h);
/* mpc.h:190-214; defaults eps = 0.01, max_iterations = 10000 (mpc.h:103-104).  max_iterations saturates at 2^32-1. */
int bmpc_set_options(bmpc_t* h, double eps, uint64_t max_iterations);
int bmpc_get_options(bmpc_t* h, double* eps, uint64_t* max_iterations);
/* Test observability (dlib keeps these private, mpc.h:360-361): lambda and Q_diag[N*I] of one instance's model. */
int bmpc_get_derived(bmpc_t* h, int64_t instance, double* lambda, double* q_diag);

/* ---- targets  (mpc.h:142-170) ------------------------------------------------------------------------------ */
/* time = -1: set_target(val) (all horizon steps); 0..N-1: set_target(val, time); time = N-1 is set_last_target.
 * per_instance = 0: val[S] for every instance; 1: val[count][S]. */
int bmpc_set_target(bmpc_t* h, const double* val, int time, int per_instance, int64_t count);
int bmpc_get_target(bmpc_t* h, int64_t instance, int time, double* val /* [S] */);

/* ---- controller state ------------------------------------------------------------------------------------------ */
/* Back to freshly constructed controllers: controls = 0, targets = 0 (mpc.h:105-112), v = 0. */
int bmpc_reset(bmpc_t* h);
/* controls[count][N][I] of instances first..first+count-1 (what dlib keeps in `controls`, mpc.h:362). */
int bmpc_get_controls(bmpc_t* h, int64_t first, int64_t count, double* controls);
/* EXTENSION (dlib::mpc has no such setter): seed the warm start. */
int bmpc_set_controls(bmpc_t* h, int64_t first, int64_t count, const double* controls);
/* Copy-construction / assignment of controller objects (dlib::mpc is a by-value type, mpc.h:349-362): controls,
 * targets, the stale-v storage and the options of `src` into `dst` (device-to-device).  Both handles must have the
 * same shape, precision and capacity and live on the same device.  The model is not copied (call bmpc_set_model). */
int bmpc_copy_state(bmpc_t* dst, bmpc_t* src);

/* ---- solve  (mpc.h:216-240 operator(), :253-347 solve_linear_mpc) --------------------------------------------- */
/* Host buffers: x0[count][S] in; u0[count][I], iters[count] (may be NULL), status[count] (may be NULL) out.
 * Includes the host<->device copies. */
int bmpc_solve(bmpc_t* h, int64_t count, const double* x0, double* u0, uint32_t* iters, uint8_t* status,
               unsigned flags);
/* Device-resident pieces of the same call, for callers that keep the batch in HBM: */
int bmpc_upload_x0(bmpc_t* h, int64_t count, const double* x0);            /* H2D + layout change */
int bmpc_solve_resident_async(bmpc_t* h, int64_t count, unsigned flags);   /* kernel only, on the handle's stream */
int bmpc_download(bmpc_t* h, int64_t count, double* u0, uint32_t* iters, uint8_t* status);
int bmpc_sync(bmpc_t* h);
/* Device pointers of the last solve's results, for callers that keep them in HBM (a multi-GPU host gathers them with
 * one collective): u0 as [I][capacity] of the handle's compute type (double or float), iters as uint32[capacity],
 * status as uint8[capacity].  Valid until bmpc_destroy; written by every solve on the handle's stream.  Any pointer
 * argument may be NULL. */
int bmpc_device_results(bmpc_t* h, void** u0, void** iters, void** status);
/* The three arrays above live in ONE device allocation, u0 first: `base` = its address (== u0), `bytes` its size,
 * iters at base + iters_offset, status at base + status_offset (offsets are multiples of 256).  A multi-GPU host
 * gathers a whole job's results with one collective and one device-to-host copy of this block per rank. */
int bmpc_device_results_block(bmpc_t* h, void** base, uint64_t* bytes, uint64_t* iters_offset, uint64_t* status_offset);
/* Run every later call of this handle on the caller's CUDA stream (a cudaStream_t passed as void*; NULL = back to
 * the handle's own stream).  Lets a host application order the engine with its own copies/events, and lets a
 * benchmark bracket launches with its own events.  The handle never destroys a borrowed stream. */
int bmpc_set_stream(bmpc_t* h, void* cuda_stream);
/* CUDA-event time of the last solve kernel (ms), the number of kernels this handle has launched so far, and
 * the launch geometry of the solve kernel (for roofline bookkeeping). */
int bmpc_last_kernel_ms(bmpc_t* h, float* ms);
/* Two-stage solve: device time of stage A (set-up + coordinate phase) and stage B (projected-gradient kernel) of
 * the last solve; stage_b_ms = 0 when the solve ran as one kernel or was a small call (bmpc_solve of up to 16
 * instances records no stage event).  Launch geometry of the stage-B kernel. */
int bmpc_last_stage_ms(bmpc_t* h, float* stage_a_ms, float* stage_b_ms);
int bmpc_pg_kernel_info(bmpc_t* h, int* grid, int* block, int* smem_bytes, int* regs);
int64_t bmpc_launch_count(bmpc_t* h);
int bmpc_kernel_info(bmpc_t* h, int* grid, int* block, int* smem_bytes, int* regs);
/* Measured arithmetic peak of the handle's device, the roofline denominator of the solver (it is bound by the FP64 /
 * FP32 CUDA-core pipe, not by HBM or tensor cores): register-resident chains; fused = 1: FMA (2 flop/instruction),
 * fused = 0: separate multiplies and adds as EXACT mode must issue them (1 flop/instruction). */
int bmpc_probe_peak(bmpc_t* h, int precision, int fused, double* tflops, double* ginstr_per_s);

/* ---- trajectory point  (src/trajectory_point_follower.cpp:392-443) ---------------------------------------------- */
#define BMPC_PT_INTERPOLATED 0  /* :428-437 */
#define BMPC_PT_LAST 1          /* :439-442 */
#define BMPC_PT_EMPTY 2         /* :394-407 */
/* Polylines as five float arrays; element k of instance b at base[b*stride_instance + k*stride_point]
 * (stride_instance = 0: one trajectory shared by all instances).  npoints: NULL (all P) or [count].
 * dist[count] in; idx/flag/pos/dir/vel [count] out.  Index selection and outputs are bit-exact w.r.t. the
 * sequential-float restatement in oracle/extract_oracle.c. */
int bmpc_extract_points(bmpc_t* h, int64_t count, int P, const int32_t* npoints, const float* px, const float* py,
                        const float* dx, const float* dy, const float* vel, int64_t stride_instance,
                        int64_t stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* out_px,
                        float* out_py, float* out_dx, float* out_dy, float* out_vel);

/* getTrajectoryPoint for a whole horizon ("sample the trajectory at N lookahead distances and
 * set_target(val, t) per step"; dlib's per-step targets, mpc.h:142-170, contract mpc_abstract.h:180-206).  Step t of
 * instance b asks for the distance d_t = d0[b] + (float)t * step[b] (float multiply and add, each rounded on its own);
 * the result for step t is exactly what bmpc_extract_points returns for that distance — the device produces all
 * `steps` points in one sequential scan of the polyline.  step[b] must be >= 0.  Polyline arguments as for
 * bmpc_extract_points; outputs are instance-major [count][steps]. */
int bmpc_extract_horizon(bmpc_t* h, int64_t count, int P, const int32_t* npoints, const float* px, const float* py,
                         const float* dx, const float* dy, const float* vel, int64_t stride_instance,
                         int64_t stride_point, const float* d0, const float* step, int steps, int32_t* idx,
                         uint8_t* flag, float* out_px, float* out_py, float* out_dx, float* out_dy, float* out_vel);

/* bmpc_extract_points without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code beside a CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  No handle: `device` is the CUDA device index, `cuda_stream` a cudaStream_t (NULL: default stream).
 * Arguments, results and bit-exactness as for bmpc_extract_points (host arrays in, host arrays out); the device staging
 * lives in ONE file-scope device array of BMPC_STATIC_EXTRACT_BYTES inside the library (every GPU that loads libbmpc.so
 * holds it until the process exits), and the call makes no cudaMalloc / cudaFree / allocating library call.
 * A problem whose staging — five polyline arrays of (count-1)*stride_instance + (P-1)*stride_point + 1 floats, dist,
 * npoints and seven outputs of `count` elements, each rounded up to 256 bytes — exceeds the array is rejected with
 * BMPC_E_INVALID (bmpc_last_error() names the sizes): use bmpc_extract_points for it.
 * All calls on a device share the array, which keeps the last call's contents (its output window is cleared at the
 * start of every call).  Synchronous on return, and concurrent callers are serialised inside the library, so calls made
 * through this entry point never overlap on the GPU; work that did overlap on the array would race. */
#define BMPC_STATIC_EXTRACT_BYTES (4u << 20)
int bmpc_extract_points_static(int device, void* cuda_stream, int64_t count, int P, const int32_t* npoints,
                               const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                               int64_t stride_instance, int64_t stride_point, const float* dist, int32_t* idx,
                               uint8_t* flag, float* out_px, float* out_py, float* out_dx, float* out_dy, float* out_vel);

/* ---- closed loop on the device  (BASELINE.json configs[4]) ------------------------------------
 * Every tick, for every vehicle, without a host round trip:
 *   sample the world path into a vehicle-frame polyline of P float points (what the TRAJECTORY channel carries)
 *   -> getTrajectoryPoint(lookahead)                       src/trajectory_point_follower.cpp:392-443
 *   -> target = (point.y, atan(dir.y/dir.x), 0, 0) for every horizon step      src/...:84-85, :368-371
 *   -> persistent-controller solve from x0 = (0, 0, delta, delta_dot)          mpc.h:216-240 (BMPC_WARM)
 *   -> plant step with u0.
 * The reference has no plant and no path sampler: those two are HARNESS code, defined here so that the CPU restatement
 * (oracle/closed_loop_oracle.c) and the kernels compute the same numbers bit for bit — only + - * / sqrt, each
 * rounded on its own, in this order:
 *   k0      = first index with path_x[k] > X, clamped to [0, npath - P]
 *   point i = world point k0 + i: (dx, dy) = (path_x - X, path_y - Y) in double;
 *             position = ((float)(c*dx + s*dy), (float)(c*dy - s*dx)); the direction of the selected point is the
 *             world tangent rotated the same way; velocity = (float)v
 *   phi     = bmpc_harness_atan(dir.y / dir.x) in double if dir.x > 0, else +-pi/2 by the sign of dir.y
 *   plant   : delta' = (A22*delta + A23*ddelta) + B2*u;  ddelta' = (A32*delta + A33*ddelta) + B3*u;
 *             dpsi = ((T*v)/l)*delta;  (c', s') = (c - s*dpsi, s + c*dpsi) / sqrt(c'^2 + s'^2);
 *             X' = X + (T*v)*c;  Y' = Y + (T*v)*s          (old c, s, delta on the right-hand sides)
 *   bmpc_harness_atan(t): t1 = t/(1+sqrt(1+t*t)); t2 = t1/(1+sqrt(1+t1*t1)); z = t2*t2;
 *             4 * t2 * (1 - z/3 + z^2/5 - ... - z^11/23) by Horner from the last term.
 * Requires a shared (not per-instance) model with S = 4, I = 1 (the survey's bicycle model). */
typedef struct bmpc_cl_config {
    double v;            /* speed [m/s] */
    double T;            /* plant step [s] (src/trajectory_point_follower.cpp:96) */
    double l;            /* wheelbase [m] (include/trajectory_point_follower.h:47) */
    double lookahead;    /* d = regelpunktMin + v*regelpunktSlope (src/...:71), evaluated by the caller; used as float */
    int32_t P;           /* polyline points per tick */
    int32_t npath;       /* world path points */
} bmpc_cl_config;
/* path_*: npath doubles each (world positions, ascending x, and unit tangents); pose0[count][4] = X, Y, cos psi, sin psi;
 * act0[count][2] = delta, delta_dot (NULL = zeros).  Resets the controllers (fresh dlib::mpc objects). */
int bmpc_cl_init(bmpc_t* h, int64_t count, const bmpc_cl_config* cfg, const double* path_x, const double* path_y,
                 const double* path_tx, const double* path_ty, const double* pose0, const double* act0);
/* A target per horizon step instead of one for all steps (§8(f) rank 2): every tick the vehicle-frame polyline is
 * sampled at d_t = d0 + (float)t * step, t = 0 .. N-1 (see bmpc_extract_horizon) and target[t] = (point_t.y,
 * atan(dir_t.y/dir_t.x), 0, 0) is written for every step (dlib: set_target(val, t), mpc.h:142-155) before the solve.
 * d0, step in metres, used as float, step >= 0; the natural choice is d0 = step = T*v (state i+1 of the prediction lies
 * (i+1)*T*v ahead).  enable = 0: back to the reference's single target at `lookahead` (src/...:368-371).  Call after
 * bmpc_cl_init; bmpc_cl_get then reports target / idx of horizon step 0. */
int bmpc_cl_set_horizon_targets(bmpc_t* h, int enable, double d0, double step);
/* `ticks` control ticks for all vehicles; kernels only, synchronous on return. */
int bmpc_cl_run(bmpc_t* h, int ticks);
/* State after the last tick; any pointer may be NULL.  pose[count][4], act[count][2], u0[count] (last command),
 * target[count][2] (last y_soll, phi_soll), idx[count] (last selected polyline index), iters_total[count] (solver
 * iterations summed over all ticks), iters_max[count]. */
int bmpc_cl_get(bmpc_t* h, int64_t count, double* pose, double* act, double* u0, double* target, int32_t* idx,
                uint64_t* iters_total, uint32_t* iters_max);

/* ---- fleet tick: TrajectoryPointController::cycle(), tobiMPC branch, for `count` vehicles in one call ------------------
 * Replaces, per vehicle, src/trajectory_point_follower.cpp:64-126 + :276-283 + mpcControllerTobi :301-389 (SURVEY.md
 * §8(f) rank 1).  The handle must be a (2,2,N) controller (the reference uses N = MPC_HORIZON = 4).  Inputs, host arrays:
 *   velocity[count]   car->velocity() (float);  |v| < 0.1 is replaced by 0.1 (:78-82)
 *   v_model[count]    m_mpcLookupVelocity.linearSearch(v) (:323) evaluated by the caller (the lookup table lives in the
 *                     un-vendored lms_math); NULL = use v itself
 *   dist[count]       look-ahead distance (:66 lookup or :71 regelpunktMin + v*regelpunktSlope) evaluated by the caller,
 *                     or NULL (see below)
 *   P, npoints, px..vel, strides   the TRAJECTORY polylines, exactly as for bmpc_extract_points
 * On the device: getTrajectoryPoint (:392-443), phi_soll = atan2(dir.y, dir.x) and y_soll = position.y (:84-85), the
 * per-vehicle model (:326-340) with bounds -+alpha_max (:16-18), a fresh dlib::mpc (constructor math mpc.h:114-123),
 * set_target (:368-371), operator()((0,0)) (:374-377), the CarCommand fields (:114-125) and the crossing rule
 * "targetSpeed < 0.5 -> steering 0" (:276-283).  Outputs, host arrays of `count`: steering_front / steering_rear
 * (radians), target_speed (trajectoryPoint.velocity), target_distance (position.length(), float), state
 * (BMPC_CAR_IDLE when the point's velocity is 0, else BMPC_CAR_DRIVING); optional (may be NULL): idx / flag of the
 * selected polyline point, target[count][2] = (y_soll, phi_soll), iters.
 * Parity: everything is bit-exact against the CPU restatement except atan2, which is CUDA's double atan2 (<= 2 ulp)
 * where the reference calls libm's; the tests bound the effect (|d phi_soll| <= 4e-16, |d steering| <= 1e-9).
 * The call replaces the handle's model, targets and controller state (fresh controllers every tick, like the
 * reference, which constructs its dlib::mpc inside mpcControllerTobi).
 *
 * Look-ups and the crossing clamp on the device:
 *   dist == NULL      the look-ahead distance is evaluated per vehicle on the device: the BMPC_LOOKUP_DISTANCE table at
 *                     car->velocity() (:66), or regelpunktMin + v*regelpunktSlope in FOH drive mode (:70-73,
 *                     bmpc_fleet_set_drive_mode)
 *   v_model == NULL   and a BMPC_LOOKUP_VELOCITY table is set: m_mpcLookupVelocity.linearSearch(v) (:323) on the device
 *   the "stop at crossing" clamp at the end of getTrajectoryPoint (:445-473) runs on the device after the extraction,
 *   with one slowDownCar PID per vehicle resident in HBM across ticks; it is ON by default with the reference's config
 *   defaults (bmpc_fleet_set_crossing).  target_speed is the clamped trajectoryPoint.velocity.
 * sensor_utils::PID and lms::math::LookupTable are not vendored by the reference (parity unpinned); the semantics adopted,
 * all in float with every operation rounded on its own, are those of the headless shim the module is built against:
 *   PID::pid(e):   integral += e*dt;  deriv = first ? 0 : (e - prev)/dt;  prev = e;  Kp*e + Ki*integral + Kd*deriv
 *   linearSearch:  n == 0 -> x;  x <= vx[0] -> vy[0];  first i with x <= vx[i] ->
 *                  vy[i-1] + (vy[i]-vy[i-1])*(x-vx[i-1])/(vx[i]-vx[i-1]);  else vy[n-1] */
#define BMPC_CAR_IDLE 0
#define BMPC_CAR_DRIVING 1
typedef struct bmpc_tick_config {
    double T;                      /* mpcParameters.stepSize, :96 (0.1) */
    double l;                      /* wheelbase, include/trajectory_point_follower.h:47 (0.21) */
    double weight_y, weight_phi;   /* :92-93 (20, 7) */
    double weight_steering_front, weight_steering_rear;   /* :94-95 (0.0005, 10) */
    double alpha_max;              /* :16 (22*M_PI/180) */
} bmpc_tick_config;
int bmpc_fleet_tick_tobi(bmpc_t* h, int64_t count, const bmpc_tick_config* cfg, const float* velocity,
                         const double* v_model, const float* dist, int P, const int32_t* npoints,
                         const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                         int64_t stride_instance, int64_t stride_point, double* steering_front,
                         double* steering_rear, float* target_speed, float* target_distance, uint8_t* state,
                         int32_t* idx, uint8_t* flag, double* target, uint32_t* iters);

/* lms::math::LookupTable<float, ASC> (include/trajectory_point_follower.h:33-34; filled by configsChanged(),
 * src/...:292-295): n samples, vx ascending.  linearSearch = piecewise-linear, clamped at both ends; n = 0 removes the
 * table (identity, like an unconfigured table). */
#define BMPC_LOOKUP_DISTANCE 0   /* m_trajectoryPointDistanceLookup, src/...:66, :216 */
#define BMPC_LOOKUP_VELOCITY 1   /* m_mpcLookupVelocity, src/...:323 */
int bmpc_set_lookup(bmpc_t* h, int which, int n, const float* vx, const float* vy);
/* Evaluates a table for `count` host floats on the device (what the fleet tick does internally; exposed for tests and
 * for callers that want the look-ahead distances back). */
int bmpc_lookup(bmpc_t* h, int which, int64_t count, const float* x, float* y);
/* phxService->driveMode() == FOH (src/...:70-73): look-ahead = regelpunkt_min + v * regelpunkt_slope (defaults 0.6, 0.1)
 * instead of the distance table. */
int bmpc_fleet_set_drive_mode(bmpc_t* h, int foh, float regelpunkt_min, float regelpunkt_slope);

/* The crossing clamp, src/...:445-473.  cfg == NULL switches it off (trajectoryPoint.velocity is the polyline's). */
typedef struct bmpc_crossing_config {
    float max_velocity_crossing;   /* "maxVelocityCrossing" 1.0   (:446, :452) */
    float stopping_distance;       /* "stoppingDistance"    0.35  (:449) */
    float distance_to_stop;        /* "distanceToStop"      1     (:450) */
    float crossing_safety_zone;    /* "crossingSaftyZone"   0.05  (:457) */
} bmpc_crossing_config;
int bmpc_fleet_set_crossing(bmpc_t* h, const bmpc_crossing_config* cfg);
/* The three sensor_utils::PID members (include/trajectory_point_follower.h:40-42), one state per vehicle each.
 * set = PID::set(Kp, Ki, Kd, dt) (src/...:296-298; defaults slowDownCar 1,0,0,0.01; front / rear 1,1,0,0.01);
 * reset = PID::reset() for every vehicle.  get/put move the states [count] (integral, prev, first) for tests and
 * checkpoints; any pointer may be NULL. */
#define BMPC_PID_SLOWDOWN 0
#define BMPC_PID_FRONT 1
#define BMPC_PID_REAR 2
int bmpc_fleet_set_pid(bmpc_t* h, int which, float kp, float ki, float kd, float dt);
int bmpc_fleet_reset_pid(bmpc_t* h, int which);
int bmpc_fleet_get_pid(bmpc_t* h, int which, int64_t count, float* integral, float* prev, uint8_t* first);
int bmpc_fleet_put_pid(bmpc_t* h, int which, int64_t count, const float* integral, const float* prev, const uint8_t* first);

/* The non-MPC branch of cycle() (src/...:214-225, config type neither tobiMPC nor mikMPC) for `count` vehicles:
 * look-ahead (dist or the table, as above) -> getTrajectoryPoint incl. the crossing clamp -> steering_front =
 * pidControllerFront.pid(position.y), steering_rear = pidControllerRear.pid(directory.angle()), target_speed,
 * target_distance, then the rule of :276-283.  Works on any handle (no MPC shape involved); the PID states persist in
 * HBM across ticks.  idx / flag may be NULL.  directory.angle() is atan2f on the device (<= 2 ulp against libm). */
int bmpc_fleet_tick_pid(bmpc_t* h, int64_t count, const float* velocity, const float* dist, int P, const int32_t* npoints,
                        const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                        int64_t stride_instance, int64_t stride_point, double* steering_front, double* steering_rear,
                        float* target_speed, float* target_distance, int32_t* idx, uint8_t* flag);

/* ---- device groups: one controller population sharded over several GPUs of this process -----------------------------
 * SURVEY.md §8(b) (`devices[]`, `ndev` at create) and §8(e): every dlib::mpc object is independent (mpc.h:216-347 touches
 * only its own members), so a population of `capacity` controllers is split along the instance axis over `ndev` devices.
 * A group owns one bmpc_t and one host worker thread per device; every group call runs on all members concurrently and
 * returns when all are done.  There is no device-to-device traffic: the only "collective" is the gather of the results
 * into the caller's host arrays, done by each member's own D2H copy.  Arguments and host layouts are those of the
 * single-handle call of the same name; `count` / `instance` are GLOBAL (0 .. capacity-1).
 *
 * Assignment of global instance i to a member is fixed at create (persistent controllers keep their state in place):
 *   BMPC_GROUP_INTERLEAVED  member i mod ndev, local slot i / ndev — every prefix [0, count) is balanced, heavy-tailed
 *                           iteration counts and blocked parameter sweeps are spread evenly (§8(e)); the members pack /
 *                           unpack their strided share of the host arrays in their worker threads
 *   BMPC_GROUP_CONTIGUOUS   member k owns [k*per, (k+1)*per), per = ceil(capacity / ndev) — the host arrays are used in
 *                           place (no packing); balanced only when count == capacity
 * `devices` may be NULL (devices 0 .. ndev-1) and may name a device more than once (two members then share it; used by
 * the single-GPU tests of this layer).  Results are bit-identical to one handle solving the same instances. */
typedef struct bmpc_group bmpc_group_t;
#define BMPC_GROUP_INTERLEAVED 0
#define BMPC_GROUP_CONTIGUOUS 1
int bmpc_group_create(bmpc_group_t** out, int S, int I, int N, int64_t capacity, int precision, int mode,
                      const int* devices, int ndev, int assignment);
void bmpc_group_destroy(bmpc_group_t* g);
int bmpc_group_size(bmpc_group_t* g);
/* Member k's handle (options, kernel info, device-resident results ...); owned by the group. */
bmpc_t* bmpc_group_member(bmpc_group_t* g, int k);
/* How many of the global instances [0, count) member k holds (its local slots 0 .. *local_count-1) and, for the
 * contiguous assignment, the global index of its local slot 0 (interleaved: k). */
int bmpc_group_shard(bmpc_group_t* g, int k, int64_t count, int64_t* first, int64_t* local_count);
int bmpc_group_set_model(bmpc_group_t* g, const double* A, const double* B, const double* C, const double* Q,
                         const double* R, const double* lower, const double* upper, int per_instance, int64_t count);
int bmpc_group_set_options(bmpc_group_t* g, double eps, uint64_t max_iterations);
int bmpc_group_set_target(bmpc_group_t* g, const double* val, int time, int per_instance, int64_t count);
int bmpc_group_reset(bmpc_group_t* g);
/* operator() for the global instances [0, count): x0[count][S] in; u0[count][I], iters[count], status[count] out (the
 * last two may be NULL) — H2D, solve and D2H of every member overlap with the others'. */
int bmpc_group_solve(bmpc_group_t* g, int64_t count, const double* x0, double* u0, uint32_t* iters, uint8_t* status,
                     unsigned flags);
/* controls[count][N][I] of the global instances first .. first+count-1 (bmpc_get_controls). */
int bmpc_group_get_controls(bmpc_group_t* g, int64_t first, int64_t count, double* controls);
/* Device time of the last solve: the maximum over the members (ms). */
int bmpc_group_last_kernel_ms(bmpc_group_t* g, float* ms);

#ifdef __cplusplus
}
#endif
#endif /* BMPC_H */
