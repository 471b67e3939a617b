// Internal definitions of libbmpc.so (host side): handle layout, error plumbing, generic layout kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/bmpc.h"
#include "arith.cuh"
#include "dispatch.h"

namespace bmpc {

void set_error(const std::string& msg);
int fail(int code, const std::string& msg);
int fail_cuda(cudaError_t e, const char* what);

#define BMPC_CUDA(call)                                                  \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) return ::bmpc::fail_cuda(e__, #call);    \
    } while (0)

// A growable device buffer (bytes).
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    bool view = false;            // a window into another DevBuf's allocation: never freed, never grown
    cudaError_t reserve(size_t n);
    cudaError_t alloc_zero(size_t n, cudaStream_t st);
    void release();
};

// Layout helpers (layout_kernels.cu).  AoS = instance-major host layout [count][len] of doubles;
// SoA = device layout [len][stride] of T (T = double or float, selected by elem_size 8 / 4).
cudaError_t aos_to_soa(const double* d_aos, void* d_soa, long long count, int len, long long stride, int elem_size,
                       cudaStream_t st);
cudaError_t soa_to_aos(const void* d_soa, double* d_aos, long long count, int len, long long stride, int elem_size,
                       cudaStream_t st);
// Ragged warm solve with per-step targets: keeps the logical targets of the instances [first, first+count) in place
// while the shared ring head advances by one (see k_rotate_targets).
cudaError_t rotate_targets(void* d_target, long long first, long long count, int S, int N, long long stride, int elem_size,
                           cudaStream_t st);
// Broadcast one host vector of `len` doubles (passed by value through a small device buffer) into [len][stride].
cudaError_t broadcast_rows(const double* d_vec, void* d_soa, long long count, int len, long long stride, int elem_size,
                           cudaStream_t st);
cudaError_t launch_extract(long long count, int P, const int32_t* npoints, const float* px, const float* py,
                           const float* dx, const float* dy, const float* vel, long long stride_instance,
                           long long stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* opx,
                           float* opy, float* odx, float* ody, float* ovel, cudaStream_t st);

cudaError_t launch_extract_horizon(long long count, int P, const int32_t* npoints, const float* px, const float* py,
                                   const float* dx, const float* dy, const float* vel, long long stride_instance,
                                   long long stride_point, const float* d0, const float* step, int steps, int32_t* idx,
                                   uint8_t* flag, float* opx, float* opy, float* odx, float* ody, float* ovel, cudaStream_t st);

// Small-call paths (layout_kernels.cu): values passed as kernel arguments.
constexpr int kSmallVals = 64;
struct SmallVals { double x[kSmallVals]; };
struct PutPlan { int n; int off[8]; int len[8]; void* dst[8]; };
cudaError_t put_vals(const SmallVals& v, const PutPlan& plan, long long first, long long count, long long stride, int elem_size,
                     cudaStream_t st, unsigned long long* zero8 = nullptr);
constexpr int kSmallSolve = 16;    // solves of up to this many instances use the packed one-copy result path
__host__ __device__ inline size_t small_results_err_offset(int count, int I) { return (((size_t)count * I * 8 + (size_t)count * 5) + 7) & ~(size_t)7; }
cudaError_t pack_results(const void* u0, const uint32_t* iters, const uint8_t* status, const unsigned long long* queue, int count,
                         int I, long long stride, int elem_size, unsigned char* out, cudaStream_t st,
                         unsigned long long* flag = nullptr, unsigned long long seq = 0);

// bmpc_solve_resident_async on the instances [first, first + count) of the handle (cold solves with one target per
// instance only: the fleet tick's chunks).
int solve_range_async(bmpc_handle* h, long long first, long long count, unsigned flags);

const ShapeOps* find_shape(int S, int I, int N);
const ShapeOps* select_pattern(int S, int I, int N, const double* A, const double* B, const double* C, const double* Q,
                               long long count, bool as_float);

}  // namespace bmpc

struct bmpc_handle {
    int S, I, N, NI;
    int precision, mode, device;
    int elem;                     // sizeof(T)
    long long cap;
    const bmpc::ShapeOps* ops;         // dense unit of the shape (constructor kernel)
    const bmpc::ShapeOps* solve_ops;   // pattern-specialised unit chosen by bmpc_set_model for the current model
    cudaStream_t stream = nullptr;       // the stream calls run on (own_stream unless bmpc_set_stream borrowed one)
    cudaStream_t own_stream = nullptr;
    cudaStream_t copy_stream = nullptr;  // fleet tick: uploads of the next chunk while this one computes
    cudaEvent_t ev_chunk[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_mid = nullptr;   // solve start / end / between stage A and stage B
    bool timed_split = false;
    bool timed = false;
    int sm_count = 0;
    int64_t launches = 0;

    // options (mpc.h:103-104)
    double eps = 0.01;
    uint32_t max_iter = 10000;

    // model
    bool have_model = false, per_instance = false, specialize = true;
    long long model_count = 0;
    std::vector<double> hA, hB, hC, hQ, hR, hlo, hhi, hqdiag;   // shared model (host copies, already rounded to T)
    double hlambda = 0;
    bmpc::DevBuf mA, mB, mC, mQ, mR, mlo, mhi, mlambda, mqdiag;  // SoA [elem][cap]

    // state
    bmpc::DevBuf results;         // ONE allocation: u0 | iters | status (each 256-byte aligned) — u0 / iters / status below are views
    bmpc::DevBuf x0, u0, iters, status, controls, vstate, target_const, target_full, queue;
    bmpc::DevBuf pg_list, park, park_iter;   // split solve: hand-over ring and the by-slot state of parked instances (pg_kernel.cuh)
    int small_cfg = 15;              // small-call shortcuts (developer knob BMPC_SMALL, bits: 1 counters zeroed by the x0 kernel,
                                     // 2 no stage event, 4 results written to pinned host memory + polled, 8 one kernel where no pair kernel)
    int small_now = 0;               // the bits in force for the solve being launched (small_solve sets and clears it)
    unsigned long long small_seq = 0;
    bool ring_dirty = true;          // pg_list may hold entries of an earlier launch (bmpc_solve_resident_async)
    size_t pg_ring = 0;              // ring capacity (a power of two >= capacity + lanes in flight)
    bool sym_ok = false;             // the model meets the preconditions of stage B's symmetric-bound kernel (bmpc_set_model)
    bool slice = true;               // stage B may time-slice small batches
    bool retire = true;              // stage B may retire surplus warps near the end of a launch (bmpc_set_split(h, 2) turns it off)
    bool sym = true;                 // stage B may use its symmetric-bound kernel (bmpc_set_split(h, 3) turns it off)
    bool split = true;               // bmpc_set_split; the two-stage solve is used whenever the shape offers it
    bool pair = true;                // stage C (pair_kernel.cuh) may be used: tails, small batches, per-instance models, long horizons
    long long pair_threshold = 0;    // stage B hands over to stage C once this many instances or fewer are unfinished
    long long pair_direct_max = 0;   // batches up to this size skip stage B
    bmpc::DevBuf c_list, c_park, c_iter;   // stage C's list (by-slot rows of stride pair_threshold)
    bool full_targets = false;
    int head = 0;

    // staging
    bmpc::DevBuf stage;           // device, doubles in host (AoS) layout
    bmpc::DevBuf small;           // device, a few hundred doubles
    void* pinned = nullptr;       // host pinned mirror for small transfers
    size_t pinned_bytes = 0;
    bmpc::DevBuf ex[16];          // trajectory-point extraction buffers

    // fleet tick (fleet_tick.cu): look-up tables, crossing clamp, per-vehicle PID controllers
    bmpc::DevBuf lut[2];          // 2n floats: vx | vy   (BMPC_LOOKUP_DISTANCE, BMPC_LOOKUP_VELOCITY)
    int lut_n[2] = {0, 0};
    bool foh = false;             // drive mode FOH: look-ahead = foh_min + v * foh_slope (src/...:70-73)
    float foh_min = 0.6f, foh_slope = 0.1f;
    bool crossing_on = true;      // the reference always applies the clamp (src/...:445-473)
    bmpc_crossing_config crossing{1.0f, 0.35f, 1.0f, 0.05f};
    float pid_gain[3][4] = {{1, 0, 0, 0.01f}, {1, 1, 0, 0.01f}, {1, 1, 0, 0.01f}};   // src/...:296-298 defaults
    bmpc::DevBuf pid_state[3];    // float integral[cap] | float prev[cap] | uint8 first[cap]
    bool pid_ready[3] = {false, false, false};

    // closed loop (closed_loop.cu)
    bool cl_ready = false;
    long long cl_count = 0;
    bmpc_cl_config cl_cfg{};
    bmpc::DevBuf cl_path;         // 4 x npath doubles: x, y, tx, ty
    bmpc::DevBuf cl_pose;         // [4][cap] doubles: X, Y, c, s
    bmpc::DevBuf cl_tgt;          // [2][cap] doubles: last y_soll, phi_soll
    bmpc::DevBuf cl_idx;          // int32 [cap]
    bmpc::DevBuf cl_itot;         // uint64 [cap]
    bmpc::DevBuf cl_imax;         // uint32 [cap]
    bool cl_horizon = false;      // a target per horizon step (bmpc_cl_set_horizon_targets)
    float cl_h_d0 = 0.f, cl_h_step = 0.f;
};
