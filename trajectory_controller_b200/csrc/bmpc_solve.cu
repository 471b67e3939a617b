// libbmpc.so — C ABI part 2: targets, controller state, solve, trajectory-point extraction.
#include <cstring>

#include "bmpc_internal.h"

using namespace bmpc;

namespace {

// Physical ring slot of horizon step `time` (targets shift by one per warm solve, mpc.h:236-237).
inline int slot_of(const bmpc_handle* h, int time) { int s = h->head + time; return s >= h->N ? s - h->N : s; }

int ensure_full_targets(bmpc_handle* h)
{
    if (h->full_targets) return BMPC_OK;
    const size_t row = (size_t)h->cap * h->S * h->elem;
    BMPC_CUDA(h->target_full.reserve(row * h->N));
    for (int t = 0; t < h->N; ++t)   // every horizon step starts as the constant target
        BMPC_CUDA(cudaMemcpyAsync((char*)h->target_full.p + row * t, h->target_const.p, row, cudaMemcpyDeviceToDevice, h->stream));
    h->full_targets = true;
    h->head = 0;
    return BMPC_OK;
}

int write_target_rows(bmpc_handle* h, void* dst_rows, const double* val, int per_instance, long long count)
{
    const int S = h->S;
    if (per_instance) {
        const size_t bytes = (size_t)count * S * 8;
        BMPC_CUDA(h->stage.reserve(bytes));
        BMPC_CUDA(cudaMemcpyAsync(h->stage.p, val, bytes, cudaMemcpyHostToDevice, h->stream));
        BMPC_CUDA(aos_to_soa((const double*)h->stage.p, dst_rows, count, S, h->cap, h->elem, h->stream));
    } else if (S <= kSmallVals) {
        // one vector for every instance: passed as a kernel argument (no staging copy to wait for)
        SmallVals v; PutPlan plan{};
        plan.n = 1; plan.off[0] = 0; plan.len[0] = S; plan.dst[0] = dst_rows;
        for (int e = 0; e < S; ++e) v.x[e] = val[e];
        BMPC_CUDA(put_vals(v, plan, 0, h->cap, h->cap, h->elem, h->stream));
    } else {
        std::memcpy(h->pinned, val, (size_t)S * 8);
        BMPC_CUDA(cudaMemcpyAsync(h->small.p, h->pinned, (size_t)S * 8, cudaMemcpyHostToDevice, h->stream));
        BMPC_CUDA(broadcast_rows((const double*)h->small.p, dst_rows, h->cap, S, h->cap, h->elem, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));           // the pinned mirror is free again
    }
    h->launches++;
    return BMPC_OK;
}

}  // namespace

extern "C" {

int bmpc_set_target(bmpc_t* h, const double* val, int time, int per_instance, int64_t count)
{
    if (!h || !val) return fail(BMPC_E_INVALID, "bmpc_set_target: null argument");
    if (time < -1 || time >= h->N) return fail(BMPC_E_INVALID, "bmpc_set_target: time must be -1 or < horizon (mpc.h:146 requires)");
    if (per_instance && (count < 1 || count > h->cap)) return fail(BMPC_E_INVALID, "bmpc_set_target: count out of range");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t row = (size_t)h->cap * h->S * h->elem;
    int rc;
    if (time < 0) {                                           // set_target(val): every horizon step (mpc.h:157-163)
        if ((rc = write_target_rows(h, h->target_const.p, val, per_instance, count))) return rc;
        if (h->full_targets)
            for (int t = 0; t < h->N; ++t) {
                if (per_instance)   // only `count` instances change: row-wise 2D copy of the first `count` columns
                    BMPC_CUDA(cudaMemcpy2DAsync((char*)h->target_full.p + row * t, (size_t)h->cap * h->elem, h->target_const.p,
                                                (size_t)h->cap * h->elem, (size_t)count * h->elem, h->S,
                                                cudaMemcpyDeviceToDevice, h->stream));
                else
                    BMPC_CUDA(cudaMemcpyAsync((char*)h->target_full.p + row * t, h->target_const.p, row, cudaMemcpyDeviceToDevice, h->stream));
            }
    } else {                                                  // set_target(val, time) (mpc.h:142-155)
        if ((rc = ensure_full_targets(h))) return rc;
        if ((rc = write_target_rows(h, (char*)h->target_full.p + row * slot_of(h, time), val, per_instance, count))) return rc;
    }
    // per-instance values went through the staging buffer, which the next call may reuse; a shared vector travelled
    // as a kernel argument and needs no wait (the call is stream-ordered like every other)
    if (per_instance) BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_get_target(bmpc_t* h, int64_t instance, int time, double* val)
{
    if (!h || !val) return fail(BMPC_E_INVALID, "bmpc_get_target: null argument");
    if (time < 0 || time >= h->N) return fail(BMPC_E_INVALID, "bmpc_get_target: time must be < horizon (mpc.h:133 requires)");
    if (instance < 0 || instance >= h->cap) return fail(BMPC_E_INVALID, "bmpc_get_target: instance out of range");
    BMPC_CUDA(cudaSetDevice(h->device));
    const int es = h->elem;
    const size_t row = (size_t)h->cap * h->S * es;
    const char* base = h->full_targets ? (const char*)h->target_full.p + row * slot_of(h, time) : (const char*)h->target_const.p;
    BMPC_CUDA(cudaMemcpy2DAsync(h->pinned, es, base + instance * es, (size_t)h->cap * es, es, h->S, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    for (int r = 0; r < h->S; ++r) val[r] = es == 8 ? ((double*)h->pinned)[r] : (double)((float*)h->pinned)[r];
    return BMPC_OK;
}

int bmpc_reset(bmpc_t* h)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_reset: null handle");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(cudaMemsetAsync(h->controls.p, 0, h->controls.bytes, h->stream));
    BMPC_CUDA(cudaMemsetAsync(h->vstate.p, 0, h->vstate.bytes, h->stream));
    BMPC_CUDA(cudaMemsetAsync(h->target_const.p, 0, h->target_const.bytes, h->stream));
    if (h->full_targets) BMPC_CUDA(cudaMemsetAsync(h->target_full.p, 0, h->target_full.bytes, h->stream));
    h->head = 0;
    return BMPC_OK;                     // stream-ordered: nothing of the caller's is referenced, nothing to wait for
}

int bmpc_get_controls(bmpc_t* h, int64_t first, int64_t count, double* controls)
{
    if (!h || !controls) return fail(BMPC_E_INVALID, "bmpc_get_controls: null argument");
    if (first < 0 || count < 1 || first + count > h->cap) return fail(BMPC_E_INVALID, "bmpc_get_controls: range out of capacity");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t bytes = (size_t)count * h->NI * 8;
    BMPC_CUDA(h->stage.reserve(bytes));
    BMPC_CUDA(soa_to_aos((const char*)h->controls.p + first * h->elem, (double*)h->stage.p, count, h->NI, h->cap, h->elem, h->stream));
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(controls, h->stage.p, bytes, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_set_controls(bmpc_t* h, int64_t first, int64_t count, const double* controls)
{
    if (!h || !controls) return fail(BMPC_E_INVALID, "bmpc_set_controls: null argument");
    if (first < 0 || count < 1 || first + count > h->cap) return fail(BMPC_E_INVALID, "bmpc_set_controls: range out of capacity");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t bytes = (size_t)count * h->NI * 8;
    BMPC_CUDA(h->stage.reserve(bytes));
    BMPC_CUDA(cudaMemcpyAsync(h->stage.p, controls, bytes, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(aos_to_soa((const double*)h->stage.p, (char*)h->controls.p + first * h->elem, count, h->NI, h->cap, h->elem, h->stream));
    h->launches++;
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_copy_state(bmpc_t* dst, bmpc_t* src)
{
    if (!dst || !src) return fail(BMPC_E_INVALID, "bmpc_copy_state: null handle");
    if (dst == src) return BMPC_OK;
    if (dst->S != src->S || dst->I != src->I || dst->N != src->N || dst->precision != src->precision ||
        dst->cap != src->cap || dst->device != src->device)
        return fail(BMPC_E_INVALID, "bmpc_copy_state: handles differ in shape, precision, capacity or device");
    BMPC_CUDA(cudaSetDevice(dst->device));
    BMPC_CUDA(cudaStreamSynchronize(src->stream));
    BMPC_CUDA(cudaMemcpyAsync(dst->controls.p, src->controls.p, src->controls.bytes, cudaMemcpyDeviceToDevice, dst->stream));
    BMPC_CUDA(cudaMemcpyAsync(dst->vstate.p, src->vstate.p, src->vstate.bytes, cudaMemcpyDeviceToDevice, dst->stream));
    BMPC_CUDA(cudaMemcpyAsync(dst->target_const.p, src->target_const.p, src->target_const.bytes, cudaMemcpyDeviceToDevice, dst->stream));
    if (src->full_targets) {
        BMPC_CUDA(dst->target_full.reserve(src->target_full.bytes));
        BMPC_CUDA(cudaMemcpyAsync(dst->target_full.p, src->target_full.p, src->target_full.bytes, cudaMemcpyDeviceToDevice, dst->stream));
    }
    dst->full_targets = src->full_targets;
    dst->head = src->head;
    dst->eps = src->eps;
    dst->max_iter = src->max_iter;
    BMPC_CUDA(cudaStreamSynchronize(dst->stream));
    return BMPC_OK;
}

static int upload_x0_async(bmpc_t* h, int64_t count, const double* x0)
{
    if (!h || !x0) return fail(BMPC_E_INVALID, "bmpc_upload_x0: null argument");
    if (count < 1 || count > h->cap) return fail(BMPC_E_INVALID, "bmpc_upload_x0: count out of capacity");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t bytes = (size_t)count * h->S * 8;
    BMPC_CUDA(h->stage.reserve(bytes + (size_t)count * h->I * 8));
    BMPC_CUDA(cudaMemcpyAsync(h->stage.p, x0, bytes, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(aos_to_soa((const double*)h->stage.p, h->x0.p, count, h->S, h->cap, h->elem, h->stream));
    h->launches++;
    return BMPC_OK;
}

int bmpc_upload_x0(bmpc_t* h, int64_t count, const double* x0)
{
    int rc = upload_x0_async(h, count, x0);
    if (rc) return rc;
    BMPC_CUDA(cudaStreamSynchronize(h->stream));   // synchronous on return (bmpc.h): x0 may be reused by the caller
    return BMPC_OK;
}

int bmpc_solve_resident_async(bmpc_t* h, int64_t count, unsigned flags)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_solve: null handle");
    return bmpc::solve_range_async(h, 0, count, flags);
}

}  // extern "C"

int bmpc::solve_range_async(bmpc_handle* h, long long first, long long count, unsigned flags)
{
    if (!h->have_model) return fail(BMPC_E_STATE, "bmpc_solve: bmpc_set_model has not been called");
    if (first < 0 || count < 1 || first + count > h->cap) return fail(BMPC_E_INVALID, "bmpc_solve: count out of capacity");
    if (h->per_instance && first + count > h->model_count) return fail(BMPC_E_INVALID, "bmpc_solve: more instances than per-instance models");
    if (first != 0 && ((flags & BMPC_WARM) || h->full_targets)) return fail(BMPC_E_INVALID, "bmpc_solve: a sub-range solve must be cold with one target per instance");
    BMPC_CUDA(cudaSetDevice(h->device));
    const bool warm = (flags & BMPC_WARM) != 0;
    GenericSolve g{};
    g.A = h->hA.data(); g.B = h->hB.data(); g.C = h->hC.data(); g.Q = h->hQ.data(); g.R = h->hR.data();
    g.lo = h->hlo.data(); g.hi = h->hhi.data(); g.lambda = h->hlambda; g.qdiag = h->hqdiag.data();
    g.eps = h->eps; g.max_iter = h->max_iter;
    g.warm = warm; g.keep_controls = warm || (flags & BMPC_KEEP_CONTROLS);
    g.target_full = h->full_targets; g.head = h->head;
    g.count = count; g.stride = h->cap; g.first = first;
    g.x0 = h->x0.p; g.controls = h->controls.p; g.vstate = h->vstate.p;
    g.target = h->full_targets ? h->target_full.p : h->target_const.p;
    g.u0 = h->u0.p; g.iters = (uint32_t*)h->iters.p; g.status = (uint8_t*)h->status.p;
    g.mA = h->mA.p; g.mB = h->mB.p; g.mC = h->mC.p; g.mQ = h->mQ.p; g.mR = h->mR.p; g.mlo = h->mlo.p; g.mhi = h->mhi.p;
    g.mlambda = h->mlambda.p; g.mqdiag = h->mqdiag.p;
    g.queue = (unsigned long long*)h->queue.p;
    // Two-stage solve (pg_kernel.cuh): needs the stage-B kernel for this shape, iterations beyond the 50 coordinate
    // steps, and eps > 0 (stage B tests "|df| >= eps" instead of carrying max_df; bmpc_set_options enforces eps > 0).
    // (a small call of a shape without a lane-pair kernel runs as ONE kernel: a launch and a hand-over less, same results)
    const bool one_kernel_small = (h->small_now & 8) && !(h->solve_ops->has_pair && h->c_list.p != nullptr && h->pair && h->sym && h->sym_ok);
    const bool split = h->split && h->solve_ops->has_pg && h->max_iter > 50 && h->eps > 0 && !one_kernel_small;
    g.split = split; g.pg_list = (uint32_t*)h->pg_list.p;
    g.sym = 0;
    if (split && h->sym_ok && h->sym) {   // the symmetric-bound stage-B kernel, where the shape has one (register-resident horizons)
        KernelInfo kp0;
        BMPC_CUDA(h->solve_ops->pg_info(h->precision, h->mode, h->per_instance, 0, &kp0));
        g.sym = kp0.sym_ok;
    }
    // Stage C (pair_kernel.cuh): one instance per lane pair.  Same preconditions as the symmetric-bound kernel (the model
    // check of bmpc_set_model; stage A keeps non-finite instances to itself when g.sym is set).  It replaces stage B
    // where the per-lane layout is slow or the batch is small — per-instance models (model block in shared memory),
    // horizons whose u / MM do not fit a lane's registers, batches up to pair_direct_max — and takes over the last
    // pair_threshold instances of every other launch.
    const bool pair_ok = split && h->pair && h->sym && h->sym_ok && h->solve_ops->has_pair && h->c_list.p != nullptr;
    const bool lane_regs = (long long)h->NI * h->elem <= 256;          // PgLayout::UREG: stage B runs from registers
    const bool pair_direct = pair_ok && (count <= h->pair_direct_max || h->per_instance || !lane_regs);
    const bool pair_tail = pair_ok && !pair_direct && g.sym && h->pair_threshold > 0;
    if (pair_direct) g.sym = 1;
    g.c_direct = pair_direct; g.c_threshold = pair_tail ? (unsigned)h->pair_threshold : 0u;
    g.c_list = (uint32_t*)h->c_list.p; g.c_park = h->c_park.p; g.c_iter = (uint32_t*)h->c_iter.p;
    g.c_stride = (unsigned)h->pair_threshold;
    g.pg_mask = (unsigned)(h->pg_ring - 1);
    g.park = h->park.p; g.park_iter = (uint32_t*)h->park_iter.p;
    // Stage-B scheduling (pg_kernel.cuh).  Slot reuse (parking) is what retirement and fairness slices need; it costs a
    // fence per resumed instance, so very large batches, whose tail is a small share of the launch, run without it.
    // Fairness slices pay when lanes get only a few instances each (measured, scripts/slice_threshold.py: +4..10 %
    // between 1 and 4 instances per lane, -4 % below, neutral above); 8 warps x 32 lanes per SM are resident.
    const long long lanes = 1ll * h->sm_count * 256;
    const bool small_enough = count <= 8 * lanes;
    g.pg_slice = h->slice && small_enough && count > lanes && count <= 4 * lanes;
    // retirement needs more than one warp per SM sub-partition's worth of instances to have anything to retire
    const bool retire = h->retire && small_enough && count > 1ll * h->sm_count * 128;
    g.pg_reuse = (g.pg_slice || retire) ? (1 | (retire ? 2 : 0)) : 0;   // bit 0: parking, bit 1: retirement
    if (pair_direct) { g.pg_slice = 0; g.pg_reuse = 0; }                // stage B does not run: stage A's ring entries stay behind
    // A launch that recycles ring slots needs every slot empty at the start and leaves it so; one that does not leaves
    // stage A's entries behind.  So the ring is wiped only when a recycling launch follows a non-recycling one.
    if (split && g.pg_reuse && h->ring_dirty) BMPC_CUDA(cudaMemsetAsync(h->pg_list.p, 0xff, h->pg_ring * 4, h->stream));
    if (split) h->ring_dirty = !g.pg_reuse;
    if (!(h->small_now & 1)) BMPC_CUDA(cudaMemsetAsync(h->queue.p, 0, 64, h->stream));   // (small calls: zeroed by the x0 kernel)
    BMPC_CUDA(cudaEventRecord(h->ev0, h->stream));
    KernelInfo ki;
    BMPC_CUDA(h->solve_ops->info(h->precision, h->mode, h->per_instance, &ki));
    // Persistent grid: one CTA per SM.  A batch that does not fill every warp is spread over all SMs, one warp per
    // SM sub-partition first (a warp alone on its scheduler runs a pass in ~4 200 cycles, two share it at ~6 050 for
    // both): each CTA lets only its first `warp_limit` warps take work.
    const long long need_warps = (count + 31) / 32;
    int grid = (int)(need_warps < h->sm_count ? need_warps : h->sm_count);
    auto limit_for = [&](int warps_per_cta) {
        const long long per = (need_warps + grid - 1) / grid;
        return (int)(per < warps_per_cta ? per : warps_per_cta);
    };
    g.warp_limit_a = limit_for(ki.warps);
    BMPC_CUDA(h->solve_ops->solve(h->precision, h->mode, h->per_instance, g, grid, h->stream));
    h->launches++;
    const bool stage_event = split && !(h->small_now & 2);             // small calls: bmpc_last_stage_ms reports stage B as 0
    h->timed_split = stage_event;
    if (split) {
        if (stage_event) BMPC_CUDA(cudaEventRecord(h->ev_mid, h->stream));
        if (!pair_direct) {
            KernelInfo kp;
            BMPC_CUDA(h->solve_ops->pg_info(h->precision, h->mode, h->per_instance, g.sym, &kp));
            g.warp_limit_b = limit_for(kp.warps);
            BMPC_CUDA(h->solve_ops->pg(h->precision, h->mode, h->per_instance, g, grid, h->stream));
            h->launches++;
        }
        if (pair_direct || pair_tail) {
            // 16 instances per warp; the kernel itself lets only as many warps per CTA work as its list needs
            const long long most = pair_direct ? count : (count < h->pair_threshold ? count : h->pair_threshold);
            const long long pw = (most + 15) / 16;
            const int grid_c = (int)(pw < h->sm_count ? pw : h->sm_count);
            BMPC_CUDA(h->solve_ops->pair(h->precision, h->mode, h->per_instance, g, grid_c, h->stream));
            h->launches++;
        }
    }
    BMPC_CUDA(cudaEventRecord(h->ev1, h->stream));
    h->timed = true;
    if (warm && h->full_targets) {                                                  // mpc.h:236-237
        h->head = h->head + 1 == h->N ? 0 : h->head + 1;
        // the head is shared by all slots of the handle, but only the first `count` controllers ran operator():
        // the others keep their logical targets (their rows follow the head)
        if (count < h->cap) {
            BMPC_CUDA(rotate_targets(h->target_full.p, count, h->cap - count, h->S, h->N, h->cap, h->elem, h->stream));
            h->launches++;
        }
    }
    return BMPC_OK;
}

extern "C" {

int bmpc_download(bmpc_t* h, int64_t count, double* u0, uint32_t* iters, uint8_t* status)
{
    if (!h || !u0) return fail(BMPC_E_INVALID, "bmpc_download: null argument");
    if (count < 1 || count > h->cap) return fail(BMPC_E_INVALID, "bmpc_download: count out of capacity");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t off = (size_t)count * h->S * 8, bytes = (size_t)count * h->I * 8;
    BMPC_CUDA(h->stage.reserve(off + bytes));
    double* d_out = (double*)((char*)h->stage.p + off);
    BMPC_CUDA(soa_to_aos(h->u0.p, d_out, count, h->I, h->cap, h->elem, h->stream));
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(u0, d_out, bytes, cudaMemcpyDeviceToHost, h->stream));
    if (iters) BMPC_CUDA(cudaMemcpyAsync(iters, h->iters.p, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    if (status) BMPC_CUDA(cudaMemcpyAsync(status, h->status.p, (size_t)count, cudaMemcpyDeviceToHost, h->stream));
    unsigned long long* err = (unsigned long long*)h->pinned;
    BMPC_CUDA(cudaMemcpyAsync(err, (char*)h->queue.p + 32, 8, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    if (*err) h->ring_dirty = true;
    if (*err) return fail(BMPC_E_STATE, "bmpc_download: the last solve gave up (hand-over ring overrun); results are incomplete");
    return BMPC_OK;
}

// A solve of a few instances (the facade's operator(), count = 1): x0 travels as a kernel argument and the results come
// back in ONE packed copy — two launches and one synchronisation around the solve kernels.
static int small_solve(bmpc_t* h, int count, const double* x0, double* u0, uint32_t* iters, uint8_t* status, unsigned flags)
{
    BMPC_CUDA(cudaSetDevice(h->device));
    const int opts = h->small_cfg;
    SmallVals v; PutPlan plan{};
    for (int b = 0; b < count; ++b) {        // one launch per instance keeps the argument block small; count is tiny
        plan.n = 1; plan.off[0] = 0; plan.len[0] = h->S; plan.dst[0] = h->x0.p;
        for (int e = 0; e < h->S; ++e) v.x[e] = x0[(size_t)b * h->S + e];
        // the first of them also zeroes the launch counters of the solve (instead of a memset of its own)
        BMPC_CUDA(put_vals(v, plan, b, 1, h->cap, h->elem, h->stream, (b == 0 && (opts & 1)) ? (unsigned long long*)h->queue.p : nullptr));
        h->launches++;
    }
    h->small_now = opts;
    int rc = bmpc_solve_resident_async(h, count, flags);
    h->small_now = 0;
    if (rc) return rc;
    const size_t err_off = small_results_err_offset(count, h->I), bytes = err_off + 8;
    unsigned char* pin = (unsigned char*)h->pinned;
    if (opts & 4) {
        // Results straight into the pinned host block (device-visible: unified addressing), then a sequence word the
        // host polls: no copy operation and no stream synchronisation on the way back.
        volatile unsigned long long* flag = (volatile unsigned long long*)(pin + bytes);
        const unsigned long long seq = 0x5EED000000000000ull | (++h->small_seq & 0xffffffffffffull);   // never a plausible number another call leaves there
        BMPC_CUDA(pack_results(h->u0.p, (const uint32_t*)h->iters.p, (const uint8_t*)h->status.p, (const unsigned long long*)h->queue.p,
                               count, h->I, h->cap, h->elem, pin, h->stream, (unsigned long long*)flag, seq));
        h->launches++;
        for (unsigned spins = 1; *flag != seq; ++spins) {
#if defined(__x86_64__) || defined(__i386__)
            __builtin_ia32_pause();
#endif
            if ((spins & 0x3ffu) == 0) {     // a failed launch never writes the flag: ask the stream now and then
                const cudaError_t q = cudaStreamQuery(h->stream);
                if (q != cudaSuccess && q != cudaErrorNotReady) return bmpc::fail_cuda(q, "bmpc_solve (small call)");
            }
        }
    } else {
        BMPC_CUDA(pack_results(h->u0.p, (const uint32_t*)h->iters.p, (const uint8_t*)h->status.p, (const unsigned long long*)h->queue.p,
                               count, h->I, h->cap, h->elem, (unsigned char*)h->small.p, h->stream));
        h->launches++;
        BMPC_CUDA(cudaMemcpyAsync(pin, h->small.p, bytes, cudaMemcpyDeviceToHost, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
    }
    std::memcpy(u0, pin, (size_t)count * h->I * 8);
    if (iters) std::memcpy(iters, pin + (size_t)count * h->I * 8, (size_t)count * 4);
    if (status) std::memcpy(status, pin + (size_t)count * h->I * 8 + (size_t)count * 4, (size_t)count);
    unsigned long long err; std::memcpy(&err, pin + err_off, 8);
    if (err) { h->ring_dirty = true; return fail(BMPC_E_STATE, "bmpc_solve: the solve gave up (hand-over ring overrun); results are incomplete"); }
    return BMPC_OK;
}

int bmpc_solve(bmpc_t* h, int64_t count, const double* x0, double* u0, uint32_t* iters, uint8_t* status, unsigned flags)
{
    int rc;
    if (h && x0 && u0 && count >= 1 && count <= kSmallSolve && count <= h->cap && (long long)h->S <= kSmallVals &&
        small_results_err_offset((int)count, h->I) + 16 <= h->pinned_bytes)
        return small_solve(h, (int)count, x0, u0, iters, status, flags);
    if ((rc = upload_x0_async(h, count, x0))) return rc;   // x0 stays valid: bmpc_download synchronises before returning
    if ((rc = bmpc_solve_resident_async(h, count, flags))) return rc;
    return bmpc_download(h, count, u0, iters, status);
}

int bmpc_sync(bmpc_t* h)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_sync: null handle");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_device_results(bmpc_t* h, void** u0, void** iters, void** status)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_device_results: null handle");
    if (u0) *u0 = h->u0.p;
    if (iters) *iters = h->iters.p;
    if (status) *status = h->status.p;
    return BMPC_OK;
}

int bmpc_device_results_block(bmpc_t* h, void** base, uint64_t* bytes, uint64_t* iters_offset, uint64_t* status_offset)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_device_results_block: null handle");
    if (base) *base = h->results.p;
    if (bytes) *bytes = h->results.bytes;
    if (iters_offset) *iters_offset = (uint64_t)((char*)h->iters.p - (char*)h->results.p);
    if (status_offset) *status_offset = (uint64_t)((char*)h->status.p - (char*)h->results.p);
    return BMPC_OK;
}

int bmpc_last_kernel_ms(bmpc_t* h, float* ms)
{
    if (!h || !ms) return fail(BMPC_E_INVALID, "bmpc_last_kernel_ms: null argument");
    if (!h->timed) return fail(BMPC_E_STATE, "bmpc_last_kernel_ms: no solve launched yet");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(cudaEventSynchronize(h->ev1));
    BMPC_CUDA(cudaEventElapsedTime(ms, h->ev0, h->ev1));
    return BMPC_OK;
}

int bmpc_last_stage_ms(bmpc_t* h, float* stage_a_ms, float* stage_b_ms)
{
    if (!h || !stage_a_ms || !stage_b_ms) return fail(BMPC_E_INVALID, "bmpc_last_stage_ms: null argument");
    if (!h->timed) return fail(BMPC_E_STATE, "bmpc_last_stage_ms: no solve launched yet");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(cudaEventSynchronize(h->ev1));
    if (h->timed_split) {
        BMPC_CUDA(cudaEventElapsedTime(stage_a_ms, h->ev0, h->ev_mid));
        BMPC_CUDA(cudaEventElapsedTime(stage_b_ms, h->ev_mid, h->ev1));
    } else {
        BMPC_CUDA(cudaEventElapsedTime(stage_a_ms, h->ev0, h->ev1));
        *stage_b_ms = 0.f;
    }
    return BMPC_OK;
}

int bmpc_pg_kernel_info(bmpc_t* h, int* grid, int* block, int* smem_bytes, int* regs)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_pg_kernel_info: null handle");
    if (!h->solve_ops->has_pg) return fail(BMPC_E_UNSUPPORTED, "bmpc_pg_kernel_info: this shape has no stage-B kernel (one-kernel solve)");
    BMPC_CUDA(cudaSetDevice(h->device));
    KernelInfo ki;
    BMPC_CUDA(h->solve_ops->pg_info(h->precision, h->mode, h->per_instance, h->sym_ok && h->sym, &ki));
    if (grid) *grid = h->sm_count;
    if (block) *block = ki.block;
    if (smem_bytes) *smem_bytes = ki.smem_bytes;
    if (regs) *regs = ki.regs;
    return BMPC_OK;
}

int64_t bmpc_launch_count(bmpc_t* h) { return h ? h->launches : 0; }

int bmpc_kernel_info(bmpc_t* h, int* grid, int* block, int* smem_bytes, int* regs)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_kernel_info: null handle");
    BMPC_CUDA(cudaSetDevice(h->device));
    KernelInfo ki;
    BMPC_CUDA(h->solve_ops->info(h->precision, h->mode, h->per_instance, &ki));
    if (grid) *grid = h->sm_count;
    if (block) *block = ki.block;
    if (smem_bytes) *smem_bytes = ki.smem_bytes;
    if (regs) *regs = ki.regs;
    return BMPC_OK;
}

int bmpc_extract_points(bmpc_t* h, int64_t count, int P, const int32_t* npoints, const float* px, const float* py,
                        const float* dx, const float* dy, const float* vel, int64_t stride_instance,
                        int64_t stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* out_px,
                        float* out_py, float* out_dx, float* out_dy, float* out_vel)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_extract_points: null handle");
    if (count < 1 || P < 0 || !dist || !idx || !flag || !out_px || !out_py || !out_dx || !out_dy || !out_vel)
        return fail(BMPC_E_INVALID, "bmpc_extract_points: bad argument");
    if (P > 0 && (!px || !py || !dx || !dy || !vel)) return fail(BMPC_E_INVALID, "bmpc_extract_points: null trajectory array");
    if (stride_instance < 0 || stride_point < 0) return fail(BMPC_E_INVALID, "bmpc_extract_points: negative stride");
    BMPC_CUDA(cudaSetDevice(h->device));
    if (count == 1 && stride_point == 1 && (size_t)P * 20 + 64 <= h->pinned_bytes / 2) {
        // One polyline (the module's getTrajectoryPoint): the five arrays and the distance are packed into the pinned
        // mirror and cross in ONE copy each way instead of six in and seven out.
        int n = npoints ? npoints[0] : P;
        if (n > P) n = P;
        if (n < 0) n = 0;
        float* pin = (float*)h->pinned;
        const float* src1[5] = {px, py, dx, dy, vel};
        for (int a = 0; a < 5; ++a)
            if (n) std::memcpy(pin + (size_t)a * n, src1[a], (size_t)n * 4);
        pin[(size_t)5 * n] = dist[0];
        const size_t in_bytes = ((size_t)5 * n + 1) * 4;
        BMPC_CUDA(h->ex[0].reserve(in_bytes));
        BMPC_CUDA(h->ex[7].reserve(32));
        BMPC_CUDA(cudaMemcpyAsync(h->ex[0].p, pin, in_bytes, cudaMemcpyHostToDevice, h->stream));
        const float* d = (const float*)h->ex[0].p;
        unsigned char* o = (unsigned char*)h->ex[7].p;
        BMPC_CUDA(launch_extract(1, n, nullptr, d, d + n, d + 2 * (size_t)n, d + 3 * (size_t)n, d + 4 * (size_t)n, 0, 1, d + 5 * (size_t)n,
                                 (int32_t*)o, (uint8_t*)(o + 4), (float*)(o + 8), (float*)(o + 12), (float*)(o + 16), (float*)(o + 20),
                                 (float*)(o + 24), h->stream));
        h->launches++;
        unsigned char* back = (unsigned char*)h->pinned + h->pinned_bytes / 2;
        BMPC_CUDA(cudaMemcpyAsync(back, o, 32, cudaMemcpyDeviceToHost, h->stream));
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
        std::memcpy(idx, back, 4); *flag = back[4];
        std::memcpy(out_px, back + 8, 4); std::memcpy(out_py, back + 12, 4); std::memcpy(out_dx, back + 16, 4);
        std::memcpy(out_dy, back + 20, 4); std::memcpy(out_vel, back + 24, 4);
        return BMPC_OK;
    }
    const size_t nflt = P > 0 ? (size_t)((count - 1) * stride_instance + (int64_t)(P - 1) * stride_point + 1) : 0;
    const float* src[5] = {px, py, dx, dy, vel};
    for (int a = 0; a < 5; ++a) {
        BMPC_CUDA(h->ex[a].reserve(nflt * 4 + 4));
        if (nflt) BMPC_CUDA(cudaMemcpyAsync(h->ex[a].p, src[a], nflt * 4, cudaMemcpyHostToDevice, h->stream));
    }
    BMPC_CUDA(h->ex[5].reserve((size_t)count * 4));
    BMPC_CUDA(cudaMemcpyAsync(h->ex[5].p, dist, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    if (npoints) {
        BMPC_CUDA(h->ex[6].reserve((size_t)count * 4));
        BMPC_CUDA(cudaMemcpyAsync(h->ex[6].p, npoints, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    }
    for (int a = 7; a < 14; ++a) BMPC_CUDA(h->ex[a].reserve((size_t)count * 4));
    BMPC_CUDA(launch_extract(count, P, npoints ? (const int32_t*)h->ex[6].p : nullptr, (const float*)h->ex[0].p,
                             (const float*)h->ex[1].p, (const float*)h->ex[2].p, (const float*)h->ex[3].p,
                             (const float*)h->ex[4].p, stride_instance, stride_point, (const float*)h->ex[5].p,
                             (int32_t*)h->ex[7].p, (uint8_t*)h->ex[8].p, (float*)h->ex[9].p, (float*)h->ex[10].p,
                             (float*)h->ex[11].p, (float*)h->ex[12].p, (float*)h->ex[13].p, h->stream));
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(idx, h->ex[7].p, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaMemcpyAsync(flag, h->ex[8].p, (size_t)count, cudaMemcpyDeviceToHost, h->stream));
    float* outs[5] = {out_px, out_py, out_dx, out_dy, out_vel};
    for (int a = 0; a < 5; ++a)
        BMPC_CUDA(cudaMemcpyAsync(outs[a], h->ex[9 + a].p, (size_t)count * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_extract_horizon(bmpc_t* h, int64_t count, int P, const int32_t* npoints, const float* px, const float* py,
                         const float* dx, const float* dy, const float* vel, int64_t stride_instance,
                         int64_t stride_point, const float* d0, const float* step, int steps, int32_t* idx,
                         uint8_t* flag, float* out_px, float* out_py, float* out_dx, float* out_dy, float* out_vel)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_extract_horizon: null handle");
    if (count < 1 || P < 0 || steps < 1 || !d0 || !step || !idx || !flag || !out_px || !out_py || !out_dx || !out_dy || !out_vel)
        return fail(BMPC_E_INVALID, "bmpc_extract_horizon: bad argument");
    if (P > 0 && (!px || !py || !dx || !dy || !vel)) return fail(BMPC_E_INVALID, "bmpc_extract_horizon: null trajectory array");
    if (stride_instance < 0 || stride_point < 0) return fail(BMPC_E_INVALID, "bmpc_extract_horizon: negative stride");
    for (int64_t b = 0; b < count; ++b)
        if (!(step[b] >= 0.0f)) return fail(BMPC_E_INVALID, "bmpc_extract_horizon: step must be >= 0 (ascending distances)");
    BMPC_CUDA(cudaSetDevice(h->device));
    const size_t nflt = P > 0 ? (size_t)((count - 1) * stride_instance + (int64_t)(P - 1) * stride_point + 1) : 0;
    const float* src[5] = {px, py, dx, dy, vel};
    for (int a = 0; a < 5; ++a) {
        BMPC_CUDA(h->ex[a].reserve(nflt * 4 + 4));
        if (nflt) BMPC_CUDA(cudaMemcpyAsync(h->ex[a].p, src[a], nflt * 4, cudaMemcpyHostToDevice, h->stream));
    }
    BMPC_CUDA(h->ex[5].reserve((size_t)count * 8));                       // d0 | step
    BMPC_CUDA(cudaMemcpyAsync(h->ex[5].p, d0, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(cudaMemcpyAsync((float*)h->ex[5].p + count, step, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    if (npoints) {
        BMPC_CUDA(h->ex[6].reserve((size_t)count * 4));
        BMPC_CUDA(cudaMemcpyAsync(h->ex[6].p, npoints, (size_t)count * 4, cudaMemcpyHostToDevice, h->stream));
    }
    const size_t n = (size_t)count * steps;
    for (int a = 7; a < 14; ++a) BMPC_CUDA(h->ex[a].reserve(n * 4));
    BMPC_CUDA(launch_extract_horizon(count, P, npoints ? (const int32_t*)h->ex[6].p : nullptr, (const float*)h->ex[0].p,
                                     (const float*)h->ex[1].p, (const float*)h->ex[2].p, (const float*)h->ex[3].p,
                                     (const float*)h->ex[4].p, stride_instance, stride_point, (const float*)h->ex[5].p,
                                     (const float*)h->ex[5].p + count, steps, (int32_t*)h->ex[7].p, (uint8_t*)h->ex[8].p,
                                     (float*)h->ex[9].p, (float*)h->ex[10].p, (float*)h->ex[11].p, (float*)h->ex[12].p,
                                     (float*)h->ex[13].p, h->stream));
    h->launches++;
    BMPC_CUDA(cudaMemcpyAsync(idx, h->ex[7].p, n * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaMemcpyAsync(flag, h->ex[8].p, n, cudaMemcpyDeviceToHost, h->stream));
    float* outs[5] = {out_px, out_py, out_dx, out_dy, out_vel};
    for (int a = 0; a < 5; ++a)
        BMPC_CUDA(cudaMemcpyAsync(outs[a], h->ex[9 + a].p, n * 4, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

}  // extern "C"
