// libbmpc.so — C ABI part 1: errors, lifetime, model, options.  See include/bmpc.h for the contract.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "bmpc_internal.h"
#include "shapes.h"

namespace bmpc {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int fail(int code, const std::string& msg) { g_last_error = msg; return code; }
int fail_cuda(cudaError_t e, const char* what)
{
    g_last_error = std::string("CUDA error '") + cudaGetErrorString(e) + "' in " + what;
    cudaGetLastError();   // clear the sticky-less error state
    return e == cudaErrorMemoryAllocation ? BMPC_E_NOMEM : BMPC_E_CUDA;
}

cudaError_t DevBuf::reserve(size_t n)
{
    if (n <= bytes) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc(&p, n);
    if (e == cudaSuccess) bytes = n; else p = nullptr;
    return e;
}
cudaError_t DevBuf::alloc_zero(size_t n, cudaStream_t st)
{
    cudaError_t e = reserve(n);
    if (e != cudaSuccess) return e;
    return cudaMemsetAsync(p, 0, n, st);
}
void DevBuf::release()
{
    if (p && !view) cudaFree(p);
    p = nullptr; bytes = 0;
}

#define X(s, i, n, pt) extern const ShapeOps bmpc_ops_##s##_##i##_##n##_##pt;
BMPC_SHAPES(X)
#undef X

static const ShapeOps* const g_all_ops[] = {
#define X(s, i, n, pt) &bmpc_ops_##s##_##i##_##n##_##pt,
    BMPC_SHAPES(X)
#undef X
};

// The dense instantiation of a shape: accepts every model, owns the constructor kernel.
const ShapeOps* find_shape(int S, int I, int N)
{
    for (const ShapeOps* o : g_all_ops)
        if (o->S == S && o->I == I && o->N == N && o->ctor && std::strcmp(o->pattern, "dense") == 0) return o;
    return nullptr;
}

// A named pattern unit of the shape (callers guarantee that the model satisfies it), or null.
const ShapeOps* find_pattern(int S, int I, int N, const char* name)
{
    for (const ShapeOps* o : g_all_ops)
        if (o->S == S && o->I == I && o->N == N && std::strcmp(o->pattern, name) == 0) return o;
    return nullptr;
}

static int popcount_u(unsigned x) { int n = 0; while (x) { n += x & 1u; x >>= 1; } return n; }

// A model (count instances, instance-major doubles already rounded to the compute type by `rnd`) satisfies a pattern
// iff every non-zero entry is allowed by the NZ bits and every ONE-bit entry equals exactly 1.
static bool entries_fit(const double* v, long long count, int len, unsigned nz, unsigned one, bool as_float)
{
    for (long long b = 0; b < count; ++b)
        for (int e = 0; e < len; ++e) {
            double x = v[b * len + e];
            if (as_float) x = (double)(float)x;
            if (x != 0.0 && !((nz >> e) & 1u)) return false;
            if (((one >> e) & 1u) && x != 1.0) return false;
        }
    return true;
}

// The sparsest compiled pattern of the shape that the model satisfies (the dense one always does).
const ShapeOps* select_pattern(int S, int I, int N, const double* A, const double* B, const double* C, const double* Q,
                               long long count, bool as_float)
{
    const ShapeOps* best = nullptr;
    int best_cost = 1 << 30;
    for (const ShapeOps* o : g_all_ops) {
        if (o->S != S || o->I != I || o->N != N) continue;
        if (!entries_fit(A, count, S * S, o->anz, o->aone, as_float) || !entries_fit(B, count, S * I, o->bnz, o->bone, as_float) ||
            !entries_fit(C, count, S, o->cnz, 0u, as_float) || !entries_fit(Q, count, S, o->qnz, 0u, as_float))
            continue;
        // cost ~ multiplications + additions per horizon step that survive
        const int cost = 2 * (popcount_u(o->anz) * 2 - popcount_u(o->aone)) + 2 * (popcount_u(o->bnz) * 2 - popcount_u(o->bone)) +
                         popcount_u(o->cnz) + 2 * popcount_u(o->qnz);
        if (cost < best_cost) { best_cost = cost; best = o; }
    }
    return best;
}

// dlib::mpc's constructor contract — min(Q) >= 0, min(R) > 0, min(upper - lower) >= 0 (mpc.h:92-94) — and finite
// entries, checked for every instance (dlib leaves a violation undefined; the engine rejects it).  *sym_ok reports
// whether stage B's symmetric-bound kernel (pg_kernel.cuh, SYM) applies: lower == -upper with upper > 0, and a model
// tame enough that no horizon recursion can overflow (the kernel's integer-pipe compares assume non-NaN values).
static const char* validate_model(int S, int I, int N, const double* A, const double* B, const double* C, const double* Q,
                                  const double* R, const double* lo, const double* hi, long long count, bool as_float,
                                  bool* sym_ok)
{
    bool sym = true;
    double worst = 0;   // log10 of the growth bound, max over instances
    auto val = [&](const double* v, long long i) { return as_float ? (double)(float)v[i] : v[i]; };
    for (long long b = 0; b < count; ++b) {
        double arow = 0, acol[16] = {0}, bmax = 0, qmax = 0, rmax = 0, hmax = 0;
        for (int r = 0; r < S; ++r) {
            double row = 0;
            for (int k = 0; k < S; ++k) {
                const double x = val(A, b * S * S + r * S + k);
                if (!std::isfinite(x)) return "A has a non-finite entry";
                row += std::fabs(x); if (k < 16) acol[k] += std::fabs(x);
            }
            arow = std::max(arow, row);
        }
        double anorm = std::max(1.0, arow);
        for (int k = 0; k < S && k < 16; ++k) anorm = std::max(anorm, acol[k]);
        for (int e = 0; e < S * I; ++e) {
            const double x = val(B, b * S * I + e);
            if (!std::isfinite(x)) return "B has a non-finite entry";
            bmax = std::max(bmax, std::fabs(x));
        }
        for (int e = 0; e < S; ++e) {
            const double c = val(C, b * S + e), q = val(Q, b * S + e);
            if (!std::isfinite(c)) return "C has a non-finite entry";
            if (!(q >= 0) || !std::isfinite(q)) return "Q must be finite and >= 0 (mpc.h:92 requires min(Q) >= 0)";
            qmax = std::max(qmax, q);
        }
        for (int e = 0; e < I; ++e) {
            const double r = val(R, b * I + e), l = val(lo, b * I + e), u = val(hi, b * I + e);
            if (!(r > 0) || !std::isfinite(r)) return "R must be finite and > 0 (mpc.h:93 requires min(R) > 0)";
            if (!std::isfinite(l) || !std::isfinite(u)) return "control bounds must be finite";
            if (!(u >= l)) return "upper must be >= lower (mpc.h:94 requires min(upper-lower) >= 0)";
            if (!(u > 0 && l == -u)) sym = false;
            rmax = std::max(rmax, r); hmax = std::max(hmax, std::max(std::fabs(l), std::fabs(u)));
        }
        // |M| <= N g^N |B| I h, |Q M| and the adjoint chain add another N g^N |Q|, the gradient |B| S on top (mpc.h:275-283)
        const double lg = 2.0 * N * std::log10(anorm) + 2.0 * std::log10((double)N) + 2.0 * std::log10(bmax * I + 1.0) +
                          std::log10(qmax + 1.0) + std::log10(rmax + 1.0) + std::log10(hmax + 1.0) + std::log10((double)S);
        worst = std::max(worst, lg);
    }
    *sym_ok = sym && worst < (as_float ? 30.0 : 250.0);
    return nullptr;
}

}  // namespace bmpc

using namespace bmpc;

extern "C" {

const char* bmpc_last_error(void) { return g_last_error.c_str(); }
#ifndef BMPC_BUILD_ID
#define BMPC_BUILD_ID "unknown"
#endif
const char* bmpc_version(void) { return "bmpc 0.2 sm_100a build " BMPC_BUILD_ID " (batched dlib::mpc-compatible linear MPC, CUDA)"; }
int bmpc_shape_supported(int S, int I, int N) { return find_shape(S, I, N) != nullptr; }

int bmpc_create(bmpc_t** out, int S, int I, int N, int64_t capacity, int precision, int mode, int device)
{
    if (!out) return fail(BMPC_E_INVALID, "bmpc_create: out is null");
    *out = nullptr;
    if (capacity < 1) return fail(BMPC_E_INVALID, "bmpc_create: capacity must be >= 1");
    if (precision != BMPC_FP64 && precision != BMPC_FP32) return fail(BMPC_E_UNSUPPORTED, "bmpc_create: precision must be 64 or 32");
    if (mode != BMPC_MODE_EXACT && mode != BMPC_MODE_FAST) return fail(BMPC_E_UNSUPPORTED, "bmpc_create: unknown mode");
    const ShapeOps* ops = find_shape(S, I, N);
    if (!ops) {
        char b[160];
        snprintf(b, sizeof b, "bmpc_create: shape (S=%d,I=%d,N=%d) is not instantiated in this build (see csrc/shapes.h)", S, I, N);
        return fail(BMPC_E_UNSUPPORTED, b);
    }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(BMPC_E_CUDA, "bmpc_create: no CUDA device available (this engine has no CPU fallback)");
    }
    if (device < 0) BMPC_CUDA(cudaGetDevice(&device));
    if (device >= ndev) return fail(BMPC_E_INVALID, "bmpc_create: device index out of range");
    BMPC_CUDA(cudaSetDevice(device));

    bmpc_handle* h = new (std::nothrow) bmpc_handle();
    if (!h) return fail(BMPC_E_NOMEM, "bmpc_create: out of host memory");
    h->S = S; h->I = I; h->N = N; h->NI = N * I;
    h->precision = precision; h->mode = mode; h->device = device;
    h->elem = precision == BMPC_FP64 ? 8 : 4;
    h->cap = capacity; h->ops = ops; h->solve_ops = ops;
    int rc = BMPC_OK;
    auto chk = [&](cudaError_t ce, const char* what) { if (rc == BMPC_OK && ce != cudaSuccess) rc = fail_cuda(ce, what); };
    chk(cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device), "cudaDeviceGetAttribute");
    chk(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking), "cudaStreamCreate");
    h->stream = h->own_stream;
    chk(cudaEventCreate(&h->ev0), "cudaEventCreate");
    chk(cudaEventCreate(&h->ev1), "cudaEventCreate");
    chk(cudaEventCreate(&h->ev_mid), "cudaEventCreate");
    const size_t c = (size_t)capacity, es = (size_t)h->elem;
    chk(h->x0.alloc_zero(c * S * es, h->stream), "alloc x0");
    {   // the results of a solve live in one allocation, so that a multi-GPU host gathers them with ONE collective
        // and one copy (bmpc_device_results_block)
        auto up = [](size_t n) { return (n + 255) & ~(size_t)255; };
        const size_t o_it = up(c * I * es), o_st = o_it + up(c * 4), total = o_st + up(c);
        chk(h->results.alloc_zero(total, h->stream), "alloc results");
        if (h->results.p) {
            h->u0.p = h->results.p; h->u0.bytes = c * I * es; h->u0.view = true;
            h->iters.p = (char*)h->results.p + o_it; h->iters.bytes = c * 4; h->iters.view = true;
            h->status.p = (char*)h->results.p + o_st; h->status.bytes = c; h->status.view = true;
        }
    }
    chk(h->controls.alloc_zero(c * h->NI * es, h->stream), "alloc controls");
    chk(h->vstate.alloc_zero(c * h->NI * es, h->stream), "alloc v");
    chk(h->target_const.alloc_zero(c * S * es, h->stream), "alloc target");
    chk(h->queue.alloc_zero(64, h->stream), "alloc queue");
    // Hand-over ring: every instance can be parked at once, plus one open ticket per lane in flight (a launch of
    // `count` instances keeps fewer than count + 32 * (grid + 1) lanes busy, bmpc_solve_resident_async).
    {
        const size_t lanes_max = (size_t)h->sm_count * 8 * 32, lanes_small = 2 * (c + 32);
        const size_t want = c + (lanes_small < lanes_max ? lanes_small : lanes_max);
        h->pg_ring = 64; while (h->pg_ring < want) h->pg_ring <<= 1;
    }
    chk(h->pg_list.alloc_zero(h->pg_ring * 4, h->stream), "alloc pg_list");
    chk(h->park.reserve(h->pg_ring * 3 * h->NI * es), "alloc park");
    chk(h->park_iter.reserve(h->pg_ring * 4), "alloc park_iter");
    // Stage C (pair_kernel.cuh): 16 instances per warp; one warp per SM sub-partition is the fast configuration, so stage B
    // hands over when that many instances are left and batches up to that size go to stage C directly.
    // BMPC_PAIR_THRESHOLD / BMPC_PAIR_DIRECT: developer overrides for A/B measurements.
    h->pair_threshold = (long long)h->sm_count * 4 * 16;
    h->pair_direct_max = h->pair_threshold;
    if (const char* e = std::getenv("BMPC_PAIR_THRESHOLD")) h->pair_threshold = std::atoll(e);
    if (const char* e = std::getenv("BMPC_PAIR_DIRECT")) h->pair_direct_max = std::atoll(e);
    if (const char* e = std::getenv("BMPC_SMALL")) h->small_cfg = std::atoi(e);
    if (h->pair_threshold < 0) h->pair_threshold = 0;
    if (ops->S == 4 && ops->I == 1 && h->pair_threshold > 0) {
        const size_t ct = (size_t)h->pair_threshold;
        chk(h->c_list.reserve(ct * 4), "alloc c_list");
        chk(h->c_park.reserve(ct * 3 * h->NI * es), "alloc c_park");
        chk(h->c_iter.reserve(ct * 4), "alloc c_iter");
    }
    chk(h->small.alloc_zero(8 * 4096, h->stream), "alloc small");
    h->pinned_bytes = 8 * 4096;
    chk(cudaHostAlloc(&h->pinned, h->pinned_bytes, cudaHostAllocDefault), "cudaHostAlloc");
    chk(cudaStreamSynchronize(h->stream), "cudaStreamSynchronize");
    if (rc != BMPC_OK) { std::string keep = g_last_error; bmpc_destroy(h); g_last_error = keep; return rc; }
    *out = h;
    return BMPC_OK;
}

void bmpc_destroy(bmpc_t* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    DevBuf* bufs[] = {&h->mA, &h->mB, &h->mC, &h->mQ, &h->mR, &h->mlo, &h->mhi, &h->mlambda, &h->mqdiag, &h->x0, &h->results, &h->u0,
                      &h->iters, &h->status, &h->controls, &h->vstate, &h->target_const, &h->target_full, &h->queue,
                      &h->pg_list, &h->park, &h->park_iter, &h->c_list, &h->c_park, &h->c_iter, &h->stage, &h->small, &h->cl_path, &h->cl_pose, &h->cl_tgt, &h->cl_idx, &h->cl_itot, &h->cl_imax};
    for (DevBuf* b : bufs) b->release();
    for (DevBuf& b : h->ex) b.release();
    for (DevBuf& b : h->lut) b.release();
    for (DevBuf& b : h->pid_state) b.release();
    if (h->pinned) cudaFreeHost(h->pinned);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->ev_mid) cudaEventDestroy(h->ev_mid);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
    for (cudaEvent_t e : h->ev_chunk) if (e) cudaEventDestroy(e);
    delete h;
}

int bmpc_set_stream(bmpc_t* h, void* cuda_stream)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_set_stream: null handle");
    BMPC_CUDA(cudaSetDevice(h->device));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));      // nothing of ours may still be in flight on the old stream
    h->stream = cuda_stream ? (cudaStream_t)cuda_stream : h->own_stream;
    h->timed = false;
    return BMPC_OK;
}

int bmpc_set_specialization(bmpc_t* h, int enable)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_set_specialization: null handle");
    h->specialize = enable != 0;
    return BMPC_OK;
}
int bmpc_set_split(bmpc_t* h, int enable)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_set_split: null handle");
    h->split = enable != 0;
    if (enable >= 8) {   // developer A/B: 8 | 1 (fairness slices) | 2 (retirement of surplus warps) | 4 (symmetric-bound kernel) | 16 (stage C)
        h->slice = (enable & 1) != 0; h->retire = (enable & 2) != 0; h->sym = (enable & 4) != 0; h->pair = (enable & 16) != 0;
    } else {
        h->slice = enable != 2;
        h->retire = enable != 2;
        h->sym = enable != 3;
        h->pair = enable == 1;
    }
    return BMPC_OK;
}
const char* bmpc_pattern_name(bmpc_t* h) { return h && h->solve_ops ? h->solve_ops->pattern : ""; }

int bmpc_set_options(bmpc_t* h, double eps, uint64_t max_iterations)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_set_options: null handle");
    if (!(eps > 0)) return fail(BMPC_E_INVALID, "bmpc_set_options: eps must be > 0 (mpc.h:203 requires)");
    h->eps = eps;
    h->max_iter = max_iterations > 0xffffffffull ? 0xffffffffu : (uint32_t)max_iterations;
    return BMPC_OK;
}
int bmpc_get_options(bmpc_t* h, double* eps, uint64_t* max_iterations)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_get_options: null handle");
    if (eps) *eps = h->eps;
    if (max_iterations) *max_iterations = h->max_iter;
    return BMPC_OK;
}

// Upload `count` instance-major double vectors of length len and store them as [len][cap] of T.
static int upload_soa(bmpc_t* h, const double* host, long long count, int len, DevBuf& dst)
{
    const size_t bytes = (size_t)count * len * 8;
    BMPC_CUDA(h->stage.reserve(bytes));
    BMPC_CUDA(dst.reserve((size_t)h->cap * len * h->elem));
    BMPC_CUDA(cudaMemcpyAsync(h->stage.p, host, bytes, cudaMemcpyHostToDevice, h->stream));
    BMPC_CUDA(aos_to_soa((const double*)h->stage.p, dst.p, count, len, h->cap, h->elem, h->stream));
    h->launches++;
    // the staging buffer is reused by the next upload: wait (uploads are not on the solve hot path)
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    return BMPC_OK;
}

int bmpc_set_model(bmpc_t* h, const double* A, const double* B, const double* C, const double* Q, const double* R,
                   const double* lower, const double* upper, int per_instance, int64_t count)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_set_model: null handle");
    if (!A || !B || !C || !Q || !R || !lower || !upper) return fail(BMPC_E_INVALID, "bmpc_set_model: null matrix pointer");
    const int S = h->S, I = h->I, N = h->N;
    BMPC_CUDA(cudaSetDevice(h->device));
    const long long cnt = per_instance ? count : 1;
    if (cnt < 1 || cnt > h->cap) return fail(BMPC_E_INVALID, "bmpc_set_model: count out of range");
    {   // dlib's requires clause (mpc.h:92-94, mpc_abstract.h:90-97) plus finiteness, for every instance
        const char* bad = validate_model(S, I, N, A, B, C, Q, R, lower, upper, cnt, h->elem == 4, &h->sym_ok);
        if (bad) return fail(BMPC_E_INVALID, std::string("bmpc_set_model: ") + bad);
    }
    int rc;
    const int lens[7] = {S * S, S * I, S, S, I, I, I};
    const double* srcs[7] = {A, B, C, Q, R, lower, upper};
    DevBuf* bufs[7] = {&h->mA, &h->mB, &h->mC, &h->mQ, &h->mR, &h->mlo, &h->mhi};
    int total = 0; for (int a = 0; a < 7; ++a) total += lens[a];
    if (cnt == 1 && total <= kSmallVals) {
        // one model (the facade's case, a fresh dlib::mpc per cycle() of the reference module): the numbers travel as
        // kernel arguments — no staging copy, no synchronisation before the constructor kernel
        SmallVals v; PutPlan plan; plan.n = 7;
        int at = 0;
        for (int a = 0; a < 7; ++a) {
            BMPC_CUDA(bufs[a]->reserve((size_t)h->cap * lens[a] * h->elem));
            plan.off[a] = at; plan.len[a] = lens[a]; plan.dst[a] = bufs[a]->p;
            for (int e = 0; e < lens[a]; ++e) v.x[at + e] = srcs[a][e];
            at += lens[a];
        }
        plan.off[7] = 0; plan.len[7] = 0; plan.dst[7] = nullptr;
        BMPC_CUDA(put_vals(v, plan, 0, 1, h->cap, h->elem, h->stream));
        h->launches++;
    } else if ((rc = upload_soa(h, A, cnt, S * S, h->mA)) || (rc = upload_soa(h, B, cnt, S * I, h->mB)) ||
               (rc = upload_soa(h, C, cnt, S, h->mC)) || (rc = upload_soa(h, Q, cnt, S, h->mQ)) ||
               (rc = upload_soa(h, R, cnt, I, h->mR)) || (rc = upload_soa(h, lower, cnt, I, h->mlo)) ||
               (rc = upload_soa(h, upper, cnt, I, h->mhi)))
        return rc;
    BMPC_CUDA(h->mlambda.reserve((size_t)h->cap * h->elem));
    BMPC_CUDA(h->mqdiag.reserve((size_t)h->cap * h->NI * h->elem));
    GenericCtor g{cnt, h->cap, h->mA.p, h->mB.p, h->mQ.p, h->mR.p, h->mlambda.p, h->mqdiag.p};
    BMPC_CUDA(h->ops->ctor(h->precision, h->mode, g, h->stream));   // mpc.h:114-123 on the device
    h->launches++;
    h->per_instance = per_instance != 0;
    h->model_count = cnt;
    h->solve_ops = h->specialize ? select_pattern(S, I, N, A, B, C, Q, cnt, h->elem == 4) : h->ops;
    if (!h->solve_ops) return fail(BMPC_E_UNSUPPORTED, "bmpc_set_model: no compiled pattern accepts this model (dense unit missing)");
    if (!per_instance) {
        // Keep the shared model on the host, rounded to the compute type, for the kernel-parameter block; read the
        // device-computed lambda / Q_diag back (a few hundred bytes).
        auto rnd = [&](const double* src, int n, std::vector<double>& dst) {
            dst.resize(n);
            for (int k = 0; k < n; ++k) dst[k] = h->elem == 8 ? src[k] : (double)(float)src[k];
        };
        rnd(A, S * S, h->hA); rnd(B, S * I, h->hB); rnd(C, S, h->hC); rnd(Q, S, h->hQ);
        rnd(R, I, h->hR); rnd(lower, I, h->hlo); rnd(upper, I, h->hhi);
        h->hqdiag.resize(h->NI);
        int r2 = bmpc_get_derived(h, 0, &h->hlambda, h->hqdiag.data());
        if (r2) return r2;
    } else {
        BMPC_CUDA(cudaStreamSynchronize(h->stream));
    }
    h->have_model = true;
    (void)N;
    return BMPC_OK;
}

int bmpc_get_derived(bmpc_t* h, int64_t instance, double* lambda, double* q_diag)
{
    if (!h) return fail(BMPC_E_INVALID, "bmpc_get_derived: null handle");
    if (!h->mlambda.p) return fail(BMPC_E_STATE, "bmpc_get_derived: no model set");
    if (instance < 0 || instance >= h->cap) return fail(BMPC_E_INVALID, "bmpc_get_derived: instance out of range");
    BMPC_CUDA(cudaSetDevice(h->device));
    const int es = h->elem;
    unsigned char* pin = (unsigned char*)h->pinned;
    BMPC_CUDA(cudaMemcpyAsync(pin, (char*)h->mlambda.p + instance * es, es, cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaMemcpy2DAsync(pin + 8, es, (char*)h->mqdiag.p + instance * es, (size_t)h->cap * es, es, h->NI,
                                cudaMemcpyDeviceToHost, h->stream));
    BMPC_CUDA(cudaStreamSynchronize(h->stream));
    if (lambda) *lambda = es == 8 ? ((double*)pin)[0] : (double)((float*)pin)[0];
    if (q_diag)
        for (int k = 0; k < h->NI; ++k) q_diag[k] = es == 8 ? ((double*)(pin + 8))[k] : (double)((float*)(pin + 8))[k];
    return BMPC_OK;
}

}  // extern "C"
