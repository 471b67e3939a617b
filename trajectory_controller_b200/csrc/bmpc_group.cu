// libbmpc.so — C ABI part 5: device groups (include/bmpc.h "device groups").  One controller population sharded by
// instance over several GPUs of ONE process: a bmpc_t and a host worker thread per device, no device-to-device traffic,
// the results gathered into the caller's host arrays by every member's own D2H copy.
// Host code only: every byte of device work goes through the single-handle entry points of this library.
#include <condition_variable>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <thread>

#include "bmpc_internal.h"

using namespace bmpc;

namespace {

// A host thread that runs one job at a time for its member (CUDA's current device is per thread; the members' calls
// overlap because each blocks only its own thread).
struct Worker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::function<int()> job;
    bool pending = false, finished = false, quit = false;
    int rc = 0;
    std::string err;

    void loop()
    {
        std::unique_lock<std::mutex> lk(m);
        for (;;) {
            cv.wait(lk, [&] { return pending || quit; });
            if (quit) return;
            std::function<int()> j = std::move(job);
            pending = false;
            lk.unlock();
            const int r = j();
            std::string e = r ? std::string(bmpc_last_error()) : std::string();
            lk.lock();
            rc = r; err = std::move(e); finished = true;
            cv.notify_all();
        }
    }
    void post(std::function<int()> j)
    {
        std::lock_guard<std::mutex> lk(m);
        job = std::move(j); pending = true; finished = false;
        cv.notify_all();
    }
    int wait()
    {
        std::unique_lock<std::mutex> lk(m);
        cv.wait(lk, [&] { return finished; });
        return rc;
    }
};

// Page-locked host staging of one member (interleaved assignment: its packed share of the caller's arrays).
struct Pinned {
    void* p = nullptr;
    size_t bytes = 0;
    int reserve(size_t n)
    {
        if (n <= bytes) return BMPC_OK;
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
        BMPC_CUDA(cudaHostAlloc(&p, n, cudaHostAllocDefault));
        bytes = n;
        return BMPC_OK;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};

}  // namespace

struct bmpc_group {
    int S = 0, I = 0, N = 0, NI = 0, ndev = 0, assignment = 0;
    long long cap = 0, per = 0;                 // per: block length of the contiguous assignment
    std::vector<int> dev;
    std::vector<bmpc_t*> h;
    std::vector<std::unique_ptr<Worker>> w;
    std::vector<Pinned> pin_in, pin_out;        // per member

    // global instances [0, count) held by member k: local slots 0 .. n-1
    long long local_count(int k, long long count) const
    {
        if (assignment == BMPC_GROUP_INTERLEAVED) return count > k ? (count - k + ndev - 1) / ndev : 0;
        const long long first = (long long)k * per;
        long long room = cap - first; if (room > per) room = per;
        long long n = count - first;
        if (n > room) n = room;
        return n > 0 ? n : 0;
    }
    long long first_global(int k) const { return assignment == BMPC_GROUP_INTERLEAVED ? k : (long long)k * per; }
    // global index of member k's local slot j
    long long global_of(int k, long long j) const { return assignment == BMPC_GROUP_INTERLEAVED ? j * ndev + k : (long long)k * per + j; }

    // Runs f(k) on every member's worker concurrently; first failure wins.
    int run(const std::function<int(int)>& f)
    {
        for (int k = 0; k < ndev; ++k) w[k]->post([&f, k] { return f(k); });
        int rc = BMPC_OK; std::string msg;
        for (int k = 0; k < ndev; ++k) {
            const int r = w[k]->wait();
            if (r && !rc) { rc = r; msg = "member " + std::to_string(k) + " (device " + std::to_string(dev[k]) + "): " + w[k]->err; }
        }
        return rc ? fail(rc, msg) : BMPC_OK;
    }
};

namespace {

// member k's share of an instance-major host array src[count][len] -> dst[n][len] (interleaved assignment)
void pack(const bmpc_group* g, int k, long long n, int len, const double* src, double* dst)
{
    for (long long j = 0; j < n; ++j) std::memcpy(dst + j * len, src + g->global_of(k, j) * len, (size_t)len * 8);
}

}  // namespace

extern "C" {

int bmpc_group_create(bmpc_group_t** out, int S, int I, int N, int64_t capacity, int precision, int mode,
                      const int* devices, int ndev, int assignment)
{
    if (!out) return fail(BMPC_E_INVALID, "bmpc_group_create: out is null");
    *out = nullptr;
    if (assignment != BMPC_GROUP_INTERLEAVED && assignment != BMPC_GROUP_CONTIGUOUS)
        return fail(BMPC_E_INVALID, "bmpc_group_create: unknown assignment");
    int visible = 0;
    if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
        cudaGetLastError();
        return fail(BMPC_E_CUDA, "bmpc_group_create: no CUDA device available (this engine has no CPU fallback)");
    }
    if (ndev <= 0) { if (devices) return fail(BMPC_E_INVALID, "bmpc_group_create: ndev must be >= 1"); ndev = visible; }
    if (ndev > 64) return fail(BMPC_E_INVALID, "bmpc_group_create: at most 64 members");
    if (capacity < ndev) return fail(BMPC_E_INVALID, "bmpc_group_create: capacity must be >= the number of members");
    for (int k = 0; k < ndev; ++k) {
        const int d = devices ? devices[k] : k;
        if (d < 0 || d >= visible) return fail(BMPC_E_INVALID, "bmpc_group_create: device index out of range");
    }
    bmpc_group* g = new (std::nothrow) bmpc_group();
    if (!g) return fail(BMPC_E_NOMEM, "bmpc_group_create: out of host memory");
    g->S = S; g->I = I; g->N = N; g->NI = N * I; g->ndev = ndev; g->assignment = assignment; g->cap = capacity;
    g->per = (capacity + ndev - 1) / ndev;
    g->dev.resize(ndev); g->h.assign(ndev, nullptr); g->pin_in.resize(ndev); g->pin_out.resize(ndev);
    for (int k = 0; k < ndev; ++k) {
        g->dev[k] = devices ? devices[k] : k;
        g->w.emplace_back(new Worker());
        Worker* wk = g->w.back().get();
        wk->th = std::thread([wk] { wk->loop(); });
    }
    // members are created on their own threads: the caller's current device is left alone
    const int rc = g->run([&](int k) {
        long long mcap = g->local_count(k, capacity);
        if (mcap < 1) mcap = 1;
        return bmpc_create(&g->h[k], S, I, N, mcap, precision, mode, g->dev[k]);
    });
    if (rc) { std::string keep = bmpc_last_error(); bmpc_group_destroy(g); set_error(keep); return rc; }
    *out = g;
    return BMPC_OK;
}

void bmpc_group_destroy(bmpc_group_t* g)
{
    if (!g) return;
    for (int k = 0; k < (int)g->w.size(); ++k) {
        Worker* wk = g->w[k].get();
        wk->post([g, k] {
            if (g->h[k]) bmpc_destroy(g->h[k]);
            g->h[k] = nullptr;
            cudaSetDevice(g->dev[k]);
            g->pin_in[k].release(); g->pin_out[k].release();
            return 0;
        });
        wk->wait();
        { std::lock_guard<std::mutex> lk(wk->m); wk->quit = true; wk->cv.notify_all(); }
        wk->th.join();
    }
    delete g;
}

int bmpc_group_size(bmpc_group_t* g) { return g ? g->ndev : 0; }
bmpc_t* bmpc_group_member(bmpc_group_t* g, int k) { return (g && k >= 0 && k < g->ndev) ? g->h[k] : nullptr; }

int bmpc_group_shard(bmpc_group_t* g, int k, int64_t count, int64_t* first, int64_t* local_count)
{
    if (!g || k < 0 || k >= g->ndev) return fail(BMPC_E_INVALID, "bmpc_group_shard: bad group or member index");
    if (count < 0 || count > g->cap) return fail(BMPC_E_INVALID, "bmpc_group_shard: count out of capacity");
    if (first) *first = g->first_global(k);
    if (local_count) *local_count = g->local_count(k, count);
    return BMPC_OK;
}

int bmpc_group_set_model(bmpc_group_t* g, const double* A, const double* B, const double* C, const double* Q,
                         const double* R, const double* lower, const double* upper, int per_instance, int64_t count)
{
    if (!g) return fail(BMPC_E_INVALID, "bmpc_group_set_model: null group");
    if (!A || !B || !C || !Q || !R || !lower || !upper) return fail(BMPC_E_INVALID, "bmpc_group_set_model: null matrix pointer");
    if (!per_instance) return g->run([&](int k) { return bmpc_set_model(g->h[k], A, B, C, Q, R, lower, upper, 0, 1); });
    if (count < 1 || count > g->cap) return fail(BMPC_E_INVALID, "bmpc_group_set_model: count out of range");
    const int S = g->S, I = g->I;
    const int len[7] = {S * S, S * I, S, S, I, I, I};
    const double* src[7] = {A, B, C, Q, R, lower, upper};
    return g->run([&](int k) -> int {
        const long long n = g->local_count(k, count);
        if (n == 0) return BMPC_OK;                              // this member holds none of the models
        if (g->assignment == BMPC_GROUP_CONTIGUOUS) {
            const long long f = g->first_global(k);
            return bmpc_set_model(g->h[k], A + f * len[0], B + f * len[1], C + f * len[2], Q + f * len[3], R + f * len[4],
                                  lower + f * len[5], upper + f * len[6], 1, n);
        }
        cudaSetDevice(g->dev[k]);
        int tot = 0; for (int a = 0; a < 7; ++a) tot += len[a];
        int rc = g->pin_in[k].reserve((size_t)n * tot * 8);
        if (rc) return rc;
        double* p[7]; double* at = (double*)g->pin_in[k].p;
        for (int a = 0; a < 7; ++a) { p[a] = at; at += n * len[a]; pack(g, k, n, len[a], src[a], p[a]); }
        return bmpc_set_model(g->h[k], p[0], p[1], p[2], p[3], p[4], p[5], p[6], 1, n);
    });
}

int bmpc_group_set_options(bmpc_group_t* g, double eps, uint64_t max_iterations)
{
    if (!g) return fail(BMPC_E_INVALID, "bmpc_group_set_options: null group");
    for (int k = 0; k < g->ndev; ++k) {
        const int rc = bmpc_set_options(g->h[k], eps, max_iterations);   // host-side state only
        if (rc) return rc;
    }
    return BMPC_OK;
}

int bmpc_group_set_target(bmpc_group_t* g, const double* val, int time, int per_instance, int64_t count)
{
    if (!g || !val) return fail(BMPC_E_INVALID, "bmpc_group_set_target: null argument");
    if (!per_instance) return g->run([&](int k) { return bmpc_set_target(g->h[k], val, time, 0, 1); });
    if (count < 1 || count > g->cap) return fail(BMPC_E_INVALID, "bmpc_group_set_target: count out of range");
    return g->run([&](int k) -> int {
        const long long n = g->local_count(k, count);
        if (n == 0) return BMPC_OK;
        if (g->assignment == BMPC_GROUP_CONTIGUOUS) return bmpc_set_target(g->h[k], val + g->first_global(k) * g->S, time, 1, n);
        cudaSetDevice(g->dev[k]);
        int rc = g->pin_in[k].reserve((size_t)n * g->S * 8);
        if (rc) return rc;
        pack(g, k, n, g->S, val, (double*)g->pin_in[k].p);
        return bmpc_set_target(g->h[k], (const double*)g->pin_in[k].p, time, 1, n);
    });
}

int bmpc_group_reset(bmpc_group_t* g)
{
    if (!g) return fail(BMPC_E_INVALID, "bmpc_group_reset: null group");
    return g->run([&](int k) { return bmpc_reset(g->h[k]); });
}

int bmpc_group_solve(bmpc_group_t* g, int64_t count, const double* x0, double* u0, uint32_t* iters, uint8_t* status,
                     unsigned flags)
{
    if (!g || !x0 || !u0) return fail(BMPC_E_INVALID, "bmpc_group_solve: null argument");
    if (count < 1 || count > g->cap) return fail(BMPC_E_INVALID, "bmpc_group_solve: count out of capacity");
    const int S = g->S, I = g->I;
    if (g->assignment == BMPC_GROUP_CONTIGUOUS)
        return g->run([&](int k) -> int {
            const long long n = g->local_count(k, count);
            if (n == 0) return BMPC_OK;
            const long long f = g->first_global(k);
            return bmpc_solve(g->h[k], n, x0 + f * S, u0 + f * I, iters ? iters + f : nullptr, status ? status + f : nullptr, flags);
        });
    const int rc1 = g->run([&](int k) -> int {
        const long long n = g->local_count(k, count);
        if (n == 0) return BMPC_OK;
        cudaSetDevice(g->dev[k]);
        int rc;
        if ((rc = g->pin_in[k].reserve((size_t)n * S * 8))) return rc;
        // u0 | iters | status of this member, gathered into the caller's arrays below
        const size_t ob = (size_t)n * I * 8, ib = (size_t)n * 4;
        if ((rc = g->pin_out[k].reserve(ob + ib + (size_t)n))) return rc;
        double* px = (double*)g->pin_in[k].p;
        double* pu = (double*)g->pin_out[k].p;
        uint32_t* pi = (uint32_t*)((char*)g->pin_out[k].p + ob);
        uint8_t* ps = (uint8_t*)g->pin_out[k].p + ob + ib;
        pack(g, k, n, S, x0, px);
        return bmpc_solve(g->h[k], n, px, pu, iters ? pi : nullptr, status ? ps : nullptr, flags);
    });
    if (rc1) return rc1;
    // The gather into the caller's arrays is split by OUTPUT range, not by member: worker k writes the contiguous block
    // [k*count/ndev, (k+1)*count/ndev) from every member's staging (neighbouring instances belong to different members,
    // and eight threads writing into the same cache lines cost 4 ms per million instances).
    const int nd = g->ndev;
    std::vector<long long> held(nd);
    for (int m = 0; m < nd; ++m) held[m] = g->local_count(m, count);
    return g->run([&](int k) -> int {
        const long long lo = count * k / nd, hi = count * (k + 1) / nd;
        for (long long i = lo; i < hi; ++i) {
            const int m = (int)(i % nd);
            const long long j = i / nd, nm = held[m];
            const char* base = (const char*)g->pin_out[m].p;
            const double* pu = (const double*)base;
            for (int c = 0; c < I; ++c) u0[i * I + c] = pu[j * I + c];
            if (iters) iters[i] = ((const uint32_t*)(base + (size_t)nm * I * 8))[j];
            if (status) status[i] = ((const uint8_t*)(base + (size_t)nm * I * 8 + (size_t)nm * 4))[j];
        }
        return BMPC_OK;
    });
}

int bmpc_group_get_controls(bmpc_group_t* g, int64_t first, int64_t count, double* controls)
{
    if (!g || !controls) return fail(BMPC_E_INVALID, "bmpc_group_get_controls: null argument");
    if (first < 0 || count < 1 || first + count > g->cap) return fail(BMPC_E_INVALID, "bmpc_group_get_controls: range out of capacity");
    const int NI = g->NI;
    return g->run([&](int k) -> int {
        // local slots [lo, hi) of member k that fall into the global range
        const long long lo = g->local_count(k, first), hi = g->local_count(k, first + count);
        if (hi <= lo) return BMPC_OK;
        cudaSetDevice(g->dev[k]);
        int rc = g->pin_out[k].reserve((size_t)(hi - lo) * NI * 8);
        if (rc) return rc;
        double* p = (double*)g->pin_out[k].p;
        if ((rc = bmpc_get_controls(g->h[k], lo, hi - lo, p))) return rc;
        for (long long j = lo; j < hi; ++j)
            std::memcpy(controls + (g->global_of(k, j) - first) * NI, p + (j - lo) * NI, (size_t)NI * 8);
        return BMPC_OK;
    });
}

int bmpc_group_last_kernel_ms(bmpc_group_t* g, float* ms)
{
    if (!g || !ms) return fail(BMPC_E_INVALID, "bmpc_group_last_kernel_ms: null argument");
    std::vector<float> t(g->ndev, -1.f);
    g->run([&](int k) { float v = 0.f; if (bmpc_last_kernel_ms(g->h[k], &v) == BMPC_OK) t[k] = v; return 0; });   // members that never solved are skipped
    float best = -1.f;
    for (float v : t) if (v > best) best = v;
    if (best < 0.f) return fail(BMPC_E_STATE, "bmpc_group_last_kernel_ms: no member has solved yet");
    *ms = best;
    return BMPC_OK;
}

}  // extern "C"
