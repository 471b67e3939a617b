// TEST INFRASTRUCTURE — NOT PRODUCT CODE.
//
// Thin C-ABI wrapper over the REAL reference solver, dlib::mpc, compiled from the
// sources where they lie: /root/reference/dlib_files/dlib/control/mpc.h (header-only).
// No reference source is copied into this repository; this file only #includes it.
// Output goes to oracle/_ref/libdlibref.so (git-ignored, travels to the GPU box).
//
// Used by: tests/ (to pin oracle/mpc_oracle.c bit-for-bit), tests/golden/make_golden.py
// (fixture generation) and bench.py's cpu_baseline / --impl reference legs (one
// dlib::mpc per host thread).  Nothing under trajectory_controller_b200/ may link it.
//
// Canonical flags: g++ -O2 -std=c++11 -ffp-contract=off (no -march=native: FMA
// contraction changes low bits).
//
// Determinism note: dlib::mpc leaves v[]/v_old[]/M[]/MM[]/df[] uninitialised
// (matrix_data_layout.h:137) and reads a stale v[] when the `continue` at mpc.h:322
// fires in the last SMO iteration.  Every controller here is placement-constructed
// on zeroed storage so that this corner is deterministic (v starts at zero), which is
// the single de-undefinition the whole project adopts (DESIGN.md §"Undefined corner").

#include <dlib/matrix.h>
#include <dlib/control/lspi.h>   // pulled in by dlib/control.h before mpc.h; keeps include order identical
#define private public           // expose controls/target/v/lambda/Q_diag of dlib::mpc ONLY
#include <dlib/control/mpc.h>
#undef private

#include <chrono>
#include <cstdlib>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

namespace {

struct ctrl_base {
    virtual ~ctrl_base() {}
    virtual void set_target(const double* val, long time) = 0;
    virtual void set_options(double eps, unsigned long max_iter) = 0;
    virtual void step(const double* x0, double* u0) = 0;
    virtual void get_state(double* controls, double* target, double* v) const = 0;
    virtual void set_state(const double* controls, const double* target, const double* v) = 0;
    virtual void get_derived(double* lambda, double* q_diag) const = 0;
};

template <long S, long I, unsigned long N>
struct ctrl : ctrl_base {
    typedef dlib::mpc<S, I, N> mpc_t;
    // zeroed raw storage, then placement-new: uninitialised members read as 0.0
    typename std::aligned_storage<sizeof(mpc_t), alignof(mpc_t)>::type raw;
    mpc_t* m;

    ctrl(const double* A, const double* B, const double* C, const double* Q, const double* R,
         const double* lo, const double* hi)
    {
        std::memset(&raw, 0, sizeof(raw));
        dlib::matrix<double, S, S> a; dlib::matrix<double, S, I> b;
        dlib::matrix<double, S, 1> c, q; dlib::matrix<double, I, 1> r, l, u;
        for (long i = 0; i < S; ++i) for (long j = 0; j < S; ++j) a(i, j) = A[i * S + j];
        for (long i = 0; i < S; ++i) for (long j = 0; j < I; ++j) b(i, j) = B[i * I + j];
        for (long i = 0; i < S; ++i) { c(i) = C[i]; q(i) = Q[i]; }
        for (long j = 0; j < I; ++j) { r(j) = R[j]; l(j) = lo[j]; u(j) = hi[j]; }
        m = new (&raw) mpc_t(a, b, c, q, r, l, u);
    }
    ~ctrl() { m->~mpc_t(); }

    void set_target(const double* val, long time) override {
        dlib::matrix<double, S, 1> t; for (long i = 0; i < S; ++i) t(i) = val[i];
        if (time < 0) m->set_target(t); else m->set_target(t, (unsigned long)time);
    }
    void set_options(double eps, unsigned long max_iter) override {
        m->set_epsilon(eps); m->set_max_iterations(max_iter);
    }
    void step(const double* x0, double* u0) override {
        dlib::matrix<double, S, 1> x; for (long i = 0; i < S; ++i) x(i) = x0[i];
        dlib::matrix<double, I, 1> u = (*m)(x);
        for (long j = 0; j < I; ++j) u0[j] = u(j);
    }
    void get_state(double* controls, double* target, double* v) const override {
        for (unsigned long t = 0; t < N; ++t) {
            for (long j = 0; j < I; ++j) {
                if (controls) controls[t * I + j] = m->controls[t](j);
                if (v) v[t * I + j] = m->v[t](j);
            }
            if (target) for (long i = 0; i < S; ++i) target[t * S + i] = m->target[t](i);
        }
    }
    void set_state(const double* controls, const double* target, const double* v) override {
        for (unsigned long t = 0; t < N; ++t) {
            for (long j = 0; j < I; ++j) {
                if (controls) m->controls[t](j) = controls[t * I + j];
                if (v) m->v[t](j) = v[t * I + j];
            }
            if (target) for (long i = 0; i < S; ++i) m->target[t](i) = target[t * S + i];
        }
    }
    void get_derived(double* lambda, double* q_diag) const override {
        *lambda = m->lambda;
        for (unsigned long t = 0; t < N; ++t) for (long j = 0; j < I; ++j) q_diag[t * I + j] = m->Q_diag[t](j);
    }
};

#define DLIBREF_SHAPES(X) \
    X(2, 2, 4) X(2, 1, 30) X(4, 1, 10) X(4, 1, 30) X(4, 1, 60) X(4, 1, 100) X(4, 2, 10) X(3, 2, 7)

ctrl_base* make(int S, int I, int N, const double* A, const double* B, const double* C, const double* Q,
                const double* R, const double* lo, const double* hi)
{
#define X(s, i, n) if (S == s && I == i && N == n) return new ctrl<s, i, n>(A, B, C, Q, R, lo, hi);
    DLIBREF_SHAPES(X)
#undef X
    return nullptr;
}

}  // namespace

extern "C" {

int dlibref_supported(int S, int I, int N)
{
#define X(s, i, n) if (S == s && I == i && N == n) return 1;
    DLIBREF_SHAPES(X)
#undef X
    return 0;
}

void* dlibref_new(int S, int I, int N, const double* A, const double* B, const double* C, const double* Q,
                  const double* R, const double* lo, const double* hi)
{ return make(S, I, N, A, B, C, Q, R, lo, hi); }

void dlibref_free(void* h) { delete static_cast<ctrl_base*>(h); }
void dlibref_set_target(void* h, const double* val, long time) { static_cast<ctrl_base*>(h)->set_target(val, time); }
void dlibref_set_options(void* h, double eps, unsigned long max_iter) { static_cast<ctrl_base*>(h)->set_options(eps, max_iter); }
void dlibref_step(void* h, const double* x0, double* u0) { static_cast<ctrl_base*>(h)->step(x0, u0); }
void dlibref_get_state(void* h, double* controls, double* target, double* v) { static_cast<ctrl_base*>(h)->get_state(controls, target, v); }
void dlibref_set_state(void* h, const double* controls, const double* target, const double* v) { static_cast<ctrl_base*>(h)->set_state(controls, target, v); }
void dlibref_get_derived(void* h, double* lambda, double* q_diag) { static_cast<ctrl_base*>(h)->get_derived(lambda, q_diag); }

// Batched COLD solves, the reference's own usage pattern (a fresh dlib::mpc per solve,
// src/trajectory_point_follower.cpp:366-380): construct, set_target(all steps), solve.
//   model arrays: per_instance_model==0 -> one model shared; ==1 -> arrays of `count` models,
//   instance-major (A: [count][S*S], B: [count][S*I], C,Q: [count][S], R,lo,hi: [count][I]).
//   x0: [count][S]; target: [count][S] or NULL (zeros); u0: [count][I]; controls: [count][N*I] or NULL.
// Work is split into contiguous slices over `nthreads` std::threads (one dlib::mpc alive per thread).
// Returns wall seconds of the parallel region (steady_clock), or -1 if the shape is not instantiated.
double dlibref_batch_cold(int S, int I, int N, int per_instance_model, const double* A, const double* B,
                          const double* C, const double* Q, const double* R, const double* lo, const double* hi,
                          long count, const double* x0, const double* target, double eps, unsigned long max_iter,
                          double* u0, double* controls, int nthreads)
{
    if (!dlibref_supported(S, I, N)) return -1.0;
    if (nthreads < 1) nthreads = 1;
    auto work = [&](long lo_i, long hi_i) {
        std::vector<double> zeros(S, 0.0);
        for (long k = lo_i; k < hi_i; ++k) {
            const long m = per_instance_model ? k : 0;
            ctrl_base* c = make(S, I, N, A + m * S * S, B + m * S * I, C + m * S, Q + m * S, R + m * I,
                                lo + m * I, hi + m * I);
            c->set_options(eps, max_iter);
            c->set_target(target ? target + k * S : zeros.data(), -1);
            c->step(x0 + k * S, u0 + k * I);
            if (controls) c->get_state(controls + k * (long)N * I, nullptr, nullptr);
            delete c;
        }
    };
    auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) {
        long a = count * t / nthreads, b = count * (t + 1) / nthreads;
        th.emplace_back(work, a, b);
    }
    for (auto& t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    return std::chrono::duration<double>(t1 - t0).count();
}

// Closed loop of ONE persistent controller (dlib's own usage, dlib/test/mpc.cpp:295-316):
// for k in 0..ticks-1: u = ctrl(x); record; x = A x + B u + C (plain row-major dot products).
// If last_targets != NULL ([ticks][S]) it calls set_last_target(last_targets[k]) before each solve.
// u_hist: [ticks][I]; x_hist: [ticks][S] (state BEFORE the tick) or NULL.
void dlibref_closed_loop(void* h, int S, int I, const double* A, const double* B, const double* C, double* x,
                         long ticks, const double* last_targets, int N, double* u_hist, double* x_hist)
{
    ctrl_base* c = static_cast<ctrl_base*>(h);
    std::vector<double> xn(S), u(I);
    for (long k = 0; k < ticks; ++k) {
        if (last_targets) c->set_target(last_targets + k * S, N - 1);
        if (x_hist) std::memcpy(x_hist + k * S, x, sizeof(double) * S);
        c->step(x, u.data());
        std::memcpy(u_hist + k * I, u.data(), sizeof(double) * I);
        for (int r = 0; r < S; ++r) {
            double t = A[r * S] * x[0];
            for (int j = 1; j < S; ++j) t += A[r * S + j] * x[j];
            double b = B[r * I] * u[0];
            for (int j = 1; j < I; ++j) b += B[r * I + j] * u[j];
            xn[r] = (t + b) + C[r];
        }
        std::memcpy(x, xn.data(), sizeof(double) * S);
    }
}

}  // extern "C"
