// bmpc::mpc<S,I,horizon> — host C++ facade with dlib::mpc's method set, solving on the GPU through the C ABI.
//
// Mirrors /root/reference/dlib_files/dlib/control/mpc.h:13-240 (contract: mpc_abstract.h:70-269):
//   statics S, I, horizon; default ctor; 7-argument ctor (A,B,C,Q,R,lower,upper);
//   get_A/B/C/Q/R, get_lower_constraints, get_upper_constraints;
//   set_target(val,time), set_target(val), set_last_target(val), get_target(time);
//   get/set_max_iterations, get/set_epsilon;  operator()(current_state) -> first control.
// Same defaults (eps = 0.01, max_iterations = 10000 — mpc.h:103-104; a default-constructed object has
// max_iterations = 0 — mpc.h:38), same state evolution (controls shift before the solve, targets shift after it —
// mpc.h:229-237), by-value copy semantics.  There is no set_last_control in dlib 18.18; `last_iterations()` and
// `seed_controls()` are EXTENSIONS and say so.
//
// Each object owns one capacity-1 engine slot (a bmpc_t handle).  Handles are expensive to create (device
// allocations, a stream), and the reference module constructs a fresh controller every cycle
// (src/trajectory_point_follower.cpp:366), so released handles go to a per-thread, per-shape pool and are reset on
// reuse: steady-state construction costs a few small launches, no allocation.
//
// Errors: dlib::mpc cannot fail; a GPU can.  Any ABI error throws bmpc::error (there is NO CPU fallback).
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "bmpc.h"          // the C ABI (repo: include/bmpc.h)
#include "bmpc/matrix.h"

namespace bmpc {

class error : public std::runtime_error {
public:
    int code;
    error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

namespace detail {
inline void check(int rc, const char* where)
{
    if (rc != BMPC_OK) throw error(rc, std::string(where) + ": " + bmpc_last_error());
}

// Per-thread free list of capacity-1 handles of one shape.  Destroyed with the thread.
template <long S, long I, unsigned long N>
struct handle_pool {
    std::vector<bmpc_t*> free_;
    ~handle_pool() { for (bmpc_t* h : free_) bmpc_destroy(h); }
    static handle_pool& get() { static thread_local handle_pool p; return p; }
    bmpc_t* acquire()
    {
        if (!free_.empty()) {
            bmpc_t* h = free_.back(); free_.pop_back();
            check(bmpc_reset(h), "bmpc_reset");                       // controls = 0, targets = 0 (mpc.h:105-112)
            check(bmpc_set_options(h, 0.01, 10000), "bmpc_set_options");   // mpc.h:103-104
            return h;
        }
        bmpc_t* h = nullptr;
        check(bmpc_create(&h, (int)S, (int)I, (int)N, 1, BMPC_FP64, BMPC_MODE_EXACT, -1), "bmpc_create");
        return h;
    }
    void release(bmpc_t* h) { if (h) free_.push_back(h); }
};
}  // namespace detail

template <long S_, long I_, unsigned long horizon_>
class mpc {
public:
    const static long S = S_;
    const static long I = I_;
    const static unsigned long horizon = horizon_;

    typedef matrix<double, S_, S_> mat_A;
    typedef matrix<double, S_, I_> mat_B;
    typedef matrix<double, S_, 1> vec_S;
    typedef matrix<double, I_, 1> vec_I;

    // mpc.h:27-49: everything zero, max_iterations = 0.  No engine slot is taken until a model exists.
    mpc() : h_(nullptr), eps_(0.01), max_iterations_(0), last_iters_(0) {}

    // mpc.h:51-125
    mpc(const mat_A& A, const mat_B& B, const vec_S& C, const vec_S& Q, const vec_I& R, const vec_I& lower,
        const vec_I& upper)
        : h_(nullptr), A_(A), B_(B), C_(C), Q_(Q), R_(R), lower_(lower), upper_(upper), eps_(0.01),
          max_iterations_(10000), last_iters_(0)
    {
        h_ = pool().acquire();
        try {
            detail::check(bmpc_set_model(h_, A_.data(), B_.data(), C_.data(), Q_.data(), R_.data(), lower_.data(),
                                         upper_.data(), 0, 1), "bmpc_set_model");
        } catch (...) { pool().release(h_); h_ = nullptr; throw; }
    }

    mpc(const mpc& o) : h_(nullptr) { copy_from(o); }
    mpc& operator=(const mpc& o)
    {
        if (this != &o) { drop(); copy_from(o); }
        return *this;
    }
    ~mpc() { drop(); }

    const mat_A& get_A() const { return A_; }
    const mat_B& get_B() const { return B_; }
    const vec_S& get_C() const { return C_; }
    const vec_S& get_Q() const { return Q_; }
    const vec_I& get_R() const { return R_; }
    const vec_I& get_lower_constraints() const { return lower_; }
    const vec_I& get_upper_constraints() const { return upper_; }

    // mpc.h:127-140
    const vec_S get_target(const unsigned long time) const
    {
        vec_S t;
        if (time >= horizon_) throw error(BMPC_E_INVALID, "mpc::get_target: time must be < horizon (mpc.h:133)");
        if (h_) detail::check(bmpc_get_target(h_, 0, (int)time, t.data()), "bmpc_get_target");
        return t;
    }
    // mpc.h:142-155
    void set_target(const vec_S& val, const unsigned long time)
    {
        if (time >= horizon_) throw error(BMPC_E_INVALID, "mpc::set_target: time must be < horizon (mpc.h:146)");
        if (h_) detail::check(bmpc_set_target(h_, val.data(), (int)time, 0, 0), "bmpc_set_target");
    }
    // mpc.h:157-163
    void set_target(const vec_S& val)
    {
        if (h_) detail::check(bmpc_set_target(h_, val.data(), -1, 0, 0), "bmpc_set_target");
    }
    // mpc.h:165-170
    void set_last_target(const vec_S& val) { set_target(val, horizon_ - 1); }

    // mpc.h:172-214
    unsigned long get_max_iterations() const { return max_iterations_; }
    void set_max_iterations(unsigned long max_iter)
    {
        max_iterations_ = max_iter;
        if (h_) detail::check(bmpc_set_options(h_, eps_, max_iterations_), "bmpc_set_options");
    }
    void set_epsilon(double eps)
    {
        if (!(eps > 0)) throw error(BMPC_E_INVALID, "mpc::set_epsilon: eps must be > 0 (mpc.h:203)");
        eps_ = eps;
        if (h_) detail::check(bmpc_set_options(h_, eps_, max_iterations_), "bmpc_set_options");
    }
    double get_epsilon() const { return eps_; }

    // mpc.h:216-240: shift the stored controls, solve, shift the targets, return controls[0].
    vec_I operator()(const vec_S& current_state)
    {
        vec_I u;                                   // default-constructed object: controls are zero and stay zero
        if (!h_) return u;
        uint32_t it = 0;
        detail::check(bmpc_solve(h_, 1, current_state.data(), u.data(), &it, nullptr, BMPC_WARM), "bmpc_solve");
        last_iters_ = it;
        return u;
    }

    // ---- EXTENSIONS (not part of dlib::mpc) ------------------------------------------------------------------
    // The value of solve_linear_mpc's local `iter` (mpc.h:270) when the last operator() returned.
    unsigned long last_iterations() const { return last_iters_; }
    // Overwrite the stored control sequence (the warm start the next operator() shifts and starts from).
    void seed_controls(const matrix<double, (long)horizon_ * I_, 1>& controls)
    {
        if (h_) detail::check(bmpc_set_controls(h_, 0, 1, controls.data()), "bmpc_set_controls");
    }
    bmpc_t* native_handle() const { return h_; }

private:
    typedef detail::handle_pool<S_, I_, horizon_> pool_t;
    static pool_t& pool() { return pool_t::get(); }

    void drop()
    {
        if (h_) { pool().release(h_); h_ = nullptr; }
    }
    void copy_from(const mpc& o)
    {
        A_ = o.A_; B_ = o.B_; C_ = o.C_; Q_ = o.Q_; R_ = o.R_; lower_ = o.lower_; upper_ = o.upper_;
        eps_ = o.eps_; max_iterations_ = o.max_iterations_; last_iters_ = o.last_iters_;
        if (!o.h_) return;
        h_ = pool().acquire();
        try {
            detail::check(bmpc_set_model(h_, A_.data(), B_.data(), C_.data(), Q_.data(), R_.data(), lower_.data(),
                                         upper_.data(), 0, 1), "bmpc_set_model");
            detail::check(bmpc_copy_state(h_, o.h_), "bmpc_copy_state");
        } catch (...) { pool().release(h_); h_ = nullptr; throw; }
    }

    bmpc_t* h_;
    mat_A A_;
    mat_B B_;
    vec_S C_, Q_;
    vec_I R_, lower_, upper_;
    double eps_;
    unsigned long max_iterations_;
    unsigned long last_iters_;
};

}  // namespace bmpc
