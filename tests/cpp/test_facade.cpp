// dlib::mpc-compatible facade (bmpc/mpc.h via dlib_compat) exercised the way dlib's own test and the reference
// module use dlib::mpc.  Needs a GPU (the facade has no CPU path).
//   1. dlib's known-answer scenario (dlib_files/dlib/test/mpc.cpp:257-318): mpc<2,1,30>, eps 1e-8, 30 closed-loop
//      ticks, every u0 within 1e-7 of an INDEPENDENT dense box-QP solve (here: condensed Hessian + projected
//      Gauss-Seidel in long double — not dlib's SMO, not the engine's recursion).
//   2. the same scenario at default eps: golden values recorded from the real header (tests/golden/make_golden.py).
//   3. the reference module's controller mpc<2,2,4> (src/trajectory_point_follower.cpp:326-380): golden u0.
//   4. object semantics: defaults, accessors, target shifting, copy construction / assignment, default ctor.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include <dlib/control.h>

#include "check.h"

using dlib::matrix;

namespace {

// min 1/2 u'Hu + g'u, lo <= u <= hi, by cyclic projected coordinate descent (H positive definite).
std::vector<long double> box_qp(const std::vector<long double>& H, const std::vector<long double>& g, int n,
                                long double lo, long double hi)
{
    std::vector<long double> u(n, 0.0L);
    for (int sweep = 0; sweep < 200000; ++sweep) {
        long double change = 0;
        for (int k = 0; k < n; ++k) {
            long double r = g[k];
            for (int j = 0; j < n; ++j) if (j != k) r += H[k * n + j] * u[j];
            long double x = -r / H[k * n + k];
            x = x < lo ? lo : (x > hi ? hi : x);
            change = fmaxl(change, fabsl(x - u[k]));
            u[k] = x;
        }
        if (change < 1e-15L) break;
    }
    return u;
}

// Condensed form of dlib::mpc's objective for S=2, I=1: states x_{i+1} = A x_i + B u_i + C, cost
// sum_i Q.*(x_{i+1} - target_i)^2 + R u_i^2  (mpc_abstract.h:24-49).
double reference_u0(const matrix<double, 2, 2>& A, const matrix<double, 2, 1>& B, const matrix<double, 2, 1>& C,
                    const matrix<double, 2, 1>& Q, double R, double lo, double hi, const matrix<double, 2, 1>& x0, int N)
{
    typedef long double ld;
    auto simulate = [&](const std::vector<ld>& u, std::vector<ld>& xs) {
        ld a = x0(0), b = x0(1);
        for (int i = 0; i < N; ++i) {
            const ld na = A(0, 0) * a + A(0, 1) * b + B(0) * u[i] + C(0);
            const ld nb = A(1, 0) * a + A(1, 1) * b + B(1) * u[i] + C(1);
            a = na; b = nb;
            xs[2 * i] = a; xs[2 * i + 1] = b;
        }
    };
    std::vector<ld> zero(N, 0), f(2 * N), col(2 * N);
    simulate(zero, f);                                     // free response (target is 0 in the scenario)
    std::vector<ld> K(2 * N * N);
    for (int k = 0; k < N; ++k) {
        std::vector<ld> e(N, 0); e[k] = 1;
        simulate(e, col);
        for (int r = 0; r < 2 * N; ++r) K[r * N + k] = col[r] - f[r];
    }
    std::vector<ld> H(N * N, 0), g(N, 0);
    for (int a = 0; a < N; ++a) {
        for (int b = 0; b < N; ++b) {
            ld s = 0;
            for (int r = 0; r < 2 * N; ++r) s += K[r * N + a] * (ld)Q(r % 2) * K[r * N + b];
            H[a * N + b] = s + (a == b ? (ld)R : 0);
        }
        ld s = 0;
        for (int r = 0; r < 2 * N; ++r) s += K[r * N + a] * (ld)Q(r % 2) * f[r];
        g[a] = s;
    }
    return (double)box_qp(H, g, N, lo, hi)[0];
}

uint64_t fnv1a(const std::vector<double>& v)
{
    uint64_t h = 0xCBF29CE484222325ull;
    for (double d : v) {
        unsigned char b[8]; std::memcpy(b, &d, 8);
        for (int k = 0; k < 8; ++k) { h ^= b[k]; h *= 0x100000001B3ull; }
    }
    return h;
}

void kat(double eps, bool check_qp, bool check_golden)
{
    matrix<double, 2, 2> A; A = 1, 1, 0, 1;
    matrix<double, 2, 1> B, C, Q; B = 0, 1; C = 0.02, 0.1; Q = 2, 0;
    matrix<double, 1, 1> R, lower, upper; R = 1; lower = -0.2; upper = 0.2;
    dlib::mpc<2, 1, 30> solver(A, B, C, Q, R, lower, upper);
    CHECK(solver.get_epsilon() == 0.01 && solver.get_max_iterations() == 10000);      // mpc.h:103-104
    solver.set_epsilon(eps);
    solver.set_max_iterations(10000);
    matrix<double, 2, 1> x; x = 0; x(0) = 5;
    std::vector<double> u0s;
    for (int i = 0; i < 30; ++i) {
        const matrix<double, 1, 1> control = solver(x);
        if (check_qp) {
            const double want = reference_u0(A, B, C, Q, 1.0, -0.2, 0.2, x, 30);
            CHECK_MSG(std::fabs(control(0) - want) < 1e-7, "tick %d: u0 %.12g vs dense QP %.12g", i, control(0), want);
        }
        CHECK(solver.last_iterations() <= 10000);
        u0s.push_back(control(0));
        x = A * x + B * control + C;
    }
    if (check_golden) {
        CHECK_MSG(u0s[29] == -0.10000002376964306, "u0[29] = %.17g", u0s[29]);
        // FNV-1a-64 over the little-endian bytes of the 30 u0 of the REAL dlib::mpc (tests/golden/make_golden.py prints it)
        CHECK_MSG(fnv1a(u0s) == 0xce44159b7ae1643cull, "hash %016llx", (unsigned long long)fnv1a(u0s));
    }
}

void tobi_spot()
{
    const double T = 0.1, v = 1.0, l = 0.21, amax = 22 * M_PI / 180;
    matrix<double, 2, 2> A; A = 1, T * v, 0, 1;
    matrix<double, 2, 2> B; B = 0, T * v, T * v / l, -T * v / l;
    matrix<double, 2, 1> C, Q, R, lo, hi, target, x0;
    C = 0, 0; Q = 20, 7; R = 0.0005, 10; lo = -amax, -amax; hi = amax, amax; target = 0.05, 0.1; x0 = 0, 0;
    dlib::mpc<2, 2, 4> controller(A, B, C, Q, R, lo, hi);
    controller.set_target(target);
    const matrix<double, 2, 1> u = controller(x0);
    CHECK_MSG(u(0) == 0.27357241694602841 && u(1) == 0.023830246920932288, "u = (%.17g, %.17g)", u(0), u(1));
    // a second fresh controller (handle comes back from the pool) must start from zero state again
    dlib::mpc<2, 2, 4> again(A, B, C, Q, R, lo, hi);
    again.set_target(target);
    const matrix<double, 2, 1> u2 = again(x0);
    CHECK(u2(0) == u(0) && u2(1) == u(1));
}

void object_semantics()
{
    dlib::mpc<2, 1, 30> empty;                                   // mpc.h:27-49
    CHECK(empty.get_max_iterations() == 0 && empty.get_epsilon() == 0.01);
    matrix<double, 2, 1> x; x = 1, 2;
    CHECK(empty(x)(0) == 0.0);
    CHECK((dlib::mpc<2, 1, 30>::S == 2) && (dlib::mpc<2, 1, 30>::I == 1) && (dlib::mpc<2, 1, 30>::horizon == 30));

    matrix<double, 2, 2> A; A = 1, 1, 0, 1;
    matrix<double, 2, 1> B, C, Q; B = 0, 1; C = 0.02, 0.1; Q = 2, 0;
    matrix<double, 1, 1> R, lower, upper; R = 1; lower = -0.2; upper = 0.2;
    dlib::mpc<2, 1, 30> a(A, B, C, Q, R, lower, upper);
    CHECK(a.get_A()(0, 1) == 1 && a.get_B()(1) == 1 && a.get_C()(1) == 0.1 && a.get_Q()(0) == 2 && a.get_R()(0) == 1);
    CHECK(a.get_lower_constraints()(0) == -0.2 && a.get_upper_constraints()(0) == 0.2);
    // targets: set_target(val,time), set_last_target, and the shift after each operator() (mpc.h:236-237)
    matrix<double, 2, 1> t5, t29; t5 = 0.5, -0.5; t29 = 2.9, -2.9;
    a.set_target(t5, 5);
    a.set_last_target(t29);
    CHECK(a.get_target(5)(0) == 0.5 && a.get_target(29)(1) == -2.9 && a.get_target(4)(0) == 0);
    x = 5, 0;
    a(x);
    CHECK(a.get_target(4)(0) == 0.5 && a.get_target(5)(0) == 0 && a.get_target(28)(0) == 2.9 && a.get_target(29)(0) == 2.9);
    // copies are independent objects with the same state
    dlib::mpc<2, 1, 30> b(a);
    dlib::mpc<2, 1, 30> c;
    c = a;
    CHECK(b.get_target(4)(0) == 0.5 && c.get_target(28)(1) == -2.9);
    const double ua = a(x)(0), ub = b(x)(0), uc = c(x)(0);
    CHECK(ua == ub && ua == uc);
    CHECK(a.last_iterations() == b.last_iterations() && a.last_iterations() > 0);
    b(x);                                                          // advancing b must not disturb a or c
    const double ua2 = a(x)(0), uc2 = c(x)(0);
    CHECK(ua2 == uc2);
    bool threw = false;
    try { a.set_target(t5, 30); } catch (const bmpc::error&) { threw = true; }
    CHECK(threw);
}

}  // namespace

int main()
{
    try {
        kat(1e-8, true, false);
        kat(0.01, false, true);
        tobi_spot();
        object_semantics();
    } catch (const std::exception& e) {
        std::fprintf(stderr, "exception: %s\n", e.what());
        return 2;
    }
    return finish("test_facade");
}
