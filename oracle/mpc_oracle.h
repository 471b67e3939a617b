/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement (plain C, double precision, no FMA contraction) of the reference's
 * linear-MPC solver, dlib::mpc<S,I,horizon> as vendored in
 * /root/reference/dlib_files/dlib/control/mpc.h (dlib 18.18).  It exists so that tests can
 * observe what dlib hides (the iteration count, the full control sequence) and so that the
 * GPU box — where /root/reference does not exist — still has a checker.
 *
 * PARITY IS PINNED: tests/test_oracle_vs_dlib.py checks this file bit-for-bit (controls[],
 * u0, lambda, Q_diag) against the real header compiled in oracle/_ref/libdlibref.so, against
 * dlib's own known-answer test (dlib/test/mpc.cpp:257-318) and against the golden fixtures in
 * tests/golden/ that were generated from the real header.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * use anything under oracle/.  The product (trajectory_controller_b200/) never does.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (never -march=native / -ffast-math).
 */
#ifndef MPC_ORACLE_H
#define MPC_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_S 8
#define ORC_MAX_I 4
#define ORC_MAX_N 128

typedef struct orc_mpc {
    int S, I, N;
    /* mpc.h:349-358 — parameters, row-major like dlib::matrix (matrix_data_layout.h:131-155) */
    double A[ORC_MAX_S * ORC_MAX_S];
    double B[ORC_MAX_S * ORC_MAX_I];
    double C[ORC_MAX_S];
    double Q[ORC_MAX_S];
    double R[ORC_MAX_I];
    double lower[ORC_MAX_I];
    double upper[ORC_MAX_I];
    unsigned long max_iterations; /* mpc.h:103 */
    double eps;                   /* mpc.h:104 */
    double lambda;                /* mpc.h:360 */
    double Q_diag[ORC_MAX_N * ORC_MAX_I];   /* mpc.h:361 */
    /* persistent state: mpc.h:358,362 and the scratch member v (mpc.h:249) that can be read stale */
    double target[ORC_MAX_N * ORC_MAX_S];
    double controls[ORC_MAX_N * ORC_MAX_I];
    double v[ORC_MAX_N * ORC_MAX_I];
    /* observability that dlib does not offer */
    unsigned long last_iters;     /* value of `iter` when solve_linear_mpc returned (mpc.h:270-271,310) */
    double last_max_df;           /* max_df of the last stopping test evaluated */
} orc_mpc;

/* mpc.h:51-125 (7-argument constructor incl. the lambda / Q_diag recursion at :114-123).
 * Returns 0, or -1 when (S,I,N) exceed the ORC_MAX_* bounds. */
int  orc_mpc_init(orc_mpc* m, int S, int I, int N, const double* A, const double* B, const double* C,
                  const double* Q, const double* R, const double* lower, const double* upper);
/* mpc.h:142-170; time < 0 means "all steps" (set_target(val)). */
void orc_mpc_set_target(orc_mpc* m, const double* val, long time);
/* mpc.h:216-240: shift controls, solve_linear_mpc (mpc.h:253-347), shift targets, return controls[0]. */
void orc_mpc_step(orc_mpc* m, const double* x0, double* u0);

/* Batched cold solves with the same argument convention as dlibref_batch_cold (oracle/ref/dlib_ref.cpp).
 * iters: [count] or NULL.  nthreads pthreads over contiguous slices.  Returns wall seconds. */
double orc_batch_cold(int S, int I, int N, int per_instance_model, const double* A, const double* B,
                      const double* C, const double* Q, const double* R, const double* lo, const double* hi,
                      long count, const double* x0, const double* target, double eps, unsigned long max_iter,
                      double* u0, double* controls, unsigned int* iters, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
