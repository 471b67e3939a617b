/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See mpc_oracle.h for the contract and how parity is pinned.
 *
 * Every function cites the reference lines it restates: /root/reference/dlib_files/dlib/control/mpc.h
 * ("mpc.h") and the dlib::matrix pieces it instantiates:
 *   - products are per-element dot products, k = 0..n-1, first term initialises (matrix/matrix.h:43-60)
 *   - diagm(q)*x is a DENSE product with literal zeros (matrix/matrix_utilities.h:879-925)
 *   - sums associate left; x += e is x(r) + e(r) (matrix/matrix.h:1370-1398)
 *   - clamp(val,lo,hi) = val<=hi ? (lo<=val ? val : lo) : hi (matrix/matrix_utilities.h:2829-2846)
 *   - put_in_range(a,b,val) (algs.h:716-748)
 * Compile with -ffp-contract=off: the reference build has no FMA contraction.
 */
#include "mpc_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* y = X z, X is n x n row-major.  matrix.h:43-60 */
static void mat_vec(int n, const double* X, const double* z, double* y)
{
    for (int r = 0; r < n; ++r) {
        double t = X[r * n] * z[0];
        for (int k = 1; k < n; ++k) t += X[r * n + k] * z[k];
        y[r] = t;
    }
}
/* y = trans(X) z, X is n x n row-major.  matrix_utilities.h:375-410 feeding matrix.h:43-60 */
static void matT_vec(int n, const double* X, const double* z, double* y)
{
    for (int r = 0; r < n; ++r) {
        double t = X[r] * z[0];
        for (int k = 1; k < n; ++k) t += X[k * n + r] * z[k];
        y[r] = t;
    }
}
/* y = Bm u, Bm is S x I */
static void B_vec(int S, int I, const double* Bm, const double* u, double* y)
{
    for (int r = 0; r < S; ++r) {
        double t = Bm[r * I] * u[0];
        for (int k = 1; k < I; ++k) t += Bm[r * I + k] * u[k];
        y[r] = t;
    }
}
/* y = trans(Bm) m, result I x 1 */
static void BT_vec(int S, int I, const double* Bm, const double* m, double* y)
{
    for (int j = 0; j < I; ++j) {
        double t = Bm[j] * m[0];
        for (int k = 1; k < S; ++k) t += Bm[k * I + j] * m[k];
        y[j] = t;
    }
}
/* y = diagm(q) z evaluated as the dense product dlib performs (zeros multiplied and added). */
static void diagm_vec(int n, const double* q, const double* z, double* y)
{
    for (int r = 0; r < n; ++r) {
        double t = (r == 0 ? q[0] : 0.0) * z[0];
        for (int k = 1; k < n; ++k) t += (r == k ? q[k] : 0.0) * z[k];
        y[r] = t;
    }
}
/* matrix_utilities.h:2829-2846 (op_clamp_m) */
static double clamp_m(double val, double lo, double hi) { return val <= hi ? (lo <= val ? val : lo) : hi; }
/* algs.h:716-748 */
static double put_in_range(double a, double b, double val)
{
    if (a < b) { if (val < a) return a; else if (val > b) return b; }
    else       { if (val < b) return b; else if (val > a) return a; }
    return val;
}

/* mpc.h:51-125 */
int orc_mpc_init(orc_mpc* m, int S, int I, int N, const double* A, const double* B, const double* C,
                 const double* Q, const double* R, const double* lower, const double* upper)
{
    if (S < 1 || I < 1 || N < 1 || S > ORC_MAX_S || I > ORC_MAX_I || N > ORC_MAX_N) return -1;
    memset(m, 0, sizeof(*m));   /* target, controls = 0 (mpc.h:105-112); v = 0 is OUR definition of the stale read */
    m->S = S; m->I = I; m->N = N;
    memcpy(m->A, A, sizeof(double) * S * S);
    memcpy(m->B, B, sizeof(double) * S * I);
    memcpy(m->C, C, sizeof(double) * S);
    memcpy(m->Q, Q, sizeof(double) * S);
    memcpy(m->R, R, sizeof(double) * I);
    memcpy(m->lower, lower, sizeof(double) * I);
    memcpy(m->upper, upper, sizeof(double) * I);
    m->max_iterations = 10000;  /* mpc.h:103 */
    m->eps = 0.01;              /* mpc.h:104 */

    /* mpc.h:116: lambda = sum(R)*horizon.  sum() accumulates from 0 (matrix_utilities.h:1420-1437). */
    double sumR = 0;
    for (int j = 0; j < I; ++j) sumR += R[j];
    m->lambda = sumR * (double)N;

    /* mpc.h:117: temp = diagm(Q) */
    double T[ORC_MAX_S * ORC_MAX_S], BtT[ORC_MAX_I * ORC_MAX_S], AtT[ORC_MAX_S * ORC_MAX_S], Tn[ORC_MAX_S * ORC_MAX_S];
    for (int r = 0; r < S; ++r) for (int c = 0; c < S; ++c) T[r * S + c] = (r == c) ? Q[r] : 0.0;

    for (int c = 0; c < N; ++c) {   /* mpc.h:118-123 */
        /* (trans(B)*temp): I x S */
        for (int i = 0; i < I; ++i) for (int k = 0; k < S; ++k) {
            double t = B[i] * T[k];                                  /* l = 0: B(0,i)*T(0,k) */
            for (int l = 1; l < S; ++l) t += B[l * I + i] * T[l * S + k];
            BtT[i * S + k] = t;
        }
        /* diagonal of (trans(B)*temp)*B — evaluated twice by the reference (trace, diag) with identical results */
        double tr = 0;
        for (int j = 0; j < I; ++j) {
            double t = BtT[j * S] * B[j];
            for (int k = 1; k < S; ++k) t += BtT[j * S + k] * B[k * I + j];
            tr += t;                                                  /* trace = sum(diag()) matrix_la.h:1662-1680 */
            m->Q_diag[(N - 1 - c) * I + j] = t;                       /* mpc.h:121 */
        }
        m->lambda += tr;                                              /* mpc.h:120 */
        /* temp = trans(A)*temp*A + diagm(Q)   mpc.h:122 */
        for (int r = 0; r < S; ++r) for (int k = 0; k < S; ++k) {
            double t = A[r] * T[k];
            for (int l = 1; l < S; ++l) t += A[l * S + r] * T[l * S + k];
            AtT[r * S + k] = t;
        }
        for (int r = 0; r < S; ++r) for (int k = 0; k < S; ++k) {
            double t = AtT[r * S] * A[k];
            for (int l = 1; l < S; ++l) t += AtT[r * S + l] * A[l * S + k];
            Tn[r * S + k] = t + ((r == k) ? Q[r] : 0.0);
        }
        memcpy(T, Tn, sizeof(double) * S * S);
    }
    return 0;
}

/* mpc.h:142-170 */
void orc_mpc_set_target(orc_mpc* m, const double* val, long time)
{
    if (time < 0) { for (int i = 0; i < m->N; ++i) memcpy(m->target + i * m->S, val, sizeof(double) * m->S); }
    else memcpy(m->target + time * m->S, val, sizeof(double) * m->S);
}

/* mpc.h:253-347 */
static void solve_linear_mpc(orc_mpc* m, const double* x0)
{
    const int S = m->S, I = m->I, N = m->N;
    double M[ORC_MAX_N * ORC_MAX_S], MM[ORC_MAX_N * ORC_MAX_I], df[ORC_MAX_N * ORC_MAX_I];
    double t1[ORC_MAX_S], t2[ORC_MAX_S];
    double* u = m->controls;

    /* mpc.h:258-266: MM = trans(K)*Q*(M - target) */
    mat_vec(S, m->A, x0, t1);
    for (int r = 0; r < S; ++r) M[r] = t1[r] + m->C[r];
    for (int i = 1; i < N; ++i) {
        mat_vec(S, m->A, M + (i - 1) * S, t1);
        for (int r = 0; r < S; ++r) M[i * S + r] = t1[r] + m->C[r];
    }
    for (int i = 0; i < N; ++i) {
        for (int r = 0; r < S; ++r) t1[r] = M[i * S + r] - m->target[i * S + r];
        diagm_vec(S, m->Q, t1, M + i * S);
    }
    for (int i = N - 2; i >= 0; --i) {
        matT_vec(S, m->A, M + (i + 1) * S, t1);
        for (int r = 0; r < S; ++r) M[i * S + r] = M[i * S + r] + t1[r];
    }
    for (int i = 0; i < N; ++i) BT_vec(S, I, m->B, M + i * S, MM + i * I);

    const double lambda = m->lambda;
    unsigned long iter = 0;
    m->last_max_df = 0;
    for (; iter < m->max_iterations; ++iter) {
        /* mpc.h:275-283: df = H*controls + MM without forming H */
        B_vec(S, I, m->B, u, M);
        for (int i = 1; i < N; ++i) {
            mat_vec(S, m->A, M + (i - 1) * S, t1);
            B_vec(S, I, m->B, u + i * I, t2);
            for (int r = 0; r < S; ++r) M[i * S + r] = t1[r] + t2[r];
        }
        for (int i = 0; i < N; ++i) {
            memcpy(t1, M + i * S, sizeof(double) * S);
            diagm_vec(S, m->Q, t1, M + i * S);
        }
        for (int i = N - 2; i >= 0; --i) {
            matT_vec(S, m->A, M + (i + 1) * S, t1);
            for (int r = 0; r < S; ++r) M[i * S + r] = M[i * S + r] + t1[r];
        }
        for (int i = 0; i < N; ++i) {
            double bt[ORC_MAX_I];
            BT_vec(S, I, m->B, M + i * S, bt);
            for (int j = 0; j < I; ++j) {
                /* diagm(R)*controls[i] as dense product */
                double ru = (j == 0 ? m->R[0] : 0.0) * u[i * I];
                for (int k = 1; k < I; ++k) ru += (j == k ? m->R[k] : 0.0) * u[i * I + k];
                df[i * I + j] = (MM[i * I + j] + bt[j]) + ru;
            }
        }

        /* mpc.h:289-311: stopping test on the largest free gradient component */
        double max_df = 0; int max_t = 0, max_v = 0;
        for (int i = 0; i < N; ++i) for (int j = 0; j < I; ++j) {
            const double uij = u[i * I + j], d = df[i * I + j];
            if (!((uij <= m->lower[j] && d > 0) || (uij >= m->upper[j] && d < 0))) {
                if (fabs(d) > max_df) { max_df = fabs(d); max_t = i; max_v = j; }
            }
        }
        m->last_max_df = max_df;
        if (max_df < m->eps) break;

        if (iter < 50) {   /* mpc.h:319-335: greedy coordinate step */
            const double qd = m->Q_diag[max_t * I + max_v];
            if (qd == 0) continue;
            double* c = &u[max_t * I + max_v];
            *c = -(df[max_t * I + max_v] - qd * (*c)) / qd;
            *c = put_in_range(m->lower[max_v], m->upper[max_v], *c);
            if (iter + 1 == 50) memcpy(m->v, u, sizeof(double) * N * I);
        } else {           /* mpc.h:336-345: accelerated projected gradient */
            const double inv = 1.0 / lambda;
            const double beta = (sqrt(lambda) - 1) / (sqrt(lambda) + 1);
            for (int i = 0; i < N; ++i) for (int j = 0; j < I; ++j) {
                const int k = i * I + j;
                const double v_old = m->v[k];
                m->v[k] = clamp_m(u[k] - inv * df[k], m->lower[j], m->upper[j]);
                u[k] = clamp_m(m->v[k] + beta * (m->v[k] - v_old), m->lower[j], m->upper[j]);
            }
        }
    }
    m->last_iters = iter;
}

/* mpc.h:216-240 */
void orc_mpc_step(orc_mpc* m, const double* x0, double* u0)
{
    const int S = m->S, I = m->I, N = m->N;
    for (int i = 1; i < N; ++i) memcpy(m->controls + (i - 1) * I, m->controls + i * I, sizeof(double) * I);
    solve_linear_mpc(m, x0);
    for (int i = 1; i < N; ++i) memcpy(m->target + (i - 1) * S, m->target + i * S, sizeof(double) * S);
    memcpy(u0, m->controls, sizeof(double) * I);
}

/* ---- batched cold runner (harness, mirrors dlibref_batch_cold) ------------------------------------ */
typedef struct {
    int S, I, N, per_model; const double *A, *B, *C, *Q, *R, *lo, *hi;
    long a, b; const double *x0, *target; double eps; unsigned long max_iter;
    double* u0; double* controls; unsigned int* iters;
} slice_t;

static void* slice_run(void* p)
{
    slice_t* s = (slice_t*)p;
    orc_mpc* m = (orc_mpc*)malloc(sizeof(orc_mpc));
    double zeros[ORC_MAX_S] = {0};
    for (long k = s->a; k < s->b; ++k) {
        const long q = s->per_model ? k : 0;
        orc_mpc_init(m, s->S, s->I, s->N, s->A + q * s->S * s->S, s->B + q * s->S * s->I, s->C + q * s->S,
                     s->Q + q * s->S, s->R + q * s->I, s->lo + q * s->I, s->hi + q * s->I);
        m->eps = s->eps; m->max_iterations = s->max_iter;
        orc_mpc_set_target(m, s->target ? s->target + k * s->S : zeros, -1);
        orc_mpc_step(m, s->x0 + k * s->S, s->u0 + k * s->I);
        if (s->controls) memcpy(s->controls + k * (long)s->N * s->I, m->controls, sizeof(double) * s->N * s->I);
        if (s->iters) s->iters[k] = (unsigned int)m->last_iters;
    }
    free(m);
    return 0;
}

double orc_batch_cold(int S, int I, int N, int per_instance_model, const double* A, const double* B,
                      const double* C, const double* Q, const double* R, const double* lo, const double* hi,
                      long count, const double* x0, const double* target, double eps, unsigned long max_iter,
                      double* u0, double* controls, unsigned int* iters, int nthreads)
{
    if (S > ORC_MAX_S || I > ORC_MAX_I || N > ORC_MAX_N) return -1.0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 1024) nthreads = 1024;
    slice_t* sl = (slice_t*)malloc(sizeof(slice_t) * nthreads);
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * nthreads);
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int t = 0; t < nthreads; ++t) {
        slice_t s = {S, I, N, per_instance_model, A, B, C, Q, R, lo, hi,
                     count * t / nthreads, count * (t + 1) / nthreads, x0, target, eps, max_iter, u0, controls, iters};
        sl[t] = s;
        pthread_create(&th[t], 0, slice_run, &sl[t]);
    }
    for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    free(sl); free(th);
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}
