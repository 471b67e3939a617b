/* TEST INFRASTRUCTURE — see closed_loop_oracle.h. */
#include "closed_loop_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>

#include "extract_oracle.h"
#include "mpc_oracle.h"

double orc_harness_atan(double t)
{
    const double t1 = t / (1.0 + sqrt(1.0 + t * t));
    const double t2 = t1 / (1.0 + sqrt(1.0 + t1 * t1));
    const double z = t2 * t2;
    double p = 1.0 / 23.0;
    for (int k = 21; k >= 1; k -= 2) p = 1.0 / (double)k - z * p;
    return 4.0 * (t2 * p);
}

typedef struct {
    long first, last;
    int N, P, npath, ticks;
    const double *A, *B, *C, *Q, *R, *lo, *hi, *path_x, *path_y, *path_tx, *path_ty;
    double eps, v, T, l, lookahead;
    unsigned long max_iter;
    double *pose, *act, *u0, *target;
    int32_t* idx;
    uint64_t* iters_total;
    uint32_t* iters_max;
    int horizon_targets;          /* 1: a target per horizon step (orc_extract_horizon), 0: one target for all steps */
    float h_d0, h_step;
} job_t;

static void* worker(void* arg)
{
    job_t* j = (job_t*)arg;
    orc_mpc* m = (orc_mpc*)malloc(sizeof(orc_mpc));
    float* buf = (float*)malloc(sizeof(float) * 5 * (size_t)j->P);
    float *px = buf, *py = buf + j->P, *dx = buf + 2 * j->P, *dy = buf + 3 * j->P, *vl = buf + 4 * j->P;
    const double Tv = j->T * j->v, Tv_l = Tv / j->l;
    const float d = (float)j->lookahead;
    for (long b = j->first; b < j->last; ++b) {
        orc_mpc_init(m, 4, 1, j->N, j->A, j->B, j->C, j->Q, j->R, j->lo, j->hi);   /* one persistent controller per vehicle */
        m->eps = j->eps; m->max_iterations = j->max_iter;
        double X = j->pose[4 * b], Y = j->pose[4 * b + 1], c = j->pose[4 * b + 2], s = j->pose[4 * b + 3];
        double de = j->act[2 * b], dd = j->act[2 * b + 1];
        uint64_t tot = 0; uint32_t mx = 0;
        for (int t = 0; t < j->ticks; ++t) {
            /* world path -> vehicle-frame polyline (what the TRAJECTORY channel would carry) */
            int k0 = 0;
            while (k0 < j->npath && !(j->path_x[k0] > X)) ++k0;
            if (k0 > j->npath - j->P) k0 = j->npath - j->P;
            if (k0 < 0) k0 = 0;
            const int n = j->P < j->npath - k0 ? j->P : j->npath - k0;
            for (int i = 0; i < n; ++i) {
                const double ex = j->path_x[k0 + i] - X, ey = j->path_y[k0 + i] - Y;
                const double tx = j->path_tx[k0 + i], ty = j->path_ty[k0 + i];
                px[i] = (float)(c * ex + s * ey); py[i] = (float)(c * ey - s * ex);
                dx[i] = (float)(c * tx + s * ty); dy[i] = (float)(c * ty - s * tx);
                vl[i] = (float)j->v;
            }
            orc_point pt;
            double y_soll = 0.0, phi = 0.0;
            if (!j->horizon_targets) {
                orc_extract_point(px, py, dx, dy, vl, 1, n, d, &pt);
                y_soll = (double)pt.pos_y;
                if (pt.dir_x > 0.0f) phi = orc_harness_atan((double)pt.dir_y / (double)pt.dir_x);
                else phi = pt.dir_y >= 0.0f ? 1.5707963267948966 : -1.5707963267948966;
                const double tgt[4] = {y_soll, phi, 0.0, 0.0};
                orc_mpc_set_target(m, tgt, -1);
            } else {
                /* set_target(val, t) per horizon step (mpc.h:142-155) from the look-ahead points at
                 * d_t = h_d0 + t * h_step; the reported target / idx are those of step 0 */
                orc_point pts[ORC_MAX_N];
                orc_extract_horizon(px, py, dx, dy, vl, 1, n, j->h_d0, j->h_step, j->N, pts);
                for (int h = j->N - 1; h >= 0; --h) {
                    const double ys = (double)pts[h].pos_y;
                    double ph;
                    if (pts[h].dir_x > 0.0f) ph = orc_harness_atan((double)pts[h].dir_y / (double)pts[h].dir_x);
                    else ph = pts[h].dir_y >= 0.0f ? 1.5707963267948966 : -1.5707963267948966;
                    const double tgt[4] = {ys, ph, 0.0, 0.0};
                    orc_mpc_set_target(m, tgt, h);
                    y_soll = ys; phi = ph;
                }
                pt = pts[0];
            }
            const double x0[4] = {0.0, 0.0, de, dd};
            double u;
            orc_mpc_step(m, x0, &u);
            tot += m->last_iters; if (m->last_iters > mx) mx = (uint32_t)m->last_iters;
            /* harness plant (old c, s, delta on the right-hand sides) */
            const double de2 = (j->A[10] * de + j->A[11] * dd) + j->B[2] * u;
            const double dd2 = (j->A[14] * de + j->A[15] * dd) + j->B[3] * u;
            const double dpsi = Tv_l * de;
            const double c1 = c - s * dpsi, s1 = s + c * dpsi;
            const double nrm = sqrt(c1 * c1 + s1 * s1);
            X = X + Tv * c; Y = Y + Tv * s;
            c = c1 / nrm; s = s1 / nrm;
            de = de2; dd = dd2;
            if (t == j->ticks - 1) {
                j->u0[b] = u; j->target[2 * b] = y_soll; j->target[2 * b + 1] = phi; j->idx[b] = pt.idx;
            }
        }
        j->pose[4 * b] = X; j->pose[4 * b + 1] = Y; j->pose[4 * b + 2] = c; j->pose[4 * b + 3] = s;
        j->act[2 * b] = de; j->act[2 * b + 1] = dd;
        j->iters_total[b] = tot; j->iters_max[b] = mx;
    }
    free(buf); free(m);
    return 0;
}

int orc_closed_loop(long count, int N, const double* A, const double* B, const double* C, const double* Q,
                    const double* R, const double* lo, const double* hi, double eps, unsigned long max_iter,
                    double v, double T, double l, double lookahead, int P, int npath, const double* path_x,
                    const double* path_y, const double* path_tx, const double* path_ty, int ticks, double* pose,
                    double* act, double* u0, double* target, int32_t* idx, uint64_t* iters_total,
                    uint32_t* iters_max, int nthreads)
{
    return orc_closed_loop_h(count, N, A, B, C, Q, R, lo, hi, eps, max_iter, v, T, l, lookahead, P, npath, path_x, path_y,
                             path_tx, path_ty, ticks, pose, act, u0, target, idx, iters_total, iters_max, nthreads, 0, 0.0f, 0.0f);
}

int orc_closed_loop_h(long count, int N, const double* A, const double* B, const double* C, const double* Q,
                      const double* R, const double* lo, const double* hi, double eps, unsigned long max_iter,
                      double v, double T, double l, double lookahead, int P, int npath, const double* path_x,
                      const double* path_y, const double* path_tx, const double* path_ty, int ticks, double* pose,
                      double* act, double* u0, double* target, int32_t* idx, uint64_t* iters_total,
                      uint32_t* iters_max, int nthreads, int horizon_targets, float h_d0, float h_step)
{
    if (count < 1 || N < 1 || N > ORC_MAX_N || P < 1 || npath < P || nthreads < 1) return -1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256]; job_t jobs[256];
    for (int k = 0; k < nthreads; ++k) {
        job_t j = {count * k / nthreads, count * (k + 1) / nthreads, N, P, npath, ticks, A, B, C, Q, R, lo, hi,
                   path_x, path_y, path_tx, path_ty, eps, v, T, l, lookahead, max_iter, pose, act, u0, target, idx,
                   iters_total, iters_max, horizon_targets, h_d0, h_step};
        jobs[k] = j;
        pthread_create(&th[k], 0, worker, &jobs[k]);
    }
    for (int k = 0; k < nthreads; ++k) pthread_join(th[k], 0);
    return 0;
}
