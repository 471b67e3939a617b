/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See extract_oracle.h (parity unpinned for vertex2f math). */
#include "extract_oracle.h"
#include <math.h>

/* src/trajectory_point_follower.cpp:392-443 */
void orc_extract_point(const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                       long stride, int P, float d, orc_point* out)
{
    /* :394-403 default: idle forward */
    out->pos_x = d; out->pos_y = 0.0f; out->dir_x = 1.0f; out->dir_y = 0.0f; out->vel = 0.0f;
    out->idx = -1; out->flag = ORC_PT_EMPTY;
    if (P <= 0) return;                                   /* :404-407 */

    float cur = 0.0f;                                     /* :423 */
    for (int i = 1; i < P; ++i) {                         /* :424 */
        const float bx = px[(i - 1) * stride], by = py[(i - 1) * stride];
        const float tx = px[i * stride], ty = py[i * stride];
        /* bot.position.distance(top.position) = (bot - top).length()   :427 */
        const float ex = bx - tx, ey = by - ty;
        const float len = sqrtf(ex * ex + ey * ey);
        cur = cur + len;
        if (cur > d) {                                    /* :428 strict */
            const float delta = cur - d;                  /* :431 */
            /* (bot - top).normalize() * delta            :432 */
            const float nx = ex / len, ny = ey / len;
            const float ax = nx * delta, ay = ny * delta;
            out->pos_x = tx + ax; out->pos_y = ty + ay;   /* :433-434 */
            out->dir_x = dx[i * stride]; out->dir_y = dy[i * stride]; out->vel = vel[i * stride];
            out->idx = i; out->flag = ORC_PT_INTERPOLATED;
            return;                                       /* :435-436 */
        }
    }
    /* :439-442 */
    const long l = (long)(P - 1) * stride;
    out->pos_x = px[l]; out->pos_y = py[l]; out->dir_x = dx[l]; out->dir_y = dy[l]; out->vel = vel[l];
    out->idx = P - 1; out->flag = ORC_PT_LAST;
}

void orc_extract_batch(long count, int P, const int32_t* npoints, const float* px, const float* py,
                       const float* dx, const float* dy, const float* vel, long stride_instance,
                       long stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* out_px,
                       float* out_py, float* out_dx, float* out_dy, float* out_vel)
{
    for (long b = 0; b < count; ++b) {
        orc_point p;
        const long o = b * stride_instance;
        int n = npoints ? npoints[b] : P;
        if (n > P) n = P;
        orc_extract_point(px + o, py + o, dx + o, dy + o, vel + o, stride_point, n, dist[b], &p);
        idx[b] = p.idx; flag[b] = p.flag;
        out_px[b] = p.pos_x; out_py[b] = p.pos_y; out_dx[b] = p.dir_x; out_dy[b] = p.dir_y; out_vel[b] = p.vel;
    }
}

void orc_extract_horizon(const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                         long stride, int P, float d0, float step, int steps, orc_point* out)
{
    int t = 0;                                            /* next horizon step without a point */
    if (P <= 0) {                                         /* :394-407 for every step */
        for (; t < steps; ++t) {
            const float d = d0 + (float)t * step;
            out[t].pos_x = d; out[t].pos_y = 0.0f; out[t].dir_x = 1.0f; out[t].dir_y = 0.0f; out[t].vel = 0.0f;
            out[t].idx = -1; out[t].flag = ORC_PT_EMPTY;
        }
        return;
    }
    float cur = 0.0f;                                     /* :423 */
    for (int i = 1; i < P && t < steps; ++i) {            /* :424 */
        const float bx = px[(i - 1) * stride], by = py[(i - 1) * stride];
        const float tx = px[i * stride], ty = py[i * stride];
        const float ex = bx - tx, ey = by - ty;
        const float len = sqrtf(ex * ex + ey * ey);       /* :427 */
        cur = cur + len;
        for (; t < steps; ++t) {                          /* every step whose distance this segment is the first to exceed */
            const float d = d0 + (float)t * step;
            if (!(cur > d)) break;                        /* :428 strict */
            const float delta = cur - d;                  /* :431 */
            const float nx = ex / len, ny = ey / len;     /* :432 */
            const float ax = nx * delta, ay = ny * delta;
            out[t].pos_x = tx + ax; out[t].pos_y = ty + ay;   /* :433-434 */
            out[t].dir_x = dx[i * stride]; out[t].dir_y = dy[i * stride]; out[t].vel = vel[i * stride];
            out[t].idx = i; out[t].flag = ORC_PT_INTERPOLATED;
        }
    }
    const long l = (long)(P - 1) * stride;                /* :439-442 for the steps the polyline is too short for */
    for (; t < steps; ++t) {
        out[t].pos_x = px[l]; out[t].pos_y = py[l]; out[t].dir_x = dx[l]; out[t].dir_y = dy[l]; out[t].vel = vel[l];
        out[t].idx = P - 1; out[t].flag = ORC_PT_LAST;
    }
}
