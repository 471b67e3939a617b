/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.  See pid_oracle.h (parity unpinned: sensor_utils / lms_math are absent). */
#include "pid_oracle.h"
#include <math.h>

void orc_pid_reset(orc_pid* c) { c->integral = 0.0f; c->prev = 0.0f; c->first = 1; }

float orc_pid_step(orc_pid* c, float error)
{
    c->integral = c->integral + error * c->dt;
    const float deriv = c->first ? 0.0f : (error - c->prev) / c->dt;
    c->prev = error; c->first = 0;
    return c->kp * error + c->ki * c->integral + c->kd * deriv;
}

float orc_lookup(const float* vx, const float* vy, int n, float x)
{
    if (n <= 0) return x;
    if (x <= vx[0]) return vy[0];
    for (int i = 1; i < n; ++i)
        if (x <= vx[i]) return vy[i - 1] + (vy[i] - vy[i - 1]) * (x - vx[i - 1]) / (vx[i] - vx[i - 1]);
    return vy[n - 1];
}

/* src/trajectory_point_follower.cpp:445-473 */
void orc_crossing_slowdown(const float* px, const float* py, const float* vel, long stride, int P,
                           const orc_crossing_config* cfg, orc_pid* pid, float* point_velocity)
{
    float minVelocity = cfg->max_velocity_crossing;                              /* :446 */
    for (int k = 0; k < P; ++k) {                                                /* :447 */
        if (vel[k * stride] == 0) {                                              /* :448 */
            const float x = px[k * stride], y = py[k * stride];
            const int s = (0.0f < x) - (x < 0.0f);
            const float distanceToStop = (float)s * sqrtf(x * x + y * y) - cfg->stopping_distance;   /* :449 */
            if (distanceToStop < cfg->distance_to_stop) {                        /* :450 */
                const float cap = cfg->max_velocity_crossing;                    /* :452 */
                float velocity = orc_pid_step(pid, distanceToStop);              /* :453 */
                if (isnan(velocity) || velocity >= cap) velocity = cap;          /* :454-456 */
                if (distanceToStop <= cfg->crossing_safety_zone || velocity < 0) velocity = 0.0f;   /* :457-459 */
                if (velocity < minVelocity) minVelocity = velocity;              /* :460-462 */
                *point_velocity = minVelocity;                                   /* :464 */
                if (minVelocity == 0) break;                                     /* :466-468 */
            } else {
                orc_pid_reset(pid);                                              /* :469-471 */
            }
        }
    }
}

void orc_crossing_batch(long count, int P, const int32_t* npoints, const float* px, const float* py, const float* vel,
                        long stride_instance, long stride_point, const orc_crossing_config* cfg, float kp, float ki,
                        float kd, float dt, float* integral, float* prev, uint8_t* first, float* point_velocity)
{
    for (long b = 0; b < count; ++b) {
        int n = npoints ? npoints[b] : P;
        if (n > P) n = P;
        if (n <= 0) continue;                                                    /* :404-407 returns before the clamp */
        orc_pid c = {kp, ki, kd, dt, integral[b], prev[b], first[b]};
        const long o = b * stride_instance;
        orc_crossing_slowdown(px + o, py + o, vel + o, stride_point, n, cfg, &c, &point_velocity[b]);
        integral[b] = c.integral; prev[b] = c.prev; first[b] = (uint8_t)c.first;
    }
}

/* src/trajectory_point_follower.cpp:214-225 */
void orc_pid_branch_batch(long count, const float* pos_x, const float* pos_y, const float* dir_x, const float* dir_y,
                          const float gf[4], const float gr[4], float* f_integral, float* f_prev, uint8_t* f_first,
                          float* r_integral, float* r_prev, uint8_t* r_first, double* steering_front,
                          double* steering_rear, float* target_distance)
{
    for (long b = 0; b < count; ++b) {
        orc_pid f = {gf[0], gf[1], gf[2], gf[3], f_integral[b], f_prev[b], f_first[b]};
        orc_pid r = {gr[0], gr[1], gr[2], gr[3], r_integral[b], r_prev[b], r_first[b]};
        steering_front[b] = (double)orc_pid_step(&f, pos_y[b]);                  /* :219 */
        steering_rear[b] = (double)orc_pid_step(&r, atan2f(dir_y[b], dir_x[b])); /* :220 directory.angle() */
        target_distance[b] = sqrtf(pos_x[b] * pos_x[b] + pos_y[b] * pos_y[b]);   /* :224 */
        f_integral[b] = f.integral; f_prev[b] = f.prev; f_first[b] = (uint8_t)f.first;
        r_integral[b] = r.integral; r_prev[b] = r.prev; r_first[b] = (uint8_t)r.first;
    }
}
