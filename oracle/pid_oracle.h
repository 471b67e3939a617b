/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement of the parts of TrajectoryPointController that sit around the MPC solve and carry per-vehicle
 * state or table look-ups (all of /root/reference/src/trajectory_point_follower.cpp):
 *   :66, :216, :323   lms::math::LookupTable<float, ASC>::linearSearch
 *   :445-473          the "stop at crossing" velocity clamp at the end of getTrajectoryPoint (stateful PID slowDownCar)
 *   :214-225          the non-MPC branch: steering = pid(position.y), pid(directory.angle())
 *
 * PARITY UNPINNED.  sensor_utils::PID, lms::math::LookupTable, lms::math::sgn and vertex2f::length/angle live in
 * libraries that are absent from /root/reference and from this machine (sensor_utils, lms_math — no
 * version pinned, README.md:9-13), and the reference holds no test or fixture for these branches.  The semantics
 * adopted are the ones of the headless shim the reference source files are compiled against for the module golden
 * (trajectory_controller_b200/cpp/shim/sensor_utils/pid.h, shim/lms/math/lookup_table.h):
 *   PID (float):   integral += e*dt;  deriv = first ? 0 : (e - prev)/dt;  prev = e;  out = Kp*e + Ki*integral + Kd*deriv
 *   LookupTable:   n == 0 -> x;  x <= vx[0] -> vy[0];  first i with x <= vx[i] ->
 *                  vy[i-1] + (vy[i]-vy[i-1])*(x-vx[i-1])/(vx[i]-vx[i-1]);  else vy[n-1]        (float, no FMA)
 *   sgn(x) = (0 < x) - (x < 0);   length() = sqrtf(x*x + y*y);   angle() = atan2f(y, x)
 * The control flow (which points are visited, the clamps and their order, reset(), break) is pinned by the reference
 * source.  "Parity" for these rows therefore means: GPU == this restatement (and the module golden, which runs the
 * reference's own source over the same shim, agrees with it for the single-vehicle case).
 */
#ifndef PID_ORACLE_H
#define PID_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_pid {
    float kp, ki, kd, dt;       /* set(Kp, Ki, Kd, dt)   src/...:296-298 */
    float integral, prev;
    int first;
} orc_pid;
void orc_pid_reset(orc_pid* c);
float orc_pid_step(orc_pid* c, float error);

float orc_lookup(const float* vx, const float* vy, int n, float x);

typedef struct orc_crossing_config {
    float max_velocity_crossing;   /* "maxVelocityCrossing", 1.0   :446,:452 */
    float stopping_distance;       /* "stoppingDistance",   0.35   :449 */
    float distance_to_stop;        /* "distanceToStop",     1      :450 */
    float crossing_safety_zone;    /* "crossingSaftyZone",  0.05   :457 */
} orc_crossing_config;

/* :445-473 over one polyline (element k of a component at base[k*stride]); *point_velocity is
 * trajectoryPoint.velocity (in/out), `pid` is the vehicle's slowDownCar. */
void orc_crossing_slowdown(const float* px, const float* py, const float* vel, long stride, int P,
                           const orc_crossing_config* cfg, orc_pid* pid, float* point_velocity);

/* Batch: per-vehicle PID states as three arrays (integral, prev, first), gains shared. */
void orc_crossing_batch(long count, int P, const int32_t* npoints, const float* px, const float* py, const float* vel,
                        long stride_instance, long stride_point, const orc_crossing_config* cfg, float kp, float ki,
                        float kd, float dt, float* integral, float* prev, uint8_t* first, float* point_velocity);

/* :214-225 for a batch, given the extracted points: steering_front = pidFront(pos.y), steering_rear =
 * pidRear(directory.angle()), target_distance = position.length(). */
void orc_pid_branch_batch(long count, const float* pos_x, const float* pos_y, const float* dir_x, const float* dir_y,
                          const float gains_front[4], const float gains_rear[4], float* f_integral, float* f_prev,
                          uint8_t* f_first, float* r_integral, float* r_prev, uint8_t* r_first, double* steering_front,
                          double* steering_rear, float* target_distance);
#ifdef __cplusplus
}
#endif
#endif
