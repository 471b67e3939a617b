/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement of TrajectoryPointController::getTrajectoryPoint's arc-length scan and
 * interpolation, /root/reference/src/trajectory_point_follower.cpp:392-443, over a POD polyline.
 *
 * PARITY UNPINNED for the float geometry: lms::math::vertex2f (distance / normalize / operator* /
 * operator+,-) lives in the un-vendored `lms_math` library (no version pinned: README.md:9-13,
 * CMakeLists.txt:19), which is absent from /root/reference and from this machine, and the
 * reference holds no test or fixture for this function.  The float semantics adopted here are the
 * ones reconstructed in SURVEY.md Appendix B:
 *     length()   = sqrtf(x*x + y*y)          distance(o) = (*this - o).length()
 *     normalize()= (x/len, y/len)            operator*(f) = (x*f, y*f)      + and - component-wise
 * all in float, no FMA contraction, sequential float accumulation of the arc length, strict `>`.
 * "Bit-exact index selection" is therefore defined as: GPU == this restatement.
 * The control flow (loop bounds, break, fallback, default point) is pinned by the reference source.
 */
#ifndef EXTRACT_ORACLE_H
#define EXTRACT_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* flags: how the point was produced */
#define ORC_PT_INTERPOLATED 0   /* :428-437  cumulative distance exceeded d at segment (idx-1, idx) */
#define ORC_PT_LAST         1   /* :439-442  never exceeded: last point returned, idx = P-1 */
#define ORC_PT_EMPTY        2   /* :394-407  empty trajectory: default point (d,0),(1,0),vel 0, idx = -1 */

typedef struct orc_point {
    float pos_x, pos_y, dir_x, dir_y, vel;
    int32_t idx;
    uint8_t flag;
} orc_point;

/* One polyline of P points given as five strided float arrays (element k at base[k*stride]). */
void orc_extract_point(const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                       long stride, int P, float d, orc_point* out);

/* Batch: instance b reads element k of component c at c[b*stride_instance + k*stride_point];
 * npoints may be NULL (all P) or hold per-instance counts in [0, P].  Outputs are arrays of `count`. */
void orc_extract_batch(long count, int P, const int32_t* npoints, const float* px, const float* py,
                       const float* dx, const float* dy, const float* vel, long stride_instance,
                       long stride_point, const float* dist, int32_t* idx, uint8_t* flag, float* out_px,
                       float* out_py, float* out_dx, float* out_dy, float* out_vel);

/* Look-ahead points for `steps` horizon steps of one polyline ("sample the trajectory at N
 * lookahead distances and set_target(val, t) per step"; dlib's per-step targets, mpc.h:142-170).  Step t asks
 * getTrajectoryPoint (:392-443) for the distance d_t = d0 + (float)t * step — one float multiply and one float add,
 * each rounded on its own.  DEFINITION: out[t] == orc_extract_point(..., d_t).  orc_extract_horizon restates that as
 * the single sequential scan the device runs (needs step >= 0 so that the d_t ascend; the arc length `cur` is the
 * same float sum for every t, so the first segment with cur > d_t is the same one); tests/test_oracle.py checks the
 * two against each other. */
void orc_extract_horizon(const float* px, const float* py, const float* dx, const float* dy, const float* vel,
                         long stride, int P, float d0, float step, int steps, orc_point* out /* [steps] */);
#ifdef __cplusplus
}
#endif
#endif
