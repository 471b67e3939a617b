/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * CPU restatement of the closed control loop the engine runs on the device (include/bmpc.h "closed loop"):
 * per tick and vehicle   world path -> vehicle-frame polyline -> getTrajectoryPoint
 * (/root/reference/src/trajectory_point_follower.cpp:392-443, via orc_extract_point) -> target
 * (src/...:84-85,368-371) -> one PERSISTENT dlib::mpc<4,1,N> per vehicle (orc_mpc_step = mpc.h:216-240)
 * -> harness plant.  The reference has no plant and no path sampler; those are harness code whose arithmetic is
 * specified in include/bmpc.h and restated here independently (compile with -ffp-contract=off).
 * Parity status: the solver part is pinned (mpc_oracle.h); extraction float geometry is unpinned
 * (extract_oracle.h); the harness has no reference counterpart by construction.
 */
#ifndef CLOSED_LOOP_ORACLE_H
#define CLOSED_LOOP_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
double orc_harness_atan(double t);
/* model: the 7 dlib::mpc constructor arrays for S = 4, I = 1.  pose [count][4] = X, Y, cos, sin and act [count][2]
 * are updated in place; u0/target[count][2]/idx/iters_total/iters_max receive the last tick's values / totals.
 * Returns 0, or -1 on bad arguments. */
int orc_closed_loop(long count, int N, const double* A, const double* B, const double* C, const double* Q,
                    const double* R, const double* lo, const double* hi, double eps, unsigned long max_iter,
                    double v, double T, double l, double lookahead, int P, int npath, const double* path_x,
                    const double* path_y, const double* path_tx, const double* path_ty, int ticks, double* pose,
                    double* act, double* u0, double* target, int32_t* idx, uint64_t* iters_total,
                    uint32_t* iters_max, int nthreads);
/* The same loop with a target per horizon step (horizon_targets = 1): every tick the polyline is sampled at
 * d_t = h_d0 + t * h_step (orc_extract_horizon) and target[t] = (pos.y, atan(dir.y/dir.x), 0, 0) is set with
 * set_target(val, t) (mpc.h:142-155) before the solve; `target` / `idx` report step 0 of the last tick. */
int orc_closed_loop_h(long count, int N, const double* A, const double* B, const double* C, const double* Q,
                      const double* R, const double* lo, const double* hi, double eps, unsigned long max_iter,
                      double v, double T, double l, double lookahead, int P, int npath, const double* path_x,
                      const double* path_y, const double* path_tx, const double* path_ty, int ticks, double* pose,
                      double* act, double* u0, double* target, int32_t* idx, uint64_t* iters_total,
                      uint32_t* iters_max, int nthreads, int horizon_targets, float h_d0, float h_step);
#ifdef __cplusplus
}
#endif
#endif
