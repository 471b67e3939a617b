/* HEADLESS SHIM of the `control-systems` submodule's andromeda.h (commit 586f4e0c, NOT in the reference tree:
 * .SUBMODULES.json "missing").  Only the names the module's mikMPC branch uses
 * (src/trajectory_point_follower.cpp:129-206); the stub solver returns zeros.  Exists so that the reference's own
 * module source can be compiled for the drop-in check; the mikMPC branch itself is out of scope. */
#ifndef ANDROMEDA_SHIM_H
#define ANDROMEDA_SHIM_H
#define HORIZON_LEN 10
#define CHAIN_NUM_NODES 11
#define NUM_STATES 4
#define NUM_INPUTS 2
void call_andromeda(double* currentCarState, double* q_diag, double* r_diag, double* p_diag, double* nodes_x,
                    double* nodes_y, double link_length, double* nodes_vMin, double* nodes_vMax,
                    double max_lateral_acc, double max_num_iter, double alpha, double beta_1, double beta_2,
                    double u_1_lb, double u_1_ub, double u_2_lb, double u_2_ub, double* v_star, double* u_1_star,
                    double* u_2_star);
#endif
