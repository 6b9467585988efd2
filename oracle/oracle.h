/*
 * oracle.h -- C ABI of the CPU oracle (TEST INFRASTRUCTURE, not product code).
 *
 * A plain, dependency-free CPU restatement of the reference's hot path
 * (MotionPlanner::solve_trajectory -> polympc SQP + boxADMM + pinocchio RNEA;
 * /root/reference/mpc_solver/motionPlanner.cpp:177-208).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library.
 *
 * PARITY UNPINNED: polympc, pinocchio, ruckig and Eigen are absent from /root/reference and from
 * this image, so the SQP/ADMM part restates the published algorithms from
 * SURVEY.md Appendix A.  What IS pinned against reference-embedded numbers (tests/golden/):
 * forward kinematics, the world-aligned Jacobian and 402 RNEA samples.
 *
 * Variable layout of one OCP (polympc's, motionPlanner.cpp:162-165):
 *     z = [ x_0 .. x_{N-1} (14 each: q(7), v(7)) | u_0 .. u_{N-1} (7 each) | p ]      N = PO*NS+1
 * index 0 is t=0.  Dual layout: [ eq (14*(N-1)) | ineq (8*N) | box (nvar) ].
 */
#ifndef MPCB_ORACLE_H
#define MPCB_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int poly_order;          /* robot_ocp.hpp:31  (3) */
    int num_seg;             /* robot_ocp.hpp:32  (6) */
    int sqp_max_iter;        /* motionPlanner.cpp:15 (2) */
    int qp_max_iter;         /* motionPlanner.cpp:16 (700) */
    int line_search_max_iter;/* motionPlanner.cpp:17 (10 -> 9 trials, polympc_redef.hpp:101) */
    int qp_check_termination;/* boxADMM default 25 [MEM] */
    int qp_adaptive_rho;     /* boxADMM default off */
    int qp_adaptive_rho_interval; /* 25 */
    double qp_eps_abs;       /* motionPlanner.cpp:20 (1e-3) */
    double qp_eps_rel;       /* motionPlanner.cpp:19 (1e-3) */
    double qp_rho;           /* 0.1 */
    double qp_sigma;         /* 1e-6 */
    double qp_alpha;         /* 1.0 */
    double qp_adaptive_rho_tolerance; /* 5 */
    double sqp_eps_prim;     /* 1e-3 */
    double sqp_eps_dual;     /* 1e-3 */
    double sqp_tau;          /* 0.5 */
    double sqp_eta;          /* 0.25 */
    double hess_reg;         /* polympc_redef.hpp:68 (0.001, sparse variant) */
} orc_settings;

typedef struct {
    double qmin[7], qmax[7], vmax[7], amax[7], taumax[7], jmax[7]; /* pandaWrapper.hpp:29-34 */
    double margins[5];       /* position, velocity, acceleration, torque, jerk (motionPlanner.cpp:56-62) */
    double min_height;       /* pandaWrapper.hpp:40 */
    double final_eps;        /* motionPlanner.hpp:44 */
    double p_lb, p_ub;       /* motionPlanner.cpp:78-79 */
} orc_limits;

typedef struct {
    int status;              /* 1 = SOLVED, 2 = MAX_ITER_EXCEEDED */
    int sqp_iters;
    int qp_iters_total;
    int qp_status[4];        /* per SQP iteration (first 4) */
    int qp_iters[4];
    double alpha[4];
    int ls_trials[4];
    double primal_step_norm, dual_step_norm, max_violation, l1_violation, cost;
} orc_stats;

typedef struct {
    int flags[4];            /* jerk, linear vel, angular vel, collision; 1 = ok (benchmark.cpp:181-189) */
    double mins[28], maxs[28];/* q, v, a, tau extrema over the 201 samples (benchmark.cpp:171-172) */
    double final_err[14];    /* last sample - target (benchmark.cpp:177-178) */
} orc_account;

void orc_default_settings(orc_settings* s);
void orc_default_limits(orc_limits* l);

/* ---- rigid-body pieces (pinocchio restatement) ---- */
void orc_fk(const double* q, double* R_out /*7x9*/, double* p_out /*7x3*/);
void orc_rnea(const double* q, const double* v, const double* a, double* tau);
void orc_rnea_zero_gravity(const double* q, const double* v, const double* a, double* tau);
void orc_rnea_derivatives(const double* q, const double* v, const double* a, double* tau,
                          double* dq /*7x7 row-major*/, double* dv, double* M /*symmetric*/);
void orc_crba(const double* q, double* M /*7x7 symmetric*/);
void orc_tool_height(const double* q, double* z, double* Jz /*7, nullable*/);
void orc_frame_jacobian_world_aligned(const double* q, const double* offset /*3, in joint-7 frame*/, double* J /*6x7*/);
void orc_node_constraints(const double* q, const double* v, const double* a, double* g /*8*/, double* Jg /*8x22, nullable*/);

/* ---- transcription ---- */
int  orc_num_nodes(const orc_settings* s);
int  orc_num_var(const orc_settings* s);
int  orc_num_eq(const orc_settings* s);
int  orc_num_ineq(const orc_settings* s);
void orc_time_nodes(const orc_settings* s, double* t /*N*/);
void orc_diff_matrix(const orc_settings* s, double* D /*(PO+1)^2 row-major, ascending nodes*/);
void orc_bounds(const orc_settings* s, const orc_limits* lim, const double* x0, const double* xT,
                double* lbx, double* ubx /*nvar*/, double* lbg, double* ubg /*8*/);
void orc_eval_constraints(const orc_settings* s, const double* z, double* c_eq, double* g);
void orc_linearise(const orc_settings* s, const double* z, const double* lam,
                   double* c_eq, double* g, double* Jg /*N x 8 x 22*/, double* Hdiag /*nvar*/, double* Harrow /*nvar*/);
/* dense export of the QP at (z, lam): A is (neq+nineq) x nvar row-major, H nvar x nvar */
void orc_qp_dense(const orc_settings* s, const double* z, const double* lam, double* A, double* H);

/* ---- QP (boxADMM restatement).  l/u: (neq+nineq) rows then nvar box rows ---- */
int orc_qp_admm(const orc_settings* s, const double* z_lin, const double* Jg, const double* Hdiag, const double* Harrow,
                const double* h, const double* lA, const double* uA, const double* lb, const double* ub,
                double* x_out /*nvar*/, double* y_out /*ndual*/, int* iters_out, double* res_out /*r_p, r_d, eps_p, eps_d*/);

/* ---- whole solve ---- */
int orc_solve(const orc_settings* s, const orc_limits* lim, const double* x0, const double* xT,
              const double* guess_x, const double* guess_u, double guess_T, const double* guess_lam /*nullable*/,
              double* sol_x, double* sol_u, double* sol_T, double* sol_lam /*nullable*/, orc_stats* stats);
/* batch over independent OCPs with a simple thread pool (timed CPU baseline); arrays are problem-major */
int orc_solve_batch(const orc_settings* s, const orc_limits* lim, int n, int nthreads,
                    const double* x0, const double* xT, const double* guess_x, const double* guess_u, const double* guess_T,
                    double* sol_x, double* sol_u, double* sol_T, orc_stats* stats);

/* ---- read-out + accounting (motionPlanner.hpp:99-116, benchmark.cpp:118-194) ---- */
void orc_sample_solution(const orc_settings* s, const double* sol_x, const double* sol_u, double T, int npts,
                         double* traj /* (npts+1) x 29: t, q, v, a, tau */);
void orc_account_trajectory(const orc_limits* lim, const double* traj, int npts, const double* target /*14*/, orc_account* acc);
void orc_interp_state(const orc_settings* s, const double* sol_x, const double* sol_u, double t01, double* q, double* v, double* a);

#ifdef __cplusplus
}
#endif
#endif
