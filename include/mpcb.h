/*
 * mpcb.h -- C ABI of the B200 batched minimum-time Panda OCP solver (libmpcb.so).
 *
 * This is the drop-in boundary for the reference's hot path: everything MotionPlanner::solve_trajectory does
 * behind `mpc.solve()` (/root/reference/mpc_solver/motionPlanner.cpp:177-208) -- polympc's SQP + boxADMM over
 * minTime_ocp (mpc_solver/robot_ocp.hpp) with MySolver's regularisation and line search
 * (mpc_solver/polympc_redef.hpp) and pinocchio's RNEA -- for a BATCH of independent start->target problems.
 * Plain C types only (no torch, no Eigen); every call returns 0 on success and a negative mpcb_error otherwise.
 * There is no CPU fallback: without a CUDA device mpcb_create fails with MPCB_ERR_NO_DEVICE.
 *
 * Variable layout of one OCP (polympc's; motionPlanner.cpp:162-165, motionPlanner.hpp:159-162), N = PO*NS+1 nodes:
 *     z      = [ x_0 .. x_{N-1} (14 each: q(7), v(7)) | u_0 .. u_{N-1} (7 each) | p ]        index 0 <-> t = 0
 *     lambda = [ eq (14*(N-1)) | ineq (8*N) | box (21*N+1, same order as z) ]
 * Batch arrays are problem-major and contiguous (problem i at offset i*stride).
 */
#ifndef MPCB_H
#define MPCB_H

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    MPCB_OK = 0,
    MPCB_ERR_NO_DEVICE = -1,     /* no CUDA device / driver: the product has no CPU path */
    MPCB_ERR_BAD_ARG = -2,
    MPCB_ERR_UNSUPPORTED = -3,   /* (poly_order, num_seg) not one of the instantiated shapes (3,6), (3,12), (5,6), (7,4), or a setting out of range */
    MPCB_ERR_CUDA = -4,          /* a CUDA runtime call failed; see mpcb_last_error() */
    MPCB_ERR_ALLOC = -5
} mpcb_error;

/* Solver settings.  Defaults = what MotionPlanner's constructor sets (motionPlanner.cpp:15-20) on top of
 * polympc's defaults. */
typedef struct {
    int poly_order;               /* robot_ocp.hpp:31  POLY_ORDER (3) */
    int num_seg;                  /* robot_ocp.hpp:32  NUM_SEG (6) */
    int sqp_max_iter;             /* motionPlanner.cpp:15 (2) */
    int qp_max_iter;              /* motionPlanner.cpp:16 (700) */
    int line_search_max_iter;     /* motionPlanner.cpp:17 (10 -> at most 9 trials, polympc_redef.hpp:101) */
    int qp_check_termination;     /* boxADMM: residual test every 25 iterations */
    int qp_adaptive_rho;          /* boxADMM adaptive rho: 0 = fixed rho (default); 1 = rho re-estimated at the residual checks every
                                   * qp_adaptive_rho_interval iterations, KKT re-factorised when it moved by more than the tolerance factor */
    int qp_adaptive_rho_interval;
    double qp_eps_abs;            /* motionPlanner.cpp:20 (1e-3) */
    double qp_eps_rel;            /* motionPlanner.cpp:19 (1e-3) */
    double qp_rho;                /* 0.1 */
    double qp_sigma;              /* 1e-6 */
    double qp_alpha;              /* 1.0 */
    double qp_adaptive_rho_tolerance;
    double sqp_eps_prim;          /* 1e-3 */
    double sqp_eps_dual;          /* 1e-3 */
    double sqp_tau;               /* line-search contraction 0.5 */
    double sqp_eta;               /* Armijo constant 0.25 */
    double hess_reg;              /* polympc_redef.hpp:68 (0.001) */
} mpcb_settings;

/* Robot limits (robot_utils/pandaWrapper.hpp:29-40) and the planner's margins / tolerances
 * (motionPlanner.cpp:56-100, motionPlanner.hpp:44). */
typedef struct {
    double qmin[7], qmax[7], vmax[7], amax[7], taumax[7], jmax[7];
    double margins[5];            /* position, velocity, acceleration, torque, jerk */
    double min_height;            /* tool height lower bound */
    double final_eps;             /* final state box half-width (1e-2) */
    double p_lb, p_ub;            /* bounds on the final time (0, 10) */
} mpcb_limits;

/* Per-problem solver report (polympc's sqp info + the quantities the termination test uses). */
typedef struct {
    int status;                   /* 1 = SOLVED, 2 = MAX_ITER_EXCEEDED */
    int sqp_iters;
    int qp_iters_total;
    int qp_status[4];             /* per SQP iteration (first 4): 1 = solved, 2 = max iter */
    int qp_iters[4];
    double alpha[4];              /* accepted (or last) line-search step */
    int ls_trials[4];
    double primal_step_norm, dual_step_norm, max_violation, l1_violation, cost;
} mpcb_stats;

/* Per-trajectory accounting of examples/benchmark.cpp:118-194 (MPC half). */
typedef struct {
    int flags[4];                 /* jerk, linear velocity, angular velocity, collision; 1 = ok */
    double mins[28], maxs[28];    /* q, v, a, tau extrema over the samples */
    double final_err[14];         /* last sample - target */
} mpcb_account;

typedef struct mpcb_handle mpcb_handle;

enum { MPCB_MEM_HOST = 0, MPCB_MEM_DEVICE = 1 };

void mpcb_default_settings(mpcb_settings* s);
void mpcb_default_limits(mpcb_limits* l);            /* Panda limits, margins 1 (motionPlanner.cpp:24) */

/* Sizes for a (poly_order, num_seg) pair. */
int mpcb_num_nodes(const mpcb_settings* s);
int mpcb_num_var(const mpcb_settings* s);
int mpcb_num_dual(const mpcb_settings* s);
/* Normalised node times in [0,1], ascending (polympc `time_nodes`; motionPlanner.cpp:156). */
int mpcb_time_nodes(const mpcb_settings* s, double* t /*N*/);

/* One handle per CUDA device; not thread-safe; replaces the `mpc_t mpc` member (motionPlanner.hpp:28-29).
 * A handle owns one workspace: calls that pass different streams are serialised on the device (a call on a new stream first
 * waits for the handle's previous call), so one handle never has two solves in flight; use one handle per concurrent stream. */
int mpcb_create(const mpcb_settings* s, int device, mpcb_handle** out);
void mpcb_destroy(mpcb_handle* h);
const char* mpcb_last_error(const mpcb_handle* h);   /* h may be NULL: error of the last failed mpcb_create */

/* Replaces set_constraint_margins / set_min_height / the limits table (motionPlanner.cpp:56-100). */
int mpcb_set_limits(mpcb_handle* h, const mpcb_limits* lim);

/*
 * Solve n independent OCPs (replaces n calls of mpc.solve(), motionPlanner.cpp:184, with the bounds
 * set_current_state / set_target_state would install, motionPlanner.cpp:27-54).
 *   x0, xT      n x 14       initial state (pinned) and target state (box +- final_eps)
 *   guess_x     n x 14N      initial guess at the nodes (what warm_start_RK writes, motionPlanner.cpp:156-174)
 *   guess_u     n x 7N
 *   guess_T     n
 *   guess_lam   n x ndual or NULL (zeros)
 *   sol_x/sol_u/sol_T          outputs, same layouts;  sol_lam  n x ndual or NULL;  stats  n or NULL
 *   mem         MPCB_MEM_HOST: plain host pointers, copies are made inside the call (synchronous);
 *               MPCB_MEM_DEVICE: device pointers on the handle's device, the call only enqueues work on `stream`
 *   stream      cudaStream_t (NULL = default stream)
 * Environment (read by mpcb_create): MPCB_PARTS = 2..4 lets the call split every chunk of at least two waves of CTAs over internal
 * streams of the handle (fork / join with events behind `stream`); default 1 = off.
 */
int mpcb_solve_batch(mpcb_handle* h, int n, const double* x0, const double* xT, const double* guess_x,
                     const double* guess_u, const double* guess_T, const double* guess_lam, double* sol_x,
                     double* sol_u, double* sol_T, double* sol_lam, mpcb_stats* stats, int mem, void* stream);

/*
 * mpcb_solve_batch for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Same kernels, arguments and results as mpcb_solve_batch with MPCB_MEM_DEVICE (every pointer is a device pointer on the handle's
 * device; the call only enqueues work on `stream`), but the per-OCP solver workspace and the work counters live in file-scope
 * static __device__ arrays of libmpcb instead of buffers the handle allocates, and the call makes no cudaMalloc / cudaFree / event
 * creation.  Limit: n <= MPCB_NOALLOC_MAX_OCPS for every collocation shape; the static workspace is sized for that many problems
 * of the base shape (3,6), which has the largest per-OCP workspace (8 x 539 KB = 4.3 MB of device memory, held by every GPU that
 * loads the library until the process exits).  n beyond the limit returns
 * MPCB_ERR_BAD_ARG with the limit in mpcb_last_error; nothing is enqueued.  The scratch is shared by ALL handles and calls of
 * the process on that device and keeps the last call's contents (the call resets what must start at zero: the work counters):
 * calls that overlap on the GPU (different streams or handles, no ordering between them) race -- order them on one stream or with
 * events.  The per-kernel timers (mpcb_last_solve_ms, mpcb_last_kernel_ms) are not updated by this entry point.
 */
#define MPCB_NOALLOC_MAX_OCPS 8
int mpcb_solve_batch_noalloc(mpcb_handle* h, int n, const double* x0, const double* xT, const double* guess_x,
                             const double* guess_u, const double* guess_T, const double* guess_lam, double* sol_x,
                             double* sol_u, double* sol_T, double* sol_lam, mpcb_stats* stats, void* stream);

/*
 * Read-out + accounting (get_MPC_trajectory<npts>, motionPlanner.hpp:99-116; benchmark.cpp:118-194):
 *   traj   n x (npts+1) x 29  rows [t, q(7), v(7), a(7), tau(7)]  (may be NULL)
 *   acc    n                  (may be NULL);   target n x 14 (needed when acc != NULL)
 */
int mpcb_sample_batch(mpcb_handle* h, int n, const double* sol_x, const double* sol_u, const double* sol_T,
                      const double* target, int npts, double* traj, mpcb_account* acc, int mem, void* stream);

/* Device time of the last mpcb_solve_batch kernel on this handle, in milliseconds (CUDA events on the launch
 * stream); < 0 if none.  Synchronises the stream. */
double mpcb_last_solve_ms(mpcb_handle* h);
/* Number of kernel launches issued by this handle so far. */
long long mpcb_launch_count(const mpcb_handle* h);
/* The last mpcb_solve_batch split by kernel kind: 0 = SQP stage kernels (linearise, factorise, line search, step),
 * 1 = ADMM kernel (the QP iterations), 2 = node kernels (RNEA values / derivatives, one thread per node of the batch).
 * Summed device time in ms (CUDA events around each launch) and launch count. */
enum { MPCB_KERNEL_SQP_STAGE = 0, MPCB_KERNEL_ADMM = 1, MPCB_KERNEL_NODE = 2 };
double mpcb_last_kernel_ms(mpcb_handle* h, int kind);
int mpcb_last_kernel_launches(const mpcb_handle* h, int kind);
/* Debug: cumulative SM cycles per ADMM phase of CTA 0 (all zero unless libmpcb was built with -DMPCB_PROFILE).
 * ADMM kernel slots: 0 load, 1 rhs, 2 interior W/F^T, 3 separator rhs, 4 separator solve, 5 interior back-substitution,
 * 6 row/box update, 7 residual test, 8 store.  SQP stage slots: 16 load, 17 linearise, 18 KKT assembly, 19 interior sweeps,
 * 20 separator, 21 export, 22 merit + line search, 23 step + termination, 24 store, 25 factorise tail. */
int mpcb_debug_phase_cycles(mpcb_handle* h, long long* out512, int reset);   /* 512 slots; 64.. = per-warp stamps */

/* Debug / per-function parity tests: minTime_ocp::evalConstraints<double> and <ad_scalar_t> (robot_ocp.hpp:80-163) exactly as the
 * node kernels evaluate them, for n arbitrary states.  qva: n x 21 (q, v, a); g_val: n x 8 from the value pass; g_jac: n x 8 and
 * Jg: n x 8 x 22 from the derivative pass.  Host pointers. */
int mpcb_debug_node_constraints(mpcb_handle* h, int n, const double* qva, double* g_val, double* g_jac, double* Jg);

/* FP64 FMA peak micro-benchmark on the handle's device (roofline denominator); returns TFLOP/s. */
double mpcb_measure_fp64_tflops(mpcb_handle* h);

/* ---- host-side helpers of the planner (no GPU work) ------------------------------------------------------ */

/* MotionPlanner::sample_random_state (motionPlanner.cpp:102-114) + the benchmark's target generation
 * (examples/benchmark.cpp:19-42), with a counter-based RNG instead of std::rand.
 *   kind 0: mpc_benchmark -- start fixed at mid-range at rest, random target pose, target velocity from a random
 *           Cartesian tool velocity through the damped pseudo-inverse, with the benchmark's scale-downs
 *   kind 1: offline_trajectory -- random start and target states (examples/offline_trajectory.cpp:18-41)
 * Outputs x0, xT: n x 14. */
int mpcb_generate_problems(const mpcb_limits* lim, int kind, unsigned long long seed, int n, double* x0, double* xT);

/* Jerk-limited, time-synchronised joint profiles: the role of Ruckig<7> in the reference (motionPlanner.hpp:35-37,
 * motionPlanner.cpp:36-38,50-52,86-88,149-160).  Per joint 7 constant-jerk phases (pulse, cruise, pulse) from
 * (q0, v0, a = 0) to (qT, vT, a = 0) under margins x (v_max, a_max, j_max); all joints end at T.
 *   mpcb_otg_calculate   Ruckig::calculate   (0 = ok, 1 = ok with T moved past a blocked interval, < 0 = no profile, e.g.
 *                                             a velocity outside the limits: Ruckig rejects that input too)
 *   mpcb_otg_at_time     Trajectory::at_time (motionPlanner.cpp:160; motionPlanner.hpp:85,135)
 * Pinned on the Ruckig trajectory the reference embeds (tests/test_otg.py). */
typedef struct {
    double T;                     /* Trajectory::get_duration() */
    int status;
    double q0[7], v0[7], vc[7];   /* start state and cruise velocity per joint */
    double t[7][7], j[7][7];      /* per joint: phase durations and jerks */
} mpcb_otg_traj;
int mpcb_otg_calculate(const mpcb_limits* lim, const double* x0, const double* xT, mpcb_otg_traj* traj);
int mpcb_otg_at_time(const mpcb_otg_traj* traj, double t, double* q, double* v, double* a);

/* Initial guess at the collocation nodes = MotionPlanner::warm_start_RK (motionPlanner.cpp:146-175): the jerk-limited
 * profile of mpcb_otg_calculate sampled at the normalised node times, guess_T = its duration.  Problems the generator
 * rejects (Ruckig would too, and the reference then keeps a stale trajectory, motionPlanner.cpp:149) get a time-scaled
 * cubic Hermite profile that respects the velocity and acceleration margins instead. */
int mpcb_guess_batch(const mpcb_settings* s, const mpcb_limits* lim, int n, const double* x0, const double* xT,
                     double* guess_x, double* guess_u, double* guess_T);

/* The same warm start computed on the device of handle `h` with its current limits (otg_guess_kernel: 8 lanes per problem, one
 * joint per lane), so that a batch can be generated, warm-started and solved without the host touching it.  `mem` says whether the
 * five arrays are host or device pointers.  Same per-joint routines as mpcb_guess_batch; results agree to rounding (the device
 * contracts multiply-adds). */
int mpcb_guess_batch_gpu(mpcb_handle* h, int n, const double* x0, const double* xT, double* guess_x, double* guess_u, double* guess_T,
                         int mem, void* cuda_stream);

/* The warm-start ("Ruckig") half of the benchmark on the device: get_ruckig_trajectory<npts> (motionPlanner.hpp:73-96) of the jerk-limited
 * profile from x0 to xT under the handle's limits, with the per-trajectory checks of examples/benchmark.cpp:72-114 (same flags, extrema and
 * final-state error as mpcb_sample_batch computes for the MPC half).  traj n x (npts+1) x 29, acc n, T_out n (profile durations); each may
 * be NULL.  sample_profile_kernel: one CTA per problem, the profile by the routines the warm start uses. */
int mpcb_sample_profile_batch(mpcb_handle* h, int n, const double* x0, const double* xT, int npts, double* traj, mpcb_account* acc,
                              double* T_out, int mem, void* cuda_stream);

/* mpcb_generate_problems on the device of handle `h` with its current limits (gen_problems_kernel, one problem per thread): problem i of
 * the call draws from stream (seed, first + i), so a rank generates its shard [first, first + n) of a batch in place.  Same routines as
 * the host version; results agree to rounding (device sin/cos and contracted multiply-adds). */
int mpcb_generate_problems_gpu(mpcb_handle* h, int kind, unsigned long long seed, unsigned long long first, int n, double* x0, double* xT,
                               int mem, void* cuda_stream);

/* Single-state host helpers used by the C++ MotionPlanner mirror (include/motion_planner.hpp):
 *   mpcb_sample_random_state  MotionPlanner::sample_random_state (motionPlanner.cpp:102-114), counter-based RNG (seed, stream)
 *   mpcb_forward_velocities   PandaWrapper::forward_velocities (robot_utils/pandaWrapper.cpp:90-107): tool twist [lin; ang]
 *   mpcb_inverse_velocities   PandaWrapper::inverse_velocities (pandaWrapper.cpp:60-88): damped pseudo-inverse, damping 1e-5
 *   mpcb_rnea                 pinocchio::rnea on the Panda model (motionPlanner.hpp:91-94,110-113,127)
 *   mpcb_frame_height         z of joint 7 (frame 0; motionPlanner.cpp:111) or of panda_tool (frame 1; robot_ocp.hpp:91) */
int mpcb_sample_random_state(const mpcb_limits* lim, unsigned long long seed, unsigned long long stream, double* q, double* v);
int mpcb_forward_velocities(const double* q, const double* qd, double* task6);
int mpcb_inverse_velocities(const double* q, const double* linear3, const double* angular3, double* qd);
int mpcb_rnea(const double* q, const double* v, const double* a, double* tau);
/*   mpcb_tool_pose            oMf[panda_tool] after forwardKinematics + updateFramePlacement (examples/benchmark.cpp:108-110):
 *                             rotation column-major (Eigen Matrix3d::data()) and position
 *   mpcb_inverse_kinematic    PandaWrapper::inverse_kinematic (robot_utils/pandaWrapper.cpp:14-58): damped closed-loop IK of the
 *                             tool frame from a random configuration (seeded); 0 = converged, 1 = iteration limit */
int mpcb_tool_pose(const double* q, double* R_colmajor, double* pos);
int mpcb_inverse_kinematic(const mpcb_limits* lim, const double* R_des_colmajor, const double* p_des, unsigned long long seed, double* q_out, int* iters);
int mpcb_frame_height(const double* q, int frame, double* z);

#ifdef __cplusplus
}
#endif
#endif
