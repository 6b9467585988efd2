// Host-side planner helpers behind the C ABI: problem generation and the initial guess (no GPU work).
//
//   mpcb_generate_problems : MotionPlanner::sample_random_state (/root/reference/mpc_solver/motionPlanner.cpp:102-114)
//                            + the target generation of examples/benchmark.cpp:19-42 (kind 0) and
//                            examples/offline_trajectory.cpp:18-41 (kind 1), using PandaWrapper::inverse_velocities /
//                            forward_velocities (robot_utils/pandaWrapper.cpp:60-107).
//   mpcb_guess_batch       : MotionPlanner::warm_start_RK (motionPlanner.cpp:146-175) on the jerk-limited profiles of mpcb_otg.h (the role of Ruckig<7>),
//                            sampled at polympc's node times; inputs the generator rejects get a time-scaled cubic Hermite profile.
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "mpcb.h"
#include "mpcb_gen.h"
#include "mpcb_host_util.h"
#include "mpcb_otg.h"
#include "mpcb_rbd.h"
#include "mpcb_shape.h"

namespace {

using namespace mpcb;

using namespace mpcb::gen;

}  // namespace

extern "C" {

int mpcb_generate_problems(const mpcb_limits* lim, int kind, unsigned long long seed, int n, double* x0, double* xT) {
    if (!lim || !x0 || !xT || n < 0 || (kind != 0 && kind != 1)) return MPCB_ERR_BAD_ARG;
    if (!(lim->min_height <= MPCB_GEN_MAX_HEIGHT)) return MPCB_ERR_BAD_ARG;      // out of reach: the rejection sampler would not terminate
    for (int i = 0; i < n; i++) generate_one(lim, kind, seed, (uint64_t)i, x0 + 14 * (size_t)i, xT + 14 * (size_t)i);
    return MPCB_OK;
}

// ---- single-state helpers for the MotionPlanner mirror (include/motion_planner.hpp) -----------------------------------
int mpcb_sample_random_state(const mpcb_limits* lim, unsigned long long seed, unsigned long long stream, double* q, double* v) {
    if (!lim || !q || !v || !(lim->min_height <= MPCB_GEN_MAX_HEIGHT)) return MPCB_ERR_BAD_ARG;
    Rng rng(seed, stream);
    sample_random_state(lim, rng, q, v);
    return MPCB_OK;
}
int mpcb_forward_velocities(const double* q, const double* qd, double* task6) {
    if (!q || !qd || !task6) return MPCB_ERR_BAD_ARG;
    forward_velocities(q, qd, task6);
    return MPCB_OK;
}
int mpcb_inverse_velocities(const double* q, const double* linear3, const double* angular3, double* qd) {
    if (!q || !linear3 || !angular3 || !qd) return MPCB_ERR_BAD_ARG;
    const double task[6] = {linear3[0], linear3[1], linear3[2], angular3[0], angular3[1], angular3[2]};
    inverse_velocities(q, task, qd);
    return MPCB_OK;
}
// oMf[panda_tool] (pinocchio::forwardKinematics + updateFramePlacement, examples/benchmark.cpp:108-110): rotation (column-major,
// as Eigen's Matrix3d::data()) and position.  panda_joint8 / panda_tool carry no rotation (panda_arm.urdf:134-152).
int mpcb_tool_pose(const double* q, double* R_colmajor, double* pos) {
    if (!q || !R_colmajor || !pos) return MPCB_ERR_BAD_ARG;
    double oR[7][9], op[7][3];
    rbd::fk_world(q, oR, op);
    rbd::frame_jacobian(oR, op, PANDA(PANDA_TOOL_OFFSET), nullptr, pos);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R_colmajor[3 * c + r] = oR[6][3 * r + c];
    return MPCB_OK;
}
// PandaWrapper::inverse_kinematic (robot_utils/pandaWrapper.cpp:14-58): damped closed-loop inverse kinematics of the tool
// frame from a random configuration; eps 1e-4, <= 1000 iterations, step 0.1, damping 1e-2; the error is log6 of
// oMdes^-1 oMf (linear part first), the Jacobian is the LOCAL frame Jacobian, exactly as the reference does.
// Returns 0 when converged, 1 when the iteration limit was hit (the reference returns q in both cases).
int mpcb_inverse_kinematic(const mpcb_limits* lim, const double* R_des_colmajor, const double* p_des, unsigned long long seed, double* q_out, int* iters) {
    if (!lim || !R_des_colmajor || !p_des || !q_out) return MPCB_ERR_BAD_ARG;
    Rng rng(seed, 0x1C);
    double q[7], Rd[9];
    for (int j = 0; j < 7; j++) q[j] = 0.5 * (lim->qmin[j] + lim->qmax[j]) + 0.5 * rng.sym() * (lim->qmax[j] - lim->qmin[j]);   // randomConfiguration
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Rd[3 * r + c] = R_des_colmajor[3 * c + r];
    int rc = 1, i = 0;
    for (;; i++) {
        double oR[7][9], op[7][3], Jw[42], pf[3];
        rbd::fk_world(q, oR, op);
        rbd::frame_jacobian(oR, op, PANDA(PANDA_TOOL_OFFSET), Jw, pf);
        const double* Rf = oR[6];
        double dR[9], dp[3], d[3] = {pf[0] - p_des[0], pf[1] - p_des[1], pf[2] - p_des[2]};
        for (int r = 0; r < 3; r++) {
            dp[r] = Rd[r] * d[0] + Rd[3 + r] * d[1] + Rd[6 + r] * d[2];
            for (int c = 0; c < 3; c++) dR[3 * r + c] = Rd[r] * Rf[c] + Rd[3 + r] * Rf[3 + c] + Rd[6 + r] * Rf[6 + c];
        }
        // log6
        double w[3] = {0.5 * (dR[7] - dR[5]), 0.5 * (dR[2] - dR[6]), 0.5 * (dR[3] - dR[1])};
        double ct = 0.5 * (dR[0] + dR[4] + dR[8] - 1.0);
        ct = ct > 1 ? 1 : (ct < -1 ? -1 : ct);
        const double th = acos(ct), st = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
        const double sc = th < 1e-8 ? 1.0 : (st > 1e-12 ? th / st : 0.0);
        for (int k = 0; k < 3; k++) w[k] *= sc;
        const double beta = th < 1e-4 ? 1.0 / 12.0 : (1.0 - th * sin(th) / (2.0 * (1.0 - ct))) / (th * th);
        double wxp[3] = {w[1] * dp[2] - w[2] * dp[1], w[2] * dp[0] - w[0] * dp[2], w[0] * dp[1] - w[1] * dp[0]};
        double wxwxp[3] = {w[1] * wxp[2] - w[2] * wxp[1], w[2] * wxp[0] - w[0] * wxp[2], w[0] * wxp[1] - w[1] * wxp[0]};
        double err[6];
        for (int k = 0; k < 3; k++) { err[k] = dp[k] - 0.5 * wxp[k] + beta * wxwxp[k]; err[3 + k] = w[k]; }
        double en = 0;
        for (int k = 0; k < 6; k++) en += err[k] * err[k];
        if (sqrt(en) < 1e-4) { rc = 0; break; }
        if (i >= 1000) break;
        // LOCAL frame Jacobian = blockdiag(Rf^T, Rf^T) x world-aligned Jacobian
        double J[42], A[36], Lm[36], D[6], y[6];
        for (int j = 0; j < 7; j++)
            for (int r = 0; r < 3; r++) {
                J[r * 7 + j] = Rf[r] * Jw[j] + Rf[3 + r] * Jw[7 + j] + Rf[6 + r] * Jw[14 + j];
                J[(3 + r) * 7 + j] = Rf[r] * Jw[21 + j] + Rf[3 + r] * Jw[28 + j] + Rf[6 + r] * Jw[35 + j];
            }
        for (int r = 0; r < 6; r++)
            for (int c = 0; c < 6; c++) { double s2 = 0; for (int j = 0; j < 7; j++) s2 += J[r * 7 + j] * J[c * 7 + j]; A[r * 6 + c] = s2 + (r == c ? 1e-2 : 0.0); }
        memset(Lm, 0, sizeof Lm);
        for (int j = 0; j < 6; j++) {
            double dd = A[j * 6 + j];
            for (int k = 0; k < j; k++) dd -= Lm[j * 6 + k] * Lm[j * 6 + k] * D[k];
            D[j] = dd; Lm[j * 6 + j] = 1;
            for (int r = j + 1; r < 6; r++) {
                double s2 = A[r * 6 + j];
                for (int k = 0; k < j; k++) s2 -= Lm[r * 6 + k] * Lm[j * 6 + k] * D[k];
                Lm[r * 6 + j] = s2 / dd;
            }
        }
        for (int r = 0; r < 6; r++) { double s2 = err[r]; for (int k = 0; k < r; k++) s2 -= Lm[r * 6 + k] * y[k]; y[r] = s2; }
        for (int r = 0; r < 6; r++) y[r] /= D[r];
        for (int r = 5; r >= 0; r--) { double s2 = y[r]; for (int k = r + 1; k < 6; k++) s2 -= Lm[k * 6 + r] * y[k]; y[r] = s2; }
        for (int j = 0; j < 7; j++) { double s2 = 0; for (int r = 0; r < 6; r++) s2 += J[r * 7 + j] * y[r]; q[j] -= 0.1 * s2; }
    }
    for (int j = 0; j < 7; j++) q_out[j] = q[j];
    if (iters) *iters = i;
    return rc;
}
int mpcb_rnea(const double* q, const double* v, const double* a, double* tau) {
    if (!q || !v || !a || !tau) return MPCB_ERR_BAD_ARG;
    double g[8];
    rbd::node_constraints_value(q, v, a, g);
    for (int j = 0; j < 7; j++) tau[j] = g[j];
    return MPCB_OK;
}
int mpcb_frame_height(const double* q, int frame, double* z) {
    if (!q || !z || frame < 0 || frame > 1) return MPCB_ERR_BAD_ARG;
    const double zero[3] = {0, 0, 0};
    double pos[3];
    rbd::fk_point7(q, frame == 0 ? zero : PANDA(PANDA_TOOL_OFFSET), pos);
    *z = pos[2];
    return MPCB_OK;
}

int mpcb_num_nodes(const mpcb_settings* s) { return s->poly_order * s->num_seg + 1; }
int mpcb_num_var(const mpcb_settings* s) { return 21 * mpcb_num_nodes(s) + 1; }
int mpcb_num_dual(const mpcb_settings* s) { const int n = mpcb_num_nodes(s); return 14 * (n - 1) + 8 * n + 21 * n + 1; }
int mpcb_time_nodes(const mpcb_settings* s, double* t) {
    if (!s || !t || s->poly_order < 1 || s->poly_order > 7 || s->num_seg < 1) return MPCB_ERR_BAD_ARG;
    ChebConst cc;
    mpcb_fill_cheb(s->poly_order, s->num_seg, &cc);
    for (int sg = 0; sg < s->num_seg; sg++)
        for (int j = 0; j <= s->poly_order; j++) t[sg * s->poly_order + j] = ((double)sg + 0.5 * (1.0 + cc.xi[j])) / s->num_seg;
    return MPCB_OK;
}
void mpcb_default_settings(mpcb_settings* s) { default_settings(s); }
void mpcb_default_limits(mpcb_limits* l) { default_limits(l); }

int mpcb_otg_calculate(const mpcb_limits* lim, const double* x0, const double* xT, mpcb_otg_traj* traj) {
    if (!lim || !x0 || !xT || !traj) return MPCB_ERR_BAD_ARG;
    return otg::calculate(lim, x0, xT, traj);
}
int mpcb_otg_at_time(const mpcb_otg_traj* traj, double t, double* q, double* v, double* a) {
    if (!traj || !q || !v || !a) return MPCB_ERR_BAD_ARG;
    otg::at_time(traj, t, q, v, a);
    return MPCB_OK;
}

// Fallback of mpcb_guess_batch: cubic Hermite joint profile q(s), s = t/T, between (q0, v0) and (q1, v1); T = first value of the geometric grid
// T_j = 0.05 * 1.05^j (capped at p_ub) for which |v| <= m_v v_max over [0,T] and |a| <= m_a a_max (the acceleration is
// linear in s, so its extrema are at the ends; the velocity is quadratic, extrema at the ends or at the vertex).
int mpcb_guess_batch(const mpcb_settings* s, const mpcb_limits* lim, int n, const double* x0, const double* xT, double* gx,
                     double* gu, double* gT) {
    if (!s || !lim || !x0 || !xT || !gx || !gu || !gT || n < 0) return MPCB_ERR_BAD_ARG;
    const int N = mpcb_num_nodes(s);
    if (N > 128) return MPCB_ERR_BAD_ARG;
    double tn[128];
    if (mpcb_time_nodes(s, tn) != MPCB_OK) return MPCB_ERR_BAD_ARG;
    for (int i = 0; i < n; i++) {
        const double *q0 = x0 + 14 * (size_t)i, *v0 = q0 + 7, *q1 = xT + 14 * (size_t)i, *v1 = q1 + 7;
        {   // motionPlanner.cpp:146-175: jerk-limited synchronised profile sampled at the node times
            mpcb_otg_traj tr;
            if (otg::calculate(lim, q0, q1, &tr) >= 0 && tr.T > 0 && tr.T <= lim->p_ub) {
                gT[i] = tr.T;
                for (int k = 0; k < N; k++)
                    otg::at_time(&tr, tn[k] * tr.T, gx + ((size_t)i * N + k) * 14, gx + ((size_t)i * N + k) * 14 + 7, gu + ((size_t)i * N + k) * 7);
                continue;
            }
        }
        double T = 0.05;
        for (;;) {
            bool ok = true;
            for (int j = 0; j < 7 && ok; j++)
                ok = otg::hermite_ok(q1[j] - q0[j], v0[j], v1[j], T, lim->margins[1] * lim->vmax[j], lim->margins[2] * lim->amax[j]);
            if (ok || T >= lim->p_ub) break;
            T *= 1.05;
            if (T > lim->p_ub) T = lim->p_ub;
        }
        gT[i] = T;
        for (int k = 0; k < N; k++) {
            double* xk = gx + ((size_t)i * N + k) * 14;
            double* uk = gu + ((size_t)i * N + k) * 7;
            for (int j = 0; j < 7; j++) otg::hermite_eval(q0[j], v0[j], q1[j], v1[j], T, tn[k], xk + j, xk + 7 + j, uk + j);
        }
    }
    return MPCB_OK;
}

}  // extern "C"
