/*
 * oracle.cpp -- CPU oracle: TEST INFRASTRUCTURE, NOT PRODUCT CODE.  PARITY UNPINNED for the SQP/ADMM part
 * (see oracle.h).  Plain C++17, no dependencies; build with `make -C oracle` (-O2 -ffp-contract=off).
 *
 * What each block restates (reference file:line under /root/reference, or SURVEY.md appendix when the
 * arithmetic lives in an absent third-party library):
 *   rigid body  : pinocchio forwardKinematics/rnea/crba/computeRNEADerivatives/computeFrameJacobian as called
 *                 from mpc_solver/robot_ocp.hpp:86-91,111-155 and robot_utils/pandaWrapper.cpp:65-73  (App. A.1, A.9)
 *   transcription: polympc ContinuousOCP<minTime_ocp, Spline<Chebyshev<3,GAUSS_LOBATTO>,6>, SPARSE>
 *                 parameterised at robot_ocp.hpp:31-42,55-73,193-213                                     (App. A.2, A.3)
 *   regularise  : mpc_solver/polympc_redef.hpp:57-70
 *   line search : mpc_solver/polympc_redef.hpp:73-121
 *   SQP loop    : polympc SQPBase::solve, settings at mpc_solver/motionPlanner.cpp:15-20                 (App. A.4)
 *   QP          : polympc boxADMM::solve                                                                 (App. A.5)
 *   bounds      : mpc_solver/motionPlanner.cpp:27-100
 *   read-out    : mpc_solver/motionPlanner.hpp:99-116 ; accounting: examples/benchmark.cpp:118-194
 */
#include "oracle.h"
#include "panda_model_constants.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace {

constexpr int NJ = 7;
constexpr double INF = std::numeric_limits<double>::infinity();

// ------------------------------------------------------------------------------------------------
// 3-vectors / 3x3 (row-major) helpers
// ------------------------------------------------------------------------------------------------
inline void cross3(const double* a, const double* b, double* o) {
    double x = a[1] * b[2] - a[2] * b[1];
    double y = a[2] * b[0] - a[0] * b[2];
    double z = a[0] * b[1] - a[1] * b[0];
    o[0] = x; o[1] = y; o[2] = z;
}
inline void matvec3(const double* R, const double* v, double* o) {
    double x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
    double y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
    double z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
    o[0] = x; o[1] = y; o[2] = z;
}
inline void matTvec3(const double* R, const double* v, double* o) {
    double x = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
    double y = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
    double z = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
    o[0] = x; o[1] = y; o[2] = z;
}
inline void matmul3(const double* A, const double* B, double* C) {
    double t[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) t[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
    std::memcpy(C, t, sizeof t);
}
inline void skew3(const double* v, double* S) {
    S[0] = 0; S[1] = -v[2]; S[2] = v[1];
    S[3] = v[2]; S[4] = 0; S[5] = -v[0];
    S[6] = -v[1]; S[7] = v[0]; S[8] = 0;
}

// joint transform liMi(q_i) = Placement_i * Rz(q_i)
inline void joint_local(int i, double qi, double* R, double* p) {
    double c = std::cos(qi), s = std::sin(qi);
    double Rz[9] = {c, -s, 0, s, c, 0, 0, 0, 1};
    matmul3(PANDA_PLACEMENT_R[i], Rz, R);
    p[0] = PANDA_PLACEMENT_P[i][0]; p[1] = PANDA_PLACEMENT_P[i][1]; p[2] = PANDA_PLACEMENT_P[i][2];
}

// pinocchio::forwardKinematics (robot_ocp.hpp:87): world placements oMi
void fk(const double* q, double oR[NJ][9], double op[NJ][3]) {
    double Rp[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, pp[3] = {0, 0, 0};
    for (int i = 0; i < NJ; i++) {
        double Rl[9], pl[3], t[3];
        joint_local(i, q[i], Rl, pl);
        matmul3(Rp, Rl, oR[i]);
        matvec3(Rp, pl, t);
        op[i][0] = pp[0] + t[0]; op[i][1] = pp[1] + t[1]; op[i][2] = pp[2] + t[2];
        std::memcpy(Rp, oR[i], sizeof Rp);
        std::memcpy(pp, op[i], sizeof pp);
    }
}

// ------------------------------------------------------------------------------------------------
// spatial vectors: (linear, angular) -- pinocchio's ordering
// ------------------------------------------------------------------------------------------------
struct Sp { double l[3], a[3]; };

// SE3(R,p).actInv on a motion
inline Sp motion_actInv(const double* R, const double* p, const Sp& m) {
    Sp o; double t[3], d[3];
    cross3(p, m.a, t);
    d[0] = m.l[0] - t[0]; d[1] = m.l[1] - t[1]; d[2] = m.l[2] - t[2];
    matTvec3(R, d, o.l);
    matTvec3(R, m.a, o.a);
    return o;
}
// SE3(R,p).act on a force
inline Sp force_act(const double* R, const double* p, const Sp& f) {
    Sp o; double t[3];
    matvec3(R, f.l, o.l);
    matvec3(R, f.a, o.a);
    cross3(p, o.l, t);
    o.a[0] += t[0]; o.a[1] += t[1]; o.a[2] += t[2];
    return o;
}
// Y * v for a body inertia (mass, com c, I about com), body frame
inline Sp inertia_mul(double m, const double* c, const double* I, const Sp& v) {
    Sp h; double t[3];
    cross3(v.a, c, t);
    h.l[0] = m * (v.l[0] + t[0]); h.l[1] = m * (v.l[1] + t[1]); h.l[2] = m * (v.l[2] + t[2]);
    matvec3(I, v.a, h.a);
    cross3(c, h.l, t);
    h.a[0] += t[0]; h.a[1] += t[1]; h.a[2] += t[2];
    return h;
}
inline Sp motion_cross_motion(const Sp& v, const Sp& m) {
    Sp o; double t1[3], t2[3];
    cross3(v.a, m.l, t1); cross3(v.l, m.a, t2);
    o.l[0] = t1[0] + t2[0]; o.l[1] = t1[1] + t2[1]; o.l[2] = t1[2] + t2[2];
    cross3(v.a, m.a, o.a);
    return o;
}
inline Sp motion_cross_force(const Sp& v, const Sp& f) {
    Sp o; double t1[3], t2[3];
    cross3(v.a, f.l, o.l);
    cross3(v.a, f.a, t1); cross3(v.l, f.l, t2);
    o.a[0] = t1[0] + t2[0]; o.a[1] = t1[1] + t2[1]; o.a[2] = t1[2] + t2[2];
    return o;
}

// pinocchio::rnea (robot_ocp.hpp:91,120), body-frame recursion of SURVEY A.1
void rnea(const double* q, const double* qd, const double* qdd, double* tau, double gravity) {
    double Rl[NJ][9], pl[NJ][3];
    Sp v[NJ], a[NJ], f[NJ];
    Sp vp = {{0, 0, 0}, {0, 0, 0}};
    Sp ap = {{0, 0, gravity}, {0, 0, 0}};
    for (int i = 0; i < NJ; i++) {
        joint_local(i, q[i], Rl[i], pl[i]);
        v[i] = motion_actInv(Rl[i], pl[i], vp);
        v[i].a[2] += qd[i];
        Sp sq = {{0, 0, 0}, {0, 0, qd[i]}};
        Sp vxs = motion_cross_motion(v[i], sq);
        a[i] = motion_actInv(Rl[i], pl[i], ap);
        a[i].a[2] += qdd[i];
        for (int k = 0; k < 3; k++) { a[i].l[k] += vxs.l[k]; a[i].a[k] += vxs.a[k]; }
        Sp h = inertia_mul(PANDA_MASS[i], PANDA_COM[i], PANDA_ICOM[i], v[i]);
        Sp ya = inertia_mul(PANDA_MASS[i], PANDA_COM[i], PANDA_ICOM[i], a[i]);
        Sp vxh = motion_cross_force(v[i], h);
        for (int k = 0; k < 3; k++) { f[i].l[k] = ya.l[k] + vxh.l[k]; f[i].a[k] = ya.a[k] + vxh.a[k]; }
        vp = v[i]; ap = a[i];
    }
    for (int i = NJ - 1; i >= 0; i--) {
        tau[i] = f[i].a[2];
        if (i > 0) {
            Sp fp = force_act(Rl[i], pl[i], f[i]);
            for (int k = 0; k < 3; k++) { f[i - 1].l[k] += fp.l[k]; f[i - 1].a[k] += fp.a[k]; }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// 6x6 helpers for the world-frame derivative recursion
// ------------------------------------------------------------------------------------------------
typedef double M6[6][6];
typedef double V6[6];

inline void crm(const double* v, M6 X) {  // motion cross-product matrix [[w x, v x],[0, w x]]
    double Sw[9], Sv[9];
    skew3(v + 3, Sw); skew3(v, Sv);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            X[i][j] = Sw[3 * i + j]; X[i][j + 3] = Sv[3 * i + j];
            X[i + 3][j] = 0; X[i + 3][j + 3] = Sw[3 * i + j];
        }
}
inline void crf(const double* v, M6 X) {  // force cross-product matrix = -crm(v)^T
    M6 T; crm(v, T);
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) X[i][j] = -T[j][i];
}
inline void mv6(const M6 A, const double* x, double* y) {
    double t[6];
    for (int i = 0; i < 6; i++) { double s = 0; for (int j = 0; j < 6; j++) s += A[i][j] * x[j]; t[i] = s; }
    std::memcpy(y, t, sizeof t);
}
inline void mm6(const M6 A, const M6 B, M6 C) {
    M6 T;
    for (int i = 0; i < 6; i++) for (int j = 0; j < 6; j++) { double s = 0; for (int k = 0; k < 6; k++) s += A[i][k] * B[k][j]; T[i][j] = s; }
    std::memcpy(C, T, sizeof T);
}
inline double dot6(const double* a, const double* b) { double s = 0; for (int i = 0; i < 6; i++) s += a[i] * b[i]; return s; }

// world-frame spatial inertia of body i about the world origin
void world_inertia(int i, const double* R, const double* p, M6 Y) {
    double m = PANDA_MASS[i];
    double cw[3], t[3];
    matvec3(R, PANDA_COM[i], t);
    cw[0] = t[0] + p[0]; cw[1] = t[1] + p[1]; cw[2] = t[2] + p[2];
    double RI[9], Rt[9], Iw[9], C[9], CC[9];
    matmul3(R, PANDA_ICOM[i], RI);
    for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) Rt[3 * r + c] = R[3 * c + r];
    matmul3(RI, Rt, Iw);
    skew3(cw, C);
    matmul3(C, C, CC);
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) {
            Y[r][c] = (r == c) ? m : 0.0;
            Y[r][c + 3] = -m * C[3 * r + c];
            Y[r + 3][c] = m * C[3 * r + c];
            Y[r + 3][c + 3] = Iw[3 * r + c] - m * CC[3 * r + c];
        }
}

// pinocchio::computeRNEADerivatives + rnea + crba as used at robot_ocp.hpp:118-122.
// dq, dv full 7x7; M symmetric (the reference symmetrises crba's upper triangle, robot_ocp.hpp:122;
// computeRNEADerivatives' own `da` output is the upper triangle of the same matrix).
void rnea_derivatives(const double* q, const double* qd, const double* qdd, double* tau, double* dq, double* dv, double* M) {
    double oR[NJ][9], op[NJ][3];
    fk(q, oR, op);
    V6 J[NJ], ov[NJ], oa[NJ], oh[NJ], of[NJ], dJ[NJ], dVdq[NJ], dAdq[NJ], dAdv[NJ];
    static thread_local M6 oY[NJ], B[NJ];
    V6 ov0 = {0, 0, 0, 0, 0, 0}, oa0 = {0, 0, PANDA_GRAVITY, 0, 0, 0};
    for (int i = 0; i < NJ; i++) {
        const double* vpar = (i == 0) ? ov0 : ov[i - 1];
        const double* apar = (i == 0) ? oa0 : oa[i - 1];
        double z[3] = {oR[i][2], oR[i][5], oR[i][8]};
        double pxz[3]; cross3(op[i], z, pxz);
        for (int k = 0; k < 3; k++) { J[i][k] = pxz[k]; J[i][k + 3] = z[k]; }
        V6 Jqd;
        for (int k = 0; k < 6; k++) { Jqd[k] = J[i][k] * qd[i]; ov[i][k] = vpar[k] + Jqd[k]; }
        M6 Xv; crm(ov[i], Xv);
        V6 t; mv6(Xv, Jqd, t);
        for (int k = 0; k < 6; k++) oa[i][k] = apar[k] + J[i][k] * qdd[i] + t[k];
        world_inertia(i, oR[i], op[i], oY[i]);
        mv6(oY[i], ov[i], oh[i]);
        M6 Xf; crf(ov[i], Xf);
        V6 ya, vxh; mv6(oY[i], oa[i], ya); mv6(Xf, oh[i], vxh);
        for (int k = 0; k < 6; k++) of[i][k] = ya[k] + vxh[k];
        mv6(Xv, J[i], dJ[i]);
        if (i == 0) { for (int k = 0; k < 6; k++) dVdq[i][k] = 0; }
        else { M6 Xp; crm(vpar, Xp); mv6(Xp, J[i], dVdq[i]); }
        {
            M6 Xa; crm(apar, Xa); mv6(Xa, J[i], dAdq[i]);
            if (i > 0) { M6 Xp; crm(vpar, Xp); V6 t2; mv6(Xp, dVdq[i], t2); for (int k = 0; k < 6; k++) dAdq[i][k] += t2[k]; }
        }
        for (int k = 0; k < 6; k++) dAdv[i][k] = dJ[i][k] + dVdq[i][k];
        // B_i = crf(ov) oY - oY crm(ov) + Fx(oh)
        M6 T1, T2; mm6(Xf, oY[i], T1); mm6(oY[i], Xv, T2);
        double Sf[9], Sn[9]; skew3(oh[i], Sf); skew3(oh[i] + 3, Sn);
        for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) B[i][r][c] = T1[r][c] - T2[r][c];
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) {
                B[i][r][c + 3] -= Sf[3 * r + c];
                B[i][r + 3][c] -= Sf[3 * r + c];
                B[i][r + 3][c + 3] -= Sn[3 * r + c];
            }
    }
    V6 dFda[NJ], dFdv[NJ], dFdq[NJ];
    for (int i = 0; i < NJ * NJ; i++) { dq[i] = 0; dv[i] = 0; M[i] = 0; }
    for (int i = NJ - 1; i >= 0; i--) {
        tau[i] = dot6(J[i], of[i]);
        mv6(oY[i], J[i], dFda[i]);
        V6 t1, t2;
        mv6(B[i], J[i], t1); mv6(oY[i], dAdv[i], t2);
        for (int k = 0; k < 6; k++) dFdv[i][k] = t1[k] + t2[k];
        mv6(oY[i], dAdq[i], t1); mv6(B[i], dVdq[i], t2);
        for (int k = 0; k < 6; k++) dFdq[i][k] = t1[k] + t2[k];
        for (int j = i; j < NJ; j++) {
            M[i * NJ + j] = dot6(J[i], dFda[j]);
            dv[i * NJ + j] = dot6(J[i], dFdv[j]);
            dq[i * NJ + j] = dot6(J[i], dFdq[j]);
        }
        {   // dFdq_i += crf(J_i) of_i   (after row i used it)
            M6 XJ; crf(J[i], XJ); mv6(XJ, of[i], t1);
            for (int k = 0; k < 6; k++) dFdq[i][k] += t1[k];
        }
        // ancestors: rows (J_i^T oY_i) and (J_i^T B_i)
        V6 JtY, JtB;
        for (int c = 0; c < 6; c++) { double s1 = 0, s2 = 0; for (int r = 0; r < 6; r++) { s1 += J[i][r] * oY[i][r][c]; s2 += J[i][r] * B[i][r][c]; } JtY[c] = s1; JtB[c] = s2; }
        for (int j = 0; j < i; j++) {
            dq[i * NJ + j] = dot6(JtY, dAdq[j]) + dot6(JtB, dVdq[j]);
            dv[i * NJ + j] = dot6(JtY, dAdv[j]) + dot6(JtB, J[j]);
        }
        if (i > 0) {
            for (int r = 0; r < 6; r++) {
                for (int c = 0; c < 6; c++) { oY[i - 1][r][c] += oY[i][r][c]; B[i - 1][r][c] += B[i][r][c]; }
                of[i - 1][r] += of[i][r];
            }
        }
    }
    for (int i = 0; i < NJ; i++) for (int j = 0; j < i; j++) M[i * NJ + j] = M[j * NJ + i];
}

// pinocchio::crba, via composite world-frame inertias
void crba(const double* q, double* M) {
    double oR[NJ][9], op[NJ][3];
    fk(q, oR, op);
    static thread_local M6 Y[NJ];
    V6 J[NJ];
    for (int i = 0; i < NJ; i++) {
        world_inertia(i, oR[i], op[i], Y[i]);
        double z[3] = {oR[i][2], oR[i][5], oR[i][8]}, pxz[3];
        cross3(op[i], z, pxz);
        for (int k = 0; k < 3; k++) { J[i][k] = pxz[k]; J[i][k + 3] = z[k]; }
    }
    for (int i = NJ - 1; i > 0; i--) for (int r = 0; r < 6; r++) for (int c = 0; c < 6; c++) Y[i - 1][r][c] += Y[i][r][c];
    for (int j = 0; j < NJ; j++) {
        V6 F; mv6(Y[j], J[j], F);
        for (int i = 0; i <= j; i++) { M[i * NJ + j] = dot6(J[i], F); M[j * NJ + i] = M[i * NJ + j]; }
    }
}

// LOCAL_WORLD_ALIGNED frame Jacobian of a frame rigidly attached to joint 7 at `off`
// (computeFrameJacobian + blockdiag(R,R) rotation, robot_ocp.hpp:147-155, pandaWrapper.cpp:65-73)
void frame_jacobian(const double* q, const double* off, double* Jout /*6x7*/, double* pos /*3*/) {
    double oR[NJ][9], op[NJ][3], t[3], pf[3];
    fk(q, oR, op);
    matvec3(oR[6], off, t);
    pf[0] = op[6][0] + t[0]; pf[1] = op[6][1] + t[1]; pf[2] = op[6][2] + t[2];
    if (pos) { pos[0] = pf[0]; pos[1] = pf[1]; pos[2] = pf[2]; }
    if (!Jout) return;
    for (int i = 0; i < NJ; i++) {
        double z[3] = {oR[i][2], oR[i][5], oR[i][8]};
        double d[3] = {pf[0] - op[i][0], pf[1] - op[i][1], pf[2] - op[i][2]}, zxd[3];
        cross3(z, d, zxd);
        for (int k = 0; k < 3; k++) { Jout[k * NJ + i] = zxd[k]; Jout[(k + 3) * NJ + i] = z[k]; }
    }
}

// minTime_ocp::evalConstraints<double> (robot_ocp.hpp:80-96) and <ad_scalar_t> (robot_ocp.hpp:98-163)
void node_constraints(const double* q, const double* v, const double* a, double* g, double* Jg) {
    double pos[3];
    if (!Jg) {
        rnea(q, v, a, g, PANDA_GRAVITY);
        frame_jacobian(q, PANDA_TOOL_OFFSET, nullptr, pos);
        g[7] = pos[2];
        return;
    }
    double tau[NJ], dq[49], dv[49], M[49], J[42];
    rnea_derivatives(q, v, a, tau, dq, dv, M);
    frame_jacobian(q, PANDA_TOOL_OFFSET, J, pos);
    for (int i = 0; i < NJ; i++) {
        g[i] = tau[i];
        double* row = Jg + i * 22;
        double dt = 0;   // djoint_torque_dtime_f = dv*qdot + triu(M)*qddot   (robot_ocp.hpp:116,124)
        for (int j = 0; j < NJ; j++) {
            row[j] = dq[i * NJ + j];
            row[7 + j] = dv[i * NJ + j];
            row[14 + j] = M[i * NJ + j];
            dt += dv[i * NJ + j] * v[j];
        }
        for (int j = i; j < NJ; j++) dt += M[i * NJ + j] * a[j];
        row[21] = dt;
    }
    g[7] = pos[2];
    double* row = Jg + 7 * 22;
    for (int j = 0; j < 22; j++) row[j] = 0;
    for (int j = 0; j < NJ; j++) row[j] = J[2 * NJ + j];
}

// ------------------------------------------------------------------------------------------------
// Transcription
// ------------------------------------------------------------------------------------------------
struct Dims {
    int po, ns, nn, ncol, nvar, neq, nineq, na, ndual;
    double ts;
    double D[8][8];     // differentiation matrix, ascending Lobatto nodes
    double xi[8];       // local nodes in [-1,1], ascending
    explicit Dims(const orc_settings* s) {
        po = s->poly_order; ns = s->num_seg;
        nn = po * ns + 1; ncol = nn - 1; nvar = nn * 21 + 1; neq = 14 * ncol; nineq = 8 * nn; na = neq + nineq; ndual = na + nvar;
        ts = 1.0 / (2.0 * ns);
        const double pi = 3.14159265358979323846;
        for (int j = 0; j <= po; j++) xi[j] = -std::cos(pi * j / po);
        if (po == 3) { xi[0] = -1.0; xi[1] = -0.5; xi[2] = 0.5; xi[3] = 1.0; }   // exact values
        for (int i = 0; i <= po; i++) {
            double ci = (i == 0 || i == po) ? 2.0 : 1.0;
            double rowsum = 0;
            for (int j = 0; j <= po; j++) {
                if (i == j) continue;
                double cj = (j == 0 || j == po) ? 2.0 : 1.0;
                double sgn = ((i + j) % 2) ? -1.0 : 1.0;
                D[i][j] = (ci / cj) * sgn / (xi[i] - xi[j]);
                rowsum += D[i][j];
            }
            D[i][i] = -rowsum;
        }
        if (po == 3) {   // closed-form entries (reversed to ascending order)
            const double d3[4][4] = {{-19.0 / 6, 4, -4.0 / 3, 0.5}, {-1, 1.0 / 3, 1, -1.0 / 3}, {1.0 / 3, -1, -1.0 / 3, 1}, {-0.5, 4.0 / 3, -4, 19.0 / 6}};
            for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) D[i][j] = d3[i][j];
        }
    }
    int ix(int k, int c) const { return 14 * k + c; }
    int iu(int k, int j) const { return 14 * nn + 7 * k + j; }
    int ip() const { return nvar - 1; }
    // node-major permuted index: w_k = (q,v,u) 21 each, then p
    int nm_x(int k, int c) const { return 21 * k + c; }
    int nm_u(int k, int j) const { return 21 * k + 14 + j; }
};

void eval_constraints(const Dims& d, const double* z, double* c_eq, double* g) {
    const double p = z[d.ip()];
    for (int s = 0; s < d.ns; s++)
        for (int j = 0; j < d.po; j++) {
            int k = s * d.po + j;
            for (int c = 0; c < 14; c++) {
                double acc = 0;
                for (int l = 0; l <= d.po; l++) acc += d.D[j][l] * z[d.ix(s * d.po + l, c)];
                double f = (c < 7) ? z[d.ix(k, 7 + c)] : z[d.iu(k, c - 7)];
                c_eq[14 * k + c] = acc - d.ts * p * f;
            }
        }
    if (g)
        for (int k = 0; k < d.nn; k++) node_constraints(z + d.ix(k, 0), z + d.ix(k, 7), z + d.iu(k, 0), g + 8 * k, nullptr);
}

// linearisation_sparse_impl + hessian_regularisation_sparse_impl (polympc_redef.hpp:57-70), SURVEY A.3
void linearise(const Dims& d, double reg, const double* z, const double* lam, double* c_eq, double* g, double* Jg, double* Hd, double* Ha) {
    eval_constraints(d, z, c_eq, nullptr);
    for (int k = 0; k < d.nn; k++) node_constraints(z + d.ix(k, 0), z + d.ix(k, 7), z + d.iu(k, 0), g + 8 * k, Jg + k * 8 * 22);
    for (int i = 0; i < d.nvar; i++) { Hd[i] = 0; Ha[i] = 0; }
    // Lagrangian Hessian: only d2/dp dv and d2/dp du of the collocation rows (inequality 2nd derivatives are zero,
    // robot_ocp.hpp:165-168; cost Hessian zero, robot_ocp.hpp:193-213)
    for (int k = 0; k < d.ncol; k++)
        for (int j = 0; j < 7; j++) {
            Ha[d.ix(k, 7 + j)] = -d.ts * lam[14 * k + j];
            Ha[d.iu(k, j)] = -d.ts * lam[14 * k + 7 + j];
        }
    // Gershgorin (polympc_redef.hpp:57-70): column i off-diagonal sum is |arrow_i| (or sum of |arrow| for p)
    double sum = 0;
    for (int i = 0; i < d.nvar - 1; i++) {
        double ri = std::fabs(Ha[i]), aii = Hd[i];
        if (aii - ri <= 0) Hd[i] += (ri - aii) + reg;
        sum += std::fabs(Ha[i]);
    }
    { int i = d.ip(); double aii = Hd[i]; if (aii - sum <= 0) Hd[i] += (sum - aii) + reg; }
}

// A * x for the QP Jacobian at linearisation point z (rows: eq then ineq), SURVEY A.3
void A_mul(const Dims& d, const double* z, const double* Jg, const double* x, double* y) {
    const double p = z[d.ip()], xp = x[d.ip()];
    for (int s = 0; s < d.ns; s++)
        for (int j = 0; j < d.po; j++) {
            int k = s * d.po + j;
            for (int c = 0; c < 14; c++) {
                double acc = 0;
                for (int l = 0; l <= d.po; l++) acc += d.D[j][l] * x[d.ix(s * d.po + l, c)];
                int fi = (c < 7) ? d.ix(k, 7 + c) : d.iu(k, c - 7);
                acc += -d.ts * p * x[fi];
                acc += -d.ts * z[fi] * xp;
                y[14 * k + c] = acc;
            }
        }
    for (int k = 0; k < d.nn; k++)
        for (int r = 0; r < 8; r++) {
            const double* row = Jg + (k * 8 + r) * 22;
            double acc = 0;
            for (int c = 0; c < 14; c++) acc += row[c] * x[d.ix(k, c)];
            for (int c = 0; c < 7; c++) acc += row[14 + c] * x[d.iu(k, c)];
            acc += row[21] * xp;
            y[d.neq + 8 * k + r] = acc;
        }
}

// A^T * y
void AT_mul(const Dims& d, const double* z, const double* Jg, const double* y, double* x) {
    const double p = z[d.ip()];
    for (int i = 0; i < d.nvar; i++) x[i] = 0;
    double xp = 0;
    for (int s = 0; s < d.ns; s++)
        for (int j = 0; j < d.po; j++) {
            int k = s * d.po + j;
            for (int c = 0; c < 14; c++) {
                double yy = y[14 * k + c];
                for (int l = 0; l <= d.po; l++) x[d.ix(s * d.po + l, c)] += d.D[j][l] * yy;
                int fi = (c < 7) ? d.ix(k, 7 + c) : d.iu(k, c - 7);
                x[fi] += -d.ts * p * yy;
                xp += -d.ts * z[fi] * yy;
            }
        }
    for (int k = 0; k < d.nn; k++)
        for (int r = 0; r < 8; r++) {
            const double* row = Jg + (k * 8 + r) * 22;
            double yy = y[d.neq + 8 * k + r];
            for (int c = 0; c < 14; c++) x[d.ix(k, c)] += row[c] * yy;
            for (int c = 0; c < 7; c++) x[d.iu(k, c)] += row[14 + c] * yy;
            xp += row[21] * yy;
        }
    x[d.ip()] += xp;
}

// dense A (na x nvar) -- for S assembly and for tests
void A_dense(const Dims& d, const double* z, const double* Jg, std::vector<double>& A) {
    A.assign((size_t)d.na * d.nvar, 0.0);
    std::vector<double> e(d.nvar, 0.0), col(d.na);
    for (int i = 0; i < d.nvar; i++) {
        e[i] = 1.0;
        A_mul(d, z, Jg, e.data(), col.data());
        for (int r = 0; r < d.na; r++) A[(size_t)r * d.nvar + i] = col[r];
        e[i] = 0.0;
    }
}

// ------------------------------------------------------------------------------------------------
// Envelope (skyline) Cholesky of the reduced KKT matrix in node-major order
// ------------------------------------------------------------------------------------------------
struct Skyline {
    int n = 0;
    std::vector<int> first;
    std::vector<double> L;   // dense n x n storage, only the envelope is touched
    void factor(const std::vector<double>& S, int n_) {
        n = n_; L = S; first.assign(n, 0);
        for (int i = 0; i < n; i++) { int f = i; for (int j = 0; j < i; j++) if (S[(size_t)i * n + j] != 0.0) { f = j; break; } first[i] = f; }
        for (int i = 0; i < n; i++) {
            double* Li = &L[(size_t)i * n];
            for (int j = first[i]; j < i; j++) {
                const double* Lj = &L[(size_t)j * n];
                double s = Li[j];
                int k0 = std::max(first[i], first[j]);
                for (int k = k0; k < j; k++) s -= Li[k] * Lj[k];
                Li[j] = s / Lj[j];
            }
            double s = Li[i];
            for (int k = first[i]; k < i; k++) s -= Li[k] * Li[k];
            Li[i] = std::sqrt(s);
        }
    }
    void solve(double* b) const {
        for (int i = 0; i < n; i++) {
            const double* Li = &L[(size_t)i * n];
            double s = b[i];
            for (int k = first[i]; k < i; k++) s -= Li[k] * b[k];
            b[i] = s / Li[i];
        }
        for (int i = n - 1; i >= 0; i--) {
            const double* Li = &L[(size_t)i * n];
            double xi = b[i] / Li[i];
            b[i] = xi;
            for (int k = first[i]; k < i; k++) b[k] -= Li[k] * xi;
        }
    }
};

struct QPResult { int status; int iters; double rp, rd, ep, ed; };

// polympc boxADMM::solve.  KKT is solved in its reduced SPD form
//   (H + sigma I + diag(rho_box) + A^T diag(rho_A) A) x~ = sigma x - h + rho_b z_b - y_b + A^T(rho_A z_A - y_A)
//   nu = rho_A (A x~ - z_A) + y_A ,  z~_A = z_A + (nu - y_A)/rho_A
// which is the block elimination of the quasi-definite KKT system the reference factorises with SimplicialLDLT.
QPResult qp_admm(const Dims& d, const orc_settings* st, const double* zlin, const double* Jg, const double* Hd, const double* Ha,
                 const double* h, const double* lA, const double* uA, const double* lb, const double* ub, double* xout, double* yout) {
    const int n = d.nvar, m = d.na, ip = d.ip();
    // ORC_EXP_RHO_EQ: experiment knob of tools/calibrate_admm.py only (equality-row penalty factor); default 1e3
    static const double RHO_EQ = std::getenv("ORC_EXP_RHO_EQ") ? std::atof(std::getenv("ORC_EXP_RHO_EQ")) : 1e3;
    const double RHO_TOL = 1e-4, RHO_MIN = 1e-6, LOOSE = 1e16;
    double rho = st->qp_rho;
    const double sigma = st->qp_sigma, alpha = st->qp_alpha;
    std::vector<double> rhoA(m), rhoB(n);
    auto set_rho = [&](double r) {
        for (int i = 0; i < m; i++) rhoA[i] = (lA[i] < -LOOSE && uA[i] > LOOSE) ? RHO_MIN : (std::fabs(uA[i] - lA[i]) < RHO_TOL ? RHO_EQ * r : r);
        for (int i = 0; i < n; i++) rhoB[i] = (lb[i] < -LOOSE && ub[i] > LOOSE) ? RHO_MIN : (std::fabs(ub[i] - lb[i]) < RHO_TOL ? RHO_EQ * r : r);
    };
    set_rho(rho);
    // node-major permutation
    std::vector<int> perm(n);
    for (int k = 0; k < d.nn; k++) { for (int c = 0; c < 14; c++) perm[d.ix(k, c)] = d.nm_x(k, c); for (int j = 0; j < 7; j++) perm[d.iu(k, j)] = d.nm_u(k, j); }
    perm[ip] = n - 1;
    std::vector<double> Ad;
    A_dense(d, zlin, Jg, Ad);
    // sparse row lists of A for the assembly
    std::vector<std::vector<std::pair<int, double>>> rows(m);
    for (int r = 0; r < m; r++) for (int c = 0; c < n; c++) { double v = Ad[(size_t)r * n + c]; if (v != 0.0) rows[r].push_back({c, v}); }
    Skyline chol;
    std::vector<double> S((size_t)n * n);
    auto assemble = [&]() {
        std::fill(S.begin(), S.end(), 0.0);
        for (int i = 0; i < n; i++) S[(size_t)perm[i] * n + perm[i]] = Hd[i] + sigma + rhoB[i];
        for (int i = 0; i < n - 1; i++) if (Ha[i] != 0.0) { S[(size_t)perm[ip] * n + perm[i]] += Ha[i]; S[(size_t)perm[i] * n + perm[ip]] += Ha[i]; }
        for (int r = 0; r < m; r++)
            for (auto& a : rows[r]) for (auto& b : rows[r]) S[(size_t)perm[a.first] * n + perm[b.first]] += rhoA[r] * a.second * b.second;
        chol.factor(S, n);
    };
    assemble();
    std::vector<double> x(n, 0.0), zA(m, 0.0), zB(n, 0.0), yA(m, 0.0), yB(n, 0.0);
    std::vector<double> rhs(n), w(m), t(n), xt(n), Ax(m), b(n), Hx(n), Aty(n);
    QPResult res{2, 0, 0, 0, 0, 0};
    int iter;
    for (iter = 1; iter <= st->qp_max_iter; iter++) {
        for (int i = 0; i < m; i++) w[i] = rhoA[i] * zA[i] - yA[i];
        AT_mul(d, zlin, Jg, w.data(), t.data());
        for (int i = 0; i < n; i++) rhs[i] = sigma * x[i] - h[i] + rhoB[i] * zB[i] - yB[i] + t[i];
        for (int i = 0; i < n; i++) b[perm[i]] = rhs[i];
        chol.solve(b.data());
        for (int i = 0; i < n; i++) xt[i] = b[perm[i]];
        A_mul(d, zlin, Jg, xt.data(), Ax.data());
        // z~_A = z_A + (nu - y_A)/rho_A with nu = rho_A (A x~ - z_A) + y_A   ->  A x~
        for (int i = 0; i < n; i++) x[i] = alpha * xt[i] + (1 - alpha) * x[i];
        for (int i = 0; i < m; i++) {
            double zh = alpha * Ax[i] + (1 - alpha) * zA[i];
            double zn = std::min(std::max(zh + yA[i] / rhoA[i], lA[i]), uA[i]);
            yA[i] += rhoA[i] * (zh - zn);
            zA[i] = zn;
        }
        for (int i = 0; i < n; i++) {
            double zh = alpha * xt[i] + (1 - alpha) * zB[i];
            double zn = std::min(std::max(zh + yB[i] / rhoB[i], lb[i]), ub[i]);
            yB[i] += rhoB[i] * (zh - zn);
            zB[i] = zn;
        }
        if (st->qp_check_termination > 0 && iter % st->qp_check_termination == 0) {
            A_mul(d, zlin, Jg, x.data(), Ax.data());
            AT_mul(d, zlin, Jg, yA.data(), Aty.data());
            double nAx = 0, nx = 0, nz = 0, rp = 0, rd = 0, nHx = 0, nAty = 0, nh = 0;
            for (int i = 0; i < n; i++) Hx[i] = Hd[i] * x[i];
            for (int i = 0; i < n; i++) if (i != ip) { Hx[i] += Ha[i] * x[ip]; Hx[ip] += Ha[i] * x[i]; }   // H = diag + p-arrow
            for (int i = 0; i < m; i++) { nAx = std::max(nAx, std::fabs(Ax[i])); nz = std::max(nz, std::fabs(zA[i])); rp = std::max(rp, std::fabs(Ax[i] - zA[i])); }
            for (int i = 0; i < n; i++) {
                nx = std::max(nx, std::fabs(x[i])); nz = std::max(nz, std::fabs(zB[i])); rp = std::max(rp, std::fabs(x[i] - zB[i]));
                double aty = Aty[i] + yB[i];
                nAty = std::max(nAty, std::fabs(aty)); nHx = std::max(nHx, std::fabs(Hx[i])); nh = std::max(nh, std::fabs(h[i]));
                rd = std::max(rd, std::fabs(Hx[i] + h[i] + aty));
            }
            double ep = st->qp_eps_abs + st->qp_eps_rel * std::max(std::max(nAx, nx), nz);
            double ed = st->qp_eps_abs + st->qp_eps_rel * std::max(std::max(nHx, nAty), nh);
            res.rp = rp; res.rd = rd; res.ep = ep; res.ed = ed;
            if (rp <= ep && rd <= ed) { res.status = 1; break; }
            if (st->qp_adaptive_rho && iter % st->qp_adaptive_rho_interval == 0) {
                double np_ = std::max(std::max(nAx, nx), nz), nd_ = std::max(std::max(nHx, nAty), nh);
                double rn = rho * std::sqrt((rp / (np_ + 1e-10)) / (rd / (nd_ + 1e-10)));
                rn = std::min(std::max(rn, 1e-6), 1e6);
                if (rn > rho * st->qp_adaptive_rho_tolerance || rn < rho / st->qp_adaptive_rho_tolerance) { rho = rn; set_rho(rho); assemble(); }
            }
        }
    }
    res.iters = std::min(iter, st->qp_max_iter);
    for (int i = 0; i < n; i++) xout[i] = x[i];
    for (int i = 0; i < m; i++) yout[i] = yA[i];
    for (int i = 0; i < n; i++) yout[m + i] = yB[i];
    return res;
}

// MotionPlanner bounds plumbing (motionPlanner.cpp:27-100; quirk Q7 dropped: pins applied last)
void make_bounds(const Dims& d, const orc_limits* lim, const double* x0, const double* xT, double* lbx, double* ubx, double* lbg, double* ubg) {
    const double mp = lim->margins[0], mv = lim->margins[1], ma = lim->margins[2], mt = lim->margins[3];
    for (int k = 0; k < d.nn; k++)
        for (int j = 0; j < 7; j++) {
            double safety = (1 - mp) * (lim->qmax[j] - lim->qmin[j]) / 2;
            lbx[d.ix(k, j)] = lim->qmin[j] + safety; ubx[d.ix(k, j)] = lim->qmax[j] - safety;
            lbx[d.ix(k, 7 + j)] = -mv * lim->vmax[j]; ubx[d.ix(k, 7 + j)] = mv * lim->vmax[j];
            lbx[d.iu(k, j)] = -ma * lim->amax[j]; ubx[d.iu(k, j)] = ma * lim->amax[j];
        }
    lbx[d.ip()] = lim->p_lb; ubx[d.ip()] = lim->p_ub;
    for (int c = 0; c < 14; c++) {
        lbx[d.ix(0, c)] = x0[c]; ubx[d.ix(0, c)] = x0[c];
        lbx[d.ix(d.nn - 1, c)] = xT[c] - lim->final_eps; ubx[d.ix(d.nn - 1, c)] = xT[c] + lim->final_eps;
    }
    for (int j = 0; j < 7; j++) { lbg[j] = -mt * lim->taumax[j]; ubg[j] = mt * lim->taumax[j]; }
    lbg[7] = lim->min_height; ubg[7] = INF;
}

struct Viol { double l1, linf; };
Viol violation(const Dims& d, const double* z, const double* lbx, const double* ubx, const double* lbg, const double* ubg, double* c_eq, double* g) {
    eval_constraints(d, z, c_eq, g);
    Viol v{0, 0};
    auto add = [&](double e) { if (e > 0) { v.l1 += e; v.linf = std::max(v.linf, e); } };
    for (int i = 0; i < d.neq; i++) add(std::fabs(c_eq[i]));
    for (int k = 0; k < d.nn; k++) for (int r = 0; r < 8; r++) { add(lbg[r] - g[8 * k + r]); add(g[8 * k + r] - ubg[r]); }
    for (int i = 0; i < d.nvar; i++) { add(lbx[i] - z[i]); add(z[i] - ubx[i]); }
    return v;
}

int solve(const orc_settings* st, const orc_limits* lim, const double* x0, const double* xT, const double* gx, const double* gu, double gT,
          const double* glam, double* sol_x, double* sol_u, double* sol_T, double* sol_lam, orc_stats* stats) {
    Dims d(st);
    const int n = d.nvar, m = d.na;
    std::vector<double> z(n), lam(d.ndual, 0.0), lbx(n), ubx(n), c_eq(d.neq), g(d.nineq), Jg((size_t)d.nn * 8 * 22), Hd(n), Ha(n), h(n, 0.0);
    std::vector<double> lA(m), uA(m), lb(n), ub(n), p(n), lamqp(d.ndual), dlam(d.ndual), zs(n), ce2(d.neq), g2(d.nineq);
    double lbg[8], ubg[8];
    std::memcpy(z.data(), gx, sizeof(double) * 14 * d.nn);
    std::memcpy(z.data() + 14 * d.nn, gu, sizeof(double) * 7 * d.nn);
    z[d.ip()] = gT;
    if (glam) std::memcpy(lam.data(), glam, sizeof(double) * d.ndual);
    make_bounds(d, lim, x0, xT, lbx.data(), ubx.data(), lbg, ubg);
    h[d.ip()] = 1.0;   // gradient of the Mayer cost p (robot_ocp.hpp:206)
    orc_stats S; std::memset(&S, 0, sizeof S);
    S.status = 2;
    linearise(d, st->hess_reg, z.data(), lam.data(), c_eq.data(), g.data(), Jg.data(), Hd.data(), Ha.data());
    for (int it = 1; it <= st->sqp_max_iter; it++) {
        for (int i = 0; i < d.neq; i++) { lA[i] = -c_eq[i]; uA[i] = -c_eq[i]; }
        for (int k = 0; k < d.nn; k++) for (int r = 0; r < 8; r++) { lA[d.neq + 8 * k + r] = lbg[r] - g[8 * k + r]; uA[d.neq + 8 * k + r] = ubg[r] - g[8 * k + r]; }
        for (int i = 0; i < n; i++) { lb[i] = lbx[i] - z[i]; ub[i] = ubx[i] - z[i]; }
        QPResult qr = qp_admm(d, st, z.data(), Jg.data(), Hd.data(), Ha.data(), h.data(), lA.data(), uA.data(), lb.data(), ub.data(), p.data(), lamqp.data());
        S.qp_iters_total += qr.iters;
        if (it <= 4) { S.qp_status[it - 1] = qr.status; S.qp_iters[it - 1] = qr.iters; }
        for (int i = 0; i < d.ndual; i++) dlam[i] = lamqp[i] - lam[i];
        // step_size_selection_impl (polympc_redef.hpp:73-121); mu = |lambda_qp|_inf
        double mu = 0; for (int i = 0; i < d.ndual; i++) mu = std::max(mu, std::fabs(lamqp[i]));
        Viol v0 = violation(d, z.data(), lbx.data(), ubx.data(), lbg, ubg, ce2.data(), g2.data());
        double cost = z[d.ip()];
        double phi = cost + mu * v0.l1;
        double gradp = 0; for (int i = 0; i < n; i++) gradp += h[i] * p[i];
        double Dphi = gradp - mu * v0.l1;
        double alpha = 1.0; int trials = 0;
        for (int i = 1; i < st->line_search_max_iter; i++) {
            trials++;
            for (int k = 0; k < n; k++) zs[k] = z[k] + alpha * p[k];
            Viol vs = violation(d, zs.data(), lbx.data(), ubx.data(), lbg, ubg, ce2.data(), g2.data());
            double phis = zs[d.ip()] + mu * vs.l1;
            if (phis <= phi + alpha * st->sqp_eta * Dphi) break;
            alpha = st->sqp_tau * alpha;
        }
        for (int i = 0; i < n; i++) z[i] += alpha * p[i];
        for (int i = 0; i < d.ndual; i++) lam[i] += alpha * dlam[i];
        double pn = 0, dn = 0;
        for (int i = 0; i < n; i++) pn = std::max(pn, std::fabs(p[i]));
        for (int i = 0; i < d.ndual; i++) dn = std::max(dn, std::fabs(dlam[i]));
        Viol vn = violation(d, z.data(), lbx.data(), ubx.data(), lbg, ubg, ce2.data(), g2.data());
        S.sqp_iters = it; S.primal_step_norm = alpha * pn; S.dual_step_norm = alpha * dn; S.max_violation = vn.linf; S.l1_violation = vn.l1; S.cost = z[d.ip()];
        if (it <= 4) { S.alpha[it - 1] = alpha; S.ls_trials[it - 1] = trials; }
        if (alpha * pn <= st->sqp_eps_prim && alpha * dn <= st->sqp_eps_dual && vn.linf <= st->sqp_eps_prim) { S.status = 1; break; }
        if (it < st->sqp_max_iter)   // update_linearisation_sparse_impl (polympc_redef.hpp:141-147); skipped after the last step (A.8-U5)
            linearise(d, st->hess_reg, z.data(), lam.data(), c_eq.data(), g.data(), Jg.data(), Hd.data(), Ha.data());
    }
    std::memcpy(sol_x, z.data(), sizeof(double) * 14 * d.nn);
    std::memcpy(sol_u, z.data() + 14 * d.nn, sizeof(double) * 7 * d.nn);
    *sol_T = z[d.ip()];
    if (sol_lam) std::memcpy(sol_lam, lam.data(), sizeof(double) * d.ndual);
    if (stats) *stats = S;
    return 0;
}

// polympc solution_x_at / solution_u_at: piecewise Lagrange interpolation on the segment containing t (motionPlanner.hpp:106-108)
void interp(const Dims& d, const double* sol_x, const double* sol_u, double t, double* q, double* v, double* a) {
    t = std::min(std::max(t, 0.0), 1.0);
    int s = std::min((int)std::floor(t * d.ns), d.ns - 1);
    double xi = 2.0 * (t * d.ns - s) - 1.0;
    double w[8];
    for (int j = 0; j <= d.po; j++) {
        double num = 1, den = 1;
        for (int l = 0; l <= d.po; l++) if (l != j) { num *= (xi - d.xi[l]); den *= (d.xi[j] - d.xi[l]); }
        w[j] = num / den;
    }
    for (int c = 0; c < 7; c++) { q[c] = 0; v[c] = 0; a[c] = 0; }
    for (int j = 0; j <= d.po; j++) {
        int k = s * d.po + j;
        for (int c = 0; c < 7; c++) { q[c] += w[j] * sol_x[14 * k + c]; v[c] += w[j] * sol_x[14 * k + 7 + c]; a[c] += w[j] * sol_u[7 * k + c]; }
    }
}

}  // namespace

// ================================================================================================
extern "C" {

void orc_default_settings(orc_settings* s) {
    s->poly_order = 3; s->num_seg = 6; s->sqp_max_iter = 2; s->qp_max_iter = 700; s->line_search_max_iter = 10;
    s->qp_check_termination = 25; s->qp_adaptive_rho = 0; s->qp_adaptive_rho_interval = 25;
    s->qp_eps_abs = 1e-3; s->qp_eps_rel = 1e-3; s->qp_rho = 0.1; s->qp_sigma = 1e-6; s->qp_alpha = 1.0; s->qp_adaptive_rho_tolerance = 5;
    s->sqp_eps_prim = 1e-3; s->sqp_eps_dual = 1e-3; s->sqp_tau = 0.5; s->sqp_eta = 0.25; s->hess_reg = 0.001;
}
void orc_default_limits(orc_limits* l) {   // pandaWrapper.hpp:29-40, margins 1.0 (motionPlanner.cpp:24)
    const double qmin[7] = {-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973};
    const double qmax[7] = {2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973};
    const double vmax[7] = {2.1750, 2.1750, 2.1750, 2.1750, 2.6100, 2.6100, 2.6100};
    const double amax[7] = {15.0, 7.5, 10.0, 12.5, 15.0, 20.0, 20.0};
    const double jmax[7] = {7500, 3750, 5000, 6250, 7500, 10000, 10000};
    const double tmax[7] = {87, 87, 87, 87, 12, 12, 12};
    for (int i = 0; i < 7; i++) { l->qmin[i] = qmin[i]; l->qmax[i] = qmax[i]; l->vmax[i] = vmax[i]; l->amax[i] = amax[i]; l->jmax[i] = jmax[i]; l->taumax[i] = tmax[i]; }
    for (int i = 0; i < 5; i++) l->margins[i] = 1.0;
    l->min_height = 0.05; l->final_eps = 1e-2; l->p_lb = 0.0; l->p_ub = 10.0;
}

void orc_fk(const double* q, double* R_out, double* p_out) {
    double R[NJ][9], p[NJ][3];
    fk(q, R, p);
    std::memcpy(R_out, R, sizeof R); std::memcpy(p_out, p, sizeof p);
}
void orc_rnea(const double* q, const double* v, const double* a, double* tau) { rnea(q, v, a, tau, PANDA_GRAVITY); }
void orc_rnea_zero_gravity(const double* q, const double* v, const double* a, double* tau) { rnea(q, v, a, tau, 0.0); }
void orc_rnea_derivatives(const double* q, const double* v, const double* a, double* tau, double* dq, double* dv, double* M) { rnea_derivatives(q, v, a, tau, dq, dv, M); }
void orc_crba(const double* q, double* M) { crba(q, M); }
void orc_tool_height(const double* q, double* z, double* Jz) {
    double J[42], pos[3];
    frame_jacobian(q, PANDA_TOOL_OFFSET, Jz ? J : nullptr, pos);
    *z = pos[2];
    if (Jz) for (int j = 0; j < 7; j++) Jz[j] = J[2 * 7 + j];
}
void orc_frame_jacobian_world_aligned(const double* q, const double* offset, double* J) { frame_jacobian(q, offset, J, nullptr); }
void orc_node_constraints(const double* q, const double* v, const double* a, double* g, double* Jg) { node_constraints(q, v, a, g, Jg); }

int orc_num_nodes(const orc_settings* s) { return Dims(s).nn; }
int orc_num_var(const orc_settings* s) { return Dims(s).nvar; }
int orc_num_eq(const orc_settings* s) { return Dims(s).neq; }
int orc_num_ineq(const orc_settings* s) { return Dims(s).nineq; }
void orc_time_nodes(const orc_settings* s, double* t) {
    Dims d(s);
    for (int sg = 0; sg < d.ns; sg++) for (int j = 0; j <= d.po; j++) t[sg * d.po + j] = ((double)sg + 0.5 * (1.0 + d.xi[j])) / d.ns;
}
void orc_diff_matrix(const orc_settings* s, double* D) {
    Dims d(s);
    for (int i = 0; i <= d.po; i++) for (int j = 0; j <= d.po; j++) D[i * (d.po + 1) + j] = d.D[i][j];
}
void orc_bounds(const orc_settings* s, const orc_limits* lim, const double* x0, const double* xT, double* lbx, double* ubx, double* lbg, double* ubg) {
    Dims d(s); make_bounds(d, lim, x0, xT, lbx, ubx, lbg, ubg);
}
void orc_eval_constraints(const orc_settings* s, const double* z, double* c_eq, double* g) { Dims d(s); eval_constraints(d, z, c_eq, g); }
void orc_linearise(const orc_settings* s, const double* z, const double* lam, double* c_eq, double* g, double* Jg, double* Hd, double* Ha) {
    Dims d(s); linearise(d, s->hess_reg, z, lam, c_eq, g, Jg, Hd, Ha);
}
void orc_qp_dense(const orc_settings* s, const double* z, const double* lam, double* A, double* H) {
    Dims d(s);
    std::vector<double> c_eq(d.neq), g(d.nineq), Jg((size_t)d.nn * 8 * 22), Hd(d.nvar), Ha(d.nvar), Ad;
    linearise(d, s->hess_reg, z, lam, c_eq.data(), g.data(), Jg.data(), Hd.data(), Ha.data());
    A_dense(d, z, Jg.data(), Ad);
    std::memcpy(A, Ad.data(), sizeof(double) * Ad.size());
    for (size_t i = 0; i < (size_t)d.nvar * d.nvar; i++) H[i] = 0;
    for (int i = 0; i < d.nvar; i++) H[(size_t)i * d.nvar + i] = Hd[i];
    for (int i = 0; i < d.nvar - 1; i++) { H[(size_t)d.ip() * d.nvar + i] = Ha[i]; H[(size_t)i * d.nvar + d.ip()] = Ha[i]; }
}
int orc_qp_admm(const orc_settings* s, const double* z_lin, const double* Jg, const double* Hdiag, const double* Harrow, const double* h,
                const double* lA, const double* uA, const double* lb, const double* ub, double* x_out, double* y_out, int* iters_out, double* res_out) {
    Dims d(s);
    QPResult r = qp_admm(d, s, z_lin, Jg, Hdiag, Harrow, h, lA, uA, lb, ub, x_out, y_out);
    if (iters_out) *iters_out = r.iters;
    if (res_out) { res_out[0] = r.rp; res_out[1] = r.rd; res_out[2] = r.ep; res_out[3] = r.ed; }
    return r.status;
}
int orc_solve(const orc_settings* s, const orc_limits* lim, const double* x0, const double* xT, const double* guess_x, const double* guess_u,
              double guess_T, const double* guess_lam, double* sol_x, double* sol_u, double* sol_T, double* sol_lam, orc_stats* stats) {
    return solve(s, lim, x0, xT, guess_x, guess_u, guess_T, guess_lam, sol_x, sol_u, sol_T, sol_lam, stats);
}
int orc_solve_batch(const orc_settings* s, const orc_limits* lim, int n, int nthreads, const double* x0, const double* xT, const double* gx,
                    const double* gu, const double* gT, double* sol_x, double* sol_u, double* sol_T, orc_stats* stats) {
    Dims d(s);
    std::atomic<int> next(0);
    auto work = [&]() {
        for (;;) {
            int i = next.fetch_add(1);
            if (i >= n) break;
            solve(s, lim, x0 + 14 * (size_t)i, xT + 14 * (size_t)i, gx + (size_t)14 * d.nn * i, gu + (size_t)7 * d.nn * i, gT[i], nullptr,
                  sol_x + (size_t)14 * d.nn * i, sol_u + (size_t)7 * d.nn * i, sol_T + i, nullptr, stats ? stats + i : nullptr);
        }
    };
    if (nthreads <= 1) { work(); return 0; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(work);
    for (auto& t : th) t.join();
    return 0;
}

// get_MPC_trajectory<N> (motionPlanner.hpp:99-116): time(i) = 1.0/N * i, torque by RNEA, time *= T
void orc_sample_solution(const orc_settings* s, const double* sol_x, const double* sol_u, double T, int npts, double* traj) {
    Dims d(s);
    for (int i = 0; i <= npts; i++) {
        double t = 1.0 / npts * i;
        double* row = traj + 29 * i;
        interp(d, sol_x, sol_u, t, row + 1, row + 8, row + 15);
        rnea(row + 1, row + 8, row + 15, row + 22, PANDA_GRAVITY);
        row[0] = t * T;
    }
}
void orc_interp_state(const orc_settings* s, const double* sol_x, const double* sol_u, double t01, double* q, double* v, double* a) {
    Dims d(s); interp(d, sol_x, sol_u, t01, q, v, a);
}
// benchmark.cpp:118-160 (checks) and :164-194 (row contents) for one trajectory
void orc_account_trajectory(const orc_limits* lim, const double* traj, int npts, const double* target, orc_account* acc) {
    for (int f = 0; f < 4; f++) acc->flags[f] = 1;
    for (int c = 0; c < 28; c++) { acc->mins[c] = INF; acc->maxs[c] = -INF; }
    const double dT = traj[29 * npts] / npts;                      // benchmark.cpp:128
    for (int i = 0; i <= npts; i++) {
        const double* row = traj + 29 * i;
        if (i >= 1) {
            const double* prev = row - 29;
            for (int j = 0; j < 7; j++) {
                double jerk = (row[15 + j] - prev[15 + j]) / dT;
                if (std::fabs(jerk) > 10 * lim->jmax[j]) acc->flags[0] = 0;   // benchmark.cpp:135
            }
        }
        double J[42], pos[3], tv[6];
        frame_jacobian(row + 1, PANDA_TOOL_OFFSET, J, pos);
        for (int r = 0; r < 6; r++) { double sum = 0; for (int j = 0; j < 7; j++) sum += J[r * 7 + j] * row[8 + j]; tv[r] = sum; }
        if (std::sqrt(tv[0] * tv[0] + tv[1] * tv[1] + tv[2] * tv[2]) > 1.7) acc->flags[1] = 0;   // benchmark.cpp:144
        if (std::sqrt(tv[3] * tv[3] + tv[4] * tv[4] + tv[5] * tv[5]) > 2.5) acc->flags[2] = 0;   // benchmark.cpp:148
        if (pos[2] < 0) acc->flags[3] = 0;                                                        // benchmark.cpp:156
        for (int c = 0; c < 28; c++) { acc->mins[c] = std::min(acc->mins[c], row[1 + c]); acc->maxs[c] = std::max(acc->maxs[c], row[1 + c]); }
    }
    for (int c = 0; c < 14; c++) acc->final_err[c] = traj[29 * npts + 1 + c] - target[c];
}

}  // extern "C"
