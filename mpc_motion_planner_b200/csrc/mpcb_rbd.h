// Rigid-body dynamics of the 7-DoF Panda for the OCP's inequality constraints.
//
// Replaces the pinocchio calls the reference makes from minTime_ocp::evalConstraints
// (/root/reference/mpc_solver/robot_ocp.hpp:80-96 value path: forwardKinematics + updateFramePlacement + rnea;
//  :98-163 derivative path: computeRNEADerivatives + rnea + crba + computeFrameJacobian rotated to the
//  world-aligned frame).  Model constants are baked from robot_utils/panda-model/panda_arm.urdf
//  (tools/bake_panda_model.py): 7 revolute-Z joints, link8 + 1 kg tool merged into body 7.
//
// All functions are __host__ __device__: the CUDA kernels call them per collocation node; the host planner
// code (problem generator, MotionPlanner helpers) calls the same code on the CPU side of the API.
#pragma once
#include <math.h>
#include "mpcb_common.h"

namespace mpcb {
namespace rbd {

constexpr int NJ = 7;

MPCB_HD void cross3(const double* a, const double* b, double* o) {
    double x = a[1] * b[2] - a[2] * b[1];
    double y = a[2] * b[0] - a[0] * b[2];
    double z = a[0] * b[1] - a[1] * b[0];
    o[0] = x; o[1] = y; o[2] = z;
}
MPCB_HD void mat3_vec(const double* R, const double* v, double* o) {
    double x = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
    double y = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
    double z = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
    o[0] = x; o[1] = y; o[2] = z;
}
MPCB_HD void mat3T_vec(const double* R, const double* v, double* o) {
    double x = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
    double y = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
    double z = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
    o[0] = x; o[1] = y; o[2] = z;
}

// Local joint rotation  Placement_R[i] * Rz(q)  (translation is the constant PLACEMENT_P[i]).
MPCB_HD void joint_rot(int i, double c, double s, double* R) {
    const double* P = PANDA(PANDA_PLACEMENT_R)[i];
    MPCB_UNROLL
    for (int r = 0; r < 3; r++) {
        R[3 * r + 0] = P[3 * r] * c + P[3 * r + 1] * s;
        R[3 * r + 1] = -P[3 * r] * s + P[3 * r + 1] * c;
        R[3 * r + 2] = P[3 * r + 2];
    }
}

// World placements oMi of the 7 joint frames (pinocchio::forwardKinematics).
MPCB_HD void fk_world(const double* q, double (*oR)[9], double (*op)[3]) {
    double Rp[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, pp[3] = {0, 0, 0};
    for (int i = 0; i < NJ; i++) {
        double s, c, Rl[9], t[3];
        sincos(q[i], &s, &c);
        joint_rot(i, c, s, Rl);
        MPCB_UNROLL
        for (int r = 0; r < 3; r++)
            MPCB_UNROLL
            for (int cc = 0; cc < 3; cc++) oR[i][3 * r + cc] = Rp[3 * r] * Rl[cc] + Rp[3 * r + 1] * Rl[3 + cc] + Rp[3 * r + 2] * Rl[6 + cc];
        mat3_vec(Rp, PANDA(PANDA_PLACEMENT_P)[i], t);
        MPCB_UNROLL
        for (int k = 0; k < 3; k++) op[i][k] = pp[k] + t[k];
        MPCB_UNROLL
        for (int k = 0; k < 9; k++) Rp[k] = oR[i][k];
        MPCB_UNROLL
        for (int k = 0; k < 3; k++) pp[k] = op[i][k];
    }
}

// Position of a point rigidly attached to joint 7 (offset in joint-7 frame) -- only the last placement is kept.
MPCB_HD void fk_point7(const double* q, const double* off, double* pos) {
    double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, p[3] = {0, 0, 0};
    for (int i = 0; i < NJ; i++) {
        double s, c, Rl[9], Rn[9], t[3];
        sincos(q[i], &s, &c);
        joint_rot(i, c, s, Rl);
        MPCB_UNROLL
        for (int r = 0; r < 3; r++)
            MPCB_UNROLL
            for (int cc = 0; cc < 3; cc++) Rn[3 * r + cc] = R[3 * r] * Rl[cc] + R[3 * r + 1] * Rl[3 + cc] + R[3 * r + 2] * Rl[6 + cc];
        mat3_vec(R, PANDA(PANDA_PLACEMENT_P)[i], t);
        MPCB_UNROLL
        for (int k = 0; k < 3; k++) p[k] += t[k];
        MPCB_UNROLL
        for (int k = 0; k < 9; k++) R[k] = Rn[k];
    }
    double t[3];
    mat3_vec(R, off, t);
    pos[0] = p[0] + t[0]; pos[1] = p[1] + t[1]; pos[2] = p[2] + t[2];
}

// Spatial inertia times motion, body frame: h = Y (vl; va) with Y = (m, com c, I about com)
MPCB_HD void inertia_mul(int i, const double* vl, const double* va, double* hl, double* ha) {
    const double m = PANDA(PANDA_MASS)[i];
    const double* c = PANDA(PANDA_COM)[i];
    double t[3];
    cross3(va, c, t);
    hl[0] = m * (vl[0] + t[0]); hl[1] = m * (vl[1] + t[1]); hl[2] = m * (vl[2] + t[2]);
    mat3_vec(PANDA(PANDA_ICOM)[i], va, ha);
    cross3(c, hl, t);
    ha[0] += t[0]; ha[1] += t[1]; ha[2] += t[2];
}

// Inverse dynamics tau = RNEA(q, qd, qdd) (pinocchio::rnea; robot_ocp.hpp:91), body-frame recursion,
// plus the tool height z (robot_ocp.hpp:87-91) from the same joint rotations.  g[0..6] = tau, g[7] = z_tool.
MPCB_HD void node_constraints_value(const double* q, const double* qd, const double* qdd, double* g) {
    double cs[NJ][2];
    double fl[NJ][3], fa[NJ][3];
    double vl[3] = {0, 0, 0}, va[3] = {0, 0, 0};
    double al[3] = {0, 0, PANDA(PANDA_GRAVITY)}, aa[3] = {0, 0, 0};
    // world pose accumulation for the tool height
    double oR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, opz[3] = {0, 0, 0};
    for (int i = 0; i < NJ; i++) {
        double s, c, R[9], t[3], d[3];
        sincos(q[i], &s, &c);
        cs[i][0] = c; cs[i][1] = s;
        joint_rot(i, c, s, R);
        const double* p = PANDA(PANDA_PLACEMENT_P)[i];
        // tool height chain
        {
            double Rn[9], tt[3];
            MPCB_UNROLL
            for (int r = 0; r < 3; r++)
                MPCB_UNROLL
                for (int cc = 0; cc < 3; cc++) Rn[3 * r + cc] = oR[3 * r] * R[cc] + oR[3 * r + 1] * R[3 + cc] + oR[3 * r + 2] * R[6 + cc];
            mat3_vec(oR, p, tt);
            MPCB_UNROLL
            for (int k = 0; k < 3; k++) opz[k] += tt[k];
            MPCB_UNROLL
            for (int k = 0; k < 9; k++) oR[k] = Rn[k];
        }
        // v_i = liMi^-1 o v_parent + S qd
        double nvl[3], nva[3], nal[3], naa[3];
        cross3(p, va, t);
        d[0] = vl[0] - t[0]; d[1] = vl[1] - t[1]; d[2] = vl[2] - t[2];
        mat3T_vec(R, d, nvl);
        mat3T_vec(R, va, nva);
        nva[2] += qd[i];
        // a_i = liMi^-1 o a_parent + S qdd + v_i x (S qd)
        cross3(p, aa, t);
        d[0] = al[0] - t[0]; d[1] = al[1] - t[1]; d[2] = al[2] - t[2];
        mat3T_vec(R, d, nal);
        mat3T_vec(R, aa, naa);
        naa[2] += qdd[i];
        {   // v x (0;0,0,qd): linear = vl x (0,0,qd), angular = va x (0,0,qd)
            double sq[3] = {0, 0, qd[i]}, t1[3], t2[3];
            cross3(nvl, sq, t1); cross3(nva, sq, t2);
            nal[0] += t1[0]; nal[1] += t1[1]; nal[2] += t1[2];
            naa[0] += t2[0]; naa[1] += t2[1]; naa[2] += t2[2];
        }
        // f_i = Y a + v x* (Y v)
        double hl[3], ha[3], yl[3], ya[3], t1[3], t2[3];
        inertia_mul(i, nvl, nva, hl, ha);
        inertia_mul(i, nal, naa, yl, ya);
        cross3(nva, hl, t1);
        fl[i][0] = yl[0] + t1[0]; fl[i][1] = yl[1] + t1[1]; fl[i][2] = yl[2] + t1[2];
        cross3(nva, ha, t1); cross3(nvl, hl, t2);
        fa[i][0] = ya[0] + (t1[0] + t2[0]); fa[i][1] = ya[1] + (t1[1] + t2[1]); fa[i][2] = ya[2] + (t1[2] + t2[2]);
        MPCB_UNROLL
        for (int k = 0; k < 3; k++) { vl[k] = nvl[k]; va[k] = nva[k]; al[k] = nal[k]; aa[k] = naa[k]; }
    }
    for (int i = NJ - 1; i >= 0; i--) {
        g[i] = fa[i][2];
        if (i > 0) {
            double R[9], l[3], a[3], t[3];
            joint_rot(i, cs[i][0], cs[i][1], R);
            mat3_vec(R, fl[i], l);
            mat3_vec(R, fa[i], a);
            cross3(PANDA(PANDA_PLACEMENT_P)[i], l, t);
            MPCB_UNROLL
            for (int k = 0; k < 3; k++) { fl[i - 1][k] += l[k]; fa[i - 1][k] += a[k] + t[k]; }
        }
    }
    double t[3];
    mat3_vec(oR, PANDA(PANDA_TOOL_OFFSET), t);
    g[7] = opz[2] + t[2];
}

// ---- 6x6 helpers for the world-frame derivative recursion ----------------------------------------------------
// spatial vectors are (linear[3]; angular[3]).
MPCB_HD void crm_mul(const double* v, const double* m, double* o) {   // v x m   (motion cross motion)
    double t1[3], t2[3], t3[3];
    cross3(v + 3, m, t1); cross3(v, m + 3, t2); cross3(v + 3, m + 3, t3);
    o[0] = t1[0] + t2[0]; o[1] = t1[1] + t2[1]; o[2] = t1[2] + t2[2];
    o[3] = t3[0]; o[4] = t3[1]; o[5] = t3[2];
}
MPCB_HD void crf_mul(const double* v, const double* f, double* o) {   // v x* f  (motion cross force)
    double t1[3], t2[3], t3[3];
    cross3(v + 3, f, t1); cross3(v + 3, f + 3, t2); cross3(v, f, t3);
    o[0] = t1[0]; o[1] = t1[1]; o[2] = t1[2];
    o[3] = t2[0] + t3[0]; o[4] = t2[1] + t3[1]; o[5] = t2[2] + t3[2];
}
MPCB_HD double dot6(const double* a, const double* b) {
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2] + a[3] * b[3] + a[4] * b[4] + a[5] * b[5];
}
MPCB_HD void mv6(const double* A /*36 row-major*/, const double* x, double* y) {
    double t[6];
    MPCB_UNROLL
    for (int i = 0; i < 6; i++) t[i] = dot6(A + 6 * i, x);
    MPCB_UNROLL
    for (int i = 0; i < 6; i++) y[i] = t[i];
}
MPCB_HD void vm6(const double* x, const double* A, double* y) {   // y = x^T A
    double t[6];
    MPCB_UNROLL
    for (int c = 0; c < 6; c++) {
        double s = 0;
        MPCB_UNROLL
        for (int r = 0; r < 6; r++) s += x[r] * A[6 * r + c];
        t[c] = s;
    }
    MPCB_UNROLL
    for (int i = 0; i < 6; i++) y[i] = t[i];
}

// World-frame spatial inertia (about the world origin) of body i placed at (R, p): 6x6 row-major.
MPCB_HD void world_inertia(int i, const double* R, const double* p, double* Y) {
    const double m = PANDA(PANDA_MASS)[i];
    double cw[3], t[3];
    mat3_vec(R, PANDA(PANDA_COM)[i], t);
    cw[0] = t[0] + p[0]; cw[1] = t[1] + p[1]; cw[2] = t[2] + p[2];
    const double* I = PANDA(PANDA_ICOM)[i];
    double RI[9], Iw[9];
    MPCB_UNROLL
    for (int r = 0; r < 3; r++)
        MPCB_UNROLL
        for (int c = 0; c < 3; c++) RI[3 * r + c] = R[3 * r] * I[c] + R[3 * r + 1] * I[3 + c] + R[3 * r + 2] * I[6 + c];
    MPCB_UNROLL
    for (int r = 0; r < 3; r++)
        MPCB_UNROLL
        for (int c = 0; c < 3; c++) Iw[3 * r + c] = RI[3 * r] * R[3 * c] + RI[3 * r + 1] * R[3 * c + 1] + RI[3 * r + 2] * R[3 * c + 2];
    const double C[9] = {0, -cw[2], cw[1], cw[2], 0, -cw[0], -cw[1], cw[0], 0};
    MPCB_UNROLL
    for (int r = 0; r < 3; r++)
        MPCB_UNROLL
        for (int c = 0; c < 3; c++) {
            double cc = C[3 * r] * C[c] + C[3 * r + 1] * C[3 + c] + C[3 * r + 2] * C[6 + c];
            Y[6 * r + c] = (r == c) ? m : 0.0;
            Y[6 * r + c + 3] = -m * C[3 * r + c];
            Y[6 * (r + 3) + c] = m * C[3 * r + c];
            Y[6 * (r + 3) + c + 3] = Iw[3 * r + c] - m * cc;
        }
}

// B = crf(v) Y - Y crm(v) + Fx(h)   with h = Y v   (pinocchio's doYcrb term)
MPCB_HD void body_B(const double* v, const double* Y, const double* h, double* B) {
    // crm(v) = [[W, V],[0, W]], crf(v) = [[W, 0],[V, W]] with W = skew(angular), V = skew(linear)
    const double W[9] = {0, -v[5], v[4], v[5], 0, -v[3], -v[4], v[3], 0};
    const double V[9] = {0, -v[2], v[1], v[2], 0, -v[0], -v[1], v[0], 0};
    double X[36], Xf[36];
    MPCB_UNROLL
    for (int r = 0; r < 3; r++)
        MPCB_UNROLL
        for (int c = 0; c < 3; c++) {
            X[6 * r + c] = W[3 * r + c]; X[6 * r + c + 3] = V[3 * r + c]; X[6 * (r + 3) + c] = 0; X[6 * (r + 3) + c + 3] = W[3 * r + c];
            Xf[6 * r + c] = W[3 * r + c]; Xf[6 * r + c + 3] = 0; Xf[6 * (r + 3) + c] = V[3 * r + c]; Xf[6 * (r + 3) + c + 3] = W[3 * r + c];
        }
    for (int r = 0; r < 6; r++)
        MPCB_UNROLL
        for (int c = 0; c < 6; c++) {
            double s1 = 0, s2 = 0;
            MPCB_UNROLL
            for (int k = 0; k < 6; k++) { s1 += Xf[6 * r + k] * Y[6 * k + c]; s2 += Y[6 * r + k] * X[6 * k + c]; }
            B[6 * r + c] = s1 - s2;
        }
    const double Sf[9] = {0, -h[2], h[1], h[2], 0, -h[0], -h[1], h[0], 0};
    const double Sn[9] = {0, -h[5], h[4], h[5], 0, -h[3], -h[4], h[3], 0};
    MPCB_UNROLL
    for (int r = 0; r < 3; r++)
        MPCB_UNROLL
        for (int c = 0; c < 3; c++) {
            B[6 * r + c + 3] -= Sf[3 * r + c];
            B[6 * (r + 3) + c] -= Sf[3 * r + c];
            B[6 * (r + 3) + c + 3] -= Sn[3 * r + c];
        }
}

// Scratch for the derivative pass (per evaluation).  The CUDA kernel places it in shared/global scratch; the host
// code on the stack.
struct DerivScratch {
    double oR[NJ][9], op[NJ][3];
    double J[NJ][6], ov[NJ][6], oa[NJ][6], of[NJ][6];
    double dVdq[NJ][6], dAdq[NJ][6], dAdv[NJ][6];
    double dFda[NJ][6], dFdv[NJ][6], dFdq[NJ][6];
    double oY[NJ][36], B[NJ][36];
};

// tau, dtau/dq, dtau/dv (7x7 row-major, full) and the symmetric joint-space inertia M
// (pinocchio::computeRNEADerivatives + rnea + crba as used at robot_ocp.hpp:118-122), world-frame recursion.
MPCB_HD_NOINLINE void rnea_derivatives(const double* q, const double* qd, const double* qdd, DerivScratch& S,
                                       double* tau, double* dq, double* dv, double* M) {
    fk_world(q, S.oR, S.op);
    for (int i = 0; i < NJ; i++) {
        double vpar[6], apar[6];
        if (i == 0) {
            for (int k = 0; k < 6; k++) { vpar[k] = 0; apar[k] = 0; }
            apar[2] = PANDA(PANDA_GRAVITY);
        } else {
            for (int k = 0; k < 6; k++) { vpar[k] = S.ov[i - 1][k]; apar[k] = S.oa[i - 1][k]; }
        }
        const double z[3] = {S.oR[i][2], S.oR[i][5], S.oR[i][8]};
        double pxz[3];
        cross3(S.op[i], z, pxz);
        double* J = S.J[i];
        J[0] = pxz[0]; J[1] = pxz[1]; J[2] = pxz[2]; J[3] = z[0]; J[4] = z[1]; J[5] = z[2];
        double Jqd[6], t[6];
        for (int k = 0; k < 6; k++) { Jqd[k] = J[k] * qd[i]; S.ov[i][k] = vpar[k] + Jqd[k]; }
        crm_mul(S.ov[i], Jqd, t);
        for (int k = 0; k < 6; k++) S.oa[i][k] = apar[k] + J[k] * qdd[i] + t[k];
        world_inertia(i, S.oR[i], S.op[i], S.oY[i]);
        double oh[6], ya[6], vxh[6];
        mv6(S.oY[i], S.ov[i], oh);
        mv6(S.oY[i], S.oa[i], ya);
        crf_mul(S.ov[i], oh, vxh);
        for (int k = 0; k < 6; k++) S.of[i][k] = ya[k] + vxh[k];
        double dJ[6];
        crm_mul(S.ov[i], J, dJ);
        if (i == 0) { for (int k = 0; k < 6; k++) S.dVdq[i][k] = 0; }
        else crm_mul(vpar, J, S.dVdq[i]);
        crm_mul(apar, J, S.dAdq[i]);
        if (i > 0) { double t2[6]; crm_mul(vpar, S.dVdq[i], t2); for (int k = 0; k < 6; k++) S.dAdq[i][k] += t2[k]; }
        for (int k = 0; k < 6; k++) S.dAdv[i][k] = dJ[k] + S.dVdq[i][k];
        body_B(S.ov[i], S.oY[i], oh, S.B[i]);
    }
    for (int i = 0; i < NJ * NJ; i++) { dq[i] = 0; dv[i] = 0; M[i] = 0; }
    for (int i = NJ - 1; i >= 0; i--) {
        const double* J = S.J[i];
        tau[i] = dot6(J, S.of[i]);
        mv6(S.oY[i], J, S.dFda[i]);
        double t1[6], t2[6];
        mv6(S.B[i], J, t1); mv6(S.oY[i], S.dAdv[i], t2);
        for (int k = 0; k < 6; k++) S.dFdv[i][k] = t1[k] + t2[k];
        mv6(S.oY[i], S.dAdq[i], t1); mv6(S.B[i], S.dVdq[i], t2);
        for (int k = 0; k < 6; k++) S.dFdq[i][k] = t1[k] + t2[k];
        for (int j = i; j < NJ; j++) {
            M[i * NJ + j] = dot6(J, S.dFda[j]);
            dv[i * NJ + j] = dot6(J, S.dFdv[j]);
            dq[i * NJ + j] = dot6(J, S.dFdq[j]);
        }
        // dFdq_i += J_i x* of_i  (after row i used it)
        crf_mul(J, S.of[i], t1);
        for (int k = 0; k < 6; k++) S.dFdq[i][k] += t1[k];
        double JtY[6], JtB[6];
        vm6(J, S.oY[i], JtY);
        vm6(J, S.B[i], JtB);
        for (int j = 0; j < i; j++) {
            dq[i * NJ + j] = dot6(JtY, S.dAdq[j]) + dot6(JtB, S.dVdq[j]);
            dv[i * NJ + j] = dot6(JtY, S.dAdv[j]) + dot6(JtB, S.J[j]);
        }
        if (i > 0) {
            for (int k = 0; k < 36; k++) { S.oY[i - 1][k] += S.oY[i][k]; S.B[i - 1][k] += S.B[i][k]; }
            for (int k = 0; k < 6; k++) S.of[i - 1][k] += S.of[i][k];
        }
    }
    for (int i = 0; i < NJ; i++)
        for (int j = 0; j < i; j++) M[i * NJ + j] = M[j * NJ + i];
}

// World-aligned Jacobian (6x7 row-major: linear rows then angular rows) and position of a frame rigidly attached to
// joint 7 at `off` (computeFrameJacobian + blockdiag(R,R); robot_ocp.hpp:147-155, pandaWrapper.cpp:65-73).
MPCB_HD void frame_jacobian(const double (*oR)[9], const double (*op)[3], const double* off, double* J, double* pos) {
    double t[3], pf[3];
    mat3_vec(oR[6], off, t);
    pf[0] = op[6][0] + t[0]; pf[1] = op[6][1] + t[1]; pf[2] = op[6][2] + t[2];
    if (pos) { pos[0] = pf[0]; pos[1] = pf[1]; pos[2] = pf[2]; }
    if (!J) return;
    for (int i = 0; i < NJ; i++) {
        const double z[3] = {oR[i][2], oR[i][5], oR[i][8]};
        const double d[3] = {pf[0] - op[i][0], pf[1] - op[i][1], pf[2] - op[i][2]};
        double zxd[3];
        cross3(z, d, zxd);
        for (int k = 0; k < 3; k++) { J[k * NJ + i] = zxd[k]; J[(k + 3) * NJ + i] = z[k]; }
    }
}

// Full derivative evaluation of one node's 8 inequality outputs (robot_ocp.hpp:98-163):
//   g[0..6] = tau, g[7] = z_tool
//   Jg (8 x 22 row-major): rows 0-6 = [dtau/dq | dtau/dv | M | dtau/dv*qd + triu(M)*qdd], row 7 = [Jz_tool | 0]
// The time column reproduces the reference's hand-filled heuristic (robot_ocp.hpp:116,124,138).
MPCB_HD_NOINLINE void node_constraints_jac(const double* q, const double* qd, const double* qdd, DerivScratch& S,
                                           double* g, double* Jg) {
    double tau[NJ], dq[49], dv[49], M[49];
    rnea_derivatives(q, qd, qdd, S, tau, dq, dv, M);
    for (int i = 0; i < NJ; i++) {
        g[i] = tau[i];
        double* row = Jg + i * 22;
        double dt = 0;
        for (int j = 0; j < NJ; j++) {
            row[j] = dq[i * NJ + j];
            row[7 + j] = dv[i * NJ + j];
            row[14 + j] = M[i * NJ + j];
            dt += dv[i * NJ + j] * qd[j];
        }
        for (int j = i; j < NJ; j++) dt += M[i * NJ + j] * qdd[j];
        row[21] = dt;
    }
    // S.oR / S.op still hold the world placements (rnea_derivatives does not modify them)
    double J[42], pos[3];
    frame_jacobian(S.oR, S.op, PANDA(PANDA_TOOL_OFFSET), J, pos);
    g[7] = pos[2];
    double* row = Jg + 7 * 22;
    for (int j = 0; j < 22; j++) row[j] = 0;
    for (int j = 0; j < NJ; j++) row[j] = J[2 * NJ + j];
}

}  // namespace rbd
}  // namespace mpcb
