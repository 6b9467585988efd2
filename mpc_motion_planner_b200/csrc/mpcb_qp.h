// Linearisation, KKT factorisation and the box-ADMM iteration of one OCP (phases; see mpcb_ctx.h).
//
// What it replaces in the reference stack (all inside `mpc.solve()`, /root/reference/mpc_solver/motionPlanner.cpp:184):
//   * polympc ContinuousOCP linearisation over minTime_ocp (mpc_solver/robot_ocp.hpp:55-73 dynamics, :98-163 inequality
//     constraints with hand-filled Jacobian, :165-168 zero second derivatives) + MySolver's Gershgorin regularisation
//     (mpc_solver/polympc_redef.hpp:57-70)                                                            -> lin_* phases
//   * polympc boxADMM::solve (OSQP iteration with implicit box rows, sparse LDLT of the KKT matrix)
//                                                                                       -> fac_* and admm_* phases
// The KKT system is solved through its reduced SPD form S = H + sigma I + rho_box + A^T rho_A A with a one-level nested
// dissection (mpcb_shape.h): explicit inverses of the NS interior blocks, F_s = S_II^-1 S_IB and the inverse of the
// separator Schur complement, all produced by symmetric sweeps; one ADMM iteration is then three dependent stages of
// dense mat-vecs instead of 2*NVAR dependent substitution steps.
#pragma once
#include "mpcb_ctx.h"

namespace mpcb {

// =====================================================================================================================
// Linearisation
// =====================================================================================================================
template <class SH>
MPCB_DEV int node_thread(int k) { return (k % 16) * 32 + k / 16; }   // spread the heavy per-node work over warps

template <class SH>
MPCB_DEV void lin_a(Ctx<SH>& c, TS<SH>& ts, int tid) {
    const double p = c.z[SH::IP], tsc = c.cheb->ts, reg = c.st->hess_reg, rho = c.st->qp_rho;
    if (tid == 0) { c.scal[SC_PBAR] = p; c.scal[SC_MTP] = -tsc * p; }
    // inequality rows: one node per thread (value + Jacobian), scratch in the (dead) matrix area
    for (int k = 0; k < SH::NN; k++) {
        if (node_thread<SH>(k) != tid) continue;
        double q[7], v[7], u[7], g[8];
        for (int j = 0; j < 7; j++) { q[j] = c.z[SH::kvar(k, j)]; v[j] = c.z[SH::kvar(k, 7 + j)]; u[j] = c.z[SH::kvar(k, 14 + j)]; }
        rbd::DerivScratch& S = *reinterpret_cast<rbd::DerivScratch*>(c.sm + Layout<SH>::MAT + k * Layout<SH>::DERIV_STRIDE);
        rbd::node_constraints_jac(q, v, u, S, g, c.Jg + k * 176);
        for (int r = 0; r < 8; r++) c.w[SH::NEQ + 8 * k + r] = g[r];
    }
    // collocation rows:  sum_l D[j,l] x_{3s+l} - ts p f = 0   (robot_ocp.hpp:64-70 for f)
    if (tid < SH::NEQ) {
        double acc = 0;
        MPCB_UNROLL
        for (int l = 0; l <= SH::PO; l++) acc += ts.dco[l] * c.z[ts.ridx[l]];
        const double f = c.z[ts.ridx[SH::PO + 1]];
        const double ceq = acc - tsc * p * f;
        ts.l = -ceq; ts.u = -ceq;
        ts.pc = -tsc * f;
        c.pcol[tid] = ts.pc;
        ts.rho = row_rho(ts.l, ts.u, rho);
        c.xt[tid] = ts.rho;
    }
    // Lagrangian Hessian arrow: H[p, v_k] = -ts lam_eq[k, q-row], H[p, u_k] = -ts lam_eq[k, v-row]; box bounds
    if (tid < SH::NVAR) {
        const double zi = c.z[tid];
        double lo, hi;
        var_bounds(c, ts.vk, ts.vcomp, lo, hi);
        ts.lb = lo - zi; ts.ub = hi - zi;
        ts.rhoB = row_rho(ts.lb, ts.ub, rho);
        ts.ha = (ts.gfrow >= 0) ? -tsc * c.lam[ts.gfrow] : 0.0;
        ts.hd = fabs(ts.ha) + reg;          // Gershgorin: a_ii = 0, r_i = |arrow_i|  (polympc_redef.hpp:57-70)
        c.Ha[tid] = (tid == SH::IP) ? 0.0 : ts.ha;
    }
}
template <class SH>
MPCB_DEV void lin_b(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid >= SH::NEQ && tid < SH::NA) {
        const int rr = (tid - SH::NEQ) % 8;
        const double g = c.w[tid];
        double lo, hi;
        ineq_bounds(c, rr, lo, hi);
        ts.l = lo - g; ts.u = hi - g;
        ts.rho = row_rho(ts.l, ts.u, c.st->qp_rho);
        c.xt[tid] = ts.rho;
    }
    if (tid >= NT - 32) {
        double s = 0;
        for (int i = tid - (NT - 32); i < SH::NVAR - 1; i += 32) s += fabs(c.Ha[i]);
        c.psum[tid - (NT - 32)] = s;
    }
}
template <class SH>
MPCB_DEV void lin_c(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid == SH::IP) {
        double s = 0;
        for (int j = 0; j < 32; j++) s += c.psum[j];
        ts.hd = s + c.st->hess_reg;
        ts.ha = 0;
        c.scal[SC_HDP] = ts.hd;
    }
}

// =====================================================================================================================
// Factorisation: assemble S block-wise, sweep the interiors (-> -S_II^-1, F, Schur contributions), sweep the separator
// =====================================================================================================================
template <class SH>
MPCB_DEV double* aug_ptr(const Ctx<SH>& c, int s, int m, int nb, int i, int j) {   // i >= j, augmented indices
    if (i < m) return c.W + SH::woff(s) + i * (i + 1) / 2 + j;
    if (j < m) return c.F + SH::foff(s) + j * nb + (i - m);
    return c.BB + SH::bboff(s) + (i - m) * (i - m + 1) / 2 + (j - m);
}

template <class SH>
MPCB_DEV void fac_zero(Ctx<SH>& c, int tid) {
    for (int e = tid; e < SH::MATSIZE + SH::BBSIZE; e += NT) c.W[e] = 0.0;
}
template <class SH>
MPCB_DEV void fac_diag(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NVAR) {
        const double d = ts.hd + c.st->qp_sigma + ts.rhoB;
        c.rhs[tid] = d;
        if (ts.vseg >= 0) {
            const int s = ts.vseg, l = ts.vloc, nb = SH::nb(s);
            c.W[SH::woff(s) + l * (l + 1) / 2 + l] = d;
            c.F[SH::foff(s) + l * nb + (nb - 1)] = ts.ha;
        }
    }
}
// one collocation row (node s*PO+j, component cc) of every segment: S += rho a a^T
template <class SH>
MPCB_DEV void fac_eq_row(Ctx<SH>& c, int tid, int j, int cc) {
    constexpr int GS = NT / SH::NS, NV = SH::PO + 3, NP = NV * (NV + 1) / 2;
    const int s = tid / GS, lane = tid % GS;
    if (s >= SH::NS || lane >= NP) return;
    const int k = s * SH::PO + j, r = 14 * k + cc, m = SH::m(s), nb = SH::nb(s);
    int a = 0, rem = lane;
    while (rem > a) { rem -= a + 1; a++; }
    const int b = rem;
    int idx[2]; double val[2];
    const int sel[2] = {a, b};
    for (int t = 0; t < 2; t++) {
        const int v = sel[t];
        if (v <= SH::PO) { idx[t] = SH::aug_index(s, s * SH::PO + v, cc); val[t] = c.cheb->D[j][v]; }
        else if (v == SH::PO + 1) { idx[t] = SH::aug_index(s, k, (cc < 7) ? 7 + cc : 14 + (cc - 7)); val[t] = c.scal[SC_MTP]; }
        else { idx[t] = m + nb - 1; val[t] = c.pcol[r]; }
    }
    const int i = idx[0] > idx[1] ? idx[0] : idx[1], jj = idx[0] > idx[1] ? idx[1] : idx[0];
    *aug_ptr(c, s, m, nb, i, jj) += c.xt[r] * val[0] * val[1];
}
// the 8 inequality rows of one node of every segment
template <class SH>
MPCB_DEV void fac_ineq_node(Ctx<SH>& c, int tid, int jn) {
    constexpr int GS = NT / SH::NS;
    const int s = tid / GS, lane = tid % GS;
    if (s >= SH::NS) return;
    if (jn == SH::PO && s != SH::NS - 1) return;
    const int k = (jn < SH::PO) ? s * SH::PO + jn : SH::NN - 1, m = SH::m(s), nb = SH::nb(s);
    const double* G = c.Jg + k * 176;
    const double* rho = c.xt + SH::NEQ + 8 * k;
    for (int e = lane; e < 253; e += GS) {
        int a = 0, rem = e;
        while (rem > a) { rem -= a + 1; a++; }
        const int b = rem;
        double acc = 0;
        for (int rr = 0; rr < 8; rr++) acc += rho[rr] * G[rr * 22 + a] * G[rr * 22 + b];
        const int ia = (a < 21) ? SH::aug_index(s, k, a) : m + nb - 1;
        const int ib = (b < 21) ? SH::aug_index(s, k, b) : m + nb - 1;
        const int i = ia > ib ? ia : ib, jj = ia > ib ? ib : ia;
        *aug_ptr(c, s, m, nb, i, jj) += acc;
    }
}
// symmetric sweep of pivot kp on every segment's augmented matrix [[S_II, S_IB],[., S_BB]]
template <class SH>
MPCB_DEV void fac_sweep_copy(Ctx<SH>& c, int tid, int kp) {
    constexpr int GS = NT / SH::NS;
    const int s = tid / GS, lane = tid % GS;
    if (s >= SH::NS) return;
    const int m = SH::m(s), nb = SH::nb(s);
    if (kp >= m) return;
    double* col = c.colbuf + s * Layout<SH>::COLSTRIDE;
    for (int e = lane; e < m + nb; e += GS) col[e] = *aug_ptr(c, s, m, nb, e > kp ? e : kp, e > kp ? kp : e);
}
template <class SH>
MPCB_DEV void fac_sweep_update(Ctx<SH>& c, int tid, int kp) {
    constexpr int GS = NT / SH::NS, TJ = 8, TI = GS / TJ;
    const int s = tid / GS, lane = tid % GS;
    if (s >= SH::NS || lane >= TI * TJ) return;
    const int m = SH::m(s), nb = SH::nb(s), M = m + nb;
    if (kp >= m) return;
    const double* col = c.colbuf + s * Layout<SH>::COLSTRIDE;
    const double dinv = 1.0 / col[kp];
    for (int i = lane / TJ; i < M; i += TI) {
        const double ci = col[i] * dinv;
        for (int j = lane % TJ; j <= i; j += TJ) {
            double* a = aug_ptr(c, s, m, nb, i, j);
            if (i == kp) *a = (j == kp) ? -dinv : col[j] * dinv;
            else if (j == kp) *a = ci;
            else *a -= ci * col[j];
        }
    }
}
// local boundary column of separator variable b in segment s (or -1)
template <class SH>
MPCB_DEV int sep_local(int s, int b) {
    if (b == SH::NB - 1) return SH::nb(s) - 1;
    const int sj = b / 14 + 1, cc = b % 14;
    if (sj == s) return cc;
    if (sj == s + 1) return (s > 0 ? 14 : 0) + cc;
    return -1;
}
template <class SH>
MPCB_DEV void fac_c_assemble(Ctx<SH>& c, int tid) {
    for (int e = tid; e < SH::NB * SH::NB; e += NT) {
        const int i = e / SH::NB, j = e % SH::NB;
        if (j > i) continue;
        double v = (i == j) ? c.rhs[SH::NI + i] : 0.0;
        if (i == SH::NB - 1 && j < SH::NB - 1) v += c.Ha[SH::NI + j];
        for (int s = 0; s < SH::NS; s++) {
            const int li = sep_local<SH>(s, i), lj = sep_local<SH>(s, j);
            if (li < 0 || lj < 0) continue;
            const int a = li > lj ? li : lj, b = li > lj ? lj : li;
            v += c.BB[SH::bboff(s) + a * (a + 1) / 2 + b];
        }
        c.C[i * (i + 1) / 2 + j] = v;
    }
}
template <class SH>
MPCB_DEV void fac_c_copy(Ctx<SH>& c, int tid, int kp) {
    for (int e = tid; e < SH::NB; e += NT) {
        const int i = e > kp ? e : kp, j = e > kp ? kp : e;
        c.colbuf[e] = c.C[i * (i + 1) / 2 + j];
    }
}
template <class SH>
MPCB_DEV void fac_c_update(Ctx<SH>& c, int tid, int kp) {
    const double* col = c.colbuf;
    const double dinv = 1.0 / col[kp];
    for (int i = tid / 16; i < SH::NB; i += NT / 16) {
        const double ci = col[i] * dinv;
        for (int j = tid % 16; j <= i; j += 16) {
            double* a = c.C + i * (i + 1) / 2 + j;
            if (i == kp) *a = (j == kp) ? -dinv : col[j] * dinv;
            else if (j == kp) *a = ci;
            else *a -= ci * col[j];
        }
    }
}

template <class SH>
MPCB_DEV void factorise(MPCB_DRIVER_PARAMS(SH)) {
    MPCB_PHASE(fac_zero(c, tid));
    MPCB_PHASE(fac_diag(c, ts, tid));
    for (int j = 0; j < SH::PO; j++)
        for (int cc = 0; cc < 14; cc++) MPCB_PHASE(fac_eq_row(c, tid, j, cc));
    for (int jn = 0; jn <= SH::PO; jn++) MPCB_PHASE(fac_ineq_node(c, tid, jn));
    for (int kp = 0; kp < SH::MMAX; kp++) {
        MPCB_PHASE(fac_sweep_copy(c, tid, kp));
        MPCB_PHASE(fac_sweep_update(c, tid, kp));
    }
    MPCB_PHASE(fac_c_assemble(c, tid));
    for (int kp = 0; kp < SH::NB; kp++) {
        MPCB_PHASE(fac_c_copy(c, tid, kp));
        MPCB_PHASE(fac_c_update(c, tid, kp));
    }
    // W now holds -S_II^-1 (packed lower, per segment), F = S_II^-1 S_IB, C = -(Schur complement)^-1
}

}  // namespace mpcb
