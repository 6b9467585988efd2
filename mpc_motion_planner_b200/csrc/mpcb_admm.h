// The box-ADMM iteration (polympc boxADMM::solve restated in SURVEY.md A.5) on top of the factorisation of mpcb_qp.h.
//
//   rhs   = sigma x - h + rho_b z_b - y_b + A^T (rho_A z_A - y_A)                           admm_rhs
//   x~    = S^-1 rhs   via   t = W rhs_I ; r = rhs_B - F^T rhs_I ; x~_B = C^-1 r ; x~_I = t - F x~_B
//                                                                                    admm_solve1/2/3
//   z~_A  = A x~ ;  x = a x~ + (1-a) x ;  z = clip(a z~ + (1-a) z + y/rho) ;  y += rho (a z~ + (1-a) z_old - z)   admm_update
// Every `check_termination` iterations the OSQP residual test runs (admm_check_*).
// Row r's (z_A, y_A, l, u, rho) and variable i's (x, z_b, y_b, lb, ub, rho_b) live in the registers of thread r / i for
// the whole QP; shared memory only carries w = rho_A z_A - y_A, rhs and x~ between the stages.
#pragma once
#include "mpcb_qp.h"

namespace mpcb {

template <class SH>
MPCB_DEV double row_times(const Ctx<SH>& c, const TS<SH>& ts, int r, const double* x) {   // (A x)_r
    if (r < SH::NEQ) {
        double acc = 0;
        MPCB_UNROLL
        for (int l = 0; l <= SH::PO; l++) acc += ts.dco[l] * x[ts.ridx[l]];
        acc += c.scal[SC_MTP] * x[ts.ridx[SH::PO + 1]];
        acc += ts.pc * x[SH::IP];
        return acc;
    }
    const double* G = c.Jg + (r - SH::NEQ) * 22;
    const double* xs = x + ts.ridx[0];
    const double* us = x + ts.ridx[1];
    double acc = 0;
    MPCB_UNROLL
    for (int cc = 0; cc < 14; cc++) acc += G[cc] * xs[cc];
    MPCB_UNROLL
    for (int cc = 0; cc < 7; cc++) acc += G[14 + cc] * us[cc];
    acc += G[21] * x[SH::IP];
    return acc;
}
template <class SH>
MPCB_DEV double col_times(const Ctx<SH>& c, const TS<SH>& ts, const double* y) {   // (A^T y)_i for i != p
    double g = 0;
    MPCB_UNROLL
    for (int n = 0; n < 2 * SH::PO; n++) g += ts.gco[n] * y[ts.grow[n]];     // padded with zero coefficients
    if (ts.gfrow >= 0) g += c.scal[SC_MTP] * y[ts.gfrow];
    if (ts.gk >= 0) {
        const double* G = c.Jg + ts.gk * 176 + ts.gcol;
        const double* yy = y + SH::NEQ + 8 * ts.gk;
        MPCB_UNROLL
        for (int rr = 0; rr < 8; rr++) g += G[rr * 22] * yy[rr];
    }
    return g;
}
// partial sums of (A^T y)_p = sum_r A[r,p] y_r  by the last 32 threads
template <class SH>
MPCB_DEV void pcol_partial(const Ctx<SH>& c, int tid, const double* y, double* out) {
    if (tid < NT - 32) return;
    const int j = tid - (NT - 32);
    double s = 0;
    for (int r = j; r < SH::NEQ; r += 32) s += c.pcol[r] * y[r];
    for (int r = j; r < SH::NINEQ; r += 32) s += c.Jg[r * 22 + 21] * y[SH::NEQ + r];
    out[j] = s;
}

template <class SH>
MPCB_DEV void admm_init(Ctx<SH>& c, TS<SH>& ts, int tid) {
    ts.zA = 0; ts.yA = 0; ts.x = 0; ts.zB = 0; ts.yB = 0;
    if (tid < SH::VP) { c.w[tid] = 0; c.rhs[tid] = 0; c.xt[tid] = 0; }
    if (tid < UM_COUNT) c.umax[tid] = 0;
    if (tid == 0) { c.isc[IS_QP_DONE] = 0; c.isc[IS_QP_STATUS] = 2; c.isc[IS_QP_ITERS] = 0; }
}
template <class SH>
MPCB_DEV void admm_rhs(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NVAR) {
        const double g = col_times(c, ts, c.w);
        const double h = (tid == SH::IP) ? 1.0 : 0.0;     // gradient of the Mayer cost p (robot_ocp.hpp:206)
        c.rhs[tid] = c.st->qp_sigma * ts.x - h + ts.rhoB * ts.zB - ts.yB + g;
    }
    pcol_partial(c, tid, c.w, c.psum);
}
template <class SH>
MPCB_DEV void admm_solve1(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NI) {      // t = -(Wn rhs_I), symmetric packed row/column walk
        const int s = ts.vseg, l = ts.vloc, m = SH::m(s);
        const double* Wn = c.W + SH::woff(s);
        const double* b = c.rhs + SH::ioff(s);
        double acc = 0;
        const double* row = Wn + l * (l + 1) / 2;
        for (int j = 0; j <= l; j++) acc += row[j] * b[j];
        for (int j = l + 1; j < m; j++) acc += Wn[j * (j + 1) / 2 + l] * b[j];
        c.xt[tid] = -acc;
    }
    for (int e = tid - SH::NI; e >= 0 && e < SH::PARTSIZE; e += NT) {      // F_s^T rhs_I, one boundary column per thread
        int s = 0, col = e;
        while (col >= SH::nb(s)) { col -= SH::nb(s); s++; }
        const int m = SH::m(s), nb = SH::nb(s);
        const double* Fs = c.F + SH::foff(s) + col;
        const double* b = c.rhs + SH::ioff(s);
        double acc = 0;
        for (int i = 0; i < m; i++) acc += Fs[i * nb] * b[i];
        c.part[e] = acc;
    }
    if (tid == NT - 1) {
        double s = 0;
        for (int j = 0; j < 32; j++) s += c.psum[j];
        c.rhs[SH::IP] += s;
    }
}
template <class SH>
MPCB_DEV double sep_rhs(const Ctx<SH>& c, int b) {     // r_b = rhs_B[b] - sum_s (F_s^T rhs_I)[b]
    double r = c.rhs[SH::NI + b];
    if (b == SH::NB - 1) {
        for (int s = 0; s < SH::NS; s++) r -= c.part[SH::poff(s) + SH::nb(s) - 1];
    } else {
        const int sj = b / 14 + 1, cc = b % 14;      // junction x_{sj*PO}: second block of segment sj-1, first block of sj
        r -= c.part[SH::poff(sj - 1) + (sj - 1 > 0 ? 14 : 0) + cc];
        r -= c.part[SH::poff(sj) + cc];
    }
    return r;
}
template <class SH>
MPCB_DEV void admm_solve2(Ctx<SH>& c, int tid) {       // x~_B = -(Cn r), 4 lanes per row
    const bool valid = tid < 4 * SH::NB;
    const int b = valid ? (tid >> 2) : 0, sub = tid & 3;
#ifdef MPCB_EMUL
    if (!valid || sub) return;
    double acc = 0;
    for (int j = 0; j < SH::NB; j++) {
        const int hi = b > j ? b : j, lo = b > j ? j : b;
        acc += c.C[hi * (hi + 1) / 2 + lo] * sep_rhs(c, j);
    }
    c.xt[SH::NI + b] = -acc;
#else
    double acc = 0;
    if (valid)
        for (int j = sub; j < SH::NB; j += 4) {
            const int hi = b > j ? b : j, lo = b > j ? j : b;
            acc += c.C[hi * (hi + 1) / 2 + lo] * sep_rhs(c, j);
        }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);      // executed by every thread of the CTA (no early exit above)
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (valid && sub == 0) c.xt[SH::NI + b] = -acc;
#endif
}
template <class SH>
MPCB_DEV void admm_solve3(Ctx<SH>& c, TS<SH>& ts, int tid) {     // x~_I = t - F x~_B
    if (tid >= SH::NI) return;
    const int s = ts.vseg, nb = SH::nb(s);
    const double* Fr = c.F + SH::foff(s) + ts.vloc * nb;
    const double* xb = c.xt + SH::NI;
    double acc = 0;
    int col = 0;
    if (s > 0) { const double* x1 = xb + 14 * (s - 1); for (int cc = 0; cc < 14; cc++) acc += Fr[col + cc] * x1[cc]; col += 14; }
    if (s < SH::NS - 1) { const double* x2 = xb + 14 * s; for (int cc = 0; cc < 14; cc++) acc += Fr[col + cc] * x2[cc]; col += 14; }
    acc += Fr[col] * xb[SH::NB - 1];
    c.xt[tid] -= acc;
}
template <class SH>
MPCB_DEV void admm_update(Ctx<SH>& c, TS<SH>& ts, int tid) {
    const double al = c.st->qp_alpha;
    if (tid < SH::NA) {
        const double ax = row_times(c, ts, tid, c.xt);
        const double zh = al * ax + (1 - al) * ts.zA;
        const double zn = dmin(dmax(zh + ts.yA / ts.rho, ts.l), ts.u);
        ts.yA += ts.rho * (zh - zn);
        ts.zA = zn;
        c.w[tid] = ts.rho * ts.zA - ts.yA;
    }
    if (tid < SH::NVAR) {
        const double xt = c.xt[tid];
        ts.x = al * xt + (1 - al) * ts.x;
        const double zh = al * xt + (1 - al) * ts.zB;
        const double zn = dmin(dmax(zh + ts.yB / ts.rhoB, ts.lb), ts.ub);
        ts.yB += ts.rhoB * (zh - zn);
        ts.zB = zn;
    }
}
// ---- residual test (every check_termination iterations) -----------------------------------------------------------
template <class SH>
MPCB_DEV void admm_check_a(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NA) c.rhs[tid] = ts.yA;       // rhs slot is free until the next admm_rhs
    if (tid < SH::NVAR) c.xt[tid] = ts.x;
}
template <class SH>
MPCB_DEV void admm_check_b(Ctx<SH>& c, TS<SH>& ts, int tid) {
    double rp = 0, nax = 0, nz = 0, nx = 0, rd = 0, nhx = 0, naty = 0;
    if (tid < SH::NA) {
        const double ax = row_times(c, ts, tid, c.xt);
        rp = fabs(ax - ts.zA); nax = fabs(ax); nz = fabs(ts.zA);
    }
    if (tid < SH::NVAR) {
        nx = fabs(ts.x); nz = dmax(nz, fabs(ts.zB)); rp = dmax(rp, fabs(ts.x - ts.zB));
        if (tid != SH::IP) {
            const double aty = col_times(c, ts, c.rhs) + ts.yB;
            const double hx = ts.hd * ts.x + ts.ha * c.xt[SH::IP];
            naty = fabs(aty); nhx = fabs(hx); rd = fabs(hx + aty);
        }
    }
    pcol_partial(c, tid, c.rhs, c.psum);
    if (tid >= NT - 32) {
        double s = 0;
        for (int i = tid - (NT - 32); i < SH::NVAR - 1; i += 32) s += c.Ha[i] * c.xt[i];
        c.psum[32 + tid - (NT - 32)] = s;
    }
    umax_accum(c.umax + UM_RP, rp); umax_accum(c.umax + UM_NAX, nax); umax_accum(c.umax + UM_NZ, nz); umax_accum(c.umax + UM_NX, nx);
    umax_accum(c.umax + UM_RD, rd); umax_accum(c.umax + UM_NHX, nhx); umax_accum(c.umax + UM_NATY, naty);
}
template <class SH>
MPCB_DEV void admm_check_c(Ctx<SH>& c, TS<SH>& ts, int tid, int iter) {
    if (tid != SH::IP) return;
    double s1 = 0, s2 = 0;
    for (int j = 0; j < 32; j++) { s1 += c.psum[j]; s2 += c.psum[32 + j]; }
    const double aty = s1 + ts.yB, hx = ts.hd * ts.x + s2;
    const double rd = dmax(umax_get(c.umax + UM_RD), fabs(hx + 1.0 + aty));
    const double nhx = dmax(umax_get(c.umax + UM_NHX), fabs(hx)), naty = dmax(umax_get(c.umax + UM_NATY), fabs(aty));
    const double rp = umax_get(c.umax + UM_RP);
    const double np_ = dmax(dmax(umax_get(c.umax + UM_NAX), umax_get(c.umax + UM_NX)), umax_get(c.umax + UM_NZ));
    const double nd_ = dmax(dmax(nhx, naty), 1.0);      // |h|_inf = 1
    const double ep = c.st->qp_eps_abs + c.st->qp_eps_rel * np_, ed = c.st->qp_eps_abs + c.st->qp_eps_rel * nd_;
    c.scal[SC_RP] = rp; c.scal[SC_RD] = rd; c.scal[SC_EP] = ep; c.scal[SC_ED] = ed;
    if (rp <= ep && rd <= ed) { c.isc[IS_QP_DONE] = 1; c.isc[IS_QP_STATUS] = 1; c.isc[IS_QP_ITERS] = iter; }
    for (int j = 0; j < UM_COUNT; j++) c.umax[j] = 0;
}
template <class SH>
MPCB_DEV void admm_finish(Ctx<SH>& c, TS<SH>& ts, int tid) {    // the factor is dead: park the QP result on top of it
    if (tid < SH::NA) c.qp_lam[tid] = ts.yA;
    if (tid < SH::NVAR) { c.qp_lam[SH::NA + tid] = ts.yB; c.qp_p[tid] = ts.x; }
}

// One QP: factorise at the current linearisation and iterate.  Result in c.qp_p / c.qp_lam, status in c.isc.
template <class SH>
MPCB_DEV void solve_qp(MPCB_DRIVER_PARAMS(SH)) {
    factorise<SH>(MPCB_DRIVER_ARGS);
    MPCB_PHASE(admm_init(c, ts, tid));
    const int max_iter = c.st->qp_max_iter, chk = c.st->qp_check_termination;
    for (int iter = 1; iter <= max_iter; iter++) {
        MPCB_PHASE(admm_rhs(c, ts, tid));
        MPCB_PHASE(admm_solve1(c, ts, tid));
        MPCB_PHASE(admm_solve2(c, tid));
        MPCB_PHASE(admm_solve3(c, ts, tid));
        MPCB_PHASE(admm_update(c, ts, tid));
        if (chk > 0 && iter % chk == 0) {
            MPCB_PHASE(admm_check_a(c, ts, tid));
            MPCB_PHASE(admm_check_b(c, ts, tid));
            MPCB_PHASE(admm_check_c(c, ts, tid, iter));
            if (c.isc[IS_QP_DONE]) break;
        }
    }
    // (no barrier needed between the flag read above and the next writer: admm_finish does not touch isc)
    MPCB_PHASE(admm_finish(c, ts, tid));
    MPCB_PHASE(if (tid == 0 && !c.isc[IS_QP_DONE]) c.isc[IS_QP_ITERS] = max_iter);
}

}  // namespace mpcb
