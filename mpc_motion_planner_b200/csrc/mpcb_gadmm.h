// Shape-generic box-ADMM iteration (polympc boxADMM::solve restated in SURVEY.md A.5) on top of the factorisation of
// mpcb_qp.h, for every collocation shape Shape<PO, NS> (BASELINE config 4: /root/reference/mpc_solver/robot_ocp.hpp:31-42
// makes POLY_ORDER / NUM_SEG compile-time macros) and for the adaptive-rho variant of the solver (polympc's
// adaptive_rho / adaptive_rho_interval / adaptive_rho_tolerance settings).
//
// It runs inside the SQP-stage kernel right after the factorisation (same CTA, same per-thread roles: thread r owns
// constraint row r, thread i owns variable i of the kernel order), so the row / box state stays in registers and the
// operators W = -S_II^-1 (packed lower, per segment), F = S_II^-1 S_IB and C = -(Schur complement)^-1 are used exactly
// where the sweeps left them: in shared memory for the base shape, in the CTA's block of global memory (L2-resident,
// Layout::GMAT) for the larger shapes whose inverses do not fit on one SM.  The (3,6)-specialised register/TMEM kernel
// of mpcb_admm.h stays the fast path for the reference's own shape with a fixed rho.
//
//   rhs   = sigma x - h + rho_b z_b - y_b + A^T (rho_A z_A - y_A)                                              g_rhs
//   x~    = S^-1 rhs   via   t = W rhs_I ; r = rhs_B - F^T rhs_I ; x~_B = C^-1 r ; x~_I = t - F x~_B   g_solve1/2/3
//   z~_A  = A x~ ;  x = a x~ + (1-a) x ;  z = clip(a z~ + (1-a) z + y/rho) ;  y += rho (a z~ + (1-a) z_old - z)   g_update
// Every `check_termination` iterations the OSQP residual test runs (g_check_*); with adaptive rho the penalty is
// re-estimated there, and when it moved by more than the tolerance factor the rows / boxes get their new penalties and
// the KKT matrix is factorised again (x, z, y are kept).
#pragma once
#include "mpcb_qp.h"

namespace mpcb {

enum { SC_NP = SC_ED + 1, SC_ND, SC_RHO };      // scalars of the adaptive-rho update (slots behind SC_ED, < SC_COUNT)
enum { IS_REFAC = IS_SQP_DONE + 1 };
static_assert((int)SC_RHO < (int)SC_COUNT && (int)IS_REFAC < (int)IS_COUNT, "scalar slots");

template <class SH>
MPCB_DEV double g_row_times(const Ctx<SH>& c, const TS<SH>& ts, int r, const double* x) {   // (A x)_r
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
MPCB_DEV double g_col_times(const Ctx<SH>& c, const TS<SH>& ts, const double* y) {   // (A^T y)_i for i != p
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
// 32 partial sums of (A^T y)_p = sum_r A[r,p] y_r  by the last 32 threads (fixed order: the result does not depend on scheduling)
template <class SH>
MPCB_DEV void g_pcol_partial(const Ctx<SH>& c, int tid, const double* y, double* out) {
    if (tid < SH::NT - 32) return;
    const int j = tid - (SH::NT - 32);
    double s = 0;
    for (int r = j; r < SH::NEQ; r += 32) s += c.pcol[r] * y[r];
    for (int r = j; r < SH::NINEQ; r += 32) s += c.Jg[r * 22 + 21] * y[SH::NEQ + r];
    out[j] = s;
}

template <class SH>
MPCB_DEV void g_init(Ctx<SH>& c, TS<SH>& ts, int tid) {
    ts.zA = 0; ts.yA = 0; ts.x = 0; ts.zB = 0; ts.yB = 0;        // boxADMM starts cold (warm_start = false)
    if (tid < SH::VP) { c.w[tid] = 0; c.rhs[tid] = 0; c.xt[tid] = 0; }
    if (tid < UM_COUNT) c.umax[tid] = 0;
    if (tid == 0) { c.isc[IS_QP_DONE] = 0; c.isc[IS_QP_STATUS] = 2; c.isc[IS_QP_ITERS] = 0; c.isc[IS_REFAC] = 0; c.scal[SC_RHO] = c.st->qp_rho; }
}
template <class SH>
MPCB_DEV void g_rhs(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NVAR) {
        const double g = g_col_times(c, ts, c.w);
        const double h = (tid == SH::IP) ? 1.0 : 0.0;     // gradient of the Mayer cost p (robot_ocp.hpp:206)
        c.rhs[tid] = c.st->qp_sigma * ts.x - h + ts.rhoB * ts.zB - ts.yB + g;
    }
    g_pcol_partial(c, tid, c.w, c.psum);
}
// Larger shapes (Layout::GMAT): the packed factors are expanded once per factorisation into full, coalesced copies -- W_s as a
// square m x m array (thread = row l reads Wf[j m + l]: consecutive threads, consecutive addresses), F_s column-major (thread = row
// reads Fc[col m + i]), C square -- so that every operator load of the iteration is a full 256-byte warp request.
template <class SH>
MPCB_DEV void g_expand(Ctx<SH>& c, int tid) {
    if constexpr (Layout<SH>::GMAT) {
        for (int s = 0; s < SH::NS; s++) {
            const int m = SH::m(s), nb = SH::nb(s);
            const double* Wn = c.W + SH::woff(s);
            double* Wf = c.Wf + SH::wfoff(s);
            for (int e = tid; e < m * m; e += SH::NT) {
                const int j = e / m, l = e - j * m, hi = l > j ? l : j, lo = l > j ? j : l;
                Wf[e] = Wn[hi * (hi + 1) / 2 + lo];
            }
            const double* Fs = c.F + SH::foff(s);
            double* Fc = c.Fc + SH::foff(s);
            for (int e = tid; e < m * nb; e += SH::NT) {
                const int col = e / m, i = e - col * m;
                Fc[e] = Fs[i * nb + col];
            }
        }
        for (int e = tid; e < SH::NB * SH::NB; e += SH::NT) {
            const int j = e / SH::NB, b = e - j * SH::NB, hi = b > j ? b : j, lo = b > j ? j : b;
            c.Cf[e] = c.C[hi * (hi + 1) / 2 + lo];
        }
    }
}
template <class SH>
MPCB_DEV void g_solve1(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if constexpr (Layout<SH>::GMAT) {
        if (tid < SH::NI) {      // t = -(Wn rhs_I) from the full copy, four accumulators
            const int s = ts.vseg, l = ts.vloc, m = SH::m(s);
            const double* Wf = c.Wf + SH::wfoff(s) + l;
            const double* b = c.rhs + SH::ioff(s);
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            int j = 0;
            for (; j + 8 <= m; j += 8) {                 // eight operator loads in flight per thread: the stream is latency-bound otherwise
                const double* wj = Wf + j * m;
                const double w0 = wj[0], w1 = wj[m], w2 = wj[2 * m], w3 = wj[3 * m], w4 = wj[4 * m], w5 = wj[5 * m], w6 = wj[6 * m], w7 = wj[7 * m];
                a0 += w0 * b[j]; a1 += w1 * b[j + 1]; a2 += w2 * b[j + 2]; a3 += w3 * b[j + 3];
                a0 += w4 * b[j + 4]; a1 += w5 * b[j + 5]; a2 += w6 * b[j + 6]; a3 += w7 * b[j + 7];
            }
            for (; j < m; j++) a0 += Wf[j * m] * b[j];
            c.xt[tid] = -((a0 + a1) + (a2 + a3));
        }
        for (int e = tid - SH::NI; e >= 0 && e < SH::PARTSIZE; e += SH::NT) {      // F_s^T rhs_I, one boundary column per thread
            int s = 0, col = e;
            while (col >= SH::nb(s)) { col -= SH::nb(s); s++; }
            const int m = SH::m(s), nb = SH::nb(s);
            const double* Fs = c.F + SH::foff(s) + col;          // row-major F: the threads of a segment read consecutive columns
            const double* b = c.rhs + SH::ioff(s);
            double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
            int i = 0;
            for (; i + 4 <= m; i += 4) {
                const double f0 = Fs[i * nb], f1 = Fs[(i + 1) * nb], f2 = Fs[(i + 2) * nb], f3 = Fs[(i + 3) * nb];
                a0 += f0 * b[i]; a1 += f1 * b[i + 1]; a2 += f2 * b[i + 2]; a3 += f3 * b[i + 3];
            }
            for (; i < m; i++) a0 += Fs[i * nb] * b[i];
            c.part[e] = (a0 + a1) + (a2 + a3);
        }
        if (tid == SH::NT - 1) {
            double s = 0;
            for (int j = 0; j < 32; j++) s += c.psum[j];
            c.rhs[SH::IP] += s;
        }
        return;
    }
    if (tid < SH::NI) {      // t = -(Wn rhs_I), symmetric packed row/column walk
        const int s = ts.vseg, l = ts.vloc, m = SH::m(s);
        const double* Wn = c.W + SH::woff(s);
        const double* b = c.rhs + SH::ioff(s);
        double a0 = 0, a1 = 0;
        const double* row = Wn + l * (l + 1) / 2;
        int j = 0;
        for (; j + 1 <= l; j += 2) { a0 += row[j] * b[j]; a1 += row[j + 1] * b[j + 1]; }
        for (; j <= l; j++) a0 += row[j] * b[j];
        for (j = l + 1; j + 1 < m; j += 2) { a0 += Wn[j * (j + 1) / 2 + l] * b[j]; a1 += Wn[(j + 1) * (j + 2) / 2 + l] * b[j + 1]; }
        for (; j < m; j++) a0 += Wn[j * (j + 1) / 2 + l] * b[j];
        c.xt[tid] = -(a0 + a1);
    }
    for (int e = tid - SH::NI; e >= 0 && e < SH::PARTSIZE; e += SH::NT) {      // F_s^T rhs_I, one boundary column per thread
        int s = 0, col = e;
        while (col >= SH::nb(s)) { col -= SH::nb(s); s++; }
        const int m = SH::m(s), nb = SH::nb(s);
        const double* Fs = c.F + SH::foff(s) + col;
        const double* b = c.rhs + SH::ioff(s);
        double a0 = 0, a1 = 0;
        int i = 0;
        for (; i + 1 < m; i += 2) { a0 += Fs[i * nb] * b[i]; a1 += Fs[(i + 1) * nb] * b[i + 1]; }
        for (; i < m; i++) a0 += Fs[i * nb] * b[i];
        c.part[e] = a0 + a1;
    }
    if (tid == SH::NT - 1) {
        double s = 0;
        for (int j = 0; j < 32; j++) s += c.psum[j];
        c.rhs[SH::IP] += s;
    }
}
template <class SH>
MPCB_DEV double g_sep_rhs(const Ctx<SH>& c, int b) {     // r_b = rhs_B[b] - sum_s (F_s^T rhs_I)[b]
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
// r_B into the (free) tail of the `part` buffer's neighbour: the w slots are busy, so it goes to psum's sibling -- rB lives in
// colbuf, which is idle outside the sweeps
template <class SH>
MPCB_DEV void g_solve2a(Ctx<SH>& c, int tid) {
    if (tid < SH::NB) c.colbuf[tid] = g_sep_rhs(c, tid);
}
template <class SH>
MPCB_DEV void g_solve2b(Ctx<SH>& c, int tid) {       // x~_B = -(Cn r), one thread per row
    if (tid >= SH::NB) return;
    const int b = tid;
    const double* r = c.colbuf;
    if constexpr (Layout<SH>::GMAT) {                  // full copy: thread b reads Cf[j NB + b]
        const double* Cf = c.Cf + b;
        double a0 = 0, a1 = 0, a2 = 0, a3 = 0;
        int j = 0;
        for (; j + 4 <= SH::NB; j += 4) {
            const double c0 = Cf[j * SH::NB], c1 = Cf[(j + 1) * SH::NB], c2 = Cf[(j + 2) * SH::NB], c3 = Cf[(j + 3) * SH::NB];
            a0 += c0 * r[j]; a1 += c1 * r[j + 1]; a2 += c2 * r[j + 2]; a3 += c3 * r[j + 3];
        }
        for (; j < SH::NB; j++) a0 += Cf[j * SH::NB] * r[j];
        c.xt[SH::NI + b] = -((a0 + a1) + (a2 + a3));
        return;
    }
    const double* row = c.C + b * (b + 1) / 2;
    double a0 = 0, a1 = 0;
    int j = 0;
    for (; j + 1 <= b; j += 2) { a0 += row[j] * r[j]; a1 += row[j + 1] * r[j + 1]; }
    for (; j <= b; j++) a0 += row[j] * r[j];
    for (j = b + 1; j + 1 < SH::NB; j += 2) { a0 += c.C[j * (j + 1) / 2 + b] * r[j]; a1 += c.C[(j + 1) * (j + 2) / 2 + b] * r[j + 1]; }
    for (; j < SH::NB; j++) a0 += c.C[j * (j + 1) / 2 + b] * r[j];
    c.xt[SH::NI + b] = -(a0 + a1);
}
template <class SH>
MPCB_DEV void g_solve3(Ctx<SH>& c, TS<SH>& ts, int tid) {     // x~_I = t - F x~_B
    if (tid >= SH::NI) return;
    const int s = ts.vseg, nb = SH::nb(s);
    const double* xb = c.xt + SH::NI;
    if constexpr (Layout<SH>::GMAT) {                  // column-major copy: thread = row i reads Fc[col m + i]
        const int m = SH::m(s);
        const double* Fc = c.Fc + SH::foff(s) + ts.vloc;
        double a0 = 0, a1 = 0;
        int col = 0;
        if (s > 0) { const double* x1 = xb + 14 * (s - 1); for (int cc = 0; cc < 14; cc += 2) { a0 += Fc[(col + cc) * m] * x1[cc]; a1 += Fc[(col + cc + 1) * m] * x1[cc + 1]; } col += 14; }
        if (s < SH::NS - 1) { const double* x2 = xb + 14 * s; for (int cc = 0; cc < 14; cc += 2) { a0 += Fc[(col + cc) * m] * x2[cc]; a1 += Fc[(col + cc + 1) * m] * x2[cc + 1]; } col += 14; }
        c.xt[tid] -= (a0 + a1) + Fc[col * m] * xb[SH::NB - 1];
        return;
    }
    const double* Fr = c.F + SH::foff(s) + ts.vloc * nb;
    double acc = 0;
    int col = 0;
    if (s > 0) { const double* x1 = xb + 14 * (s - 1); for (int cc = 0; cc < 14; cc++) acc += Fr[col + cc] * x1[cc]; col += 14; }
    if (s < SH::NS - 1) { const double* x2 = xb + 14 * s; for (int cc = 0; cc < 14; cc++) acc += Fr[col + cc] * x2[cc]; col += 14; }
    acc += Fr[col] * xb[SH::NB - 1];
    c.xt[tid] -= acc;
}
template <class SH>
MPCB_DEV void g_update(Ctx<SH>& c, TS<SH>& ts, int tid) {
    const double al = c.st->qp_alpha;
    if (tid < SH::NA) {
        const double ax = g_row_times(c, ts, tid, c.xt);
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
MPCB_DEV void g_check_a(Ctx<SH>& c, TS<SH>& ts, int tid) {
    if (tid < SH::NA) c.rhs[tid] = ts.yA;       // rhs slot is free until the next g_rhs
    if (tid < SH::NVAR) c.xt[tid] = ts.x;
}
template <class SH>
MPCB_DEV void g_check_b(Ctx<SH>& c, TS<SH>& ts, int tid) {
    double rp = 0, nax = 0, nz = 0, nx = 0, rd = 0, nhx = 0, naty = 0;
    if (tid < SH::NA) {
        const double ax = g_row_times(c, ts, tid, c.xt);
        rp = fabs(ax - ts.zA); nax = fabs(ax); nz = fabs(ts.zA);
    }
    if (tid < SH::NVAR) {
        nx = fabs(ts.x); nz = dmax(nz, fabs(ts.zB)); rp = dmax(rp, fabs(ts.x - ts.zB));
        if (tid != SH::IP) {
            const double aty = g_col_times(c, ts, c.rhs) + ts.yB;
            const double hx = ts.hd * ts.x + ts.ha * c.xt[SH::IP];
            naty = fabs(aty); nhx = fabs(hx); rd = fabs(hx + aty);
        }
    }
    g_pcol_partial(c, tid, c.rhs, c.psum);
    if (tid >= SH::NT - 64 && tid < SH::NT - 32) {
        double s = 0;
        for (int i = tid - (SH::NT - 64); i < SH::NVAR - 1; i += 32) s += c.Ha[i] * c.xt[i];
        c.psum[32 + tid - (SH::NT - 64)] = s;
    }
    umax_accum(c.umax + UM_RP, rp); umax_accum(c.umax + UM_NAX, nax); umax_accum(c.umax + UM_NZ, nz); umax_accum(c.umax + UM_NX, nx);
    umax_accum(c.umax + UM_RD, rd); umax_accum(c.umax + UM_NHX, nhx); umax_accum(c.umax + UM_NATY, naty);
}
// thread IP: the p row of the dual residual, the test itself and -- with adaptive rho -- the new penalty (polympc boxADMM /
// OSQP: rho * sqrt((r_p / max-primal-norm) / (r_d / max-dual-norm)), clipped to [1e-6, 1e6], taken when it left [rho / tol, rho * tol])
template <class SH>
MPCB_DEV void g_check_c(Ctx<SH>& c, TS<SH>& ts, int tid, int iter) {
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
    c.isc[IS_REFAC] = 0;
    if (rp <= ep && rd <= ed) { c.isc[IS_QP_DONE] = 1; c.isc[IS_QP_STATUS] = 1; c.isc[IS_QP_ITERS] = iter; }
    else if (c.st->qp_adaptive_rho && c.st->qp_adaptive_rho_interval > 0 && iter % c.st->qp_adaptive_rho_interval == 0) {
        const double rho = c.scal[SC_RHO], tol = c.st->qp_adaptive_rho_tolerance;
        double rn = rho * sqrt((rp / (np_ + 1e-10)) / (rd / (nd_ + 1e-10)));
        rn = dmin(dmax(rn, 1e-6), 1e6);
        if (rn > rho * tol || rn < rho / tol) { c.scal[SC_RHO] = rn; c.isc[IS_REFAC] = 1; }
    }
    for (int j = 0; j < UM_COUNT; j++) c.umax[j] = 0;
}
// new penalties for every row / box (boxADMM's per-row rule, row_rho) and the vectors the factorisation reads
template <class SH>
MPCB_DEV void g_rerho(Ctx<SH>& c, TS<SH>& ts, int tid) {
    const double rho = c.scal[SC_RHO];
    if (tid < SH::NA) { ts.rho = row_rho(ts.l, ts.u, rho); c.xt[tid] = ts.rho; c.w[tid] = ts.rho * ts.zA - ts.yA; }
    if (tid < SH::NVAR) ts.rhoB = row_rho(ts.lb, ts.ub, rho);
}
// the QP result goes to the OCP's result block in global memory (kernel order; read by node_val_kernel and import_qp)
template <class SH>
MPCB_DEV void g_finish(Ctx<SH>& c, TS<SH>& ts, int tid, double* res, int max_iter) {
    typedef QBL<SH> Q;
    if (tid < SH::NA) res[Q::R_LAM + tid] = ts.yA;
    if (tid < SH::NVAR) { res[Q::R_LAM + SH::NA + tid] = ts.yB; res[tid] = ts.x; }
    if (tid == 0) {
        res[Q::R_INFO] = (double)c.isc[IS_QP_STATUS];
        res[Q::R_INFO + 1] = (double)(c.isc[IS_QP_DONE] ? c.isc[IS_QP_ITERS] : max_iter);
    }
}

// One QP at the current linearisation: factorise and iterate.  Result in `res` (layout QBL<SH>).
template <class SH>
MPCB_DEV void solve_qp_generic(MPCB_DRIVER_PARAMS(SH), double* res) {
    MPCB_PROF_DECL
    factorise<SH>(MPCB_DRIVER_ARGS);
    MPCB_PHASE(g_expand(c, tid));
    MPCB_PROF(25)
    MPCB_PHASE(g_init(c, ts, tid));
    const int max_iter = c.st->qp_max_iter, chk = c.st->qp_check_termination;
    for (int iter = 1; iter <= max_iter; iter++) {
        MPCB_PHASE(g_rhs(c, ts, tid));
        MPCB_PHASE(g_solve1(c, ts, tid));
        MPCB_PHASE(g_solve2a(c, tid));
        MPCB_PHASE(g_solve2b(c, tid));
        MPCB_PHASE(g_solve3(c, ts, tid));
        MPCB_PHASE(g_update(c, ts, tid));
        if (chk > 0 && iter % chk == 0) {
            MPCB_PHASE(g_check_a(c, ts, tid));
            MPCB_PHASE(g_check_b(c, ts, tid));
            MPCB_PHASE(g_check_c(c, ts, tid, iter));
            if (c.isc[IS_QP_DONE]) break;
            if (c.isc[IS_REFAC]) {
                MPCB_PHASE(g_rerho(c, ts, tid));
                factorise<SH>(MPCB_DRIVER_ARGS);
                MPCB_PHASE(g_expand(c, tid));
            }
        }
    }
    MPCB_PHASE(g_finish(c, ts, tid, res, max_iter));
    MPCB_PROF(6)
}

}  // namespace mpcb
