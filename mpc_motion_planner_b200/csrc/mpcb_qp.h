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
MPCB_DEV void lin_a(Ctx<SH>& c, TS<SH>& ts, int tid) {
    const double p = c.z[SH::IP], tsc = c.cheb->ts, reg = c.st->hess_reg, rho = c.st->qp_rho;
    if (tid == 0) { c.scal[SC_PBAR] = p; c.scal[SC_MTP] = -tsc * p; }
    // inequality rows: values + Jacobians were evaluated by node_jac_kernel (mpcb_node.h), one thread per node of the batch
    for (int e = tid; e < SH::NN * 176; e += SH::NT) c.Jg[e] = c.gJg[e];
    if (tid < SH::NINEQ) c.w[SH::NEQ + tid] = c.gGj[tid];
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
    if (tid >= SH::NT - 32) {
        double s = 0;
        for (int i = tid - (SH::NT - 32); i < SH::NVAR - 1; i += 32) s += fabs(c.Ha[i]);
        c.psum[tid - (SH::NT - 32)] = s;
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
    for (int e = tid; e < SH::MATSIZE + SH::BBSIZE; e += SH::NT) c.W[e] = 0.0;
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
    constexpr int GS = SH::NT / SH::NS, NV = SH::PO + 3, NP = NV * (NV + 1) / 2;
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
    constexpr int GS = SH::NT / SH::NS;
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
// ---- symmetric sweeps with register-resident tiles ---------------------------------------------------------------------
// The augmented matrix [[S_II, S_IB],[., S_BB]] of a segment (M = m + nb <= 85) is cut into 5 x 5 tiles of its lower
// triangle (diagonal tiles are kept full); one thread keeps one tile in registers for all m pivots.  Per pivot the threads
// that hold the pivot column publish it (and 1/pivot) in shared memory, one barrier, and every thread applies
//     a_ij -= a_ik a_kj / a_kk  (i, j != k) ;  a_ik = a_ik / a_kk ;  a_kk = -1 / a_kk
// to its tile: 25 FMAs against 10 shared-memory loads, no index arithmetic in the inner loop.
// Round 2: whole 5 x 5 BLOCKS of pivots are swept at a time (sweep_block_*): with K the diagonal tile of block P,
//     A_PP <- -K^-1 ;  A_iP <- A_iP K^-1 ;  A_ij <- A_ij - (A_iP K^-1) A_jP^T        (i, j outside P)
// which is the same map as the five single sweeps but costs three barriers instead of five and moves the five dependent
// divisions onto one thread (the holder of the diagonal tile inverts it in registers); the pivots behind the last full block
// (m mod 5 of them) go one by one.  MEASURED SLOWER on B200 (interior sweeps 768 k cycles per OCP against 451 k for the
// pivot-by-pivot sweep, profiles/r2p_phase_cycles.txt): the rank-5 update triples the shared-memory traffic per tile and the
// in-register inversion is a 900-cycle serial chain every block.  Kept behind MPCB_SWEEP_BLOCKED (default 0) as a record.  Sweep<SH>::PAR segments are swept
// at a time (three for the base shape, 153 threads each); the separator Schur complement C (kind == NS) goes through the same
// code afterwards.
// Result: W = -S_II^-1, F = S_II^-1 S_IB, S_BB - S_BI S_II^-1 S_IB  (resp. -C^-1), written back in place.
constexpr int TILE = 5;

template <class SH>
MPCB_DEV int sweep_m(int kind) { return kind < SH::NS ? SH::m(kind) : SH::NB; }            // number of pivots
template <class SH>
MPCB_DEV int sweep_M(int kind) { return kind < SH::NS ? SH::m(kind) + SH::nb(kind) : SH::NB; }
template <class SH>
#ifndef MPCB_SWEEP_BLOCKED
#define MPCB_SWEEP_BLOCKED 0
#endif
MPCB_DEV int sweep_nblocks(int kind) { return MPCB_SWEEP_BLOCKED ? sweep_m<SH>(kind) / TILE : 0; }   // full 5-pivot blocks (blocked sweep); the rest goes one by one
template <class SH>
MPCB_DEV double* sweep_ptr(const Ctx<SH>& c, int kind, int i, int j) {     // entry (i, j), i >= j, of matrix `kind`
    if (kind < SH::NS) return aug_ptr(c, kind, SH::m(kind), SH::nb(kind), i, j);
    return c.C + i * (i + 1) / 2 + j;
}
// which matrix (or -1) and which tile a thread works on in a round; kinds first..first+count-1 share the CTA, `slot` threads each
template <class SH>
MPCB_DEV void sweep_assign(TS<SH>& ts, int tid, int first, int count, int slot) {
    const int sl = tid / slot, t = tid - sl * slot;
    ts.tseg = -1; ts.tI = 0; ts.tJ = 0;
    if (sl >= count) return;
    const int kind = first + sl, nt = (sweep_M<SH>(kind) + TILE - 1) / TILE;
    if (t >= nt * (nt + 1) / 2) return;
    int I = 0;
    while ((I + 1) * (I + 2) / 2 <= t) I++;
    ts.tseg = kind; ts.tI = I; ts.tJ = t - I * (I + 1) / 2;
}
template <class SH>
MPCB_DEV double* sweep_col(const Ctx<SH>& c, int kind, int first, int buf) {
    return c.colbuf + (buf * Sweep<SH>::PAR + (kind - first)) * Sweep<SH>::COLS;
}
// publish column kn of the thread's matrix (entries it holds) into buffer buf; rn = kn - 5 * (kn / 5) is static
template <class SH, int RN>
MPCB_DEV void sweep_publish(const Ctx<SH>& c, const TS<SH>& ts, int first, int kn, int buf) {
    const int kind = ts.tseg;
    if (kind < 0 || kn >= sweep_m<SH>(kind)) return;
    const int Pn = kn / TILE;
    double* col = sweep_col(c, kind, first, buf);
    if (ts.tJ == Pn) {
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++) col[TILE * ts.tI + x] = ts.tile[TILE * x + RN];
        if (ts.tI == Pn) col[Sweep<SH>::COLS - 1] = 1.0 / ts.tile[TILE * RN + RN];
    } else if (ts.tI == Pn) {
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) col[TILE * ts.tJ + y] = ts.tile[TILE * RN + y];
    }
}
template <class SH>
MPCB_DEV void sweep_load_tiles(const Ctx<SH>& c, TS<SH>& ts, int tid, int first, int count, int slot) {
    sweep_assign(ts, tid, first, count, slot);
    const int kind = ts.tseg;
    if (kind < 0) return;
    const int M = sweep_M<SH>(kind);
    MPCB_UNROLL
    for (int x = 0; x < TILE; x++)
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) {
            const int i = TILE * ts.tI + x, j = TILE * ts.tJ + y;
            const int hi = i > j ? i : j, lo = i > j ? j : i;
            ts.tile[TILE * x + y] = (hi < M) ? *sweep_ptr(c, kind, hi, lo) : 0.0;
        }
}
// pivot kp = 5 P + R: update the tile from buffer (kp & 1), then publish column kp + 1 into the other buffer
template <class SH, int R>
MPCB_DEV void sweep_step(const Ctx<SH>& c, TS<SH>& ts, int first, int P) {
    const int kind = ts.tseg, kp = TILE * P + R;
    if (kind < 0 || kp >= sweep_m<SH>(kind) || P < sweep_nblocks<SH>(kind)) return;      // the full blocks went through sweep_block_*
    const double* col = sweep_col(c, kind, first, kp & 1);
    const double dinv = col[Sweep<SH>::COLS - 1];
    double ci[TILE], cj[TILE];
    MPCB_UNROLL
    for (int x = 0; x < TILE; x++) { ci[x] = col[TILE * ts.tI + x] * dinv; cj[x] = col[TILE * ts.tJ + x]; }
    MPCB_UNROLL
    for (int x = 0; x < TILE; x++)
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) ts.tile[TILE * x + y] -= ci[x] * cj[y];
    if (ts.tJ == P) {
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++) ts.tile[TILE * x + R] = ci[x];
    }
    if (ts.tI == P) {
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) ts.tile[TILE * R + y] = cj[y] * dinv;
        if (ts.tJ == P) ts.tile[TILE * R + R] = -dinv;
    }
    sweep_publish<SH, (R + 1) % TILE>(c, ts, first, kp + 1, (kp + 1) & 1);
}
template <class SH>
MPCB_DEV void sweep_store(const Ctx<SH>& c, const TS<SH>& ts) {
    const int kind = ts.tseg;
    if (kind < 0) return;
    const int M = sweep_M<SH>(kind);
    MPCB_UNROLL
    for (int x = 0; x < TILE; x++)
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) {
            const int i = TILE * ts.tI + x, j = TILE * ts.tJ + y;
            if (i < M && j <= i) *sweep_ptr(c, kind, i, j) = ts.tile[TILE * x + y];
        }
}
// ---- blocked sweep ------------------------------------------------------------------------------------------------------------
// scratch of matrix `kind`: raw panel [M5][5] | scaled panel [M5][5] | K^-1 [25]   (M5 = M rounded up to the tile size)
template <class SH>
MPCB_DEV int sweep_M5(int kind) { return (sweep_M<SH>(kind) + TILE - 1) / TILE * TILE; }
template <class SH>
MPCB_DEV double* sweep_scratch(const Ctx<SH>& c, int kind, int first) {
    if (Layout<SH>::GMAT) return c.colbuf + 2 * Sweep<SH>::PAR * Sweep<SH>::COLS + (kind - first) * Sweep<SH>::PANEL;
    return kind < SH::NS ? c.F + SH::foff(kind) : c.C;       // free while the sweep runs: the content sits in the register tiles
}
// (panels are stored k-major: entry (i, k) of a panel at [k * M5 + i], so that the five entries a tile row / column needs are
// contiguous; the loops below touch one k at a time to keep the register footprint at tile + 10 values)
// phase 1 of block P: the panel tiles publish their raw content, the diagonal tile is inverted in registers
template <class SH>
MPCB_DEV void sweep_block_a(const Ctx<SH>& c, TS<SH>& ts, int first, int P) {
    const int kind = ts.tseg;
    if (kind < 0 || P >= sweep_nblocks<SH>(kind)) return;
    double* pan = sweep_scratch(c, kind, first);
    const int M5 = sweep_M5<SH>(kind);
    if (ts.tJ == P && ts.tI > P) {                         // A_IP: entry (5 tI + x, y)
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++)
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++) pan[y * M5 + TILE * ts.tI + x] = ts.tile[TILE * x + y];
    } else if (ts.tI == P && ts.tJ < P) {                  // A_PJ = A_JP^T: entry (5 tJ + y, x)
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++)
            MPCB_UNROLL
            for (int y = 0; y < TILE; y++) pan[x * M5 + TILE * ts.tJ + y] = ts.tile[TILE * x + y];
    } else if (ts.tI == P && ts.tJ == P) {                 // K -> -K^-1 by five in-register sweeps; K^-1 goes to the scratch
        MPCB_UNROLL
        for (int k = 0; k < TILE; k++) {
            const double d = 1.0 / ts.tile[TILE * k + k];
            double ci[TILE];
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++) ci[x] = ts.tile[TILE * x + k] * d;
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++)
                MPCB_UNROLL
                for (int y = 0; y < TILE; y++)
                    if (x != k && y != k) ts.tile[TILE * x + y] -= ci[x] * ts.tile[TILE * k + y];
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++)
                if (x != k) { ts.tile[TILE * x + k] = ci[x]; ts.tile[TILE * k + x] = ci[x]; }
            ts.tile[TILE * k + k] = -d;
        }
        double* kinv = pan + 2 * TILE * M5;
        MPCB_UNROLL
        for (int e = 0; e < TILE * TILE; e++) kinv[e] = -ts.tile[e];
    }
}
// phase 2: the panel tiles are scaled by K^-1 (their final value) and published
template <class SH>
MPCB_DEV void sweep_block_b(const Ctx<SH>& c, TS<SH>& ts, int first, int P) {
    const int kind = ts.tseg;
    if (kind < 0 || P >= sweep_nblocks<SH>(kind)) return;
    const bool colp = ts.tJ == P && ts.tI > P, rowp = ts.tI == P && ts.tJ < P;
    if (!colp && !rowp) return;
    double* pan = sweep_scratch(c, kind, first);
    const int M5 = sweep_M5<SH>(kind);
    double* span = pan + TILE * M5;
    const double* kinv = pan + 2 * TILE * M5;
    if (colp) {                                            // row x of A_IP K^-1 only needs row x of A_IP
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++) {
            double r[TILE];
            MPCB_UNROLL
            for (int y = 0; y < TILE; y++) {
                double a = 0;
                MPCB_UNROLL
                for (int k = 0; k < TILE; k++) a += ts.tile[TILE * x + k] * kinv[TILE * k + y];
                r[y] = a;
            }
            MPCB_UNROLL
            for (int y = 0; y < TILE; y++) { ts.tile[TILE * x + y] = r[y]; span[y * M5 + TILE * ts.tI + x] = r[y]; }
        }
    } else {                                               // column y of K^-1 A_PJ only needs column y of A_PJ
        MPCB_UNROLL
        for (int y = 0; y < TILE; y++) {
            double r[TILE];
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++) {
                double a = 0;
                MPCB_UNROLL
                for (int k = 0; k < TILE; k++) a += kinv[TILE * x + k] * ts.tile[TILE * k + y];
                r[x] = a;
            }
            MPCB_UNROLL
            for (int x = 0; x < TILE; x++) { ts.tile[TILE * x + y] = r[x]; span[x * M5 + TILE * ts.tJ + y] = r[x]; }
        }
    }
}
// phase 3: every tile outside block row / column P takes the rank-5 update, one k at a time
template <class SH>
MPCB_DEV void sweep_block_c(const Ctx<SH>& c, TS<SH>& ts, int first, int P) {
    const int kind = ts.tseg;
    if (kind < 0 || P >= sweep_nblocks<SH>(kind) || ts.tI == P || ts.tJ == P) return;
    const double* pan = sweep_scratch(c, kind, first);
    const int M5 = sweep_M5<SH>(kind);
    const double* si = pan + TILE * M5 + TILE * ts.tI;     // scaled entries of block tI
    const double* pj = pan + TILE * ts.tJ;                 // raw entries of block tJ
    MPCB_UNROLL
    for (int k = 0; k < TILE; k++) {
        double a[TILE], b[TILE];
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++) { a[x] = si[k * M5 + x]; b[x] = pj[k * M5 + x]; }
        MPCB_UNROLL
        for (int x = 0; x < TILE; x++)
            MPCB_UNROLL
            for (int y = 0; y < TILE; y++) ts.tile[TILE * x + y] -= a[x] * b[y];
    }
}
// publish the first single pivot column behind the full blocks (the start of the per-pivot tail)
template <class SH>
MPCB_DEV void sweep_publish_tail(const Ctx<SH>& c, const TS<SH>& ts, int first) {
    const int kind = ts.tseg;
    if (kind < 0) return;
    const int kn = TILE * sweep_nblocks<SH>(kind);
    sweep_publish<SH, 0>(c, ts, first, kn, kn & 1);
}
// sweep matrices first .. first+count-1 (count <= Sweep::PAR, `slot` threads each) over all their pivots
template <class SH>
MPCB_DEV void sweep_round(MPCB_DRIVER_PARAMS(SH), int first, int count, int slot) {
    int nbmax = 0, tail_lo = 1 << 30, tail_hi = 0;
    for (int k = first; k < first + count; k++) {
        const int nbk = sweep_nblocks<SH>(k), mk = sweep_m<SH>(k);
        nbmax = nbk > nbmax ? nbk : nbmax;
        if (mk > TILE * nbk) { const int hi = (mk + TILE - 1) / TILE; tail_lo = nbk < tail_lo ? nbk : tail_lo; tail_hi = hi > tail_hi ? hi : tail_hi; }
    }
    MPCB_PHASE(sweep_load_tiles(c, ts, tid, first, count, slot));
    for (int P = 0; P < nbmax; P++) {
        MPCB_PHASE(sweep_block_a(c, ts, first, P));
        MPCB_PHASE(sweep_block_b(c, ts, first, P));
        MPCB_PHASE(sweep_block_c(c, ts, first, P));
    }
    if (tail_hi > 0) {
        MPCB_PHASE(sweep_publish_tail(c, ts, first));
        for (int P = tail_lo; P < tail_hi; P++) {
            MPCB_PHASE(sweep_step<SH, 0>(c, ts, first, P));
            MPCB_PHASE(sweep_step<SH, 1>(c, ts, first, P));
            MPCB_PHASE(sweep_step<SH, 2>(c, ts, first, P));
            MPCB_PHASE(sweep_step<SH, 3>(c, ts, first, P));
            MPCB_PHASE(sweep_step<SH, 4>(c, ts, first, P));
        }
    }
    MPCB_PHASE(sweep_store(c, ts));
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
    for (int e = tid; e < SH::NB * SH::NB; e += SH::NT) {
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
// KKT assembly: S block-wise into W (S_II, packed lower), F (S_IB, row-major) and BB (S_BB contributions)
template <class SH>
MPCB_DEV void fac_assemble(MPCB_DRIVER_PARAMS(SH)) {
    MPCB_PROF_DECL
    MPCB_PHASE(fac_zero(c, tid));
    MPCB_PHASE(fac_diag(c, ts, tid));
    for (int j = 0; j < SH::PO; j++)
        for (int cc = 0; cc < 14; cc++) MPCB_PHASE(fac_eq_row(c, tid, j, cc));
    for (int jn = 0; jn <= SH::PO; jn++) MPCB_PHASE(fac_ineq_node(c, tid, jn));
    MPCB_PROF(18)
}
// sweeps: interiors, then the separator Schur complement
template <class SH>
MPCB_DEV void fac_sweeps(MPCB_DRIVER_PARAMS(SH)) {
    MPCB_PROF_DECL
    constexpr int PAR = Sweep<SH>::PAR;
    for (int first = 0; first < SH::NS; first += PAR) sweep_round<SH>(MPCB_DRIVER_ARGS, first, SH::NS - first < PAR ? SH::NS - first : PAR, Sweep<SH>::SLOT_I);
    MPCB_PROF(19)
    MPCB_PHASE(fac_c_assemble(c, tid));
    sweep_round<SH>(MPCB_DRIVER_ARGS, SH::NS, 1, Sweep<SH>::SLOT_C);
    MPCB_PROF(20)
    // W now holds -S_II^-1 (packed lower, per segment), F = S_II^-1 S_IB, C = -(Schur complement)^-1
}
template <class SH>
MPCB_DEV void factorise(MPCB_DRIVER_PARAMS(SH)) {
    fac_assemble<SH>(MPCB_DRIVER_ARGS);
    fac_sweeps<SH>(MPCB_DRIVER_ARGS);
}

}  // namespace mpcb
