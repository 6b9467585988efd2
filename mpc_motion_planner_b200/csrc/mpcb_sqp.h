// SQP outer loop of one OCP: polympc SQPBase::solve with MySolver's customisation.
//   * l1-merit Armijo line search, mu = |lambda_qp|_inf  (/root/reference/mpc_solver/polympc_redef.hpp:73-121)
//   * full exact re-linearisation after every step          (polympc_redef.hpp:141-147)
//   * settings from MotionPlanner's constructor            (mpc_solver/motionPlanner.cpp:15-20)
// and the guess/solution plumbing of MotionPlanner::warm_start_RK / solve_trajectory (motionPlanner.cpp:156-207).
#pragma once
#include "mpcb_export.h"
#include "mpcb_gadmm.h"
#include "mpcb_node.h"

namespace mpcb {

// One problem of the batch (all pointers: global memory on the device, plain host memory in the emulation).
struct Problem {
    const double* limtab;   // 40 doubles: qlo(7) qhi(7) vlim(7) alim(7) taulim(7) minh eps plb pub pad
    const double *x0, *xT, *gx, *gu, *glam;   // glam may be null
    double gT;
    double *sx, *su, *sT, *slam;              // slam may be null
    mpcb_stats* stats;                        // may be null
};

template <class SH>
MPCB_DEV int ext_index(int k, int comp) {   // polympc layout index of kernel variable (k, comp); k < 0 -> p
    if (k < 0) return SH::NVAR - 1;
    return (comp < 14) ? 14 * k + comp : 14 * SH::NN + 7 * k + (comp - 14);
}

template <class SH>
MPCB_DEV void sqp_load_a(Ctx<SH>& c, int tid, const Problem& pb) {
    if (tid < 40) c.lims[tid] = pb.limtab[tid];
    if (tid < 14) { c.lims[LM_X0 + tid] = pb.x0[tid]; c.lims[LM_XT + tid] = pb.xT[tid]; }
    if (tid < UM_COUNT) c.umax[tid] = 0;
    if (tid == 0) c.isc[IS_SQP_DONE] = 0;
}
template <class SH>
MPCB_DEV void sqp_load_b(Ctx<SH>& c, TS<SH>& ts, int tid, const Problem& pb) {
    if (tid < SH::NVAR) {
        const int k = ts.vk, comp = ts.vcomp;
        c.z[tid] = (k < 0) ? pb.gT : (comp < 14 ? pb.gx[14 * k + comp] : pb.gu[7 * k + comp - 14]);
        c.lam[SH::NA + tid] = pb.glam ? pb.glam[SH::NA + ext_index<SH>(k, comp)] : 0.0;
    }
    if (tid < SH::NA) c.lam[tid] = pb.glam ? pb.glam[tid] : 0.0;
}

// ---- constraint violation at z + alpha p  (polympc constraints_violation, l1 and max) ------------------
template <class SH>
MPCB_DEV void vio_a(Ctx<SH>& c, TS<SH>& ts, int tid, double alpha, bool use_p, int trial) {
    const double* P = c.qp_p;
    // inequality values at this trial point: evaluated by node_val_kernel (mpcb_node.h)
    if (tid < SH::NINEQ) c.w[SH::NEQ + tid] = c.gval[trial * SH::NINEQ + tid];
    if (tid < SH::NEQ) {
        double acc = 0;
        MPCB_UNROLL
        for (int l = 0; l <= SH::PO; l++) {
            const int i = ts.ridx[l];
            acc += ts.dco[l] * (use_p ? c.z[i] + alpha * P[i] : c.z[i]);
        }
        const int fi = ts.ridx[SH::PO + 1];
        const double f = use_p ? c.z[fi] + alpha * P[fi] : c.z[fi];
        const double p = use_p ? c.z[SH::IP] + alpha * P[SH::IP] : c.z[SH::IP];
        c.rhs[tid] = fabs(acc - c.cheb->ts * p * f);
    }
    if (tid < SH::NVAR) {
        const double zi = use_p ? c.z[tid] + alpha * P[tid] : c.z[tid];
        double lo, hi;
        var_bounds(c, ts.vk, ts.vcomp, lo, hi);
        c.xt[tid] = dmax(0.0, lo - zi) + dmax(0.0, zi - hi);
    }
}
template <class SH>
MPCB_DEV void vio_b(Ctx<SH>& c, int tid) {
    if (tid >= SH::NEQ && tid < SH::NA) {
        const double g = c.w[tid];
        double lo, hi;
        ineq_bounds(c, (tid - SH::NEQ) % 8, lo, hi);
        c.rhs[tid] = dmax(0.0, lo - g) + dmax(0.0, g - hi);
    }
}
template <class SH>
MPCB_DEV void vio_c(Ctx<SH>& c, int tid) {
    double e = 0;
    if (tid < SH::NA) e = c.rhs[tid];
    if (tid < SH::NVAR) e = dmax(e, c.xt[tid]);
    umax_accum(c.umax + UM_V, e);
    if (tid >= SH::NT - 32) {
        double s = 0;
        for (int r = tid - (SH::NT - 32); r < SH::NA; r += 32) s += c.rhs[r];
        for (int i = tid - (SH::NT - 32); i < SH::NVAR; i += 32) s += c.xt[i];
        c.psum[tid - (SH::NT - 32)] = s;
    }
}
template <class SH>
MPCB_DEV void vio_d(Ctx<SH>& c, int tid) {
    if (tid != 0) return;
    double s = 0;
    for (int j = 0; j < 32; j++) s += c.psum[j];
    c.scal[SC_V_L1] = s;
    c.scal[SC_V_LINF] = umax_get(c.umax + UM_V);
    c.umax[UM_V] = 0;
}
template <class SH>
MPCB_DEV void violation(MPCB_DRIVER_PARAMS(SH), double alpha, bool use_p, int trial) {
    MPCB_PHASE(vio_a(c, ts, tid, alpha, use_p, trial));
    MPCB_PHASE(vio_b(c, tid));
    MPCB_PHASE(vio_c(c, tid));
    MPCB_PHASE(vio_d(c, tid));
}

template <class SH>
MPCB_DEV void linearise(MPCB_DRIVER_PARAMS(SH)) {
    MPCB_PHASE(lin_a(c, ts, tid));
    MPCB_PHASE(lin_b(c, ts, tid));
    MPCB_PHASE(lin_c(c, ts, tid));
}

template <class SH>
MPCB_DEV void sqp_mu(Ctx<SH>& c, int tid) {
    double m = 0;
    for (int i = tid; i < SH::NDUAL; i += SH::NT) m = dmax(m, fabs(c.qp_lam[i]));
    umax_accum(c.umax + UM_B, m);
}
template <class SH>
MPCB_DEV void sqp_step(Ctx<SH>& c, int tid, double alpha) {
    double pn = 0, dn = 0;
    if (tid < SH::NVAR) { const double p = c.qp_p[tid]; c.z[tid] += alpha * p; pn = fabs(p); }
    for (int i = tid; i < SH::NDUAL; i += SH::NT) {
        const double d = c.qp_lam[i] - c.lam[i];
        c.lam[i] += alpha * d;
        dn = dmax(dn, fabs(d));
    }
    umax_accum(c.umax + UM_A, pn);
    umax_accum(c.umax + UM_B, dn);
}
template <class SH>
MPCB_DEV void sqp_store(Ctx<SH>& c, TS<SH>& ts, int tid, const Problem& pb) {
    if (tid < SH::NVAR) {
        const int k = ts.vk, comp = ts.vcomp;
        const double v = c.z[tid];
        if (k < 0) *pb.sT = v;
        else if (comp < 14) pb.sx[14 * k + comp] = v;
        else pb.su[7 * k + comp - 14] = v;
        if (pb.slam) pb.slam[SH::NA + ext_index<SH>(k, comp)] = c.lam[SH::NA + tid];
    }
    if (tid < SH::NA && pb.slam) pb.slam[tid] = c.lam[tid];
}

// The solve of one OCP is cut into stages that run as separate kernels (each with its own register budget and grid):
//   stage_load    : guess -> iterate z, multipliers, flags, stats                                      (once)
//   node_jac      : inequality values + Jacobians at z, one thread per node of the batch (mpcb_node.h)  (per linearisation)
//   stage_linfac  : rest of the linearisation, KKT factorisation, export of the QP block               (per linearisation)
//   admm_run      : the box-ADMM iterations of the QP (mpcb_admm.h)                                    (per SQP iteration)
//   node_val      : inequality values at every line-search trial point z + alpha_t p                  (per SQP iteration)
//   stage_lsstep  : merit line search, step, termination test of SQP iteration `it`; final store      (per SQP iteration)
// `flags[0]` (global, per OCP) is set to 1 when the OCP is finished; later stages skip it.  `stats` is written by thread 0.
template <class SH>
MPCB_DEV void stage_load(MPCB_DRIVER_PARAMS(SH), const Problem& pb, int* flags) {
    mpcb_stats* S = pb.stats;
    MPCB_PROF_DECL
    MPCB_PHASE(sqp_load_a(c, tid, pb));
    MPCB_PHASE(sqp_load_b(c, ts, tid, pb));
    MPCB_PHASE(if (tid == 0) { flags[0] = 0; if (S) { memset(S, 0, sizeof(mpcb_stats)); S->status = 2; } });
    MPCB_PROF(16)
}

template <class SH>
MPCB_DEV void stage_linfac(MPCB_DRIVER_PARAMS(SH), const Problem& pb, double* qb) {
    MPCB_PROF_DECL
    MPCB_PHASE(sqp_load_a(c, tid, pb));
    linearise<SH>(MPCB_DRIVER_ARGS);
    MPCB_PROF(17)
    fac_assemble<SH>(MPCB_DRIVER_ARGS);
    export_pre<SH>(MPCB_DRIVER_ARGS, qb);          // S_BI and the row of p, before the sweeps turn S_IB into F
    fac_sweeps<SH>(MPCB_DRIVER_ARGS);
    MPCB_PROF(25)
    export_qp<SH>(MPCB_DRIVER_ARGS, qb);
    MPCB_PROF(21)
}

// shape-generic / adaptive-rho variant of stage_linfac: the QP is solved in the same kernel (mpcb_gadmm.h), the result goes to
// the OCP's result block
template <class SH>
MPCB_DEV void stage_qp_generic(MPCB_DRIVER_PARAMS(SH), const Problem& pb, double* qb) {
    MPCB_PROF_DECL
    MPCB_PHASE(sqp_load_a(c, tid, pb));
    linearise<SH>(MPCB_DRIVER_ARGS);
    MPCB_PROF(17)
    solve_qp_generic<SH>(MPCB_DRIVER_ARGS, qb + QBL<SH>::G_RES);
}

template <class SH>
MPCB_DEV void stage_lsstep(MPCB_DRIVER_PARAMS(SH), const Problem& pb, const double* qb, int* flags, const int it) {
    const mpcb_settings& st = *c.st;
    mpcb_stats* S = pb.stats;
    MPCB_PROF_DECL
    MPCB_PHASE(sqp_load_a(c, tid, pb));
    MPCB_PHASE(import_qp(c, tid, qb));
    MPCB_PHASE(sqp_mu(c, tid));
    violation<SH>(MPCB_DRIVER_ARGS, 0.0, false, 0);
    MPCB_PHASE(if (tid == 0) {
        const double mu = umax_get(c.umax + UM_B), l1 = c.scal[SC_V_L1];
        c.umax[UM_B] = 0;
        c.scal[SC_MU] = mu;
        c.scal[SC_PHI0] = c.z[SH::IP] + mu * l1;
        c.scal[SC_DPHI] = c.qp_p[SH::IP] - mu * l1;      // grad f . p = p_p  (cost = p)
        c.scal[SC_ALPHA] = 1.0;
        c.isc[IS_LS_ACCEPT] = 0;
        if (S) {
            S->qp_iters_total += (int)c.isc[IS_QP_ITERS];
            if (it <= 4) { S->qp_status[it - 1] = (int)c.isc[IS_QP_STATUS]; S->qp_iters[it - 1] = (int)c.isc[IS_QP_ITERS]; }
        }
    });
    // step_size_selection_impl: at most line_search_max_iter-1 trials; the last alpha is returned even if rejected
    int trials = 0, trial_used = st.line_search_max_iter;          // index of the step size finally taken (see trial_alpha)
    for (int i = 1; i < st.line_search_max_iter; i++) {
        trials++;
        const double alpha = c.scal[SC_ALPHA];
        violation<SH>(MPCB_DRIVER_ARGS, alpha, true, i);
        MPCB_PHASE(if (tid == 0) {
            const double a = c.scal[SC_ALPHA];
            const double phis = (c.z[SH::IP] + a * c.qp_p[SH::IP]) + c.scal[SC_MU] * c.scal[SC_V_L1];
            if (phis <= c.scal[SC_PHI0] + a * st.sqp_eta * c.scal[SC_DPHI]) c.isc[IS_LS_ACCEPT] = 1;
            else c.scal[SC_ALPHA] = st.sqp_tau * a;
        });
        if (c.isc[IS_LS_ACCEPT]) { trial_used = i; break; }
    }
    const double alpha = c.scal[SC_ALPHA];
    MPCB_PROF(22)
    MPCB_PHASE(sqp_step(c, tid, alpha));
    violation<SH>(MPCB_DRIVER_ARGS, 0.0, false, trial_used);      // at the new iterate = the trial point of the step taken
    MPCB_PHASE(if (tid == 0) {
        const double pn = umax_get(c.umax + UM_A), dn = umax_get(c.umax + UM_B);
        c.umax[UM_A] = 0; c.umax[UM_B] = 0;
        const bool done = alpha * pn <= st.sqp_eps_prim && alpha * dn <= st.sqp_eps_dual && c.scal[SC_V_LINF] <= st.sqp_eps_prim;
        c.isc[IS_SQP_DONE] = done ? 1 : 0;
        if (S) {
            S->sqp_iters = it; S->primal_step_norm = alpha * pn; S->dual_step_norm = alpha * dn;
            S->max_violation = c.scal[SC_V_LINF]; S->l1_violation = c.scal[SC_V_L1]; S->cost = c.z[SH::IP];
            if (it <= 4) { S->alpha[it - 1] = alpha; S->ls_trials[it - 1] = trials; }
            if (done) S->status = 1;
        }
    });
    MPCB_PROF(23)
    if (c.isc[IS_SQP_DONE] || it >= st.sqp_max_iter) {      // no re-linearisation after the last step
        MPCB_PHASE({ sqp_store(c, ts, tid, pb); if (tid == 0) flags[0] = 1; });
        MPCB_PROF(24)
    }
}

}  // namespace mpcb
