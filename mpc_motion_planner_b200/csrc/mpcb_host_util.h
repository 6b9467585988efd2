// Host-side helpers shared by the C-ABI implementation and the test-only emulation: defaults and the per-joint
// bound table the kernel reads (MotionPlanner::set_constraint_margins / set_min_height,
// /root/reference/mpc_solver/motionPlanner.cpp:56-100; limits from robot_utils/pandaWrapper.hpp:29-40).
#pragma once
#include "mpcb.h"

namespace mpcb {

inline void default_settings(mpcb_settings* s) {
    s->poly_order = 3; s->num_seg = 6; s->sqp_max_iter = 2; s->qp_max_iter = 700; s->line_search_max_iter = 10;
    s->qp_check_termination = 25; s->qp_adaptive_rho = 0; s->qp_adaptive_rho_interval = 25;
    s->qp_eps_abs = 1e-3; s->qp_eps_rel = 1e-3; s->qp_rho = 0.1; s->qp_sigma = 1e-6; s->qp_alpha = 1.0;
    s->qp_adaptive_rho_tolerance = 5;
    s->sqp_eps_prim = 1e-3; s->sqp_eps_dual = 1e-3; s->sqp_tau = 0.5; s->sqp_eta = 0.25; s->hess_reg = 0.001;
}
inline void default_limits(mpcb_limits* l) {
    const double qmin[7] = {-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973};
    const double qmax[7] = {2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973};
    const double vmax[7] = {2.1750, 2.1750, 2.1750, 2.1750, 2.6100, 2.6100, 2.6100};
    const double amax[7] = {15.0, 7.5, 10.0, 12.5, 15.0, 20.0, 20.0};
    const double jmax[7] = {7500, 3750, 5000, 6250, 7500, 10000, 10000};
    const double tmax[7] = {87, 87, 87, 87, 12, 12, 12};
    for (int i = 0; i < 7; i++) {
        l->qmin[i] = qmin[i]; l->qmax[i] = qmax[i]; l->vmax[i] = vmax[i]; l->amax[i] = amax[i]; l->jmax[i] = jmax[i]; l->taumax[i] = tmax[i];
    }
    for (int i = 0; i < 5; i++) l->margins[i] = 1.0;
    l->min_height = 0.05; l->final_eps = 1e-2; l->p_lb = 0.0; l->p_ub = 10.0;
}
// 40 doubles: qlo(7) qhi(7) vlim(7) alim(7) taulim(7) minh eps plb pub pad   (layout = LM_* in mpcb_ctx.h)
inline void fill_limtab(const mpcb_limits* lim, double* t) {
    const double mp = lim->margins[0], mv = lim->margins[1], ma = lim->margins[2], mt = lim->margins[3];
    for (int j = 0; j < 7; j++) {
        const double safety = (1 - mp) * (lim->qmax[j] - lim->qmin[j]) / 2;     // motionPlanner.cpp:69-71
        t[j] = lim->qmin[j] + safety; t[7 + j] = lim->qmax[j] - safety;
        t[14 + j] = mv * lim->vmax[j]; t[21 + j] = ma * lim->amax[j]; t[28 + j] = mt * lim->taumax[j];
    }
    t[35] = lim->min_height; t[36] = lim->final_eps; t[37] = lim->p_lb; t[38] = lim->p_ub; t[39] = 0;
}
inline int settings_supported(const mpcb_settings* s, const char** why) {
    if (s->qp_adaptive_rho && !(s->qp_adaptive_rho_tolerance > 1.0)) { *why = "qp_adaptive_rho_tolerance must be > 1"; return 0; }
    if (s->sqp_max_iter < 1 || s->qp_max_iter < 1 || s->line_search_max_iter < 1) { *why = "iteration limits must be >= 1"; return 0; }
    if (s->line_search_max_iter > 15) { *why = "line_search_max_iter must be <= 15 (trial-point workspace)"; return 0; }
    const int po = s->poly_order, ns = s->num_seg;
    if (!((po == 3 && (ns == 6 || ns == 12)) || (po == 5 && ns == 6) || (po == 7 && ns == 4))) {
        *why = "(poly_order, num_seg) must be one of the instantiated shapes (3,6), (3,12), (5,6), (7,4)";
        return 0;
    }
    return 1;
}

// which solve path a settings struct takes: the (3,6) register/TMEM ADMM kernel, or the shape-generic fused stage (mpcb_gadmm.h)
inline int settings_generic_path(const mpcb_settings* s) { return !(s->poly_order == 3 && s->num_seg == 6) || s->qp_adaptive_rho != 0; }

}  // namespace mpcb
