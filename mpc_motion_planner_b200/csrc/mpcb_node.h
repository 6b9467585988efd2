// Per-node rigid-body work of the OCP, one thread per (OCP, node[, line-search trial]) across the WHOLE batch.
//
// The inequality constraints of minTime_ocp are per node (robot_ocp.hpp:80-163): 7 RNEA torques + the tool height and their
// hand-filled Jacobian.  One evaluation is a long serial recursion over the 7 joints (~2 kFLOP values, ~15 kFLOP
// derivatives) -- far too serial for a CTA that owns one OCP, but embarrassingly parallel over the 19 nodes x n OCPs
// (x 11 trial step sizes of the line search, polympc_redef.hpp:101-118).  These kernels run between the SQP-stage kernels
// and hand their results over through the per-OCP workspace.
#pragma once
#include "mpcb_rbd.h"
#include "mpcb_shape.h"

namespace mpcb {

// derivative pass at node k of iterate z (kernel order): Jg (8 x 22) and g (8)
template <class SH>
MPCB_HD void node_jac(const double* z, int k, rbd::DerivScratch& S, double* Jg, double* g) {
    double q[7], v[7], u[7];
    for (int j = 0; j < 7; j++) { q[j] = z[SH::kvar(k, j)]; v[j] = z[SH::kvar(k, 7 + j)]; u[j] = z[SH::kvar(k, 14 + j)]; }
    rbd::node_constraints_jac(q, v, u, S, g, Jg);
}
// step size of line-search trial t >= 1: 1, tau, tau^2, ..  (the products are formed exactly as the search forms them)
MPCB_HD double trial_alpha(int t, double tau) {
    double a = 1.0;
    for (int i = 1; i < t; i++) a = tau * a;
    return a;
}
// values at node k of z (t == 0) or of z + alpha_t p (t >= 1)
template <class SH>
MPCB_HD void node_val(const double* z, const double* p, int t, double tau, int k, double* g) {
    double q[7], v[7], u[7];
    const double alpha = trial_alpha(t, tau);
    for (int j = 0; j < 7; j++) {
        const int iq = SH::kvar(k, j), iv = SH::kvar(k, 7 + j), iu = SH::kvar(k, 14 + j);
        q[j] = z[iq]; v[j] = z[iv]; u[j] = z[iu];
        if (t > 0) { q[j] += alpha * p[iq]; v[j] += alpha * p[iv]; u[j] += alpha * p[iu]; }
    }
    rbd::node_constraints_value(q, v, u, g);
}

}  // namespace mpcb
