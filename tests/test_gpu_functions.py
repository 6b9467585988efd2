"""Device-side per-function parity tests (layers L1 / L2): the pieces of the path one at a time against the oracle,
so that a regression localises instead of only showing up in a whole-solve comparison.
  (a) minTime_ocp::evalConstraints<double> / <ad_scalar_t> (/root/reference/mpc_solver/robot_ocp.hpp:80-163) as the node kernels evaluate
      them, on 1000 random states                                                                        -- <= 1e-11 relative
  (b) linearisation + KKT assembly + factorisation + one KKT solve: the QP iterate after ONE ADMM iteration is S^-1 (-h)     -- 1e-9
  (c) the ADMM iterate after 1, 25 and 100 iterations against the oracle's boxADMM on the oracle's own linearisation         -- 1e-9
(b) and (c) go through the public entry point with sqp_max_iter = 1 and qp_max_iter = k: the step the solver took is recovered from
the returned iterate and the reported step size."""
import numpy as np
import pytest

import mpc_motion_planner_b200 as M

pytestmark = pytest.mark.gpu

BENCH_MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]


def test_node_constraints_value_and_jacobian(oracle):
    lim = M.default_limits(BENCH_MARGINS)
    rng = np.random.default_rng(7)
    n = 1000
    q = np.array(lim.qmin) + rng.random((n, 7)) * (np.array(lim.qmax) - np.array(lim.qmin))
    v = (2 * rng.random((n, 7)) - 1) * np.array(lim.vmax)
    a = (2 * rng.random((n, 7)) - 1) * np.array(lim.amax)
    sv = M.Solver(M.default_settings(), lim)
    try:
        gv, gj, Jg = sv.debug_node_constraints(np.hstack([q, v, a]))
    finally:
        sv.close()
    worst_g = worst_J = 0.0
    for i in range(n):
        og, oJ = oracle.node_constraints(q[i], v[i], a[i])
        sg = max(1.0, np.abs(og).max()); sJ = max(1.0, np.abs(oJ).max())
        worst_g = max(worst_g, np.abs(gv[i] - og).max() / sg, np.abs(gj[i] - og).max() / sg)
        worst_J = max(worst_J, np.abs(Jg[i] - oJ).max() / sJ)
    assert worst_g <= 1e-11 and worst_J <= 1e-11, (worst_g, worst_J)


def _qp_step_on_gpu(po, ns, k, x0, xT, gx, gu, gT, adaptive=0):
    s = M.default_settings(poly_order=po, num_seg=ns, sqp_max_iter=1, qp_max_iter=k, qp_adaptive_rho=adaptive)
    sv = M.Solver(s, M.default_limits(BENCH_MARGINS))
    try:
        sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
    finally:
        sv.close()
    z0 = np.concatenate([gx, gu, [gT]]); z1 = np.concatenate([sx[0], su[0], sT])
    alpha = float(st[0]["alpha"][0])
    return (z1 - z0) / alpha, alpha, int(st[0]["qp_iters"][0])


def _qp_step_oracle(O, po, ns, k, x0, xT, gx, gu, gT, adaptive=0):
    s = O.default_settings(poly_order=po, num_seg=ns, sqp_max_iter=1, qp_max_iter=k, qp_adaptive_rho=adaptive)
    lim = O.default_limits(BENCH_MARGINS)
    d = O.dims(s)
    z = np.concatenate([gx, gu, [gT]]); lam = np.zeros(d["ndual"])
    c_eq, g, Jg, Hd, Ha = O.linearise(s, z, lam)
    lbx, ubx, lbg, ubg = O.bounds(s, lim, x0, xT)
    lA = np.concatenate([-c_eq, np.tile(lbg, d["nn"]) - g]); uA = np.concatenate([-c_eq, np.tile(ubg, d["nn"]) - g])
    h = np.zeros(d["nvar"]); h[-1] = 1.0
    st, it, x, y, res = O.qp_admm(s, z, Jg, Hd, Ha, h, lA, uA, lbx - z, ubx - z)
    return x, it


@pytest.mark.parametrize("po,ns", [(3, 6), (5, 6)])
@pytest.mark.parametrize("k", [1, 25, 100])
def test_admm_iterate_trace(oracle, po, ns, k):
    """(b) for k = 1 -- x = S^-1 (-h): linearisation, KKT assembly, nested-dissection factorisation and one solve -- and (c) for k = 25,
    100: the ADMM iterate itself, both code paths (the (3,6) register/TMEM kernel and the shape-generic fused stage)."""
    lim = M.default_limits(BENCH_MARGINS)
    s = M.default_settings(poly_order=po, num_seg=ns)
    x0, xT = M.generate_problems(lim, 0, 3, 3)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    for i in range(3):
        p_gpu, alpha, its = _qp_step_on_gpu(po, ns, k, x0[i], xT[i], gx[i], gu[i], gT[i])
        p_orc, oits = _qp_step_oracle(oracle, po, ns, k, x0[i], xT[i], gx[i], gu[i], gT[i])
        assert its == oits == k
        scale = max(1.0, np.abs(p_orc).max())
        # the recovered step carries the rounding of (z1 - z0) / alpha: 1e-9 relative to the step's size
        assert np.abs(p_gpu - p_orc).max() <= 1e-9 * scale / min(1.0, alpha), (i, k, np.abs(p_gpu - p_orc).max(), alpha)


def test_admm_iterate_trace_adaptive_rho(oracle):
    """the adaptive-rho branch at the first update (iteration 25) and beyond"""
    lim = M.default_limits(BENCH_MARGINS)
    s = M.default_settings()
    x0, xT = M.generate_problems(lim, 0, 3, 2)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    for k in (26, 60):
        for i in range(2):
            p_gpu, alpha, its = _qp_step_on_gpu(3, 6, k, x0[i], xT[i], gx[i], gu[i], gT[i], adaptive=1)
            p_orc, oits = _qp_step_oracle(oracle, 3, 6, k, x0[i], xT[i], gx[i], gu[i], gT[i], adaptive=1)
            assert its == oits
            assert np.abs(p_gpu - p_orc).max() <= 1e-8 * max(1.0, np.abs(p_orc).max()) / min(1.0, alpha)
