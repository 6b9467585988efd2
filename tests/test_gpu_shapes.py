"""GPU parity tests of BASELINE config 4 -- other collocation shapes (POLY_ORDER, NUM_SEG), /root/reference/mpc_solver/robot_ocp.hpp:31-42 --
and of the adaptive-rho variant of boxADMM (instantiated at /root/reference/mpc_solver/motionPlanner.hpp:10-11):
the shape-generic CUDA path (mpcb_gadmm.h) through the C ABI against the CPU oracle on the same seeded inputs.
Tolerances: identical status / SQP and ADMM iteration counts / line-search record, final time 1e-6 relative, nodes 1e-5 absolute."""
import numpy as np
import pytest

import mpc_motion_planner_b200 as M

pytestmark = pytest.mark.gpu

BENCH_MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]      # examples/benchmark.cpp:9
OFFLINE_MARGINS = [0.9, 0.9, 0.5, 0.9, 0.1]    # examples/offline_trajectory.cpp:9


def _stats_tuple(st):
    return (int(st["status"]), int(st["sqp_iters"]), list(st["qp_iters"]), list(st["qp_status"]), list(st["ls_trials"]), list(st["alpha"]))


def _ostats_tuple(st):
    return (st.status, st.sqp_iters, list(st.qp_iters), list(st.qp_status), list(st.ls_trials), list(st.alpha))


def _run(oracle, kind, seed, margins, n, **kw):
    s = M.default_settings(**kw)
    lim = M.default_limits(margins)
    x0, xT = M.generate_problems(lim, kind, seed, n)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    sv = M.Solver(s, lim)
    try:
        sx, su, sT, slam, st = sv.solve(x0, xT, gx, gu, gT, want_lam=True)
        traj, acc = sv.sample(sx, su, sT, target=xT, npts=200)
    finally:
        sv.close()
    so = oracle.default_settings(**kw); lo = oracle.default_limits(margins)
    osx, osu, oT, ost = oracle.solve_batch(so, lo, x0, xT, gx, gu, gT, nthreads=8)
    assert np.all(np.abs(sT - oT) <= 1e-6 * np.abs(oT)), np.abs(sT - oT).max()
    assert np.abs(sx - osx).max() < 1e-5 and np.abs(su - osu).max() < 1e-5, (np.abs(sx - osx).max(), np.abs(su - osu).max())
    bad = [i for i in range(n) if _stats_tuple(st[i]) != _ostats_tuple(ost[i])]
    assert not bad, (bad[:10], _stats_tuple(st[bad[0]]), _ostats_tuple(ost[bad[0]]))
    o1 = oracle.solve(so, lo, x0[0], xT[0], gx[0], gu[0], gT[0], want_lam=True)
    assert np.abs(o1[3] - slam[0]).max() < 1e-5 * max(1.0, np.abs(o1[3]).max())
    # read-out + accounting on the shape's own interpolation nodes (get_MPC_trajectory<200>, benchmark.cpp:118-160)
    otraj = oracle.sample_solution(so, osx[0], osu[0], oT[0], npts=200)
    oacc = oracle.account(lo, otraj, xT[0])
    assert np.abs(traj[0] - otraj).max() < 1e-4        # torques amplify the 1e-5 node tolerance
    assert list(acc[0]["flags"]) == list(oacc.flags)
    return st


@pytest.mark.parametrize("po,ns", [(3, 12), (5, 6), (7, 4)])
@pytest.mark.parametrize("kind,seed,margins,n", [(0, 1, BENCH_MARGINS, 40), (1, 0, OFFLINE_MARGINS, 24)])
def test_shape_matches_oracle(oracle, po, ns, kind, seed, margins, n):
    _run(oracle, kind, seed, margins, n, poly_order=po, num_seg=ns)


@pytest.mark.parametrize("po,ns", [(3, 6), (5, 6)])
def test_adaptive_rho_matches_oracle(oracle, po, ns):
    """qp_adaptive_rho = 1: rho re-estimated at the residual checks, KKT re-factorised when it moved by more than the tolerance factor."""
    st = _run(oracle, 0, 1, BENCH_MARGINS, 32, poly_order=po, num_seg=ns, qp_adaptive_rho=1)
    st2 = _run(oracle, 1, 0, OFFLINE_MARGINS, 16, poly_order=po, num_seg=ns, qp_adaptive_rho=1)
    assert len(st) == 32 and len(st2) == 16


def test_adaptive_rho_changes_the_result(oracle):
    """the adaptive branch is really taken: same problems, different final times than with the fixed rho"""
    lim = M.default_limits(BENCH_MARGINS)
    x0, xT = M.generate_problems(lim, 0, 1, 8)
    res = []
    for ad in (0, 1):
        s = M.default_settings(qp_adaptive_rho=ad)
        gx, gu, gT = M.guess_batch(s, lim, x0, xT)
        sv = M.Solver(s, lim)
        res.append(sv.solve(x0, xT, gx, gu, gT)[2])
        sv.close()
    assert np.abs(res[0] - res[1]).max() > 1e-4


def test_unsupported_shape_is_rejected():
    with pytest.raises(M.MpcbError):
        M.Solver(M.default_settings(poly_order=4, num_seg=5), M.default_limits(BENCH_MARGINS))
