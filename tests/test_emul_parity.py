"""CPU checks of the KERNEL LOGIC through the test-only lockstep emulation (tests/emul): the same phase code the CUDA
kernel runs, executed thread by thread on the host, compared with the oracle.  Tolerances: north_star's (final time
1e-6 relative, nodes 1e-5 absolute, identical status); the observed agreement is ~1e-12."""
import numpy as np
import pytest

import emul_py as E
import mpc_motion_planner_b200 as M


@pytest.fixture(scope="module")
def setup(oracle):
    s = oracle.default_settings()
    return oracle, s


def _problems(kind, seed, n, margins):
    lim = M.default_limits(margins)
    s = M.default_settings()
    x0, xT = M.generate_problems(lim, kind, seed, n)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    return x0, xT, gx, gu, gT


def test_kkt_nested_dissection_matches_dense_solve(setup):
    """S^-1 rhs from the sweep-based block factorisation == numpy dense solve of the oracle's exported QP matrices."""
    O, s = setup
    lim = O.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
    d = O.dims(s)
    x0, xT, gx, gu, gT = _problems(0, 5, 1, [0.8, 0.8, 0.6, 0.9, 0.1])
    rng = np.random.default_rng(0)
    z = np.concatenate([gx[0], gu[0], [gT[0]]])
    lam = 0.1 * rng.standard_normal(d["ndual"])
    rhs = rng.standard_normal(d["nvar"])
    out, rhoA, rhoB = E.kkt_apply(s, lim, x0[0], xT[0], z, lam, rhs)
    A, H = O.qp_dense(s, z, lam)
    S = H + s.qp_sigma * np.eye(d["nvar"]) + np.diag(rhoB) + A.T @ np.diag(rhoA) @ A
    ref = np.linalg.solve(S, rhs)
    assert set(np.unique(rhoA)) == {0.1, 100.0} and rhoB[0] == 100.0 and rhoB[-1] == 0.1
    assert np.abs(out - ref).max() < 1e-9 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("kind,seed,margins", [(0, 1, [0.8, 0.8, 0.6, 0.9, 0.1]), (1, 0, [0.9, 0.9, 0.5, 0.9, 0.1])])
def test_emulated_kernel_matches_oracle(setup, kind, seed, margins):
    O, s = setup
    lim = O.default_limits(margins)
    x0, xT, gx, gu, gT = _problems(kind, seed, 3, margins)
    esx, esu, eT, elam, est = E.solve(s, lim, x0, xT, gx, gu, gT, want_lam=True)
    for i in range(3):
        osx, osu, oT, olam, ost = O.solve(s, lim, x0[i], xT[i], gx[i], gu[i], gT[i], want_lam=True)
        assert abs(oT - eT[i]) <= 1e-6 * abs(oT)
        assert np.abs(osx - esx[i]).max() < 1e-5 and np.abs(osu - esu[i]).max() < 1e-5
        assert np.abs(olam - elam[i]).max() < 1e-5 * max(1.0, np.abs(olam).max())
        assert (ost.status, ost.sqp_iters, list(ost.qp_iters), list(ost.qp_status), list(ost.ls_trials)) == \
               (est[i].status, est[i].sqp_iters, list(est[i].qp_iters), list(est[i].qp_status), list(est[i].ls_trials))
        assert list(ost.alpha) == list(est[i].alpha)
        assert abs(ost.max_violation - est[i].max_violation) < 1e-8 and abs(ost.l1_violation - est[i].l1_violation) < 1e-8


def test_emulated_kernel_warm_started_duals(setup):
    """guess_lam path (receding-horizon config): a second solve seeded with the first solve's primal + dual solution."""
    O, s = setup
    m = [0.9, 0.9, 0.5, 0.9, 0.1]
    lim = O.default_limits(m)
    x0, xT, gx, gu, gT = _problems(1, 7, 1, m)
    sx, su, sT, slam, _ = E.solve(s, lim, x0, xT, gx, gu, gT, want_lam=True)
    e2 = E.solve(s, lim, x0, xT, sx, su, sT, glam=slam, want_lam=True)
    o2 = O.solve(s, lim, x0[0], xT[0], sx[0], su[0], sT[0], glam=slam[0], want_lam=True)
    assert abs(o2[2] - e2[2][0]) <= 1e-6 * abs(o2[2])
    assert np.abs(o2[0] - e2[0][0]).max() < 1e-5 and np.abs(o2[1] - e2[1][0]).max() < 1e-5


@pytest.mark.parametrize("po,ns,adaptive", [(3, 12, 0), (5, 6, 0), (7, 4, 0), (3, 6, 1)])
def test_emulated_generic_path_matches_oracle(oracle, po, ns, adaptive):
    """BASELINE config 4 (other collocation shapes, robot_ocp.hpp:31-42) and adaptive rho: the shape-generic stage (mpcb_gadmm.h,
    factorise + ADMM in one kernel) in the lockstep emulation against the oracle."""
    O = oracle
    m = [0.8, 0.8, 0.6, 0.9, 0.1]
    s = O.default_settings(poly_order=po, num_seg=ns, qp_adaptive_rho=adaptive)
    lim = O.default_limits(m)
    ms = M.default_settings(poly_order=po, num_seg=ns)
    x0, xT = M.generate_problems(M.default_limits(m), 0, 1, 2)
    gx, gu, gT = M.guess_batch(ms, M.default_limits(m), x0, xT)
    esx, esu, eT, elam, est = E.solve(s, lim, x0, xT, gx, gu, gT, want_lam=True)
    for i in range(2):
        osx, osu, oT, olam, ost = O.solve(s, lim, x0[i], xT[i], gx[i], gu[i], gT[i], want_lam=True)
        assert abs(oT - eT[i]) <= 1e-6 * abs(oT)
        assert np.abs(osx - esx[i]).max() < 1e-5 and np.abs(osu - esu[i]).max() < 1e-5
        assert np.abs(olam - elam[i]).max() < 1e-5 * max(1.0, np.abs(olam).max())
        assert (ost.status, ost.sqp_iters, list(ost.qp_iters), list(ost.qp_status), list(ost.ls_trials), list(ost.alpha)) == \
               (est[i].status, est[i].sqp_iters, list(est[i].qp_iters), list(est[i].qp_status), list(est[i].ls_trials), list(est[i].alpha))
