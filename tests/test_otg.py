"""The jerk-limited synchronised profile generator (csrc/mpcb_otg.h, the role of Ruckig<7> in the reference:
mpc_solver/motionPlanner.cpp:146-175) against the one Ruckig trajectory the reference embeds
(analysis/data_analysis.ipynb, extracted by tools/make_fixtures.py into tests/golden/embedded_trajectory.npz) and against
the properties any such profile must have.  Host code only: no GPU needed."""
import os

import numpy as np
import pytest

import mpc_motion_planner_b200 as M

OFFLINE_MARGINS = [0.9, 0.9, 0.5, 0.9, 0.1]    # examples/offline_trajectory.cpp:9
BENCH_MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]      # examples/benchmark.cpp:9


def test_reproduces_the_embedded_ruckig_trajectory():
    emb = np.load(os.path.join(os.path.dirname(__file__), "golden", "embedded_trajectory.npz"))
    lim = M.default_limits(OFFLINE_MARGINS)
    x0 = np.concatenate([emb["q_rk"][:, 0], emb["v_rk"][:, 0]])
    xT = np.concatenate([emb["target_q"], emb["target_v"]])
    rc, tr = M.otg_calculate(lim, x0, xT)
    assert rc == 0
    T_ref = emb["t_rk"][-1]                      # printed with 6 significant digits
    assert abs(tr.T - T_ref) <= 5e-6, (tr.T, T_ref)
    eq = ev = ea = 0.0
    for i in range(201):
        q, v, a = M.otg_at_time(tr, tr.T * i / 200)
        eq = max(eq, np.abs(q - emb["q_rk"][:, i]).max()); ev = max(ev, np.abs(v - emb["v_rk"][:, i]).max())
        ea = max(ea, np.abs(a - emb["a_rk"][:, i]).max())
    # 6 printed digits of values up to ~3 -> 5e-6 rounding on q, v (plus the rounding of the sample time); the acceleration
    # ramps have slope j = 375..1000 rad/s^3, so the 6-digit sample time alone moves a by up to ~1e-3
    assert eq < 2e-5 and ev < 2e-5 and ea < 2e-3, (eq, ev, ea)


@pytest.mark.parametrize("kind,seed,margins", [(0, 1, BENCH_MARGINS), (1, 0, OFFLINE_MARGINS)])
def test_profile_properties(kind, seed, margins):
    lim = M.default_limits(margins)
    x0, xT = M.generate_problems(lim, kind, seed, 64)
    vl = margins[1] * np.array(lim.vmax); al = margins[2] * np.array(lim.amax); jl = margins[4] * np.array(lim.jmax)
    n_ok = 0
    for i in range(64):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        if rc < 0:
            assert np.any(np.abs(x0[i, 7:]) > vl) or np.any(np.abs(xT[i, 7:]) > vl)      # only invalid input is rejected
            continue
        n_ok += 1
        t = np.array(tr.t); j = np.array(tr.j)
        assert np.all(t >= 0) and np.all(np.abs(j) <= jl[:, None] * (1 + 1e-12))
        assert np.allclose(t.sum(axis=1), tr.T, rtol=0, atol=1e-9)          # all joints end together
        q, v, a = M.otg_at_time(tr, 0.0)
        assert np.allclose(q, x0[i, :7], atol=1e-12) and np.allclose(v, x0[i, 7:], atol=1e-12) and np.all(a == 0)
        q, v, a = M.otg_at_time(tr, tr.T)
        assert np.allclose(q, xT[i, :7], atol=1e-8) and np.allclose(v, xT[i, 7:], atol=1e-8) and np.allclose(a, 0, atol=1e-6)
        for s in np.linspace(0, 1, 101):
            q, v, a = M.otg_at_time(tr, s * tr.T)
            assert np.all(np.abs(v) <= vl * (1 + 1e-9)) and np.all(np.abs(a) <= al * (1 + 1e-9))
        # at least one joint is at its own minimum time: shortening T must be infeasible for it -> status 0 means no growth
        assert tr.status in (0, 1)
    assert n_ok >= 48


def test_blocked_intervals_and_near_zero_cruise():
    """Synchronisation beyond the plain case: (i) a joint whose only way to last T is a cruise velocity far below the grid
    spacing (tiny displacement, non-zero end velocity: problems 106 and 400 of the benchmark batch), (ii) a duration inside
    a joint's blocked interval, where T must move to the end of the interval (status 1).  Every valid input gets a profile."""
    lim = M.default_limits(BENCH_MARGINS)
    x0, xT = M.generate_problems(lim, 0, 1, 1000)
    for i in (106, 400):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        assert rc == 0
        q, v, a = M.otg_at_time(tr, tr.T)
        assert np.allclose(q, xT[i, :7], atol=1e-9) and np.allclose(v, xT[i, 7:], atol=1e-9)
    assert all(M.otg_calculate(lim, x0[i], xT[i])[0] == 0 for i in range(0, 1000, 7))
    lim = M.default_limits(OFFLINE_MARGINS)
    vl = OFFLINE_MARGINS[1] * np.array(lim.vmax); al = OFFLINE_MARGINS[2] * np.array(lim.amax)
    x0, xT = M.generate_problems(lim, 1, 7, 4000)
    grown = 0
    for i in range(4000):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        assert rc in (0, 1)
        if rc == 1:
            grown += 1
            q, v, a = M.otg_at_time(tr, tr.T)
            assert np.allclose(q, xT[i, :7], atol=1e-9) and np.allclose(v, xT[i, 7:], atol=1e-9) and np.allclose(a, 0, atol=1e-6)
            assert np.allclose(np.array(tr.t).sum(axis=1), tr.T, rtol=0, atol=1e-9)
            for s in np.linspace(0, 1, 57):
                q, v, a = M.otg_at_time(tr, s * tr.T)
                assert np.all(np.abs(v) <= vl * (1 + 1e-9)) and np.all(np.abs(a) <= al * (1 + 1e-9))
    assert grown >= 1


def test_guess_batch_samples_the_profile():
    lim = M.default_limits(BENCH_MARGINS); s = M.default_settings()
    x0, xT = M.generate_problems(lim, 0, 1, 8)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    tn = M.time_nodes(s)
    for i in range(8):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        assert rc >= 0 and gT[i] == tr.T
        for k, tk in enumerate(tn):
            q, v, a = M.otg_at_time(tr, tk * tr.T)         # motionPlanner.cpp:156-165
            assert np.array_equal(gx[i, 14 * k:14 * k + 7], q) and np.array_equal(gx[i, 14 * k + 7:14 * k + 14], v)
            assert np.array_equal(gu[i, 7 * k:7 * k + 7], a)


def test_stationary_joints_are_hold_profiles():
    """Joints that do not move (same position, at rest at both ends) get a 'stay' profile and do not block the others -- moving
    a subset of the joints is the common real-world input and Ruckig accepts it (motionPlanner.cpp:146-160)."""
    m = [0.9, 0.9, 0.5, 0.9, 0.1]
    lim = M.default_limits(m)
    s = M.default_settings()
    x0, xT = M.generate_problems(lim, 1, 11, 1)
    x0 = x0[0].copy(); xT = xT[0].copy()
    x0[7:] = 0.0; xT[7:] = 0.0
    ref_rc, ref = M.otg_calculate(lim, x0, xT)
    assert ref_rc >= 0
    for held in ([3], [1, 2, 3, 4, 5, 6], [0, 6]):
        a = x0.copy(); b = xT.copy()
        for j in held:
            b[j] = a[j]
        rc, tr = M.otg_calculate(lim, a, b)
        assert rc >= 0, (held, rc)
        moving = [j for j in range(7) if j not in held]
        assert tr.T > 0
        for t in np.linspace(0.0, tr.T, 23):
            q, v, acc = M.otg_at_time(tr, t)
            for j in held:
                assert q[j] == a[j] and v[j] == 0.0 and acc[j] == 0.0
            assert np.all(np.abs(v) <= m[1] * np.array(lim.vmax) * (1 + 1e-9)) and np.all(np.abs(acc) <= m[2] * np.array(lim.amax) * (1 + 1e-9))
        q, v, acc = M.otg_at_time(tr, tr.T)
        assert np.abs(q - b[:7]).max() < 1e-9 and np.abs(v).max() < 1e-9
        # the warm start of the OCP samples the same profile (no cubic-Hermite fallback)
        gx, gu, gT = M.guess_batch(s, lim, a[None], b[None])
        assert abs(gT[0] - tr.T) < 1e-12
        assert np.abs(gx[0].reshape(19, 14)[:, held] - a[held]).max() == 0.0
    # nothing moves at all: duration 0 is not a usable guess for the final time, the fallback profile takes over
    rc, tr = M.otg_calculate(lim, x0, x0)
    assert rc >= 0 and tr.T == 0.0
    gx, gu, gT = M.guess_batch(s, lim, x0[None], x0[None])
    assert gT[0] > 0 and np.abs(gx[0].reshape(19, 14) - x0).max() < 1e-12
