"""Pin the oracle's rigid-body part on the reference-embedded known answers
and its derivatives on finite differences.  CPU only."""
import json
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def kat():
    return json.load(open(os.path.join(G, "kat.json")))


@pytest.fixture(scope="module")
def emb():
    return np.load(os.path.join(G, "embedded_trajectory.npz"))


def test_fk_kat(oracle, kat):
    k = kat["fk_oMi7"]
    R, p = oracle.fk(k["q"])
    assert np.abs(R[6] - np.array(k["R"])).max() < 1e-6      # 6 printed digits
    assert np.abs(p[6] - np.array(k["p"])).max() < 1e-6


def test_jacobian_kat(oracle, kat):
    k = kat["jac_joint7_world_aligned_times_ones"]
    J = oracle.frame_jacobian(k["q"], [0, 0, 0])
    assert np.abs(J @ np.ones(7) - np.array(k["Jqd"])).max() < 1e-8


def test_rnea_embedded_samples(oracle, emb):
    """402 (q,v,a)->tau samples printed with 6 significant digits; only matches with the 1 kg tool merged in."""
    errs = []
    for who in ("mpc", "rk"):
        for i in range(201):
            tau = oracle.rnea(emb["q_" + who][:, i], emb["v_" + who][:, i], emb["a_" + who][:, i])
            errs.append(np.abs(tau - emb["tau_" + who][:, i]).max())
    assert max(errs) < 5e-4 and np.median(errs) < 1e-4


def test_link8_fk_embedded(oracle, emb):
    for i in range(0, 201, 10):
        R, p = oracle.fk(emb["q_mpc"][:, i])
        pos = p[6] + R[6] @ np.array([0, 0, 0.107])
        assert np.abs(pos - emb["link8_xyz_mpc"][:, i]).max() < 1e-12


def test_rnea_derivatives_fd(oracle):
    rng = np.random.default_rng(1)
    h = 1e-6
    for _ in range(5):
        q = rng.uniform(-1.5, 1.5, 7); v = rng.uniform(-2, 2, 7); a = rng.uniform(-5, 5, 7)
        tau, dq, dv, M = oracle.rnea_derivatives(q, v, a)
        assert np.abs(tau - oracle.rnea(q, v, a)).max() < 1e-12
        for mat, idx in ((dq, 0), (dv, 1), (M, 2)):
            fd = np.zeros((7, 7))
            for j in range(7):
                e = np.zeros(7); e[j] = h
                args_p = [q, v, a]; args_m = [q, v, a]
                args_p[idx] = args_p[idx] + e; args_m[idx] = args_m[idx] - e
                fd[:, j] = (oracle.rnea(*args_p) - oracle.rnea(*args_m)) / (2 * h)
            assert np.abs(mat - fd).max() < 1e-7
        assert np.abs(M - oracle.crba(q)).max() < 1e-13
        assert np.abs(M - M.T).max() == 0.0


def test_node_constraints_layout(oracle):
    """robot_ocp.hpp:129-160: rows 0-6 = [dtau/dq | dtau/dv | M | dv*v + triu(M)*a], row 7 = [Jz | 0]."""
    rng = np.random.default_rng(2)
    q = rng.uniform(-1, 1, 7); v = rng.uniform(-1, 1, 7); a = rng.uniform(-3, 3, 7)
    g, Jg = oracle.node_constraints(q, v, a)
    tau, dq, dv, M = oracle.rnea_derivatives(q, v, a)
    z, Jz = oracle.tool_height(q)
    assert np.allclose(g[:7], tau, atol=1e-13) and abs(g[7] - z) < 1e-15
    assert np.array_equal(Jg[:7, :7], dq) and np.array_equal(Jg[:7, 7:14], dv) and np.array_equal(Jg[:7, 14:21], M)
    assert np.allclose(Jg[:7, 21], dv @ v + np.triu(M) @ a, atol=1e-12)
    assert np.array_equal(Jg[7, :7], Jz) and not Jg[7, 7:].any()
    g2, _ = oracle.node_constraints(q, v, a, jac=False)
    assert np.allclose(g, g2, atol=1e-12)


def test_chebyshev_nodes_and_D(oracle):
    s = oracle.default_settings()
    t = oracle.time_nodes(s)
    assert np.allclose(t * 24, [0, 1, 3, 4, 5, 7, 8, 9, 11, 12, 13, 15, 16, 17, 19, 20, 21, 23, 24])
    D = oracle.diff_matrix(s)
    ref = np.array([[19 / 6, -4, 4 / 3, -1 / 2], [1, -1 / 3, -1, 1 / 3], [-1 / 3, 1, 1 / 3, -1], [1 / 2, -4 / 3, 4, -19 / 6]])
    assert np.allclose(D, ref[::-1, ::-1], atol=1e-15)          # SURVEY A.2 table (descending) reversed
    # general order: D differentiates polynomials of degree <= PO exactly
    for po in (2, 4, 5, 7):
        s2 = oracle.default_settings(poly_order=po, num_seg=2)
        D2 = oracle.diff_matrix(s2)
        xi = -np.cos(np.pi * np.arange(po + 1) / po)
        for deg in range(po + 1):
            assert np.allclose(D2 @ xi ** deg, deg * xi ** max(deg - 1, 0) if deg else 0 * xi, atol=1e-10)


def test_oracle_embedded_end_to_end_band(oracle):
    """The reference's only embedded end-to-end result (analysis/data_analysis.ipynb cell 1, '1SQP_700QP'): one SQP iteration of the
    oracle from the embedded Ruckig samples lands within |dT| <= 0.01 of T = 1.55469 and within 0.05 rad of the recovered node
    positions -- the band tools/calibrate_admm.py shows the data support (parity with polympc itself stays unpinned)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import calibrate_admm as Cal
    emb, t, x0, xT, gx, gu, T = Cal.embedded_problem(oracle)
    o = oracle.solve(oracle.default_settings(sqp_max_iter=1), oracle.default_limits(Cal.OFFLINE_MARGINS), x0, xT, gx, gu, T)
    assert abs(o[2] - Cal.T_REF) <= 0.01
    ref = Cal.reference_nodes(emb, t)
    assert np.abs(o[0].reshape(19, 14)[:, :7] - ref[:, :7]).max() <= 0.05
    assert o[4].qp_iters[0] == 700 and o[4].qp_status[0] == 2          # the QP runs into its cap
