"""CPU checks of the C-ABI library: it loads, exports every symbol include/mpcb.h declares, refuses to run without a
GPU (no CPU fallback), and its host-side planner helpers behave like the reference's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import mpc_motion_planner_b200 as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "mpcb.h")).read()
    names = set(re.findall(r"\b(mpcb_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 15
    L = M.load_library()
    for n in sorted(names):
        assert hasattr(L, n), n


def test_struct_layouts_match_oracle(oracle):
    assert C.sizeof(M.Settings) == C.sizeof(oracle.Settings) and C.sizeof(M.Limits) == C.sizeof(oracle.Limits)
    assert C.sizeof(M.Stats) == C.sizeof(oracle.Stats) and C.sizeof(M.Account) == C.sizeof(oracle.Account)
    s, so = M.default_settings(), oracle.default_settings()
    for f, _ in M.Settings._fields_:
        assert getattr(s, f) == getattr(so, f), f
    l, lo = M.default_limits(), oracle.default_limits()
    for f, _ in M.Limits._fields_:
        a, b = getattr(l, f), getattr(lo, f)
        assert (list(a) == list(b)) if hasattr(a, "__len__") else (a == b), f


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(M.MpcbError, match="no CUDA device"):
        M.Solver()


def test_unsupported_settings_fail_loudly():
    L = M.load_library()
    h = C.c_void_p()
    for kw in (dict(poly_order=4), dict(num_seg=5), dict(poly_order=5, num_seg=12), dict(qp_adaptive_rho=1, qp_adaptive_rho_tolerance=1.0),
               dict(line_search_max_iter=16)):
        rc = L.mpcb_create(C.byref(M.default_settings(**kw)), 0, C.byref(h))
        assert rc == -3 and not h.value
        assert L.mpcb_last_error(None)


@pytest.mark.parametrize("po,ns", [(3, 6), (3, 12), (5, 6), (7, 4)])
def test_time_nodes_match_oracle(oracle, po, ns):
    t = M.time_nodes(M.default_settings(poly_order=po, num_seg=ns))
    assert len(t) == po * ns + 1 and np.abs(t - oracle.time_nodes(oracle.default_settings(poly_order=po, num_seg=ns))).max() < 1e-15


def test_generator_is_deterministic_and_in_bounds(oracle):
    lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
    x0, xT = M.generate_problems(lim, 0, 1, 200)
    x0b, xTb = M.generate_problems(lim, 0, 1, 200)
    assert np.array_equal(xT, xTb) and np.array_equal(x0, x0b)
    qmin, qmax, vmax = np.array(lim.qmin), np.array(lim.qmax), np.array(lim.vmax)
    mid = (qmin + qmax) / 2
    assert np.allclose(x0[:, :7], mid) and not x0[:, 7:].any()          # motionPlanner.cpp:5-8
    safety = 0.2 * (qmax - qmin) / 2
    assert (xT[:, :7] >= qmin + safety - 1e-12).all() and (xT[:, :7] <= qmax - safety + 1e-12).all()
    assert (np.abs(xT[:, 7:]) <= 0.8 * vmax + 1e-12).all()               # benchmark.cpp:38-42
    for i in range(0, 200, 20):
        R, p = oracle.fk(xT[i, :7])
        assert p[6][2] >= lim.min_height                                  # joint-7 height rejection, motionPlanner.cpp:111
        J = oracle.frame_jacobian(xT[i, :7], [0, 0, 0.257])
        tv = J @ xT[i, 7:]
        assert np.linalg.norm(tv[:3]) <= 1.7 + 1e-9                       # benchmark.cpp:25-36
    # different problems differ; kind 1 has random starts
    assert np.abs(xT[0] - xT[1]).max() > 1e-3
    a0, aT = M.generate_problems(M.default_limits([0.9, 0.9, 0.5, 0.9, 0.1]), 1, 0, 10)
    assert np.abs(a0[:, 7:]).max() > 0


def test_guess_respects_margins_and_endpoints():
    m = [0.8, 0.8, 0.6, 0.9, 0.1]
    lim = M.default_limits(m); s = M.default_settings()
    x0, xT = M.generate_problems(lim, 0, 3, 50)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    N = 19
    gx = gx.reshape(50, N, 14); gu = gu.reshape(50, N, 7)
    assert np.allclose(gx[:, 0], x0, atol=1e-12) and np.allclose(gx[:, -1], xT, atol=1e-12)
    assert (gT > 0).all() and (gT <= lim.p_ub).all()
    ok = gT < lim.p_ub
    assert ok.sum() >= 45
    assert (np.abs(gx[ok][:, :, 7:]) <= 0.8 * np.array(lim.vmax) + 1e-9).all()
    assert (np.abs(gu[ok]) <= 0.6 * np.array(lim.amax) + 1e-9).all()


def test_tool_pose_and_inverse_kinematic():
    """PandaWrapper::inverse_kinematic (robot_utils/pandaWrapper.cpp:14-58) and the FK frame placement it iterates on: the
    pose of a random configuration is reached again (tolerance 1e-4 on log6, i.e. ~1e-4 on rotation entries and position)
    whenever the reference's 1000-iteration budget suffices."""
    import ctypes as C
    import numpy as np
    import mpc_motion_planner_b200 as M
    L = M.load_library(); lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])

    def pose(q):
        R = np.zeros(9); p = np.zeros(3)
        assert L.mpcb_tool_pose(q.ctypes.data_as(C.c_void_p), R.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p)) == 0
        return R, p

    rng = np.random.default_rng(0); converged = 0
    for t in range(10):
        qs = np.array(lim.qmin) + rng.random(7) * (np.array(lim.qmax) - np.array(lim.qmin))
        R, p = pose(qs)
        assert abs(np.linalg.det(R.reshape(3, 3)) - 1) < 1e-12
        q = np.zeros(7); it = C.c_int()
        rc = L.mpcb_inverse_kinematic(C.byref(lim), R.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p), C.c_ulonglong(t),
                                      q.ctypes.data_as(C.c_void_p), C.byref(it))
        assert rc in (0, 1) and 0 <= it.value <= 1000
        if rc == 0:
            R2, p2 = pose(q)
            assert np.abs(R2 - R).max() < 2e-4 and np.abs(p2 - p).max() < 2e-4
            converged += 1
    assert converged >= 6


def test_no_miscompiled_shared_base():
    """ptxas 12.9 once rematerialised the dynamic shared-memory base of the 1024-thread stage kernels as `ULEA URa, URb, URb, 0x18`
    (both operands the same register: CTA-rank bits taken from the base itself), which faults on the device with 'illegal memory
    access' (found in round 2, csrc/mpcb_kernels.cu sqp_stage_kernel).  The shipped SASS must not contain that pattern."""
    import re
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", M.capi.lib_path()], capture_output=True, text=True, check=True).stdout
    assert "sqp_stage_kernel" in sass and "admm_kernel" in sass
    assert not re.search(r"ULEA UR(\d+), UR(\d+), UR\2, 0x18", sass)
