"""The C++ MotionPlanner mirror (include/motion_planner.hpp) and the ported example mains (examples/*.cpp)."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build():
    subprocess.check_call(["make", "-s", "-C", ROOT, "examples"])
    return os.path.join(ROOT, "build", "offline_trajectory"), os.path.join(ROOT, "build", "mpc_benchmark")


def test_examples_build_and_fail_loudly_without_a_gpu():
    import torch
    off, _ = _build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    r = subprocess.run([off, "/tmp/should_not_exist.txt"], capture_output=True, text=True)
    assert r.returncode != 0 and "no CUDA device" in (r.stderr + r.stdout)
    assert not os.path.exists("/tmp/should_not_exist.txt")


@pytest.mark.gpu
def test_offline_trajectory_file_format(tmp_path):
    off, _ = _build()
    out = tmp_path / "optimal_solution.txt"
    r = subprocess.run([off, str(out), "5"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    a = np.loadtxt(out)
    assert a.shape == (403, 29)                      # offline_trajectory.cpp:69-105, data_analysis.ipynb cell 4
    rk, mpc = a[1:202], a[202:]
    assert rk[0, 0] == 0 and mpc[0, 0] == 0 and np.all(np.diff(mpc[:, 0]) > 0)
    # final state near the +-eps = 1e-2 box (motionPlanner.hpp:44).  The QP is truncated at 700 ADMM iterations without
    # converging, so the box is met only to the ADMM residual: allow one more box width
    assert np.abs(mpc[-1, 1:15] - a[0, 1:15]).max() <= 2e-2
    assert np.abs(rk[0, 1:15] - mpc[0, 1:15]).max() < 1e-5             # both start at the pinned initial state
    T_rk, T_mpc = rk[-1, 0], mpc[-1, 0]
    assert 0 < T_mpc <= 10 and 0 < T_rk <= 10


@pytest.mark.gpu
def test_mpc_benchmark_rows(tmp_path):
    _, ben = _build()
    out = tmp_path / "benchmark_data.txt"
    r = subprocess.run([ben, "64", str(out), "1"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    a = np.loadtxt(out)
    assert a.shape == (64, 162)                      # benchmark.cpp:164-194
    flags = a[:, 140:148]
    assert set(np.unique(flags)) <= {0.0, 1.0}
    assert np.all(a[:, 56:84] <= a[:, 84:112] + 1e-12)              # MPC minima <= maxima
    # the same problems through the Python binding give the same final-state errors (6 printed digits)
    import mpc_motion_planner_b200 as M
    lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1]); s = M.default_settings()
    x0, xT = M.generate_problems(lim, 0, 1, 64)
    assert np.abs(a[:, 148:162] - xT).max() < 1e-5
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    sv = M.Solver(s, lim, device=0)
    sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
    _, acc = sv.sample(sx, su, sT, target=xT, npts=200, want_traj=False)
    assert np.abs(a[:, 126:140] - acc["final_err"]).max() < 1e-5
    sv.close()


# ---- the reference's OWN mains, compiled unchanged against shim/ (SURVEY hard part H5; `make examples-unchanged`) ----------------------
REF = "/root/reference"
REF_OFF = os.path.join(ROOT, "build", "ref_offline_trajectory")
REF_BEN = os.path.join(ROOT, "build", "ref_mpc_benchmark")


def test_reference_mains_compile_unchanged():
    """/root/reference/examples/{offline_trajectory,benchmark}.cpp compile without edits against shim/motionPlanner.hpp (a minimal
    Eigen / pinocchio surface + an Eigen-typed adapter over include/motion_planner.hpp).  Needs the reference tree (this container)."""
    if not os.path.isdir(os.path.join(REF, "examples")):
        pytest.skip("reference tree not mounted")
    subprocess.check_call(["make", "-s", "-C", ROOT, "examples-unchanged"])
    assert os.path.exists(REF_OFF) and os.path.exists(REF_BEN)
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([REF_OFF], capture_output=True, text=True, cwd=ROOT)
        assert r.returncode != 0 and "no CUDA device" in (r.stderr + r.stdout)


@pytest.mark.gpu
def test_reference_offline_main_matches_ported_main(tmp_path):
    """The unchanged reference main writes analysis/optimal_solution.txt (403 x 29); with the same seed the ported main writes the same
    numbers to the 6 digits both print."""
    if not os.path.exists(REF_OFF):
        pytest.skip("build/ref_offline_trajectory not built (make examples-unchanged needs the reference tree)")
    off, _ = _build()
    (tmp_path / "analysis").mkdir()
    r = subprocess.run([REF_OFF], capture_output=True, text=True, timeout=300, cwd=tmp_path)
    if r.returncode != 0 and "outside bounds" in (r.stderr + r.stdout):
        pytest.skip("the seed-0 states fall outside the margins (the reference main throws there)")
    assert r.returncode == 0, r.stderr
    a = np.loadtxt(tmp_path / "analysis" / "optimal_solution.txt")
    assert a.shape == (403, 29)
    out = tmp_path / "ported.txt"
    r = subprocess.run([off, str(out), "0"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    b = np.loadtxt(out)
    assert np.abs(a - b).max() <= 1e-5 * max(1.0, np.abs(b).max())


@pytest.mark.gpu
def test_reference_benchmark_main_rows(tmp_path):
    """The unchanged reference mpc_benchmark main: 1000 sequential solves through one planner, one 162-column row each."""
    if not os.path.exists(REF_BEN):
        pytest.skip("build/ref_mpc_benchmark not built (make examples-unchanged needs the reference tree)")
    (tmp_path / "analysis").mkdir()
    r = subprocess.run([REF_BEN], capture_output=True, text=True, timeout=600, cwd=tmp_path)
    assert r.returncode == 0, r.stderr[-2000:]
    a = np.loadtxt(tmp_path / "analysis" / "benchmark_data.txt")
    assert a.shape == (1000, 162)
    assert set(np.unique(a[:, 140:148])) <= {0.0, 1.0}
    assert np.all(a[:, 56:84] <= a[:, 84:112] + 1e-12) and np.all(a[:, 0:28] <= a[:, 28:56] + 1e-12)
