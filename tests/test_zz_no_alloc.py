"""mpcb_solve_batch_noalloc (include/mpcb.h): the solve over the library's file-scope static __device__ scratch, for callers that cannot
call cudaMalloc.  No counterpart in the reference, so the check is against the existing
entry point, mpcb_solve_batch with device pointers, on the same inputs, through the C ABI.  Collected last on purpose
(file name).

The kernels are the same and each OCP is solved by one CTA with a fixed reduction order, so the results are expected to be bitwise
equal; the assertion uses the tolerance of the device-function parity tests (1e-9) on the solution and exact equality on the
integer statistics."""
import numpy as np
import pytest

import mpc_motion_planner_b200 as M

pytestmark = pytest.mark.gpu
BENCH_MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]
TOL = 1e-9


def _device_problem_set(sv, lim, n, seed):
    import torch
    f64 = dict(dtype=torch.float64, device=torch.device("cuda:0"))
    x0, xT = M.generate_problems(lim, 0, seed, n)
    gx, gu, gT = M.guess_batch(sv.settings, lim, x0, xT)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(**f64)
    return t(x0), t(xT), t(gx), t(gu), t(gT)


def _outputs(sv, n):
    import torch
    dev = torch.device("cuda:0")
    f64 = dict(dtype=torch.float64, device=dev)
    N = sv.N
    # NaN-filled: an entry point that did not write its outputs cannot pass
    return (torch.full((n, 14 * N), float("nan"), **f64), torch.full((n, 7 * N), float("nan"), **f64), torch.full((n,), float("nan"), **f64),
            torch.zeros(n * M.capi.STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev))


def _stats(dst):
    return np.frombuffer(dst.cpu().numpy().tobytes(), dtype=M.capi.STATS_DTYPE)


def _run_pair(sv, lim, n, seed):
    import torch
    dx0, dxT, dgx, dgu, dgT = _device_problem_set(sv, lim, n, seed)
    rsx, rsu, rsT, rst = _outputs(sv, n)
    sv.solve_device(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, rsx.data_ptr(), rsu.data_ptr(),
                    rsT.data_ptr(), 0, rst.data_ptr())
    torch.cuda.synchronize()
    nsx, nsu, nsT, nst = _outputs(sv, n)
    sv.solve_device_noalloc(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, nsx.data_ptr(), nsu.data_ptr(),
                            nsT.data_ptr(), 0, nst.data_ptr())
    torch.cuda.synchronize()
    a, b = rsx.cpu().numpy(), nsx.cpu().numpy()
    assert np.isfinite(b).all() and np.isfinite(nsu.cpu().numpy()).all() and np.isfinite(nsT.cpu().numpy()).all()
    assert np.abs(a - b).max() <= TOL
    assert np.abs(rsu.cpu().numpy() - nsu.cpu().numpy()).max() <= TOL
    assert np.abs(rsT.cpu().numpy() - nsT.cpu().numpy()).max() <= TOL
    ra, na = _stats(rst), _stats(nst)
    for k in ("status", "sqp_iters", "qp_iters_total", "qp_status", "qp_iters", "ls_trials"):
        assert np.array_equal(ra[k], na[k]), k
    assert (na["sqp_iters"] >= 1).all() and (nsT.cpu().numpy() > 0).all()
    return nsT.cpu().numpy()


def test_noalloc_matches_solve_batch_twice():
    """Two calls in a row on different problem sets (the second one sees the first one's leftovers in the shared scratch), each
    against the allocating entry point; then the first set again: same answer as the first time."""
    s = M.default_settings()
    lim = M.default_limits(BENCH_MARGINS)
    sv = M.Solver(s, lim, device=0)
    n = M.capi.NOALLOC_MAX_OCPS
    T1 = _run_pair(sv, lim, n, seed=21)
    T2 = _run_pair(sv, lim, 3, seed=22)          # a smaller batch after a full one: stale blocks of 5 OCPs stay in the scratch
    assert not np.array_equal(T1[:3], T2)
    T3 = _run_pair(sv, lim, n, seed=21)
    assert np.array_equal(T1, T3)


def test_noalloc_two_handles_share_the_scratch():
    """A second handle (same device) uses the same static arrays: ordered calls (default stream) on either handle agree with the
    allocating path."""
    lim = M.default_limits(BENCH_MARGINS)
    a = M.Solver(M.default_settings(), lim, device=0)
    b = M.Solver(M.default_settings(), lim, device=0)
    _run_pair(a, lim, 4, seed=31)
    _run_pair(b, lim, 5, seed=32)
    _run_pair(a, lim, 2, seed=33)


def test_noalloc_rejects_sizes_beyond_the_maximum():
    import torch
    s = M.default_settings()
    lim = M.default_limits(BENCH_MARGINS)
    sv = M.Solver(s, lim, device=0)
    n = M.capi.NOALLOC_MAX_OCPS + 1
    dx0, dxT, dgx, dgu, dgT = _device_problem_set(sv, lim, n, 5)
    sx, su, sT, st = _outputs(sv, n)
    before = sv.launch_count()
    with pytest.raises(M.MpcbError, match="exceeds the static scratch"):
        sv.solve_device_noalloc(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, sx.data_ptr(), su.data_ptr(),
                                sT.data_ptr(), 0, st.data_ptr())
    torch.cuda.synchronize()
    assert sv.launch_count() == before and torch.isnan(sT).all()      # nothing was enqueued


def test_noalloc_other_shape():
    """The shape-generic kernels (BASELINE config 4, here (7,4)) run over the same static scratch, with the same limit."""
    import torch
    s = M.default_settings(poly_order=7, num_seg=4)
    lim = M.default_limits(BENCH_MARGINS)
    sv = M.Solver(s, lim, device=0)
    _run_pair(sv, lim, 3, seed=41)
    _run_pair(sv, lim, M.capi.NOALLOC_MAX_OCPS, seed=42)
    n = M.capi.NOALLOC_MAX_OCPS + 1
    dx0, dxT, dgx, dgu, dgT = _device_problem_set(sv, lim, n, 5)
    sx, su, sT, st = _outputs(sv, n)
    with pytest.raises(M.MpcbError, match="exceeds the static scratch"):
        sv.solve_device_noalloc(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, sx.data_ptr(), su.data_ptr(),
                                sT.data_ptr(), 0, st.data_ptr())
    torch.cuda.synchronize()
