"""GPU parity tests: the CUDA path, called through the C ABI (libmpcb.so), against the CPU oracle on the same seeded
inputs.  Tolerances are north_star's: identical status classification, final time within 1e-6 relative, trajectories
within 1e-5 absolute per collocation node, identical violation accounting."""
import numpy as np
import pytest

import mpc_motion_planner_b200 as M

pytestmark = pytest.mark.gpu

BENCH_MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]      # examples/benchmark.cpp:9
OFFLINE_MARGINS = [0.9, 0.9, 0.5, 0.9, 0.1]    # examples/offline_trajectory.cpp:9


def _problems(kind, seed, n, margins):
    lim = M.default_limits(margins)
    s = M.default_settings()
    x0, xT = M.generate_problems(lim, kind, seed, n)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    return s, lim, x0, xT, gx, gu, gT


def _stats_tuple(st):
    return (int(st["status"]), int(st["sqp_iters"]), list(st["qp_iters"]), list(st["qp_status"]), list(st["ls_trials"]), list(st["alpha"]))


def _ostats_tuple(st):
    return (st.status, st.sqp_iters, list(st.qp_iters), list(st.qp_status), list(st.ls_trials), list(st.alpha))


@pytest.fixture(scope="module")
def solver():
    sv = M.Solver(M.default_settings(), M.default_limits(BENCH_MARGINS))
    yield sv
    sv.close()


@pytest.mark.parametrize("kind,seed,margins,n", [(0, 1, BENCH_MARGINS, 24), (1, 0, OFFLINE_MARGINS, 12)])
def test_solve_matches_oracle(oracle, solver, kind, seed, margins, n):
    s, lim, x0, xT, gx, gu, gT = _problems(kind, seed, n, margins)
    solver.set_limits(lim)
    sx, su, sT, slam, st = solver.solve(x0, xT, gx, gu, gT, want_lam=True)
    so = oracle.default_settings(); lo = oracle.default_limits(margins)
    osx, osu, oT, ost = oracle.solve_batch(so, lo, x0, xT, gx, gu, gT, nthreads=8)
    assert np.all(np.abs(sT - oT) <= 1e-6 * np.abs(oT)), np.abs(sT - oT).max()
    assert np.abs(sx - osx).max() < 1e-5 and np.abs(su - osu).max() < 1e-5
    for i in range(n):
        assert _stats_tuple(st[i]) == _ostats_tuple(ost[i]), i
        assert abs(st[i]["max_violation"] - ost[i].max_violation) < 1e-7
        assert abs(st[i]["l1_violation"] - ost[i].l1_violation) < 1e-7
    # duals of one problem
    o1 = oracle.solve(so, lo, x0[0], xT[0], gx[0], gu[0], gT[0], want_lam=True)
    assert np.abs(o1[3] - slam[0]).max() < 1e-5 * max(1.0, np.abs(o1[3]).max())


def test_full_benchmark_batch_matches_oracle(oracle, solver):
    """BASELINE config 2 at its full size: all 1000 problems of the mpc_benchmark batch (seed 1) against the oracle --
    identical status / iteration counts / line-search record for every problem, T within 1e-6 relative, nodes within 1e-5."""
    import os
    n = 1000
    s, lim, x0, xT, gx, gu, gT = _problems(0, 1, n, BENCH_MARGINS)
    solver.set_limits(lim)
    sx, su, sT, _, st = solver.solve(x0, xT, gx, gu, gT)
    so = oracle.default_settings(); lo = oracle.default_limits(BENCH_MARGINS)
    osx, osu, oT, ost = oracle.solve_batch(so, lo, x0, xT, gx, gu, gT, nthreads=os.cpu_count() or 8)
    assert np.all(np.abs(sT - oT) <= 1e-6 * np.abs(oT)), np.abs(sT - oT).max()
    assert np.abs(sx - osx).max() < 1e-5 and np.abs(su - osu).max() < 1e-5, (np.abs(sx - osx).max(), np.abs(su - osu).max())
    bad = [i for i in range(n) if _stats_tuple(st[i]) != _ostats_tuple(ost[i])]
    assert not bad, bad[:10]


def test_embedded_reference_problem(oracle):
    """The one start/target pair the reference's notebook embeds (analysis/data_analysis.ipynb cell 1, title
    1SQP_700QP_10accel_90_pos), with its Ruckig samples as the guess and ONE SQP iteration as the title says.  GPU == oracle to the
    parity contract; against the reference's own embedded result the band is what tools/calibrate_admm.py shows the data support
    (DESIGN.md section 1): |T - 1.55469| <= 0.01 and node positions within 0.05 rad (parity with polympc itself is unpinned)."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import calibrate_admm as Cal
    emb, t, x0, xT, gx, gu, T = Cal.embedded_problem(oracle)
    sv = M.Solver(M.default_settings(sqp_max_iter=1), M.default_limits(OFFLINE_MARGINS))
    try:
        sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, T)
    finally:
        sv.close()
    o = oracle.solve(oracle.default_settings(sqp_max_iter=1), oracle.default_limits(OFFLINE_MARGINS), x0, xT, gx, gu, T)
    assert abs(sT[0] - o[2]) <= 1e-6 * o[2] and np.abs(sx[0] - o[0]).max() < 1e-5 and np.abs(su[0] - o[1]).max() < 1e-5
    assert _stats_tuple(st[0]) == _ostats_tuple(o[4])
    assert abs(sT[0] - Cal.T_REF) <= 0.01
    ref = Cal.reference_nodes(emb, t)
    assert np.abs(sx[0].reshape(19, 14)[:, :7] - ref[:, :7]).max() <= 0.05


@pytest.mark.parametrize("kind,seed,margins", [(0, 1, BENCH_MARGINS), (1, 0, OFFLINE_MARGINS)])
def test_device_warm_start_matches_host(solver, kind, seed, margins):
    """otg_guess_kernel (8 lanes per problem, one joint per lane) against the host generator it shares its per-joint
    routines with: same duration and node samples up to multiply-add contraction, for accepted inputs and for the
    cubic-Hermite fallback of rejected ones."""
    n = 300
    s, lim, x0, xT, gx, gu, gT = _problems(kind, seed, n, margins)
    x0 = x0.copy(); xT = xT.copy()
    xT[::17, 7] = 1.01 * margins[1] * lim.vmax[0]          # a target velocity outside the margin: the generator rejects it
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    solver.set_limits(lim)
    dx, du, dT = solver.guess(x0, xT)
    assert np.abs(dT - gT).max() <= 1e-12 * np.abs(gT).max()
    assert np.abs(dx - gx).max() < 1e-10 and np.abs(du - gu).max() < 1e-8      # a(t) has slope j ~ 1e3: 1e-12 in t -> 1e-9 in a
    # the device guess drives the solver to the same result (tolerances of the parity contract)
    a = solver.solve(x0[:8], xT[:8], gx[:8], gu[:8], gT[:8]); b = solver.solve(x0[:8], xT[:8], dx[:8], du[:8], dT[:8])
    assert np.abs(a[2] - b[2]).max() <= 1e-6 * np.abs(a[2]).max() and np.abs(a[0] - b[0]).max() < 1e-5


@pytest.mark.parametrize("kind,seed,margins", [(0, 1, BENCH_MARGINS), (1, 0, OFFLINE_MARGINS)])
def test_device_generator_matches_host(solver, kind, seed, margins):
    """gen_problems_kernel (one problem per thread) against mpcb_generate_problems: the same counter-based streams and the same
    routines, so the same problems up to device sin/cos rounding; a shard [first, first + n) equals the slice of the full batch."""
    n = 2000
    lim = M.default_limits(margins)
    x0, xT = M.generate_problems(lim, kind, seed, n)
    solver.set_limits(lim)
    d0, dT = solver.generate(kind, seed, n)
    # positions: affine in the uniform draws; velocities of kind 0 pass through (J J^T + 1e-5 I)^-1, which amplifies the
    # sin/cos rounding differences near singular poses (2.9e-12 observed)
    assert np.abs(d0 - x0).max() < 1e-13 and np.abs(dT[:, :7] - xT[:, :7]).max() < 1e-13 and np.abs(dT[:, 7:] - xT[:, 7:]).max() < 1e-9
    p0, pT = solver.generate(kind, seed, 300, first=1500)
    assert np.array_equal(p0, d0[1500:1800]) and np.array_equal(pT, dT[1500:1800])


def test_device_resident_pipeline(solver):
    """generate -> warm start -> solve with device pointers only (torch tensors; a non-default stream) gives what the
    host-pointer calls give for the same problems."""
    import torch
    n = 40
    lim = M.default_limits(BENCH_MARGINS)
    solver.set_limits(lim)
    dev = torch.device("cuda:0")
    f64 = dict(dtype=torch.float64, device=dev)
    N = solver.N
    dx0 = torch.empty(n, 14, **f64); dxT = torch.empty(n, 14, **f64)
    dgx = torch.empty(n, 14 * N, **f64); dgu = torch.empty(n, 7 * N, **f64); dgT = torch.empty(n, **f64)
    dsx = torch.empty_like(dgx); dsu = torch.empty_like(dgu); dsT = torch.empty_like(dgT)
    dst = torch.zeros(n * M.capi.STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        sp = stream.cuda_stream
        solver.generate_device(0, 5, n, dx0.data_ptr(), dxT.data_ptr(), first=100, stream=sp)
        solver.guess_device(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), stream=sp)
        solver.solve_device(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(),
                            dsT.data_ptr(), 0, dst.data_ptr(), stream=sp)
    stream.synchronize()
    x0 = dx0.cpu().numpy(); xT = dxT.cpu().numpy()
    h0, hT = M.generate_problems(lim, 0, 5, 140)
    assert np.abs(x0 - h0[100:]).max() < 1e-13 and np.abs(xT - hT[100:]).max() < 1e-9
    # the same device-made problems and guesses through the host-pointer entry point: bitwise the same solution
    sx, su, sT, _, st = solver.solve(x0, xT, dgx.cpu().numpy(), dgu.cpu().numpy(), dgT.cpu().numpy())
    assert np.array_equal(sx, dsx.cpu().numpy()) and np.array_equal(su, dsu.cpu().numpy()) and np.array_equal(sT, dsT.cpu().numpy())
    dstats = np.frombuffer(dst.cpu().numpy().tobytes(), dtype=M.capi.STATS_DTYPE)
    assert np.array_equal(dstats["status"], st["status"]) and np.array_equal(dstats["qp_iters_total"], st["qp_iters_total"])


def test_device_warm_start_stationary_joints(solver):
    """Joints that do not move get a 'stay' profile on the device too (otg_guess_kernel), not the cubic-Hermite fallback."""
    s, lim, x0, xT, gx, gu, gT = _problems(1, 11, 6, OFFLINE_MARGINS)
    x0 = x0.copy(); xT = xT.copy()
    x0[:, 7:] = 0.0; xT[:, 7:] = 0.0
    for i, held in enumerate(([3], [1, 2, 3, 4, 5, 6], [0, 6], [], [2, 5], [0, 1, 2, 3, 4, 5])):
        xT[i, held] = x0[i, held]
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    solver.set_limits(lim)
    dx, du, dT = solver.guess(x0, xT)
    assert np.abs(dT - gT).max() <= 1e-12 * np.abs(gT).max()
    assert np.abs(dx - gx).max() < 1e-10 and np.abs(du - gu).max() < 1e-8
    assert np.array_equal(dx.reshape(6, 19, 14)[1, :, 1:7], np.broadcast_to(x0[1, 1:7], (19, 6)))
    for i in range(6):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        assert rc >= 0 and abs(tr.T - dT[i]) <= 1e-12 * tr.T          # a jerk-limited profile, not the fallback


def test_receding_horizon_resolve(oracle, solver):
    """BASELINE config 5: re-solve 1 ms later, warm-started from the previous solution INCLUDING its multipliers
    (motionPlanner.cpp:199-207 re-seed + polympc's dual carry-over) -- the guess_lam / sol_lam arguments."""
    s, lim, x0, xT, gx, gu, gT = _problems(1, 3, 4, OFFLINE_MARGINS)
    solver.set_limits(lim)
    sx, su, sT, slam, st = solver.solve(x0, xT, gx, gu, gT, want_lam=True)
    so = oracle.default_settings(); lo = oracle.default_limits(OFFLINE_MARGINS)
    N = 19
    x0n = x0.copy(); gx2 = sx.copy(); gu2 = su.copy()
    for i in range(4):
        q, v, a = oracle.interp_state(so, sx[i], su[i], 1e-3 / sT[i])          # the state the robot reaches after 1 ms
        x0n[i] = np.concatenate([q, v])
        gx2[i, :14] = x0n[i]; gx2[i, 14 * (N - 1):] = xT[i]                     # motionPlanner.cpp:203-206
    rx, ru, rT, rlam, rst = solver.solve(x0n, xT, gx2, gu2, sT, glam=slam, want_lam=True)
    for i in range(4):
        o = oracle.solve(so, lo, x0n[i], xT[i], gx2[i], gu2[i], sT[i], glam=slam[i], want_lam=True)
        assert abs(rT[i] - o[2]) <= 1e-6 * abs(o[2]), (i, rT[i], o[2])
        assert np.abs(rx[i] - o[0]).max() < 1e-5 and np.abs(ru[i] - o[1]).max() < 1e-5
        assert _stats_tuple(rst[i]) == _ostats_tuple(o[4]), i
        assert np.abs(o[3] - rlam[i]).max() < 1e-5 * max(1.0, np.abs(o[3]).max())
    # the warm start matters: the re-solve must not be the cold solve of the same problem
    cold = solver.solve(x0n, xT, gx2, gu2, sT)
    assert not np.array_equal(cold[0], rx)


def test_batch_order_and_sharding_invariance(solver):
    """Per-problem outputs do not depend on the batch they are solved in (bitwise): what multi-GPU sharding relies on."""
    s, lim, x0, xT, gx, gu, gT = _problems(0, 2, 40, BENCH_MARGINS)
    solver.set_limits(lim)
    full = solver.solve(x0, xT, gx, gu, gT)
    lo = solver.solve(x0[:13], xT[:13], gx[:13], gu[:13], gT[:13])
    hi = solver.solve(x0[13:], xT[13:], gx[13:], gu[13:], gT[13:])
    assert np.array_equal(full[0], np.concatenate([lo[0], hi[0]])) and np.array_equal(full[2], np.concatenate([lo[2], hi[2]]))
    one = solver.solve(x0[5], xT[5], gx[5], gu[5], gT[5])
    assert np.array_equal(one[0][0], full[0][5])


def test_chunked_batch_matches_small_batches(solver):
    """BASELINE config 3 (100k-trajectory batches) runs in chunks of 4096 OCPs inside mpcb_solve_batch: a batch that spans
    three chunks gives, problem by problem, bitwise what small batches give (first, chunk-boundary and last problems)."""
    n = 9000
    s, lim, x0, xT, gx, gu, gT = _problems(0, 2, n, BENCH_MARGINS)
    solver.set_limits(lim)
    sx, su, sT, _, st = solver.solve(x0, xT, gx, gu, gT)
    assert np.isfinite(sT).all() and (st["sqp_iters"] >= 1).all()
    for lo in (0, 4093, 8190, n - 5):
        sl = slice(lo, lo + 5)
        rx, ru, rT, _, rst = solver.solve(x0[sl], xT[sl], gx[sl], gu[sl], gT[sl])
        assert np.array_equal(rx, sx[sl]) and np.array_equal(ru, su[sl]) and np.array_equal(rT, sT[sl]), lo
        assert np.array_equal(rst["qp_iters"], st["qp_iters"][sl])


@pytest.mark.parametrize("parts,shape", [(4, (3, 6)), (2, (3, 6)), (3, (7, 4))])
def test_internal_stream_split_is_bitwise_identical(solver, parts, shape):
    """MPCB_PARTS (opt-in): mpcb_solve_batch splits a chunk of at least two waves over internal streams of the handle (own work counters,
    workspace slices and matrix blocks per part, fork / join with events).  Same results, bit for bit, as the single-stream path, on a
    batch that spans two chunks (so that parts and chunks both occur), for the base shape and for a generic one; the per-kind kernel
    times stay defined (union of the launch spans)."""
    import os
    n = 4096 + 700 if shape == (3, 6) else 640
    s = M.default_settings(poly_order=shape[0], num_seg=shape[1])
    lim = M.default_limits(BENCH_MARGINS)
    x0, xT = M.generate_problems(lim, 0, 3, n)
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    ref = M.Solver(s, lim) if shape != (3, 6) else solver
    ref.set_limits(lim)
    rx, ru, rT, _, rst = ref.solve(x0, xT, gx, gu, gT)
    launches_ref = ref.last_kernel_launches(0) + ref.last_kernel_launches(1) + ref.last_kernel_launches(2)
    old = os.environ.get("MPCB_PARTS")
    os.environ["MPCB_PARTS"] = str(parts)
    try:
        sv = M.Solver(s, lim)
    finally:
        if old is None:
            del os.environ["MPCB_PARTS"]
        else:
            os.environ["MPCB_PARTS"] = old
    sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
    assert np.array_equal(sx, rx) and np.array_equal(su, ru) and np.array_equal(sT, rT)
    for k in ("status", "sqp_iters", "qp_iters_total", "qp_iters", "ls_trials"):
        assert np.array_equal(st[k], rst[k]), k
    launches = sv.last_kernel_launches(0) + sv.last_kernel_launches(1) + sv.last_kernel_launches(2)
    assert launches > launches_ref                       # the split really happened
    assert 0 < sv.last_kernel_ms(1) <= sv.last_solve_ms() + 1e-3
    sv.close()


def test_large_batch_properties(solver):
    """mpc_benchmark size (1000): every problem terminates with a finite solution that honours the pinned initial state,
    the final-state box up to the solver's own reported violation, and the p bounds."""
    s, lim, x0, xT, gx, gu, gT = _problems(0, 1, 1000, BENCH_MARGINS)
    solver.set_limits(lim)
    sx, su, sT, _, st = solver.solve(x0, xT, gx, gu, gT)
    assert np.isfinite(sx).all() and np.isfinite(su).all() and np.isfinite(sT).all()
    assert set(np.unique(st["status"])) <= {1, 2}
    assert (st["sqp_iters"] >= 1).all() and (st["qp_iters_total"] <= 1400).all()
    sx = sx.reshape(1000, 19, 14)
    viol = st["max_violation"]
    assert (np.abs(sx[:, 0] - x0).max(axis=1) <= viol + 1e-12).all()
    assert (np.abs(sx[:, -1] - xT).max(axis=1) <= 0.01 + viol + 1e-12).all()
    assert (sT >= -viol - 1e-12).all() and (sT <= 10 + viol + 1e-12).all()
    assert solver.last_solve_ms() > 0


def test_sample_and_accounting_match_oracle(oracle, solver):
    s, lim, x0, xT, gx, gu, gT = _problems(0, 4, 6, BENCH_MARGINS)
    solver.set_limits(lim)
    sx, su, sT, _, _ = solver.solve(x0, xT, gx, gu, gT)
    traj, acc = solver.sample(sx, su, sT, target=xT, npts=200)
    so = oracle.default_settings(); lo = oracle.default_limits(BENCH_MARGINS)
    for i in range(6):
        ot = oracle.sample_solution(so, sx[i], su[i], sT[i], 200)
        assert np.abs(traj[i] - ot).max() < 1e-9
        oa = oracle.account(lo, ot, xT[i])
        assert list(acc[i]["flags"]) == list(oa.flags)
        assert np.abs(acc[i]["mins"] - np.array(oa.mins)).max() < 1e-9 and np.abs(acc[i]["maxs"] - np.array(oa.maxs)).max() < 1e-9
        assert np.abs(acc[i]["final_err"] - np.array(oa.final_err)).max() < 1e-9


def test_bad_arguments_are_rejected(solver):
    import ctypes as C
    L = M.load_library()
    assert L.mpcb_solve_batch(solver.h, C.c_int(-1), None, None, None, None, None, None, None, None, None, None, None, 0, None) == -2
    assert L.mpcb_solve_batch(solver.h, C.c_int(1), None, None, None, None, None, None, None, None, None, None, None, 0, None) == -2
    assert L.mpcb_solve_batch(solver.h, C.c_int(0), None, None, None, None, None, None, None, None, None, None, None, 0, None) == 0


def test_profile_sampling_and_accounting_on_device(oracle, solver):
    """The warm-start ("Ruckig") half of the benchmark on the device (mpcb_sample_profile_batch, examples/benchmark.cpp:72-114): samples
    equal the host profile generator's, torques the oracle's RNEA, flags / extrema / final error the oracle's accounting."""
    s, lim, x0, xT, gx, gu, gT = _problems(0, 6, 40, BENCH_MARGINS)
    x0 = x0.copy(); xT = xT.copy()
    xT[7, 7] = 1.01 * BENCH_MARGINS[1] * lim.vmax[0]          # one rejected input: the cubic-Hermite fallback
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    solver.set_limits(lim)
    traj, acc, T = solver.sample_profile(x0, xT, npts=200)
    assert np.abs(T - gT).max() <= 1e-12 * np.abs(gT).max()
    lo = oracle.default_limits(BENCH_MARGINS)
    for i in (0, 3, 7, 21, 39):
        rc, tr = M.otg_calculate(lim, x0[i], xT[i])
        if i != 7:
            assert rc >= 0
            for p in (0, 1, 57, 133, 200):
                q, v, a = M.otg_at_time(tr, tr.T / 200 * p)
                assert np.abs(traj[i, p, 1:8] - q).max() < 1e-10 and np.abs(traj[i, p, 8:15] - v).max() < 1e-10 and np.abs(traj[i, p, 15:22] - a).max() < 1e-7
        else:
            assert rc < 0
        for p in (0, 99, 200):
            assert np.abs(traj[i, p, 22:29] - oracle.rnea(traj[i, p, 1:8], traj[i, p, 8:15], traj[i, p, 15:22])).max() < 1e-9
        oa = oracle.account(lo, traj[i], xT[i])
        assert list(acc[i]["flags"]) == list(oa.flags)
        assert np.abs(acc[i]["mins"] - np.array(oa.mins)).max() < 1e-9 and np.abs(acc[i]["maxs"] - np.array(oa.maxs)).max() < 1e-9
        assert np.abs(acc[i]["final_err"] - np.array(oa.final_err)).max() < 1e-9
