#!/usr/bin/env python3
"""Benchmark of the hot path: solved minimum-time Panda OCPs per second (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--shape PO,NS] [--global-batch G]

A "step" is one pass of the hot path (mpcb_solve_batch: the whole SQP of every OCP of the batch) over one batch of synthetic
problems: the `mpc_benchmark` configuration (examples/benchmark.cpp of the reference: margins 0.8/0.8/0.6/0.9/0.1, start at the
mid-range pose at rest, random target; 1000 trajectories) per GPU.  With N > 1 (launched by torchrun, one rank per GPU) every rank
solves its own 1000-problem shard of the global batch (no data-path collective; weak scaling) and the per-rank statistics are
gathered with one NCCL all_gather; in addition BASELINE config 3 (one 100 000-problem batch, seed 2, sharded over the ranks: strong
scaling) is timed and reported under stats.config3.

  value : OCPs/s with the inputs already resident in HBM (device pointers through the C ABI), CUDA events
  e2e   : the same through the host-pointer C ABI call (pinned host buffers, H2D + D2H inside the timed region)
  --impl reference : the CPU restatement of the reference's algorithm (oracle/, the reference itself cannot be built here:
                     polympc/pinocchio/ruckig/Eigen are absent) on all host threads, on a bounded sample.  That arm maps only
                     oracle/liboracle_cpu.so: its inputs are made by a child process.
  --shape PO,NS    : BASELINE config 4, another collocation shape (robot_ocp.hpp:31-32): (3,12), (5,6), (7,4)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MARGINS = [0.8, 0.8, 0.6, 0.9, 0.1]      # examples/benchmark.cpp:9
METRIC = "solved OCPs/sec (mpc_benchmark shape)"
UNIT = "OCP/s"
GUESS = "jerk-limited synchronised profile (csrc/mpcb_otg.h, Ruckig's role; pinned on the reference-embedded Ruckig trajectory)"
# DRAM bytes of one admm_kernel launch per OCP, from the committed ncu --set full capture of the kernel on the base shape
# (tools/gpu_round2.sh: dram__bytes_read.sum 64.78 MB + dram__bytes_write.sum 0.72 MB for a 148-OCP launch): the factorised QP block
# of every OCP is read once per QP.
NCU_ADMM_DRAM_BYTES_PER_OCP = (64.782336e6 + 0.716032e6) / 148
NCU_TRAFFIC_SOURCE = "ncu --set full, profiles/r2zc_admm_kernel_raw.csv (dram read + write of one 148-OCP admm_kernel launch, scaled to the batch)"


def workload_name(po, ns, B):
    return (f"mpc_benchmark: {B} random Panda min-time OCPs per GPU (PO={po}, NS={ns}, 2 SQP x <=700 ADMM, "
            "pos/vel/acc/torque/height limits)")


def make_config(args, world):
    po, ns = args.shape
    return {"workload": workload_name(po, ns, args.batch), "batch_per_gpu": args.batch, "global_batch": args.batch * world, "seed": 1,
            "guess": GUESS, "l2": "256 MiB buffer written between timed steps (L2 flush)", "parallelism": f"batch shards x{world}, stats all_gather"}


def work_model(po, ns):
    """Algorithmic FP64 work of one OCP: structure-exploiting counts of THIS implementation, per shape."""
    nn = po * ns + 1
    nvar, neq, nineq = 21 * nn + 1, 14 * (nn - 1), 8 * nn
    mb = 7 + 21 * (po - 1)
    m = [mb + (14 if s == 0 else 0) + (21 if s == ns - 1 else 0) for s in range(ns)]
    nb = [(14 if s > 0 else 0) + (14 if s < ns - 1 else 0) + 1 for s in range(ns)]
    NB = 14 * (ns - 1) + 1
    w = sum(x * x for x in m); f = sum(x * y for x, y in zip(m, nb)); c = NB * NB
    annz = neq * (po + 3) + nineq * 22
    return {"W_IT": 2 * (w + 2 * f + c + 2 * annz) + 8 * (neq + nineq + nvar),       # W, F^T, C, F mat-vecs + A, A^T + vector ops
            "W_FAC": sum(x * (x + y) ** 2 for x, y in zip(m, nb)) + NB ** 3 + 300 * nvar,   # sweeps (interiors + separator) + assembly
            "W_LIN": 300000.0 / 19 * nn,                                                   # per node: RNEA derivatives + CRBA + frame Jacobian
            "W_MERIT": 35000.0 / 19 * nn,                                                  # per node: RNEA + FK, + residuals
            "operator_bytes": 8 * (w + 2 * f + c)}


def flops_of(stats, wm):
    import numpy as np
    sqp = stats["sqp_iters"].astype(np.float64)
    admm = stats["qp_iters_total"].astype(np.float64)
    ls = stats["ls_trials"].sum(axis=1).astype(np.float64) + 2 * sqp
    return float((sqp * (wm["W_LIN"] + wm["W_FAC"]) + admm * wm["W_IT"] + ls * wm["W_MERIT"]).sum())


class ClockSampler:
    QUERY = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index; self.samples = []; self.stop = False; self.th = None

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.th = threading.Thread(target=self._run, daemon=True); self.th.start(); return self

    def __exit__(self, *a):
        self.stop = True; self.th.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        import statistics
        sm = [float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(s) > 2 + i and s[2 + i].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(self.samples)}


# ---- CPU legs (oracle = test infrastructure; only these two legs may touch it) ------------------------------------------------------
def make_inputs_in_child(n, po, ns):
    """Inputs of the workload (problems + warm start) made by a CHILD process through the product's host helpers and handed over as
    an .npz: the reference arm itself then maps only oracle/liboracle_cpu.so."""
    fd, path = tempfile.mkstemp(suffix=".npz", prefix="mpcb_inputs_")
    os.close(fd)
    code = (f"import sys; sys.path.insert(0, {ROOT!r}); import numpy as np; import mpc_motion_planner_b200 as M; "
            f"s = M.default_settings(poly_order={po}, num_seg={ns}); lim = M.default_limits({MARGINS!r}); "
            f"x0, xT = M.generate_problems(lim, 0, 1, {n}); gx, gu, gT = M.guess_batch(s, lim, x0, xT); "
            f"np.savez({path!r}, x0=x0, xT=xT, gx=gx, gu=gu, gT=gT)")
    subprocess.run([sys.executable, "-c", code], check=True, stdout=subprocess.DEVNULL)
    import numpy as np
    d = np.load(path)
    out = tuple(d[k] for k in ("x0", "xT", "gx", "gu", "gT"))
    os.unlink(path)
    return out


def oracle_leg(inputs, po, ns, threads, n_single=16):
    """Times the oracle (CPU restatement, kind 'port') on `threads` host threads over the given problems, then single-thread on the
    first n_single of them one at a time (OCP/s and p50 / p99 single-solve latency).  Returns the figures and the oracle's solutions."""
    from oracle import oracle_py as O
    x0, xT, gx, gu, gT = inputs
    n = len(gT)
    so = O.default_settings(poly_order=po, num_seg=ns); lo = O.default_limits(MARGINS)
    t0 = time.perf_counter()
    osx, osu, oT, ost = O.solve_batch(so, lo, x0, xT, gx, gu, gT, nthreads=threads)
    dt = time.perf_counter() - t0
    lat = []
    for i in range(min(n_single, n)):
        t1 = time.perf_counter()
        O.solve(so, lo, x0[i], xT[i], gx[i], gu[i], gT[i])
        lat.append((time.perf_counter() - t1) * 1e3)
    lat.sort()
    single = None
    if lat:
        single = {"value": len(lat) / (sum(lat) * 1e-3), "unit": UNIT, "cores": 1, "n": len(lat),
                  "p50_single_solve_latency_ms": lat[len(lat) // 2], "p99_single_solve_latency_ms": lat[min(len(lat) - 1, int(0.99 * len(lat)))]}
    return {"ocps": n / dt, "seconds": dt, "single": single, "sol": (osx, osu, oT, ost), "settings": so, "limits": lo}


def oracle_accounting(leg, xT):
    """benchmark.cpp:118-194 accounting of the oracle's own solutions: counts of failed flags [jerk, linear velocity, angular velocity, collision]."""
    from oracle import oracle_py as O
    osx, osu, oT, _ = leg["sol"]
    fails = [0, 0, 0, 0]
    for i in range(len(oT)):
        tr = O.sample_solution(leg["settings"], osx[i], osu[i], oT[i], 200)
        acc = O.account(leg["limits"], tr, xT[i])
        for k in range(4):
            fails[k] += int(acc.flags[k] == 0)
    return fails


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    po, ns = args.shape
    threads = os.cpu_count() or 1
    n_sample = max(threads * 64, 128)                    # about 2 s of CPU work per step on all host threads
    from oracle import oracle_py as O
    O.build()
    inputs = make_inputs_in_child(n_sample, po, ns)
    times = []
    leg = None
    for i in range(args.warmup + args.steps):
        leg = oracle_leg(inputs, po, ns, threads, n_single=8 if i == args.warmup + args.steps - 1 else 0)
        if i >= args.warmup:
            times.append(leg["seconds"])
    dt = sum(times) / len(times)
    value = n_sample / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(args, args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"first {n_sample} OCPs of the workload per step on {threads} threads (oracle/: CPU restatement; the reference binary cannot be built: polympc, pinocchio, ruckig, Eigen absent)",
                             "single_thread": leg["single"]},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---- the GPU arm -------------------------------------------------------------------------------------------------------------------
def lagrange_state(nodes, po, sx, su, t01):
    """State (q, v) of a solution at normalised time t01 inside the FIRST segment: polympc's piecewise Lagrange interpolation
    (solution_x_at, motionPlanner.hpp:106-108) on the segment's PO + 1 nodes.  Used to advance x0 between receding-horizon re-solves."""
    import numpy as np
    tk = nodes[:po + 1]
    w = np.ones(po + 1)
    for j in range(po + 1):
        for l in range(po + 1):
            if l != j:
                w[j] *= (t01 - tk[l]) / (tk[j] - tk[l])
    return (w[:, None] * sx.reshape(-1, 14)[:po + 1]).sum(axis=0)


def run_ours(args):
    import numpy as np
    import torch
    import mpc_motion_planner_b200 as M
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.batch
    po, ns = args.shape
    base_shape = (po, ns) == (3, 6)
    wm = work_model(po, ns)
    s = M.default_settings(poly_order=po, num_seg=ns); lim = M.default_limits(MARGINS)
    # rank r owns problems [r*B, (r+1)*B) of the global seed-1 batch (generator streams are indexed by problem)
    from mpc_motion_planner_b200.sharding import shard_range
    x0a, xTa = M.generate_problems(lim, 0, 1, B * world)
    lo, hi = shard_range(rank, world, B * world)
    x0 = x0a[lo:hi].copy(); xT = xTa[lo:hi].copy()
    gx, gu, gT = M.guess_batch(s, lim, x0, xT)
    sv = M.Solver(s, lim, device=local)
    N = sv.N
    # ---- device-resident arm -------------------------------------------------------------------------------------
    dx0 = torch.from_numpy(x0).to(dev); dxT = torch.from_numpy(xT).to(dev); dgx = torch.from_numpy(gx).to(dev)
    dgu = torch.from_numpy(gu).to(dev); dgT = torch.from_numpy(gT).to(dev)
    dsx = torch.empty_like(dgx); dsu = torch.empty_like(dgu); dsT = torch.empty_like(dgT)
    dst = torch.zeros(B * M.capi.STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream().cuda_stream

    def step_device():
        sv.solve_device(B, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(),
                        dsT.data_ptr(), 0, dst.data_ptr(), stream)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        flush.fill_(1); step_device()
    barrier()
    launches0 = sv.launch_count()
    evs = []
    admm_ms = []; stage_ms = []; node_ms = []; admm_launches = 0
    with ClockSampler(local) as clk:
        t_wall0 = time.perf_counter()
        for _ in range(args.steps):
            flush.fill_(1)                                   # L2 flush between timed iterations (outside the event pair)
            a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record(); step_device(); b.record()
            evs.append((a, b))
            admm_ms.append(sv.last_kernel_ms(1)); stage_ms.append(sv.last_kernel_ms(0)); node_ms.append(sv.last_kernel_ms(2))   # per-launch CUDA events
            admm_launches += sv.last_kernel_launches(1)
        barrier()
        t_wall = time.perf_counter() - t_wall0
    launches = sv.launch_count() - launches0
    ms_steps = [a.elapsed_time(b) for a, b in evs]
    ms_total = sum(ms_steps)
    stats = np.frombuffer(dst.cpu().numpy().tobytes(), dtype=M.capi.STATS_DTYPE)
    flops_step = flops_of(stats, wm)
    flops_admm_step = float(stats["qp_iters_total"].astype(np.float64).sum()) * wm["W_IT"]
    if not base_shape:     # the fused kernel of the generic path also linearises and factorises
        flops_admm_step += float(stats["sqp_iters"].astype(np.float64).sum()) * wm["W_FAC"]
    sol_T = dsT.cpu().numpy()
    # ---- end-to-end arm: host-pointer C-ABI call with pinned host buffers --------------------------------------------
    def pinned(a):
        t = torch.empty(a.shape, dtype=torch.float64, pin_memory=True); t.numpy()[...] = a; return t
    hx0, hxT, hgx, hgu, hgT = pinned(x0), pinned(xT), pinned(gx), pinned(gu), pinned(gT)
    hsx = torch.empty((B, 14 * N), dtype=torch.float64, pin_memory=True); hsu = torch.empty((B, 7 * N), dtype=torch.float64, pin_memory=True)
    hsT = torch.empty(B, dtype=torch.float64, pin_memory=True)
    hst = torch.empty(B * M.capi.STATS_DTYPE.itemsize, dtype=torch.uint8, pin_memory=True)
    import ctypes as C
    def step_host():
        rc = sv.L.mpcb_solve_batch(sv.h, C.c_int(B), C.c_void_p(hx0.data_ptr()), C.c_void_p(hxT.data_ptr()), C.c_void_p(hgx.data_ptr()),
                                   C.c_void_p(hgu.data_ptr()), C.c_void_p(hgT.data_ptr()), None, C.c_void_p(hsx.data_ptr()), C.c_void_p(hsu.data_ptr()),
                                   C.c_void_p(hsT.data_ptr()), None, C.c_void_p(hst.data_ptr()), C.c_int(0), None)
        assert rc == 0, sv.L.mpcb_last_error(sv.h)
    step_host()
    barrier()
    e2e_steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()                                          # synchronous: returns after the D2H copies
    barrier()
    t_e2e = (time.perf_counter() - t0) / e2e_steps
    assert np.array_equal(hsT.numpy(), sol_T), "host-pointer and device-pointer calls must give identical results"
    h2d = B * (14 + 14 + 14 * N + 7 * N + 1) * 8
    d2h = B * ((14 * N + 7 * N + 1) * 8 + M.capi.STATS_DTYPE.itemsize)
    # ---- A15: read-out + benchmark accounting of the solved batch on the device (sample_kernel, benchmark.cpp:118-194) ------------
    dacc = torch.zeros(B * M.capi.ACCOUNT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    acc_ms = []
    for _ in range(3):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); sv.sample_device(B, dsx.data_ptr(), dsu.data_ptr(), dsT.data_ptr(), dxT.data_ptr(), 200, 0, dacc.data_ptr(), stream); b.record()
        torch.cuda.synchronize(); acc_ms.append(a.elapsed_time(b))
    acc = np.frombuffer(dacc.cpu().numpy().tobytes(), dtype=M.capi.ACCOUNT_DTYPE)
    fails = [int((acc["flags"][:, k] == 0).sum()) for k in range(4)]
    # ---- BASELINE config 3: one global batch (seed 2) sharded over the ranks, generated and warm-started on the device ------------
    config3 = None
    G = args.global_batch if args.global_batch else (100000 if world > 1 else 0)
    if G:
        glo, ghi = shard_range(rank, world, G)
        nb3 = ghi - glo
        c_x0 = torch.empty((nb3, 14), dtype=torch.float64, device=dev); c_xT = torch.empty_like(c_x0)
        c_gx = torch.empty((nb3, 14 * N), dtype=torch.float64, device=dev); c_gu = torch.empty((nb3, 7 * N), dtype=torch.float64, device=dev)
        c_gT = torch.empty(nb3, dtype=torch.float64, device=dev)
        c_sx = torch.empty_like(c_gx); c_su = torch.empty_like(c_gu); c_sT = torch.empty_like(c_gT)
        c_st = torch.zeros(nb3 * M.capi.STATS_DTYPE.itemsize, dtype=torch.uint8, device=dev)
        sv.generate_device(0, 2, nb3, c_x0.data_ptr(), c_xT.data_ptr(), first=glo, stream=stream)
        sv.guess_device(nb3, c_x0.data_ptr(), c_xT.data_ptr(), c_gx.data_ptr(), c_gu.data_ptr(), c_gT.data_ptr(), stream)
        def step3():
            sv.solve_device(nb3, c_x0.data_ptr(), c_xT.data_ptr(), c_gx.data_ptr(), c_gu.data_ptr(), c_gT.data_ptr(), 0, c_sx.data_ptr(), c_su.data_ptr(),
                            c_sT.data_ptr(), 0, c_st.data_ptr(), stream)
        step3()                                              # warm-up
        barrier()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        k3 = 2
        a.record()
        for _ in range(k3):
            step3()
        b.record(); barrier()
        ms3 = a.elapsed_time(b) / k3
        st3 = np.frombuffer(c_st.cpu().numpy().tobytes(), dtype=M.capi.STATS_DTYPE)
        t3 = torch.tensor([ms3, float(st3["qp_iters_total"].sum()), float(nb3)], dtype=torch.float64, device=dev)
        if dist is not None:
            allt = [torch.empty_like(t3) for _ in range(world)]
            dist.all_gather(allt, t3)
            allt = torch.stack(allt).cpu().numpy()
        else:
            allt = t3.cpu().numpy()[None]
        config3 = {"workload": f"one {G}-problem batch of the mpc_benchmark generator (seed 2), contiguous shards over {world} GPU(s); problems and warm start made on the device",
                   "global_batch": G, "scaling": "strong", "value": G / (allt[:, 0].max() * 1e-3), "unit": UNIT, "ms_per_step": float(allt[:, 0].max()),
                   "steps": k3, "warmup": 1, "admm_iters_per_ocp": float(allt[:, 1].sum() / G), "per_rank_ocps": [int(x) for x in allt[:, 2]],
                   "l2": "per-rank shard of the QP workspace is far larger than L2"}
        del c_x0, c_xT, c_gx, c_gu, c_gT, c_sx, c_su, c_sT, c_st
    # ---- max over ranks + statistics gather ---------------------------------------------------------------------------
    from mpc_motion_planner_b200 import sharding
    loc = sharding.local_stats(ms_total, t_e2e, flops_step, stats["status"], stats["qp_iters_total"], stats["sqp_iters"], sol_T)
    allr = sharding.gather_stats(loc, dist, dev)                     # the only collective of the path: per-rank statistics
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    red = sharding.reduce_stats(allr)
    ms_total_max = red["ms_total_max"]; t_e2e_max = red["e2e_s_max"]
    total_ocps = red["n_ocp"]
    ms_per_step = ms_total_max / args.steps
    value = total_ocps / (ms_per_step * 1e-3)
    fp64_peak = sv.fp64_tflops()
    achieved_step = flops_step / (ms_total / args.steps * 1e-3) / 1e12     # rank 0, all kernels of a step
    # dominant kernel: algorithmic FLOPs per launch / average launch duration (CUDA events around each launch)
    admm_ms_launch = sum(admm_ms) / max(admm_launches, 1)
    flops_admm_launch = flops_admm_step * args.steps / max(admm_launches, 1)
    achieved = flops_admm_launch / (admm_ms_launch * 1e-3) / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    algo_bytes = h2d + d2h
    kernel_name = "admm_kernel" if base_shape else "sqp_stage_kernel<GEN> (linearise + factorise + box-ADMM fused, mpcb_gadmm.h)"
    roofline = {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s", "frac": achieved / fp64_peak if fp64_peak > 0 else None,
                "traffic": NCU_ADMM_DRAM_BYTES_PER_OCP * B if base_shape else None,
                "traffic_source": NCU_TRAFFIC_SOURCE if base_shape else "no ncu capture for this shape",
                "peak_source": "FP64 FMA micro-benchmark run in this process (MEASURED_PEAKS.json has no FP64 figure)",
                "kernel": kernel_name, "flops_per_launch": flops_admm_launch, "kernel_ms": admm_ms_launch,
                "kernel_share_of_step": sum(admm_ms) / ms_total, "flops_per_admm_iteration": wm["W_IT"],
                "step": {"flops": flops_step, "ms": ms_total / args.steps, "achieved": achieved_step,
                         "admm_ms": sum(admm_ms) / args.steps, "sqp_stage_ms": sum(stage_ms) / args.steps, "node_ms": sum(node_ms) / args.steps},
                "hbm": {"achieved_gbs": algo_bytes / (ms_total / args.steps * 1e-3) / 1e9, "peak_gbs": peaks.get("hbm_gbs"),
                        "note": "algorithmic HBM bytes per launch = inputs + outputs; the path is not HBM-bound"}}
    if not base_shape:
        # the generic kernel streams its per-OCP operators (full copies of W, F twice, C^-1: far more than L2 holds for 148 CTAs) from
        # HBM every iteration: that stream, not the FP64 pipe, bounds it.  Algorithmic bytes per launch = ADMM iterations x operator bytes.
        it_per_launch = float(stats["qp_iters_total"].astype(np.float64).sum()) * args.steps / max(admm_launches, 1)
        gbs = it_per_launch * wm["operator_bytes"] / (admm_ms_launch * 1e-3) / 1e9
        fp64 = {k: roofline[k] for k in ("achieved", "peak", "unit", "frac", "flops_per_launch", "flops_per_admm_iteration")}
        roofline.update({"bound": "hbm", "achieved": gbs, "peak": peaks.get("hbm_gbs", 6550.0), "unit": "GB/s",
                         "frac": gbs / peaks.get("hbm_gbs", 6550.0), "bytes_per_admm_iteration": wm["operator_bytes"],
                         "bytes_per_launch": it_per_launch * wm["operator_bytes"],
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (sustained copy)" if "hbm_gbs" in peaks else "fallback of B200_PROFILING.md",
                         "fp64": fp64})
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": make_config(args, world),
        "e2e": {"value": total_ocps / t_e2e_max, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        "roofline": roofline,
        "stats": {"solved": red["solved"], "admm_iters_per_ocp": red["qp_iters"] / total_ocps, "sqp_iters_per_ocp": red["sqp_iters"] / total_ocps,
                  "mean_T": red["mean_T"], "min_T": red["min_T"], "max_T": red["max_T"],
                  "wall_s_timed_region": t_wall, "p50_step_ms": float(sorted(ms_steps)[len(ms_steps) // 2]),
                  "accounting": {"what": "benchmark.cpp:118-160 flags over 201 samples of every solved OCP of rank 0's shard (sample_kernel)",
                                 "n": B, "failed": {"jerk": fails[0], "linear_velocity": fails[1], "angular_velocity": fails[2], "collision": fails[3]},
                                 "sample_kernel_ms": float(sorted(acc_ms)[len(acc_ms) // 2])}},
    }
    if config3 is not None:
        line["stats"]["config3"] = config3
    # single-solve latency (n = 1), device-resident
    lat = []
    for _ in range(9):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        sv.solve_device(1, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(), dsT.data_ptr(), 0, dst.data_ptr(), stream)
        b.record(); torch.cuda.synchronize(); lat.append(a.elapsed_time(b))
    lat.sort()
    line["stats"]["p50_single_solve_latency_ms"] = float(lat[len(lat) // 2])
    line["stats"]["p99_single_solve_latency_ms"] = float(lat[-1])
    # the warm start itself (warm_start_RK for the whole batch) on the device: otg_guess_kernel, device-resident
    wg = []
    for _ in range(4):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); sv.guess_device(B, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), stream); b.record()
        torch.cuda.synchronize(); wg.append(a.elapsed_time(b))
    line["stats"]["warm_start_kernel_ms"] = float(sorted(wg)[len(wg) // 2])
    # the problem generator on the device (gen_problems_kernel: this rank's shard of the seed-1 batch), into scratch buffers
    tx0 = torch.empty_like(dx0); txT = torch.empty_like(dxT)
    wg = []
    for _ in range(4):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); sv.generate_device(0, 1, B, tx0.data_ptr(), txT.data_ptr(), first=lo, stream=stream); b.record()
        torch.cuda.synchronize(); wg.append(a.elapsed_time(b))
    line["stats"]["problem_generator_kernel_ms"] = float(sorted(wg)[len(wg) // 2])
    line["stats"]["problem_generator_max_abs_diff_vs_host"] = float(max((tx0.cpu() - torch.from_numpy(x0)).abs().max(), (txT.cpu() - torch.from_numpy(xT)).abs().max()))
    # BASELINE config 5 (receding horizon at 1 kHz, latency bound): a batch of 8 re-solved 40 times; before every re-solve the start
    # state advances to where the previous solution is 1 ms later, and the guess is the previous solution with its multipliers
    # (motionPlanner.cpp:199-207 + the dual carry-over) -- the sequence of tests/test_gpu_parity.py::test_receding_horizon_resolve
    nd = sv.ndual
    nodes = M.time_nodes(s)
    RB = 8
    dlam_a = torch.zeros(RB * nd, dtype=torch.float64, device=dev); dlam_b = torch.zeros_like(dlam_a)
    rx0 = dx0[:RB].clone(); rxT = dxT[:RB].clone()
    rx = dgx[:RB].clone(); ru = dgu[:RB].clone(); rT = dgT[:RB].clone()
    sv.solve_device(RB, rx0.data_ptr(), rxT.data_ptr(), rx.data_ptr(), ru.data_ptr(), rT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(), dsT.data_ptr(),
                    dlam_a.data_ptr(), dst.data_ptr(), stream)
    rh = []
    for _ in range(40):
        hx = dsx[:RB].cpu().numpy(); hu = dsu[:RB].cpu().numpy(); hT = dsT[:RB].cpu().numpy()
        x0n = np.stack([lagrange_state(nodes, po, hx[i], hu[i], 1e-3 / hT[i]) for i in range(RB)])
        rx0.copy_(torch.from_numpy(x0n))
        rx.copy_(dsx[:RB]); ru.copy_(dsu[:RB]); rT.copy_(dsT[:RB])
        rx[:, :14] = rx0; rx[:, 14 * (N - 1):] = rxT                               # motionPlanner.cpp:203-206
        torch.cuda.synchronize()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        sv.solve_device(RB, rx0.data_ptr(), rxT.data_ptr(), rx.data_ptr(), ru.data_ptr(), rT.data_ptr(), dlam_a.data_ptr(), dsx.data_ptr(), dsu.data_ptr(),
                        dsT.data_ptr(), dlam_b.data_ptr(), dst.data_ptr(), stream)
        b.record(); torch.cuda.synchronize(); rh.append(a.elapsed_time(b))
        dlam_a, dlam_b = dlam_b, dlam_a
    rh.sort()
    line["stats"]["receding_horizon"] = {"batch": RB, "resolves": len(rh), "p50_ms": rh[len(rh) // 2], "p99_ms": rh[min(len(rh) - 1, int(0.99 * len(rh)))],
                                         "note": "x0 advanced by 1 ms along the previous solution before every re-solve; warm start from the previous solution and multipliers"}
    # the same batch with the library's opt-in splitting over internal streams (MPCB_PARTS = 4: the persistent kernels of one part fill the
    # idle SMs of another part's last wave); informational, the headline figures above come from the default single-stream path
    if "MPCB_PARTS" not in os.environ:
        os.environ["MPCB_PARTS"] = "4"
        try:
            sv4 = M.Solver(s, lim, device=local)
            t4 = []
            for i in range(5):
                a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
                flush.fill_(1)
                a.record()
                sv4.solve_device(B, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(),
                                 dsT.data_ptr(), 0, dst.data_ptr(), stream)
                b.record(); torch.cuda.synchronize()
                if i >= 2:
                    t4.append(a.elapsed_time(b))
            t4.sort()
            line["stats"]["split_over_4_internal_streams"] = {"ms_per_step": t4[len(t4) // 2], "value": B / (t4[len(t4) // 2] * 1e-3), "unit": UNIT,
                                                              "note": "MPCB_PARTS=4, this rank's batch, device-resident; opt-in (see INTEGRATION.md)"}
            sv4.close()
        except Exception as e:        # informational only
            line["stats"]["split_over_4_internal_streams"] = {"error": str(e)[:200]}
        del os.environ["MPCB_PARTS"]
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n_sample = min(B, max(threads * 320, 640)) if not base_shape else max(threads * 320, 640)      # about 10 s of CPU work on all host threads
        if n_sample <= B:
            inputs = (x0[:n_sample], xT[:n_sample], gx[:n_sample], gu[:n_sample], gT[:n_sample])
        else:
            cx0, cxT = M.generate_problems(lim, 0, 1, n_sample)
            inputs = (cx0, cxT) + M.guess_batch(s, lim, cx0, cxT)
        leg = oracle_leg(inputs, po, ns, threads)
        line["cpu_baseline"] = {"value": leg["ocps"], "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {n_sample} OCPs of the same batch, {leg['seconds']:.1f} s, oracle/ (CPU restatement; reference binary unbuildable here)",
                                "single_thread": leg["single"]}
        # the oracle's own accounting of its solutions of the sample against the GPU's on the same problems
        nc = min(n_sample, B)
        ofails = oracle_accounting({**leg, "sol": tuple(a[:nc] for a in leg["sol"][:3]) + (None,)}, inputs[1][:nc])
        gfails = [int((acc["flags"][:nc, k] == 0).sum()) for k in range(4)]
        line["stats"]["accounting"]["vs_oracle"] = {"n": nc, "gpu_failed": gfails, "oracle_failed": ofails, "identical": gfails == ofails}
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The one JSON line of the contract goes to the process's original stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # stdout carries exactly ONE line (the JSON): everything else that writes to fd 1 -- NCCL's "NCCL version ..." banner,
    # library chatter -- is sent to stderr
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1000, help="OCPs per GPU per step")
    ap.add_argument("--shape", default="3,6", help="collocation shape POLY_ORDER,NUM_SEG (BASELINE config 4): 3,6 | 3,12 | 5,6 | 7,4")
    ap.add_argument("--global-batch", type=int, default=0, help="BASELINE config 3: also time one batch of this many OCPs sharded over the ranks (default 100000 when N > 1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.shape = tuple(int(x) for x in args.shape.split(","))
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
