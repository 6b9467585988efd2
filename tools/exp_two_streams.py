#!/usr/bin/env python3
"""Experiment: does solving two half batches on two streams (two handles) beat one full batch on one stream?  (tail of the persistent
kernels: 1000 OCPs = 6.76 waves of 148 CTAs.)  python tools/exp_two_streams.py [n] [parts]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import mpc_motion_planner_b200 as M

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
parts = int(sys.argv[2]) if len(sys.argv) > 2 else 2
s = M.default_settings(); lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
x0, xT = M.generate_problems(lim, 0, 1, n)
gx, gu, gT = M.guess_batch(s, lim, x0, xT)
dev = torch.device("cuda:0"); f64 = dict(dtype=torch.float64, device=dev)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(**f64)
dx0, dxT, dgx, dgu, dgT = t(x0), t(xT), t(gx), t(gu), t(gT)
sv = [M.Solver(s, lim, device=0) for _ in range(parts)]
N = sv[0].N
dsx = torch.empty(n, 14 * N, **f64); dsu = torch.empty(n, 7 * N, **f64); dsT = torch.empty(n, **f64)
streams = [torch.cuda.Stream(device=dev) for _ in range(parts)]
bounds = [n * i // parts for i in range(parts + 1)]

def run_split():
    for p in range(parts):
        a, b = bounds[p], bounds[p + 1]
        sv[p].solve_device(b - a, dx0[a:].data_ptr(), dxT[a:].data_ptr(), dgx[a:].data_ptr(), dgu[a:].data_ptr(), dgT[a:].data_ptr(), 0,
                           dsx[a:].data_ptr(), dsu[a:].data_ptr(), dsT[a:].data_ptr(), 0, 0, stream=streams[p].cuda_stream)

def run_one():
    sv[0].solve_device(n, dx0.data_ptr(), dxT.data_ptr(), dgx.data_ptr(), dgu.data_ptr(), dgT.data_ptr(), 0, dsx.data_ptr(), dsu.data_ptr(),
                       dsT.data_ptr(), 0, 0, stream=streams[0].cuda_stream)

def timeit(fn, reps=5):
    for _ in range(2):
        fn(); torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        e0.record(torch.cuda.default_stream(dev))
        for st in streams: st.wait_stream(torch.cuda.default_stream(dev))
        fn()
        for st in streams: torch.cuda.default_stream(dev).wait_stream(st)
        e1.record(torch.cuda.default_stream(dev))
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]

run_one(); torch.cuda.synchronize(); ref = dsT.clone()
t1 = timeit(run_one)
run_split(); torch.cuda.synchronize()
same = bool(torch.equal(ref, dsT))
t2 = timeit(run_split)
print(f"n={n} one stream {t1:.3f} ms ({n / t1:.1f} k OCP/s) | {parts} streams {t2:.3f} ms ({n / t2:.1f} k OCP/s) | ratio {t1 / t2:.4f} | same results {same}")
