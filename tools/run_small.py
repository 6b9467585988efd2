#!/usr/bin/env python3
"""One small solve (n OCPs of the mpc_benchmark generator) through the C ABI: the command ncu captures kernels from."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mpc_motion_planner_b200 as M
n = int(sys.argv[1]) if len(sys.argv) > 1 else 148
s = M.default_settings(); lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
x0, xT = M.generate_problems(lim, 0, 1, n)
gx, gu, gT = M.guess_batch(s, lim, x0, xT)
sv = M.Solver(s, lim, device=0)
sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
print("n", n, "mean T", float(sT.mean()), "ms", sv.last_solve_ms(), "admm ms", sv.last_kernel_ms(1), "stage ms", sv.last_kernel_ms(0))
