#!/usr/bin/env python3
"""Every device entry point once on a tiny batch (generate -> warm start -> solve -> sample): the command compute-sanitizer runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mpc_motion_planner_b200 as M
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
s = M.default_settings(); lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
if len(sys.argv) > 2:
    s.qp_max_iter = int(sys.argv[2])
sv = M.Solver(s, lim, device=0)
x0, xT = sv.generate(0, 1, n)
gx, gu, gT = sv.guess(x0, xT)
sx, su, sT, lam, st = sv.solve(x0, xT, gx, gu, gT, want_lam=True)
traj, acc = sv.sample(sx, su, sT, target=xT, npts=200)
print("n", n, "T", np.round(sT, 6).tolist(), "traj", traj.shape, "launches", sv.launch_count())
