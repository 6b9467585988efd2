import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mpc_motion_planner_b200 as M
po, ns = int(sys.argv[1]), int(sys.argv[2])
kw = dict(poly_order=po, num_seg=ns)
for a in sys.argv[3:]:
    k, v = a.split("="); kw[k] = int(v)
s = M.default_settings(**kw); lim = M.default_limits([0.8, 0.8, 0.6, 0.9, 0.1])
x0, xT = M.generate_problems(lim, 0, 1, 4)
gx, gu, gT = M.guess_batch(s, lim, x0, xT)
sv = M.Solver(s, lim)
try:
    sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
    print(po, ns, kw, "T", sT, "iters", st["qp_iters"][:, :2].tolist(), "ms", sv.last_solve_ms())
except Exception as e:
    print(po, ns, kw, "FAILED", e)
