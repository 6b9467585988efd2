#!/usr/bin/env python3
"""Per-phase SM-cycle breakdown of the ADMM kernel (debug build `make prof`, -DMPCB_PROFILE).

    make prof && MPCB_LIB=mpc_motion_planner_b200/libmpcb_prof.so python tools/phase_profile.py [n]

Thread 0 of CTA 0 accumulates clock64() deltas after every phase; the numbers are cycles per ADMM iteration of one OCP
with all SMs busy (n >= SM count).  Not a bench value (the hooks cost a few cycles per phase)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import mpc_motion_planner_b200 as M

n = int(sys.argv[1]) if len(sys.argv) > 1 else 148
margins = [0.8, 0.8, 0.6, 0.9, 0.1]
s = M.default_settings(); lim = M.default_limits(margins)
x0, xT = M.generate_problems(lim, 0, 1, n)
gx, gu, gT = M.guess_batch(s, lim, x0, xT)
sv = M.Solver(s, lim, device=0)
out = (C.c_longlong * 512)()
sv.solve(x0, xT, gx, gu, gT)
sv.L.mpcb_debug_phase_cycles(sv.h, out, 1)
sx, su, sT, _, st = sv.solve(x0, xT, gx, gu, gT)
sv.L.mpcb_debug_phase_cycles(sv.h, out, 1)
names = ["load", "A rhs", "B interior W/F^T", "B' separator rhs", "C separator solve", "D interior back-subst", "E row/box update",
         "residual test", "store"]
cta0_ocps = max(1, round(n / 148))          # CTA 0 handles about n/148 OCPs per ADMM launch
iters = 2 * 700.0 * cta0_ocps
tot = sum(out[i] for i in range(1, 8))
print(f"n={n}  admm kernel ms={sv.last_kernel_ms(1):.3f}  stage ms={sv.last_kernel_ms(0):.3f}  node ms={sv.last_kernel_ms(2):.3f}  (cycles below: CTA 0, ~{cta0_ocps} OCP(s) x 2 QPs)")
for i, nm in enumerate(names):
    per = out[i] / iters if 1 <= i <= 7 else out[i] / (2.0 * cta0_ocps)
    print(f"  {nm:26s} {out[i]:14d} cycles   {per:10.1f} per {'iteration' if 1 <= i <= 7 else 'QP'}")
print(f"  iteration total            {tot / iters:10.1f} cycles")
stage = {16: "load guess", 17: "linearise (x2)", 18: "KKT assembly (x2)", 19: "interior sweeps (x2)", 20: "separator (x2)", 
         21: "export (x2)", 22: "merit + line search (x2)", 23: "step + termination (x2)", 24: "store"}
print("SQP stage kernels, cycles of CTA 0 per OCP (whole solve):")
for k, nm in stage.items():
    print(f"  {nm:28s} {out[k] / cta0_ocps:12.0f}")
print(f"  total                        {sum(out[k] for k in stage) / cta0_ocps:12.0f}")
print(f"  (factorise as a whole = assembly + sweeps + separator: {out[25] / cta0_ocps:.0f})")

names2 = ["A rhs", "B1 F^T", "B2 W + sep rhs", "C separator", "D back-subst", "E update"]
t0 = min(v for v in (out[64 + 16 * 15 + w] for w in range(16)) if v > 0)
print("per-warp end-of-work times in iteration 100 of CTA 0 (cycles after the start of phase A; then the barrier):")
for s_, nm in enumerate(names2):
    ends = [out[64 + 16 * s_ + w] - t0 for w in range(16)]
    print(f"  {nm:16s} " + " ".join(f"{e:5d}" for e in ends))
print(f"  {'after E barrier':16s} " + " ".join(f"{out[64 + 16 * 6 + w] - t0:5d}" for w in range(16)))
inner = {7: "B: vector + chunk 0 arrived", 8: "B: chunk-0 products issued", 9: "B: register products issued", 10: "B: all products issued (before the reduce)", 11: "B: transpose-reduce done", 12: "B': unit sums formed", 13: "B': first pair exchange done"}
if any(out[64 + 16 * k_ + w] for k_ in inner for w in range(16)):
    print("time stamps inside phase B (same origin):")
    for k_, nm in inner.items():
        print(f"  {nm:44s} " + " ".join(f"{(out[64 + 16 * k_ + w] - t0) if out[64 + 16 * k_ + w] else 0:5d}" for w in range(16)))
