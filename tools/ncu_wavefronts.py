#!/usr/bin/env python3
"""Shared-memory wavefronts of a kernel by CUDA source line (ncu source page joined with nvdisasm line info).
usage: ncu_wavefronts.py <report.ncu-rep> <lib.so> <kernel-substring> [iterations-per-SM]"""
import csv, collections, glob, os, re, subprocess, sys, tempfile
rep, lib, kern = sys.argv[1:4]
iters = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
cubin = [f for f in glob.glob(tmp + "/*.cubin") if "sm_100" in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of = {}; insec = False; cur = None
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m: insec = kern in m.group(1); continue
    if not insec: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m: line_of[int(m.group(1), 16)] = (cur, m.group(2))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Address")
body = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
base = int(body[0][0], 16)
iW = hdr.index("L1 Wavefronts Shared"); iI = hdr.index("L1 Wavefronts Shared Ideal"); iS = hdr.index("Source"); iE = hdr.index("Instructions Executed")
by = collections.Counter(); ideal = collections.Counter(); kinds = collections.defaultdict(collections.Counter)
tot = 0
for r in body:
    w = float(r[iW] or 0)
    if w == 0: continue
    off = int(r[0], 16) - base
    key = line_of.get(off, ((None, 0), ""))[0]
    by[key] += w; ideal[key] += float(r[iI] or 0); tot += w
    kinds[key][r[iS].split()[0] if not r[iS].startswith("@") else r[iS].split()[1]] += int(r[iE] or 0)
print(f"total shared wavefronts {tot:.0f}  (per iteration per SM: {tot / 148 / iters:.0f})")
for key, w in by.most_common(40):
    f, ln = key if key else (None, 0)
    print(f"{100 * w / tot:5.1f}%  {w / 148 / iters:8.1f}/it  ideal {ideal[key] / 148 / iters:8.1f}  {f}:{ln}  {dict(kinds[key])}")
