#!/usr/bin/env python3
"""Attribute ncu warp-stall samples of a kernel to CUDA source lines.
usage: ncu_hotlines.py <report.ncu-rep> <lib.so> <kernel-substring> [topN]
Joins `ncu --page source --csv` (SASS view: address, samples) with `nvdisasm -g` line info of the sm_100a cubin by
instruction offset inside the kernel's .text section."""
import csv, collections, glob, os, re, subprocess, sys, tempfile

rep, lib, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
cubin = [f for f in glob.glob(tmp + "/*.cubin") if "sm_100" in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of = {}          # offset -> (file, line) for the wanted kernel (inlined-at chains: keep the innermost)
insec = False; cur = None
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m:
        insec = kern in m.group(1); continue
    if not insec: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m: line_of[int(m.group(1), 16)] = (cur, m.group(2))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Address")
body = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
base = int(body[0][0], 16)
iS = hdr.index("# Samples"); iE = hdr.index("Instructions Executed")
by = collections.Counter(); ex = collections.Counter(); tot = 0
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
stalls = collections.defaultdict(collections.Counter)
for r in body:
    off = int(r[0], 16) - base
    key = line_of.get(off, ((None, 0), ""))[0]
    s = int(r[iS] or 0); tot += s
    by[key] += s; ex[key] += int(r[iE] or 0)
    for i in stall_cols: stalls[key][hdr[i][6:]] += int(r[i] or 0)
print(f"total samples {tot}, instructions {sum(ex.values())}")
srccache = {}
for key, s in by.most_common(top):
    f, ln = key if key else (None, 0)
    text = ""
    if f:
        p = [q for q in glob.glob(os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", f))]
        if p:
            srccache.setdefault(p[0], open(p[0]).read().splitlines())
            text = srccache[p[0]][ln - 1].strip()[:100]
    why = " ".join(f"{k}={v}" for k, v in stalls[key].most_common(3) if v)
    print(f"{100*s/tot:5.1f}%  inst {ex[key]:>11}  {f}:{ln}  {text[:70]}  | {why}")
