#!/usr/bin/env python3
"""Instruction mix of a kernel from an ncu --set full report: warp-level executed instructions by SASS opcode and by source line.
usage: ncu_instmix.py <report.ncu-rep> <lib.so> <kernel-substring> [per] [topN]
`per` divides the counts (e.g. 148*700 = iterations x CTAs of the captured launch) so that the figures read per iteration per SM."""
import csv, collections, glob, os, re, subprocess, sys, tempfile

rep, lib, kern = sys.argv[1:4]
per = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
cubin = [f for f in glob.glob(tmp + "/*.cubin") if "sm_100" in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
line_of = {}
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
iE = hdr.index("Instructions Executed"); iSrc = hdr.index("Source")
byop = collections.Counter(); byline = collections.Counter(); lineops = collections.defaultdict(collections.Counter)
for r in body:
    off = int(r[0], 16) - base
    n = int(r[iE] or 0)
    sass = r[iSrc].strip()
    sass = re.sub(r"^@!?U?P\d+\s+", "", sass)
    op = sass.split()[0].split(".")[0] if sass else "?"
    full = sass.split()[0] if sass else "?"
    byop[full if op in ("LDS", "STS", "LDTM", "STTM", "SHFL", "BAR", "LDG", "STG", "LDL", "STL") else op] += n
    key = line_of.get(off, ((None, 0), ""))[0]
    byline[key] += n; lineops[key][op] += n
tot = sum(byop.values())
print(f"warp-level instructions {tot}  ({tot/per:.1f} per unit)")
print("-- by opcode")
for op, n in byop.most_common(top):
    print(f"  {100*n/tot:5.1f}%  {n/per:10.1f}  {op}")
print("-- by source line")
for key, n in byline.most_common(top):
    f, ln = key if key else (None, 0)
    ops = " ".join(f"{k}={v/per:.0f}" for k, v in lineops[key].most_common(5))
    print(f"  {100*n/tot:5.1f}%  {n/per:10.1f}  {f}:{ln}  | {ops}")
