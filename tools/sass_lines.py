#!/usr/bin/env python3
"""Static SASS instruction count per CUDA source line of one kernel (nvdisasm -g line info), optionally restricted to an
address window: usage: sass_lines.py <lib.so|obj.o> <kernel-substring> [lo_hex hi_hex] [topN]"""
import collections, glob, os, re, subprocess, sys, tempfile
lib, kern = sys.argv[1:3]
lo = int(sys.argv[3], 16) if len(sys.argv) > 4 else 0
hi = int(sys.argv[4], 16) if len(sys.argv) > 4 else 1 << 30
top = int(sys.argv[5]) if len(sys.argv) > 5 else 60
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, stdout=subprocess.DEVNULL, check=True)
cubin = [f for f in glob.glob(tmp + "/*.cubin") if "sm_100" in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout.splitlines()
cnt = collections.Counter(); ops = collections.defaultdict(collections.Counter)
insec = False; cur = None
for l in dis:
    m = re.match(r"\s*\.text\.(\S+):", l)
    if m: insec = kern in m.group(1); continue
    if not insec: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m:
        a = int(m.group(1), 16)
        if lo <= a < hi:
            cnt[cur] += 1
            op = m.group(2).split()[0]
            if op.startswith("@"): op = m.group(2).split()[1]
            ops[cur][op.split(".")[0]] += 1
print("total", sum(cnt.values()))
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mpc_motion_planner_b200", "csrc")
cache = {}
for key, n in cnt.most_common(top):
    f, ln = key if key else ("?", 0)
    text = ""
    p = os.path.join(root, f)
    if os.path.exists(p):
        cache.setdefault(p, open(p).read().splitlines()); text = cache[p][ln - 1].strip()[:70]
    print(f"{n:5d}  {f}:{ln:<4d} {text:70s} | " + " ".join(f"{k}={v}" for k, v in ops[key].most_common(5)))
