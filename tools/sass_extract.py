"""SASS evidence for the tensor-core / TMA / mbarrier path: per kernel of libmnenc.so, how many of the instructions named in
/opt/skills/guides/B200_PROFILING.md appear (UTCHMMA = tcgen05.mma, LDTM = tcgen05.ld, UBLKCP = cp.async.bulk, SYNCS = mbarrier,
LDGSTS = cp.async, UTCBAR = tcgen05.commit), plus registers per thread.   python tools/sass_extract.py > profiles/r02_sass_extract.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, "imageexperiments_b200", "libmnenc.so")
WANT = ("UTCHMMA", "UTCBAR", "LDTM", "UBLKCP", "SYNCS", "LDGSTS", "ELECT", "FMNMX3", "UTCATOMSWS", "ATOMG", "REDG", "RED", "HMMA", "FFMA", "FMUL", "FADD")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
fn = None
counts = collections.OrderedDict()
total = collections.Counter()
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"\(.*", "", fn.replace("(anonymous namespace)::", "").replace("<unnamed>::", ""))
        counts[fn] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and fn:
        total[fn] += 1
        op = m.group(1)
        if op in WANT:
            counts[fn][op] += 1
res = subprocess.run(["cuobjdump", "-res-usage", lib], capture_output=True, text=True).stdout
regs = {}
cur = None
for line in res.splitlines():
    m = re.search(r"Function (\S+):", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*", "", cur.replace("(anonymous namespace)::", "").replace("<unnamed>::", ""))
    m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
    if m and cur:
        regs[cur] = (int(m.group(1)), int(m.group(2)))
print(f"# {os.path.relpath(lib, ROOT)}: cuobjdump -sass / -res-usage, sm_100a")
for fn, c in counts.items():
    r = regs.get(fn, ("?", "?"))
    ops = ", ".join(f"{k} {v}" for k, v in sorted(c.items(), key=lambda kv: -kv[1]))
    print(f"{fn}: {total[fn]} instructions, {r[0]} registers, {r[1]} B static smem; {ops}")
