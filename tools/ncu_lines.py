"""Aggregate the SASS-level stall samples of an .ncu-rep per CUDA source line (needs -lineinfo).
usage: python tools/ncu_lines.py gpurun_out/x.ncu-rep imageexperiments_b200/libmnenc.so <cubin-substring> [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, so, pat = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
cubin = [f for f in os.listdir(tmp) if pat in f][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
kernel = rows[0][1]
short = re.sub(r"\(.*", "", kernel).split("::")[-1]
cur, addr2line, infn = None, {}, False
for l in sass.splitlines():
    if l.startswith("\t.section") or l.startswith(".section"):
        infn = short in l and ".text." in l
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S.*?);", l)
    if m and infn:
        addr2line[int(m.group(1), 16)] = cur
hdr, data = rows[1], rows[2:]
ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.defaultdict(lambda: [0, 0, collections.Counter()])
base = int(data[0][ia], 16)
tot = totex = 0
for r in data:
    ln = addr2line.get(int(r[ia], 16) - base)
    s, e = int(r[isamp] or 0), int(r[iex] or 0)
    agg[ln][0] += s
    agg[ln][1] += e
    for i in stall:
        v = int(r[i] or 0)
        if v:
            agg[ln][2][hdr[i]] += v
    tot += s
    totex += e
print(f"# {kernel}: {tot} samples, {totex} warp instructions")
print("| source line | % samples | % instructions | top stall reasons |\n|---|---|---|---|")
for ln, (s, e, c) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    name = f"{ln[0]}:{ln[1]}" if ln else "?"
    print(f"| {name} | {100 * s / tot:.1f} | {100 * e / totex:.1f} | {', '.join(f'{k[6:]} {v}' for k, v in c.most_common(3))} |")
