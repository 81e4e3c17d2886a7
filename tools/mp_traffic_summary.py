"""ncu CSV (--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum over tools/profile_encode.py) ->
profiles/r02_mp_ncu_summary.json: DRAM traffic and time of the matching-pursuit launches of ONE image (the last encode in the
file), per kernel and in total. bench.py reads the JSON for roofline.traffic.

    python tools/mp_traffic_summary.py gpurun_out/<csv> [profiles/r02_mp_ncu_summary.json]
"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1]
dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r02_mp_ncu_summary.json")
rows = list(csv.reader(open(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[hi]
col = {n: hdr.index(n) for n in ("ID", "Kernel Name", "Metric Name", "Metric Unit", "Metric Value")}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3}
launches = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= col["Metric Value"]:
        continue
    lid = int(r[col["ID"]])
    d = launches.setdefault(lid, {"name": r[col["Kernel Name"]].split("(")[0].split("::")[-1]})
    d[r[col["Metric Name"]]] = float(r[col["Metric Value"]].replace(",", "")) * SCALE.get(r[col["Metric Unit"]], 1.0)
ids = list(launches)
starts = [i for i in ids if launches[i]["name"] == "waveInitKernel"]
last = [launches[i] for i in ids if i >= starts[-1]]
per = collections.OrderedDict()
for d in last:
    if not d["name"].startswith("wave"):
        continue
    k = per.setdefault(d["name"], {"launches": 0, "us": 0.0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0})
    k["launches"] += 1
    k["us"] += d.get("gpu__time_duration.sum", 0.0)
    k["dram_read_bytes"] += d.get("dram__bytes_read.sum", 0.0)
    k["dram_write_bytes"] += d.get("dram__bytes_write.sum", 0.0)
for k in per.values():
    k["us"] = round(k["us"], 1)
    k["dram_read_bytes"], k["dram_write_bytes"] = int(k["dram_read_bytes"]), int(k["dram_write_bytes"])
total = sum(k["dram_read_bytes"] + k["dram_write_bytes"] for k in per.values())
out = {"source": os.path.basename(src), "what": "matching-pursuit launches of one 4928x3264 image at quality 3.5 (ncu, per-launch, cold caches)",
       "dram_bytes_per_image": int(total), "us_per_image_serialised": round(sum(k["us"] for k in per.values()), 1), "per_kernel": per}
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps(out))
