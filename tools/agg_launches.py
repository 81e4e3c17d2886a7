"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum --csv) of tools/profile_encode.py: per-kernel totals of the LAST
encode in the file and the per-iteration times of the matching-pursuit kernels.   python tools/agg_launches.py <csv>"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
hdr = rows[hi]
kn, mv, mu = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
data = rows[hi + 1:]
starts = [i for i, r in enumerate(data) if 'waveInit' in r[kn]]
seg = data[starts[-1]:]


def us(r):
    v = float(r[mv].replace(',', ''))
    return v / 1e3 if r[mu] == 'ns' else (v * 1e3 if r[mu] == 'ms' else v)


agg = collections.OrderedDict()
for r in seg:
    n = r[kn].split('(')[0].replace('unnamed>::', '')
    a = agg.setdefault(n, [0, 0.0])
    a[0] += 1
    a[1] += us(r)
tw = 0.0
for n, (c, t) in agg.items():
    print(f"{n:60s} {c:4d} {t:10.1f} us")
    if 'wave' in n:
        tw += t
print('matching pursuit kernels', round(tw, 1), 'us; all kernels of the encode', round(sum(t for c, t in agg.values()), 1), 'us')
for key in ['waveBase', 'waveDetail', 'waveDecide', 'waveOverflow', 'waveBinFill']:
    print(key, [round(us(r), 1) for r in seg if key in r[kn]][:20])
