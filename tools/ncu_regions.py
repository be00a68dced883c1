"""tools/ncu_regions.py — summarise `ncu --page source --csv` output: instructions and stall samples per
opcode and per run of SASS lines.  Usage: ncu -i rep --page source --csv | python tools/ncu_regions.py [chunk]"""
import collections
import csv
import sys

rows = list(csv.reader(sys.stdin))
h = rows[1]
idx = {n: i for i, n in enumerate(h)}
chunk = int(sys.argv[1]) if len(sys.argv) > 1 else 300
keys = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
seg = []
for r in rows[2:]:
    if len(r) < len(h):
        continue
    try:
        inst = int(r[idx["Instructions Executed"]]); samp = int(r[idx["# Samples"]])
    except ValueError:
        continue
    src = r[idx["Source"]].split()
    op = src[0] if not src[0].startswith("@") else src[1]
    seg.append((op, inst, samp, {k: int(r[idx[k]] or 0) for k in keys}))
ti = sum(x[1] for x in seg); ts = sum(x[2] for x in seg)
print("total inst", ti, "samples", ts, "sass lines", len(seg))
by = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for op, inst, samp, st in seg:
    b = by[op.split(".")[0]]; b[0] += inst; b[1] += samp; b[2].update(st)
for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])[:10]:
    print("%-8s inst %10d (%4.1f%%) samples %6d (%4.1f%%)" % (k, v[0], 100 * v[0] / ti, v[1], 100 * v[1] / ts), dict(v[2].most_common(3)))
for i in range(0, len(seg), chunk):
    c = seg[i:i + chunk]; cs = collections.Counter()
    for x in c:
        cs.update(x[3])
    print("%5d inst %10d samples %6d (%4.1f%%) dfma %10d" % (i, sum(x[1] for x in c), sum(x[2] for x in c), 100 * sum(x[2] for x in c) / ts,
                                                     sum(x[1] for x in c if x[0].startswith("DFMA"))), dict(cs.most_common(3)))
