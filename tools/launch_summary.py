"""Summarise an `ncu --metrics gpu__time_duration.sum[,launch__grid_size,launch__block_size] --csv` launch list:
per-kernel totals, and optionally the launch sequence of one subtraction.
    python tools/launch_summary.py gpurun_out/launches.csv [--sequence N]"""
import collections
import csv
import re
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, cur = None, collections.OrderedDict()
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            ki, mi, ni, ii = r.index("Kernel Name"), r.index("Metric Value"), r.index("Metric Name"), r.index("ID")
            continue
        if hdr is None or len(r) <= mi:
            continue
        try:
            v = float(r[mi].replace(",", ""))
        except ValueError:
            continue
        k = re.sub(r"^void ", "", re.sub(r"\(.*", "", r[ki]))
        cur.setdefault((int(r[ii]), k), {})[r[ni]] = v
    return cur


def main():
    cur = load(sys.argv[1])
    agg = collections.OrderedDict()
    for (i, k), d in cur.items():
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += d.get("gpu__time_duration.sum", 0.0)
    tot = sum(a[1] for a in agg.values())
    nsub = agg.get("k_make_kernels", [1])[0]
    print("subtractions in the list: %d; kernel time per subtraction %.1f us; %d launches" % (nsub, tot / 1e3 / nsub, sum(a[0] for a in agg.values())))
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-34s n=%5d  per-subtraction %8.1f us  avg %8.1f us  share %5.1f%%" % (k[:34], a[0], a[1] / 1e3 / nsub, a[1] / 1e3 / a[0], 100 * a[1] / tot))
    if "--sequence" in sys.argv:
        n = int(sys.argv[sys.argv.index("--sequence") + 1])
        ids = sorted(cur)
        starts = [i for i, (_, k) in enumerate(ids) if k == "k_input_mask"]
        a, b = starts[n], starts[n + 1] if n + 1 < len(starts) else len(ids)
        print("\nlaunch sequence of subtraction %d (LU and small-copy launches folded):" % n)
        fold = collections.Counter()
        for id_, k in ids[a:b]:
            d = cur[(id_, k)]
            if k.startswith("k_lu_") or k in ("k_copy_words", "k_zero_words"):
                fold[k] += d["gpu__time_duration.sum"]
                continue
            print("  %-30s %8.1f us  grid %6d x %d" % (k[:30], d["gpu__time_duration.sum"] / 1e3, d.get("launch__grid_size", 0), d.get("launch__block_size", 0)))
        for k, v in fold.items():
            print("  %-30s %8.1f us  (all launches)" % (k, v / 1e3))


if __name__ == "__main__":
    main()
