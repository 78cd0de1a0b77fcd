"""Summarise one tree (tree_begin .. end of window) of an ncu launch-list CSV: per-kernel totals and, with -v, the
histogram / partition launches one by one.   python scripts/ncu_tree.py CSV [-v]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
verbose = "-v" in sys.argv
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[hi], rows[hi + 1:]
ki, mi, vi, ii, ui = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "ID", "Metric Unit"))
per = {}
for r in data:
    if len(r) <= vi:
        continue
    d = per.setdefault(int(r[ii]), {"k": r[ki]})
    v = float(r[vi].replace(",", ""))
    if r[mi].startswith("gpu__time") and r[ui] == "ns":
        v /= 1e3
    d[r[mi]] = v
ids = sorted(per)
tb = [i for i in ids if per[i]["k"].startswith("tree_begin")]
a = tb[0]
b = tb[1] if len(tb) > 1 else ids[-1] + 1
tot = collections.OrderedDict()
for i in ids:
    if a <= i < b:
        d = per[i]
        k = d["k"].split("(")[0].replace("void ", "")[:30]
        t = tot.setdefault(k, [0, 0.0, 0.0, 0.0])
        t[0] += 1
        t[1] += d.get("gpu__time_duration.sum", 0.0)
        t[2] += d.get("dram__bytes_read.sum", 0.0) / 1e6
        t[3] += d.get("dram__bytes_write.sum", 0.0) / 1e6
        if verbose and ("hist" in k or "part_" in k or "find" in k or "select" in k):
            print(i, k, "%.1f us  R %.1f MB  W %.1f MB" % (d.get("gpu__time_duration.sum", 0.0), d.get("dram__bytes_read.sum", 0) / 1e6,
                                                            d.get("dram__bytes_write.sum", 0) / 1e6))
gt = sum(t[1] for t in tot.values())
for k, t in tot.items():
    print("%-32s n=%3d total %8.1f us (%4.1f%%)  avg %6.1f us  DRAM R %8.1f MB  W %7.1f MB" % (k, t[0], t[1], 100 * t[1] / gt, t[1] / t[0], t[2], t[3]))
print("tree total %.1f us over %d launches" % (gt, sum(t[0] for t in tot.values())))
