"""Summarises gpurun_out/launches.csv (ncu --metrics gpu__time_duration.sum) and gpurun_out/hist_traffic.csv into
profiles/ (tracked).  Usage: python scripts/summarize_ncu.py r1"""
import collections
import csv
import io
import json
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"


def rows_of(path):
    lines = open(path).read().split("\n")
    start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
    return list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))


def val(r):
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    return {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(u, v)


rows = rows_of("gpurun_out/launches.csv")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    k = r["Kernel Name"].split("(")[0]
    agg[k][0] += 1
    agg[k][1] += val(r)
tot = sum(v[1] for v in agg.values())
with open("profiles/%s_launch_summary.md" % tag, "w") as fh:
    fh.write("# ncu launch list summary (%s)\n\n" % tag)
    fh.write("Command: `ncu --metrics gpu__time_duration.sum --clock-control none -s 1100 -c 2200 python bench.py --steps 2 "
             "--warmup 3 --no-cpu-baseline` (Higgs-shape 10.5M x 28, float mode).  Per-launch times under ncu are cold-cache "
             "and serialised: compare SHARES, not absolutes.\n\n| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        fh.write("| `%s` | %d | %.1f | %.2f | %.1f%% |\n" % (k, v[0], v[1], v[1] / v[0], 100 * v[1] / tot))
    fh.write("\nTotal %d launches, %.1f us.\n" % (len(rows), tot))
with open("profiles/%s_launches_head.csv" % tag, "w") as fh:
    w = csv.writer(fh)
    w.writerow(["ID", "Kernel Name", "Grid", "Block", "duration_us"])
    for r in rows[:600]:
        w.writerow([r["ID"], r["Kernel Name"].split("(")[0], r.get("Grid Size", ""), r.get("Block Size", ""), "%.3f" % val(r)])

tr = rows_of("gpurun_out/hist_traffic.csv")
per = collections.defaultdict(dict)
for r in tr:
    per[r["ID"]][r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])


def to_bytes(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


n = len(per)
rd = sum(to_bytes(*d["dram__bytes_read.sum"]) for d in per.values())
wr = sum(to_bytes(*d["dram__bytes_write.sum"]) for d in per.values())
out = {"kernel": "hist_float_kernel", "launches": n, "dram_bytes_read_per_launch": rd / n, "dram_bytes_write_per_launch": wr / n,
       "dram_bytes_per_launch": (rd + wr) / n,
       "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k regex:hist_float -s 254 -c 254 (one whole tree of "
                 "the bench workload, %s)" % tag}
json.dump(out, open("profiles/hist_traffic.json", "w"), indent=1)
print(json.dumps(out))
print(open("profiles/%s_launch_summary.md" % tag).read())
