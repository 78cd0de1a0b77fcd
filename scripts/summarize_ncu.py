"""Summarises the ncu captures a gpurun call left in gpurun_out/ into profiles/ (tracked):

  gpurun_out/launches.csv      ncu --metrics gpu__time_duration.sum            -> profiles/<tag>_launch_summary.md (+ head csv)
  gpurun_out/hist_traffic.csv  ncu --metrics dram__bytes_*,gpu__time_duration -k regex:hist_float
                                                                                -> profiles/hist_traffic.json (read by bench.py)
  gpurun_out/hist_full.ncu-rep ncu --set full -k regex:hist_float             -> profiles/<tag>_hist_float_full.md

All captures are taken with GBL_USE_GRAPH=0 (the learner then drives its waves from the host; the kernels are the
same ones the CUDA graph replays).  Usage: python scripts/summarize_ncu.py r1
"""
import collections
import csv
import io
import json
import os
import subprocess
import sys

tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
CMD_LIST = ("GBL_USE_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -s 330 -c 400 python bench.py "
            "--steps 2 --warmup 3 --no-cpu-baseline")
CMD_TRAFFIC = ("GBL_USE_GRAPH=0 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum "
               "--clock-control none -k regex:hist_float -s 54 -c 36 python bench.py --steps 2 --warmup 3 --no-cpu-baseline")
CMD_FULL = ("GBL_USE_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:hist_float -s 54 -c 3 "
            "python bench.py --steps 1 --warmup 3 --no-cpu-baseline")


def rows_of(path):
    lines = open(path).read().split("\n")
    start = [i for i, l in enumerate(lines) if l.startswith('"ID"')][0]
    return list(csv.DictReader(io.StringIO("\n".join(lines[start:]))))


def val(r):
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    return {"ns": v / 1e3, "us": v, "ms": v * 1e3, "s": v * 1e6}.get(u, v)


def short(name):
    return name.split("(")[0].replace("void ", "")


# ------------------------------------------------------------------------------------------------ launch list
rows = rows_of("gpurun_out/launches.csv")
seq = [(short(r["Kernel Name"]), val(r)) for r in rows]
starts = [i for i, (k, _) in enumerate(seq) if k.startswith("tree_begin")]
tree = seq[starts[0]:starts[1]] if len(starts) > 1 else seq
agg = collections.defaultdict(lambda: [0, 0.0])
for k, v in tree:
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
with open("profiles/%s_launch_summary.md" % tag, "w") as fh:
    fh.write("# ncu launch list summary (%s): one whole tree of the bench workload\n\n" % tag)
    fh.write("Command: `%s` (Higgs shape 10.5M x 28, float mode).  Per-launch times under ncu are cold-cache and "
             "serialised: compare SHARES, not absolutes.  The tree below is the first complete one in the capture "
             "window (`tree_begin` .. next `tree_begin`; the window also contains the score update, the next "
             "iteration's gradients and the per-learner record build of the bench's other boosters).\n\n" % CMD_LIST)
    fh.write("| kernel | launches | total us | avg us | share |\n|---|---|---|---|---|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        fh.write("| `%s` | %d | %.1f | %.2f | %.1f%% |\n" % (k, v[0], v[1], v[1] / v[0], 100 * v[1] / tot))
    fh.write("\nTotal %d launches, %.1f us.\n\n## Wave by wave (us)\n\n" % (len(tree), tot))
    fh.write("| wave | part_count | part_plan | part_scatter | histogram | scan + select |\n|---|---|---|---|---|---|\n")
    wave, cur, pre = 0, {}, []
    for k, v in tree:
        if k.startswith("find_best"):
            cur["find"] = v
        elif k.startswith("select"):
            v = v + cur.get("find", 0.0)
            if wave == 0:
                fh.write("| 0 (root; before it: %s) | - | - | - | %.1f | %.1f |\n" % (", ".join(pre), cur.get("hist", 0.0), v))
            else:
                fh.write("| %d | %.1f | %.1f | %.1f | %.1f | %.1f |\n" % (wave, cur.get("part_count_kernel", 0.0),
                                                                       cur.get("part_plan_kernel", 0.0),
                                                                       cur.get("part_scatter_kernel", 0.0), cur.get("hist", 0.0), v))
            wave, cur = wave + 1, {}
        elif k.startswith("hist_"):
            cur["hist"] = v
        elif k.startswith("part_"):
            cur[k] = v
        elif wave == 0:
            pre.append("%s %.1f" % (k.replace("_kernel", ""), v))
        else:
            cur.setdefault("post", []).append("%s %.1f" % (k.replace("_kernel", ""), v))
    if cur.get("post"):
        fh.write("\nAfter the last wave: %s.\n" % ", ".join(cur["post"]))
with open("profiles/%s_launches_head.csv" % tag, "w") as fh:
    w = csv.writer(fh)
    w.writerow(["ID", "Kernel Name", "Grid", "Block", "duration_us"])
    for r in rows[:300]:
        w.writerow([r["ID"], short(r["Kernel Name"]), r.get("Grid Size", ""), r.get("Block Size", ""), "%.3f" % val(r)])

# ------------------------------------------------------------------------------------------------ DRAM traffic
tr = rows_of("gpurun_out/hist_traffic.csv")
per = collections.defaultdict(dict)
for r in tr:
    per[r["ID"]][r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])


def to_bytes(v, u):
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)


def to_us(v, u):
    return v * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1.0)


n = len(per)
rd = sum(to_bytes(*d["dram__bytes_read.sum"]) for d in per.values())
wr = sum(to_bytes(*d["dram__bytes_write.sum"]) for d in per.values())
us = sum(to_us(*d["gpu__time_duration.sum"]) for d in per.values())
out = {"kernel": "hist_float_kernel", "launches": n, "dram_bytes_read_per_launch": rd / n, "dram_bytes_write_per_launch": wr / n,
       "dram_bytes_per_launch": (rd + wr) / n, "avg_launch_us_under_ncu": us / n,
       "dram_gbs_under_ncu": (rd + wr) / us / 1e3,
       "source": "%s (36 consecutive launches = two trees of the bench workload, %s)" % (CMD_TRAFFIC, tag)}
json.dump(out, open("profiles/hist_traffic.json", "w"), indent=1)
print(json.dumps(out))

# ------------------------------------------------------------------------------------------------ full capture
rep = "gpurun_out/hist_full.ncu-rep"
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(io.StringIO(raw)))
    hdr, units = rr[0], rr[1]
    idx = {h: i for i, h in enumerate(hdr)}
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__inst_executed.sum", "smsp__inst_executed_op_shared_atom.sum",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
            "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "launch__registers_per_thread",
            "launch__shared_mem_per_block_dynamic"]
    with open("profiles/%s_hist_float_full.md" % tag, "w") as fh:
        fh.write("# ncu --set full: hist_float_kernel (%s)\n\nCommand: `%s`.  Three consecutive launches of one tree "
                 "(gather mode: the rows of all smaller children of a wave).  rows = shared-atomic instructions / 4 "
                 "(4 ATOMS per row and 32-feature tile; Higgs has one tile).\n\n" % (tag, CMD_FULL))
        fh.write("| metric | " + " | ".join("launch %d" % (i + 1) for i in range(len(rr) - 2)) + " |\n")
        fh.write("|---|" + "---|" * (len(rr) - 2) + "\n")
        for m in want:
            if m in idx:
                fh.write("| `%s` [%s] | %s |\n" % (m, units[idx[m]], " | ".join(r[idx[m]] for r in rr[2:])))
        a, b = idx.get("smsp__inst_executed_op_shared_atom.sum"), idx.get("dram__bytes_read.sum")
        if a is not None and b is not None:
            cells = []
            for r in rr[2:]:
                nrows = float(r[a].replace(",", "")) / 4.0
                bytes_rd = to_bytes(float(r[b].replace(",", "")), units[b])
                cells.append("%.2f M rows, %.0f B/row" % (nrows / 1e6, bytes_rd / nrows))
            fh.write("| derived: rows, DRAM read bytes per row (algorithmic: 40) | %s |\n" % " | ".join(cells))
        fh.write("\nReading: with the 32 bin loads of a batch in flight together, `mio_throttle` (the LSU/shared-memory "
                 "instruction queue, kept full by 4 `ATOMS` per row and tile at 2 cycles each) has overtaken "
                 "`long_scoreboard` (waiting for the gathered row records) as the top stall; issue slots are about one "
                 "third busy.  These ~0.5 M-row launches move 132 B of DRAM traffic per row for 40 algorithmic bytes: "
                 "sparse gathers fetch whole 128-byte lines.  Bank conflicts stay under 1 % of the atom wavefronts "
                 "(lane = feature = bank).\n")
    print(open("profiles/%s_hist_float_full.md" % tag).read())
print(open("profiles/%s_launch_summary.md" % tag).read())
