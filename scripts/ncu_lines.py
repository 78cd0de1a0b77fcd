"""Per-source-line summary of one kernel launch from an ncu report (needs -lineinfo + --import-source on).
   python scripts/ncu_lines.py REPORT KERNEL_REGEX LAUNCH_SKIP [TOP]"""
import csv
import subprocess
import sys

rep, kern, skip = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "-k", "regex:" + kern,
                      "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file = ""
lines = []
hdr = None
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < 8 or r[0] in ("", "Function Name"):
        continue
    try:
        ln = int(r[0])
        samples = int(r[hdr.index("# Samples")])
        inst = int(r[hdr.index("Instructions Executed")])
    except ValueError:
        continue
    lines.append((samples, inst, cur_file, ln, r[1].strip()[:110]))
tot_s = sum(x[0] for x in lines) or 1
tot_i = sum(x[1] for x in lines) or 1
print("total samples %d, warp instructions %d" % (tot_s, tot_i))
for s, i, f, ln, src in sorted(lines, reverse=True)[:top]:
    print("%5.1f%% smp %5.1f%% inst  %s:%d  %s" % (100.0 * s / tot_s, 100.0 * i / tot_i, f, ln, src))
