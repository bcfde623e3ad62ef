"""Writes profiles/<tag>_kernels.md from an ncu report (and optionally a launch-list csv).
Run here (ncu is installed, no GPU needed):  python scripts/make_profile_summary.py <tag> <rep> [launches.csv]"""
import collections
import csv
import io
import os
import subprocess
import sys

tag, rep = sys.argv[1], sys.argv[2]
launches = sys.argv[3] if len(sys.argv) > 3 else None
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
COLS = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst"), ("smsp__inst_executed.sum", "warp insts"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"), ("launch__registers_per_thread", "regs"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long_sb"),
        ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "stall barrier"),
        ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "stall math_pipe"),
        ("l1tex__t_sector_hit_rate.pct", "L1 hit %")]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
out = ["# ncu summary `%s`  (source: %s, `ncu --set full --clock-control none`)" % (tag, os.path.basename(rep)), ""]
out.append("| kernel | grid x block | " + " | ".join(c[1] for c in COLS) + " |")
out.append("|---|---|" + "---|" * len(COLS))
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].split("::")[-1]
    cells = []
    for k, _ in COLS:
        if k in idx:
            v = r[idx[k]]
            try:
                fv = float(v.replace(",", ""))
                v = ("%.3f" % fv).rstrip("0").rstrip(".") if fv < 1e6 else "%.3g" % fv
            except Exception:
                pass
            cells.append(v + " " + units[idx[k]].replace("byte", "B"))
        else:
            cells.append("-")
    out.append("| %s | %s x %s | %s |" % (name, r[idx["Grid Size"]], r[idx["Block Size"]], " | ".join(cells)))
if launches:
    rows = list(csv.reader(l for l in open(launches) if l.startswith('"')))
    h = rows[0]
    i_name, i_val = h.index("Kernel Name"), h.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[i_val].replace(",", ""))
        except Exception:
            continue
        a = agg.setdefault(r[i_name].split("(")[0].split("::")[-1], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    out += ["", "## launch list (`ncu --metrics gpu__time_duration.sum`, cold-cache serialised: compare shares)", "",
            "| kernel | launches | avg us | share |", "|---|---|---|---|"]
    for n, (c, t) in agg.items():
        out.append("| %s | %d | %.1f | %.3f |" % (n, c, t / 1e3 / c, t / tot))
path = os.path.join(root, "profiles", tag + "_kernels.md")
open(path, "w").write("\n".join(out) + "\n")
print("wrote", path)
