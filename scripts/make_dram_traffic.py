"""Writes profiles/<tag>_dram_traffic.json (DRAM bytes per launch of every kernel) from an `ncu --set full` report.
bench.py reads the newest such file for roofline.traffic.  Usage: make_dram_traffic.py <tag> <rep> "<what was profiled>" """
import csv, io, json, os, subprocess, sys
tag, rep, what = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "KB": 1e3, "MB": 1e6, "GB": 1e9, "B": 1.0}
tscale = {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3, "nsecond": 1e-6, "usecond": 1e-3, "msecond": 1.0, "second": 1e3}
def val(r, k, table):
    return float(r[idx[k]].replace(",", "")) * table.get(units[idx[k]], 1.0)
out = {"source": "%s (ncu --set full --clock-control none), %s" % (os.path.basename(rep), what), "kernels": {}}
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].split("::")[-1].split("<")[0]
    if name in out["kernels"]:
        continue  # first launch of each kernel
    rd, wr = val(r, "dram__bytes_read.sum", scale), val(r, "dram__bytes_write.sum", scale)
    out["kernels"][name] = {"dram_bytes_read": rd, "dram_bytes_write": wr, "dram_bytes": rd + wr,
                            "time_ms_under_ncu": val(r, "gpu__time_duration.sum", tscale)}
path = os.path.join(root, "profiles", tag + "_dram_traffic.json")
json.dump(out, open(path, "w"), indent=1)
print(path, {k: round(v["dram_bytes"] / 1e9, 3) for k, v in out["kernels"].items()})
