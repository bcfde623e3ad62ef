"""Top stall-sample SASS instructions of one kernel from an ncu report (source page).
Usage: ncu_stalls.py <rep> <kernel-substring> [top]"""
import csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
his = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
hi = his[0]
end = his[1] if len(his) > 1 else len(rows)
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:end] if len(r) > 10 and r[idx["# Samples"]].isdigit()]
stall_cols = [h for h in hdr if h.startswith("stall_")]
def iv(r, h):
    try:
        return int(r[idx[h]] or 0)
    except ValueError:
        return 0
tot = sum(iv(r, "# Samples") for r in data)
print("total samples", tot, "instructions", len(data))
agg = {h: sum(iv(r, h) for r in data) for h in stall_cols}
print([(k, v) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:10]])
for r in sorted(data, key=lambda r: -iv(r, "# Samples"))[:top_n]:
    st = sorted(((h, iv(r, h)) for h in stall_cols if iv(r, h)), key=lambda x: -x[1])[:3]
    print(r[idx["Address"]][-5:], r[idx["# Samples"]].rjust(6), r[idx["Instructions Executed"]].rjust(10), r[idx["Avg. Threads Executed"]].rjust(4),
          r[idx["Source"]][:58].ljust(58), st)
