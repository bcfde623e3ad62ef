"""Executed-instruction mix (by SASS opcode) of one kernel from an ncu report.  Usage: ncu_opmix.py <rep> <kernel-substring>"""
import collections, csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
his = [i for i, r in enumerate(rows) if "Instructions Executed" in r]
hi, end = his[0], (his[1] if len(his) > 1 else len(rows))
hdr = rows[hi]; idx = {h: i for i, h in enumerate(hdr)}
mix = collections.Counter(); tot = 0
for r in rows[hi + 1:end]:
    if len(r) <= idx["Instructions Executed"] or not r[idx["Instructions Executed"]].isdigit():
        continue
    n = int(r[idx["Instructions Executed"]])
    src = r[idx["Source"]].strip()
    parts = src.split()
    op = parts[1] if parts and parts[0].startswith("@") and len(parts) > 1 else (parts[0] if parts else "?")
    mix[op.split(".")[0]] += n; tot += n
print("total warp instructions", tot)
for op, n in mix.most_common(25):
    print("%-12s %12d  %5.1f%%" % (op, n, 100.0 * n / tot))
