"""Per-source-line dynamic instruction counts for one kernel: joins `ncu --page source` (SASS rows) with
nvdisasm -g line info of the cubin inside the built .so.  Usage: ncu_lines.py <rep> <kernel-substring> [top]"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.environ.get("B200JPEG_LIB") or os.path.join(root, "nvimagecodec_b200", "lib", "libb200jpeg_ext.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "kernels.sm_100a.cubin", so], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.split("\n")
start = [i for i, l in enumerate(sass) if l.strip().startswith(".text.") and kern in l][0]
nxt = [i for i, l in enumerate(sass) if i > start and l.strip().startswith(".text.")]
end = nxt[0] if nxt else len(sass)
cur, seq = None, []
for l in sass[start:end]:
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        seq.append((int(m.group(1), 16), cur, m.group(2)))
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern] + os.environ.get("NCU_ARGS", "").split(), capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hi = [i for i, r in enumerate(rows) if "Instructions Executed" in r][0]
hdr = rows[hi]
idx = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) > idx["Avg. Threads Executed"]]
data = data[:len(seq)]  # the report may hold several launches; the first one lines up with the SASS listing


def f(x):
    try:
        return float(x)
    except Exception:
        return 0.0


by = collections.defaultdict(lambda: [0.0, 0.0, 0.0])
tot = 0.0
for (addr, line, txt), r in zip(seq, data):
    ex, th, sm = f(r[idx["Instructions Executed"]]), f(r[idx["Thread Instructions Executed"]]), f(r[idx["# Samples"]])
    by[line][0] += ex
    by[line][1] += th
    by[line][2] += sm
    tot += ex
srcdir = os.environ.get("B200JPEG_SRC") or os.path.join(root, "nvimagecodec_b200", "csrc")
files = {}


def text(key):
    if not key:
        return ""
    fn, ln = key
    if fn not in files:
        try:
            files[fn] = open(os.path.join(srcdir, fn)).read().split("\n")
        except Exception:
            files[fn] = []
    return "%s:%d %s" % (fn, ln, files[fn][ln - 1].strip()[:90] if ln - 1 < len(files[fn]) else "")
print("sass instrs", len(seq), "ncu rows", len(data), "warp instrs", int(tot))
for line, (ex, th, sm) in sorted(by.items(), key=lambda kv: -kv[1][0])[:top]:
    print("ex=%5.2f%% lanes=%4.1f samples=%6d | %s" % (100 * ex / tot, th / ex if ex else 0, sm, text(line)))
