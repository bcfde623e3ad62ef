"""Per-kernel SASS summary of the built library (no GPU needed): instruction count, memory / dot-product mnemonics, and the
lines around the widest global stores.  Usage: python scripts/sass_summary.py > profiles/<tag>_sass.md"""
import collections, os, re, subprocess, sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.environ.get("B200JPEG_LIB") or os.path.join(root, "nvimagecodec_b200", "lib", "libb200jpeg_ext.so")
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout.split("\n")
kern, cur = collections.OrderedDict(), None
for l in txt:
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = re.sub(r".*kernels_cu_[0-9a-f]+\d\d", "", m.group(1))
        cur = re.sub(r"ENS_8BatchDev.*|EPh.*", "", cur)
        kern[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m and cur:
        kern[cur].append(m.group(2).strip())
print("# SASS summary of `libb200jpeg_ext.so` (`cuobjdump -sass`, sm_100a; `scripts/sass_summary.py`)\n")
print("| kernel | instructions | global stores | global loads | shared | other |")
print("|---|---|---|---|---|---|")
for k, ins in kern.items():
    ops = collections.Counter(re.sub(r"^@!?U?P\d+\s+", "", i).split()[0] for i in ins)
    def pick(pref):
        return ", ".join("%s x%d" % (o, n) for o, n in sorted(ops.items()) if o.startswith(pref)) or "-"
    other = ", ".join("%s x%d" % (o, n) for o, n in sorted(ops.items()) if o.startswith(("IDP", "ATOM", "RED", "BAR", "LDL", "STL", "LDGSTS")))
    print("| `%s` | %d | %s | %s | %s | %s |" % (k, len(ins), pick("STG") + ("; " + pick("ST.E") if pick("ST.E") != "-" else ""), pick("LDG"), (pick("STS") + "; " + pick("LDS")), other or "-"))
print("\n## The widest global stores, in context\n")
for k, ins in kern.items():
    idx = [i for i, x in enumerate(ins) if "STG.E.128" in x]
    if not idx:
        continue
    print("### `%s` (%d x STG.E.128)\n\n```" % (k, len(idx)))
    i = idx[0]
    for x in ins[max(0, i - 6):i + 4]:
        print(x)
    print("```\n")
