"""Prints SASS of one kernel with the decoded control fields (stall, yield, write/read barrier, wait mask).
Usage: sass_ctrl.py <kernel-substring> [lib]"""
import re, subprocess, sys, os
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
kern = sys.argv[1]
so = sys.argv[2] if len(sys.argv) > 2 else os.path.join(root, "nvimagecodec_b200", "lib", "libb200jpeg_ext.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout.split("\n")
on = False
pend = None
for l in out:
    if "Function :" in l:
        on = kern in l
        continue
    if not on:
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* 0x([0-9a-f]{16}) \*/", l)
    if m:
        pend = (m.group(1), m.group(2))
        continue
    m = re.match(r"\s+/\* 0x([0-9a-f]{16}) \*/", l)
    if m and pend:
        hi = int(m.group(1), 16)
        ctrl = hi >> 41
        stall = ctrl & 15
        yld = (ctrl >> 4) & 1
        wb = (ctrl >> 5) & 7
        rb = (ctrl >> 8) & 7
        wait = (ctrl >> 11) & 63
        print("%s  st%-2d %s wb%s rb%s wait%-6s  %s" % (pend[0], stall, "Y" if yld else "-", wb if wb != 7 else "-", rb if rb != 7 else "-",
                                                     "".join(str(i) for i in range(6) if wait >> i & 1) or "-", pend[1].strip()[:90]))
        pend = None
