"""Dynamic SASS opcode mix per kernel of an .ncu-rep (source page, --print-source sass): which instructions the warps
actually executed. usage: python profiles/ncu_opmix.py <rep> [top]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 14
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
kern, hdr = None, None
agg = collections.OrderedDict()
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "Kernel Name":
        kern = r[1].split("(")[0].replace("svtme::", "").replace("void ", "")
    elif r and r[0] == "Address":
        hdr = r
    elif hdr and kern and len(r) >= len(hdr) - 2 and r[0].startswith("0x"):
        src, n = r[hdr.index("Source")].split(), int(r[hdr.index("Instructions Executed")])
        op = (src[1] if src[0].startswith("@") else src[0]).split(".")[0]
        agg.setdefault(kern, collections.Counter())[op] += n
for k, c in agg.items():
    t = sum(c.values())
    print(f"{k}: {t} warp instructions: " + ", ".join(f"{o} {100 * n / t:.1f}%" for o, n in c.most_common(top)))
