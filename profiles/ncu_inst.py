"""Instruction share per source line of one kernel of an .ncu-rep (complements ncu_lines.py, which ranks by stall samples).
usage: python profiles/ncu_inst.py <rep> <kernel-name substring> [top]"""
import csv
import subprocess
import sys

rep, want = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr = fn = fpath = None
agg, tot = {}, 0
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r and r[0] == "Function Name":
        fn = r[1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit() and fn and want in fn:
        n = int(r[hdr.index("Instructions Executed")])
        a = agg.setdefault((fpath, int(r[0])), [0, r[1].strip()[:100]])
        a[0] += n
        tot += n
print(f"kernel *{want}*: {tot} warp instructions")
for (f, l), (n, src) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f}:{l:4d} {100 * n / tot:5.1f}%  | {src}")
