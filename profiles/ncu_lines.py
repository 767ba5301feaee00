"""Per-source-line hot spots of one kernel of an .ncu-rep (needs -lineinfo and --import-source on).
usage: python profiles/ncu_lines.py <rep> <kernel-id, e.g. :::5> [top]"""
import csv
import subprocess
import sys

rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", kid, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, hdr, out = None, None, []
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr is not None and len(r) == len(hdr) and r[0].isdigit():
        out.append((cur, r))


def col(name):
    return hdr.index(name)


def num(v):
    try:
        return int(v)
    except ValueError:
        return 0


S, I, W, X = col("Warp Stall Sampling (All Samples)"), col("Instructions Executed"), col("L1 Wavefronts Shared"), col("L1 Wavefronts Shared Excessive")
ts, ti = sum(num(r[S]) for _, r in out), sum(num(r[I]) for _, r in out)
print(f"samples {ts} instructions {ti} shared wavefronts {sum(num(r[W]) for _, r in out)} excessive {sum(num(r[X]) for _, r in out)}")
for f, r in sorted(out, key=lambda fr: -num(fr[1][S]))[:top]:
    print(f"{f}:{r[0]:>4} samp {100 * num(r[S]) / max(ts, 1):5.1f}% inst {100 * num(r[I]) / max(ti, 1):5.1f}% wf {num(r[W]):>9} exc {num(r[X]):>9} | {r[1].strip()[:100]}")
