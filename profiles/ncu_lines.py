"""Per-source-line hot spots of one kernel of an .ncu-rep (needs -lineinfo and --import-source on).
usage: python profiles/ncu_lines.py <rep> <kernel-name substring> [top]
Stall samples of every launch of the matching kernel are summed per (file, line)."""
import csv
import subprocess
import sys

rep, want = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
hdr, fn, fpath, agg = None, None, None, {}
stalls = ("stall_barrier", "stall_long_sb", "stall_short_sb", "stall_math", "stall_wait", "stall_not_selected", "stall_selected",
          "stall_mio", "stall_dispatch", "stall_branch_resolving", "stall_no_inst", "stall_lg")
tot_st = dict.fromkeys(stalls, 0)
for r in csv.reader(txt.splitlines()):
    if r and r[0] == "File Path":
        fpath = r[1].split("/")[-1]
    elif r and r[0] == "Function Name":
        fn = r[1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit() and fn and want in fn:
        S, I = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Instructions Executed")
        a = agg.setdefault((fpath, int(r[0])), [0, 0, r[1].strip()[:100], dict.fromkeys(stalls, 0)])
        a[0] += int(r[S]); a[1] += int(r[I])
        for n in stalls:
            v = int(r[hdr.index(n)]); a[3][n] += v; tot_st[n] += v
ts, ti = sum(v[0] for v in agg.values()), sum(v[1] for v in agg.values())
print(f"kernel *{want}*: samples {ts} instructions {ti}")
print("stall reasons: " + ", ".join(f"{n[6:]} {100 * v / max(ts, 1):.1f}%" for n, v in sorted(tot_st.items(), key=lambda kv: -kv[1]) if v))
for (f, l), (s, i, t, st) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    main = max(st.items(), key=lambda kv: kv[1])
    print(f"{f}:{l:>4} samp {100 * s / max(ts, 1):5.1f}% inst {100 * i / max(ti, 1):5.1f}% top-stall {main[0][6:]:<14} | {t}")
