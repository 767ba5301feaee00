"""Summarise ncu outputs brought back in gpurun_out/ into small tracked files under profiles/<round>/.
usage: python profiles/summarize.py <launches.csv> <prof.ncu-rep> <out_prefix>"""
import collections
import csv
import subprocess
import sys

launch_csv, rep, out = sys.argv[1:4]
rows = [r for r in csv.reader(l for l in open(launch_csv) if l.startswith('"'))]
hdr = rows[0]
ik, iv, iu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    v = v / 1e3 if r[iu] in ("ns", "nsecond") else v * 1e3 if r[iu] in ("ms", "msecond") else v   # -> us
    name = r[ik].split("(")[0]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
tot = sum(a[1] for a in agg.values())
with open(out + "_launches.md", "w") as f:
    f.write(f"ncu launch list ({launch_csv}): per-kernel totals; cold-cache, serialised -> compare SHARES\n\n")
    f.write("| kernel | launches | total us | share |\n|---|---|---|---|\n")
    for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        f.write(f"| {k} | {n} | {us:.1f} | {100 * us / tot:.1f}% |\n")
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h, u = rr[0], rr[1]
want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_active.avg"]
with open(out + "_metrics.md", "w") as f:
    f.write(f"ncu --set full ({rep}), selected raw metrics per captured launch\n\n")
    for r in rr[2:]:
        f.write("\n```\n")
        for w in want:
            if w in h:
                i = h.index(w)
                f.write(f"{w:75s} {r[i]} {u[i]}\n")
        f.write("```\n")
print("wrote", out + "_launches.md", out + "_metrics.md")
