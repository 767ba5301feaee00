"""Profiling driver (ncu target): one 1080p mini-GOP (8 B pictures, 9 uploads) through the C-ABI, twice.
Small on purpose: `ncu --set full` replays every kernel ~40 times."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
frames = bench.synth_sequence(9, seed=1)
ctx = pkg.MeContext(bench.W, bench.H, 9)
specs = []
for poc, tl, r0, r1 in [j for j in bench.gop_jobs() if j[0] <= 8]:
    p = pkg.default_params(bench.W, bench.H, 2, tl, 3, 1 if tl < 3 else 0)
    p.same_ref_poc = 1 if r0 == r1 else 0
    p.num_pus = int(os.environ.get("SVTME_PROF_PUS", "85"))     # 209: every picture in the 209-PU mode
    specs.append((poc, r0, r1, p))
for rep in range(int(os.environ.get("SVTME_PROF_REPS", "2"))):
    for i in range(9):
        ctx.upload(i, frames[i])
    res, _ = ctx.motion_estimate(specs)
print("ok", int(res["dist0"].sum()), ctx.kernel_launches)
ctx.close()
