"""bench.py's `ois` key on a short sequence (17 pictures = two mini-GOPs, 1080p): motion estimation through the C-ABI, the rows
into a pinned tensor laid out like the bench's, then bench.measure_ois -- prints what the bench line would carry."""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
n_pics = 17
jobs = bench.gop_jobs(n_pics)
frames = bench.synth_sequence(n_pics, seed=3)
ctx = pkg.MeContext(bench.W, bench.H, n_pics)
for i in range(n_pics):
    ctx.upload(i, frames[i])
specs = []
for poc, tl, r0, r1 in jobs:
    p = pkg.default_params(bench.W, bench.H, 2, tl, 3, 1 if tl < 3 else 0)
    p.same_ref_poc = 1 if r0 == r1 else 0
    specs.append((poc, r0, r1, p))
results, _ = ctx.motion_estimate(specs)
out = torch.from_numpy(results.view(np.uint8).reshape(len(jobs), ctx.n_sb, 85, 40).copy()).pin_memory()
print(json.dumps(bench.measure_ois(pkg, ctx, jobs, out)))
print(json.dumps({k: v for k, v in ctx.kernel_times().items()}))
ctx.close()
