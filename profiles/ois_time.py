"""Time svtme_ois on one 1080p picture, one launch per search path (intra picture / layer-0 inter: all 35 modes / layer-2
inter: OIS points). Wall clock of the synchronous call (upload of the ME SADs, kernel, 3.1 MB of results back); the kernel's
own duration comes from the ncu launch list of the same command (profiles/r02/ois_launches.csv)."""
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import me_common as M  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
w, h = 1920, 1080
f = M.synth_video(w, h, 1, seed=7)[0]
ctx = pkg.MeContext(w, h, 1)
ctx.upload(0, f)
n_sb = ctx.n_sb
me = (np.random.default_rng(1).integers(0, 4000, size=(n_sb, 85))).astype(np.uint32)
for name, prm in (("intra", pkg.SvtMeOisParams(0, 0, 1, 0, 2, 0, 1)), ("layer0_all35", pkg.SvtMeOisParams(2, 0, 1, 0, 2, 0, 1)),
                  ("layer2_points", pkg.SvtMeOisParams(2, 2, 0, 0, 2, 0, 1))):
    for rep in range(3):
        t = time.perf_counter()
        cand, total = ctx.ois([(0, prm, me)])
        dt = (time.perf_counter() - t) * 1e3
    t = time.perf_counter()
    ctx.ois([(0, prm, me)] * 8)                                   # one launch, grid (510, 8)
    dt8 = (time.perf_counter() - t) * 1e3 / 8
    print(f"{name}: {dt:.3f} ms per picture (synchronous call), {dt8:.3f} in a batch of 8; total histogram {np.bincount(total.ravel(), minlength=19).tolist()}")
ctx.close()
