"""Debug helper: run one golden ME case on the GPU and list the PUs whose results differ."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import me_common as M  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
name = sys.argv[1]
g = M.load_golden(name)
prm = M.params_from_list(pkg.SvtMeParams, g["params"])
ctx = pkg.MeContext(prm.width, prm.height, 3)
for i in range(3):
    ctx.upload(i, np.ascontiguousarray(g["frames"][i]))
res, dbg = ctx.motion_estimate([(1, 0, 2, prm)], want_debug=True)
ctx.close()
nl = prm.num_lists
print("params", M.params_to_list(prm))
for sb in range(dbg[0].shape[0]):
    d = dbg[0][sb]
    print("sb", sb, "sa", d["sa_origin_x"][:nl], d["sa_origin_y"][:nl], d["sa_width"][:nl], d["sa_height"][:nl])
    for l in range(nl):
        bad = np.nonzero((d["best_sad"][l] != g["best_sad"][sb][l]) | (d["best_mv"][l] != g["best_mv"][sb][l]))[0]
        for pu in bad[:40]:
            mv, emv = int(d["best_mv"][l][pu]), int(g["best_mv"][sb][l][pu])
            f = lambda v: ((((v & 0xFFFF) ^ 0x8000) - 0x8000), ((((v >> 16) & 0xFFFF) ^ 0x8000) - 0x8000))
            print("  list", l, "pu", pu, "sad", d["best_sad"][l][pu], "want", g["best_sad"][sb][l][pu], "mv", f(mv), "want", f(emv))
