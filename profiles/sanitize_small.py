"""compute-sanitizer target: a small mixed batch (85- and 209-PU pictures, B and P, HME on/off) through the C-ABI, checked
against the oracle, plus one transform/quantisation call. Run as
    compute-sanitizer --tool {memcheck,racecheck,synccheck,initcheck} python profiles/sanitize_small.py
(In round 1 compute-sanitizer answered "closed on this pool"; see profiles/r02/sanitizer.md for round 2.)"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import me_common as M  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
meo = M.load_oracle()
w, h = 224, 136                                  # 4 x 3 SBs: a 32-wide last column and an 8-high last row
fr = M.synth_video(w, h, 4, seed=5, max_shift=24)


def prm(nl, tl, **over):
    p = pkg.default_params(w, h, nl, tl, 3, 1)
    for k, v in over.items():
        setattr(p, k, v)
    return p


specs = [(1, 0, 2, prm(2, 1)), (2, 0, 3, prm(2, 2, num_pus=209)), (3, 0, 0, prm(1, 0)), (1, 0, 2, prm(2, 3, num_pus=209, enable_hme_flag=0)),
         (2, 1, 3, prm(2, 1, search_area_w=21, search_area_h=9, use_subpel=0)),
         (1, 0, 2, prm(2, 1, search_area_w=127, search_area_h=90)),                   # Wide instantiations (sub-areas, 198-row sub-pel planes)
         (2, 1, 3, prm(2, 2, search_area_w=71, search_area_h=127, num_pus=209))]
ctx = pkg.MeContext(w, h, 4)
for i, f in enumerate(fr):
    ctx.upload(i, f)
res, dbg = ctx.motion_estimate(specs, want_debug=True)
ctx.close()
pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
for j, (c, r0, r1, p) in enumerate(specs):
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[c], pyr[r0], pyr[r1] if p.num_lists == 2 else None)
    assert M.results_diff(res[j], want, p.num_lists) == [], j
# edge SBs 16 and 24 pixels wide (any-width HME search), the allocation-free entry point, SB-granular ranges
for w2 in (208, 88):
    fr2 = M.synth_video(w2, 128, 3, seed=w2, max_shift=16)
    c2 = pkg.MeContext(w2, 128, 3)
    for i, f in enumerate(fr2):
        c2.upload(i, f)
    p2 = pkg.default_params(w2, 128, 2, 1, 3, 1)
    a, _ = c2.motion_estimate([(1, 0, 2, p2)])
    b, _ = c2.motion_estimate_static((1, 0, 2, p2))
    n2 = a[0].shape[0]
    parts = [c2.motion_estimate([(1, 0, 2, p2)], sb_range=r)[0][0] for r in ((0, 1), (1, n2))]
    c2.close()
    py2 = [M.build_pyramid(meo, "meo", f) for f in fr2]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p2), py2[1], py2[0], py2[2])
    assert M.results_diff(a[0], want) == [] and a[0].tobytes() == b[0].tobytes() == np.concatenate(parts).tobytes(), w2
print("me ok")

# ---- transform + quantisation: every size once (type 0; 64-point sizes with and without the energy output)
import txq_common as T  # noqa: E402
import txq_golden as G  # noqa: E402

q10 = next(iter(G.blocks()))["q10"]
txo = T.oracle()
rng = np.random.default_rng(3)
with pkg.TxqContext(0) as txq:
    for (tw, th) in T.TX_SIZES:
        res_ = T.synth_residual(rng, 2 * th, 5 * tw, 8, "mixed")
        got = txq.fwd_txfm_quant(res_, tw, th, 0, q10)
        want = T.txq_plane(txo, res_, tw, th, 0, q10)
        for k in ("coeff", "qcoeff", "dqcoeff", "eob"):
            assert np.array_equal(got[k], want[k]), (tw, th, k)
        if max(tw, th) == 64:
            txq.fwd_txfm_quant(res_, tw, th, 0, q10, want=("qcoeff", "dqcoeff", "eob"))
    for (tw, th) in ((4, 4), (16, 8), (64, 64)):                                  # compact return path (prefix sum + gather)
        res_ = T.synth_residual(rng, 4 * th, 6 * tw, 8, "mixed")
        eob, off, lv = txq.fwd_txfm_quant_compact(res_, tw, th, 0, q10)
        want = T.txq_plane(txo, res_, tw, th, 0, q10)
        assert np.array_equal(eob, want["eob"]) and off[-1] == len(lv) == int(eob.astype(np.int64).sum()), (tw, th)
print("txq ok")
