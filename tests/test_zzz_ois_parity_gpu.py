"""svtme_ois (include/svtme.h; kernel svt-av1-1_b200/csrc/ois.cu) against the oracle's OpenLoopIntraSearchLcu
(oracle/ois_oracle.c, pinned to the compiled reference in test_ois_oracle_vs_ref.py), through the C-ABI, bit-exact on the byte
images of OisCandidate_t and on total_intra_luma_mode. Every path of the search in one batch: intra pictures, layer-0 inter
pictures (35 modes, best 18 sorted), the other layers (OIS points steered by the ME SADs), 4K threshold sets, DC only,
8x8 CUs on and off, incomplete SBs at the right / bottom edge.

Collected after the other GPU files on purpose."""
import ctypes as C

import numpy as np
import pytest

import me_common as M
from test_ois_emul import MeoOisParams, assert_same, me_sads, ois_cases, oracle_search

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("w,h", [(416, 240), (200, 136), (64, 64), (1920, 1080)])
def test_open_loop_intra_search_matches_oracle(pkg, w, h):
    M.build_oracle(ref=False)
    meo = M.load_oracle()
    synths = [dict(), dict(noise=10.0, contrast=3.0)] if w < 1920 else [dict()]
    frames = [M.synth_video(w, h, 1, seed=w + 3 * k, **s)[0] for k, s in enumerate(synths)]
    n_sb = ((w + 63) // 64) * ((h + 63) // 64)
    rng = np.random.default_rng(w + h)
    cases = ois_cases() if w < 1920 else [(0, 0, 1, 2, 0), (2, 0, 1, 2, 0), (2, 2, 0, 2, 0)]
    specs, want = [], []
    for slot, f in enumerate(frames):
        pyr = M.build_pyramid(meo, "meo", f)                    # owns the padded planes: keep it alive while the oracle reads them
        pl = pyr.c.full
        for i, (st, tl, is_ref, res, dc_only) in enumerate(cases):
            cu8, enc = (i + slot) & 1, 3 if (tl + i) % 2 else 1
            me = me_sads(rng, n_sb)
            specs.append((slot, pkg.SvtMeOisParams(st, tl, is_ref, cu8, res, dc_only, enc), me if (st or i & 1) else None))
            want.append(oracle_search(meo, MeoOisParams(w, h, st, tl, is_ref, cu8, res, dc_only, enc), pl, me))
    ctx = pkg.MeContext(w, h, len(frames))
    try:
        for slot, f in enumerate(frames):
            ctx.upload(slot, f)
        cand, total = ctx.ois(specs)
        cand2, total2 = ctx.ois(specs[1:3])                      # buffers reused, fewer jobs
        with pytest.raises(pkg.SvtMeError):
            ctx.ois([(len(frames), specs[0][1], None)])          # slot out of range
        with pytest.raises(pkg.SvtMeError):
            ctx.ois([(0, pkg.SvtMeOisParams(2, 2, 0, 0, 1, 0, 1), None)])   # this path reads the ME SADs
    finally:
        ctx.close()
    for k, out in enumerate(want):
        assert_same(out, cand[k], total[k], (k, specs[k][0]))
    assert np.array_equal(cand2, cand[1:3]) and np.array_equal(total2, total[1:3])
    assert not cand[:, :, 0].any() and not total[:, :, 0].any()   # the 64x64 CU is never searched
    if w % 64 or h % 64:
        assert (total[0] == 0).sum() > n_sb                         # CUs outside the picture stay empty


def test_extreme_content_matches_oracle(pkg):
    """Black, white, checkerboards, stripes, full-range noise: zero SADs (OIS point 0), ties in the best-18 selection and the swap
    sort, the boundary smoothing clipping both ways. One batch of 7 pictures x 4 path settings."""
    from test_ois_oracle_vs_ref import extreme_pictures
    M.build_oracle(ref=False)
    meo = M.load_oracle()
    w, h = 200, 136
    pics = extreme_pictures(w, h)
    rng = np.random.default_rng(11)
    specs, want, names, keep = [], [], [], []
    for slot, (name, f) in enumerate(pics.items()):
        pyr = M.build_pyramid(meo, "meo", f)
        keep.append(pyr)
        for st, tl, res in ((0, 0, 1), (2, 0, 1), (2, 2, 1), (2, 3, 3)):
            me = rng.integers(0, 20000, size=(12, 85)).astype(np.uint32)
            q = (st, tl, 1 if tl < 3 else 0, 0, res, 0, 2)
            specs.append((slot, pkg.SvtMeOisParams(*q), me))
            want.append(oracle_search(meo, MeoOisParams(w, h, *q), pyr.c.full, me))
            names.append((name, st, tl, res))
    ctx = pkg.MeContext(w, h, len(pics))
    try:
        for slot, f in enumerate(pics.values()):
            ctx.upload(slot, f)
        cand, total = ctx.ois(specs)
    finally:
        ctx.close()
    for k, out in enumerate(want):
        assert_same(out, cand[k], total[k], names[k])
