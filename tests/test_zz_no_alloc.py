"""svtme_motion_estimate_static (include/svtme.h): the allocation-free variant of the one-picture ME call -- job table,
SB scratch, result rows and debug rows in a file-scope device array of the library -- against the existing entry
point svtme_motion_estimate on the same pictures, through the C-ABI, bit-exact (integer work: tolerance 0), on two
calls in a row that share the array (the second one a different picture pair, P instead of B, a row range).

Collected last on purpose: passes alone in a fresh process and after the others."""
import ctypes as C

import numpy as np
import pytest

import me_common as M

pytestmark = pytest.mark.gpu

FIELDS = ("sa_origin_x", "sa_origin_y", "sa_width", "sa_height", "search_center_x", "search_center_y", "best_sad", "best_mv")


def _same(a, b, da, db):
    assert a.tobytes() == b.tobytes()                    # whole 40-byte rows, padding included
    for f in FIELDS:
        assert np.array_equal(da[f], db[f]), f


def test_static_scratch_matches_batched_entry_point(pkg):
    w, h = 416, 240
    frames = M.synth_video(w, h, 4, seed=41)
    ctx = pkg.MeContext(w, h, 4)
    try:
        for i, f in enumerate(frames):
            ctx.upload(i, f)
        b = pkg.default_params(w, h, 2, 1, 3, 1)
        p = pkg.default_params(w, h, 1, 0, 3, 1)
        want_b, wdbg_b = ctx.motion_estimate([(1, 0, 2, b)], want_debug=True)
        want_p, wdbg_p = ctx.motion_estimate([(3, 2, None, p)], want_debug=True)
        want_rows, wdbg_rows = ctx.motion_estimate([(2, 1, 3, b)], want_debug=True, sb_rows=(1, 3))
        # two calls in a row on the shared array, then a third over a row range, then the first again
        got_b, dbg_b = ctx.motion_estimate_static((1, 0, 2, b), want_debug=True)
        got_p, dbg_p = ctx.motion_estimate_static((3, 2, None, p), want_debug=True)
        _same(got_b[0], want_b[0], dbg_b[0], wdbg_b[0])
        _same(got_p[0], want_p[0], dbg_p[0], wdbg_p[0])
        assert not got_p[0]["xMvL1"].any() and (got_p[0]["total"] == 1).all()
        got_rows, dbg_rows = ctx.motion_estimate_static((2, 1, 3, b), want_debug=True, sb_rows=(1, 3))
        _same(got_rows[0], want_rows[0], dbg_rows[0], wdbg_rows[0])
        again, _ = ctx.motion_estimate_static((1, 0, 2, b))
        assert again[0].tobytes() == want_b[0].tobytes()
        # the batched entry point is unaffected by the calls in between
        after, _ = ctx.motion_estimate([(1, 0, 2, b)])
        assert after[0].tobytes() == want_b[0].tobytes()
    finally:
        ctx.close()


def test_static_scratch_rejects_what_does_not_fit(pkg):
    w, h = 1920, 1088                                    # 510 SBs > SVTME_STATIC_MAX_SBS
    ctx = pkg.MeContext(w, h, 3)
    try:
        with pytest.raises(pkg.SvtMeError, match="SVTME_STATIC_MAX_SBS"):
            ctx.motion_estimate_static((1, 0, 2, pkg.default_params(w, h)))
    finally:
        ctx.close()
    ctx = pkg.MeContext(256, 128, 3)
    try:
        prm = pkg.default_params(256, 128)
        prm.num_pus = 209
        with pytest.raises(pkg.SvtMeError, match="85-PU"):
            ctx.motion_estimate_static((1, 0, 2, prm))
        with pytest.raises(pkg.SvtMeError, match="slot out of range"):
            ctx.motion_estimate_static((1, 0, 7, pkg.default_params(256, 128)))
    finally:
        ctx.close()
