"""209-PU mode (PIC_ALL_DEPTH_MODE pictures: enc-mode 0-2 reference pictures) of the CUDA path through the C-ABI,
bit-exact against the oracle: open_loop_me_fullpel_search_sblock + ExtSadCalculation (85 square + 124 non-square PUs,
quirks Q10 / Q11), half/quarter-pel and bi-prediction for all 209 PUs. Needs a B200."""
import numpy as np
import pytest

import me_common as M

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def meo():
    M.build_oracle(ref=False)
    return M.load_oracle()


def _params(pkg, w, h, nl, tl, over=None, num_pus=209, is_ref=1):
    p = pkg.default_params(w, h, nl, tl, 3, is_ref)
    p.num_pus = num_pus
    for k, v in (over or {}).items():
        setattr(p, k, v)
    return p


def _run(pkg, meo, frames, specs, sb_rows=None):
    """specs: list of ((cur, ref0, ref1), SvtMeParams). Returns nothing; asserts bit-exactness per job."""
    h, w = frames[0].shape
    ctx = pkg.MeContext(w, h, len(frames))
    try:
        for i, f in enumerate(frames):
            ctx.upload(i, f)
        res, dbg = ctx.motion_estimate([(c, r0, r1, p) for (c, r0, r1), p in specs], want_debug=True, sb_rows=sb_rows)
    finally:
        ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in frames]
    sb_cols = (w + 63) // 64
    for j, ((c, r0, r1), p) in enumerate(specs):
        nl = p.num_lists
        want, wdbg = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[c], pyr[r0], pyr[r1] if nl == 2 else None)
        if sb_rows is not None:
            want, wdbg = want[sb_rows[0] * sb_cols:sb_rows[1] * sb_cols], wdbg[sb_rows[0] * sb_cols:sb_rows[1] * sb_cols]
        assert res[j].shape == want.shape, (j, res[j].shape, want.shape)
        for f in ("sa_origin_x", "sa_origin_y", "sa_width", "sa_height"):
            assert np.array_equal(dbg[j][f][:, :nl], wdbg[f][:, :nl]), (j, f)
        assert np.array_equal(dbg[j]["best_sad"][:, :nl], wdbg["best_sad"][:, :nl]), j       # square PUs after sub-pel
        assert np.array_equal(dbg[j]["best_mv"][:, :nl], wdbg["best_mv"][:, :nl]), j
        assert M.results_diff(res[j], want, nl) == [], j


@pytest.mark.parametrize("w,h", [(416, 240), (64, 64), (160, 72), (320, 136)])
@pytest.mark.parametrize("nl,tl", [(2, 1), (1, 0), (2, 3)])
def test_sizes_and_layers(pkg, meo, w, h, nl, tl):
    fr = M.synth_video(w, h, 3, seed=2 * w + h + nl + tl, max_shift=30)
    _run(pkg, meo, fr, [((1, 0, 2), _params(pkg, w, h, nl, tl))])


@pytest.mark.parametrize("over", [
    {"use_subpel": 0},                                     # the full-pel kernel alone: results = integer search
    {"search_area_w": 1, "search_area_h": 1}, {"search_area_w": 7, "search_area_h": 5},
    {"search_area_w": 27, "search_area_h": 13, "use_subpel": 0}, {"search_area_w": 43, "search_area_h": 64},
    {"search_area_w": 16, "search_area_h": 16}, {"search_area_w": 33, "search_area_h": 2},
    {"enable_hme_flag": 0}, {"cu8x8_mode": 1}, {"same_ref_poc": 1}, {"is_used_as_reference": 0},
])
def test_parameter_sweep(pkg, meo, over):
    fr = M.synth_video(320, 192, 3, seed=len(str(over)), max_shift=40)
    _run(pkg, meo, fr, [((1, 0, 2), _params(pkg, 320, 192, 2, 1, over))])


@pytest.mark.parametrize("synth", [
    dict(noise=0.0, contrast=0.0),                  # flat: every SAD ties, pure raster-order tie-break (and the Q11 replay)
    dict(noise=0.0, contrast=0.05, max_shift=4),
    dict(noise=12.0, contrast=4.0, max_shift=60),   # large differences: wrapping SSD (Q5) on tall 8-wide PUs (Q10)
    dict(noise=1.0, contrast=1.0, max_shift=90),
])
def test_content_extremes(pkg, meo, synth):
    fr = M.synth_video(256, 192, 3, seed=7, **synth)
    _run(pkg, meo, fr, [((1, 0, 2), _params(pkg, 256, 192, 2, 1))])


def test_mixed_batch_of_square_and_ext_jobs(pkg, meo):
    """enc-mode 1/2: reference pictures use 209 PUs, non-reference pictures 85 -- one batch holds both."""
    w, h = 256, 128
    fr = M.synth_video(w, h, 5, seed=31, max_shift=20)
    specs = [((2, 0, 4), _params(pkg, w, h, 2, 1)), ((1, 0, 2), _params(pkg, w, h, 2, 2, num_pus=85, is_ref=0)),
             ((3, 2, 4), _params(pkg, w, h, 2, 2, num_pus=85, is_ref=0)), ((4, 0, 0), _params(pkg, w, h, 1, 0))]
    _run(pkg, meo, fr, specs)


def test_sb_row_range(pkg, meo):
    w, h = 192, 256
    fr = M.synth_video(w, h, 3, seed=41, max_shift=20)
    _run(pkg, meo, fr, [((1, 0, 2), _params(pkg, w, h, 2, 1))], sb_rows=(1, 3))


def test_720p_picture(pkg, meo):
    w, h = 1280, 720
    fr = M.synth_video(w, h, 3, seed=51, max_shift=40)
    _run(pkg, meo, fr, [((1, 0, 2), _params(pkg, w, h, 2, 1))])
