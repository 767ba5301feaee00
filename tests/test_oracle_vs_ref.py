"""Pins the plain-C restatement (oracle/me_oracle.c) bit-exactly against the reference's own
objects compiled from /root/reference (oracle/_ref/libsvtref.so) on seeded inputs and a sweep of
search parameters. CPU only; skipped where neither the reference tree nor a prebuilt _ref exists."""
import numpy as np
import pytest

import me_common as M


@pytest.fixture(scope="module")
def libs():
    M.build_oracle(ref=True)
    if not M.ref_available():
        pytest.skip("oracle/_ref/libsvtref.so not available (no /root/reference here)")
    return M.load_oracle(), M.load_ref()


def _compare(libs, w, h, seed, nl, tl, is_ref=1, over=None, synth=None, hl=3):
    meo, ref = libs
    fr = M.synth_video(w, h, 3, seed=seed, **(synth or {}))
    po = [M.build_pyramid(meo, "meo", f) for f in fr]
    pr = [M.build_pyramid(ref, "ref", f) for f in fr]
    for a, b in zip(po, pr):
        for x, y in zip(a.planes, b.planes):
            assert np.array_equal(x, y)
    p = M.default_params(meo, "meo", w, h, num_lists=nl, tl=tl, hl=hl, is_ref=is_ref)
    p2 = M.default_params(ref, "ref", w, h, num_lists=nl, tl=tl, hl=hl, is_ref=is_ref)
    assert p.as_tuple() == p2.as_tuple()
    for k, v in (over or {}).items():
        if isinstance(v, tuple):
            for i, x in enumerate(v):
                getattr(p, k)[i] = x
        else:
            setattr(p, k, v)
    ro, do = M.motion_estimate(meo, "meo", p, po[1], po[0], po[2] if nl == 2 else None)
    rr, dr = M.motion_estimate(ref, "ref", p, pr[1], pr[0], pr[2] if nl == 2 else None)
    assert M.results_diff(ro, rr, nl) == []
    assert np.array_equal(do["best_sad"][:, :nl], dr["best_sad"][:, :nl])
    assert np.array_equal(do["best_mv"][:, :nl], dr["best_mv"][:, :nl])
    assert np.array_equal(do["sa_origin_x"][:, :nl], dr["sa_origin_x"][:, :nl])
    assert np.array_equal(do["sa_origin_y"][:, :nl], dr["sa_origin_y"][:, :nl])


@pytest.mark.parametrize("w,h", [(416, 240), (256, 192), (64, 64), (160, 72), (320, 136)])
@pytest.mark.parametrize("nl,tl", [(2, 1), (1, 0), (2, 3), (2, 0)])
def test_sizes_and_layers(libs, w, h, nl, tl):
    _compare(libs, w, h, seed=w + h + nl + tl, nl=nl, tl=tl)


@pytest.mark.parametrize("over", [
    {"search_area_w": 24, "search_area_h": 17}, {"search_area_w": 7, "search_area_h": 5},
    {"search_area_w": 43, "search_area_h": 64}, {"search_area_w": 1, "search_area_h": 1},
    {"enable_hme_flag": 0}, {"enable_hme_level2": 0}, {"enable_hme_level1": 0, "enable_hme_level2": 0},
    {"enable_hme_level0": 0}, {"num_hme_regions_w": 1, "num_hme_regions_h": 1},
    {"num_hme_regions_w": 2, "num_hme_regions_h": 1}, {"same_ref_poc": 1}, {"use_subpel": 0},
    {"cu8x8_mode": 1}, {"hme_l1_w": (12, 20), "hme_l1_h": (9, 16)}, {"hme_l2_w": (5, 11), "hme_l2_h": (3, 8)},
])
def test_parameter_sweep(libs, over):
    _compare(libs, 320, 192, seed=len(str(over)), nl=2, tl=1, over=over, synth=dict(max_shift=40))


@pytest.mark.parametrize("w,h,nl,tl,over", [
    (256, 192, 2, 1, {}), (128, 128, 1, 0, {}), (160, 72, 2, 2, {}), (320, 136, 2, 0, {"same_ref_poc": 1}),
    (128, 128, 2, 1, {"use_subpel": 0}), (256, 128, 2, 1, {"cu8x8_mode": 1}), (192, 128, 2, 3, {"search_area_w": 21, "search_area_h": 9}),
    (192, 128, 2, 1, {"enable_hme_flag": 0}),
])
def test_ext_209_pu_mode(libs, w, h, nl, tl, over):
    """PIC_ALL_DEPTH_MODE pictures (enc-mode 0-2 reference pictures): open_loop_me_fullpel_search_sblock and the 124
    non-square PUs (quirks Q10, Q11), half/quarter-pel and bi-pred for all 209 PUs."""
    _compare(libs, w, h, seed=3 * w + h + nl, nl=nl, tl=tl, over=dict(over, num_pus=209), synth=dict(max_shift=30))


@pytest.mark.parametrize("synth", [dict(noise=0.0, contrast=0.0), dict(noise=12.0, contrast=4.0, max_shift=60)])
def test_ext_209_content_extremes(libs, synth):
    _compare(libs, 192, 128, seed=5, nl=2, tl=1, over={"num_pus": 209}, synth=synth)


@pytest.mark.parametrize("synth", [
    dict(noise=0.0, contrast=0.0),                  # flat picture: every SAD ties, pure tie-break order
    dict(noise=0.0, contrast=0.05, max_shift=4),    # near-flat
    dict(noise=12.0, contrast=4.0, max_shift=60),   # saturating differences (wrapping-SSD quirk, 16-bit quirk)
    dict(noise=1.0, contrast=1.0, max_shift=90),    # motion beyond the search range
])
def test_content_extremes(libs, synth):
    _compare(libs, 256, 192, seed=7, nl=2, tl=1, synth=synth)
    _compare(libs, 256, 192, seed=8, nl=2, tl=2, is_ref=0, synth=synth)


def test_black_white_checker(libs):
    """0/255 checkerboards: the worst case for the 16-bit saturating partial sums (quirk Q3)."""
    meo, ref = libs
    w, h = 256, 128
    yy, xx = np.mgrid[0:h, 0:w]
    a = (((xx // 3 + yy // 5) & 1) * 255).astype(np.uint8)
    b = (((xx // 3 + yy // 5 + 1) & 1) * 255).astype(np.uint8)
    c = np.roll(a, (3, 5), (0, 1))
    for lib, pre in ((meo, "meo"), (ref, "ref")):
        pass
    po = [M.build_pyramid(meo, "meo", f) for f in (b, a, c)]
    pr = [M.build_pyramid(ref, "ref", f) for f in (b, a, c)]
    p = M.default_params(meo, "meo", w, h, num_lists=2, tl=1)
    ro, do = M.motion_estimate(meo, "meo", p, po[1], po[0], po[2])
    rr, dr = M.motion_estimate(ref, "ref", p, pr[1], pr[0], pr[2])
    assert M.results_diff(ro, rr) == []
    assert np.array_equal(do["best_sad"], dr["best_sad"]) and np.array_equal(do["best_mv"], dr["best_mv"])


@pytest.mark.parametrize("hl,tl", [(3, 0), (4, 0), (4, 1), (5, 0), (5, 1), (0, 0), (2, 2)])
def test_hme_level0_multipliers(libs, hl, tl):
    """HME_LEVEL_0_SEARCH_AREA_MULTIPLIER_X/Y[hierarchical_levels][temporal_layer] (EbDefinitions.h:2980-2996):
    up to 5.25x wider level-0 areas, clipped at the picture borders."""
    _compare(libs, 448, 256, seed=hl * 7 + tl, nl=2, tl=tl, hl=hl, synth=dict(max_shift=70))


@pytest.mark.parametrize("w,h", [(64, 64), (32, 16), (416, 240), (1920, 34), (128, 8), (96, 6)])
def test_unpack16_matches_reference(libs, w, h):
    """10-bit input split (UnPack2D): oracle restatement against the reference's SSE2 kernel."""
    meo, ref = libs
    rng = np.random.default_rng(w + h)
    v = rng.integers(0, 1024, size=(h, w), dtype=np.uint16)
    v[0, :8] = [0, 1, 2, 3, 1020, 1021, 1022, 1023]
    a, b = M.unpack16(meo, "meo", v), M.unpack16(ref, "ref", v)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert np.array_equal(a[0], (v >> 2).astype(np.uint8))
    if w not in (32, 64):                                        # quirk Q9 lives in those two special cases only
        assert np.array_equal(a[1], ((v & 3) << 6).astype(np.uint8))
