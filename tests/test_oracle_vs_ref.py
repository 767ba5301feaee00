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


@pytest.mark.parametrize("w,h", [(80, 64), (72, 64), (88, 128), (112, 64), (144, 128), (200, 72), (720, 480), (848, 480), (1368, 128)])
@pytest.mark.parametrize("nl,tl", [(2, 1), (1, 0), (2, 0)])
def test_edge_sb_widths(libs, w, h, nl, tl):
    """Edge SBs 8 / 16 / 24 / 48 wide (every width % 64 the reference has a live NxM kernel for, EbComputeSAD.h:139-150):
    hme_mv_center_check on the narrow block, the HME levels on 2..24-wide decimated blocks (C SadLoopKernel for the widths
    the AVX2 kernel has no case for; quirks Q12 and Q13 in the 4- and 24-wide AVX2 cases)."""
    for seed in range(2 if w * h > 100000 else 4):
        _compare(libs, w, h, seed=w + h + nl + 17 * seed, nl=nl, tl=tl, synth=dict(max_shift=30))


@pytest.mark.parametrize("w", [72, 80, 88, 112])
def test_edge_sb_widths_saturating_content(libs, w):
    _compare(libs, w, 128, seed=w, nl=2, tl=1, synth=dict(noise=12.0, contrast=4.0, max_shift=60))
    _compare(libs, w, 128, seed=w + 1, nl=2, tl=1, over={"num_pus": 209}, synth=dict(noise=12.0, contrast=4.0, max_shift=60))


def test_widths_without_a_live_kernel_are_rejected(libs):
    """40- and 56-wide edge SBs: the reference's hme_mv_center_check calls NxMSadKernelVoidFunc (EbComputeSAD.h:143-147),
    its result is undefined; the oracle refuses them instead of inventing one."""
    meo, _ = libs
    for w in (104, 120):
        fr = M.synth_video(w, 64, 3, seed=1)
        po = [M.build_pyramid(meo, "meo", f) for f in fr]
        p = M.default_params(meo, "meo", w, 64)
        with pytest.raises(Exception):
            M.motion_estimate(meo, "meo", p, po[1], po[0], po[2])


@pytest.mark.parametrize("over", [
    {"search_area_w": 65, "search_area_h": 65}, {"search_area_w": 96, "search_area_h": 80},
    {"search_area_w": 127, "search_area_h": 127}, {"search_area_w": 127, "search_area_h": 33},
    {"search_area_w": 71, "search_area_h": 127}, {"search_area_w": 200, "search_area_h": 255},
    {"search_area_w": 121, "search_area_h": 3},
])
@pytest.mark.parametrize("pus", [85, 209])
def test_wide_search_areas(libs, over, pus):
    """MotionEstimateLcu searches MIN(search_area, 127) (ME.c:6644-6645); the context fields are uint8_t
    (EbMotionEstimationContext.h:403-404), so 255 is the largest value the reference can hold."""
    _compare(libs, 320, 192, seed=len(str(over)) + pus, nl=2, tl=1, over=dict(over, num_pus=pus), synth=dict(max_shift=70))
    if pus == 85:
        _compare(libs, 208, 136, seed=len(str(over)), nl=1, tl=0, over=over, synth=dict(max_shift=50))


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


@pytest.mark.parametrize("w", [256, 112, 80, 72, 88])
def test_black_white_checker(libs, w):
    """0/255 checkerboards: the worst case for the 16-bit saturating partial sums (quirks Q3, Q13)."""
    meo, ref = libs
    h = 128
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


@pytest.mark.parametrize("w,h", [(416, 240), (256, 192), (64, 64), (200, 72), (1368, 128), (720, 480)])
def test_zz_sad_matches_reference(libs, w, h):
    """ComputeDecimatedZzSad (EbMotionEstimationProcess.c:219-349): zero-motion SAD classes of every SB, complete and
    incomplete, for still, slowly and quickly moving content (all five cost classes occur)."""
    meo, ref = libs
    seen = set()
    for seed, synth in enumerate([dict(noise=0.0, max_shift=0), dict(noise=0.6, max_shift=1), dict(noise=1.5, max_shift=3),
                                  dict(noise=3.0, max_shift=24), dict(noise=12.0, contrast=4.0, max_shift=60)]):
        fr = M.synth_video(w, h, 2, seed=seed + w, **synth)
        if seed == 0:
            fr[1] = fr[0].copy()                                   # a still pair: class 0 everywhere
        po = [M.build_pyramid(meo, "meo", f) for f in fr]
        zo, no = M.zz_sad(meo, "meo", w, h, po[1], po[0])
        zr, nr = M.zz_sad(ref, "ref", w, h, po[1], po[0])
        assert np.array_equal(zo, zr) and np.array_equal(no, nr)
        seen |= set(zo.tolist())
    if w >= 256 and w % 64 == 0:
        assert {0, 30} <= seen
    if w % 64 or h % 64:
        assert 255 in seen                                         # INVALID_ZZ_COST of the incomplete SBs
