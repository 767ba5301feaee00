"""Parity tests proper: the CUDA path, called through the C-ABI (include/svtme.h), against the oracle
and the committed golden vectors, bit-exact (all of this is integer work). Needs a B200."""
import ctypes as C

import numpy as np
import pytest

import me_common as M

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def meo():
    M.build_oracle(ref=False)
    return M.load_oracle()


def _set(p, over):
    for k, v in (over or {}).items():
        if isinstance(v, tuple):
            for i, x in enumerate(v):
                getattr(p, k)[i] = x
        else:
            setattr(p, k, v)


def _check_against_oracle(pkg, meo, frames, prm_list, triples):
    """frames: list of luma arrays; prm_list[i] / triples[i] = SvtMeParams and (cur, ref0, ref1) per job."""
    h, w = frames[0].shape
    ctx = pkg.MeContext(w, h, len(frames))
    try:
        for i, f in enumerate(frames):
            ctx.upload(i, f)
        res, dbg = ctx.motion_estimate([(c, r0, r1, p) for (c, r0, r1), p in zip(triples, prm_list)], want_debug=True)
    finally:
        ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in frames]
    for j, ((c, r0, r1), p) in enumerate(zip(triples, prm_list)):
        want, wdbg = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[c], pyr[r0],
                                       pyr[r1] if p.num_lists == 2 else None)
        nl = p.num_lists
        for f in ("sa_origin_x", "sa_origin_y", "sa_width", "sa_height", "search_center_x", "search_center_y"):
            assert np.array_equal(dbg[j][f][:, :nl], wdbg[f][:, :nl]), (j, f)
        assert np.array_equal(dbg[j]["best_mv"][:, :nl], wdbg["best_mv"][:, :nl]), j
        assert np.array_equal(dbg[j]["best_sad"][:, :nl], wdbg["best_sad"][:, :nl]), j
        assert M.results_diff(res[j], want, nl) == [], j
        if nl == 1:
            assert not res[j]["xMvL1"].any() and not res[j]["yMvL1"].any()


@pytest.mark.parametrize("w,h", [(416, 240), (64, 64), (160, 72), (1280, 720)])
def test_pyramid_parity(pkg, meo, w, h):
    f = M.synth_video(w, h, 1, seed=w)[0]
    ctx = pkg.MeContext(w, h, 1)
    ctx.upload(0, f)
    got = [ctx.download_plane(0, l) for l in range(3)]
    ctx.close()
    want = M.build_pyramid(meo, "meo", f)
    for g, x in zip(got, want.planes):
        assert np.array_equal(g, x)


def test_pyramid_strided_and_device_upload(pkg, meo):
    import torch
    w, h = 320, 136
    f = M.synth_video(w, h, 1, seed=9)[0]
    wide = np.zeros((h, w + 40), dtype=np.uint8)
    wide[:, :w] = f
    ctx = pkg.MeContext(w, h, 2)
    ctx.upload_ptr(0, wide.ctypes.data, wide.strides[0])
    d = torch.from_numpy(wide).cuda()
    torch.cuda.synchronize()
    ctx.upload_device(1, d.data_ptr(), d.stride(0))
    a = [ctx.download_plane(0, l) for l in range(3)]
    b = [ctx.download_plane(1, l) for l in range(3)]
    ctx.close()
    want = M.build_pyramid(meo, "meo", f)
    for x, y, z in zip(a, b, want.planes):
        assert np.array_equal(x, z) and np.array_equal(y, z)


def test_batched_device_upload(pkg, meo):
    """svtme_pictures_upload_device: one pyramid launch for many pictures gives the planes of the per-picture calls."""
    import torch
    w, h = 416, 240
    fr = M.synth_video(w, h, 5, seed=77)
    wide = np.zeros((5, h, w + 24), dtype=np.uint8)
    wide[:, :, :w] = np.stack(fr)
    d = torch.from_numpy(wide).cuda()
    torch.cuda.synchronize()
    ctx = pkg.MeContext(w, h, 6)
    ctx.upload_device_batch([4, 0, 2, 5, 1], [d[i].data_ptr() for i in range(5)], d.stride(1))
    for slot, i in zip([4, 0, 2, 5, 1], range(5)):
        want = M.build_pyramid(meo, "meo", fr[i])
        for lvl, x in enumerate(want.planes):
            assert np.array_equal(ctx.download_plane(slot, lvl), x), (slot, lvl)
    with pytest.raises(pkg.SvtMeError, match="0x80001005"):
        ctx.upload_device_batch([0, 9], [d[0].data_ptr(), d[1].data_ptr()], d.stride(1))
    # the same from host memory (svtme_pictures_upload), strided and contiguous
    ctx.upload_ptr_batch([1, 3, 0], [wide[i].ctypes.data for i in (2, 4, 1)], wide.strides[1])
    cont = [np.ascontiguousarray(fr[i]) for i in (0, 3)]
    ctx.upload_ptr_batch([5, 2], [c.ctypes.data for c in cont], w)
    for slot, i in ((1, 2), (3, 4), (0, 1), (5, 0), (2, 3)):
        want = M.build_pyramid(meo, "meo", fr[i])
        for lvl, x in enumerate(want.planes):
            assert np.array_equal(ctx.download_plane(slot, lvl), x), (slot, lvl)
    ctx.close()


@pytest.mark.parametrize("name", M.golden_cases())
def test_cuda_matches_reference_golden(pkg, name):
    """CUDA path against vectors produced by the unmodified reference objects."""
    g = M.load_golden(name)
    prm = M.params_from_list(pkg.SvtMeParams, g["params"])
    frames = g["frames"]
    ctx = pkg.MeContext(prm.width, prm.height, 3)
    for i in range(3):
        ctx.upload(i, np.ascontiguousarray(frames[i]))
        assert [M.plane_crc(ctx.download_plane(i, l)) for l in range(3)] == [int(v) for v in g["pyr_crc"][i]]
    res, dbg = ctx.motion_estimate([(1, 0, 2, prm)], want_debug=True)
    ctx.close()
    nl = prm.num_lists
    assert np.array_equal(dbg[0]["best_sad"][:, :nl], g["best_sad"][:, :nl])
    assert np.array_equal(dbg[0]["best_mv"][:, :nl], g["best_mv"][:, :nl])
    assert M.results_diff(res[0], g["results"], nl) == []


@pytest.mark.parametrize("w,h", [(416, 240), (256, 192), (64, 64), (160, 72), (320, 136)])
@pytest.mark.parametrize("nl,tl", [(2, 1), (1, 0), (2, 3), (2, 0)])
def test_sizes_and_layers(pkg, meo, w, h, nl, tl):
    fr = M.synth_video(w, h, 3, seed=w + h + nl + tl)
    _check_against_oracle(pkg, meo, fr, [pkg.default_params(w, h, nl, tl, 3, 1)], [(1, 0, 2)])


@pytest.mark.parametrize("w,h", [(80, 64), (72, 64), (88, 128), (112, 64), (144, 128), (200, 72), (720, 480), (848, 480), (1368, 128)])
@pytest.mark.parametrize("nl,tl,pus", [(2, 1, 85), (1, 0, 85), (2, 0, 209)])
def test_edge_sb_widths(pkg, meo, w, h, nl, tl, pus):
    """Edge SBs 8 / 16 / 24 / 48 pixels wide (EbComputeSAD.h:139-150): centre check and HME levels on the narrow blocks
    (any-width search of hme.cu, quirk Q12), full-pel / sub-pel / bi-pred on the padded 64x64 SB like the reference."""
    for seed in range(2):
        fr = M.synth_video(w, h, 3, seed=w + h + nl + 17 * seed, max_shift=30)
        p = pkg.default_params(w, h, nl, tl, 3, 1)
        p.num_pus = pus
        if pus == 85:
            _check_against_oracle(pkg, meo, fr, [p], [(1, 0, 2)])
        else:
            _check_results_against_oracle(pkg, meo, fr, p)


def _check_results_against_oracle(pkg, meo, fr, p):
    h, w = fr[0].shape
    ctx = pkg.MeContext(w, h, 3)
    try:
        for i, f in enumerate(fr):
            ctx.upload(i, f)
        res, _ = ctx.motion_estimate([(1, 0, 2, p)])
    finally:
        ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[1], pyr[0], pyr[2] if p.num_lists == 2 else None)
    assert M.results_diff(res[0], want, p.num_lists) == []


def test_widths_without_a_reference_kernel_are_rejected(pkg):
    for w in (104, 120, 1960):
        with pytest.raises(pkg.SvtMeError, match="0x80001005"):
            pkg.MeContext(w, 64, 2)


WIDE = [
    {"search_area_w": 65, "search_area_h": 65}, {"search_area_w": 96, "search_area_h": 80},
    {"search_area_w": 127, "search_area_h": 127}, {"search_area_w": 127, "search_area_h": 33},
    {"search_area_w": 71, "search_area_h": 127}, {"search_area_w": 200, "search_area_h": 255},
    {"search_area_w": 121, "search_area_h": 3}, {"search_area_w": 16, "search_area_h": 100},
]


@pytest.mark.parametrize("over", WIDE)
@pytest.mark.parametrize("pus", [85, 209])
def test_wide_search_areas(pkg, meo, over, pus):
    """Search areas beyond 64 x 64 (the reference searches MIN(search_area, 127), ME.c:6644-6645): the Wide instantiations
    of the full-pel kernels (sub-areas merged in the reference's visiting order, quirks Q4 and Q11 across sub-areas) and of
    the sub-pel kernel (198-row planes)."""
    w, h = 320, 192
    fr = M.synth_video(w, h, 3, seed=len(str(over)) + pus, max_shift=70)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    _set(p, over)
    p.num_pus = pus
    if pus == 85:
        _check_against_oracle(pkg, meo, fr, [p], [(1, 0, 2)])
        q = pkg.default_params(208, 136, 1, 0, 3, 1)
        _set(q, over)
        _check_against_oracle(pkg, meo, M.synth_video(208, 136, 3, seed=len(str(over)), max_shift=50), [q], [(1, 0, 2)])
    else:
        _check_results_against_oracle(pkg, meo, fr, p)


@pytest.mark.parametrize("synth", [dict(noise=0.0, contrast=0.0), dict(noise=12.0, contrast=4.0, max_shift=60)])
def test_wide_search_content_extremes(pkg, meo, synth):
    """Flat pictures: every SAD ties, so the merge of the sub-areas must reproduce the pure visiting order."""
    w, h = 256, 192
    fr = M.synth_video(w, h, 3, seed=11, **synth)
    for pus in (85, 209):
        p = pkg.default_params(w, h, 2, 1, 3, 1)
        p.search_area_w, p.search_area_h, p.num_pus = 127, 100, pus
        _check_results_against_oracle(pkg, meo, fr, p)


def test_mixed_narrow_and_wide_batch(pkg, meo):
    """One batch holding narrow / wide-area and 85 / 209-PU pictures: four kernel classes over the same grid."""
    w, h = 256, 128
    fr = M.synth_video(w, h, 4, seed=3, max_shift=40)
    prms = []
    for sa, pus in ((64, 85), (100, 85), (127, 209), (48, 209)):
        p = pkg.default_params(w, h, 2, 1, 3, 1)
        p.search_area_w = p.search_area_h = sa
        p.num_pus = pus
        prms.append(p)
    ctx = pkg.MeContext(w, h, 4)
    try:
        for i, f in enumerate(fr):
            ctx.upload(i, f)
        res, _ = ctx.motion_estimate([(1, 0, 2, prms[0]), (2, 1, 3, prms[1]), (1, 0, 3, prms[2]), (2, 0, 3, prms[3])])
    finally:
        ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    for r, p, (c, a, b) in zip(res, prms, ((1, 0, 2), (2, 1, 3), (1, 0, 3), (2, 0, 3))):
        want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[c], pyr[a], pyr[b])
        assert M.results_diff(r, want) == []


SWEEP = [
    {"search_area_w": 24, "search_area_h": 17}, {"search_area_w": 7, "search_area_h": 5},
    {"search_area_w": 43, "search_area_h": 64}, {"search_area_w": 1, "search_area_h": 1},
    {"search_area_w": 9, "search_area_h": 3}, {"search_area_w": 16, "search_area_h": 16},
    {"enable_hme_flag": 0}, {"enable_hme_level2": 0}, {"enable_hme_level1": 0, "enable_hme_level2": 0},
    {"enable_hme_level0": 0}, {"enable_hme_level1": 0}, {"num_hme_regions_w": 1, "num_hme_regions_h": 1},
    {"num_hme_regions_w": 2, "num_hme_regions_h": 1}, {"same_ref_poc": 1}, {"use_subpel": 0},
    {"cu8x8_mode": 1}, {"hme_l1_w": (12, 20), "hme_l1_h": (9, 16)}, {"hme_l2_w": (5, 11), "hme_l2_h": (3, 8)},
    {"is_used_as_reference": 0},
]


@pytest.mark.parametrize("over", SWEEP)
def test_parameter_sweep(pkg, meo, over):
    w, h = 320, 192
    fr = M.synth_video(w, h, 3, seed=len(str(over)), max_shift=40)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    _set(p, over)
    _check_against_oracle(pkg, meo, fr, [p], [(1, 0, 2)])


@pytest.mark.parametrize("hl,tl", [(3, 0), (4, 0), (4, 1), (5, 0), (5, 1), (0, 0), (2, 2)])
def test_hme_level0_multipliers(pkg, meo, hl, tl):
    w, h = 448, 256
    fr = M.synth_video(w, h, 3, seed=hl * 7 + tl, max_shift=70)
    _check_against_oracle(pkg, meo, fr, [pkg.default_params(w, h, 2, tl, hl, 1)], [(1, 0, 2)])


@pytest.mark.parametrize("synth", [
    dict(noise=0.0, contrast=0.0), dict(noise=0.0, contrast=0.05, max_shift=4),
    dict(noise=12.0, contrast=4.0, max_shift=60), dict(noise=1.0, contrast=1.0, max_shift=90),
])
def test_content_extremes(pkg, meo, synth):
    w, h = 256, 192
    fr = M.synth_video(w, h, 3, seed=7, **synth)
    _check_against_oracle(pkg, meo, fr, [pkg.default_params(w, h, 2, 1, 3, 1), pkg.default_params(w, h, 2, 2, 3, 0)],
                          [(1, 0, 2), (1, 2, 0)])


def test_black_white_checker(pkg, meo):
    w, h = 256, 128
    yy, xx = np.mgrid[0:h, 0:w]
    a = (((xx // 3 + yy // 5) & 1) * 255).astype(np.uint8)
    b = (((xx // 3 + yy // 5 + 1) & 1) * 255).astype(np.uint8)
    c = np.roll(a, (3, 5), (0, 1))
    _check_against_oracle(pkg, meo, [b, a, c], [pkg.default_params(w, h, 2, 1, 3, 1)], [(1, 0, 2)])


def test_mixed_batch(pkg, meo):
    """One call, several pictures with different slice types / layers / references (a mini-GOP)."""
    w, h = 320, 192
    fr = M.synth_video(w, h, 9, seed=21, max_shift=16)
    jobs, prms = [], []
    # hierarchical mini-GOP 8: (poc, temporal layer, ref0, ref1), EbPredictionStructure.c:183-240
    for poc, tl, r0, r1 in [(8, 0, 0, 0), (4, 1, 0, 8), (2, 2, 0, 4), (6, 2, 4, 8), (1, 3, 0, 2), (3, 3, 2, 4), (5, 3, 4, 6), (7, 3, 6, 8)]:
        p = pkg.default_params(w, h, 2, tl, 3, 1 if tl < 3 else 0)
        p.same_ref_poc = 1 if r0 == r1 else 0
        jobs.append((poc, r0, r1)); prms.append(p)
    p = pkg.default_params(w, h, 1, 0, 3, 1)
    jobs.append((8, 0, 0)); prms.append(p)
    _check_against_oracle(pkg, meo, fr, prms, jobs)


@pytest.mark.gpu
@pytest.mark.parametrize("chunks", ["1", "3", "9"])
def test_chunked_submission(pkg, meo, chunks, monkeypatch):
    """SVTME_CHUNKS: the pictures of a batch run as 1 / 3 / 9 chunks on separate streams (api.cu); results do not change."""
    monkeypatch.setenv("SVTME_CHUNKS", chunks)
    test_mixed_batch(pkg, meo)


def test_1080p_full_size_parity(pkg, meo):
    """BASELINE config 2 geometry (1920x1080: 30x17 SBs, last SB row 56 high -> no HME there)."""
    w, h = 1920, 1080
    fr = M.synth_video(w, h, 3, seed=1080, max_shift=40)
    _check_against_oracle(pkg, meo, fr, [pkg.default_params(w, h, 2, 1, 3, 1)], [(1, 0, 2)])


def test_1080p_properties(pkg):
    """Size-independent properties at full size: (1) a picture predicted from itself has zero SAD and zero
    MVs for every PU; (2) a pure translation inside the search range is recovered by every interior SB's
    64x64 PU with zero distortion; (3) results do not depend on batch position (idempotence)."""
    w, h = 1920, 1080
    rng = np.random.default_rng(4)
    big = rng.integers(0, 256, size=(h + 64, w + 64), dtype=np.uint8)
    cur = np.ascontiguousarray(big[32:32 + h, 32:32 + w])
    dx, dy = 12, -8            # multiples of 4: the point-sampled HME levels see the same translation
    ref = np.ascontiguousarray(big[32 + dy:32 + dy + h, 32 + dx:32 + dx + w])       # ref(x,y) = cur(x+dx, y+dy)
    ctx = pkg.MeContext(w, h, 2)
    ctx.upload(0, cur); ctx.upload(1, ref)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    res, _ = ctx.motion_estimate([(0, 0, 0, p), (0, 1, 1, p), (0, 0, 0, p)])
    ctx.close()
    assert np.array_equal(res[0], res[2])
    sq = [0] + list(range(5, 85))      # the 32x32 PUs carry the reference's gcc-only MV quirk (Q4): their MV is off by +-4 px
    full = res[0].reshape(17, 30, 85)[:-1][:, :, sq]     # last SB row is 56 high: its bottom PUs lie in replicated padding (ties)
    assert not res[0]["dist0"][:, sq].any() and not full["xMvL0"].any() and not full["yMvL0"].any()
    r = res[1].reshape(17, 30, 85)[1:-1, 1:-1, 0]                                      # interior SBs, 64x64 PU
    assert not r["dist0"].any()
    assert (r["xMvL0"] == -dx * 4).all() and (r["yMvL0"] == -dy * 4).all()


def test_error_codes(pkg):
    w, h = 128, 128
    ctx = pkg.MeContext(w, h, 2)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    p.num_pus = 100                      # only 85 and 209 exist
    with pytest.raises(pkg.SvtMeError, match="0x80001005"):
        ctx.motion_estimate([(0, 1, 1, p)])
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    with pytest.raises(pkg.SvtMeError, match="0x80001005"):
        ctx.motion_estimate([(0, 5, 1, p)])
    p.search_area_w = 256                # the reference holds the area in a uint8_t: 255 is its largest value
    with pytest.raises(pkg.SvtMeError, match="0x80001005"):
        ctx.motion_estimate([(0, 1, 1, p)])
    ctx.close()


def test_pipelined_batches_and_slot_reuse(pkg, meo):
    """Several batches in flight on the three streams (upload / search / result copy), more batches than ring
    entries, and slots re-uploaded while earlier batches still read their previous pictures."""
    w, h = 320, 192
    fr = M.synth_video(w, h, 12, seed=33, max_shift=16)
    ctx = pkg.MeContext(w, h, 3)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    outs, keep = [], []
    for b in range(5):                                   # batch b uses pictures 2b, 2b+1, 2b+2 in slots 0..2
        for sl in range(3):
            ctx.upload(sl, fr[2 * b + sl])
        jobs, res, dbg = ctx.make_jobs([(1, 0, 2, p), (0, 1, 1, p)], want_debug=True)
        ctx.motion_estimate_async(jobs)
        outs.append((res, dbg)); keep.append(jobs)
    ctx.fetch(keep[-1])
    ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    mp = M.convert_params(M.MeoParams, p)
    for b, (res, dbg) in enumerate(outs):
        want0, _ = M.motion_estimate(meo, "meo", mp, pyr[2 * b + 1], pyr[2 * b], pyr[2 * b + 2])
        want1, _ = M.motion_estimate(meo, "meo", mp, pyr[2 * b], pyr[2 * b + 1], pyr[2 * b + 1])
        assert M.results_diff(res[0], want0) == [], b
        assert M.results_diff(res[1], want1) == [], b


def test_upload_16bit_matches_oracle_split(pkg, meo):
    """10-bit input: the device-side split (msb = v >> 2 feeds ME, rem = v << 6) and the pyramid built from it."""
    w, h = 416, 240
    rng = np.random.default_rng(10)
    f8 = M.synth_video(w, h, 1, seed=31)[0]
    v = (f8.astype(np.uint16) << 2) | rng.integers(0, 4, size=(h, w), dtype=np.uint16)
    big = np.zeros((h, w + 24), np.uint16)
    big[:, 8:8 + w] = v
    strided = big[:, 8:8 + w]                                   # stride 440 samples
    msb, rem = M.unpack16(meo, "meo", v)
    assert np.array_equal(msb, f8)
    got_rem = np.zeros((h, w), np.uint8)
    ctx = pkg.MeContext(w, h, 2)
    ctx.upload16(0, strided, got_rem)
    ctx.upload16(1, np.ascontiguousarray(v))                    # without the remainder plane
    ctx.synchronize()
    want = M.build_pyramid(meo, "meo", msb)
    for lvl in range(3):
        assert np.array_equal(ctx.download_plane(0, lvl), want.planes[lvl]), lvl
        assert np.array_equal(ctx.download_plane(1, lvl), want.planes[lvl]), lvl
    ctx.close()
    assert np.array_equal(got_rem, rem)


def test_sb_row_ranges_equal_whole_picture(pkg, meo):
    """SB-row sharding: searching row ranges separately returns exactly the rows of the whole-picture call."""
    w, h = 416, 240                                             # 7 x 4 SBs, last row 48 high
    fr = M.synth_video(w, h, 3, seed=77, max_shift=20)
    ctx = pkg.MeContext(w, h, 3)
    for i, f in enumerate(fr):
        ctx.upload(i, f)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    whole, wdbg = ctx.motion_estimate([(1, 0, 2, p)], want_debug=True)
    parts, dparts = [], []
    for rows in [(0, 1), (1, 3), (3, 4)]:
        r, d = ctx.motion_estimate([(1, 0, 2, p), (1, 0, 2, p)], want_debug=True, sb_rows=rows)
        assert r.shape[1] == (rows[1] - rows[0]) * 7 and np.array_equal(r[0], r[1])
        parts.append(r[0]); dparts.append(d[0])
    assert np.array_equal(np.concatenate(parts), whole[0])
    assert np.array_equal(np.concatenate(dparts), wdbg[0])
    with pytest.raises(pkg.SvtMeError):
        ctx.motion_estimate([(1, 0, 2, p)], sb_rows=(2, 5))
    with pytest.raises(pkg.SvtMeError):
        ctx.motion_estimate([(1, 0, 2, p)], sb_rows=(3, 3))
    ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[1], pyr[0], pyr[2])
    assert M.results_diff(whole[0], want) == []


def test_4k_sb_rows_against_oracle_and_properties(pkg, meo):
    """BASELINE configs[2] geometry (3840x2160: 60 x 34 SBs). The oracle checks two SB-row ranges (first rows and the
    ragged last rows) computed as a 2-way row shard would; the whole picture is checked through size-independent
    properties (self-prediction is free; a translation is recovered)."""
    w, h = 3840, 2160
    rng = np.random.default_rng(2160)
    base = M.synth_video(w // 4, h // 4, 3, seed=9, max_shift=8)
    fr = [np.ascontiguousarray(np.kron(f, np.ones((4, 4), np.uint8))[:h, :w]) for f in base]
    fr = [np.clip(f.astype(np.int16) + rng.integers(-6, 7, size=f.shape), 0, 255).astype(np.uint8) for f in fr]
    ctx = pkg.MeContext(w, h, 3)
    for i, f in enumerate(fr):
        ctx.upload(i, f)
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    top, _ = ctx.motion_estimate([(1, 0, 2, p)], sb_rows=(0, 2))
    bottom, _ = ctx.motion_estimate([(1, 0, 2, p)], sb_rows=(32, 34))
    self_res, _ = ctx.motion_estimate([(1, 1, 1, p)])
    ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[1], pyr[0], pyr[2], want_debug=False)
    assert M.results_diff(top[0], want[:2 * 60]) == []
    assert M.results_diff(bottom[0], want[32 * 60:]) == []
    sq = [0] + list(range(5, 85))
    assert not self_res[0]["dist0"][:, sq].any()
    # the WHOLE picture through the 10-bit path (16-bit samples split on the device, ME on the 8-bit plane), and the same
    # picture as three uneven raster ranges of SBs (SvtMeJob.range_in_sbs: the shard of bench.py's strong_4k)
    ctx = pkg.MeContext(w, h, 3)
    for i, f in enumerate(fr):
        ctx.upload16(i, (f.astype(np.uint16) << 2) | rng.integers(0, 4, size=f.shape, dtype=np.uint16))
    whole, _ = ctx.motion_estimate([(1, 0, 2, p)])
    parts = [ctx.motion_estimate([(1, 0, 2, p)], sb_range=r)[0][0] for r in ((0, 255), (255, 1277), (1277, 2040))]
    ctx.close()
    assert M.results_diff(whole[0], want) == []
    assert np.concatenate(parts).tobytes() == whole[0].tobytes()


@pytest.mark.parametrize("name", M.encoder_cases())
def test_real_encoder_me_results(pkg, name):
    """The CUDA path against what the unmodified reference ENCODER computed: every inter picture of a clip encoded through
    the reference's public API (tests/golden/enc_*.npz, BASELINE configs[0]), with the per-picture settings the encoder's
    own processes chose, all pictures of the clip in one batched call. No oracle in between."""
    g = M.load_golden(name)
    frames = M.encoder_clip(g)
    h, w = frames[0].shape
    jobs, lists = [], []
    for hd in g["hdr"]:
        nl = int(hd[7])
        prm, bad = M.encoder_params(pkg.SvtMeParams, hd, pkg.default_params(w, h, nl, int(hd[8]), int(hd[9]), int(hd[10])))
        assert bad == [], (name, int(hd[1]), bad)
        jobs.append((int(hd[1]), int(hd[11]), int(hd[12]) if nl == 2 else int(hd[11]), prm))
        lists.append(nl)
    ctx = pkg.MeContext(w, h, len(frames))
    try:
        for i, f in enumerate(frames):
            ctx.upload(i, f)
        res, _ = ctx.motion_estimate(jobs)
    finally:
        ctx.close()
    for j, (hd, rows) in enumerate(zip(g["hdr"], g["rows"])):
        want = M.encoder_rows(hd, rows).reshape(res[j].shape)      # mixed clips (enc-mode 1/2): 85 or 209 PUs per picture
        assert M.results_diff(res[j], want, lists[j]) == [], (name, int(hd[1]))


def test_sb_facade_call_shape(pkg, meo):
    """include/svtme_lcu.h: one call per SB, from several threads, in scrambled order, with only 3 resident slots (so
    pictures are evicted and re-uploaded): every row equals the oracle's, each picture is searched exactly once."""
    import threading
    w, h = 320, 200
    fr = M.synth_video(w, h, 5, seed=33, max_shift=10)
    n_sb = 5 * 4
    pics = [(2, 0, 4, 2), (1, 0, 2, 2), (3, 2, 4, 2), (4, 0, 0, 1)]           # (cur, ref0, ref1, lists)
    prms = []
    for cur, r0, r1, nl in pics:
        p = pkg.default_params(w, h, nl, 1, 3, 1)
        prms.append(p)
    calls = [(j, sb) for j in range(len(pics)) for sb in range(n_sb)]
    np.random.default_rng(5).shuffle(calls)
    got = np.zeros((len(pics), n_sb, 85), dtype=pkg.PU_RESULT_DTYPE)
    errors = []
    with pkg.MeLcuSession(w, h, num_slots=3) as ses:
        def worker(part):
            try:
                for j, sb in part:
                    cur, r0, r1, nl = pics[j]
                    got[j, sb] = ses.motion_estimate_sb(prms[j], (cur, fr[cur]), (r0, fr[r0]), (r1, fr[r1]) if nl == 2 else None, sb)
            except Exception as e:                                              # noqa: BLE001
                errors.append(e)
        ths = [threading.Thread(target=worker, args=(calls[i::4],)) for i in range(4)]
        [t.start() for t in ths]
        [t.join() for t in ths]
        assert errors == []
        st = ses.stats()
        assert st["rows"] == len(calls) and st["pictures"] >= len(pics) and st["uploads"] >= 5
        with pytest.raises(pkg.SvtMeError):
            ses.motion_estimate_sb(prms[0], (2, fr[2]), (0, fr[0]), (4, fr[4]), n_sb)          # SB index out of range
        with pytest.raises(pkg.SvtMeError):
            ses.motion_estimate_sb(prms[0], (2, fr[2]), (0, fr[0]), None, 0)                   # B picture without list 1
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    for j, (cur, r0, r1, nl) in enumerate(pics):
        want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, prms[j]), pyr[cur], pyr[r0], pyr[r1] if nl == 2 else None)
        assert M.results_diff(got[j], want, nl) == [], j


@pytest.mark.parametrize("w,h", [(416, 240), (256, 192), (64, 64), (200, 72), (1920, 1080)])
def test_zz_sad_matches_oracle(pkg, meo, w, h):
    """svtme_zz_sad (ComputeDecimatedZzSad): a batch of picture pairs -- still, slow, fast and saturating content -- against
    the oracle; incomplete SBs carry INVALID_ZZ_COST; a second call (buffers reused) gives the same bytes."""
    synths = [dict(noise=0.0, max_shift=0), dict(noise=0.6, max_shift=1), dict(noise=1.5, max_shift=3),
              dict(noise=12.0, contrast=4.0, max_shift=60)]
    frames = []
    for seed, synth in enumerate(synths):
        fr = M.synth_video(w, h, 2, seed=seed + w, **synth)
        if seed == 0:
            fr[1] = fr[0].copy()
        frames += fr
    ctx = pkg.MeContext(w, h, len(frames))
    try:
        for i, f in enumerate(frames):
            ctx.upload(i, f)
        pairs = [(2 * k + 1, 2 * k) for k in range(len(synths))]
        zz, nm = ctx.zz_sad(pairs)
        zz2, nm2 = ctx.zz_sad(pairs[:2])
        with pytest.raises(pkg.SvtMeError):
            ctx.zz_sad([(0, len(frames))])
    finally:
        ctx.close()
    pyr = [M.build_pyramid(meo, "meo", f) for f in frames]
    for k, (c, p) in enumerate(pairs):
        wz, wn = M.zz_sad(meo, "meo", w, h, pyr[c], pyr[p])
        assert np.array_equal(zz[k], wz) and np.array_equal(nm[k], wn), k
    assert np.array_equal(zz2, zz[:2]) and np.array_equal(nm2, nm[:2])
    assert not zz[0][zz[0] != 255].any() and not nm[0][zz[0] != 255].any()       # the still pair
    if w % 64 or h % 64:
        assert (zz == 255).any() and (nm[zz == 255] == 30).all()
