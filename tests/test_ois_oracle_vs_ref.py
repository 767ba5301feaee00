"""Oracle of the open-loop intra search of the ME process (oracle/ois_oracle.c) pinned against the compiled reference
(OpenLoopIntraSearchLcu and its intra-prediction leaves in oracle/_ref/libsvtref.so). CPU only; the CUDA side is checked
against this oracle in test_ois_emul.py (host emulation of the kernels) and test_zzz_ois_parity_gpu.py (through the C-ABI)."""
import ctypes as C

import numpy as np
import pytest

import me_common as M
from me_common import MeoOisParams


@pytest.fixture(scope="module")
def libs():
    M.build_oracle(ref=True)
    if not M.ref_available():
        pytest.skip("oracle/_ref/libsvtref.so not available (no /root/reference here)")
    meo, ref = M.load_oracle(), M.load_ref()
    for lib, pre in ((meo, "meo"), (ref, "ref")):
        f = getattr(lib, pre + "_ois_predict_cu")
        f.restype = C.c_uint32
        f.argtypes = [C.POINTER(M.MeoPlane), C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
        g = getattr(lib, pre + "_ois_picture")
        g.restype = C.c_int if pre == "ref" else None
        g.argtypes = [C.POINTER(MeoOisParams), C.POINTER(M.MeoPlane), C.c_void_p, C.c_void_p]
    return meo, ref


def _picture(w, h, seed, **synth):
    f = M.synth_video(w, h, 1, seed=seed, **synth)[0]
    return M.build_pyramid(M.load_oracle(), "meo", f)


@pytest.mark.parametrize("n", [8, 16, 32])
def test_every_predictor_matches_the_reference_leaves(libs, n):
    """All 35 modes at picture corners, edges (missing neighbours = 128), the right / bottom border where only part of the
    above-right / below-left samples exist, and interior positions; flat, textured and saturating content."""
    meo, ref = libs
    w, h = 160, 96
    for seed, synth in enumerate([dict(), dict(noise=12.0, contrast=4.0), dict(noise=0.0, contrast=0.0)]):
        pyr = _picture(w, h, seed + n, **synth)
        pos = [(0, 0), (n, 0), (0, n), (n, n), (w - n, 0), (w - n, n), (0, h - n), (n, h - n), (w - n, h - n), (w - 2 * n, h - 2 * n), (64, 32)]
        for x0, y0 in pos:
            for mode in range(35):
                a, b = np.zeros((64, 64), np.uint8), np.zeros((64, 64), np.uint8)
                sa = meo.meo_ois_predict_cu(C.byref(pyr.c.full), x0, y0, n, mode, a.ctypes.data)
                sb = ref.ref_ois_predict_cu(C.byref(pyr.c.full), x0, y0, n, mode, b.ctypes.data)
                assert np.array_equal(a[:n, :n], b[:n, :n]), (n, x0, y0, mode)
                assert sa == sb, (n, x0, y0, mode)


def _search(lib, pre, prm, pyr, me_sad):
    return M.ois_picture(lib, pre, prm, pyr.c.full, me_sad)


@pytest.mark.parametrize("w,h", [(256, 192), (200, 136), (64, 64), (328, 72), (136, 264)])
@pytest.mark.parametrize("slice_type,tl,is_ref,res", [(0, 0, 1, 1), (2, 0, 1, 1), (2, 1, 1, 1), (2, 3, 0, 1), (1, 2, 1, 1),
                                                      (2, 0, 1, 3), (2, 2, 0, 3), (2, 1, 1, 1)])
def test_open_loop_intra_search_matches_the_reference(libs, w, h, slice_type, tl, is_ref, res):
    """Every path of OpenLoopIntraSearchLcu: intra pictures (7 modes + injected lists), layer-0 inter pictures (all 35 modes,
    best 18 sorted), the other inter pictures (DC, then 3..9 stage-1 directions steered by the ME SADs -- small, comparable and
    huge ones so that every OIS point occurs), 4K threshold sets, incomplete SBs, 8x8 CUs switched off."""
    meo, ref = libs
    rng = np.random.default_rng(w + h + 10 * tl + slice_type)
    totals = set()
    for k, synth in enumerate([dict(), dict(noise=10.0, contrast=3.0)]):
        pyr = _picture(w, h, 3 * k + tl, **synth)
        n_sb = ((w + 63) // 64) * ((h + 63) // 64)
        scale = rng.choice([0.05, 0.5, 1.0, 2.0, 8.0], size=(n_sb, 85))
        me_sad = (scale * rng.integers(0, 4000, size=(n_sb, 85))).astype(np.uint32)
        me_sad[rng.random((n_sb, 85)) < 0.05] = 0
        for cu8 in (0, 1):
            prm = MeoOisParams(w, h, slice_type, tl, is_ref, cu8, res, 0, 3 if (tl + cu8) % 2 else 1)   # enc-mode: one quirk reads it
            a = _search(meo, "meo", prm, pyr, me_sad)
            b = _search(ref, "ref", prm, pyr, me_sad)
            assert np.array_equal(a["total"], b["total"])
            totals |= set(np.unique(a["total"]).tolist())
            for f in ("distortion", "mode", "valid"):
                bad = np.argwhere(a["cand"][f] != b["cand"][f])
                assert len(bad) == 0, (f, bad[:5].tolist(), a["cand"][f][tuple(bad[0])], b["cand"][f][tuple(bad[0])])
    if slice_type == 2 and tl > 0 and w >= 200:
        assert len(totals & {1, 3, 5, 7, 9}) >= 3                          # several OIS points were exercised


def test_dc_only_mode(libs):
    meo, ref = libs
    pyr = _picture(192, 128, 5)
    prm = MeoOisParams(192, 128, 2, 2, 0, 0, 1, 1, 3)
    me = np.full((6, 85), 1000, np.uint32)
    a, b = _search(meo, "meo", prm, pyr, me), _search(ref, "ref", prm, pyr, me)
    assert a.tobytes() == b.tobytes()


def extreme_pictures(w, h):
    """Content that drives the predictors into their corners: black, white, a 0 / 255 checkerboard at pixel and at 8-pixel pitch,
    vertical and horizontal stripes (the boundary smoothing of the pure modes clips both ways), full-range noise."""
    yy, xx = np.mgrid[0:h, 0:w]
    rng = np.random.default_rng(w * 3 + h)
    return {"black": np.zeros((h, w), np.uint8), "white": np.full((h, w), 255, np.uint8),
            "checker1": (((xx + yy) & 1) * 255).astype(np.uint8), "checker8": ((((xx >> 3) + (yy >> 3)) & 1) * 255).astype(np.uint8),
            "vstripes": (((xx >> 2) & 1) * 255).astype(np.uint8), "hstripes": (((yy >> 1) & 1) * 255).astype(np.uint8),
            "noise": rng.integers(0, 256, size=(h, w), dtype=np.uint8)}


@pytest.mark.parametrize("slice_type,tl,res", [(0, 0, 1), (2, 0, 1), (2, 2, 1), (2, 3, 3)])
def test_extreme_content_matches_the_reference(libs, slice_type, tl, res):
    meo, ref = libs
    w, h = 200, 136
    rng = np.random.default_rng(7 + tl)
    for name, f in extreme_pictures(w, h).items():
        pyr = M.build_pyramid(meo, "meo", f)
        me_sad = rng.integers(0, 20000, size=(12, 85)).astype(np.uint32)
        prm = MeoOisParams(w, h, slice_type, tl, 1 if tl < 3 else 0, 0, res, 0, 2)
        a, b = _search(meo, "meo", prm, pyr, me_sad), _search(ref, "ref", prm, pyr, me_sad)
        assert a.tobytes() == b.tobytes(), name
