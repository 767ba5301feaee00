"""The arithmetic of the CUDA open-loop intra search (svt-av1-1_b200/csrc/ois_core.cuh, shared by ois.cu) against the oracle
(oracle/ois_oracle.c, itself pinned to the compiled reference in test_ois_oracle_vs_ref.py) -- through a serial host emulation
of the kernel (tests/host_emul/ois_emul.cpp) that calls the same functions in the kernel's order. CPU only."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import me_common as M
from me_common import MeoOisParams, ois_assert_same as assert_same  # noqa: F401  (re-exported for the GPU test)

HERE = os.path.dirname(os.path.abspath(__file__))


class CoreParams(C.Structure):                                   # svtme::ois::Params == SvtMeOisParams
    _fields_ = MeoOisParams._fields_[2:]


def oracle_search(meo, prm, plane, me_sad):
    return M.ois_picture(meo, "meo", prm, plane, me_sad)


def ois_cases():
    """(slice_type, layer, is_ref, resolution, dc_only): every path of OpenLoopIntraSearchLcu"""
    return [(0, 0, 1, 1, 0), (2, 0, 1, 1, 0), (2, 1, 1, 1, 0), (2, 3, 0, 1, 0), (1, 2, 1, 1, 0), (2, 0, 1, 3, 0), (2, 2, 0, 3, 0),
            (2, 4, 0, 3, 0), (2, 2, 0, 1, 1)]


def me_sads(rng, n_sb):
    """ME SADs small, comparable and huge against the DC SADs, and a few zeros, so that every OIS point occurs"""
    scale = rng.choice([0.05, 0.5, 1.0, 2.0, 8.0], size=(n_sb, 85))
    me = (scale * rng.integers(0, 4000, size=(n_sb, 85))).astype(np.uint32)
    me[rng.random((n_sb, 85)) < 0.05] = 0
    return me


@pytest.fixture(scope="module")
def emul():
    so = os.path.join(HERE, "host_emul", "ois_emul.so")
    src = os.path.join(HERE, "host_emul", "ois_emul.cpp")
    core = os.path.join(HERE, "..", "svt-av1-1_b200", "csrc", "ois_core.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(core)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wall", "-o", so, src])
    lib = C.CDLL(so)
    for f in (lib.ois_emul_picture, lib.ois_emul_picture_words, lib.ois_emul_picture_lines):
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(CoreParams), C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


@pytest.mark.parametrize("w,h", [(256, 192), (200, 136), (64, 64), (328, 72)])
@pytest.mark.parametrize("slice_type,tl,is_ref,res,dc_only", ois_cases())
def test_kernel_arithmetic_matches_the_oracle(emul, w, h, slice_type, tl, is_ref, res, dc_only):
    M.build_oracle()
    meo = M.load_oracle()
    rng = np.random.default_rng(w * 7 + h + 10 * tl + slice_type)
    for k, synth in enumerate([dict(), dict(noise=10.0, contrast=3.0), dict(noise=0.0, contrast=0.0)]):
        f = M.synth_video(w, h, 1, seed=5 * k + tl, **synth)[0]
        pyr = M.build_pyramid(meo, "meo", f)
        pl = pyr.c.full
        n_sb = ((w + 63) // 64) * ((h + 63) // 64)
        me = me_sads(rng, n_sb)
        for cu8 in (0, 1):
            enc = 3 if (tl + cu8) % 2 else 1
            out = oracle_search(meo, MeoOisParams(w, h, slice_type, tl, is_ref, cu8, res, dc_only, enc), pl, me)
            prm = CoreParams(slice_type, tl, is_ref, cu8, res, dc_only, enc)
            for kernel in (emul.ois_emul_picture, emul.ois_emul_picture_words, emul.ois_emul_picture_lines):   # the three kernels
                cand, total = np.full((n_sb, 85, 18), 0xFFFFFFFF, np.uint32), np.full((n_sb, 85), 0xFF, np.uint8)
                kernel(pl.buf + pl.origin_y * pl.stride + pl.origin_x, pl.stride, w, h, C.byref(prm), me.ctypes.data,
                       cand.ctypes.data, total.ctypes.data)
                assert_same(out, cand, total, (k, cu8))


@pytest.mark.parametrize("slice_type,tl,res", [(0, 0, 1), (2, 0, 1), (2, 2, 1), (2, 3, 3)])
def test_kernel_arithmetic_on_extreme_content(emul, slice_type, tl, res):
    """Black, white, checkerboards, stripes and full-range noise (test_ois_oracle_vs_ref.extreme_pictures): the clipping of the
    boundary smoothing, zero SADs (OIS point 0), ties in the best-18 selection and in the swap sort."""
    from test_ois_oracle_vs_ref import extreme_pictures
    M.build_oracle()
    meo = M.load_oracle()
    w, h = 200, 136
    rng = np.random.default_rng(7 + tl)
    for name, f in extreme_pictures(w, h).items():
        pyr = M.build_pyramid(meo, "meo", f)
        pl = pyr.c.full
        me = rng.integers(0, 20000, size=(12, 85)).astype(np.uint32)
        out = oracle_search(meo, MeoOisParams(w, h, slice_type, tl, 1 if tl < 3 else 0, 0, res, 0, 2), pl, me)
        prm = CoreParams(slice_type, tl, 1 if tl < 3 else 0, 0, res, 0, 2)
        for kernel in (emul.ois_emul_picture, emul.ois_emul_picture_words, emul.ois_emul_picture_lines):
            cand, total = np.full((12, 85, 18), 0xFFFFFFFF, np.uint32), np.full((12, 85), 0xFF, np.uint8)
            kernel(pl.buf + pl.origin_y * pl.stride + pl.origin_x, pl.stride, w, h, C.byref(prm), me.ctypes.data, cand.ctypes.data, total.ctypes.data)
            assert_same(out, cand, total, name)
