"""Shared test plumbing: ctypes views of oracle/me_oracle.h, library loaders, synthetic video.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use
the oracle; the product path (svt_av1_1_b200) never imports this module.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

PAD_FULL, PAD_QUARTER, PAD_SIXTEENTH = 68, 32, 16
SQUARE_PU_COUNT = 85


class MeoPlane(C.Structure):
    _fields_ = [("buf", C.c_void_p), ("stride", C.c_int32), ("origin_x", C.c_int32), ("origin_y", C.c_int32),
                ("width", C.c_int32), ("height", C.c_int32)]


class MeoPyramid(C.Structure):
    _fields_ = [("full", MeoPlane), ("quarter", MeoPlane), ("sixteenth", MeoPlane)]


class MeoParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "width", "height", "num_lists", "temporal_layer_index", "hierarchical_levels", "is_used_as_reference",
        "same_ref_poc", "enable_hme_flag", "enable_hme_level0", "enable_hme_level1", "enable_hme_level2",
        "use_subpel", "num_pus", "cu8x8_mode", "num_hme_regions_w", "num_hme_regions_h",
        "hme_l0_total_w", "hme_l0_total_h")] + [
        ("hme_l0_w", C.c_int32 * 2), ("hme_l0_h", C.c_int32 * 2), ("hme_l1_w", C.c_int32 * 2),
        ("hme_l1_h", C.c_int32 * 2), ("hme_l2_w", C.c_int32 * 2), ("hme_l2_h", C.c_int32 * 2),
        ("search_area_w", C.c_int32), ("search_area_h", C.c_int32),
        ("hme_l0_mult_x", C.c_int32), ("hme_l0_mult_y", C.c_int32)]

    def as_tuple(self):
        out = []
        for n, _t in self._fields_:
            v = getattr(self, n)
            out.append(tuple(v) if hasattr(v, "__len__") else v)
        return tuple(out)


PU_RESULT_DTYPE = np.dtype([("xMvL0", "<i2"), ("yMvL0", "<i2"), ("xMvL1", "<i2"), ("yMvL1", "<i2"),
                            ("dist0", "<u4"), ("dir0", "<u4"), ("dist1", "<u4"), ("dir1", "<u4"),
                            ("dist2", "<u4"), ("dir2", "<u4"), ("total", "u1"), ("pad", "u1", (7,))])
assert PU_RESULT_DTYPE.itemsize == 40

SB_DEBUG_DTYPE = np.dtype([("search_center_x", "<i2", (2,)), ("search_center_y", "<i2", (2,)),
                           ("sa_origin_x", "<i2", (2,)), ("sa_origin_y", "<i2", (2,)),
                           ("sa_width", "<i2", (2,)), ("sa_height", "<i2", (2,)),
                           ("best_sad", "<u4", (2, 85)), ("best_mv", "<u4", (2, 85))])


def build_oracle(ref=True):
    """Compile oracle/libme_oracle.so (and oracle/_ref/libsvtref.so when the reference tree exists)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True, stdout=subprocess.DEVNULL)
    if ref and os.path.isdir("/root/reference/Source/Lib/Codec"):
        subprocess.run(["make", "-s", "-j8", "-C", ORACLE_DIR, "ref"], check=True, stdout=subprocess.DEVNULL)


def _bind(lib, prefix):
    f = getattr(lib, prefix + "_motion_estimate_picture")
    f.restype = C.c_int
    f.argtypes = [C.POINTER(MeoParams), C.POINTER(MeoPyramid), C.POINTER(MeoPyramid), C.POINTER(MeoPyramid),
                  C.c_void_p, C.c_void_p]
    g = getattr(lib, prefix + "_build_pyramid")
    g.restype = None
    g.argtypes = [C.c_void_p, C.c_int32, C.POINTER(MeoPyramid)]
    return lib


_libs = {}


def load_oracle():
    if "meo" not in _libs:
        path = os.path.join(ORACLE_DIR, "libme_oracle.so")
        if not os.path.exists(path):
            build_oracle(ref=False)
        lib = C.CDLL(path)
        lib.meo_default_params.restype = None
        lib.meo_default_params.argtypes = [C.POINTER(MeoParams)] + [C.c_int32] * 6
        _libs["meo"] = _bind(lib, "meo")
    return _libs["meo"]


def ref_available():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libsvtref.so"))


def load_ref():
    if "ref" not in _libs:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libsvtref.so"))
        lib.ref_default_params.restype = None
        lib.ref_default_params.argtypes = [C.POINTER(MeoParams)] + [C.c_int32] * 6
        _libs["ref"] = _bind(lib, "ref")
    return _libs["ref"]


class HostPyramid:
    """Padded full / quarter / sixteenth planes in numpy memory + the MeoPyramid describing them."""

    def __init__(self, width, height):
        self.width, self.height = width, height
        self.planes = []
        self.c = MeoPyramid()
        for name, sh, pad in (("full", 0, PAD_FULL), ("quarter", 1, PAD_QUARTER), ("sixteenth", 2, PAD_SIXTEENTH)):
            w, h = width >> sh, height >> sh
            arr = np.zeros((h + 2 * pad, w + 2 * pad), dtype=np.uint8)
            self.planes.append(arr)
            p = getattr(self.c, name)
            p.buf = arr.ctypes.data
            p.stride, p.origin_x, p.origin_y, p.width, p.height = w + 2 * pad, pad, pad, w, h

    @property
    def full(self):
        return self.planes[0]

    @property
    def quarter(self):
        return self.planes[1]

    @property
    def sixteenth(self):
        return self.planes[2]


def build_pyramid(lib, prefix, luma):
    h, w = luma.shape
    luma = np.ascontiguousarray(luma)
    pyr = HostPyramid(w, h)
    getattr(lib, prefix + "_build_pyramid")(luma.ctypes.data, w, C.byref(pyr.c))
    return pyr


def default_params(lib, prefix, width, height, num_lists=2, tl=1, hl=3, is_ref=1):
    p = MeoParams()
    getattr(lib, prefix + "_default_params")(C.byref(p), width, height, num_lists, tl, hl, is_ref)
    return p


def motion_estimate(lib, prefix, params, cur, ref0, ref1=None, want_debug=True):
    n_sb = ((params.width + 63) // 64) * ((params.height + 63) // 64)
    res = np.zeros((n_sb, params.num_pus), dtype=PU_RESULT_DTYPE)
    dbg = np.zeros(n_sb, dtype=SB_DEBUG_DTYPE) if want_debug else None
    rc = getattr(lib, prefix + "_motion_estimate_picture")(
        C.byref(params), C.byref(cur.c), C.byref(ref0.c), C.byref(ref1.c) if ref1 is not None else None,
        res.ctypes.data, dbg.ctypes.data if want_debug else None)
    if rc != 0:
        raise RuntimeError(f"{prefix}_motion_estimate_picture -> {rc}")
    return res, dbg


def zz_sad(lib, prefix, width, height, cur, prev):
    """ComputeDecimatedZzSad of one picture pair: (zz_cost, non_moving_index), one byte per SB."""
    n_sb = ((width + 63) // 64) * ((height + 63) // 64)
    zz, nm = np.full(n_sb, 77, np.uint8), np.full(n_sb, 77, np.uint8)
    f = getattr(lib, prefix + "_zz_sad_picture")
    f.restype = C.c_int if prefix == "ref" else None
    f.argtypes = [C.c_int32, C.c_int32, C.POINTER(MeoPyramid), C.POINTER(MeoPyramid), C.c_void_p, C.c_void_p]   # (without them ctypes passes addresses as 32-bit ints)
    rc = f(width, height, C.byref(cur.c), C.byref(prev.c), zz.ctypes.data, nm.ctypes.data)
    if prefix == "ref" and rc != 0:
        raise RuntimeError(f"ref_zz_sad_picture -> {rc}")
    return zz, nm


class MeoOisParams(C.Structure):                                 # oracle/me_oracle.h
    _fields_ = [(n, C.c_int32) for n in ("width", "height", "slice_type", "temporal_layer_index", "is_used_as_reference",
                                         "cu8x8_mode", "input_resolution", "limit_ois_to_dc", "enc_mode")]


OIS_CU = np.dtype([("cand", [("distortion", "<u4"), ("mode", "u1"), ("valid", "u1"), ("pad", "u1", (2,))], (18,)),
                   ("total", "u1"), ("pad", "u1", (3,))])          # MeoOisCu


def ois_picture(lib, prefix, prm, plane, me_sad):
    """OpenLoopIntraSearchLcu for every SB of a picture (meo_ / ref_ois_picture): [num_sb][85] OIS_CU records, zero where the
    search writes nothing. plane: the MeoPlane of the full-resolution picture; me_sad: [num_sb][85] uint32 or None."""
    n_sb = ((prm.width + 63) // 64) * ((prm.height + 63) // 64)
    out = np.zeros((n_sb, 85), OIS_CU)
    f = getattr(lib, prefix + "_ois_picture")
    f.restype = C.c_int if prefix == "ref" else None
    f.argtypes = [C.POINTER(MeoOisParams), C.POINTER(MeoPlane), C.c_void_p, C.c_void_p]
    rc = f(C.byref(prm), C.byref(plane), me_sad.ctypes.data if me_sad is not None else None, out.ctypes.data)
    if prefix == "ref" and rc != 0:
        raise RuntimeError(f"ref_ois_picture -> {rc}")
    return out


def ois_assert_same(out, cand, total, what=""):
    """OIS_CU records of the oracle against svtme_ois output (byte images of OisCandidate_t: distortion : 20, valid : 1, 3 pad
    bits, mode : 8, and the candidate counts)."""
    d, v, m = cand & 0xFFFFF, (cand >> 20) & 1, cand >> 24
    assert np.array_equal(out["total"], total), what
    for name, got in (("distortion", d), ("valid", v), ("mode", m)):
        bad = np.argwhere(out["cand"][name] != got)
        assert len(bad) == 0, (what, name, bad[:5].tolist(), out["cand"][name][tuple(bad[0])], got[tuple(bad[0])])


def synth_video(width, height, n_frames, seed=0, noise=3.0, max_shift=24, contrast=1.0):
    """Deterministic synthetic luma: smooth texture + edges, per-region translation + noise.

    Motion is piecewise (a global pan plus independently moving rectangles) so HME, the 64x64
    search and the sub-pel stages all have work to do and different PU sizes disagree."""
    rng = np.random.default_rng(seed)
    H, W = height + 4 * max_shift + 512, width + 512
    base = rng.integers(0, 256, size=(H // 8 + 2, W // 8 + 2)).astype(np.float32)
    base = np.kron(base, np.ones((8, 8), dtype=np.float32))[:H, :W]
    # separable smoothing so half/quarter-pel positions matter
    k = np.array([1, 4, 6, 4, 1], dtype=np.float32) / 16.0
    for _ in range(2):
        base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base)
        base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 0, base)
    fine = rng.normal(0, 12.0, size=base.shape).astype(np.float32)
    base = 128 + contrast * (base - 128) + fine
    frames = []
    pan = rng.integers(-max_shift // 2, max_shift // 2 + 1, size=2)
    objs = [(rng.integers(0, max(1, height - 64)), rng.integers(0, max(1, width - 64)), rng.integers(48, 160), rng.integers(48, 160),
             rng.integers(-max_shift, max_shift + 1, size=2)) for _ in range(6)]
    for t in range(n_frames):
        oy, ox = 128 + pan[0] * t // 2, 128 + pan[1] * t // 2
        f = base[oy:oy + height, ox:ox + width].copy()
        for (y0, x0, hh, ww, v) in objs:
            yy, xx = int(y0 + v[0] * t / 2), int(x0 + v[1] * t / 2)
            y1, x1 = max(0, yy), max(0, xx)
            y2, x2 = min(height, yy + hh), min(width, xx + ww)
            if y2 > y1 and x2 > x1:
                sy, sx = (int(y0) * 7 + 300) % (H - height - 160), (int(x0) * 5 + 200) % (W - width - 160)
                f[y1:y2, x1:x2] = base[sy + (y1 - yy):sy + (y2 - yy), sx + (x1 - xx):sx + (x2 - xx)] * 0.9 + 10
        f = f + rng.normal(0, noise, size=f.shape)
        frames.append(np.clip(np.rint(f), 0, 255).astype(np.uint8))
    return frames


# ---------------------------------------------------------------------------------------------
# parameter (de)serialisation, result comparison, golden fixtures
def params_to_list(p):
    """Flatten a MeoParams / SvtMeParams (identical layouts) into a list of int32."""
    out = []
    for n, _t in p._fields_:
        v = getattr(p, n)
        out.extend(list(v) if hasattr(v, "__len__") else [v])
    return out


def params_from_list(cls, vals):
    p = cls()
    it = iter(int(v) for v in vals)
    for n, _t in p._fields_:
        v = getattr(p, n)
        if hasattr(v, "__len__"):
            for i in range(len(v)):
                v[i] = next(it)
        else:
            setattr(p, n, next(it))
    return p


def convert_params(cls, p):
    return params_from_list(cls, params_to_list(p))


def results_diff(a, b, num_lists=2):
    """Human-readable list of differing fields between two PU_RESULT arrays ([] when bit-exact).

    For P pictures (num_lists == 1) the reference copies stale per-thread context memory into
    xMvL1/yMvL1 (EbMotionEstimation.c:7137-7140); those two fields are excluded there."""
    bad = []
    for n in a.dtype.names:
        if n == "pad" or (num_lists == 1 and n in ("xMvL1", "yMvL1")):
            continue
        if not np.array_equal(a[n], b[n]):
            idx = np.argwhere(a[n] != b[n])
            bad.append(f"{n}: {len(idx)} differ, first at {idx[0].tolist()}: {a[n][tuple(idx[0])]} vs {b[n][tuple(idx[0])]}")
    return bad


GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def golden_cases():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and not f.startswith(("txq_", "enc_")))


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))


def plane_crc(a):
    import zlib
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def unpack16(lib, prefix, luma16):
    """10-bit split of a uint16 (h, w) plane through the oracle ('meo') or the reference ('ref'): (msb, rem) uint8 planes."""
    h, w = luma16.shape
    src = np.ascontiguousarray(luma16)
    msb, rem = np.zeros((h, w), np.uint8), np.zeros((h, w), np.uint8)
    f = getattr(lib, prefix + "_unpack16")
    f.restype = None
    f.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    f(src.ctypes.data, w, msb.ctypes.data, w, rem.ctypes.data, w, w, h)
    return msb, rem


# ---- goldens recorded from whole-encoder runs of the reference (tests/golden/make_encoder_golden.py) ----

def encoder_cases():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz") and f.startswith("enc_"))


def encoder_rows(hd, rows):
    """MeCuResults_t rows of one dumped picture: [n_sb][hdr[18] PUs] (fixtures pad the rows of mixed clips to 209 PUs)."""
    return np.ascontiguousarray(rows[:, :int(hd[18]) * 40]).view(PU_RESULT_DTYPE)


def encoder_clip(g):
    """Regenerate the synthetic clip an enc_* golden was encoded from and check it against the stored CRCs."""
    w, h, n, _mode, seed, _hl = (int(v) for v in g["geometry"])
    frames = synth_video(w, h, n, seed=seed, **eval(str(g["synth"][0]), {"dict": dict}))
    assert [plane_crc(f) for f in frames] == [int(v) for v in g["frame_crc"]], "synthetic clip drifted from the golden"
    return frames


def encoder_params(cls, hd, defaults):
    """ME parameters of one dumped picture: `defaults` (the library's table defaults for the picture's layer) with every
    field the encoder's own processes set (dump header, oracle/ref_driver.c) written over it. Returns (params, mismatches)
    where mismatches lists the fields whose default differs from what the encoder used."""
    enc = {"width": hd[5], "height": hd[6], "num_lists": hd[7], "temporal_layer_index": hd[8], "hierarchical_levels": hd[9],
           "is_used_as_reference": hd[10], "same_ref_poc": int(hd[7] == 2 and hd[11] == hd[12]),
           "enable_hme_flag": hd[13], "enable_hme_level0": hd[14], "enable_hme_level1": hd[15], "enable_hme_level2": hd[16],
           "use_subpel": hd[17], "num_pus": hd[18], "cu8x8_mode": hd[19], "num_hme_regions_w": hd[20], "num_hme_regions_h": hd[21],
           "hme_l0_total_w": hd[22], "hme_l0_total_h": hd[23], "hme_l0_w": hd[24:26], "hme_l0_h": hd[26:28],
           "hme_l1_w": hd[28:30], "hme_l1_h": hd[30:32], "hme_l2_w": hd[32:34], "hme_l2_h": hd[34:36],
           "search_area_w": hd[36], "search_area_h": hd[37]}
    p = convert_params(cls, defaults)
    # signal_derivation_multi_processes_oq (EbPictureDecisionProcess.c:357-379): PIC_ALL(_C)_DEPTH_MODE = 209 PUs for every
    # picture of enc-mode 0 and for the reference pictures of enc-modes 1 and 2; 85 otherwise. The library default is 85.
    if len(hd) > 38 and (int(hd[38]) == 0 or (int(hd[38]) in (1, 2) and int(hd[10]))):
        p.num_pus = 209
    bad = []
    for k, v in enc.items():
        cur = getattr(p, k)
        if hasattr(cur, "__len__"):
            if [int(x) for x in cur] != [int(x) for x in v]:
                bad.append(k)
            for i in range(len(cur)):
                cur[i] = int(v[i])
        else:
            if k != "same_ref_poc" and int(cur) != int(v):
                bad.append(k)
            setattr(p, k, int(v))
    return p, bad
