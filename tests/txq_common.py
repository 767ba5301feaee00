"""ctypes views of the transform / quantisation oracle (oracle/txq_oracle.c) and of the reference accessors
(oracle/ref_txq.c). Test infrastructure only."""
import ctypes as C
import os

import numpy as np

import me_common as M

TX_SIZES = [(4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (4, 8), (8, 4), (8, 16), (16, 8), (16, 32), (32, 16), (32, 64),
            (64, 32), (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]
TX_TYPE_NAMES = ["DCT_DCT", "ADST_DCT", "DCT_ADST", "ADST_ADST", "FLIPADST_DCT", "DCT_FLIPADST", "FLIPADST_FLIPADST",
                 "ADST_FLIPADST", "FLIPADST_ADST", "IDTX", "V_DCT", "H_DCT", "V_ADST", "H_ADST", "V_FLIPADST", "H_FLIPADST"]
_V = [0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3]
_H = [0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2]


def packed(w, h):
    return min(w, 32), min(h, 32)


def log_scale(w, h):
    return int(w * h > 256) + int(w * h > 1024)


def type_supported(w, h, t):
    """(size, type) pairs with 1-D kernels on both axes: ADST only up to 16 points; 64-point only DCT."""
    v, hh = _V[t], _H[t]
    if v in (1, 2) and h > 16:
        return False
    if hh in (1, 2) and w > 16:
        return False
    if (w == 64 or h == 64) and t != 0:
        return False
    return True


def oracle():
    lib = M.load_oracle()
    lib.meo_fwd_txfm.restype = C.c_int
    lib.meo_fwd_txfm.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
    lib.meo_scan.restype = C.c_int
    lib.meo_scan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.meo_quantize_b.restype = None
    lib.meo_quantize_b.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint16)]
    return lib


def reference():
    lib = M.load_ref()
    lib.ref_txq_init()
    lib.ref_fwd_txfm.restype = C.c_int
    lib.ref_fwd_txfm.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint64)]
    lib.ref_quantize.restype = None
    lib.ref_quantize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint16)]
    lib.ref_scan.restype = C.POINTER(C.c_int16)
    lib.ref_scan.argtypes = [C.c_int] * 3
    lib.ref_max_eob.restype = C.c_int
    lib.ref_quant_params.restype = None
    lib.ref_quant_params.argtypes = [C.c_int, C.c_int, C.c_void_p]
    return lib


def aligned(n, dtype, align=64):
    """numpy array whose data pointer is `align`-byte aligned (the reference's AVX2 kernels use aligned stores)."""
    item = np.dtype(dtype).itemsize
    raw = np.zeros(n * item + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * item].view(dtype)


def fwd_txfm(lib, prefix, res, w, h, t):
    """res: int16 (h, w) C-contiguous. Returns (packed coefficients int32 (ph, pw), energy)."""
    pw, ph = packed(w, h)
    e = C.c_uint64(0)
    res = np.ascontiguousarray(res, dtype=np.int16)
    if prefix == "meo":
        out = np.zeros(pw * ph, dtype=np.int32)
        rc = lib.meo_fwd_txfm(res.ctypes.data, w, out.ctypes.data, w, h, t, C.byref(e))
        assert rc == 0
    else:
        buf = aligned(w * h, np.int32)
        rin = aligned(w * h, np.int16)
        rin[:] = res.reshape(-1)
        lib.ref_fwd_txfm(rin.ctypes.data, w, buf.ctypes.data, w, h, t, 0, C.byref(e))
        out = buf[:pw * ph].copy()
    return out.reshape(ph, pw), e.value


def scan(lib, prefix, w, h, t):
    pw, ph = packed(w, h)
    if prefix == "meo":
        s = np.zeros(pw * ph, dtype=np.int16)
        assert lib.meo_scan(w, h, t, s.ctypes.data) == pw * ph
        return s
    p = lib.ref_scan(w, h, t)
    return np.array([p[i] for i in range(pw * ph)], dtype=np.int16)


def quantize(lib, prefix, coeff, w, h, t, q10):
    pw, ph = packed(w, h)
    n = pw * ph
    coeff = np.ascontiguousarray(coeff, dtype=np.int32).reshape(-1)
    q10 = np.ascontiguousarray(q10, dtype=np.int16)
    qc, dq, eob = aligned(n, np.int32), aligned(n, np.int32), C.c_uint16(0)
    ca = aligned(n, np.int32)
    ca[:] = coeff
    coeff = ca
    if prefix == "meo":
        s = scan(lib, "meo", w, h, t)
        lib.meo_quantize_b(coeff.ctypes.data, n, q10.ctypes.data, log_scale(w, h), s.ctypes.data, qc.ctypes.data, dq.ctypes.data, C.byref(eob))
    else:
        lib.ref_quantize(coeff.ctypes.data, w, h, t, q10.ctypes.data, qc.ctypes.data, dq.ctypes.data, C.byref(eob))
    return qc, dq, eob.value


def quant_params(ref, qindex, bit_depth=8):
    q10 = np.zeros(10, dtype=np.int16)
    ref.ref_quant_params(bit_depth, qindex, q10.ctypes.data)
    return q10


def txq_plane(lib, res, w, h, t, q10):
    """Oracle over a whole residual plane (int16 (H, W), rows may be strided): dict like TxqContext.fwd_txfm_quant."""
    assert res.dtype == np.int16 and res.strides[1] == 2
    height, width = res.shape
    pw, ph = packed(w, h)
    nb = (width // w) * (height // h)
    out = {k: np.zeros((nb, pw * ph), np.int32) for k in ("coeff", "qcoeff", "dqcoeff")}
    out["eob"] = np.zeros(nb, np.uint16)
    out["energy"] = np.zeros(nb, np.uint64)
    q10 = np.ascontiguousarray(q10, dtype=np.int16)
    lib.meo_txq_plane.restype = C.c_int
    lib.meo_txq_plane.argtypes = [C.c_void_p] + [C.c_int] * 6 + [C.c_void_p] * 6
    rc = lib.meo_txq_plane(res.ctypes.data, res.strides[0] // 2, width, height, w, h, t, q10.ctypes.data,
                           out["coeff"].ctypes.data, out["qcoeff"].ctypes.data, out["dqcoeff"].ctypes.data,
                           out["eob"].ctypes.data, out["energy"].ctypes.data)
    assert rc == nb
    return out


def synth_residual(rng, height, width, bit_depth=8, kind="mixed"):
    """Residual-like int16 plane: smooth gradients + texture + sparse edges, clipped to the legal range."""
    lim = (1 << bit_depth) - 1
    if kind == "noise":
        return rng.integers(-lim, lim + 1, size=(height, width)).astype(np.int16)
    y, x = np.mgrid[0:height, 0:width]
    base = 20 * np.sin(x / 7.3) * np.cos(y / 11.1) + rng.normal(0, 6, size=(height, width))
    edges = (rng.random((height, width)) < 0.01) * rng.integers(-lim, lim + 1, size=(height, width))
    return np.clip(np.rint(base + edges), -lim, lim).astype(np.int16)


# ---- blocks recorded from the EncDec stage of whole-encoder runs (tests/golden/txq_enc_*.npz, make_encoder_golden.py) ----

def encoder_block_files():
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    return sorted(f[:-4] for f in os.listdir(d) if f.startswith("txq_enc_") and f.endswith(".npz"))


def encoder_block_groups(name):
    """Yield (w, h, tx_type, q10, residual plane (h, k*w) int16, expected) per (size, type, quantiser) group of one file:
    the k blocks of a group side by side, which is the raster the plane entry points take."""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    hdr, q10s = g["hdr"], g["q10"]
    r_off = np.concatenate([[0], np.cumsum(hdr[:, 2].astype(np.int64) * hdr[:, 3])])
    c_off = np.concatenate([[0], np.cumsum(hdr[:, 7].astype(np.int64))])
    groups = {}
    for i, hd in enumerate(hdr):
        assert int(hd[6]) == 0                                   # 8-bit encodes
        groups.setdefault((int(hd[2]), int(hd[3]), int(hd[4]), tuple(int(v) for v in q10s[i])), []).append(i)
    for (w, h, t, q10), idx in sorted(groups.items()):
        n = int(hdr[idx[0]][7])
        assert (n,) == (packed(w, h)[0] * packed(w, h)[1],)
        plane = np.concatenate([g["residual"][r_off[i]:r_off[i + 1]].reshape(h, w) for i in idx], axis=1)
        want = {k: np.stack([g[k][c_off[i]:c_off[i + 1]] for i in idx]) for k in ("coeff", "qcoeff", "dqcoeff")}
        want["eob"] = np.array([hdr[i][10] for i in idx], np.uint16)
        want["energy"] = np.array([(int(hdr[i][11]) & 0xffffffff) | (int(hdr[i][12]) << 32) for i in idx], np.uint64)
        yield w, h, t, np.array(q10, np.int16), np.ascontiguousarray(plane), want


def check_encoder_blocks(got, want, w, h, tag):
    keys = ("coeff", "qcoeff", "dqcoeff", "eob") + (("energy",) if 64 in (w, h) else ())
    for k in keys:
        assert np.array_equal(got[k], want[k]), (tag, k)
