"""ctypes binding of include/svttxq.h (forward transform + quantisation); no torch types, no oracle imports."""
import ctypes as C
import numpy as np

from .binding import SvtMeError, load_library

TXQ_EXPORTED_SYMBOLS = (
    "svttxq_create", "svttxq_destroy", "svttxq_last_error", "svttxq_supported", "svttxq_fwd_txfm_quant",
    "svttxq_fwd_txfm_quant_device", "svttxq_synchronize", "svttxq_stream", "svttxq_kernel_launches",
    "svttxq_fwd_txfm_quant_compact",
)
# TxType order of the reference (EbDefinitions.h)
TX_TYPES = ("DCT_DCT", "ADST_DCT", "DCT_ADST", "ADST_ADST", "FLIPADST_DCT", "DCT_FLIPADST", "FLIPADST_FLIPADST",
            "ADST_FLIPADST", "FLIPADST_ADST", "IDTX", "V_DCT", "H_DCT", "V_ADST", "H_ADST", "V_FLIPADST", "H_FLIPADST")


class SvtTxqQuant(C.Structure):
    _fields_ = [(n, C.c_int16 * 2) for n in ("zbin", "round", "quant", "quant_shift", "dequant")]

    @classmethod
    def from_q10(cls, q10):
        """q10 = (zbin dc,ac, round dc,ac, quant dc,ac, quant_shift dc,ac, dequant dc,ac)."""
        q = cls()
        for i, n in enumerate(("zbin", "round", "quant", "quant_shift", "dequant")):
            getattr(q, n)[0], getattr(q, n)[1] = int(q10[2 * i]), int(q10[2 * i + 1])
        return q


_configured = False


def _lib():
    global _configured
    lib = load_library()
    if not _configured:
        lib.svttxq_create.restype = C.c_int32
        lib.svttxq_create.argtypes = [C.POINTER(C.c_void_p), C.c_int32]
        lib.svttxq_destroy.restype = None
        lib.svttxq_destroy.argtypes = [C.c_void_p]
        lib.svttxq_last_error.restype = C.c_char_p
        lib.svttxq_last_error.argtypes = [C.c_void_p]
        lib.svttxq_supported.restype = C.c_int32
        lib.svttxq_supported.argtypes = [C.c_int32] * 3
        for name in ("svttxq_fwd_txfm_quant", "svttxq_fwd_txfm_quant_device"):
            f = getattr(lib, name)
            f.restype = C.c_int32
            f.argtypes = [C.c_void_p, C.c_void_p] + [C.c_int32] * 6 + [C.POINTER(SvtTxqQuant)] + [C.c_void_p] * 5
        lib.svttxq_fwd_txfm_quant_compact.restype = C.c_int32
        lib.svttxq_fwd_txfm_quant_compact.argtypes = ([C.c_void_p, C.c_void_p] + [C.c_int32] * 6 + [C.POINTER(SvtTxqQuant)] +
                                                      [C.c_void_p] * 3 + [C.c_size_t, C.POINTER(C.c_size_t)])
        lib.svttxq_synchronize.restype = C.c_int32
        lib.svttxq_synchronize.argtypes = [C.c_void_p]
        lib.svttxq_stream.restype = C.c_void_p
        lib.svttxq_stream.argtypes = [C.c_void_p]
        lib.svttxq_kernel_launches.restype = C.c_int64
        lib.svttxq_kernel_launches.argtypes = [C.c_void_p]
        _configured = True
    return lib


def supported(txw, txh, tx_type):
    return bool(_lib().svttxq_supported(txw, txh, tx_type))


def packed_dims(txw, txh):
    return min(32, txw), min(32, txh)


class TxqContext:
    """Forward transform + quantisation of the txw x txh tiles of a residual plane (mirrors
    Av1EstimateTransform + Av1QuantizeInvQuantize per transform block)."""

    def __init__(self, device=0):
        self.lib = _lib()
        self._h = C.c_void_p()
        rc = self.lib.svttxq_create(C.byref(self._h), device)
        if rc != 0:
            msg = self.lib.svttxq_last_error(self._h).decode() if self._h else "no CUDA device"
            if self._h:
                self.lib.svttxq_destroy(self._h)
            self._h = None
            raise SvtMeError(f"svttxq_create failed (0x{rc & 0xffffffff:08x}): {msg}")

    def close(self):
        if getattr(self, "_h", None):
            self.lib.svttxq_destroy(self._h)
            self._h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def _check(self, rc, what):
        if rc != 0:
            raise SvtMeError(f"{what} failed (0x{rc & 0xffffffff:08x}): {self.lib.svttxq_last_error(self._h).decode()}")

    def fwd_txfm_quant(self, residual, txw, txh, tx_type, quant, want=("coeff", "qcoeff", "dqcoeff", "eob", "energy")):
        """residual: 2-D int16 array (height, width) whose dimensions are multiples of (txh, txw); rows may be
        strided. Returns a dict of the requested outputs, one row per block in raster order."""
        assert residual.dtype == np.int16 and residual.ndim == 2 and residual.strides[1] == 2
        height, width = residual.shape
        stride = residual.strides[0] // 2
        n_blocks = (width // txw) * (height // txh)
        pw, ph = packed_dims(txw, txh)
        out = {}
        for k in ("coeff", "qcoeff", "dqcoeff"):
            out[k] = np.empty((n_blocks, ph * pw), np.int32) if k in want else None
        out["eob"] = np.empty(n_blocks, np.uint16) if "eob" in want else None
        out["energy"] = np.empty(n_blocks, np.uint64) if "energy" in want else None
        q = quant if isinstance(quant, SvtTxqQuant) else SvtTxqQuant.from_q10(quant)
        ptr = lambda a: a.ctypes.data if a is not None else None
        rc = self.lib.svttxq_fwd_txfm_quant(self._h, residual.ctypes.data, stride, width, height, txw, txh, tx_type,
                                            C.byref(q), ptr(out["coeff"]), ptr(out["qcoeff"]), ptr(out["dqcoeff"]),
                                            ptr(out["eob"]), ptr(out["energy"]))
        self._check(rc, "svttxq_fwd_txfm_quant")
        return {k: v for k, v in out.items() if v is not None}

    def fwd_txfm_quant_compact(self, residual, txw, txh, tx_type, quant, levels=None, eob=None, level_offset=None):
        """Eob-bounded return path: (eob [n_blocks] uint16, level_offset [n_blocks + 1] uint32, levels [total] int16): block b's
        quantised levels in scan order are levels[level_offset[b] : level_offset[b] + eob[b]]."""
        assert residual.dtype == np.int16 and residual.ndim == 2 and residual.strides[1] == 2
        height, width = residual.shape
        n_blocks = (width // txw) * (height // txh)
        pw, ph = packed_dims(txw, txh)
        eob = np.empty(n_blocks, np.uint16) if eob is None else eob
        level_offset = np.empty(n_blocks + 1, np.uint32) if level_offset is None else level_offset
        levels = np.empty(n_blocks * pw * ph, np.int16) if levels is None else levels
        q = quant if isinstance(quant, SvtTxqQuant) else SvtTxqQuant.from_q10(quant)
        n = C.c_size_t(0)
        rc = self.lib.svttxq_fwd_txfm_quant_compact(self._h, residual.ctypes.data, residual.strides[0] // 2, width, height, txw, txh, tx_type,
                                                    C.byref(q), eob.ctypes.data, level_offset.ctypes.data, levels.ctypes.data, levels.size, C.byref(n))
        self._check(rc, "svttxq_fwd_txfm_quant_compact")
        return eob, level_offset, levels[:n.value]

    def fwd_txfm_quant_device(self, d_residual, stride, width, height, txw, txh, tx_type, quant,
                              d_coeff=None, d_qcoeff=None, d_dqcoeff=None, d_eob=None, d_energy=None):
        """All pointers are device addresses (ints); asynchronous on stream()."""
        q = quant if isinstance(quant, SvtTxqQuant) else SvtTxqQuant.from_q10(quant)
        rc = self.lib.svttxq_fwd_txfm_quant_device(self._h, d_residual, stride, width, height, txw, txh, tx_type,
                                                   C.byref(q), d_coeff, d_qcoeff, d_dqcoeff, d_eob, d_energy)
        self._check(rc, "svttxq_fwd_txfm_quant_device")

    def synchronize(self):
        self._check(self.lib.svttxq_synchronize(self._h), "svttxq_synchronize")

    def stream(self):
        return self.lib.svttxq_stream(self._h)

    def kernel_launches(self):
        return int(self.lib.svttxq_kernel_launches(self._h))
