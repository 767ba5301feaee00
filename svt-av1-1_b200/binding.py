"""ctypes binding of include/svtme.h (no torch types, no oracle imports)."""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SQUARE_PU_COUNT = 85
MAX_PU_COUNT = 209
MAX_BATCH = 128

EXPORTED_SYMBOLS = (
    "svtme_default_params", "svtme_create", "svtme_destroy", "svtme_last_error", "svtme_picture_upload",
    "svtme_picture_upload_device", "svtme_picture_download", "svtme_motion_estimate",
    "svtme_motion_estimate_async", "svtme_fetch_results", "svtme_synchronize", "svtme_stream",
    "svtme_kernel_launches", "svtme_set_profiling", "svtme_kernel_times", "svtme_join", "svtme_picture_upload_16bit",
    "svtme_picture_upload_peer", "svtme_slot_export", "svtme_picture_upload_ipc", "svtme_synchronize_uploads",
    "svtme_pictures_upload_device", "svtme_pictures_upload", "svtme_motion_estimate_static", "svtme_rc_me_distortion",
    "svtme_zz_sad", "svtme_ois",
)
LCU_EXPORTED_SYMBOLS = ("svtme_lcu_session_create", "svtme_lcu_session_destroy", "svtme_lcu_last_error",
                        "svtme_lcu_motion_estimate", "svtme_lcu_session_stats", "svtme_lcu_motion_estimate_segment")
KERNEL_KINDS = ("pyramid", "hme", "fullpel", "subpel", "bipred_pack", "ois")


class SvtMeError(RuntimeError):
    pass


class SvtMeParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "width", "height", "num_lists", "temporal_layer_index", "hierarchical_levels", "is_used_as_reference",
        "same_ref_poc", "enable_hme_flag", "enable_hme_level0", "enable_hme_level1", "enable_hme_level2",
        "use_subpel", "num_pus", "cu8x8_mode", "num_hme_regions_w", "num_hme_regions_h",
        "hme_l0_total_w", "hme_l0_total_h")] + [
        ("hme_l0_w", C.c_int32 * 2), ("hme_l0_h", C.c_int32 * 2), ("hme_l1_w", C.c_int32 * 2),
        ("hme_l1_h", C.c_int32 * 2), ("hme_l2_w", C.c_int32 * 2), ("hme_l2_h", C.c_int32 * 2),
        ("search_area_w", C.c_int32), ("search_area_h", C.c_int32),
        ("hme_l0_mult_x", C.c_int32), ("hme_l0_mult_y", C.c_int32)]


class SvtMeIpcSlot(C.Structure):
    _fields_ = [("mem_handle", C.c_uint8 * 64), ("event_handle", C.c_uint8 * 64), ("luma_offset", C.c_int64),
                ("pitch", C.c_int32), ("device", C.c_int32), ("width", C.c_int32), ("height", C.c_int32)]


class SvtMeLcuPicture(C.Structure):
    _fields_ = [("picture_number", C.c_uint64), ("luma", C.c_void_p), ("stride", C.c_int32), ("reserved", C.c_int32)]


class SvtMeJob(C.Structure):
    _fields_ = [("cur_slot", C.c_int32), ("ref0_slot", C.c_int32), ("ref1_slot", C.c_int32),
                ("sb_row_begin", C.c_int32), ("sb_row_end", C.c_int32), ("range_in_sbs", C.c_int32),
                ("params", SvtMeParams), ("results", C.c_void_p), ("debug", C.c_void_p)]


class SvtMeOisParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("slice_type", "temporal_layer_index", "is_used_as_reference", "cu8x8_mode",
                                         "input_resolution", "limit_ois_to_dc", "enc_mode")]


class SvtMeOisJob(C.Structure):
    _fields_ = [("slot", C.c_int32), ("params", SvtMeOisParams), ("me_sad", C.c_void_p), ("cand", C.c_void_p), ("total", C.c_void_p)]


class SvtMeZzJob(C.Structure):
    _fields_ = [("cur_slot", C.c_int32), ("prev_slot", C.c_int32), ("zz_cost", C.c_void_p), ("non_moving_index", C.c_void_p)]


PU_RESULT_DTYPE = np.dtype([("xMvL0", "<i2"), ("yMvL0", "<i2"), ("xMvL1", "<i2"), ("yMvL1", "<i2"),
                            ("dist0", "<u4"), ("dir0", "<u4"), ("dist1", "<u4"), ("dir1", "<u4"),
                            ("dist2", "<u4"), ("dir2", "<u4"), ("total", "u1"), ("pad", "u1", (7,))])
SB_DEBUG_DTYPE = np.dtype([("search_center_x", "<i2", (2,)), ("search_center_y", "<i2", (2,)),
                           ("sa_origin_x", "<i2", (2,)), ("sa_origin_y", "<i2", (2,)),
                           ("sa_width", "<i2", (2,)), ("sa_height", "<i2", (2,)),
                           ("best_sad", "<u4", (2, 85)), ("best_mv", "<u4", (2, 85))])

_lib = None


def library_path():
    return os.path.join(_HERE, "libsvtme_cuda.so")


def load_library():
    """Load libsvtme_cuda.so; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        raise SvtMeError(f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                         "(there is no CPU fallback)")
    lib = C.CDLL(path)
    lib.svtme_default_params.restype = None
    lib.svtme_default_params.argtypes = [C.POINTER(SvtMeParams)] + [C.c_int32] * 6
    lib.svtme_create.restype = C.c_int32
    lib.svtme_create.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.svtme_destroy.restype = None
    lib.svtme_destroy.argtypes = [C.c_void_p]
    lib.svtme_last_error.restype = C.c_char_p
    lib.svtme_last_error.argtypes = [C.c_void_p]
    lib.svtme_picture_upload.restype = C.c_int32
    lib.svtme_picture_upload.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
    lib.svtme_picture_upload_16bit.restype = C.c_int32
    lib.svtme_picture_upload_16bit.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p]
    lib.svtme_picture_upload_peer.restype = C.c_int32
    lib.svtme_picture_upload_peer.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
    lib.svtme_slot_export.restype = C.c_int32
    lib.svtme_slot_export.argtypes = [C.c_void_p, C.c_int32, C.POINTER(SvtMeIpcSlot)]
    lib.svtme_picture_upload_ipc.restype = C.c_int32
    lib.svtme_picture_upload_ipc.argtypes = [C.c_void_p, C.c_int32, C.POINTER(SvtMeIpcSlot)]
    lib.svtme_picture_upload_device.restype = C.c_int32
    lib.svtme_picture_upload_device.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32]
    lib.svtme_pictures_upload.restype = C.c_int32
    lib.svtme_pictures_upload.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.c_int32]
    lib.svtme_pictures_upload_device.restype = C.c_int32
    lib.svtme_pictures_upload_device.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.c_int32]
    lib.svtme_picture_download.restype = C.c_int32
    lib.svtme_picture_download.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_size_t,
                                           C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    for name in ("svtme_motion_estimate", "svtme_motion_estimate_async", "svtme_fetch_results"):
        f = getattr(lib, name)
        f.restype = C.c_int32
        f.argtypes = [C.c_void_p, C.POINTER(SvtMeJob), C.c_int32]
    lib.svtme_motion_estimate_static.restype = C.c_int32
    lib.svtme_motion_estimate_static.argtypes = [C.c_void_p, C.POINTER(SvtMeJob)]
    lib.svtme_rc_me_distortion.restype = C.c_uint32
    lib.svtme_rc_me_distortion.argtypes = [C.c_void_p]
    lib.svtme_ois.restype = C.c_int32
    lib.svtme_ois.argtypes = [C.c_void_p, C.POINTER(SvtMeOisJob), C.c_int32]
    lib.svtme_zz_sad.restype = C.c_int32
    lib.svtme_zz_sad.argtypes = [C.c_void_p, C.POINTER(SvtMeZzJob), C.c_int32]
    lib.svtme_synchronize.restype = C.c_int32
    lib.svtme_synchronize.argtypes = [C.c_void_p]
    lib.svtme_synchronize_uploads.restype = C.c_int32
    lib.svtme_synchronize_uploads.argtypes = [C.c_void_p]
    lib.svtme_join.restype = C.c_int32
    lib.svtme_join.argtypes = [C.c_void_p]
    lib.svtme_stream.restype = C.c_void_p
    lib.svtme_stream.argtypes = [C.c_void_p]
    lib.svtme_kernel_launches.restype = C.c_int64
    lib.svtme_kernel_launches.argtypes = [C.c_void_p]
    lib.svtme_set_profiling.restype = C.c_int32
    lib.svtme_set_profiling.argtypes = [C.c_void_p, C.c_int32]
    lib.svtme_kernel_times.restype = C.c_int32
    lib.svtme_kernel_times.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_int32]
    lib.svtme_lcu_session_create.restype = C.c_int32
    lib.svtme_lcu_session_create.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.svtme_lcu_session_destroy.restype = None
    lib.svtme_lcu_session_destroy.argtypes = [C.c_void_p]
    lib.svtme_lcu_last_error.restype = C.c_char_p
    lib.svtme_lcu_last_error.argtypes = [C.c_void_p]
    lib.svtme_lcu_motion_estimate.restype = C.c_int32
    lib.svtme_lcu_motion_estimate.argtypes = [C.c_void_p, C.POINTER(SvtMeParams), C.POINTER(SvtMeLcuPicture), C.POINTER(SvtMeLcuPicture),
                                              C.POINTER(SvtMeLcuPicture), C.c_int32, C.c_void_p]
    lib.svtme_lcu_session_stats.restype = None
    lib.svtme_lcu_session_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    _lib = lib
    return lib


def default_params(width, height, num_lists=2, temporal_layer_index=1, hierarchical_levels=3, is_ref=1):
    p = SvtMeParams()
    load_library().svtme_default_params(C.byref(p), width, height, num_lists, temporal_layer_index,
                                        hierarchical_levels, is_ref)
    return p


class MeContext:
    """One GPU context: picture slots in HBM + batched motion estimation (mirrors the reference's
    per-picture ME stage: PA reference pictures in, me_results[sb][pu] out)."""

    def __init__(self, width, height, num_slots, device=0):
        self.lib = load_library()
        self.width, self.height, self.num_slots = width, height, num_slots
        self.n_sb = ((width + 63) // 64) * ((height + 63) // 64)
        self._h = C.c_void_p()
        rc = self.lib.svtme_create(C.byref(self._h), device, width, height, num_slots)
        if rc != 0:
            msg = self.lib.svtme_last_error(self._h).decode() if self._h else "no CUDA device / bad geometry"
            if self._h:
                self.lib.svtme_destroy(self._h)
                self._h = C.c_void_p()
            raise SvtMeError(f"svtme_create failed (0x{rc & 0xffffffff:08x}): {msg}")

    def close(self):
        if self._h:
            self.lib.svtme_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc != 0:
            raise SvtMeError(f"{what} failed (0x{rc & 0xffffffff:08x}): {self.lib.svtme_last_error(self._h).decode()}")

    def upload(self, slot, luma):
        """luma: C-contiguous uint8 numpy array (height, width) in host memory (pinned or pageable)."""
        assert luma.dtype == np.uint8 and luma.shape == (self.height, self.width)
        self._check(self.lib.svtme_picture_upload(self._h, slot, luma.ctypes.data, luma.strides[0]), "svtme_picture_upload")

    def upload16(self, slot, luma16, bit_inc=None):
        """luma16: uint16 (height, width) samples of a 10-bit picture; bit_inc: optional uint8 (height, width) array that
        receives the 2-bit remainder plane (valid after synchronize())."""
        assert luma16.dtype == np.uint16 and luma16.shape == (self.height, self.width) and luma16.strides[1] == 2
        if bit_inc is not None:
            assert bit_inc.dtype == np.uint8 and bit_inc.shape == (self.height, self.width) and bit_inc.flags.c_contiguous
        self._check(self.lib.svtme_picture_upload_16bit(self._h, slot, luma16.ctypes.data, luma16.strides[0] // 2,
                                                        bit_inc.ctypes.data if bit_inc is not None else None),
                    "svtme_picture_upload_16bit")

    def upload_peer(self, slot, peer, peer_slot):
        """Replicate picture `peer_slot` of another MeContext of this process (peer-to-peer copy + local pyramid)."""
        self._check(self.lib.svtme_picture_upload_peer(self._h, slot, peer._h, peer_slot), "svtme_picture_upload_peer")

    def export_slot(self, slot):
        """Token (bytes) another PROCESS passes to upload_ipc to replicate pictures uploaded into this slot."""
        t = SvtMeIpcSlot()
        self._check(self.lib.svtme_slot_export(self._h, slot, C.byref(t)), "svtme_slot_export")
        return bytes(t)

    def upload_ipc(self, slot, token):
        t = token if isinstance(token, SvtMeIpcSlot) else SvtMeIpcSlot.from_buffer_copy(token)
        self._check(self.lib.svtme_picture_upload_ipc(self._h, slot, C.byref(t)), "svtme_picture_upload_ipc")

    def upload_ptr(self, slot, host_ptr, stride):
        self._check(self.lib.svtme_picture_upload(self._h, slot, host_ptr, stride), "svtme_picture_upload")

    def upload_device(self, slot, dev_ptr, stride):
        self._check(self.lib.svtme_picture_upload_device(self._h, slot, dev_ptr, stride), "svtme_picture_upload_device")

    def upload_ptr_batch(self, slots, host_ptrs, stride):
        """n host-to-device copies + one pyramid launch (svtme_pictures_upload)."""
        n = len(slots)
        self._check(self.lib.svtme_pictures_upload(self._h, n, (C.c_int32 * n)(*slots), (C.c_void_p * n)(*host_ptrs), stride),
                    "svtme_pictures_upload")

    def upload_device_batch(self, slots, dev_ptrs, stride):
        """One pyramid launch for len(slots) pictures that already live in device memory (svtme_pictures_upload_device)."""
        n = len(slots)
        sl = (C.c_int32 * n)(*slots)
        pt = (C.c_void_p * n)(*dev_ptrs)
        self._check(self.lib.svtme_pictures_upload_device(self._h, n, sl, pt, stride), "svtme_pictures_upload_device")

    def download_plane(self, slot, level):
        pad = (68, 32, 16)[level]
        w, h = (self.width >> level) + 2 * pad, (self.height >> level) + 2 * pad
        out = np.empty((h, w), dtype=np.uint8)
        s, r, p = C.c_int32(), C.c_int32(), C.c_int32()
        self._check(self.lib.svtme_picture_download(self._h, slot, level, out.ctypes.data, out.nbytes,
                                                    C.byref(s), C.byref(r), C.byref(p)), "svtme_picture_download")
        assert (s.value, r.value, p.value) == (w, h, pad)
        return out

    def make_jobs(self, specs, want_debug=False, out_ptrs=None, sb_rows=None, sb_range=None):
        """specs: list of (cur_slot, ref0_slot, ref1_slot, SvtMeParams). sb_rows = (begin, end): search only those SB rows
        of every picture; sb_range = (begin, end): only those SBs in raster order (results/debug then hold just the SBs of
        that range). Returns (jobs array, results, debug)."""
        n = len(specs)
        jobs = (SvtMeJob * n)()
        sb_cols = (self.width + 63) // 64
        n_sb = (sb_range[1] - sb_range[0]) if sb_range is not None else self.n_sb if sb_rows is None else (sb_rows[1] - sb_rows[0]) * sb_cols
        # every job's rows are [n_sb][its own num_pus]: one array when the batch is uniform (85 or 209), else a list
        pus = [int(sp[3].num_pus) for sp in specs]
        if out_ptrs is not None:
            results = None
        elif len(set(pus)) <= 1:
            results = np.zeros((n, n_sb, pus[0] if pus else SQUARE_PU_COUNT), dtype=PU_RESULT_DTYPE)
        else:
            results = [np.zeros((n_sb, k), dtype=PU_RESULT_DTYPE) for k in pus]
        debug = np.zeros((n, n_sb), dtype=SB_DEBUG_DTYPE) if want_debug else None
        for i, (cur, r0, r1, prm) in enumerate(specs):
            jobs[i].cur_slot, jobs[i].ref0_slot, jobs[i].ref1_slot = cur, r0, (r1 if r1 is not None else r0)
            if sb_range is not None:
                jobs[i].sb_row_begin, jobs[i].sb_row_end, jobs[i].range_in_sbs = sb_range[0], sb_range[1], 1
            elif sb_rows is not None:
                jobs[i].sb_row_begin, jobs[i].sb_row_end = sb_rows
            jobs[i].params = prm
            jobs[i].results = results[i].ctypes.data if out_ptrs is None else out_ptrs[i]
            jobs[i].debug = debug[i].ctypes.data if want_debug else None
        return jobs, results, debug

    def motion_estimate(self, specs, want_debug=False, sb_rows=None, sb_range=None):
        jobs, results, debug = self.make_jobs(specs, want_debug, sb_rows=sb_rows, sb_range=sb_range)
        self._check(self.lib.svtme_motion_estimate(self._h, jobs, len(specs)), "svtme_motion_estimate")
        return results, debug

    def motion_estimate_static(self, spec, want_debug=False, sb_rows=None):
        """svtme_motion_estimate_static: one picture, no device allocation inside the call (static scratch array)."""
        jobs, results, debug = self.make_jobs([spec], want_debug, sb_rows=sb_rows)
        self._check(self.lib.svtme_motion_estimate_static(self._h, C.byref(jobs[0])), "svtme_motion_estimate_static")
        return results, debug

    def zz_sad(self, pairs):
        """svtme_zz_sad for (cur_slot, prev_slot) pairs: arrays [len(pairs)][num_sb] of zz cost and non-moving index."""
        n = len(pairs)
        zz, nm = np.zeros((n, self.n_sb), np.uint8), np.zeros((n, self.n_sb), np.uint8)
        jobs = (SvtMeZzJob * n)()
        for i, (c, p) in enumerate(pairs):
            jobs[i].cur_slot, jobs[i].prev_slot = c, p
            jobs[i].zz_cost, jobs[i].non_moving_index = zz[i].ctypes.data, nm[i].ctypes.data
        self._check(self.lib.svtme_zz_sad(self._h, jobs, n), "svtme_zz_sad")
        return zz, nm

    def ois(self, specs, out=None):
        """svtme_ois for (slot, SvtMeOisParams, me_sad or None) triples: (cand [n][num_sb][85][18] uint32 byte images of
        OisCandidate_t, total [n][num_sb][85] uint8). out: optional (cand, total) arrays to fill, e.g. views of pinned memory."""
        n = len(specs)
        if out is None:
            cand, total = np.zeros((n, self.n_sb, 85, 18), np.uint32), np.zeros((n, self.n_sb, 85), np.uint8)
        else:
            cand, total = out[0][:n], out[1][:n]
            assert cand.dtype == np.uint32 and cand.shape == (n, self.n_sb, 85, 18) and cand.flags.c_contiguous
            assert total.dtype == np.uint8 and total.shape == (n, self.n_sb, 85) and total.flags.c_contiguous
        jobs, keep = (SvtMeOisJob * n)(), []
        for i, (slot, prm, me) in enumerate(specs):
            jobs[i].slot, jobs[i].params = slot, prm
            if me is not None:
                me = np.ascontiguousarray(me, dtype=np.uint32)
                assert me.shape == (self.n_sb, 85)
                keep.append(me)
                jobs[i].me_sad = me.ctypes.data
            jobs[i].cand, jobs[i].total = cand[i].ctypes.data, total[i].ctypes.data
        self._check(self.lib.svtme_ois(self._h, jobs, n), "svtme_ois")
        return cand, total

    def motion_estimate_async(self, jobs):
        self._check(self.lib.svtme_motion_estimate_async(self._h, jobs, len(jobs)), "svtme_motion_estimate_async")

    def fetch(self, jobs):
        self._check(self.lib.svtme_fetch_results(self._h, jobs, len(jobs)), "svtme_fetch_results")

    def synchronize(self):
        self._check(self.lib.svtme_synchronize(self._h), "svtme_synchronize")

    def synchronize_uploads(self):
        self._check(self.lib.svtme_synchronize_uploads(self._h), "svtme_synchronize_uploads")

    def join(self):
        self._check(self.lib.svtme_join(self._h), "svtme_join")

    @property
    def stream(self):
        return self.lib.svtme_stream(self._h)

    @property
    def kernel_launches(self):
        return self.lib.svtme_kernel_launches(self._h)

    def set_profiling(self, enable):
        self._check(self.lib.svtme_set_profiling(self._h, 1 if enable else 0), "svtme_set_profiling")

    def kernel_times(self, reset=False):
        """{kind: (milliseconds, launches)} accumulated since the last reset (synchronises the stream)."""
        ms = (C.c_double * len(KERNEL_KINDS))()
        n = (C.c_int64 * len(KERNEL_KINDS))()
        self._check(self.lib.svtme_kernel_times(self._h, ms, n, 1 if reset else 0), "svtme_kernel_times")
        return {k: (ms[i], n[i]) for i, k in enumerate(KERNEL_KINDS)}


class MeLcuSession:
    """SB-granular facade (include/svtme_lcu.h): the call shape of the reference's MotionEstimateLcu
    (EbMotionEstimation.c:6152) -- one call per SB, from any thread -- served from whole-picture GPU launches."""

    def __init__(self, width, height, num_slots=16, device=0):
        self.lib = load_library()
        self.width, self.height = width, height
        self.n_sb = ((width + 63) // 64) * ((height + 63) // 64)
        self._h = C.c_void_p()
        rc = self.lib.svtme_lcu_session_create(C.byref(self._h), device, width, height, num_slots)
        if rc != 0:
            self._h = C.c_void_p()
            raise SvtMeError(f"svtme_lcu_session_create failed (0x{rc & 0xffffffff:08x}): no CUDA device / bad geometry / < 3 slots")

    def close(self):
        if self._h:
            self.lib.svtme_lcu_session_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @staticmethod
    def _pic(number, luma):
        assert luma.dtype == np.uint8 and luma.strides[1] == 1
        return SvtMeLcuPicture(int(number), luma.ctypes.data, luma.strides[0], 0)

    def motion_estimate_sb(self, params, cur, ref0, ref1, sb_index):
        """cur / ref0 / ref1: (picture_number, uint8 luma array) pairs (ref1 None for P pictures). Returns the SB's
        PU_RESULT row ([num_pus])."""
        row = np.zeros(params.num_pus, dtype=PU_RESULT_DTYPE)
        c, r0 = self._pic(*cur), self._pic(*ref0)
        r1 = self._pic(*ref1) if ref1 is not None else None
        rc = self.lib.svtme_lcu_motion_estimate(self._h, C.byref(params), C.byref(c), C.byref(r0),
                                                C.byref(r1) if r1 is not None else None, sb_index, row.ctypes.data)
        if rc != 0:
            raise SvtMeError(f"svtme_lcu_motion_estimate failed (0x{rc & 0xffffffff:08x}): {self.lib.svtme_lcu_last_error(self._h).decode()}")
        return row

    def stats(self):
        a, b, c = C.c_int64(), C.c_int64(), C.c_int64()
        self.lib.svtme_lcu_session_stats(self._h, C.byref(a), C.byref(b), C.byref(c))
        return {"pictures": a.value, "uploads": b.value, "rows": c.value}
