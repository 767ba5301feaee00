"""Encoder in the loop: the UNMODIFIED reference encoder (oracle/_ref/libsvtref.so, every process, public API) encodes a clip
twice -- once stock, once with every MotionEstimateLcu call served by the CUDA library through its SB facade
(include/svtme_lcu.h; the hook is oracle/enc_driver.c --cuda-me, the binding INTEGRATION.md section 6 describes).
The ME rows the rest of the encoder sees and the AV1 bitstream it writes must be identical. Needs a B200 and the prebuilt
oracle/_ref (it travels to the GPU box; it cannot be built there)."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import me_common as M

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.join(M.ROOT, "tests", "golden"))
ENC = os.path.join(M.ORACLE_DIR, "_ref", "enc_dump")
ENC_PATCHED = os.path.join(M.ORACLE_DIR, "_ref", "enc_dump_patched")      # the encoder with integration/svt_cuda_me.patch applied


def _env():
    env = dict(os.environ)
    extra = ["/usr/local/cuda/lib64"]
    try:
        import nvidia.cuda_runtime as rt
        extra.insert(0, os.path.join(list(rt.__path__)[0], "lib"))
    except Exception:
        pass
    env["LD_LIBRARY_PATH"] = ":".join([env.get("LD_LIBRARY_PATH", "")] + extra).strip(":")
    return env


def _encode(td, tag, yuv, w, h, n, extra, mode=3, exe=ENC, env_extra=None):
    dump, obu = os.path.join(td, tag + ".medump"), os.path.join(td, tag + ".obu")
    env = _env()
    env.update(env_extra or {})
    cmd = [exe, yuv, str(w), str(h), str(n), str(mode), dump, "--bitstream", obu] + extra
    r = subprocess.run(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=600)
    if r.returncode < 0:
        # The encoder process died on a signal. Seen ONCE in ~190 hooked encodes (glibc "stack smashing detected" in the
        # multi-threaded encoder process, first minutes of a fresh box, enc-mode 0) and not reproduced in 80 consecutive runs
        # of the same clip (profiles/stress_encoder_in_loop.py), nor in 40 stock runs. One retry, loudly, so that such an
        # event does not hide the parity tests behind it under `pytest -x`; a second death fails the test.
        import warnings
        warnings.warn(f"{os.path.basename(exe)} {tag} died on signal {-r.returncode}: {r.stderr[-300:]!r}; retrying once")
        r = subprocess.run(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return dump, open(obu, "rb").read(), r.stderr


@pytest.mark.skipif(not os.path.exists(ENC), reason="oracle/_ref/enc_dump not built (needs /root/reference at build time)")
@pytest.mark.parametrize("w,h,n,seed,mode", [
    (416, 240, 20, 7, 3), (352, 288, 12, 11, 3),
    (256, 192, 12, 13, 2),      # enc-mode 2: reference pictures searched with 209 PUs, the others with 85 (mixed batches)
    (192, 128, 10, 17, 0),      # enc-mode 0: 209 PUs everywhere
])
def test_reference_encoder_with_cuda_me_is_bit_identical(pkg, tmp_path, w, h, n, seed, mode):
    from make_encoder_golden import parse
    frames = M.synth_video(w, h, n, seed=seed, max_shift=8)
    yuv = str(tmp_path / "in.yuv")
    with open(yuv, "wb") as f:
        for x in frames:
            f.write(np.ascontiguousarray(x).tobytes())
            f.write(np.full((h // 2) * (w // 2) * 2, 128, np.uint8).tobytes())
    d0, bs0, _ = _encode(str(tmp_path), "stock", yuv, w, h, n, [], mode)
    d1, bs1, log = _encode(str(tmp_path), "cuda", yuv, w, h, n, ["--cuda-me", pkg.library_path()], mode)
    p0, p1 = parse(d0, w, h), parse(d1, w, h)
    assert {int(v["hdr"][18]) for v in p0.values()} == {3: {85}, 2: {85, 209}, 0: {209}}[mode]
    assert sorted(p0) == sorted(p1) and len(p0) == n - 1            # every picture but the key frame is inter
    n_sb = ((w + 63) // 64) * ((h + 63) // 64)
    for poc in sorted(p0):
        lists = int(p0[poc]["hdr"][7])
        assert np.array_equal(p0[poc]["hdr"][5:40], p1[poc]["hdr"][5:40])
        for sb in range(n_sb):
            a = p0[poc]["rows"][sb].view(M.PU_RESULT_DTYPE)
            b = p1[poc]["rows"][sb].view(M.PU_RESULT_DTYPE)
            assert M.results_diff(b, a, lists) == [], (poc, sb)
    m = re.search(r"cuda-me pictures (\d+) uploads (\d+) rows (\d+)", log)
    assert m, log[-500:]
    pictures, uploads, rows = (int(v) for v in m.groups())
    assert pictures == n - 1 and rows == (n - 1) * n_sb             # one whole-picture launch sequence per inter picture
    assert n <= uploads <= n + 8                                    # each picture crosses PCIe once (16 resident slots)
    assert len(bs0) > 1000 and bs0 == bs1, (len(bs0), len(bs1))     # the encode does not change


@pytest.mark.skipif(not os.path.exists(ENC), reason="oracle/_ref/enc_dump not built (needs /root/reference at build time)")
def test_reference_encoder_with_cuda_txq_and_me_is_bit_identical(pkg, tmp_path):
    """Both kernel families in the loop: every 8-bit Av1EstimateTransform call (mode decision and EncDec) and every EncDec
    Av1QuantizeInvQuantize call goes to svttxq_fwd_txfm_quant, one block per call (the granularity of the reference's call
    sites -- this checks results, not speed), and ME goes through the SB facade. Same bitstream as the stock encoder."""
    w, h, n = 416, 240, 6
    frames = M.synth_video(w, h, n, seed=19, max_shift=8)
    from make_encoder_golden import chroma_planes
    yuv = str(tmp_path / "in.yuv")
    with open(yuv, "wb") as f:
        for i, x in enumerate(frames):
            f.write(np.ascontiguousarray(x).tobytes())
            for c in chroma_planes(frames, i):
                f.write(np.ascontiguousarray(c).tobytes())
    _, bs0, _ = _encode(str(tmp_path), "stock", yuv, w, h, n, [])
    _, bs1, log1 = _encode(str(tmp_path), "txq", yuv, w, h, n, ["--cuda-txq", pkg.library_path()])
    m = re.search(r"cuda-txq transforms (\d+) quantisations (\d+)", log1)
    assert m, log1[-500:]
    transforms, quants = int(m.group(1)), int(m.group(2))
    assert quants > 1000 and transforms >= quants
    assert len(bs0) > 1000 and bs0 == bs1, (len(bs0), len(bs1))
    _, bs2, log2 = _encode(str(tmp_path), "both", yuv, w, h, n, ["--cuda-txq", pkg.library_path(), "--cuda-me", pkg.library_path()])
    assert "cuda-me pictures %d " % (n - 1) in log2 and bs2 == bs0


@pytest.mark.skipif(not os.path.exists(ENC_PATCHED), reason="oracle/_ref/enc_dump_patched not built (needs /root/reference at build time)")
@pytest.mark.parametrize("w,h,n,seed,mode", [(416, 240, 20, 7, 3), (720, 480, 10, 5, 3), (256, 192, 12, 13, 2), (192, 128, 10, 17, 0)])
def test_patched_encoder_picture_level_drop_in_is_bit_identical(pkg, tmp_path, w, h, n, seed, mode):
    """The picture-level drop-in AS CODE: integration/svt_cuda_me.patch applied to the reference's MotionEstimationKernel (a scratch
    copy compiled into oracle/_ref/libsvtref_cuda.so). With SVT_CUDA_ME_LIB set, each ME segment calls
    svtme_lcu_motion_estimate_segment instead of running the SB loop; the bitstream must equal the stock encoder's. 720x480 has
    16-pixel-wide edge SBs (the any-width HME search)."""
    frames = M.synth_video(w, h, n, seed=seed, max_shift=8)
    yuv = str(tmp_path / "in.yuv")
    with open(yuv, "wb") as f:
        for x in frames:
            f.write(np.ascontiguousarray(x).tobytes())
            f.write(np.full((h // 2) * (w // 2) * 2, 128, np.uint8).tobytes())
    _, bs0, _ = _encode(str(tmp_path), "stock", yuv, w, h, n, [], mode)
    _, bs1, log = _encode(str(tmp_path), "patched", yuv, w, h, n, [], mode, exe=ENC_PATCHED,
                          env_extra={"SVT_CUDA_ME_LIB": pkg.library_path(), "SVT_CUDA_ME_STATS": "1"})
    m = re.search(r"svt_cuda_me: pictures (\d+) uploads (\d+) rows (\d+)", log)
    assert m, log[-800:]
    n_sb = ((w + 63) // 64) * ((h + 63) // 64)
    assert int(m.group(1)) == n - 1 and int(m.group(3)) == (n - 1) * n_sb and n <= int(m.group(2)) <= n + 8
    assert "me_calls 0 " in log                                     # MotionEstimateLcu itself never ran
    assert len(bs0) > 1000 and bs0 == bs1, (len(bs0), len(bs1))
