"""Encoder in the loop, timed: the UNMODIFIED reference encoder (oracle/_ref/enc_dump, every process, public API) on a
1920x1080 clip, stock and with every MotionEstimateLcu call served by libsvtme_cuda.so through the SB facade
(include/svtme_lcu.h, oracle/enc_driver.c --cuda-me). Prints a markdown table of encode wall clock and ME-stage time
(time inside MotionEstimateLcu summed over the encoder's ME threads) and checks that the two bitstreams are identical.
usage: python profiles/encoder_timing.py [frames] [enc_mode] > profiles/r02/encoder_timing.md   (needs a B200)"""
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import bench  # noqa: E402
from test_encoder_in_loop_gpu import ENC, ENC_PATCHED, _env  # noqa: E402

frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 3
w, h = 1920, 1080
lib = os.path.join(ROOT, "svt-av1-1_b200", "libsvtme_cuda.so")
frames = bench.synth_sequence(frames_n, seed=1)
rows = []
with tempfile.TemporaryDirectory() as td:
    yuv = os.path.join(td, "in.yuv")
    with open(yuv, "wb") as f:
        chroma = np.full((h // 2) * (w // 2) * 2, 128, np.uint8).tobytes()
        for x in frames:
            f.write(np.ascontiguousarray(x).tobytes())
            f.write(chroma)
    out = {}
    for tag, exe, extra, envx in (("stock (reference AVX2 ME on the host cores)", ENC, [], {}),
                                  ("hooked per SB (MotionEstimateLcu -> svtme_lcu_motion_estimate)", ENC, ["--cuda-me", lib], {}),
                                  ("patched per segment (integration/svt_cuda_me.patch -> svtme_lcu_motion_estimate_segment)", ENC_PATCHED, [], {"SVT_CUDA_ME_LIB": lib})):
        obu = os.path.join(td, "out.obu")
        env = _env()
        env["CUDA_MODULE_LOADING"] = "EAGER"             # kernels load with the context (before the timed encode), not at first launch
        env.update(envx)
        r = subprocess.run([exe, yuv, str(w), str(h), str(frames_n), str(mode), "-", "--bitstream", obu] + extra, env=env,
                           stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=3000)
        assert r.returncode == 0, r.stderr[-2000:]
        m = re.search(r"timing wall_s ([\d.]+) me_thread_s ([\d.]+) me_calls (\d+)", r.stderr)
        assert m, r.stderr[-1000:]
        out[tag] = (float(m.group(1)), float(m.group(2)), int(m.group(3)), open(obu, "rb").read())
        rows.append((tag, float(m.group(1)), float(m.group(2)), int(m.group(3))))
    same = all(out[r[0]][3] == out[rows[0][0]][3] for r in rows)
print(f"# Reference encoder, stock vs hooked: {w}x{h}, {frames_n} frames, enc-mode {mode}, {len(os.sched_getaffinity(0))} host cores, 1 B200\n")
print("| encoder | encode wall clock (s) | frames/s | time inside MotionEstimateLcu, all ME threads (s) | MotionEstimateLcu calls |")
print("|---|---|---|---|---|")
for tag, wall, me, calls in rows:
    print(f"| {tag} | {wall:.2f} | {frames_n / wall:.2f} | {me:.2f} | {calls} |")
print(f"\nbitstreams identical: {same} ({len(out[rows[0][0]][3])} bytes)")
print("\nThe encoder's other stages (mode decision, EncDec, filters, entropy coding) are the reference's CPU code in both runs; with the hook the")
print("ME threads spend their MotionEstimateLcu time waiting for whole-picture GPU launches (one picture in flight per ME thread). The patched")
print("encoder never calls MotionEstimateLcu (0 calls): its ME threads make one library call per ME segment.")
assert same
