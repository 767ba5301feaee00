"""Repeats the encoder-in-the-loop encode of one small clip (reference encoder with every MotionEstimateLcu call served by the
CUDA library) and counts abnormal exits / differing bitstreams. usage: python profiles/stress_encoder_in_loop.py N [mode] [w h]"""
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import me_common as M  # noqa: E402
from test_encoder_in_loop_gpu import _env, ENC  # noqa: E402

N = int(sys.argv[1])
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0
w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (192, 128)
n, seed = 10, 17
frames = M.synth_video(w, h, n, seed=seed, max_shift=8)
yuv = "/tmp/stress_in.yuv"
with open(yuv, "wb") as f:
    for x in frames:
        f.write(np.ascontiguousarray(x).tobytes())
        f.write(np.full((h // 2) * (w // 2) * 2, 128, np.uint8).tobytes())
lib = os.path.join(ROOT, "svt-av1-1_b200", "libsvtme_cuda.so")
bad, sigs = 0, set()
for i in range(N):
    r = subprocess.run([ENC, yuv, str(w), str(h), str(n), str(mode), "/tmp/stress.medump", "--bitstream", "/tmp/stress.obu", "--cuda-me", lib],
                       env=_env(), stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, text=True, timeout=600)
    if r.returncode != 0:
        bad += 1
        print(i, "exit", r.returncode, r.stderr[-300:].replace("\n", " | "), flush=True)
    else:
        sigs.add(hashlib.md5(open("/tmp/stress.obu", "rb").read()).hexdigest())
print(f"SVTME_TMA={os.environ.get('SVTME_TMA', 'default')} mode {mode} {w}x{h}: runs {N} failures {bad} distinct bitstreams {len(sigs)}")
