"""Picture replication between GPUs (needs a box with >= 2 GPUs; skipped otherwise): the in-process peer path and the
inter-process IPC path, both checked against the oracle's pyramids / ME results."""
import os
import subprocess
import sys

import numpy as np
import pytest

import me_common as M

pytestmark = pytest.mark.gpu


def _gpus():
    import torch
    return torch.cuda.device_count()


def test_upload_peer_in_process(pkg):
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    meo = M.load_oracle()
    w, h = 416, 240
    fr = M.synth_video(w, h, 4, seed=55)
    a, b = pkg.MeContext(w, h, 3, device=0), pkg.MeContext(w, h, 3, device=1)
    for i in range(3):
        a.upload(i, fr[i])
        b.upload_peer(i, a, i)                          # ordered behind a's upload on the device, no host sync in between
    a.upload(0, fr[3])                                  # rewrite of a picture b may still be copying: must wait for the copy
    b.synchronize(); a.synchronize()
    for i in range(3):
        want = M.build_pyramid(meo, "meo", fr[i])
        for lvl in range(3):
            assert np.array_equal(b.download_plane(i, lvl), want.planes[lvl]), (i, lvl)
    assert np.array_equal(a.download_plane(0, 0), M.build_pyramid(meo, "meo", fr[3]).planes[0])
    p = pkg.default_params(w, h, 2, 1, 3, 1)
    res, _ = b.motion_estimate([(1, 0, 2, p)])
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr[:3]]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[1], pyr[0], pyr[2], want_debug=False)
    assert M.results_diff(res[0], want) == []
    with pytest.raises(pkg.SvtMeError):
        b.upload_peer(0, b, 1)                          # a context is not its own peer
    b.close(); a.close()


def test_upload_ipc_two_processes():
    if _gpus() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(M.ROOT, "tests", "peer_ipc_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "peer_ipc_worker ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
