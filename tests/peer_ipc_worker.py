"""Two-process check of svtme_slot_export / svtme_picture_upload_ipc (launched by test_peer_replication_gpu.py under
torch.distributed.run, one rank per GPU, gloo for the host-side token exchange). Rank r uploads pictures r, r+2, ...
from host memory, every rank imports the others over NVLink, then both ranks motion-estimate the same B picture on
different SB-row ranges; rank 0 checks the union against the oracle."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import me_common as M  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
torch.cuda.set_device(rank)
w, h = 416, 240
frames = M.synth_video(w, h, 3, seed=123)
ctx = pkg.MeContext(w, h, 3, device=rank)
tokens = [None] * world
dist.all_gather_object(tokens, [ctx.export_slot(i) for i in range(3)])
for rep in range(2):                                    # second round rewrites the slots (ordering across rounds by barriers)
    for i in range(3):
        if i % world == rank:
            ctx.upload(i, frames[i] if rep == 0 else np.ascontiguousarray(frames[2 - i]))
    dist.barrier()                                      # every owner has ISSUED its uploads
    for i in range(3):
        if i % world != rank:
            ctx.upload_ipc(i, tokens[i % world][i])
    ctx.synchronize()
    dist.barrier()                                      # importers are done: owners may rewrite
meo = M.load_oracle()
for i in range(3):                                      # after round two slot i holds frames[2 - i] on every rank
    want = M.build_pyramid(meo, "meo", frames[2 - i])
    for lvl in range(3):
        assert np.array_equal(ctx.download_plane(i, lvl), want.planes[lvl]), (rank, i, lvl)
p = pkg.default_params(w, h, 2, 1, 3, 1)
rows = (0, 2) if rank == 0 else (2, 4)
res, _ = ctx.motion_estimate([(1, 0, 2, p)], sb_rows=rows)
parts = [None] * world
dist.all_gather_object(parts, res[0].tobytes())
ctx.close()
if rank == 0:
    got = np.concatenate([np.frombuffer(b, dtype=pkg.PU_RESULT_DTYPE).reshape(-1, 85) for b in parts])
    pyr = [M.build_pyramid(meo, "meo", frames[2 - i]) for i in range(3)]
    want, _ = M.motion_estimate(meo, "meo", M.convert_params(M.MeoParams, p), pyr[1], pyr[0], pyr[2], want_debug=False)
    assert M.results_diff(got, want) == []
    print("peer_ipc_worker ok")
dist.barrier()
dist.destroy_process_group()
