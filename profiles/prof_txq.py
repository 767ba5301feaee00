"""Profiling driver (ncu target): a few representative transform + quantisation launches over 2 stacked 1080p
residual planes, device buffers. argv: list of WxH/type cases (default: 4x4/0 8x8/0 16x16/3 32x32/0 64x64/0 16x64/0)."""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")
cases = sys.argv[1:] or ["4x4/0", "8x8/0", "16x16/3", "32x32/0", "64x64/0", "16x64/0"]
planes = int(os.environ.get("TXQ_PLANES", "8"))
res = bench.txq_residual(11, planes)
dev = torch.from_numpy(res).cuda()
npix = res.size
out = [torch.empty(npix, dtype=torch.int32, device="cuda") for _ in range(3)]
eob = torch.empty(npix // 16, dtype=torch.int16, device="cuda")
en = torch.empty(npix // 16, dtype=torch.int64, device="cuda")
ctx = pkg.TxqContext(0)
q = pkg.SvtTxqQuant.from_q10(bench.txq_quant_params())
for rep in range(2):
    for c in cases:
        wh, t = c.split("/")
        w, h = (int(v) for v in wh.split("x"))
        ctx.fwd_txfm_quant_device(dev.data_ptr(), bench.TXQ_W, bench.TXQ_W, planes * bench.TXQ_H, w, h, int(t), q,
                                  out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), eob.data_ptr(), en.data_ptr())
    ctx.synchronize()
print("ok", int(out[1].sum().item()), ctx.kernel_launches())
ctx.close()
