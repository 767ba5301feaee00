"""Generates tests/golden/txq_golden.npz from the REFERENCE itself (oracle/_ref/libsvtref.so built from
/root/reference by oracle/Makefile): for every transform size x every tx type the reference has kernels for,
two residual blocks through av1_fwd_txfm2d_WxH (AVX2/SSE4.1 build, via Av1EstimateTransform) and
aom_highbd_quantize_b* with the reference's own quantiser tables (av1_build_quantizer) at a cycling qindex.
Run in the build container only:  python tests/golden/make_txq_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import txq_common as T  # noqa: E402

QINDEX = (12, 60, 124, 200, 255)


def main():
    ref = T.reference()
    rng = np.random.default_rng(20261104)
    cases, res_all, coeff_all, q_all, dq_all, eob_all, en_all, q10_all = [], [], [], [], [], [], [], []
    k = 0
    for (w, h) in T.TX_SIZES:
        for t in range(16):
            if not T.type_supported(w, h, t):
                continue
            bd = 10 if k % 3 == 2 else 8
            q10 = T.quant_params(ref, QINDEX[k % len(QINDEX)], bd)
            plane = T.synth_residual(rng, h, 2 * w, bd, "noise" if k % 4 == 3 else "mixed")
            for b in range(2):
                res = np.ascontiguousarray(plane[:, b * w:(b + 1) * w])
                c, e = T.fwd_txfm(ref, "ref", res, w, h, t)
                qc, dq, eob = T.quantize(ref, "ref", c, w, h, t, q10)
                cases.append((w, h, t, bd, QINDEX[k % len(QINDEX)]))
                res_all.append(res.reshape(-1)); coeff_all.append(c.reshape(-1)); q_all.append(qc.copy()); dq_all.append(dq.copy())
                eob_all.append(eob); en_all.append(e); q10_all.append(q10)
            k += 1
    out = os.path.join(HERE, "txq_golden.npz")
    np.savez_compressed(out, cases=np.array(cases, np.int32), residual=np.concatenate(res_all), coeff=np.concatenate(coeff_all),
                        qcoeff=np.concatenate(q_all), dqcoeff=np.concatenate(dq_all), eob=np.array(eob_all, np.uint16),
                        energy=np.array(en_all, np.uint64), q10=np.array(q10_all, np.int16))
    print("wrote", out, len(cases), "blocks", os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()


def qtables():
    """tests/golden/txq_qtables.npz: the luma quantiser lanes of EVERY qindex (0..255) for 8- and 10-bit, straight from the
    reference's av1_build_quantizer tables: q10[depth][qindex] = zbin dc,ac, round dc,ac, quant dc,ac, quant_shift dc,ac, dequant dc,ac."""
    ref = T.reference()
    tab = np.zeros((2, 256, 10), np.int16)
    for d, bd in enumerate((8, 10)):
        for qi in range(256):
            tab[d, qi] = T.quant_params(ref, qi, bd)
    out = os.path.join(HERE, "txq_qtables.npz")
    np.savez_compressed(out, q10=tab)
    print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__" and "--qtables" in sys.argv:
    qtables()
