"""The transform/quantisation oracle against the committed golden vectors (generated from the reference itself).
Runs anywhere: needs neither /root/reference nor a GPU."""
import numpy as np

import txq_common as T
import txq_golden as G


def test_oracle_matches_golden_blocks():
    meo = T.oracle()
    n = 0
    for b in G.blocks():
        c, e = T.fwd_txfm(meo, "meo", b["residual"], b["w"], b["h"], b["t"])
        assert np.array_equal(c.reshape(-1), b["coeff"]), (b["w"], b["h"], b["t"])
        assert e == b["energy"]
        qc, dq, eob = T.quantize(meo, "meo", c, b["w"], b["h"], b["t"], b["q10"])
        assert np.array_equal(qc, b["qcoeff"]) and np.array_equal(dq, b["dqcoeff"]) and eob == b["eob"], (b["w"], b["h"], b["t"])
        n += 1
    assert n == 370


def test_oracle_plane_equals_per_block():
    meo = T.oracle()
    rng = np.random.default_rng(5)
    res = T.synth_residual(rng, 64, 128)
    q10 = next(G.blocks())["q10"]
    for (w, h, t) in [(8, 8, 3), (16, 32, 0), (64, 64, 0), (4, 16, 9), (32, 8, 11)]:
        out = T.txq_plane(meo, res, w, h, t, q10)
        k = 0
        for by in range(64 // h):
            for bx in range(128 // w):
                blk = np.ascontiguousarray(res[by * h:(by + 1) * h, bx * w:(bx + 1) * w])
                c, e = T.fwd_txfm(meo, "meo", blk, w, h, t)
                qc, dq, eob = T.quantize(meo, "meo", c, w, h, t, q10)
                assert np.array_equal(out["coeff"][k], c.reshape(-1)) and np.array_equal(out["qcoeff"][k], qc)
                assert np.array_equal(out["dqcoeff"][k], dq) and out["eob"][k] == eob and out["energy"][k] == e
                k += 1


def test_qtables_fixture_shape_and_monotone_steps():
    import os
    tab = np.load(os.path.join(os.path.dirname(G.PATH), "txq_qtables.npz"))["q10"]
    assert tab.shape == (2, 256, 10)
    assert (np.diff(tab[:, :, 9].astype(np.int32), axis=1) >= 0).all()        # AC dequant step grows with qindex
    assert (tab[:, :, 8] > 0).all() and (tab[:, :, 6] > 0).all()              # DC dequant step, quant_shift positive
