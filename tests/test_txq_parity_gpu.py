"""CUDA forward transform + quantisation (svttxq_* through the C-ABI) against the oracle and the golden vectors.
Bit-exact: integer work."""
import numpy as np
import pytest

import txq_common as T
import txq_golden as G

pytestmark = pytest.mark.gpu
KEYS = ("coeff", "qcoeff", "dqcoeff", "eob", "energy")


@pytest.fixture(scope="module")
def txq(pkg):
    with pkg.TxqContext(0) as ctx:
        yield ctx


@pytest.fixture(scope="module")
def q10s():
    """distinct quantiser parameter sets taken from the golden file (the reference's own tables)."""
    seen = {}
    for b in G.blocks():
        seen.setdefault((b["bit_depth"], b["qindex"]), b["q10"])
    return seen


def assert_same(got, want, ctx=""):
    for k in KEYS:
        if not np.array_equal(got[k], want[k]):
            bad = np.argwhere(got[k] != want[k])
            raise AssertionError(f"{ctx}: {k} differs at {bad[:4].tolist()} ({len(bad)} places)")


def test_golden_blocks(txq):
    n = 0
    for b in G.blocks():
        got = txq.fwd_txfm_quant(np.ascontiguousarray(b["residual"]), b["w"], b["h"], b["t"], b["q10"])
        ctx = f"{b['w']}x{b['h']} type {b['t']}"
        assert np.array_equal(got["coeff"][0], b["coeff"]), ctx
        assert np.array_equal(got["qcoeff"][0], b["qcoeff"]), ctx
        assert np.array_equal(got["dqcoeff"][0], b["dqcoeff"]), ctx
        assert got["eob"][0] == b["eob"] and got["energy"][0] == b["energy"], ctx
        n += 1
    assert n == 370


@pytest.mark.parametrize("size", T.TX_SIZES, ids=lambda s: f"{s[0]}x{s[1]}")
def test_every_type_against_oracle(txq, q10s, size):
    w, h = size
    meo = T.oracle()
    rng = np.random.default_rng(w * 100 + h)
    keys = sorted(q10s)
    for t in range(16):
        if not T.type_supported(w, h, t):
            with pytest.raises(Exception):
                txq.fwd_txfm_quant(np.zeros((h, w), np.int16), w, h, t, q10s[keys[0]])
            continue
        bd, qi = keys[(t + w) % len(keys)]
        res = T.synth_residual(rng, 3 * h, 7 * w, bd, "noise" if t % 2 else "mixed")     # 21 blocks: ragged last CTA
        assert_same(txq.fwd_txfm_quant(res, w, h, t, q10s[(bd, qi)]), T.txq_plane(meo, res, w, h, t, q10s[(bd, qi)]),
                    f"{w}x{h} type {t} bd {bd} q {qi}")


def _rebuild(eob, off, levels, scan, n_coef, dequant, ls):
    """What a consumer of the compact form does: levels back to raster qcoeff, dqcoeff = (|q| * dequant) >> log_scale with q's sign."""
    nb = len(eob)
    q = np.zeros((nb, n_coef), np.int32)
    for b in range(nb):
        q[b, scan[:eob[b]]] = levels[off[b]:off[b] + eob[b]]
    dq_mag = (np.abs(q).astype(np.int64) * np.where(np.arange(n_coef) == 0, dequant[0], dequant[1])) >> ls
    return q, (np.sign(q) * dq_mag).astype(np.int32)


@pytest.mark.parametrize("size", T.TX_SIZES, ids=lambda s: f"{s[0]}x{s[1]}")
def test_compact_return_path(txq, q10s, size):
    """svttxq_fwd_txfm_quant_compact: eob + the eob leading levels in scan order (int16) + their offsets reproduce the
    oracle's qcoeff, dqcoeff and eob for every size; several chunks (the plane exceeds 2 Mpixel for the larger cases)."""
    w, h = size
    meo = T.oracle()
    rng = np.random.default_rng(w * 7 + h)
    keys = [k for k in sorted(q10s) if k[0] == 8]
    for t in (0, 9, 3, 10, 11):
        if not T.type_supported(w, h, t):
            continue
        q10 = q10s[keys[(t + w + h) % len(keys)]]
        rows, cols = (9 * h, 11 * w) if t else (max(9 * h, 1088), 2048)          # DCT_DCT case: > 2 Mpixel -> 2 chunks
        res = T.synth_residual(rng, rows // h * h, cols // w * w, 8, "noise" if t % 2 else "mixed")
        want = T.txq_plane(meo, res, w, h, t, q10)
        eob, off, levels = txq.fwd_txfm_quant_compact(res, w, h, t, q10)
        assert np.array_equal(eob, want["eob"]), (w, h, t)
        assert off[0] == 0 and np.array_equal(np.diff(off.astype(np.int64)), eob.astype(np.int64)) and off[-1] == len(levels)
        q, dq = _rebuild(eob, off, levels, T.scan(meo, "meo", w, h, t), want["qcoeff"].shape[1], (q10[8], q10[9]), T.log_scale(w, h))
        assert np.array_equal(q, want["qcoeff"]) and np.array_equal(dq, want["dqcoeff"]), (w, h, t)


def test_compact_return_path_limits(txq, q10s, pkg):
    q10 = q10s[sorted(q10s)[0]]
    res = T.synth_residual(np.random.default_rng(1), 64, 64, 8, "noise")
    eob = np.empty(64, np.uint16)
    with pytest.raises(pkg.SvtMeError, match="0x80001000"):                  # levels buffer too small
        txq.fwd_txfm_quant_compact(res, 8, 8, 0, q10, levels=np.empty(3, np.int16), eob=eob)
    z = np.zeros((64, 64), np.int16)                                          # all-zero residual: eob 0 everywhere, no levels
    e, off, lv = txq.fwd_txfm_quant_compact(z, 16, 16, 0, q10)
    assert not e.any() and not off.any() and len(lv) == 0


def test_strided_plane_and_optional_outputs(txq, q10s):
    meo = T.oracle()
    rng = np.random.default_rng(77)
    big = T.synth_residual(rng, 128, 256 + 24)
    res = big[:, 8:8 + 256]                                   # stride 280, offset start
    q10 = q10s[sorted(q10s)[1]]
    want = T.txq_plane(meo, res, 16, 16, 3, q10)
    got = txq.fwd_txfm_quant(res, 16, 16, 3, q10)
    assert_same(got, want, "strided")
    only = txq.fwd_txfm_quant(res, 16, 16, 3, q10, want=("qcoeff", "eob"))
    assert set(only) == {"qcoeff", "eob"}
    assert np.array_equal(only["qcoeff"], want["qcoeff"]) and np.array_equal(only["eob"], want["eob"])


def test_edge_inputs(txq, q10s):
    meo = T.oracle()
    q10 = q10s[sorted(q10s)[0]]
    for (w, h) in [(4, 4), (16, 8), (32, 32), (64, 64), (16, 64)]:
        zero = np.zeros((h, 2 * w), np.int16)
        got = txq.fwd_txfm_quant(zero, w, h, 0, q10)
        assert not got["coeff"].any() and not got["qcoeff"].any() and not got["eob"].any() and not got["energy"].any()
        for v in (255, -255, 1023, -1023):                    # saturated flat residual: DC only
            flat = np.full((h, 2 * w), v, np.int16)
            assert_same(txq.fwd_txfm_quant(flat, w, h, 0, q10), T.txq_plane(meo, flat, w, h, 0, q10), f"flat {v} {w}x{h}")
        rng = np.random.default_rng(w + h)
        ext = rng.choice(np.array([-1023, 1023], np.int16), size=(h, 2 * w))     # worst-case 10-bit checkerboard noise
        assert_same(txq.fwd_txfm_quant(ext, w, h, 0, q10), T.txq_plane(meo, ext, w, h, 0, q10), f"extreme {w}x{h}")


def test_bad_parameters(txq, q10s, pkg):
    q10 = q10s[sorted(q10s)[0]]
    with pytest.raises(pkg.SvtMeError):
        txq.fwd_txfm_quant(np.zeros((10, 16), np.int16), 8, 8, 0, q10)       # height not a multiple of txh
    with pytest.raises(pkg.SvtMeError):
        txq.fwd_txfm_quant(np.zeros((64, 64), np.int16), 64, 8, 0, q10)      # not an AV1 transform size
    with pytest.raises(pkg.SvtMeError):
        txq.fwd_txfm_quant(np.zeros((32, 32), np.int16), 32, 32, 1, q10)     # no 32-point ADST


def test_1080p_plane_properties(txq, q10s):
    """Full-size planes (config 4 of BASELINE.json): size-independent properties + a sampled oracle check."""
    meo = T.oracle()
    rng = np.random.default_rng(1080)
    res = T.synth_residual(rng, 1088, 1920)
    q10 = q10s[(8, 124)] if (8, 124) in q10s else q10s[sorted(q10s)[0]]
    for (w, h, t) in [(8, 8, 0), (16, 16, 3), (32, 32, 0), (64, 64, 0), (4, 4, 9), (16, 8, 6)]:
        got = txq.fwd_txfm_quant(res, w, h, t, q10)
        nb = (1920 // w) * (1088 // h)
        assert got["coeff"].shape[0] == nb
        # (1) sampled block rows against the oracle
        for by in (0, (1088 // h) // 2, 1088 // h - 1):
            strip = res[by * h:(by + 1) * h]
            want = T.txq_plane(meo, strip, w, h, t, q10)
            bx = 1920 // w
            for k in KEYS:
                assert np.array_equal(got[k][by * bx:(by + 1) * bx], want[k]), (w, h, t, by, k)
        # (2) dequantised = quantised * dequant >> log_scale, sign-symmetric, DC/AC steps
        ls = T.log_scale(w, h)
        step = np.full(got["qcoeff"].shape[1], int(q10[9]), np.int64)
        step[0] = int(q10[8])
        mag = (np.abs(got["qcoeff"]).astype(np.int64) * step) >> ls
        assert np.array_equal(np.abs(got["dqcoeff"]).astype(np.int64), mag)
        assert np.array_equal(np.sign(got["dqcoeff"]) * (mag != 0), np.sign(got["qcoeff"]) * (mag != 0))
        # (3) eob = 1 + last scan position with a non-zero dequantised coefficient
        scan = T.scan(meo, "meo", w, h, t).astype(np.int64)
        nz = got["dqcoeff"][:, scan] != 0
        last = np.where(nz.any(axis=1), nz.shape[1] - np.argmax(nz[:, ::-1], axis=1), 0)
        assert np.array_equal(last.astype(np.uint16), got["eob"])
        # (4) flipping the residual turns FLIPADST into ADST (same coefficients)
    a = txq.fwd_txfm_quant(res[:16, :], 16, 16, 6, q10)                      # FLIPADST_FLIPADST
    flipped = np.ascontiguousarray(res[:16, :].reshape(16, -1, 16)[::-1, :, ::-1].reshape(16, -1))
    b = txq.fwd_txfm_quant(flipped, 16, 16, 3, q10)                          # ADST_ADST
    assert np.array_equal(a["coeff"], b["coeff"])


def test_device_entry_point_and_launch_count(txq, q10s, pkg):
    import torch
    meo = T.oracle()
    rng = np.random.default_rng(9)
    res = T.synth_residual(rng, 256, 512)
    q10 = q10s[sorted(q10s)[2]]
    w, h, t = 32, 16, 0
    nb, n = (512 // w) * (256 // h), 32 * 16
    d_res = torch.from_numpy(res).cuda()
    d_q = torch.empty((nb, n), dtype=torch.int32, device="cuda")
    d_eob = torch.empty(nb, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    before = txq.kernel_launches()
    txq.fwd_txfm_quant_device(d_res.data_ptr(), 512, 512, 256, w, h, t, q10, d_qcoeff=d_q.data_ptr(), d_eob=d_eob.data_ptr())
    txq.synchronize()
    assert txq.kernel_launches() == before + 1
    want = T.txq_plane(meo, res, w, h, t, q10)
    assert np.array_equal(d_q.cpu().numpy(), want["qcoeff"])
    assert np.array_equal(d_eob.cpu().numpy().view(np.uint16), want["eob"])


def test_every_qindex(txq):
    """BASELINE configs[3] "all qindex": every luma quantiser (qindex 0..255, 8- and 10-bit; tables taken from the reference,
    tests/golden/txq_qtables.npz) on three size classes (log_scale 0, 1, 2) against the oracle."""
    import os
    meo = T.oracle()
    tab = np.load(os.path.join(os.path.dirname(G.PATH), "txq_qtables.npz"))["q10"]
    rng = np.random.default_rng(255)
    planes = {8: T.synth_residual(rng, 64, 128, 8), 10: T.synth_residual(rng, 64, 128, 10, "noise")}
    for d, bd in enumerate((8, 10)):
        for qi in range(256):
            q10 = tab[d, qi]
            w, h, t = [(8, 8, 0), (32, 16, 0), (64, 64, 0), (16, 16, 3), (4, 4, 9)][qi % 5]
            assert_same(txq.fwd_txfm_quant(planes[bd], w, h, t, q10), T.txq_plane(meo, planes[bd], w, h, t, q10), f"qindex {qi} bd {bd} {w}x{h}")


@pytest.mark.parametrize("size", [(64, 64), (64, 32), (32, 64), (16, 64), (64, 16)], ids=lambda s: f"{s[0]}x{s[1]}")
def test_64_point_sizes_without_energy(txq, q10s, size):
    """Callers that do not want three_quad_energy (EncDec) take the pruned kernels: same coefficients, levels and eob."""
    w, h = size
    meo = T.oracle()
    rng = np.random.default_rng(64 + w + h)
    for kind, key in (("mixed", sorted(q10s)[0]), ("noise", sorted(q10s)[-1])):
        res = T.synth_residual(rng, 2 * h, 5 * w, key[0], kind)                  # 10 blocks: ragged last CTA
        want = T.txq_plane(meo, res, w, h, 0, q10s[key])
        got = txq.fwd_txfm_quant(res, w, h, 0, q10s[key], want=("coeff", "qcoeff", "dqcoeff", "eob"))
        assert set(got) == {"coeff", "qcoeff", "dqcoeff", "eob"}
        for k in got:
            assert np.array_equal(got[k], want[k]), (w, h, kind, k)
        only = txq.fwd_txfm_quant(res, w, h, 0, q10s[key], want=("qcoeff", "eob"))
        assert np.array_equal(only["qcoeff"], want["qcoeff"]) and np.array_equal(only["eob"], want["eob"])


@pytest.mark.parametrize("name", T.encoder_block_files())
def test_real_encoder_encdec_blocks(txq, name):
    """The CUDA path against what the unmodified reference ENCODER computed in its EncDec stage (tests/golden/txq_enc_*.npz,
    BASELINE configs[0]): residual blocks in; coefficients, levels, dequantised levels, eob, three-quad energy out. No oracle
    in between."""
    blocks = 0
    for w, h, t, q10, plane, want in T.encoder_block_groups(name):
        got = txq.fwd_txfm_quant(plane, w, h, t, q10)
        T.check_encoder_blocks(got, want, w, h, (name, w, h, t))
        blocks += len(want["eob"])
    assert blocks > 250


def test_64_point_sizes_on_a_second_device(pkg, q10s):
    """The dynamic shared-memory attribute of the 64-point kernels is per device: contexts on two GPUs of one process
    (the first 64-point launch of the process on device 1) must both work and agree."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    rng = np.random.default_rng(64)
    bd, qi = sorted(q10s)[0]
    with pkg.TxqContext(1) as c1, pkg.TxqContext(0) as c0:
        for w, h in ((64, 64), (32, 64), (64, 16)):
            res = T.synth_residual(rng, 2 * h, 3 * w, bd, "mixed")
            a = c1.fwd_txfm_quant(res, w, h, 0, q10s[(bd, qi)])
            b = c0.fwd_txfm_quant(res, w, h, 0, q10s[(bd, qi)])
            assert_same(a, b, f"{w}x{h} device 1 vs 0")
            eob, off, lv = c1.fwd_txfm_quant_compact(res, w, h, 0, q10s[(bd, qi)])
            assert np.array_equal(eob, b["eob"])
