"""Pins oracle/txq_oracle.c (forward transform, scan orders, quantisation) bit-exactly against the reference's
own objects (oracle/_ref/libsvtref.so). CPU only; skipped where the compiled reference is absent."""
import numpy as np
import pytest

import me_common as M
import txq_common as T


@pytest.fixture(scope="module")
def libs():
    M.build_oracle(ref=True)
    if not M.ref_available():
        pytest.skip("oracle/_ref/libsvtref.so not available")
    return T.oracle(), T.reference()


@pytest.mark.parametrize("w,h", T.TX_SIZES)
def test_scan_orders(libs, w, h):
    meo, ref = libs
    assert ref.ref_max_eob(w, h) == min(w, 32) * min(h, 32)
    for t in range(16):
        assert np.array_equal(T.scan(meo, "meo", w, h, t), T.scan(ref, "ref", w, h, t)), (w, h, t)


@pytest.mark.parametrize("w,h", T.TX_SIZES)
def test_fwd_txfm_all_types(libs, w, h):
    meo, ref = libs
    rng = np.random.default_rng(w * 100 + h)
    for t in range(16):
        if not T.type_supported(w, h, t):
            continue
        for amp in (1, 17, 255):                       # 8-bit residual range is +-255
            for trial in range(3):
                res = rng.integers(-amp, amp + 1, size=(h, w)).astype(np.int16)
                if trial == 2:
                    res[:] = amp * (1 if t % 2 else -1)   # flat block: DC only
                a, ea = T.fwd_txfm(meo, "meo", res, w, h, t)
                b, eb = T.fwd_txfm(ref, "ref", res, w, h, t)
                assert np.array_equal(a, b), (w, h, T.TX_TYPE_NAMES[t], amp, trial)
                assert ea == eb, (w, h, t)


@pytest.mark.parametrize("w,h", T.TX_SIZES)
def test_quantize_qindex_sweep(libs, w, h):
    meo, ref = libs
    rng = np.random.default_rng(w * 7 + h)
    res = (rng.normal(0, 40, size=(h, w))).clip(-255, 255).astype(np.int16)
    coeff, _ = T.fwd_txfm(meo, "meo", res, w, h, 0)
    for qindex in list(range(0, 256, 5)) + [255]:
        q10 = T.quant_params(ref, qindex)
        for t in (0, 10, 11):
            if not T.type_supported(w, h, t):
                continue
            a = T.quantize(meo, "meo", coeff, w, h, t, q10)
            b = T.quantize(ref, "ref", coeff, w, h, t, q10)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2], (w, h, qindex, t)
