"""Oracle against the ME stage of REAL encodes by the unmodified reference encoder (BASELINE configs[0]: the reference's
own CPU-runnable case). tests/golden/enc_*.npz hold, for every inter picture of a clip encoded through the reference's public
API, the settings its processes handed to MotionEstimateLcu and the MeCuResults_t rows that call produced
(tests/golden/make_encoder_golden.py). CPU only; the CUDA twin of this test is in tests/test_me_parity_gpu.py."""
import numpy as np
import pytest

import me_common as M


@pytest.fixture(scope="module")
def meo():
    M.build_oracle(ref=False)
    return M.load_oracle()


@pytest.mark.parametrize("name", M.encoder_cases())
def test_oracle_reproduces_encoder_me(meo, name):
    g = M.load_golden(name)
    frames = M.encoder_clip(g)
    pyr = {}
    n_b = n_p = 0
    for hd, rows in zip(g["hdr"], g["rows"]):
        poc, lists = int(hd[1]), int(hd[7])
        dflt = M.default_params(meo, "meo", int(hd[5]), int(hd[6]), lists, int(hd[8]), int(hd[9]), int(hd[10]))
        prm, bad = M.encoder_params(M.MeoParams, hd, dflt)
        assert bad == [], (name, poc, "library defaults differ from the encoder's per-picture settings", bad)
        refs = [int(hd[11]), int(hd[12])][:lists]
        for k in [poc] + refs:
            if k not in pyr:
                pyr[k] = M.build_pyramid(meo, "meo", frames[k])
        res, _ = M.motion_estimate(meo, "meo", prm, pyr[poc], pyr[refs[0]], pyr[refs[1]] if lists == 2 else None, want_debug=False)
        want = M.encoder_rows(hd, rows).reshape(res.shape)
        assert M.results_diff(res, want, lists) == [], (name, poc, refs)
        n_b += lists == 2
        n_p += lists == 1
    assert n_b > 0 and n_p > 0                      # the hierarchy has B pictures and the trailing P picture(s)


def test_encoder_goldens_cover_every_layer():
    for name in M.encoder_cases():
        g = M.load_golden(name)
        layers = {int(h[8]) for h in g["hdr"]}
        assert layers == set(range(int(g["geometry"][5]) + 1)), (name, layers)
        assert {int(h[38]) for h in g["hdr"]} == {int(g["geometry"][3])}
