"""Oracle (oracle/me_oracle.c) against the committed golden vectors that were produced by the
unmodified reference objects (tests/golden/make_golden.py). CPU only."""
import numpy as np
import pytest

import me_common as M


@pytest.fixture(scope="module")
def meo():
    M.build_oracle(ref=False)
    return M.load_oracle()


@pytest.mark.parametrize("name", M.golden_cases())
def test_oracle_matches_golden(meo, name):
    g = M.load_golden(name)
    prm = M.params_from_list(M.MeoParams, g["params"])
    pyr = [M.build_pyramid(meo, "meo", f) for f in g["frames"]]
    for i, py in enumerate(pyr):                                   # pyramid: pad + decimate
        assert [M.plane_crc(p) for p in py.planes] == [int(v) for v in g["pyr_crc"][i]]
    res, dbg = M.motion_estimate(meo, "meo", prm, pyr[1], pyr[0], pyr[2] if prm.num_lists == 2 else None)
    assert M.results_diff(res, g["results"], prm.num_lists) == []
    nl = prm.num_lists
    assert np.array_equal(dbg["best_sad"][:, :nl], g["best_sad"][:, :nl])
    assert np.array_equal(dbg["best_mv"][:, :nl], g["best_mv"][:, :nl])
    assert np.array_equal(dbg["sa_origin_x"][:, :nl], g["sa_origin_x"][:, :nl])
    assert np.array_equal(dbg["sa_origin_y"][:, :nl], g["sa_origin_y"][:, :nl])


def test_default_params_tables(meo):
    # M0 column of EbDefinitions.h:3320-3469 as pinned by the golden parameter dumps
    for name in M.golden_cases():
        g = M.load_golden(name)
        want = M.params_from_list(M.MeoParams, g["params"])
        got = M.default_params(meo, "meo", want.width, want.height, want.num_lists, want.temporal_layer_index,
                               want.hierarchical_levels, want.is_used_as_reference)
        for f in ("num_hme_regions_w", "num_hme_regions_h", "hme_l0_total_w", "hme_l0_total_h", "hme_l0_mult_x",
                  "hme_l0_mult_y", "enable_hme_level0", "enable_hme_level1", "enable_hme_level2"):
            assert getattr(got, f) == getattr(want, f), (name, f)
        assert got.num_pus == 85 and want.num_pus in (85, 209)     # 209 = the *_ext209 fixtures (PIC_ALL_DEPTH_MODE pictures)


def test_oracle_rejects_unsupported(meo):
    prm = M.default_params(meo, "meo", 128, 128)
    fr = M.synth_video(128, 128, 2, seed=3)
    pyr = [M.build_pyramid(meo, "meo", f) for f in fr]
    prm.num_pus = 100                                             # only 85 (square) and 209 (all) exist
    with pytest.raises(RuntimeError):
        M.motion_estimate(meo, "meo", prm, pyr[1], pyr[0], pyr[0])
