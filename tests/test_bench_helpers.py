"""Host-side helpers of bench.py (no GPU): the algorithmic work model behind `roofline` / `roofline_step` and the mini-GOP job list."""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

pkg = importlib.import_module("svt-av1-1_b200")


def test_fullpel_work_model():
    # 64 x 64 positions x 64 x 32 source bytes (even rows only): DESIGN.md section 3.1
    assert bench.fullpel_byte_sads_per_sb_list() == 4096 * 2048
    bench.SEARCH_AREA = 200                                # --search-area: the reference searches MIN(area, 127)
    try:
        assert bench.fullpel_byte_sads_per_sb_list() == 127 * 127 * 2048
    finally:
        bench.SEARCH_AREA = None


def test_hme_work_model_follows_the_layer_multipliers():
    # 1080p, hierarchical levels 3: level-0 regions 48 x 24 scaled by 200 / 140 / 100 / 70 % (EbDefinitions.h:2980-2996)
    want = {0: 4 * (96 * 48 * 128 + 16 * 16 * 512 + 8 * 8 * 2048), 1: 4 * (67 * 33 * 128 + 16 * 16 * 512 + 8 * 8 * 2048),
            2: 4 * (48 * 24 * 128 + 16 * 16 * 512 + 8 * 8 * 2048), 3: 4 * (33 * 16 * 128 + 16 * 16 * 512 + 8 * 8 * 2048)}
    for tl, w in want.items():
        p = pkg.default_params(1920, 1080, 2, tl, 3, 1)
        assert bench.hme_byte_sads_per_sb_list(p) == w, tl


def test_gop_jobs_cover_every_picture_once():
    jobs = bench.gop_jobs()
    assert len(jobs) == bench.N_PICS - 1
    assert sorted(j[0] for j in jobs) == list(range(1, bench.N_PICS))
    for poc, tl, r0, r1 in jobs:
        assert 0 <= r0 < bench.N_PICS and 0 <= r1 < bench.N_PICS and r0 != poc and r1 != poc
        assert tl == (0 if poc % 8 == 0 else 1 if poc % 4 == 0 else 2 if poc % 2 == 0 else 3)


def test_fullpel_ext_lane_tables():
    """The transpose-reduction network of fullpel_ext.cu replayed on sets (tools/check_fx_lanes.py): every PU of every
    class is issued once per tile by a lane that holds all 16 quads' keys of exactly that PU, and lane_pu() names it."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location(
        "check_fx_lanes", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "check_fx_lanes.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.check()


def test_ois_specs_follow_the_job_list(pkg):
    """bench.py's `ois` key: one svtme_ois spec per picture of a mini-GOP -- its own slot, its temporal layer (layer 3 is not a
    reference), the ME SADs of the 85 square PUs taken from the me_results rows (also when the rows hold 209 PUs)."""
    import numpy as np
    jobs = bench.gop_jobs()
    rows = np.zeros((len(jobs), 6, 209), pkg.PU_RESULT_DTYPE)
    rows["dist0"] = np.arange(len(jobs) * 6 * 209, dtype=np.uint32).reshape(len(jobs), 6, 209)
    specs = bench.ois_specs(pkg, jobs, rows, 8, 8)
    assert [s[0] for s in specs] == [j[0] for j in jobs[8:16]]
    layers = sorted(s[1].temporal_layer_index for s in specs)
    assert layers == [0, 1, 2, 2, 3, 3, 3, 3]
    for (slot, prm, me), j in zip(specs, range(8, 16)):
        assert prm.slice_type == 2 and prm.is_used_as_reference == (1 if prm.temporal_layer_index < 3 else 0) and prm.input_resolution == 2
        assert me.dtype == np.uint32 and me.shape == (6, 85) and me.flags.c_contiguous and np.array_equal(me, rows[j]["dist0"][:, :85])
    assert len(bench.ois_specs(pkg, jobs, rows, 112, 8)) == 8 and len(bench.ois_specs(pkg, jobs, rows, 120, 8)) == 0


def test_ois_cpu_side_runs_the_reference_search():
    """bench.py's ois.cpu: one layer-0 and one layer-2 picture through the reference's OpenLoopIntraSearchLcu (or the oracle port)."""
    import numpy as np
    old = bench.W, bench.H
    bench.W, bench.H = 256, 192
    try:
        jobs = bench.gop_jobs(17)
        frames = bench.synth_sequence(17, seed=2, w=256, h=192)
        rows = np.zeros((len(jobs), 12, 85), pkg.PU_RESULT_DTYPE)
        rows["dist0"] = 700
        cpu = bench.ois_cpu_ms(frames, jobs, rows)
    finally:
        bench.W, bench.H = old
    assert cpu["kind"] in ("reference", "port") and cpu["cores"] == 1
    assert set(cpu["ms_per_picture"]) == {"layer0", "layer2"} and all(v > 0 for v in cpu["ms_per_picture"].values())
