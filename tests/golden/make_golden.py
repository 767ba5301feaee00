"""Generate tests/golden/*.npz from the UNMODIFIED reference objects (oracle/_ref/libsvtref.so).

Run in the build container (needs /root/reference to compile the reference):
    python tests/golden/make_golden.py
Each fixture holds the seeded input luma planes, the SvtMeParams/MeoParams tuple, the reference's
padded pyramid checksums and its MotionEstimateLcu outputs (me_results image + per-SB
p_sb_best_sad / p_sb_best_mv), so the oracle and the CUDA path can be checked where the reference
tree is absent (the GPU box).
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import me_common as M  # noqa: E402

# name, width, height, seed, synth kwargs, list of (num_lists, temporal_layer, hier_levels, is_ref, overrides)
CASES = [
    ("g416x240_b_tl1", 416, 240, 11, dict(noise=3.0, max_shift=24), (2, 1, 3, 1, {})),
    ("g416x240_p_tl0", 416, 240, 12, dict(noise=2.0, max_shift=40), (1, 0, 3, 1, {})),
    ("g288x192_b_tl3_nonref", 288, 192, 13, dict(noise=4.0, max_shift=12), (2, 3, 3, 0, {})),
    ("g320x192_b_sameref", 320, 192, 14, dict(noise=3.0, max_shift=30), (2, 0, 3, 1, {"same_ref_poc": 1})),
    ("g192x128_b_flat", 192, 128, 15, dict(noise=0.0, max_shift=6, contrast=0.05), (2, 2, 3, 1, {})),
    ("g256x192_b_contrast", 256, 192, 16, dict(noise=6.0, max_shift=48, contrast=3.0), (2, 1, 3, 1, {})),
    ("g160x64_p_small", 160, 64, 17, dict(noise=3.0, max_shift=8), (1, 0, 3, 1, {})),
    ("g256x128_b_nohme", 256, 128, 18, dict(noise=3.0, max_shift=20), (2, 1, 3, 1, {"enable_hme_flag": 0})),
    ("g256x128_b_sa24", 256, 128, 19, dict(noise=3.0, max_shift=20), (2, 1, 3, 1, {"search_area_w": 24, "search_area_h": 17})),
    ("g256x128_b_nosubpel", 256, 128, 20, dict(noise=3.0, max_shift=20), (2, 1, 3, 1, {"use_subpel": 0})),
    # 209-PU mode (PIC_ALL_DEPTH_MODE: enc-mode 0-2 reference pictures): open_loop_me_fullpel_search_sblock + non-square PUs
    ("g416x240_b_ext209", 416, 240, 21, dict(noise=3.0, max_shift=24), (2, 1, 3, 1, {"num_pus": 209})),
    ("g256x192_p_ext209", 256, 192, 22, dict(noise=2.0, max_shift=40), (1, 0, 3, 1, {"num_pus": 209})),
    ("g256x128_b_ext209_fullpel", 256, 128, 23, dict(noise=5.0, max_shift=30, contrast=2.0), (2, 2, 3, 1, {"num_pus": 209, "use_subpel": 0})),
    ("g192x128_b_ext209_flat", 192, 128, 24, dict(noise=0.0, max_shift=6, contrast=0.05), (2, 0, 3, 1, {"num_pus": 209})),
    ("g160x136_b_ext209_sa", 160, 136, 25, dict(noise=3.0, max_shift=20), (2, 1, 3, 1, {"num_pus": 209, "search_area_w": 27, "search_area_h": 13})),
]


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).tobytes()) & 0xFFFFFFFF


def main():
    M.build_oracle(ref=True)
    ref = M.load_ref()
    only = sys.argv[1:]                   # optional: generate just the named cases (substring match)
    for name, w, h, seed, kw, (nl, tl, hl, is_ref, over) in CASES:
        if only and not any(o in name for o in only):
            continue
        frames = M.synth_video(w, h, 3, seed=seed, **kw)
        pyr = [M.build_pyramid(ref, "ref", f) for f in frames]
        prm = M.default_params(ref, "ref", w, h, num_lists=nl, tl=tl, hl=hl, is_ref=is_ref)
        for k, v in over.items():
            setattr(prm, k, v)
        res, dbg = M.motion_estimate(ref, "ref", prm, pyr[1], pyr[0], pyr[2] if nl == 2 else None)
        if nl == 1:                       # L1 MVs of a P picture are stale context memory in the reference
            res["xMvL1"] = 0
            res["yMvL1"] = 0
        out = os.path.join(HERE, name + ".npz")
        np.savez_compressed(
            out, frames=np.stack(frames), params=np.array(M.params_to_list(prm), dtype=np.int32),
            pyr_crc=np.array([[crc(p) for p in py.planes] for py in pyr], dtype=np.uint64),
            results=res, best_sad=dbg["best_sad"], best_mv=dbg["best_mv"],
            sa_origin_x=dbg["sa_origin_x"], sa_origin_y=dbg["sa_origin_y"])
        print(name, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
