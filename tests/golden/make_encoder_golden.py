"""tests/golden/enc_*.npz: the ME stage of REAL encodes by the unmodified reference encoder (BASELINE configs[0]).

oracle/_ref/enc_dump (oracle/enc_driver.c) drives the whole reference encoder through its public API
(eb_init_handle / eb_svt_enc_set_parameter / eb_svt_enc_send_picture / eb_svt_get_packet) on a synthetic yuv420p clip;
the MotionEstimateLcu wrapper in oracle/ref_driver.c records, for every SB of every inter picture, what the encoder's own
processes set up (slice type, layers, reference POCs, HME flags, MeContext search tables) and the MeCuResults_t rows it
produced. The clip itself is regenerated from its seed by the tests (CRCs stored here).
Run in the build container only:  python tests/golden/make_encoder_golden.py
"""
import os
import subprocess
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import me_common as M  # noqa: E402

CASES = [  # name, width, height, frames, enc_mode, hierarchical_levels, seed, synth kwargs
    # enc_mode 3 is the default preset and the highest this reference version accepts (MAX_ENC_PRESET 3); modes 0-2 switch
    # reference pictures to the 209-PU search, which is row (h) of DESIGN.md
    ("enc_416x240_m3", 416, 240, 30, 3, 3, 416, dict(max_shift=12)),
    # (this reference version accepts hierarchical_levels 3 only); CIF: the last SB row is half height
    ("enc_352x288_m3", 352, 288, 19, 3, 3, 352, dict(max_shift=6, noise=4.0)),
    # enc-mode 2: reference pictures are PIC_ALL_C_DEPTH_MODE (209 PUs), non-reference ones 85 -- a mixed clip;
    # enc-mode 0: every picture is PIC_ALL_DEPTH_MODE (EbPictureDecisionProcess.c:357-379)
    ("enc_256x192_m2", 256, 192, 19, 2, 3, 256, dict(max_shift=10)),
    ("enc_192x128_m0", 192, 128, 19, 0, 3, 192, dict(max_shift=8, noise=4.0)),
]
HDR = 48


def parse(path, w, h):
    raw = np.fromfile(path, dtype=np.uint8)
    pos, pics = 0, {}
    while pos < raw.size:
        hd = raw[pos:pos + 4 * HDR].view(np.int32)
        assert hd[0] == 0x4D454447, "bad record"
        pos += 4 * HDR
        poc, sb, first, num_pus, rowsz = int(hd[1]), int(hd[2]), int(hd[40]), int(hd[18]), int(hd[41])
        assert rowsz == 40 and (hd[5], hd[6]) == (w, h)
        pic = pics.setdefault(poc, {"hdr": hd.copy(), "rows": {}, "luma": None})
        assert np.array_equal(pic["hdr"][5:40], hd[5:40]), "per-picture settings changed between SBs"
        if first:
            pic["luma"] = raw[pos:pos + w * h].reshape(h, w).copy()
            pos += w * h
        pic["rows"][sb] = raw[pos:pos + num_pus * 40].copy()
        pos += num_pus * 40
    return pics


def chroma_planes(frames, i):
    """Non-flat synthetic chroma (ME reads luma only; this only gives the EncDec transform/quantisation dump chroma work)."""
    a = frames[i][::2, ::2].astype(np.int32)
    b = frames[(i + 1) % len(frames)][1::2, 1::2].astype(np.int32)
    cb = np.clip(128 + (a - 128) // 3, 0, 255).astype(np.uint8)
    cr = np.clip(128 - (b - 128) // 4, 0, 255).astype(np.uint8)
    return cb, cr


TXQ_HDR = 16


def parse_txq(path):
    """Records of oracle/ref_txq.c's EncDec wrappers -> flat arrays."""
    raw = np.fromfile(path, dtype=np.uint8)
    pos, hdr, q10, res, coeff, qc, dq = 0, [], [], [], [], [], []
    while pos < raw.size:
        hd = raw[pos:pos + 4 * TXQ_HDR].view(np.int32).copy()
        assert hd[0] == 0x54584647, "bad txq record"
        pos += 4 * TXQ_HDR
        w, h, n = int(hd[2]), int(hd[3]), int(hd[7])
        q10.append(raw[pos:pos + 20].view(np.int16).copy()); pos += 20
        res.append(raw[pos:pos + 2 * w * h].view(np.int16).copy()); pos += 2 * w * h
        for dst in (coeff, qc, dq):
            dst.append(raw[pos:pos + 4 * n].view(np.int32).copy()); pos += 4 * n
        hdr.append(hd)
    return dict(hdr=np.stack(hdr), q10=np.stack(q10), residual=np.concatenate(res), coeff=np.concatenate(coeff),
                qcoeff=np.concatenate(qc), dqcoeff=np.concatenate(dq))


def main():
    M.build_oracle(ref=True)
    exe = os.path.join(M.ORACLE_DIR, "_ref", "enc_dump")
    only = sys.argv[1:]                   # optional: generate just the named cases (substring match)
    for name, w, h, n, mode, hl, seed, kw in CASES:
        if only and not any(o in name for o in only):
            continue
        frames = M.synth_video(w, h, n, seed=seed, **kw)
        with tempfile.TemporaryDirectory() as td:
            yuv, dump = os.path.join(td, "in.yuv"), os.path.join(td, "out.medump")
            with open(yuv, "wb") as f:
                for i, x in enumerate(frames):
                    f.write(np.ascontiguousarray(x).tobytes())
                    for c in chroma_planes(frames, i):
                        f.write(np.ascontiguousarray(c).tobytes())
            txq_dump = os.path.join(td, "out.txqdump")
            subprocess.run([exe, yuv, str(w), str(h), str(n), str(mode), dump, "--txq-dump", txq_dump], check=True,
                           stderr=subprocess.DEVNULL, stdout=subprocess.DEVNULL)
            pics = parse(dump, w, h)
            txq = parse_txq(txq_dump)
        n_sb = ((w + 63) // 64) * ((h + 63) // 64)
        pocs = sorted(pics)
        hdrs, rows = [], []
        for poc in pocs:
            p = pics[poc]
            assert len(p["rows"]) == n_sb and p["luma"] is not None
            assert np.array_equal(p["luma"], frames[poc]), "the encoder's input picture is not the clip frame"
            hdrs.append(p["hdr"])
            r = np.stack([p["rows"][sb] for sb in range(n_sb)])
            if int(p["hdr"][7]) == 1:
                # P pictures: xMvL1 / yMvL1 are copies of whatever the calling ME thread's context held from an earlier SB
                # (EbMotionEstimation.c:7137-7140, DESIGN Q8) -- they change from run to run with thread scheduling, so
                # they are zeroed here to keep the fixture reproducible (the tests ignore them for P pictures anyway)
                v = r.view(M.PU_RESULT_DTYPE)
                v["xMvL1"] = 0
                v["yMvL1"] = 0
            rows.append(r)
        num_pus = {int(hd[18]) for hd in hdrs}
        # a picture's rows are [n_sb][its own num_pus (hdr[18])] x 40 bytes, zero-padded to the widest picture of the clip
        width = max(num_pus) * 40
        rows = [np.pad(r, ((0, 0), (0, width - r.shape[1]))) for r in rows]
        out = os.path.join(HERE, name + ".npz")
        np.savez_compressed(out, geometry=np.array([w, h, n, mode, seed, hl], np.int32), synth=np.array([repr(kw)]),
                            frame_crc=np.array([zlib.crc32(np.ascontiguousarray(x).tobytes()) & 0xFFFFFFFF for x in frames], np.uint32),
                            hdr=np.stack(hdrs), rows=np.stack(rows))
        print("wrote", out, len(pocs), "inter pictures,", sorted(num_pus), "PUs per SB,", os.path.getsize(out), "bytes")
        # EncDec forward transform + quantisation blocks of the same encode ("txq_" prefix: read by tests/test_txq_encoder_golden.py)
        out = os.path.join(HERE, "txq_" + name + ".npz")
        np.savez_compressed(out, **txq)
        combos = {(int(a[2]), int(a[3]), int(a[4]), int(a[5])) for a in txq["hdr"]}
        print("wrote", out, len(txq["hdr"]), "blocks,", len(combos), "(size, type, plane) combinations,", os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
