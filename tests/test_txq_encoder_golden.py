"""Transform + quantisation oracle against blocks recorded from REAL encodes by the unmodified reference encoder (BASELINE
configs[0], "txfm/quant coeffs as golden"): tests/golden/txq_enc_*.npz hold residual blocks the encoder's EncDec stage
handed to Av1EstimateTransform (EbCodingLoop.c:661) together with the quantiser Av1QuantizeInvQuantize picked
(EbFullLoop.c:721-800) and what both calls wrote. CPU only; the CUDA twin is in tests/test_txq_parity_gpu.py."""
import pytest

import txq_common as T


@pytest.mark.parametrize("name", T.encoder_block_files())
def test_oracle_reproduces_encdec_blocks(name):
    lib = T.oracle()
    blocks, groups, nonzero, planes = 0, 0, 0, set()
    for w, h, t, q10, plane, want in T.encoder_block_groups(name):
        assert T.type_supported(w, h, t)
        got = T.txq_plane(lib, plane, w, h, t, q10)
        T.check_encoder_blocks(got, want, w, h, (name, w, h, t))
        blocks += len(want["eob"]); groups += 1; nonzero += int((want["eob"] > 0).sum())
        planes.add(tuple(q10))
    assert blocks > 250 and groups > 40 and nonzero > 100 and len(planes) >= 2       # luma and chroma quantisers
