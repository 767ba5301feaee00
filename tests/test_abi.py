"""The C-ABI library builds, loads without a GPU and exports every symbol include/svtme.h declares.
No compute calls here (CPU box); error behaviour without a device is checked instead."""
import ctypes as C
import os
import re

import pytest

import me_common as M

HEADER = os.path.join(M.ROOT, "include", "svtme.h")


@pytest.fixture(scope="module")
def pkg_built(pkg):
    import __graft_entry__ as g
    if not os.path.exists(pkg.library_path()):
        g.build()
    return pkg


def declared_symbols(header=HEADER, prefix="svtme_"):
    txt = open(header).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(%s\w+)\s*\(" % prefix, txt)))


def test_header_symbols_exported(pkg_built):
    lib = pkg_built.load_library()
    names = declared_symbols()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/svtme.h but not exported"
    assert set(names) == set(pkg_built.EXPORTED_SYMBOLS)


def test_lcu_header_symbols_exported(pkg_built):
    lib = pkg_built.load_library()
    names = declared_symbols(os.path.join(M.ROOT, "include", "svtme_lcu.h"), "svtme_lcu_")
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/svtme_lcu.h but not exported"
    assert set(names) == set(pkg_built.LCU_EXPORTED_SYMBOLS)
    assert C.sizeof(pkg_built.SvtMeLcuPicture) == 24
    import torch
    if not torch.cuda.is_available():                      # no device: the facade fails too, it never falls back
        with pytest.raises(pkg_built.SvtMeError):
            pkg_built.MeLcuSession(128, 128)


def test_txq_header_symbols_exported(pkg_built):
    lib = pkg_built.load_library()
    names = declared_symbols(os.path.join(M.ROOT, "include", "svttxq.h"), "svttxq_")
    assert len(names) >= 9
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/svttxq.h but not exported"
    assert set(names) == set(pkg_built.TXQ_EXPORTED_SYMBOLS)
    assert C.sizeof(pkg_built.SvtTxqQuant) == 20


def test_txq_supported_matches_reference_kernels(pkg_built):
    import txq_common as T
    for (w, h) in T.TX_SIZES:
        for t in range(16):
            assert pkg_built.txq_supported(w, h, t) == T.type_supported(w, h, t), (w, h, t)
    assert not pkg_built.txq_supported(64, 8, 0) and not pkg_built.txq_supported(12, 12, 0)
    assert not pkg_built.txq_supported(8, 8, 16)


def test_txq_no_cpu_fallback(pkg_built):
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(pkg_built.SvtMeError):
            pkg_built.TxqContext(0)


def test_struct_layouts(pkg_built):
    # byte image of MeCuResults_t (EbMotionEstimationLcuResults.h:62-75) and the parameter block
    assert pkg_built.PU_RESULT_DTYPE.itemsize == 40
    assert C.sizeof(pkg_built.SvtMeParams) == 4 * 34
    assert C.sizeof(pkg_built.SvtMeJob) == 6 * 4 + 4 * 34 + 2 * 8 and pkg_built.SvtMeJob.params.offset == 24
    assert C.sizeof(M.MeoParams) == C.sizeof(pkg_built.SvtMeParams)
    assert pkg_built.SB_DEBUG_DTYPE.itemsize == 6 * 4 + 2 * 2 * 85 * 4


def test_default_params_match_oracle(pkg_built):
    meo = M.load_oracle()
    for (w, h) in [(416, 240), (1280, 720), (1920, 1080), (3840, 2160), (832, 480)]:
        for hl, tl in [(3, 0), (3, 1), (3, 3), (4, 0), (5, 2)]:
            a = pkg_built.default_params(w, h, 2, tl, hl, 1)
            b = M.default_params(meo, "meo", w, h, 2, tl, hl, 1)
            assert M.params_to_list(a) == M.params_to_list(b)


def test_bad_parameters_and_no_cpu_fallback(pkg_built):
    lib = pkg_built.load_library()
    h = C.c_void_p()
    assert lib.svtme_create(C.byref(h), 0, 100, 64, 2) & 0xFFFFFFFF == 0x80001005     # width not a multiple of 8
    assert lib.svtme_create(None, 0, 128, 64, 2) & 0xFFFFFFFF == 0x80001005
    import torch
    if not torch.cuda.is_available():
        # no device: creation fails loudly (EB_ErrorUndefined image), nothing is computed on the CPU
        with pytest.raises(pkg_built.SvtMeError):
            pkg_built.MeContext(128, 64, 2)
