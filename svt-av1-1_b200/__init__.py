"""svt-av1-1_b200 -- B200-native (sm_100a CUDA) open-loop motion estimation of SVT-AV1 behind a C-ABI, plus
the forward transform + quantisation kernels as the secondary family.

The product is ``libsvtme_cuda.so`` (sources in ``csrc/``, interface in ``include/svtme.h`` and
``include/svttxq.h``); this
package only carries a thin ctypes binding used by the tests and by ``bench.py``. There is no CPU
fallback: loading fails loudly when the library has not been built, and ``MeContext`` raises when no
CUDA device is present.
"""
from .binding import (MeContext, SvtMeError, SvtMeParams, SvtMeJob, SvtMeIpcSlot, SvtMeOisParams, PU_RESULT_DTYPE, SB_DEBUG_DTYPE, KERNEL_KINDS,  # noqa: F401
                      default_params, load_library, library_path, EXPORTED_SYMBOLS, LCU_EXPORTED_SYMBOLS, MeLcuSession,
                      SvtMeLcuPicture)
from .txq_binding import TxqContext, SvtTxqQuant, TXQ_EXPORTED_SYMBOLS, TX_TYPES, supported as txq_supported, packed_dims  # noqa: F401,E402
