"""Reader of tests/golden/txq_golden.npz (made by tests/golden/make_txq_golden.py from the compiled reference)."""
import os

import numpy as np

import txq_common as T

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "txq_golden.npz")


def blocks():
    """Yields dicts: w, h, t, bit_depth, qindex, q10, residual (h, w), coeff/qcoeff/dqcoeff (ph*pw,), eob, energy."""
    g = np.load(PATH)
    ro = co = 0
    for i, (w, h, t, bd, qi) in enumerate(g["cases"]):
        w, h, t = int(w), int(h), int(t)
        pw, ph = T.packed(w, h)
        n = pw * ph
        yield dict(w=w, h=h, t=t, bit_depth=int(bd), qindex=int(qi), q10=g["q10"][i],
                   residual=g["residual"][ro:ro + w * h].reshape(h, w), coeff=g["coeff"][co:co + n],
                   qcoeff=g["qcoeff"][co:co + n], dqcoeff=g["dqcoeff"][co:co + n], eob=int(g["eob"][i]),
                   energy=int(g["energy"][i]))
        ro += w * h
        co += n
