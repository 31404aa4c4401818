"""Integer acceptance thresholds: the Metropolis test of ``MCstep`` (py_mc/mc.py:175-184) in the
form the kernels evaluate.

The reference accepts a trial iff ``dE < 0.0`` or ``random() <= exp(-dE/kT)``.  The kernels draw a
32-bit integer ``x32`` per site, i.e. ``u = x32 * 2**-32``.  Scaling by 2**32 is exact in float64, so
``u <= p  <=>  x32 <= floor(p * 2**32)``; ``dE < 0`` (and ``p >= 1``) map to 2**32-1 = always accept.
``p`` is computed here, on the host, with the same ``math.exp`` call as the reference, which is what
makes device acceptance bit-exact with the reference for equal draws.
"""
import math

import numpy as np

ALWAYS = 0xFFFFFFFF


def threshold(dE, kT):
    if dE < 0.0:
        return ALWAYS
    p = math.exp(-dE / kT)
    return min(int(math.floor(p * 4294967296.0)), ALWAYS)


def table(dEs, kT):
    """uint32 thresholds for a sequence of class energies dEs at temperature kT."""
    return np.array([threshold(float(d), float(kT)) for d in dEs], dtype=np.uint32)
