"""GPU-backed mirror of randomfield.random (reference: randomfield/random.py)."""
from __future__ import print_function, division

import os

import numpy as np

from . import _lib
from . import transform


def _entropy_seed():
    return int.from_bytes(os.urandom(8), 'little')


def randomize(data, seed=None):
    """Randomize data by multiplying existing sigma values by normal deviates
    (random.py:12-29): imag <- real, then every scalar is scaled by an independent N(0,1)
    deviate.  The deviates come from a counter-based Philox4x32-10 stream keyed by ``seed``
    (the reference draws from MT19937; the two agree in distribution, and like the reference
    the same seed always gives the same array).  ``seed=None`` seeds from OS entropy."""
    transform.scalar_type(data.dtype)
    if seed is None:
        seed = _entropy_seed()
    work = np.ascontiguousarray(data)
    _lib.check(_lib.lib().rf_stage_randomize(work.size, _lib.dtype_code(work.dtype),
                                             int(seed) & 0xFFFFFFFFFFFFFFFF, _lib.ptr(work)))
    if work is not data:
        data[...] = work
    return data
