"""GPU-backed mirror of randomfield.powertools (reference: randomfield/powertools.py).

Table-sized work (validation, smoothing filter, k bounds: O(500) values) stays in Python as
in the reference; every full-array pass runs in librf_b200.so.
"""
from __future__ import print_function, division

import os.path

import numpy as np

from . import _lib
from . import transform


def get_k_bounds(data, spacing, packed=True):
    """powertools.py:16-24."""
    nx, ny, nz = transform.expanded_shape(data, packed=packed)
    k0 = (2 * np.pi) / spacing
    k_min = k0 / max(nx, ny, nz)
    k_max = k0 * np.sqrt(3) / 2
    return k_min, k_max


def create_ksq_grids(data, spacing, packed):
    """powertools.py:27-37 (sparse O(nx+ny+nz) host grids)."""
    nx, ny, nz = transform.expanded_shape(data, packed=packed)
    lambda0 = spacing / (2 * np.pi)
    kx = np.fft.fftfreq(nx, lambda0)
    ky = np.fft.fftfreq(ny, lambda0)
    kz = np.fft.fftfreq(nz, lambda0)
    if packed:
        kz = kz[:nz // 2 + 1]
    return np.meshgrid(kx ** 2, ky ** 2, kz ** 2, sparse=True, indexing='ij')


def _inplace(data, fn):
    work = np.ascontiguousarray(data)
    fn(work)
    if work is not data:
        data[...] = work
    return data


def fill_with_log10k(data, spacing, packed=True):
    """Fill an array with values of log10(k) (powertools.py:40-61); [0,0,0] is -inf."""
    nx, ny, nz = transform.expanded_shape(data, packed=packed)
    return _inplace(data, lambda w: _lib.check(_lib.lib().rf_stage_fill_log10k(
        nx, ny, nz, int(bool(packed)), _lib.dtype_code(w.dtype), float(spacing), _lib.ptr(w))))


def validate_power(power):
    """powertools.py:64-82."""
    if not isinstance(power, np.ndarray):
        raise ValueError('Invalid type for power: {0}.'.format(type(power)))
    if power.dtype.names is None or 'k' not in power.dtype.names or 'Pk' not in power.dtype.names:
        raise ValueError('Missing required fields "k", "Pk" in power.')
    if not np.all(np.isfinite(power['k'])):
        raise ValueError('Power spectrum has some invalid values of k.')
    if not np.all(np.isfinite(power['Pk'])):
        raise ValueError('Power spectrum has some invalid values of P(k).')
    if not np.array_equal(power['k'], np.unique(power['k'])):
        raise ValueError('Power spectrum k values are not strictly increasing.')
    if power['k'][0] <= 0:
        raise ValueError('Power spectrum min(k) is <= 0.')
    if np.any(power['Pk'] < 0):
        raise ValueError('Power values P(k) are not all non-negative.')
    return power


def filter_power(power, sigma, out=None):
    """Gaussian filtering P(k) -> P(k) exp(-(k sigma)^2) (powertools.py:85-122)."""
    if sigma < 0:
        raise ValueError('Invalid smoothing sigma: {0}.'.format(sigma))
    if out is None:
        out = np.copy(power)
    elif out is not power:
        validate_power(power)
        if out.shape != power.shape:
            raise ValueError('Output power has wrong shape: {0}.'.format(out.shape))
        out[:] = power
    if sigma > 0:
        out['Pk'] *= np.exp(-(power['k'] * sigma) ** 2)
    return out


def sigma_table(power, shape, spacing):
    """The (k, sigma) knots of powertools.py:139-154 for an expanded grid shape, with the
    reference's range check.  sigma = N3 * sqrt(Pk / (2 Vbox))."""
    validate_power(power)
    nx, ny, nz = shape
    N3 = nx * ny * nz
    Vbox = N3 * spacing ** 3
    power_k_min, power_k_max = np.min(power['k']), np.max(power['k'])
    if power_k_min <= 0:
        raise ValueError('Power uses min(k) <= 0: {0}.'.format(power_k_min))
    k0 = (2 * np.pi) / spacing
    data_k_min, data_k_max = k0 / max(nx, ny, nz), k0 * np.sqrt(3) / 2
    if power_k_min > data_k_min or power_k_max < data_k_max:
        raise ValueError(
            'Power k range [{0}:{1}] does not cover data k range [{2}:{3}].'
            .format(power_k_min, power_k_max, data_k_min, data_k_max))
    return np.ascontiguousarray(power['k'], dtype=np.float64), N3 * np.sqrt(power['Pk'] / (2 * Vbox))


def tabulate_sigmas(data, power, spacing, packed=True):
    """Replace log10(k) values with the corresponding sigmas (powertools.py:125-164):
    linear interpolation in log10 k, evaluated in float64 on the GPU, zero outside the table."""
    shape = transform.expanded_shape(data, packed=packed)
    k, sigma = sigma_table(power, shape, spacing)
    log10_k = np.ascontiguousarray(np.log10(k))
    sigma = np.ascontiguousarray(sigma, dtype=np.float64)
    return _inplace(data, lambda w: _lib.check(_lib.lib().rf_stage_tabulate(
        w.size, _lib.dtype_code(w.dtype), len(k), _lib.ptr(log10_k), _lib.ptr(sigma), _lib.ptr(w))))


def load_default_power(scaled_by_h=True):
    """The bundled P(k, z=0) (powertools.py:167-197): 500 rows, 1e-4 <= k <= 22 h/Mpc,
    stored as data/default_power.npy (float64 copy of the reference's table)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'data', 'default_power.npy')
    try:
        tab = np.load(path)
    except (IOError, OSError):
        raise RuntimeError('Unable to load default_power.dat')
    power = np.empty(len(tab), dtype=[('k', float), ('Pk', float)])
    power['k'], power['Pk'] = tab[:, 0], tab[:, 1]
    if scaled_by_h is False:
        from .cosmotools import create_cosmology
        cosmology = create_cosmology()
        power['k'] *= cosmology.h
        power['Pk'] /= cosmology.h ** 3
    return power
