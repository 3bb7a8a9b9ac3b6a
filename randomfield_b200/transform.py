"""GPU-backed mirror of randomfield.transform (reference: randomfield/transform.py).

Same names, arguments and error behaviour; the arithmetic of ``symmetrize`` and
``Plan.execute`` runs in librf_b200.so on the B200 (no numpy.fft / pyFFTW, no CPU fallback).
"""
from __future__ import print_function, division

import numpy as np

from . import _lib


def allocate(shape, dtype, use_pyfftw=True):
    """Un-initialised C-order array (transform.py:10-43).  ``use_pyfftw`` is accepted for
    signature compatibility; alignment comes from numpy's allocator (64 bytes for large
    arrays), transfers to the GPU do not need SIMD alignment."""
    return np.empty(shape, dtype, order='C')


def expanded_shape(data, packed=False):
    """transform.py:46-60."""
    nx, ny, nz = data.shape
    if nx % 2 or ny % 2:
        raise ValueError('First two dimensions of array must be even.')
    if packed:
        if nz % 2 == 0:
            raise ValueError('Last dimension of packed array must be odd.')
        nz = 2 * (nz - 1)
    else:
        if nz % 2:
            raise ValueError('Last dimension of unpacked array must be even.')
    return nx, ny, nz


def _sctype(rep):
    if isinstance(rep, np.ndarray):
        return rep.dtype.type
    try:
        return np.dtype(rep).type
    except TypeError:
        return None


def scalar_type(complex_type):
    """transform.py:63-75."""
    ct = _sctype(complex_type)
    if ct == np.csingle:
        return np.single
    elif ct == np.complex128:
        return np.float64
    elif ct == np.clongdouble:
        return np.longdouble
    raise ValueError('Invalid complex_type: {0}.'.format(ct))


def complex_type(scalar_type):
    """transform.py:78-90."""
    st = _sctype(scalar_type)
    if st == np.single:
        return np.csingle
    elif st == np.float64:
        return np.complex128
    elif st == np.longdouble:
        return np.clongdouble
    raise ValueError('Invalid scalar_type: {0}.'.format(st))


def is_hermitian(data, packed=False, rtol=1e-08, atol=1e-08):
    """transform.py:93-111 (a checker; vectorised on the host)."""
    nx, ny, nz = expanded_shape(data, packed=packed)
    z_idx = np.array([0, nz // 2]) if packed else np.arange(nz // 2 + 1)
    ix = np.arange(nx // 2 + 1)
    iy = np.arange(ny // 2 + 1)
    a = data[np.ix_(ix, iy, z_idx)]
    b = np.conj(data[np.ix_((nx - ix) % nx, (ny - iy) % ny, (nz - z_idx) % nz)])
    return bool(np.allclose(a, b, rtol, atol))


def symmetrize(data, packed=False):
    """Symmetrize a complex 3D field so that its inverse FFT is real valued
    (transform.py:114-158), in place, on the GPU."""
    nx, ny, nz = expanded_shape(data, packed=packed)
    work = np.ascontiguousarray(data)
    _lib.check(_lib.lib().rf_stage_symmetrize(nx, ny, nz, int(bool(packed)),
                                              _lib.dtype_code(work.dtype), _lib.ptr(work)))
    if work is not data:
        data[...] = work
    return data


class Plan(object):
    """A plan for fast Fourier transforms on a single buffer (transform.py:161-315).

    ``data_in`` / ``data_out`` are host arrays with the reference's shapes, dtypes and
    memory-sharing rules; ``execute`` moves ``data_in`` to the device, transforms it with the
    CUDA kernels and writes ``data_out``.  numpy.fft normalisation.
    """

    def __init__(self, shape, dtype_in=None, data_in=None,
                 overwrite=True, inverse=True, packed=True, use_pyfftw=True):
        try:
            nx, ny, nz = shape
        except (TypeError, ValueError):
            raise ValueError('Expected 3D shape.')
        if nx % 2 or ny % 2 or nz % 2:
            raise ValueError('All shape dimensions must be even.')
        if data_in is not None:
            if not isinstance(data_in, np.ndarray):
                raise ValueError('Invalid type for data_in: {0}.'.format(type(data_in)))
            dtype_in = data_in.dtype
        dtype_in = _sctype(dtype_in) if dtype_in is not None else None
        if dtype_in is None:
            raise ValueError('Invalid dtype_in: {0}.'.format(dtype_in))

        if packed:
            if inverse:
                shape_in = (nx, ny, nz // 2 + 1)
                if not issubclass(dtype_in, np.complexfloating):
                    raise ValueError('Invalid dtype_in for inverse packed transform ' +
                                     '(should be complex): {0}.'.format(dtype_in))
                dtype_out = scalar_type(dtype_in)
                shape_out = (nx, ny, nz + 2) if overwrite else shape
            else:
                shape_in = (nx, ny, nz + 2) if overwrite else shape
                if not issubclass(dtype_in, np.floating):
                    raise ValueError('Invalid dtype_in for forward packed transform ' +
                                     '(should be floating): {0}.'.format(dtype_in))
                dtype_out = complex_type(dtype_in)
                shape_out = (nx, ny, nz // 2 + 1)
        else:
            if not issubclass(dtype_in, np.complexfloating):
                raise ValueError('Expected complex dtype_in for transform: {0}.'.format(dtype_in))
            shape_in = shape_out = shape
            dtype_out = dtype_in
        _lib.dtype_code(dtype_in)   # rejects long double: no such arithmetic on the device

        if data_in is not None:
            if data_in.shape != shape_in:
                raise ValueError('data_in has wrong shape {0}, expected {1}.'
                                 .format(data_in.shape, shape_in))
            self.data_in = data_in
            self.nbytes_allocated = 0
        else:
            self.data_in = allocate(shape_in, dtype_in, use_pyfftw=use_pyfftw)
            self.nbytes_allocated = self.data_in.nbytes
        if overwrite:
            if packed:
                self.data_out = self.data_in.view(dtype_out).reshape(shape_out)
                if inverse:
                    self.data_out_padded = self.data_out
                    self.data_out = self.data_out[:, :, :nz]
                else:
                    self.data_in_padded = self.data_in
                    self.data_in = self.data_in[:, :, :nz]
            else:
                self.data_out = self.data_in
        else:
            self.data_out = allocate(shape_out, dtype_out, use_pyfftw=use_pyfftw)
            self.nbytes_allocated += self.data_out.nbytes

        self.use_pyfftw = False
        self.shape = shape
        self.inverse = inverse
        self.packed = packed
        self.overwrite = overwrite

    def create_reverse_plan(self, reuse_output=True, overwrite=True):
        """transform.py:278-301."""
        inverse = not self.inverse
        if reuse_output:
            if self.packed and self.inverse and overwrite:
                if not self.overwrite:
                    raise RuntimeError('Cannot re-use output for reverse plan.')
                data_in = self.data_out_padded
            else:
                data_in = self.data_out
            dtype_in = None
        else:
            data_in = None
            dtype_in = self.data_out.dtype
        return Plan(shape=self.shape, dtype_in=dtype_in, data_in=data_in,
                    overwrite=overwrite, inverse=inverse, packed=self.packed,
                    use_pyfftw=self.use_pyfftw)

    def execute(self):
        """transform.py:303-315, on the GPU."""
        nx, ny, nz = self.shape
        src = np.ascontiguousarray(self.data_in)
        dst = np.empty(self.data_out.shape, dtype=self.data_out.dtype)
        _lib.check(_lib.lib().rf_plan_execute(
            nx, ny, nz, _lib.dtype_code(src.dtype), int(bool(self.inverse)), int(bool(self.packed)),
            _lib.ptr(src), _lib.ptr(dst)))
        self.data_out[:] = dst
        return self.data_out
