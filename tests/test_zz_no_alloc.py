"""rf_density_noalloc (include/rf_b200.h): the density epilogue on device arrays with its scratch in
a static __device__ array, checked through the C-ABI against the existing entry point
rf_stage_density on the same data.  Collected last on purpose.
"""
import ctypes

import numpy as np
import pytest

from randomfield_b200 import _lib

pytestmark = pytest.mark.gpu

RF_NOALLOC_MAX_NZ = 4096    # include/rf_b200.h


def _tables(nz, seed):
    rng = np.random.RandomState(seed)
    growth = np.ascontiguousarray(np.linspace(1.0, 0.45, nz) * (1 + 0.01 * rng.rand(nz)))
    rho = np.ascontiguousarray(2.7e11 * (1 + 0.1 * rng.rand(nz)))
    return growth, rho


def _existing(delta, growth, rho, lognormal, sigma):
    """rf_stage_density: host array in / out, allocates its own device memory."""
    L = _lib.lib()
    want = delta.copy()
    _lib.check(L.rf_stage_density(want.size, want.shape[-1], _lib.dtype_code(want.dtype), int(lognormal),
                                  float(sigma), _lib.ptr(growth), _lib.ptr(rho), _lib.ptr(want)))
    return want


def _noalloc(torch, d_in, d_out, nz, dtype, growth, rho, lognormal, sigma, stream):
    L = _lib.lib()
    st = L.rf_density_noalloc(ctypes.c_void_p(d_in.data_ptr()), ctypes.c_void_p(d_out.data_ptr()),
                              d_in.numel(), nz, _lib.dtype_code(dtype), int(lognormal), float(sigma),
                              _lib.ptr(growth), _lib.ptr(rho), ctypes.c_void_p(stream.cuda_stream))
    _lib.check(st)


CASES = [
    # shape, dtype, lognormal, sigma (<= 0: measured on the device)
    ((24, 20, 96), np.float32, True, 0.0),
    ((10, 6, 14), np.float64, True, 1.9),      # smaller nz right after a larger one: stale table tails
    ((24, 20, 96), np.float32, False, 0.0),
    ((8, 4, RF_NOALLOC_MAX_NZ), np.float32, True, 0.0),
    ((12, 10, 6), np.float64, True, 0.0),
]


def test_noalloc_density_matches_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    dev = torch.device("cuda:0")
    stream = torch.cuda.Stream(device=dev)
    for rep in range(2):                       # the whole sequence twice: every call sees used scratch
        for ci, (shape, dtype, lognormal, sigma) in enumerate(CASES):
            rng = np.random.RandomState(100 * rep + ci)
            delta = (1.3 * rng.normal(size=shape)).astype(dtype)
            growth, rho = _tables(shape[-1], ci)
            want = _existing(delta, growth, rho, lognormal, sigma)
            d_in = torch.from_numpy(delta).to(dev)
            d_out = torch.empty_like(d_in)
            torch.cuda.synchronize()
            # two calls in a row on one stream: out of place, then in place on a copy
            _noalloc(torch, d_in, d_out, shape[-1], dtype, growth, rho, lognormal, sigma, stream)
            d_inplace = d_in.clone()
            torch.cuda.synchronize()
            _noalloc(torch, d_inplace, d_inplace, shape[-1], dtype, growth, rho, lognormal, sigma, stream)
            stream.synchronize()
            tol = 2e-5 if dtype == np.float32 else 1e-12     # tests/test_gpu_stage_api.py's density tolerance
            for got in (d_out.cpu().numpy(), d_inplace.cpu().numpy()):
                assert got.shape == want.shape and got.dtype == want.dtype
                assert np.all(np.isfinite(got))
                assert np.all(np.abs(got - want) <= tol * np.abs(want)), (rep, ci, np.max(np.abs(got / want - 1)))
            assert np.array_equal(d_in.cpu().numpy(), delta)   # the input array is not written


def test_noalloc_density_rejects_sizes_beyond_the_static_scratch():
    torch = pytest.importorskip("torch")
    L = _lib.lib()
    nz = RF_NOALLOC_MAX_NZ + 2
    d = torch.zeros(2 * nz, dtype=torch.float32, device="cuda:0")
    growth, rho = _tables(nz, 0)
    st = L.rf_density_noalloc(ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(d.data_ptr()), d.numel(), nz,
                              _lib.RF_F32, 1, 0.0, _lib.ptr(growth), _lib.ptr(rho), None)
    assert st == _lib.RF_ERR_ARG
    assert "RF_NOALLOC_MAX_NZ" in _lib.last_error()
    torch.cuda.synchronize()
    assert float(d.abs().sum()) == 0.0          # nothing ran
    # bad arguments: nelem not a multiple of nz, null arrays
    assert L.rf_density_noalloc(ctypes.c_void_p(d.data_ptr()), ctypes.c_void_p(d.data_ptr()), 7, 4, _lib.RF_F32,
                                0, 0.0, _lib.ptr(growth), _lib.ptr(rho), None) == _lib.RF_ERR_ARG
    assert L.rf_density_noalloc(None, ctypes.c_void_p(d.data_ptr()), 8, 4, _lib.RF_F32, 0, 0.0,
                                _lib.ptr(growth), _lib.ptr(rho), None) == _lib.RF_ERR_ARG
