"""ctypes binding to librf_b200.so (include/rf_b200.h).

The product path has no CPU implementation: if the shared library is missing, or no CUDA
device is visible when a context is created, the calls below raise.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RF_LIB: load another build of the same library (A/B timing of compile-time variants)
LIB_PATH = os.environ.get("RF_LIB") or os.path.join(_HERE, "librf_b200.so")

RF_F32, RF_F64 = 0, 1
RF_BUF_FIELD, RF_BUF_RECV, RF_BUF_PARTIALS = 0, 1, 2
RF_BUF_AUX = RF_BUF_RECV
RF_ERR_ARG, RF_ERR_CUDA, RF_ERR_STATE = -1, -2, -3

# every symbol include/rf_b200.h declares: (name, restype, argtypes)
_c = ctypes
_P = _c.c_void_p
SIGNATURES = {
    "rf_last_error": (_c.c_char_p, []),
    "rf_version": (_c.c_int, []),
    "rf_create": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.POINTER(_P)]),
    "rf_destroy": (_c.c_int, [_P]),
    "rf_set_stream": (_c.c_int, [_P, _P]),
    "rf_sync": (_c.c_int, [_P]),
    "rf_set_power": (_c.c_int, [_P, _c.c_int, _P, _P, _c.c_double]),
    "rf_set_noise": (_c.c_int, [_P, _P]),
    "rf_set_noise_rows": (_c.c_int, [_P, _P, _c.c_int, _c.c_int]),
    "rf_generate_xpass": (_c.c_int, [_P, _c.c_uint64, _c.c_int]),
    "rf_finish_yz": (_c.c_int, [_P, _P, _c.c_int]),
    "rf_generate_delta": (_c.c_int, [_P, _c.c_uint64, _c.c_int, _P, _c.c_int]),
    "rf_stats": (_c.c_int, [_P, _c.POINTER(_c.c_double), _c.POINTER(_c.c_double), _c.POINTER(_c.c_longlong)]),
    "rf_density": (_c.c_int, [_P, _c.c_int, _c.c_double, _P, _P]),
    "rf_copy_field": (_c.c_int, [_P, _P]),
    "rf_copy_rows": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _P, _c.c_int]),
    "rf_copy_wait": (_c.c_int, [_P]),
    "rf_finish_yz_streamed": (_c.c_int, [_P, _P, _c.c_int, _P, _c.c_int]),
    "rf_field_elems": (_c.c_longlong, [_P]),
    "rf_kspace_public": (_c.c_int, [_P, _c.c_uint64, _c.c_int, _P, _P]),
    "rf_plan_execute": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P, _P]),
    "rf_stage_fill_log10k": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _P]),
    "rf_stage_tabulate": (_c.c_int, [_c.c_longlong, _c.c_int, _c.c_int, _P, _P, _P]),
    "rf_stage_randomize": (_c.c_int, [_c.c_longlong, _c.c_int, _c.c_uint64, _P]),
    "rf_stage_symmetrize": (_c.c_int, [_c.c_int, _c.c_int, _c.c_int, _c.c_int, _c.c_int, _P]),
    "rf_stage_density": (_c.c_int, [_c.c_longlong, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _P, _P, _P]),
    "rf_density_noalloc": (_c.c_int, [_P, _P, _c.c_longlong, _c.c_int, _c.c_int, _c.c_int, _c.c_double, _P, _P, _P]),
    "rf_host_alloc": (_P, [_c.c_longlong]),
    "rf_host_free": (None, [_P]),
    "rf_set_timing": (_c.c_int, [_P, _c.c_int]),
    "rf_get_timing": (_c.c_int, [_P, _c.POINTER(_c.c_double), _c.POINTER(_c.c_double), _c.POINTER(_c.c_double)]),
    "rf_device_ptr": (_P, [_P, _c.c_int]),
    "rf_buffer_bytes": (_c.c_longlong, [_P, _c.c_int]),
    "rf_ipc_export": (_c.c_int, [_P, _c.c_int, _P]),
    "rf_ipc_open": (_c.c_int, [_P, _c.POINTER(_P)]),
    "rf_ipc_close": (_c.c_int, [_P]),
    "rf_set_peer_recv": (_c.c_int, [_P, _c.c_int, _P]),
    "rf_set_direct": (_c.c_int, [_P, _c.c_int]),
    "rf_lensing": (_c.c_int, [_P, _c.c_int, _P]),
    "rf_r2c_power": (_c.c_int, [_P, _c.c_int, _P, _P, _P]),
    "rf_kernel_names": (_c.c_int, [_P, _c.c_char_p, _c.c_int]),
    "rf_field_state": (_c.c_int, [_P]),
    "rf_chunk_config": (_c.c_int, [_P, _c.c_int]),
    "rf_comm_stream": (_c.c_void_p, [_P]),
    "rf_chunk_begin_step": (_c.c_int, [_P]),
    "rf_chunk_xpass": (_c.c_int, [_P, _c.c_uint64, _c.c_int, _c.c_int]),
    "rf_chunk_exchange": (_c.c_int, [_P, _c.c_int]),
    "rf_chunk_arrived": (_c.c_int, [_P, _c.c_int]),
    "rf_chunk_pass2": (_c.c_int, [_P, _c.c_int]),
    "rf_chunk_exchange_x": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "rf_chunk_pass2_x": (_c.c_int, [_P, _c.c_int, _c.c_int, _c.c_int, _c.c_int]),
    "rf_chunk_pass3_x": (_c.c_int, [_P, _c.c_int, _c.c_int, _P, _c.c_int, _c.c_int]),
    "rf_is_fast": (_c.c_int, [_P]),
    "rf_is_pipelined": (_c.c_int, [_P]),
    "rf_is_fused": (_c.c_int, [_P]),
    "rf_launch_count": (_c.c_longlong, [_P]),
}

_LIB = None


def lib():
    """Load librf_b200.so (once).  Raises if it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "librf_b200.so is missing (%s): build it with `python -m randomfield_b200.build`; "
                "randomfield_b200 has no CPU fallback." % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = handle
    return _LIB


def last_error() -> str:
    msg = lib().rf_last_error()
    return msg.decode() if msg else ""


def check(status: int, value_error_ok: bool = True):
    """Map a C status to the exception types the reference raises."""
    if status == 0:
        return
    msg = last_error()
    if status == RF_ERR_ARG and value_error_ok:
        raise ValueError(msg)
    raise RuntimeError(msg or "librf_b200 call failed with status %d" % status)


def dtype_code(dtype) -> int:
    dt = np.dtype(dtype)
    if dt in (np.dtype(np.float32), np.dtype(np.complex64)):
        return RF_F32
    if dt in (np.dtype(np.float64), np.dtype(np.complex128)):
        return RF_F64
    raise ValueError("Unsupported dtype for the B200 path: {0}.".format(dt))


def ptr(arr: np.ndarray):
    return arr.ctypes.data_as(ctypes.c_void_p)


class PinnedArray:
    """numpy array over page-locked host memory owned by the library."""

    def __init__(self, shape, dtype):
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self._p = lib().rf_host_alloc(self.nbytes)
        if not self._p:
            raise MemoryError(last_error())
        buf = (ctypes.c_char * self.nbytes).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=dtype).reshape(shape)

    def __del__(self):
        try:
            if getattr(self, "_p", None):
                lib().rf_host_free(self._p)
                self._p = None
        except Exception:
            pass


class Engine:
    """One rf_ctx: the device side of a Generator (one rank of a slab decomposition)."""

    def __init__(self, nx, ny, nz, dtype=np.float32, rank=0, world=1):
        self.nx, self.ny, self.nz = int(nx), int(ny), int(nz)
        self.real_dtype = np.dtype(np.float32 if dtype_code(dtype) == RF_F32 else np.float64)
        self.complex_dtype = np.dtype(np.complex64 if self.real_dtype == np.float32 else np.complex128)
        self.rank, self.world = int(rank), int(world)
        self._ctx = ctypes.c_void_p()
        check(lib().rf_create(self.nx, self.ny, self.nz, dtype_code(dtype), self.rank, self.world,
                              ctypes.byref(self._ctx)))
        self.nxl, self.nyl = self.nx // self.world, self.ny // self.world

    # -- lifetime ---------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx:
            lib().rf_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration ------------------------------------------------------------------
    def set_stream(self, cuda_stream_handle):
        check(lib().rf_set_stream(self._ctx, ctypes.c_void_p(cuda_stream_handle or 0)))

    def set_power(self, k, sigma, spacing):
        k = np.ascontiguousarray(k, dtype=np.float64)
        sigma = np.ascontiguousarray(sigma, dtype=np.float64)
        check(lib().rf_set_power(self._ctx, len(k), ptr(k), ptr(sigma), float(spacing)))

    def set_noise(self, noise):
        if noise is None:
            check(lib().rf_set_noise(self._ctx, None))
            return
        noise = np.ascontiguousarray(noise, dtype=np.float64).reshape(-1)
        expect = 2 * self.nx * self.ny * (self.nz // 2 + 1)
        if noise.size != expect:
            raise ValueError("noise has {0} values, expected {1}.".format(noise.size, expect))
        check(lib().rf_set_noise(self._ctx, ptr(noise)))

    def set_noise_rows(self, rows, x0):
        """Upload x rows [x0, x0 + len(rows)) of the white-noise array (rows: float64, shape
        (n, ny, nz//2+1, 2) or flat)."""
        rows = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1)
        per_row = 2 * self.ny * (self.nz // 2 + 1)
        if rows.size == 0 or rows.size % per_row:
            raise ValueError("noise rows have {0} values, expected a multiple of {1}.".format(rows.size, per_row))
        check(lib().rf_set_noise_rows(self._ctx, ptr(rows), int(x0), int(x0) + rows.size // per_row))

    # -- hot path -------------------------------------------------------------------------
    def generate_xpass(self, seed, potential=False):
        check(lib().rf_generate_xpass(self._ctx, int(seed) & 0xFFFFFFFFFFFFFFFF, int(bool(potential))))

    def finish_yz(self, zscale=None, want_stats=True):
        zs = None if zscale is None else np.ascontiguousarray(zscale, dtype=np.float64)
        check(lib().rf_finish_yz(self._ctx, None if zs is None else ptr(zs), int(bool(want_stats))))

    def generate_delta(self, seed, potential=False, zscale=None, want_stats=True):
        zs = None if zscale is None else np.ascontiguousarray(zscale, dtype=np.float64)
        check(lib().rf_generate_delta(self._ctx, int(seed) & 0xFFFFFFFFFFFFFFFF, int(bool(potential)),
                                      None if zs is None else ptr(zs), int(bool(want_stats))))

    def finish_yz_streamed(self, out, zscale=None, want_stats=True, nranges=8):
        """rf_finish_yz with the result streamed into the host array `out` (page-locked for full
        overlap) range by range; call copy_wait() before reading it."""
        if out.dtype != self.real_dtype or out.shape != self.field_shape or not out.flags.c_contiguous:
            raise ValueError("bad output array for finish_yz_streamed")
        zs = None if zscale is None else np.ascontiguousarray(zscale, dtype=np.float64)
        check(lib().rf_finish_yz_streamed(self._ctx, None if zs is None else ptr(zs), int(bool(want_stats)),
                                          ptr(out), int(nranges)))

    def copy_rows(self, out_rows, x0, which=RF_BUF_FIELD, wait=True):
        """Copy local x rows [x0, x0 + len(out_rows)) of the real field (or of RF_BUF_AUX, the
        lensing potential) into `out_rows` on the copy stream."""
        if out_rows.dtype != self.real_dtype or out_rows.shape[1:] != (self.ny, self.nz) or \
                not out_rows.flags.c_contiguous:
            raise ValueError("bad output array for copy_rows")
        check(lib().rf_copy_rows(self._ctx, int(which), int(x0), int(x0) + out_rows.shape[0], ptr(out_rows),
                                 0 if wait else 1), value_error_ok=False)
        return out_rows

    def copy_wait(self):
        check(lib().rf_copy_wait(self._ctx))

    def iter_rows(self, rows=None, which=RF_BUF_FIELD, stage_bytes=1 << 29):
        """Stream this rank's real field (or RF_BUF_AUX) to the host in chunks of x rows through two
        page-locked staging buffers: yields (x0, array) with array a (n, ny, nz) view that stays valid
        until the next iteration; the copy of chunk k+1 is under way while the caller consumes chunk
        k.  For slabs too large to page-lock whole (34 GB per rank at 4096^3 on 8 GPUs)."""
        row_bytes = self.ny * self.nz * self.real_dtype.itemsize
        if rows is None:
            rows = max(1, min(self.nxl, stage_bytes // row_bytes))
        if getattr(self, "_stage_rows", 0) != rows:
            self._stage = [PinnedArray((rows, self.ny, self.nz), self.real_dtype) for _ in range(2)]
            self._stage_rows = rows
        starts = list(range(0, self.nxl, rows))
        if not starts:
            return
        n0 = min(rows, self.nxl - starts[0])
        self.copy_rows(self._stage[0].array[:n0], starts[0], which=which, wait=False)
        for k, x0 in enumerate(starts):
            self.copy_wait()
            if k + 1 < len(starts):
                n1 = min(rows, self.nxl - starts[k + 1])
                self.copy_rows(self._stage[(k + 1) & 1].array[:n1], starts[k + 1], which=which, wait=False)
            yield x0, self._stage[k & 1].array[:min(rows, self.nxl - x0)]

    def r2c_power(self, k_edges):
        """Shell-averaged P(k) of the real-space field (rf_r2c_power): returns (Pk, n_modes)."""
        edges = np.ascontiguousarray(k_edges, dtype=np.float64)
        nb = len(edges) - 1
        pk = np.zeros(nb, dtype=np.float64)
        cnt = np.zeros(nb, dtype=np.int64)
        check(lib().rf_r2c_power(self._ctx, nb, ptr(edges), ptr(pk), ptr(cnt)))
        return pk, cnt

    @property
    def kernel_names(self):
        buf = ctypes.create_string_buffer(256)
        check(lib().rf_kernel_names(self._ctx, buf, 256))
        return tuple(buf.value.decode().split(";"))

    @property
    def field_is_real(self):
        return int(lib().rf_field_state(self._ctx)) == 2

    def stats(self):
        s1, s2, n = ctypes.c_double(), ctypes.c_double(), ctypes.c_longlong()
        check(lib().rf_stats(self._ctx, ctypes.byref(s1), ctypes.byref(s2), ctypes.byref(n)),
              value_error_ok=False)
        return s1.value, s2.value, n.value

    def density(self, lognormal, sigma, growth, rho):
        g = np.ascontiguousarray(np.broadcast_to(np.asarray(growth, dtype=np.float64).reshape(-1), (self.nz,)))
        r = np.ascontiguousarray(np.broadcast_to(np.asarray(rho, dtype=np.float64).reshape(-1), (self.nz,)))
        check(lib().rf_density(self._ctx, int(bool(lognormal)), float(sigma), ptr(g), ptr(r)))

    def set_timing(self, enabled=True):
        check(lib().rf_set_timing(self._ctx, int(bool(enabled))))

    def get_timing(self):
        """(pass1, pass2, pass3) milliseconds of the last step (synchronises)."""
        t = [ctypes.c_double() for _ in range(3)]
        check(lib().rf_get_timing(self._ctx, *[ctypes.byref(x) for x in t]), value_error_ok=False)
        return tuple(x.value for x in t)

    def sync(self):
        check(lib().rf_sync(self._ctx))

    # -- data movement ------------------------------------------------------------------------
    @property
    def field_shape(self):
        return (self.nxl, self.ny, self.nz)

    def copy_field(self, out=None):
        if out is None:
            out = np.empty(self.field_shape, dtype=self.real_dtype)
        if out.dtype != self.real_dtype or out.shape != self.field_shape or not out.flags.c_contiguous:
            raise ValueError("bad output array for copy_field")
        check(lib().rf_copy_field(self._ctx, ptr(out)), value_error_ok=False)
        return out

    def kspace_public(self, seed, potential=False, with_bins=False):
        shape = (self.nx, self.nyl, self.nz // 2 + 1)
        out = np.empty(shape, dtype=self.complex_dtype)
        bins = np.empty(shape, dtype=np.int32) if with_bins else None
        check(lib().rf_kspace_public(self._ctx, int(seed) & 0xFFFFFFFFFFFFFFFF, int(bool(potential)),
                                     ptr(out), None if bins is None else ptr(bins)))
        return (out, bins) if with_bins else out

    def device_ptr(self, which=RF_BUF_FIELD):
        return lib().rf_device_ptr(self._ctx, which)

    def buffer_bytes(self, which=RF_BUF_FIELD):
        return lib().rf_buffer_bytes(self._ctx, which)

    # -- direct exchange over peer memory ---------------------------------------------------
    def ipc_export(self, which=RF_BUF_RECV):
        buf = ctypes.create_string_buffer(64)
        check(lib().rf_ipc_export(self._ctx, which, ctypes.cast(buf, ctypes.c_void_p)), value_error_ok=False)
        return bytes(buf.raw)

    @staticmethod
    def ipc_open(handle):
        buf = ctypes.create_string_buffer(bytes(handle), 64)
        out = ctypes.c_void_p()
        check(lib().rf_ipc_open(ctypes.cast(buf, ctypes.c_void_p), ctypes.byref(out)), value_error_ok=False)
        return out.value

    def set_peer_recv(self, rank, ptr):
        check(lib().rf_set_peer_recv(self._ctx, int(rank), ctypes.c_void_p(ptr)))

    def set_direct(self, enabled=True):
        check(lib().rf_set_direct(self._ctx, int(bool(enabled))))

    def lensing(self, i_min, tables5):
        """In-place lensing-potential scan of the real field; tables5: (5, nz) float64."""
        t = np.ascontiguousarray(tables5, dtype=np.float64)
        if t.shape != (5, self.nz):
            raise ValueError("tables5 must have shape (5, nz)")
        check(lib().rf_lensing(self._ctx, int(i_min), ptr(t)))

    # ---- chunked exchange (include/rf_b200.h: rf_chunk_*) ----
    def chunk_config(self, nchunks):
        check(lib().rf_chunk_config(self._ctx, int(nchunks)))

    @property
    def comm_stream(self):
        return int(lib().rf_comm_stream(self._ctx) or 0)

    def chunk_begin_step(self):
        check(lib().rf_chunk_begin_step(self._ctx))

    def chunk_xpass(self, seed, potential, chunk):
        check(lib().rf_chunk_xpass(self._ctx, int(seed) & 0xFFFFFFFFFFFFFFFF, int(bool(potential)), int(chunk)))

    def chunk_exchange(self, chunk):
        check(lib().rf_chunk_exchange(self._ctx, int(chunk)))

    def chunk_arrived(self, chunk):
        check(lib().rf_chunk_arrived(self._ctx, int(chunk)))

    def chunk_pass2(self, chunk):
        check(lib().rf_chunk_pass2(self._ctx, int(chunk)))

    # split schedule: the exchange also cut along the destination's local x axis
    def chunk_exchange_x(self, chunk, x0, x1, wait_pass1):
        check(lib().rf_chunk_exchange_x(self._ctx, int(chunk), int(x0), int(x1), int(bool(wait_pass1))))

    def chunk_pass2_x(self, chunk, x0, x1, wait_index=-1):
        check(lib().rf_chunk_pass2_x(self._ctx, int(chunk), int(x0), int(x1), int(wait_index)))

    def chunk_pass3_x(self, x0, x1, zscale=None, want_stats=True, last=False):
        zs = None if zscale is None else np.ascontiguousarray(zscale, dtype=np.float64)
        check(lib().rf_chunk_pass3_x(self._ctx, int(x0), int(x1), None if zs is None else ptr(zs),
                                     int(bool(want_stats)), int(bool(last))))

    @property
    def is_fast(self):
        return bool(lib().rf_is_fast(self._ctx))

    @property
    def is_pipelined(self):
        return bool(lib().rf_is_pipelined(self._ctx))

    @property
    def fused_yz(self):
        """0: separate pass-2 / pass-3 kernels, 2: k23q_fused_yz."""
        return int(lib().rf_is_fused(self._ctx))

    @property
    def launch_count(self):
        return int(lib().rf_launch_count(self._ctx))

    def torch_view(self, which=RF_BUF_FIELD, nbytes=None):
        """Zero-copy uint8 torch tensor over a library-owned device buffer."""
        import torch

        class _Holder:
            pass

        h = _Holder()
        n = int(nbytes if nbytes is not None else self.buffer_bytes(which))
        h.__cuda_array_interface__ = {
            "shape": (n,), "typestr": "|u1", "data": (int(self.device_ptr(which)), False), "version": 2,
        }
        return torch.as_tensor(h, device="cuda")

    def field_tensor(self):
        """The real field [nx/world, ny, nz] as a torch tensor sharing device memory."""
        import torch
        if not self.field_is_real:
            raise RuntimeError("the device buffer does not hold a real-space field")
        n = self.nxl * self.ny * self.nz
        t = self.torch_view(RF_BUF_FIELD, n * self.real_dtype.itemsize)
        tdt = torch.float32 if self.real_dtype == np.float32 else torch.float64
        return t.view(tdt).view(self.nxl, self.ny, self.nz)
