"""GPU-backed mirror of randomfield.generate.Generator (reference: randomfield/generate.py).

Constructor and method signatures, attributes, return shapes/dtypes and exceptions follow the
reference; the seed -> delta(x) pipeline, the density conversion and the Newtonian potential
run as CUDA kernels in librf_b200.so.  Differences a user can observe:

* deviates come from Philox4x32-10 instead of MT19937 (same statistics, same
  reproducibility-by-seed), unless white noise is injected with :meth:`set_white_noise`;
* ``save_potential`` does not allocate a second k-space buffer: the potential source field is
  regenerated from (seed, smoothing) when :meth:`calculate_newtonian_potential` runs;
  ``generator.potential`` materialises the reference's array on first access;
* returned arrays are views of one page-locked host buffer that later calls overwrite (the
  reference returns views of its single work buffer with the same caveat); the device copy
  is available zero-copy through :meth:`device_field`.  The lensing potential is a separate
  array, as in the reference (generate.py:398-399);
* extensions: ``dtype=`` (float64 arithmetic), :meth:`set_white_noise`, :meth:`measure_power`
  (the forward r2c transform the reference plans but never runs + shell-binned P(k)).
"""
from __future__ import print_function, division

import os

import numpy as np
import scipy.interpolate

from . import _lib
from . import cosmotools
from . import powertools
from . import transform

_H0_PER_S = 100.0 * 1.0e3 / 3.0856775814913673e22   # 100 km/s/Mpc in 1/s
_C_KM_S = 299792.458


class _PlanView(object):
    """What the reference exposes as ``generator.plan_c2r`` / ``plan_r2c``: shapes, byte
    counts and lazily materialised host arrays (transform.py:190-244)."""

    def __init__(self, generator, inverse, nbytes):
        self._g = generator
        self.shape = (generator._engine.nx, generator._engine.ny, generator._engine.nz)
        self.inverse, self.packed, self.overwrite, self.use_pyfftw = inverse, True, True, False
        self.nbytes_allocated = nbytes

    @property
    def data_out(self):
        if self.inverse:
            return self._g._host_field()
        return self._g._host_kspace()

    @property
    def data_in(self):
        if self.inverse:
            return self._g._host_kspace()
        return self._g._host_field()


def cot_k(DC, DA, Ok0):
    """cotK(D) = cosK(D) / D_A(D), 1 at the observer (generate.py:380-394)."""
    H0, clight_km_s = 100., 299792.458
    K = -Ok0 * (H0 / clight_km_s) ** 2
    if K < 0:
        cosK = np.cosh(np.sqrt(-K) * DC)
    elif K > 0:
        cosK = np.cos(np.sqrt(K) * DC)
    else:
        cosK = np.ones_like(DA, dtype=float)
    cotK = np.ones_like(cosK)
    cotK[1:] = cosK[1:] / DA[1:]
    return cotK


def lensing_tables(DC, cotK):
    """Coefficient tables of rf_lensing (include/rf_b200.h), shape (5, nz) float64: cotK; the
    weights (a, b, d) scipy.integrate.simps gives the three samples (j-2, j-1, j) of a Simpson
    panel on the abscissae DC -- hsum/6 * (2 - h1/h0, hsum^2 / (h0 h1), 2 - h0/h1) with
    h0 = DC[j-1] - DC[j-2], h1 = DC[j] - DC[j-1]; and the trapezoid half-width (DC[j] - DC[j-1]) / 2."""
    x = np.asarray(DC, dtype=np.float64)
    nz = x.size
    t = np.zeros((5, nz))
    t[0] = cotK
    h = np.diff(x)
    if nz >= 3:
        h0, h1 = h[:-1], h[1:]
        hsum = h0 + h1
        t[1, 2:] = hsum / 6.0 * (2 - h1 / h0)
        t[2, 2:] = hsum / 6.0 * hsum * hsum / (h0 * h1)
        t[3, 2:] = hsum / 6.0 * (2 - h0 / h1)
    t[4, 1:] = 0.5 * h
    return t


def lensing_scan_host(dphi, tables, i_min):
    """numpy restatement of the device algorithm (rf_lens.cuh): parity-split running sums of the
    Simpson panels instead of one integral per slice.  Host logic for the CPU tests."""
    cotk, pa, pb, pd, pt = tables
    f = np.asarray(dphi, dtype=np.float64)
    g = f * cotk
    nz = f.shape[-1]
    psi = np.zeros_like(f)
    aE = [np.zeros(f.shape[:-1]), np.zeros(f.shape[:-1])]
    aO = [np.zeros(f.shape[:-1]), np.zeros(f.shape[:-1])]
    T0 = [0.0, 0.0]
    if i_min + 1 < nz:
        T0 = [pt[i_min + 1] * (q[..., i_min] + q[..., i_min + 1]) for q in (f, g)]
    for j in range(i_min, nz):
        odd = (j - i_min) & 1
        if j >= i_min + 2:
            for n, q in enumerate((f, g)):
                pan = pa[j] * q[..., j - 2] + pb[j] * q[..., j - 1] + pd[j] * q[..., j]
                (aO if odd else aE)[n] = (aO if odd else aE)[n] + pan
        if not odd:
            I = [aE[0], aE[1]]
        else:
            I = [0.5 * (aE[n] + pt[j] * (q[..., j - 1] + q[..., j])) + 0.5 * (T0[n] + aO[n])
                 for n, q in enumerate((f, g))]
        psi[..., j] = -2.0 * (I[1] - cotk[j] * I[0])
    return psi


class Generator(object):
    """Manage random field generation for a specified geometry (generate.py:21-142).

    Parameters are those of the reference: ``nx, ny, nz`` grid sizes (all even), uniform
    ``grid_spacing_Mpc_h``, ``num_plot_sections`` (must divide nz), optional ``cosmology``
    (anything with the astropy FLRW attributes used here), optional tabulated ``power`` and
    ``verbose``.  ``dtype`` (extension, default float32 like the reference's complex64 plan)
    selects float64 arithmetic.

    ``group`` (extension) slab-decomposes the grid over the ranks of a ``torch.distributed``
    process group (``True``: the default group; one process per GPU, launched with torchrun):
    every rank constructs the Generator with the same arguments, the methods are collective, and
    each returns this rank's x-slab ``[rank * nx/G, (rank + 1) * nx/G)`` of the reference's array
    -- concatenating the ranks' results along axis 0 gives the single-GPU array bit for bit.
    ``delta_field_rms`` is the rms of the whole field.  :meth:`gather` collects a result on rank 0.
    """

    def __init__(self, nx, ny, nz, grid_spacing_Mpc_h, num_plot_sections=4,
                 cosmology=None, power=None, verbose=False, dtype=np.float32, group=None):
        if nx % 2 or ny % 2 or nz % 2:
            raise ValueError('All shape dimensions must be even.')
        self._slab = None
        self._group = None
        if group is not None and group is not False:
            import torch.distributed as dist
            if not dist.is_initialized():
                raise RuntimeError('group= needs an initialised torch.distributed process group.')
            self._group = None if group is True else group
            if dist.get_world_size(self._group) > 1:
                from .slab import SlabEngine
                if power is None and cosmology is None:
                    power = powertools.load_default_power()
                if power is None:
                    raise RuntimeError('power.calculate_power requires classy.')
                self._slab = SlabEngine(nx, ny, nz, grid_spacing_Mpc_h, powertools.validate_power(power),
                                        dtype=dtype, group=self._group)
                self._engine = self._slab.engine
        if self._slab is None:
            self._engine = _lib.Engine(nx, ny, nz, dtype=dtype)
        self._real_dtype = self._engine.real_dtype
        self._shape = (nx, ny, nz)
        self._local_shape = self._engine.field_shape
        itemsize = self._engine.complex_dtype.itemsize
        self.plan_c2r = _PlanView(self, True, nx * ny * (nz // 2 + 1) * itemsize)
        self.plan_r2c = _PlanView(self, False, 0)
        self.grid_spacing_Mpc_h = grid_spacing_Mpc_h
        k0 = (2 * np.pi) / grid_spacing_Mpc_h
        self.k_min, self.k_max = k0 / max(nx, ny, nz), k0 * np.sqrt(3) / 2
        self._potential_args = None
        self._potential_cache = None
        self._noise = None
        self._host = None
        self._host_psi = None
        self._field_kind = None
        self._last = None
        self._table_key = None

        if nz % num_plot_sections != 0:
            raise ValueError('Z-axis does not evenly divided into {0} plot sections.'
                             .format(num_plot_sections))
        self.num_plot_sections = num_plot_sections

        if cosmology is None:
            self.cosmology = cosmotools.create_cosmology()
            if power is None:
                power = powertools.load_default_power()
        else:
            self.cosmology = cosmology
            if power is None:
                power = cosmotools.calculate_power(self.cosmology, k_min=self.k_min, k_max=self.k_max,
                                                   scaled_by_h=True)
        self.power = powertools.validate_power(power)

        probe = np.empty((1, 1, nz), dtype=self._real_dtype)
        self.redshifts = cosmotools.get_redshifts(
            self.cosmology, probe, spacing=self.grid_spacing_Mpc_h, scaled_by_h=True,
            z_axis=2).reshape(nz)
        self.redshift_to_index = scipy.interpolate.interp1d(
            self.redshifts, np.arange(nz), kind='linear', bounds_error=False)
        self.DC = np.arange(nz) * self.grid_spacing_Mpc_h
        self.DA = cosmotools._value(
            self.cosmology.comoving_transverse_distance(self.redshifts),
            cosmotools._u.Mpc if cosmotools._u is not None else None) * self.cosmology.h
        self.angular_spacing = self.DA * (np.pi / 180.) / self.grid_spacing_Mpc_h
        self.z_max = self.redshifts.flat[-1]
        self.x_fov = nx / self.angular_spacing.flat[-1]
        self.y_fov = ny / self.angular_spacing.flat[-1]
        self.growth_function = cosmotools.get_growth_function(self.cosmology, self.redshifts)
        self.mean_matter_density = cosmotools.get_mean_matter_densities(
            self.cosmology, self.redshifts)

        self.verbose = verbose
        if self.verbose:
            Mb = (self.plan_c2r.nbytes_allocated + self.plan_r2c.nbytes_allocated) / 2.0 ** 20
            print('Allocated {0:.1f} Mb for {1} x {2} x {3} grid.'.format(Mb, nx, ny, nz))
            print('{0} Mpc/h spacing covered by k = {1:.5f} - {2:.5f} h/Mpc.'
                  .format(self.grid_spacing_Mpc_h, self.k_min, self.k_max))
            print('Grid has z < {0:.3f} with {1:.4f} deg x {2:.4f} deg'
                  .format(self.z_max, self.x_fov, self.y_fov),
                  ' = {0:.4f} deg**2 field of view.'.format(self.x_fov * self.y_fov))

    # ------------------------------------------------------------------ host/device views
    def _host_buffer(self):
        if self._host is None:
            self._host = _lib.PinnedArray(self._local_shape, self._real_dtype)
        return self._host.array

    def _host_field(self):
        """Host copy of the device real field in the generator's page-locked buffer."""
        out = self._host_buffer()
        self._engine.copy_field(out)
        return out

    def _host_kspace(self):
        if self._last is None:
            raise RuntimeError('No k-space field has been generated yet.')
        seed, smoothing = self._last
        self._load_table(smoothing)
        return self._engine.kspace_public(seed, potential=False)

    def _run(self, seed, potential, zscale, want_stats, out):
        """seed -> real-space field (this rank's x-slab) streamed into the host array `out`."""
        if self._slab is not None:
            self._slab.generate(seed, potential=potential, zscale=zscale, want_stats=want_stats)
            self._engine.copy_rows(out, 0, wait=False)
        else:
            # pass 1, then passes 2 + 3 range by range with each range's device -> host copy under
            # way while the next range is transformed (rf_finish_yz_streamed)
            self._engine.generate_xpass(seed, potential=potential)
            self._engine.finish_yz_streamed(out, zscale=zscale, want_stats=want_stats)

    @property
    def x_slab(self):
        """[lo, hi) of axis 0 held by this rank (the whole axis on one GPU)."""
        nxl = self._local_shape[0]
        return self._engine.rank * nxl, (self._engine.rank + 1) * nxl

    def gather(self, local, dst=0):
        """Collect the ranks' x-slabs of a result on rank `dst`: returns the (nx, ny, nz) array
        there and None elsewhere (the array itself on one GPU)."""
        if self._slab is None:
            return local
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(self._group), dist.get_world_size(self._group)
        mine = torch.from_numpy(np.ascontiguousarray(local)).cuda()
        parts = [torch.empty_like(mine) for _ in range(world)] if rank == dst else None
        dist.gather(mine, parts, dst=dist.get_global_rank(self._group, dst) if self._group is not None else dst,
                    group=self._group)
        if rank != dst:
            return None
        return np.concatenate([t.cpu().numpy() for t in parts], axis=0)

    def iter_field_rows(self, rows=None):
        """The current device field (this rank's x-slab) streamed to the host in chunks of x rows
        through two page-locked staging buffers: yields (global x0, (n, ny, nz) array); each array
        is valid until the next iteration.  For fields too large to hold on the host at once."""
        lo = self.x_slab[0]
        for x0, rows_ in self._engine.iter_rows(rows):
            yield lo + x0, rows_

    def device_field(self):
        """The current real field as a torch CUDA tensor sharing the library's memory."""
        return self._engine.field_tensor()

    def set_white_noise(self, noise):
        """Inject white noise: ``2*nx*ny*(nz//2+1)`` float64 standard-normal deviates in the
        order ``RandomState.normal(size)`` fills the reference's buffer (random.py:17-28).
        ``noise`` is the whole array, or an iterable of consecutive slabs of x rows (each
        ``(n, ny, nz//2+1, 2)`` or flat) that is uploaded slab by slab -- 16 bytes per mode never
        have to exist on the host at once (random.py:25-27).  ``None`` returns to the Philox
        generator."""
        if noise is None or isinstance(noise, np.ndarray):
            self._engine.set_noise(noise)
        else:
            x0, per_row = 0, 2 * self._shape[1] * (self._shape[2] // 2 + 1)
            for slab in noise:
                slab = np.ascontiguousarray(slab, dtype=np.float64).reshape(-1)
                self._engine.set_noise_rows(slab, x0)
                x0 += slab.size // per_row
            if x0 != self._shape[0]:
                raise ValueError('noise slabs cover {0} of {1} x rows.'.format(x0, self._shape[0]))
        self._noise = noise is not None

    @property
    def potential(self):
        """delta(k)/k**2 in the public (nx, ny, nz//2+1) layout (generate.py:200-215), or
        None when the last field was generated with ``save_potential=False``."""
        if self._potential_args is None:
            return None
        if self._potential_cache is None:
            # generated into scratch memory on the device: no side effect on the field
            self._load_table(self._potential_args[1])
            self._potential_cache = self._engine.kspace_public(self._potential_args[0], potential=True)
        return self._potential_cache

    def _load_table(self, smoothing):
        # the device table is rebuilt only when the smoothing (or the power array) changes
        key = (float(smoothing), id(self.power))
        if key == self._table_key:
            return
        self.smoothed_power = powertools.filter_power(self.power, smoothing)
        k, sigma = powertools.sigma_table(self.smoothed_power, self._shape, self.grid_spacing_Mpc_h)
        self._engine.set_power(k, sigma, self.grid_spacing_Mpc_h)
        self._table_key = key

    def measure_power(self, k_edges=None, nbins=24):
        """Shell-averaged power spectrum of the current real-space field, measured on the device
        with the forward r2c transform (the reference's ``plan_r2c``, generate.py:79-80): returns
        ``(k_edges, Pk, n_modes)`` with Pk in (Mpc/h)^3.  Extension of the reference API."""
        if self._field_kind is None:
            raise RuntimeError('No field in the buffer.')
        if k_edges is None:
            k_edges = np.linspace(self.k_min, np.pi / self.grid_spacing_Mpc_h, nbins + 1)
        k_edges = np.asarray(k_edges, dtype=np.float64)
        pk, cnt = self._engine.r2c_power(k_edges)
        return k_edges, pk, cnt

    # ------------------------------------------------------------------ the hot path
    def generate_delta_field(self, smoothing_length_Mpc_h=0., seed=None,
                             save_potential=True, show_plot=False, save_plot_name=None):
        """Generate a delta-field realization (generate.py:144-230).

        Returns a float32 (or float64) array of shape (nx, ny, nz): a view of this
        generator's host buffer that later calls overwrite."""
        self._check_plot(show_plot, save_plot_name)
        if seed is None:
            seed = int.from_bytes(os.urandom(8), 'little')
            if self._slab is not None:      # every rank must draw the same field
                import torch
                import torch.distributed as dist
                t = torch.tensor([seed >> 1], dtype=torch.int64, device='cuda')
                dist.broadcast(t, src=dist.get_global_rank(self._group, 0) if self._group is not None else 0,
                               group=self._group)
                seed = int(t.item())
        self._load_table(smoothing_length_Mpc_h)
        delta = self._host_buffer()
        self._run(seed, False, None, True, delta)
        if self._slab is not None:
            _, std = self._slab.global_stats()     # one 3-double all-reduce
            self.delta_field_rms = self._real_dtype.type(std)
        else:
            s1, s2, n = self._engine.stats()
            mean = s1 / n
            self.delta_field_rms = self._real_dtype.type(np.sqrt(max(s2 / n - mean * mean, 0.)))
        self._last = (seed, smoothing_length_Mpc_h)
        self._potential_args = (seed, smoothing_length_Mpc_h) if save_potential else None
        self._potential_cache = None
        self._field_kind = 'delta'
        if self.verbose:
            print('Delta field has standard deviation {0:.3f}.'.format(self.delta_field_rms))
        self._engine.copy_wait()
        if show_plot or save_plot_name is not None:
            self.plot_slice(show_plot=show_plot, save_plot_name=save_plot_name,
                            clip_symmetric=True, label='Matter inhomogeneity ' +
                            '$\\delta(r) = \\rho(r)/\\overline{\\rho} - 1$')
        return delta

    def convert_delta_to_density(self, apply_lognormal_transform=True,
                                 show_plot=False, save_plot_name=None):
        """Convert the delta field into a light-cone density field in g/cm**3
        (generate.py:232-280), in place on the device."""
        self._check_plot(show_plot, save_plot_name)
        if self._field_kind is None:
            raise RuntimeError('No delta field in the buffer.')
        self._engine.density(apply_lognormal_transform, float(self.delta_field_rms),
                             self.growth_function, self.mean_matter_density)
        self._field_kind = 'density'
        rho = self._host_field()
        if show_plot or save_plot_name is not None:
            self.plot_slice(show_plot=show_plot, save_plot_name=save_plot_name,
                            label='Matter density $\\rho(r)$ (g/cm$^3$)', clip_percent=2.)
        return rho

    def calculate_newtonian_potential(self, light_cone=True,
                                      show_plot=False, save_plot_name=None):
        """Newtonian potential Phi(r) in s**-2 (generate.py:282-350): inverse transform of
        -(3/2) H0^2 Om0 delta(k)/k^2, times G(z)/(1+z) on the light cone."""
        self._check_plot(show_plot, save_plot_name)
        if self._potential_args is None:
            raise RuntimeError('No saved potential field.')
        seed, smoothing = self._potential_args
        scale = -1.5 * _H0_PER_S ** 2 * self.cosmology.Om0
        zscale = np.full(self._shape[2], scale, dtype=np.float64)
        if light_cone:
            zscale = zscale * self.growth_function / (1 + self.redshifts)
        self._load_table(smoothing)
        field = self._host_buffer()
        self._run(seed, True, zscale, False, field)
        self._field_kind = 'potential'
        self._engine.copy_wait()
        if show_plot or save_plot_name is not None:
            self.plot_slice(clip_symmetric=True, show_plot=show_plot, save_plot_name=save_plot_name,
                            label='Newtonian potential $\\Phi(r)$ [s$^{-2}$]')
        return field

    def calculate_lensing_potential(self, i_min=None, show_plot=False, save_plot_name=None):
        """Lensing potential psi(r) (generate.py:353-418): for every slice the Simpson integral
        over comoving distance of -2 (cotK(D) - cotK(D_src)) dPhi, truncated below DC[i_min]
        (default nz // 32).  Integrates the Newtonian potential held by the device field (call
        calculate_newtonian_potential first) into a separate array, as the reference does
        (generate.py:398-399): the potential and the host array returned for it stay intact, and
        the call can be repeated with another i_min."""
        self._check_plot(show_plot, save_plot_name)
        nDC = self.DC.size
        if i_min is None:
            i_min = nDC // 32
        if i_min < 0 or i_min >= nDC:
            raise ValueError(
                'Invalid i_min {}. Expected 0 - {}.'.format(i_min, nDC - 1))
        if self._field_kind != 'potential':
            raise RuntimeError('No Newtonian potential in the buffer: call calculate_newtonian_potential first.')
        self._engine.lensing(i_min, lensing_tables(self.DC, cot_k(self.DC, self.DA, self.cosmology.Ok0)))
        if self._host_psi is None:
            self._host_psi = _lib.PinnedArray(self._local_shape, self._real_dtype)
        field = self._engine.copy_rows(self._host_psi.array, 0, which=_lib.RF_BUF_AUX)
        if show_plot or save_plot_name is not None:
            self.plot_slice(field=field, clip_percent=15.0, clip_symmetric=True, show_plot=show_plot,
                            save_plot_name=save_plot_name,
                            label='Lensing potential $\\psi(r)$ [(Mpc/h) s$^{-2}$]')
        return field

    @staticmethod
    def _check_plot(show_plot, save_plot_name):
        """Plotting (generate.py:420-516, matplotlib on a host copy) is outside the B200 path: a
        request for it fails BEFORE any field is generated, not after."""
        if show_plot or save_plot_name is not None:
            raise NotImplementedError('plot_slice is outside the B200 field-synthesis path: plot the returned '
                                      'array with the reference\'s plot_slice or matplotlib directly.')

    def plot_slice(self, *args, **kwargs):
        """generate.py:420-516: matplotlib slice plots of a host array -- not part of this path."""
        self._check_plot(True, None)
