"""Cosmology tables for the epilogue kernels (reference: randomfield/cosmotools.py).

The reference builds its length-nz tables (redshift, growth, mean density) from
astropy.cosmology.  astropy is an optional dependency here: when it is importable the same
calls are used; otherwise :class:`LiteFlatLambdaCDM` restates the flat LambdaCDM background
(matter + photons + neutrinos + Lambda) that those calls need.  The tables are O(nz) host
work and only *inputs* to the CUDA path.  ``apply_lognormal_transform`` (cosmotools.py:206-221)
is provided for host arrays; the Generator runs the fused device version (rf_density).
"""
from __future__ import print_function, division

import numpy as np
import scipy.interpolate
import scipy.integrate

try:  # pragma: no cover - astropy is absent on the build and GPU boxes
    import astropy.cosmology as _astro
    import astropy.units as _u
except Exception:  # ImportError or a broken install
    _astro = None
    _u = None

_C_KM_S = 299792.458
_MPC_CM = 3.0856775814913673e24
_G_CGS = 6.67384e-8                 # astropy 1.x CODATA-2010 value used for Planck13
_SIGMA_SB = 5.670373e-5             # erg cm^-2 s^-1 K^-4
_C_CGS = 2.99792458e10
_KB_EV_K = 8.6173324e-5


class LiteFlatLambdaCDM(object):
    """Flat LambdaCDM background with the subset of the astropy FLRW API that
    randomfield uses: h, H0, Om0, Ok0, Ode0, Om(z), critical_density0 (g/cm^3),
    comoving_distance(z) and comoving_transverse_distance(z) (Mpc)."""

    def __init__(self, H0=67.77, Om0=0.30712, Tcmb0=2.7255, Neff=3.046, m_nu=(0., 0., 0.06),
                 Ob0=0.048252, name=None):
        self.name = name
        self.H0 = float(H0)
        self.h = self.H0 / 100.
        self.Om0 = float(Om0)
        self.Ob0 = Ob0
        self.Tcmb0 = float(Tcmb0)
        self.Neff = float(Neff)
        self.m_nu = tuple(float(m) for m in np.atleast_1d(m_nu))
        self.Ok0 = 0.
        H0_s = self.H0 * 1.0e5 / _MPC_CM
        self.critical_density0 = 3. * H0_s ** 2 / (8. * np.pi * _G_CGS)
        rho_gamma = 4. * _SIGMA_SB * self.Tcmb0 ** 4 / _C_CGS ** 3
        self.Ogamma0 = rho_gamma / self.critical_density0
        self.has_massive_nu = self.Tcmb0 > 0 and any(m > 0 for m in self.m_nu)
        self.Onu0 = self.Ogamma0 * self._nu_relative_density(0.) if self.Tcmb0 > 0 else 0.
        self.Ode0 = 1. - self.Om0 - self.Ogamma0 - self.Onu0
        self.Odm0 = None if self.Ob0 is None else self.Om0 - self.Ob0

    def _nu_relative_density(self, z):
        z = np.asarray(z, dtype=float)
        prefac = 0.22710731766                      # 7/8 (4/11)^(4/3)
        if not self.has_massive_nu:
            return prefac * self.Neff * np.ones_like(z)
        p, invp, k = 1.83, 0.54644808743, 0.3173   # Komatsu et al. 2011 fit
        Tnu0 = 0.7137658555036082 * self.Tcmb0
        masses = np.array(self.m_nu)
        nu_y = masses[masses > 0] / (_KB_EV_K * Tnu0)
        n_massless = 3 - len(nu_y) if len(masses) <= 3 else 0
        curr = nu_y / (1. + z[..., None])
        rel = (1. + (k * curr) ** p) ** invp
        return prefac * (self.Neff / 3.) * (rel.sum(-1) + n_massless)

    def inv_efunc(self, z):
        z = np.asarray(z, dtype=float)
        zp1 = 1. + z
        Or = self.Ogamma0 * (1. + self._nu_relative_density(z)) if self.Tcmb0 > 0 else 0.
        return 1. / np.sqrt(zp1 ** 3 * (Or * zp1 + self.Om0) + self.Ode0)

    def Om(self, z):
        z = np.asarray(z, dtype=float)
        return self.Om0 * (1. + z) ** 3 * self.inv_efunc(z) ** 2

    def comoving_distance(self, z):
        """Line-of-sight comoving distance in Mpc (Gauss-Legendre on [0, z])."""
        z = np.asarray(z, dtype=float)
        x, w = np.polynomial.legendre.leggauss(64)
        zz = 0.5 * z[..., None] * (x + 1.)
        integral = 0.5 * z * np.sum(w * self.inv_efunc(zz), axis=-1)
        return (_C_KM_S / self.H0) * integral

    comoving_transverse_distance = comoving_distance   # flat


_NAMED = {
    'Planck13': dict(H0=67.77, Om0=0.30712, Tcmb0=2.7255, Neff=3.046, m_nu=(0., 0., 0.06), Ob0=0.048252),
    'WMAP9': dict(H0=69.32, Om0=0.2865, Tcmb0=2.725, Neff=3.04, m_nu=(0., 0., 0.), Ob0=0.04628),
    'WMAP7': dict(H0=70.4, Om0=0.272, Tcmb0=2.725, Neff=3.04, m_nu=(0., 0., 0.), Ob0=0.0455),
    'WMAP5': dict(H0=70.2, Om0=0.277, Tcmb0=2.725, Neff=3.04, m_nu=(0., 0., 0.), Ob0=0.0459),
}


def create_cosmology(*args, **kwargs):
    """Create a background cosmology model (cosmotools.py:14-45): a pre-defined name
    ('WMAP5', 'WMAP7', 'WMAP9', 'Planck13'; default 'Planck13') or FlatLambdaCDM parameters."""
    if len(args) > 0 and len(kwargs) > 0:
        raise TypeError('Cannot specify both a name and parameters.')
    elif len(args) > 1:
        raise TypeError('Invalid arguments: expected a name or parameters.')
    elif len(args) == 1 and not isinstance(args[0], str):
        raise TypeError('Invalid name: {0}.'.format(args[0]))
    if len(args) == 0 and len(kwargs) == 0:
        args = ['Planck13']
    if len(args) == 1:
        if args[0] not in _NAMED:
            raise ValueError('Not a recognized cosmology: {0}.'.format(args[0]))
        if _astro is not None:
            return getattr(_astro, args[0])
        return LiteFlatLambdaCDM(name=args[0], **_NAMED[args[0]])
    if _astro is not None:
        return _astro.FlatLambdaCDM(**kwargs)
    kwargs.setdefault('Tcmb0', 0.)
    kwargs.setdefault('m_nu', (0., 0., 0.))
    return LiteFlatLambdaCDM(**kwargs)


def get_class_parameters(cosmology):
    """CLASS parameter dictionary of a background model (cosmotools.py:48-97).  Works on astropy
    FLRW objects and on LiteFlatLambdaCDM; raises ValueError when an attribute is missing."""
    out = {}
    try:
        out['h'] = cosmology.h
        out['T_cmb'] = _value(cosmology.Tcmb0)
        out['Omega_b'] = cosmology.Ob0
        out['Omega_cdm'] = cosmology.Odm0 - cosmology.Onu0      # cold = dark - massive neutrinos
        out['Omega_k'] = cosmology.Ok0
        out['Omega_scf'] = 0.                                    # CLASS fills in the component left out
        if hasattr(cosmology, 'w0'):
            out.update(w0_fld=cosmology.w0, wa_fld=0., Omega_Lambda=0.)
        else:
            out['Omega_fld'] = 0.
        n_massive = 0
        if cosmology.has_massive_nu:
            masses = cosmology.m_nu
            if np.ndim(_value(masses)) == 0:
                n_massive, masses = 3, [3 * masses]              # one value = three degenerate species
            else:
                n_massive = len(masses)
            if n_massive > 3:
                raise RuntimeError('Cannot calculate cosmology with >3 massive neutrinos.')
            heavy = [repr(float(_value(m))) for m in masses if _value(m) > 0]
            n_massive -= len(masses) - len(heavy)
            if n_massive > 0:
                out['m_ncdm'] = ','.join(heavy)
        out['N_ncdm'] = n_massive
        out['N_ur'] = 3.00641 - n_massive + (3 - n_massive) * (cosmology.Neff - 3.00641) / 3.
    except AttributeError:
        raise ValueError('Cosmology is missing required attributes.')
    return out


def calculate_power(cosmology, k_min, k_max, z=0, num_k=500, scaled_by_h=True,
                    n_s=0.9619, logA=3.0980):
    """P(k, z) on num_k log-spaced points of [k_min, k_max] from CLASS (cosmotools.py:100-137).
    Host-side set-up, outside the device path; without the `classy` package it raises the
    reference's RuntimeError, which is what Generator(cosmology=..., power=None) reports."""
    try:
        from classy import Class
        engine = Class()
    except ImportError:
        raise RuntimeError('power.calculate_power requires classy.')
    pars = get_class_parameters(cosmology)
    pars['output'] = 'mPk'
    pars['P_k_max_h/Mpc' if scaled_by_h else 'P_k_max_1/Mpc'] = k_max
    pars['n_s'] = n_s
    pars['ln10^{10}A_s'] = logA
    engine.set(pars)
    engine.compute()
    h = engine.h() if scaled_by_h else 1.
    table = np.empty((num_k,), dtype=[('k', float), ('Pk', float)])
    table['k'][:] = np.logspace(np.log10(k_min), np.log10(k_max), num_k)
    table['Pk'][:] = [engine.pk(k * h, z) * h ** 3 for k in table['k']]
    engine.struct_cleanup()
    engine.empty()
    return table


def _value(q, unit=None):
    """Strip astropy units (converting first) or pass plain numbers through."""
    if hasattr(q, 'to') and unit is not None:
        return q.to(unit).value
    return getattr(q, 'value', q)


def get_redshifts(cosmology, data, spacing, scaled_by_h=True, z_axis=2,
                  num_interpolation_points=100):
    """Redshift of every grid plane along z_axis, plane-parallel (cosmotools.py:140-175)."""
    if len(data.shape) != 3:
        raise ValueError('Input data is not 3D.')
    try:
        nz = data.shape[z_axis]
    except (IndexError, TypeError):
        raise ValueError('Invalid z_axis: {0}.'.format(z_axis))
    output_shape = np.ones((3,), dtype=int)
    output_shape[z_axis] = nz
    comoving_distances = np.arange(nz) * float(spacing)
    if scaled_by_h:
        comoving_distances = comoving_distances / cosmology.h
    d_max = comoving_distances[-1]
    mpc = _u.Mpc if _u is not None else None

    def dist(z):
        return _value(cosmology.comoving_distance(z), mpc)

    # z_at_value(comoving_distance, d_max): bracket then bisect
    z_hi = 1.0
    while dist(z_hi) < d_max:
        z_hi *= 2.
    z_lo = 0.
    for _ in range(200):
        z_mid = 0.5 * (z_lo + z_hi)
        if dist(z_mid) < d_max:
            z_lo = z_mid
        else:
            z_hi = z_mid
    max_redshift = 0.5 * (z_lo + z_hi)
    redshift_grid = np.linspace(0., 1.05 * max_redshift, num_interpolation_points)
    comoving_distance_grid = dist(redshift_grid)
    z_interpolator = scipy.interpolate.interp1d(
        comoving_distance_grid, redshift_grid, kind='cubic', copy=False)
    redshifts = z_interpolator(comoving_distances)
    redshifts[0] = 0
    return redshifts.reshape(output_shape)


def get_mean_matter_densities(cosmology, redshifts):
    """Mean matter densities in g / cm**3 (cosmotools.py:178-184)."""
    gcm3 = (_u.gram / _u.cm ** 3) if _u is not None else None
    mean_density0 = _value(cosmology.critical_density0, gcm3) * cosmology.Om0
    return mean_density0 * (1 + redshifts) ** 3


def get_growth_function(cosmology, redshifts):
    """Linder 2005 growth approximation (cosmotools.py:187-203)."""
    z_axis = np.arange(3)[np.argsort(redshifts.shape)][-1]
    w0 = getattr(cosmology, 'w0', -1)
    gamma = 0.55 + 0.05 * (1 + w0)
    integrand = np.power(cosmology.Om(redshifts), gamma) / (1 + redshifts)
    exponent = scipy.integrate.cumulative_trapezoid(
        y=integrand, x=redshifts, axis=z_axis, initial=0)
    return np.exp(-exponent)


def apply_lognormal_transform(delta, growth, sigma=None):
    """Normal(0, sigma) -> log-normal with mean one and std growth*sigma, in place
    (cosmotools.py:206-221), evaluated on the GPU.  ``growth`` is a scalar or broadcasts
    along the last axis of ``delta``; ``sigma=None`` uses the standard deviation of delta."""
    from . import _lib
    nz = delta.shape[-1]
    g = np.ascontiguousarray(np.broadcast_to(np.asarray(growth, dtype=np.float64).reshape(-1), (nz,)))
    rho = np.ones(nz, dtype=np.float64)
    work = np.ascontiguousarray(delta)
    _lib.check(_lib.lib().rf_stage_density(work.size, nz, _lib.dtype_code(work.dtype), 1,
                                           -1.0 if sigma is None else float(sigma),
                                           _lib.ptr(g), _lib.ptr(rho), _lib.ptr(work)))
    if work is not delta:
        delta[...] = work
    return delta
