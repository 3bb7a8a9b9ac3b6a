"""TEST INFRASTRUCTURE ONLY -- loads the UNMODIFIED reference modules.

From /root/reference in the build container; on the GPU box, which has no reference tree, from
the byte-for-byte staged copy under oracle/_ref/ (git-ignored; recipe and hash manifest:
oracle/make_ref.py, run by __graft_entry__.build()).  It exists to
(1) pin the numpy restatement in oracle/oracle.py against the real reference,
(2) generate the golden fixtures under tests/golden/ (tests/golden/make_golden.py) and
(3) be the timed CPU arm of bench.py (--impl reference, cpu_baseline.kind = "reference").

The reference is Python-2 era: transform.py / powertools.py / random.py import each other
with implicit-relative imports and use numpy names removed in numpy 2.  We load them by file
path under their bare names, after defining the removed aliases.
generate.py and cosmotools.py cannot be imported (astropy is absent), so replay() restates
the nine calls of generate.py:191-219 against the loaded modules.
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

REFERENCE_ROOT = os.environ.get("RF_REFERENCE_ROOT", "/root/reference")
_STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")
if not os.path.exists(os.path.join(REFERENCE_ROOT, "randomfield", "transform.py")) and \
        os.path.exists(os.path.join(_STAGED, "randomfield", "transform.py")):
    REFERENCE_ROOT = _STAGED
_PKG = os.path.join(REFERENCE_ROOT, "randomfield")
_mods = {}


def available() -> bool:
    return os.path.exists(os.path.join(_PKG, "transform.py"))


def staged() -> bool:
    """True when the modules come from the staged copy (oracle/_ref) rather than the reference tree."""
    return REFERENCE_ROOT == _STAGED


def _shim_numpy():
    aliases = {
        "complex_": np.complex128, "float_": np.float64, "clongfloat": np.clongdouble,
        "longfloat": np.longdouble, "float": float, "int": int,
    }
    for name, val in aliases.items():
        if not hasattr(np, name):
            setattr(np, name, val)
    if not hasattr(np, "obj2sctype"):
        def obj2sctype(rep, default=None):
            try:
                if isinstance(rep, np.ndarray):
                    return rep.dtype.type
                return np.dtype(rep).type
            except Exception:
                return default
        np.obj2sctype = obj2sctype


def _load(name, filename, as_name=None):
    spec = importlib.util.spec_from_file_location(as_name or name, os.path.join(_PKG, filename))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[as_name or name] = mod
    spec.loader.exec_module(mod)
    return mod


def modules():
    """Return (transform, powertools, random) reference modules."""
    if not _mods:
        if not available():
            raise RuntimeError("reference tree not found at %s" % REFERENCE_ROOT)
        _shim_numpy()
        # the reference re-imports itself by bare name at call time (powertools.py:185), so the
        # bare names stay registered; the product lives under randomfield_b200.* and never
        # uses them
        _mods["transform"] = _load("transform", "transform.py")
        _mods["powertools"] = _load("powertools", "powertools.py")
        _mods["random"] = _load("rf_ref_random", "random.py", as_name="rf_ref_random")
    return _mods["transform"], _mods["powertools"], _mods["random"]


def replay(nx, ny, nz, spacing, seed=None, smoothing=0.0, power=None, dtype=np.complex64,
           noise=None, save_potential=False):
    """Replay generate.py:191-219 on the reference modules.

    Returns dict(delta, rms, kfield (copy before the transform), potential (or None)).
    If `noise` is given it replaces RandomState(seed).normal(size) (random.py:24-28) so
    the same white noise can be injected into the CUDA path.
    """
    transform, powertools, rrandom = modules()
    plan = transform.Plan(shape=(nx, ny, nz), dtype_in=dtype, packed=True, overwrite=True,
                          inverse=True, use_pyfftw=True)
    if power is None:
        power = powertools.load_default_power()
    powertools.fill_with_log10k(plan.data_in, spacing=spacing, packed=True)
    smoothed = powertools.filter_power(power, smoothing)
    powertools.tabulate_sigmas(plan.data_in, power=smoothed, spacing=spacing, packed=True)
    if noise is None:
        rrandom.randomize(plan.data_in, seed=seed)
    else:
        # random.py:17-28 with the deviates supplied by the caller
        plan.data_in.imag = plan.data_in.real
        real_type = transform.scalar_type(plan.data_in.dtype)
        sigmas = plan.data_in.view(real_type).reshape(2 * plan.data_in.size)
        sigmas *= np.asarray(noise, dtype=np.float64).reshape(-1)
    transform.symmetrize(plan.data_in, packed=True)
    kfield = plan.data_in.copy()
    potential = None
    if save_potential:
        # generate.py:200-215
        potential = np.empty_like(plan.data_in)
        potential.imag = 0.
        kx2, ky2, kz2 = powertools.create_ksq_grids(potential, spacing=spacing, packed=True)
        np.add(kx2, ky2, out=potential.real)
        potential.real += kz2
        old = np.seterr(divide='ignore')
        np.reciprocal(potential.real, out=potential.real)
        np.seterr(**old)
        potential[0, 0, 0] = 0.
        potential *= plan.data_in
    delta = plan.execute()
    rms = np.std(delta.flat)
    return dict(delta=np.array(delta), rms=rms, kfield=kfield, potential=potential)


def default_power():
    return modules()[1].load_default_power()


# ------------------------------------------------------------------------------------------
# calculate_lensing_potential (generate.py:353-418) -- the reference's own loop on stubs
# ------------------------------------------------------------------------------------------
def _generate_module():
    """Import the reference's generate.py unmodified.  It cannot be imported as it stands here:
    astropy is absent (generate.py:13-14, and cosmotools.py needs astropy.cosmology) and scipy >=
    1.14 has no integrate.simps (generate.py:408).  Stubs for exactly those names are installed
    first: astropy.units / astropy.constants with the one quantity the method reads (the speed of
    light in km/s), an empty `cosmotools`, and scipy.integrate.simps = oracle.simps_old, the
    restatement of the rule scipy shipped when the reference was written.  Only stage the copy
    when generate.py is there (oracle/make_ref.py copies the three modules of the hot path; this
    needs the reference tree itself)."""
    if "generate" in _mods:
        return _mods["generate"]
    import types
    import warnings
    import scipy.integrate
    path = os.path.join(_PKG, "generate.py")
    if not os.path.exists(path):
        raise RuntimeError("generate.py not found under %s" % _PKG)
    modules()
    import oracle

    class _Quantity(object):
        def __init__(self, value):
            self.value = value

        def to(self, unit):
            return self

    units = types.ModuleType("astropy.units")

    class _Unit(object):
        def __truediv__(self, other):
            return self
        __div__ = __truediv__
    units.km = units.s = units.Mpc = _Unit()
    constants = types.ModuleType("astropy.constants")
    constants.c = _Quantity(299792.458)          # km/s, what astropy.constants.c.to(u.km / u.s).value gives
    astropy = types.ModuleType("astropy")
    astropy.units, astropy.constants = units, constants
    saved = {k: sys.modules.get(k) for k in ("astropy", "astropy.units", "astropy.constants", "cosmotools")}
    sys.modules.update({"astropy": astropy, "astropy.units": units, "astropy.constants": constants,
                        "cosmotools": types.ModuleType("cosmotools")})
    if not hasattr(scipy.integrate, "simps"):
        def simps(y, x=None, dx=1, axis=-1, even='avg'):
            assert axis == -1 and even == 'avg' and x is not None
            return oracle.simps_old(y, x)
        scipy.integrate.simps = simps
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")          # invalid escape sequences in the docstrings
            mod = _load("rf_ref_generate", "generate.py", as_name="rf_ref_generate")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    _mods["generate"] = mod
    return mod


def generate_available() -> bool:
    return os.path.exists(os.path.join(_PKG, "generate.py"))


def lensing_replay(dphi, DC, DA, Ok0=0.0, i_min=None):
    """Generator.calculate_lensing_potential (generate.py:353-418) run UNMODIFIED on a stub object
    that carries the five attributes the method reads."""
    import types
    gen = _generate_module()
    stub = types.SimpleNamespace(DC=np.asarray(DC), DA=np.asarray(DA), cosmology=types.SimpleNamespace(Ok0=Ok0),
                                 plan_c2r=types.SimpleNamespace(data_out=dphi))
    return gen.Generator.calculate_lensing_potential(stub, i_min=i_min)
