"""CPU tests of the cosmology tables the epilogue kernels take as input:
the reference's analytic checks (randomfield/tests/test_cosmotools.py:11-83) re-hosted on
randomfield_b200.cosmotools, which runs without astropy (LiteFlatLambdaCDM)."""
import numpy as np
import pytest

from randomfield_b200 import cosmotools
from randomfield_b200.cosmotools import (create_cosmology, get_redshifts, get_growth_function,
                                         get_mean_matter_densities, LiteFlatLambdaCDM)

_C_KM_S = 299792.458


def test_create_cosmology_arguments():
    """tests/test_cosmotools.py:11-27: a name or parameters, never both."""
    assert create_cosmology().h == pytest.approx(0.6777)
    assert create_cosmology('WMAP9').Om0 == pytest.approx(0.2865)
    model = create_cosmology(H0=65, Om0=0.25)
    assert model.h == pytest.approx(0.65) and model.Om0 == pytest.approx(0.25) and model.Ok0 == 0
    with pytest.raises(TypeError):
        create_cosmology('Planck13', H0=65)
    with pytest.raises(TypeError):
        create_cosmology('WMAP9', 'Planck13')
    with pytest.raises(TypeError):
        create_cosmology(123)
    with pytest.raises(ValueError):
        create_cosmology('NoSuchCosmology')


def test_redshifts():
    """tests/test_cosmotools.py:55-72: matter-only flat universe, z = r (4 - r) / (r - 2)^2 with
    r = Dc(z) / (c / H0)."""
    nz = 100
    spacing = 2.5
    data = np.empty((1, 2, nz))
    model = create_cosmology(H0=70, Om0=1, Tcmb0=0)
    Dc0 = _C_KM_S / 70.0
    for scale in (1, model.h):
        r = np.arange(nz) * spacing / scale / Dc0
        analytic_redshifts = r * (4 - r) / (r - 2) ** 2
        computed_redshifts = get_redshifts(model, data, spacing=spacing, scaled_by_h=(scale != 1))
        assert computed_redshifts.shape == (1, 1, nz)
        assert np.allclose(analytic_redshifts, computed_redshifts.reshape(nz))
    with pytest.raises(ValueError):
        get_redshifts(model, np.empty((2, nz)), spacing=spacing)
    with pytest.raises(ValueError):
        get_redshifts(model, data, spacing=spacing, z_axis=3)


def test_growth():
    """tests/test_cosmotools.py:75-83: matter-only flat universe, G(z) = 1 / (1 + z)."""
    nz = 100
    z = np.linspace(0., 1., nz).reshape(1, nz, 1)
    model = create_cosmology(H0=70, Om0=1, Tcmb0=0)
    Gz = get_growth_function(model, z)
    assert Gz.shape == (1, nz, 1)
    assert np.allclose(Gz, 1 / (1 + z))


def test_comoving_distance_against_closed_forms():
    """Einstein-de Sitter: Dc = 2 c/H0 (1 - 1/sqrt(1+z)); de Sitter-like limit Om0 -> 0: Dc = c z / H0."""
    z = np.linspace(0.0, 3.0, 31)
    eds = LiteFlatLambdaCDM(H0=70, Om0=1, Tcmb0=0, m_nu=(0., 0., 0.))
    assert np.allclose(eds.comoving_distance(z), 2 * _C_KM_S / 70 * (1 - 1 / np.sqrt(1 + z)), rtol=1e-10)
    ds = LiteFlatLambdaCDM(H0=70, Om0=1e-12, Tcmb0=0, m_nu=(0., 0., 0.))
    assert np.allclose(ds.comoving_distance(z), _C_KM_S / 70 * z, rtol=1e-8)
    assert np.allclose(eds.Om(z), 1.0)


def test_mean_matter_density_and_planck13_budget():
    """rho_mean(z) = rho_crit0 Om0 (1+z)^3 in g/cm^3 (cosmotools.py:178-184); Planck13 closes:
    Om0 + Ogamma0 + Onu0 + Ode0 = 1 with the massive-neutrino term of astropy's Planck13."""
    model = create_cosmology()
    z = np.array([0.0, 1.0])
    rho = get_mean_matter_densities(model, z)
    h = model.h
    rho_crit = 1.87847e-29 * h * h               # g/cm^3
    assert rho[0] == pytest.approx(rho_crit * model.Om0, rel=2e-4)
    assert rho[1] == pytest.approx(8 * rho[0])
    if isinstance(model, LiteFlatLambdaCDM):
        assert model.Om0 + model.Ogamma0 + model.Onu0 + model.Ode0 == pytest.approx(1.0)
        assert model.Ogamma0 == pytest.approx(5.39e-5, rel=5e-3)
        # comoving distance to z = 1 for Planck13: 3395.9 Mpc (astropy 1.x / 5.x)
        assert float(model.comoving_distance(1.0)) == pytest.approx(3395.9, rel=1e-3)


def test_generator_tables_are_consistent():
    """The tables Generator.__init__ derives (generate.py:104-129): redshifts monotonic from 0,
    growth decreasing from 1, rho_mean growing, DA == DC for a flat model."""
    nz, spacing = 64, 4.0
    model = create_cosmology()
    z = get_redshifts(model, np.empty((1, 1, nz)), spacing=spacing, scaled_by_h=True, z_axis=2).reshape(nz)
    assert z[0] == 0 and np.all(np.diff(z) > 0)
    G = get_growth_function(model, z)
    assert G[0] == 1 and np.all(np.diff(G) < 0)
    rho = get_mean_matter_densities(model, z)
    assert np.all(np.diff(rho) > 0)
    dc = cosmotools._value(model.comoving_distance(z)) * model.h
    assert np.allclose(dc, np.arange(nz) * spacing, atol=2e-3 * spacing)


def test_class_parameters_of_the_default_model():
    """cosmotools.get_class_parameters (cosmotools.py:48-97) on the Planck13 stand-in: one massive
    species of 0.06 eV, the other two light; a missing attribute is a ValueError."""
    c = create_cosmology()
    if not isinstance(c, cosmotools.LiteFlatLambdaCDM):
        pytest.skip("astropy model in use")
    p = cosmotools.get_class_parameters(c)
    assert p['h'] == pytest.approx(0.6777) and p['T_cmb'] == pytest.approx(2.7255)
    assert p['Omega_b'] == pytest.approx(0.048252)
    assert p['Omega_cdm'] == pytest.approx(0.30712 - 0.048252 - c.Onu0)
    assert p['Omega_k'] == 0. and p['Omega_scf'] == 0. and p['Omega_fld'] == 0.
    assert 'w0_fld' not in p
    assert p['N_ncdm'] == 1 and p['m_ncdm'] == '0.06'
    assert p['N_ur'] == pytest.approx(3.00641 - 1 + 2 * (3.046 - 3.00641) / 3.)
    # massless neutrinos: no m_ncdm entry, three light species
    q = cosmotools.get_class_parameters(cosmotools.LiteFlatLambdaCDM(H0=70., Om0=0.3, Ob0=0.05, m_nu=(0., 0., 0.)))
    assert q['N_ncdm'] == 0 and 'm_ncdm' not in q and q['N_ur'] == pytest.approx(3.046)
    with pytest.raises(ValueError):
        cosmotools.get_class_parameters(object())


def test_calculate_power_without_class_raises_the_reference_error():
    """cosmotools.py:105-109: no classy -> RuntimeError('power.calculate_power requires classy.')."""
    try:
        import classy  # noqa: F401
        pytest.skip("classy is installed")
    except ImportError:
        pass
    with pytest.raises(RuntimeError, match='requires classy'):
        cosmotools.calculate_power(create_cosmology(), 1e-3, 1.0)
