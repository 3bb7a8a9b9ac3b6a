"""CPU tests: the oracle against the committed golden vectors (made from the unmodified
reference), against the reference itself when /root/reference is present, and the reference's
own known-answer tests re-hosted on the oracle."""
import os

import numpy as np
import pytest
from scipy.special import erf

import oracle
import ref_replay

CASES = ["ref_64_f32", "ref_64_f32_smooth", "ref_128x64x256_f32", "ref_16x8x12_f64", "ref_12x10x8_f32"]


def _load(golden_dir, name):
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    shape = tuple(int(v) for v in g["shape"])
    return g, shape, float(g["spacing"]), int(g["seed"]), float(g["smoothing"]), np.dtype(str(g["dtype"]))


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(golden_dir, name):
    g, shape, spacing, seed, smoothing, dtype = _load(golden_dir, name)
    r = oracle.generate_delta_field(shape, spacing, oracle.load_default_power(), seed=seed,
                                    smoothing=smoothing, dtype=dtype, save_potential=True)
    sd, sk = g["delta_stride"], g["kfield_stride"]
    # the k-space field is pure elementwise arithmetic + MT19937: bit-exact
    assert np.array_equal(r["kfield"][::sk[0], ::sk[1], ::sk[2]], g["kfield"])
    assert np.array_equal(r["potential"][::sk[0], ::sk[1], ::sk[2]], g["potential"])
    # the FFT is numpy's pocketfft on both sides: identical on the same numpy build, and
    # equal to rounding error on any other
    tol = 5e-6 if dtype == np.complex64 else 1e-13
    d = r["delta"][::sd[0], ::sd[1], ::sd[2]]
    assert np.sqrt(np.mean((d - g["delta"]) ** 2)) <= tol * float(g["rms"])
    assert abs(r["rms"] - g["rms"]) <= tol * float(g["rms"])


@pytest.mark.skipif(not ref_replay.available(), reason="reference tree not mounted")
@pytest.mark.parametrize("shape,smoothing,dtype", [((32, 16, 24), 0.0, np.complex64),
                                                   ((8, 6, 4), 2.0, np.complex128)])
def test_oracle_matches_reference_bitwise(shape, smoothing, dtype):
    r = ref_replay.replay(*shape, 2.0, seed=31, smoothing=smoothing, dtype=dtype, save_potential=True)
    o = oracle.generate_delta_field(shape, 2.0, oracle.load_default_power(), seed=31,
                                    smoothing=smoothing, dtype=dtype, save_potential=True)
    for key in ("kfield", "potential", "delta"):
        assert np.array_equal(r[key], o[key]), key
    assert r["rms"] == o["rms"]


def test_symmetrize_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_symmetrize.npz"))
    for key in ("packed_complex64", "packed_complex128", "full_complex64", "full_complex128"):
        a = g[key + "_in"].copy()
        packed = key.startswith("packed")
        oracle.symmetrize(a, packed=packed)
        assert np.array_equal(a, g[key + "_out"])
        assert oracle.is_hermitian(a, packed=packed)


def test_randomstate_stream(golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_misc.npz"))
    assert np.array_equal(np.random.RandomState(123).normal(size=8), g["normal123"])


def test_gaussian_variance_kat():
    """Reference tests/test_generate.py:24-62 on the oracle."""
    spacing, n, seed = 2.5, 64, 123
    kmin, kmax = (2 * np.pi) / (spacing * n), np.pi / spacing
    sigma, P0 = 2.5 * spacing, 1.23
    calc = P0 / (2 * np.pi) ** 1.5 / sigma ** 3 * (
        erf(kmax * sigma / np.sqrt(2)) ** 3 - erf(kmin * sigma / np.sqrt(2)) ** 3)
    power = np.empty(100, dtype=[('k', float), ('Pk', float)])
    power['k'] = np.linspace(kmin, np.sqrt(3) * kmax, len(power))
    power['Pk'] = P0 * np.exp(-0.5 * (power['k'] * sigma) ** 2)
    var = np.mean([np.var(oracle.generate_delta_field((n, n, n), spacing, power, seed=seed + t)["delta"])
                   for t in range(10)])
    assert abs(var - calc) < 0.01 * calc


def test_lognormal_kat():
    """Reference tests/test_cosmotools.py:86-99 on the oracle."""
    growth, sigma = 0.3, 2.5
    np.random.seed(123)
    delta = np.empty((64, 64, 128), dtype=np.float32)
    delta[:] = sigma * np.random.normal(size=delta.shape)
    rho = oracle.apply_lognormal_transform(delta, growth, sigma=2.5)
    assert np.all(rho > 0)
    assert abs(np.mean(rho) - 1.) < 1e-3
    assert abs(np.std(rho) - growth * sigma) < 1e-2 * sigma


def test_philox_kat():
    """Random123 known-answer vectors for Philox4x32-10."""
    def run(c, k):
        return [int(x[0]) for x in oracle.philox4x32_10(*[np.array([v]) for v in c], *k)]
    assert run((0, 0, 0, 0), (0, 0)) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert run((0xffffffff,) * 4, (0xffffffff,) * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert run((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0)) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_pairing_rule_equals_reference_symmetrize():
    shape = (8, 6, 4)
    rng = np.random.RandomState(2)
    a = rng.normal(size=(8, 6, 3)) + 1j * rng.normal(size=(8, 6, 3))
    b = a.copy()
    oracle.symmetrize(b, packed=True)
    mx, my, slave, vertex = oracle.hermitian_pairing(shape)
    c = a.copy()
    for iz in (0, 2):
        pl = a[mx, my, iz]
        pl = np.where(slave, np.conj(pl), pl)
        c[:, :, iz] = np.where(vertex, pl.real + 0j, pl)
    c[0, 0, 0] = 0
    assert np.array_equal(b, c)


def test_philox_field_statistics():
    """The Philox k-field has the reference's statistics: P(k)-weighted variance (KAT) and
    is Hermitian, so its transform is real."""
    shape, spacing = (32, 32, 32), 2.5
    power = oracle.load_default_power()
    kf, bins = oracle.philox_kfield(shape, spacing, power, seed=5)
    assert oracle.is_hermitian(kf, packed=True)
    full = np.fft.ifftn(np.concatenate([kf, np.conj(np.roll(np.roll(kf[:, :, 1:-1][:, :, ::-1][::-1, ::-1], 1, 0), 1, 1))], axis=2))
    assert np.abs(full.imag).max() < 1e-9 * np.abs(full.real).max()
    assert bins.min() >= -1 and bins.max() < len(power)


LENS_CASES = ["flat_f32", "open_f32", "closed_f64", "odd_f64"]


@pytest.mark.parametrize("name", LENS_CASES)
def test_lensing_oracle_matches_golden(golden_dir, name):
    """oracle.lensing_potential against psi computed by the reference's OWN loop (generate.py:353-418
    run unmodified on stub attributes, tests/golden/make_golden.py).  The reference stores its
    integrand in dPhi's dtype before integrating, the oracle evaluates it in float64: equal to
    float32 rounding for float32 fields, to 1e-13 for float64 ones."""
    g = np.load(os.path.join(golden_dir, "ref_lensing.npz"))
    imin, ok0 = g[name + "_par"]
    psi = oracle.lensing_potential(g[name + "_dphi"], g[name + "_DC"], g[name + "_DA"], Ok0=float(ok0),
                                   i_min=None if imin < 0 else int(imin))
    ref = g[name + "_psi"]
    assert psi.dtype == ref.dtype and psi.shape == ref.shape
    scale = np.abs(ref.astype(np.float64)).max()
    tol = 3e-7 if ref.dtype == np.float32 else 1e-13
    assert np.abs(psi.astype(np.float64) - ref).max() <= tol * scale


@pytest.mark.skipif(not ref_replay.generate_available(), reason="reference tree not mounted")
def test_lensing_oracle_matches_reference_loop():
    rng = np.random.RandomState(3)
    for nz, imin, ok0 in ((64, None, 0.0), (33, 3, 0.0), (40, 0, 0.01), (17, 1, -0.02), (8, 7, 0.0)):
        dphi = (1e-33 * rng.normal(size=(3, 4, nz))).astype(np.float32)
        DC = np.arange(nz) * 2.0
        DA = DC * (1 + 0.001 * np.arange(nz))
        ref = ref_replay.lensing_replay(dphi, DC, DA, Ok0=ok0, i_min=imin)
        got = oracle.lensing_potential(dphi, DC, DA, Ok0=ok0, i_min=imin)
        assert np.abs(got.astype(np.float64) - ref).max() <= 3e-7 * np.abs(ref).max()
    with pytest.raises(ValueError):
        ref_replay.lensing_replay(dphi, DC, DA, i_min=nz)
