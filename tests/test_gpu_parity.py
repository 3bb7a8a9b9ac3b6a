"""GPU parity tests (run with -m gpu on the B200 box).  Everything goes through the C ABI
(librf_b200.so via ctypes); the oracle is only the checker.

Gates (BASELINE.json north_star):
 1. injected white noise: rel RMS(delta_gpu - delta_ref)/std <= 1e-5 (f32), <= 1e-12 (f64)
 2. Hermitian pairing and sigma-table bin index bit-exact
 3. (multi-GPU identity lives in test_gpu_multi.py)
 4. measured P(k) recovers the input spectrum within cosmic variance
"""
import os

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

TOL_F32 = 1e-5
TOL_F64 = 1e-12


@pytest.fixture(scope="module")
def lib():
    from randomfield_b200 import _lib
    _lib.lib()
    return _lib


def _engine(lib, shape, spacing, power, smoothing=0.0, dtype=np.float32):
    from randomfield_b200 import powertools
    eng = lib.Engine(*shape, dtype=dtype)
    k, sig = powertools.sigma_table(powertools.filter_power(power, smoothing), shape, spacing)
    eng.set_power(k, sig, spacing)
    return eng


def _relrms(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.sqrt(np.mean((a - b) ** 2)) / np.std(b)


# ------------------------------------------------------------------------------------------
# gate 2: bit-exact integer work
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,spacing", [((64, 64, 64), 2.0), ((128, 64, 256), 2.0), ((12, 10, 8), 2.5)])
def test_bins_and_pairing_bit_exact(lib, shape, spacing):
    power = oracle.load_default_power()
    eng = _engine(lib, shape, spacing, power)
    kf, bins = eng.kspace_public(seed=77, with_bins=True)
    ksq = oracle.canonical_ksq(shape, spacing, np.float32)
    assert np.array_equal(bins, oracle.canonical_bins(ksq, power['k'], np.float32))
    # pairing: slaves are the exact conjugate of their master, vertices are real, DC is zero
    nx, ny, nz = shape
    mx, my, slave, vertex = oracle.hermitian_pairing(shape)
    for iz in (0, nz // 2):
        plane = kf[:, :, iz]
        want = np.where(slave, np.conj(plane[mx, my]), plane)
        assert np.array_equal(plane, want)
        assert np.all(plane.imag[vertex] == 0)
    assert kf[0, 0, 0] == 0
    from randomfield_b200 import transform
    assert transform.is_hermitian(kf, packed=True)


def test_bins_vs_numpy_reference_mismatch_count(lib):
    """The reference picks its bin from float32 log10 k (powertools.py:56-60,155-163); ours is
    the canonical float32 k^2 compare.  They may differ only for modes within rounding of a
    knot; count them."""
    shape, spacing = (128, 128, 128), 2.0
    power = oracle.load_default_power()
    eng = _engine(lib, shape, spacing, power)
    _, bins = eng.kspace_public(seed=1, with_bins=True)
    data = np.empty((128, 128, 65), dtype=np.complex64)
    oracle.fill_with_log10k(data, spacing)
    ref_bins = np.searchsorted(np.log10(power['k']), data.real.astype(np.float64), side='right') - 1
    ref_bins[0, 0, 0] = -1
    mism = int(np.count_nonzero(ref_bins != bins))
    print("bin mismatches vs numpy reference: %d of %d" % (mism, bins.size))
    assert mism <= 1e-4 * bins.size


# ------------------------------------------------------------------------------------------
# gate 1: injected white noise vs the reference algorithm
# ------------------------------------------------------------------------------------------
INJECT_CASES = [
    ((64, 64, 64), 2.0, 0.0, np.float32, 123),
    ((64, 64, 64), 2.0, 0.0, np.float64, 123),
    ((128, 64, 256), 2.0, 0.0, np.float32, 5),
    ((256, 256, 256), 2.0, 4.0, np.float32, 123),     # BASELINE config 2
    ((256, 256, 256), 2.0, 4.0, np.float64, 123),     # BASELINE config 2, float64 mode
    ((12, 10, 8), 2.5, 0.0, np.float32, 9),           # generic (non power-of-two) path
    ((16, 8, 12), 2.5, 1.0, np.float64, 7),
]


@pytest.mark.parametrize("shape,spacing,smoothing,dtype,seed", INJECT_CASES)
def test_injected_noise_matches_reference(lib, shape, spacing, smoothing, dtype, seed):
    power = oracle.load_default_power()
    nx, ny, nz = shape
    n = 2 * nx * ny * (nz // 2 + 1)
    noise = np.random.RandomState(seed).normal(size=n)     # what random.py:24-28 draws
    cdtype = np.complex64 if dtype == np.float32 else np.complex128
    ref = oracle.generate_delta_field(shape, spacing, power, smoothing=smoothing, dtype=cdtype,
                                      noise=noise)
    # float64 evaluation of the same k-field: the truth line for the f32 gate
    truth = np.fft.irfftn(ref["kfield"].astype(np.complex128), s=shape, axes=(0, 1, 2))
    eng = _engine(lib, shape, spacing, power, smoothing, dtype)
    eng.set_noise(noise)
    eng.generate_delta(0, want_stats=True)
    delta = eng.copy_field()
    assert delta.shape == shape and delta.dtype == dtype
    tol = TOL_F32 if dtype == np.float32 else TOL_F64
    assert _relrms(delta, ref["delta"]) <= tol
    assert _relrms(delta, truth) <= tol
    s1, s2, cnt = eng.stats()
    std = np.sqrt(s2 / cnt - (s1 / cnt) ** 2)
    assert abs(std - ref["rms"]) <= 1e-5 * ref["rms"]


@pytest.mark.parametrize("name", ["ref_64_f32", "ref_64_f32_smooth", "ref_128x64x256_f32",
                                  "ref_16x8x12_f64", "ref_12x10x8_f32"])
def test_injected_noise_matches_golden(lib, golden_dir, name):
    """Same gate against the committed fixtures made from the unmodified reference."""
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    shape = tuple(int(v) for v in g["shape"])
    cdtype = np.dtype(str(g["dtype"]))
    dtype = np.float32 if cdtype == np.complex64 else np.float64
    nx, ny, nz = shape
    noise = np.random.RandomState(int(g["seed"])).normal(size=2 * nx * ny * (nz // 2 + 1))
    eng = _engine(lib, shape, float(g["spacing"]), oracle.load_default_power(), float(g["smoothing"]), dtype)
    eng.set_noise(noise)
    eng.generate_delta(0)
    sd = g["delta_stride"]
    delta = eng.copy_field()[::sd[0], ::sd[1], ::sd[2]]
    tol = TOL_F32 if dtype == np.float32 else 5e-12
    assert np.sqrt(np.mean((delta.astype(np.float64) - g["delta"]) ** 2)) <= tol * float(g["rms"])


# ------------------------------------------------------------------------------------------
# Philox path vs its float64 restatement
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,dtype", [((64, 64, 64), np.float32), ((64, 128, 256), np.float32),
                                         ((64, 64, 2048), np.float32), ((64, 64, 4096), np.float32),
                                         ((2048, 64, 64), np.float32), ((64, 2048, 64), np.float32),
                                         ((4096, 64, 64), np.float32), ((64, 4096, 128), np.float32),
                                         ((64, 64, 64), np.float64), ((12, 10, 8), np.float32)])
def test_philox_field_matches_oracle(lib, shape, dtype):
    spacing, seed = 2.0, 2024
    power = oracle.load_default_power()
    kf64, _ = oracle.philox_kfield(shape, spacing, power, seed, smoothing=1.0)
    truth = np.fft.irfftn(kf64, s=shape, axes=(0, 1, 2))
    eng = _engine(lib, shape, spacing, power, 1.0, dtype)
    kf = eng.kspace_public(seed)
    scale = np.sqrt(np.mean(np.abs(kf64) ** 2))
    assert np.sqrt(np.mean(np.abs(kf - kf64) ** 2)) <= 2e-6 * scale
    eng.generate_delta(seed)
    delta = eng.copy_field()
    # f64 mode still takes its bins/k^2 in float64 while the oracle restates the f32 ones:
    # identical except at knot edges, where sigma is continuous
    assert _relrms(delta, truth) <= (TOL_F32 if dtype == np.float32 else 1e-6)


def test_potential_regeneration(lib):
    """delta(k)/k^2 (generate.py:200-215) regenerated from the seed equals the reference
    construction applied to our own k-field."""
    shape, spacing, seed = (64, 64, 64), 2.0, 11
    power = oracle.load_default_power()
    eng = _engine(lib, shape, spacing, power)
    kf = eng.kspace_public(seed)
    pot = eng.kspace_public(seed, potential=True)
    want = oracle.potential_from_kfield(kf, spacing)
    scale = np.sqrt(np.mean(np.abs(want) ** 2))
    assert np.sqrt(np.mean(np.abs(pot - want) ** 2)) <= 1e-6 * scale
    assert pot[0, 0, 0] == 0


# ------------------------------------------------------------------------------------------
# gate 4: statistics of the Philox field
# ------------------------------------------------------------------------------------------
def test_gaussian_variance_kat_gpu(lib):
    """The reference's known-answer test (tests/test_generate.py:24-62) on the GPU Philox path:
    Gaussian P(k) = P0 exp(-(k sigma)^2/2); mean variance within 1 % of the analytic value.  The
    reference averages its 10 MT19937 seeds (scatter of the mean ~1 %, it passes at 0.25 %); the
    Philox stream is a different sample, so 200 seeds are averaged (scatter ~0.2 %) to make the
    same 1 % gate a 4-sigma statement."""
    from scipy.special import erf
    spacing, n, seed = 2.5, 64, 123
    kmin, kmax = (2 * np.pi) / (spacing * n), np.pi / spacing
    sigma, P0 = 2.5 * spacing, 1.23
    calc = P0 / (2 * np.pi) ** 1.5 / sigma ** 3 * (
        erf(kmax * sigma / np.sqrt(2)) ** 3 - erf(kmin * sigma / np.sqrt(2)) ** 3)
    power = np.empty(100, dtype=[('k', float), ('Pk', float)])
    power['k'] = np.linspace(kmin, np.sqrt(3) * kmax, len(power))
    power['Pk'] = P0 * np.exp(-0.5 * (power['k'] * sigma) ** 2)
    eng = _engine(lib, (n, n, n), spacing, power)
    var = []
    for t in range(200):
        eng.generate_delta(seed + t, want_stats=True)
        s1, s2, cnt = eng.stats()
        var.append(s2 / cnt - (s1 / cnt) ** 2)
        if t < 3:
            assert abs(var[-1] - np.var(eng.copy_field().astype(np.float64))) < 1e-5 * var[-1]
    assert abs(np.mean(var) - calc) < 0.01 * calc


def test_power_spectrum_recovered(lib):
    """Measured P(k) of the generated field recovers the input spectrum within cosmic
    variance: |delta_k|^2 V / N^6 averaged over shells of N_modes modes scatters around P(k)
    with relative std 1/sqrt(N_modes)."""
    shape, spacing, seed = (128, 128, 128), 2.0, 2025
    power = oracle.load_default_power()
    eng = _engine(lib, shape, spacing, power)
    eng.generate_delta(seed)
    delta = eng.copy_field().astype(np.float64)
    nx, ny, nz = shape
    N3 = nx * ny * nz
    V = N3 * spacing ** 3
    dk = np.fft.rfftn(delta)
    kx2, ky2, kz2 = oracle.create_ksq_grids(shape, spacing, packed=True)
    k = np.sqrt(kx2 + ky2 + kz2)
    pk_in = np.interp(np.log10(np.where(k > 0, k, 1.0)), np.log10(power['k']), power['Pk'])
    ratio = (np.abs(dk) ** 2 * V / N3 ** 2) / pk_in
    # interior planes only: modes there are independent complex Gaussians (no Hermitian pairs)
    sel = np.zeros(k.shape, dtype=bool)
    sel[:, :, 1:nz // 2] = True
    edges = np.linspace(k[sel].min(), k[sel].max(), 25)
    which = np.digitize(k[sel], edges)
    r = ratio[sel]
    worst = 0.0
    for b in range(1, len(edges)):
        m = which == b
        nm = int(m.sum())
        if nm < 50:
            continue
        dev = abs(r[m].mean() - 1.0) * np.sqrt(nm)
        worst = max(worst, dev)
    assert worst < 5.0, worst
    assert abs(r.mean() - 1.0) < 5.0 / np.sqrt(r.size)


# ------------------------------------------------------------------------------------------
# lensing potential scan vs the reference loop (oracle, unpinned)
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape,dtype,i_min", [((16, 12, 64), np.float32, None), ((8, 8, 256), np.float32, 3),
                                               ((8, 8, 256), np.float64, 0), ((12, 10, 8), np.float32, 1),
                                               ((4, 6, 1024), np.float32, 32), ((4, 4, 4096), np.float32, None),
                                               ((6, 4, 128), np.float32, 5), ((4, 6, 384), np.float32, 2),
                                               ((4, 4, 2048), np.float32, None), ((6, 4, 512), np.float64, 7),
                                               ((40, 34, 256), np.float32, 0), ((2, 2, 4096), np.float64, 100)])
def test_lensing_potential_matches_reference_loop(lib, shape, dtype, i_min):
    import randomfield_b200 as rf
    gen = rf.Generator(*shape, 2.0, dtype=dtype)
    gen.generate_delta_field(seed=11, save_potential=True)
    dphi = gen.calculate_newtonian_potential(light_cone=True).copy()
    psi = gen.calculate_lensing_potential(i_min=i_min).copy()
    ref = oracle.lensing_potential(dphi, gen.DC, gen.DA, Ok0=gen.cosmology.Ok0, i_min=i_min)
    assert psi.shape == shape and psi.dtype == dtype
    im = shape[2] // 32 if i_min is None else i_min
    assert np.all(psi[:, :, :im] == 0)
    scale = np.abs(ref.astype(np.float64)).max()
    err = np.abs(psi.astype(np.float64) - ref.astype(np.float64)).max() / scale
    assert err <= (2e-6 if dtype == np.float32 else 1e-12), err
    with pytest.raises(ValueError):
        gen.calculate_lensing_potential(i_min=shape[2])
