"""The reference's own unit tests re-hosted on the CUDA backend (run with -m gpu).

Every module-level function of randomfield_b200.{transform,powertools,random,cosmotools} -- i.e.
every rf_stage_* / rf_plan_execute export of include/rf_b200.h -- is executed here against the
reference's assertions (randomfield/tests/test_transform.py:150-298, test_powertools.py:15-128,
test_random.py:12-31, test_cosmotools.py:86-99) and against the oracle / the golden fixtures
generated from the unmodified reference (tests/golden/make_golden.py).
"""
import os
from itertools import product

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu

shape = (4, 6, 8)
seed = 123
TF = (True, False)
complex_types = (np.complex64, np.complex128)
float_types = (np.float32, np.float64)
spacing = 2.5


@pytest.fixture(scope="module")
def rf():
    import randomfield_b200
    from randomfield_b200 import _lib
    _lib.lib()
    return randomfield_b200


def _fill_normal(plan, T):
    from randomfield_b200.transform import scalar_type
    real_size = 2 * plan.data_in.size
    plan.data_in.view(scalar_type(T)).reshape(real_size)[:] = np.random.normal(size=real_size)


# ------------------------------------------------------------------------------------------
# transform.py
# ------------------------------------------------------------------------------------------
def test_is_hermitian_of_r2c_output(rf):
    """tests/test_transform.py:150-160: the forward packed / unpacked transform of real data."""
    from randomfield_b200.transform import Plan, is_hermitian, complex_type
    np.random.seed(seed)
    for overwrite, packed, ftype in product(TF, TF, float_types):
        plan = Plan(shape=shape, dtype_in=ftype if packed else complex_type(ftype), inverse=False,
                    overwrite=overwrite, packed=packed)
        plan.data_in[:] = np.random.normal(size=plan.data_in.shape)
        x = plan.data_in.copy()          # (overwrite=True: the output shares the input's memory)
        result = plan.execute()
        assert is_hermitian(result, packed=packed)
        # and it is the right transform: numpy is the reference's own backend (transform.py:266-270)
        want = np.fft.rfftn(x.astype(np.float64)) if packed else np.fft.fftn(x.astype(np.complex128))
        assert np.allclose(result, want, atol=1e-5 if ftype == np.float32 else 1e-12)


def test_symmetrize(rf):
    """tests/test_transform.py:163-177."""
    from randomfield_b200.transform import Plan, symmetrize, is_hermitian
    np.random.seed(seed)
    for overwrite, packed, ctype in product(TF, TF, complex_types):
        plan = Plan(shape=shape, dtype_in=ctype, inverse=True, overwrite=overwrite, packed=packed)
        _fill_normal(plan, ctype)
        symmetrize(plan.data_in, packed=packed)
        assert is_hermitian(plan.data_in, packed=packed)
        result = plan.execute()
        assert np.allclose(result.imag, 0)


def test_symmetrize_matches_reference(rf, golden_dir):
    """Packed and unpacked results of the unmodified reference's symmetrize (golden fixture)."""
    from randomfield_b200.transform import symmetrize
    g = np.load(os.path.join(golden_dir, "ref_symmetrize.npz"))
    for key in ("packed", "full"):
        for ctype in ("complex64", "complex128"):
            name = key + "_" + ctype
            data = g[name + "_in"].copy()
            symmetrize(data, packed=(key == "packed"))
            assert np.array_equal(data, g[name + "_out"]), name
            want = oracle.symmetrize(g[name + "_in"].copy(), packed=(key == "packed"))
            assert np.array_equal(data, want)


def test_c2c_round_trip(rf):
    """tests/test_transform.py:180-200."""
    from randomfield_b200.transform import Plan
    np.random.seed(seed)
    for inverse_first, overwrite_f, overwrite_r, dtype in product(TF, TF, TF, complex_types):
        plan_f = Plan(shape=shape, dtype_in=dtype, inverse=inverse_first, overwrite=overwrite_f, packed=False)
        plan_r = Plan(shape=shape, dtype_in=dtype, inverse=not inverse_first, overwrite=overwrite_r, packed=False)
        _fill_normal(plan_f, dtype)
        original = np.copy(plan_f.data_in)
        want = (np.fft.ifftn if inverse_first else np.fft.fftn)(original.astype(np.complex128))
        plan_r.data_in[:] = plan_f.execute()
        assert np.allclose(plan_r.data_in, want, atol=2e-5 if dtype == np.complex64 else 1e-12)
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


def test_c2c_reverse(rf):
    """tests/test_transform.py:203-220."""
    from randomfield_b200.transform import Plan
    np.random.seed(seed)
    for inverse_first, overwrite_f, overwrite_r, reuse, dtype in product(TF, TF, TF, TF, complex_types):
        plan_f = Plan(shape=shape, dtype_in=dtype, inverse=inverse_first, overwrite=overwrite_f, packed=False)
        plan_r = plan_f.create_reverse_plan(reuse_output=reuse, overwrite=overwrite_r)
        _fill_normal(plan_f, dtype)
        original = np.copy(plan_f.data_in)
        plan_f.execute()
        if not reuse:
            plan_r.data_in[:] = plan_f.data_out
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


def test_c2r_round_trip(rf):
    """tests/test_transform.py:223-244."""
    from randomfield_b200.transform import Plan, symmetrize, scalar_type
    np.random.seed(seed)
    for overwrite_f, overwrite_r, dtype in product(TF, TF, complex_types):
        plan_f = Plan(shape=shape, dtype_in=dtype, inverse=True, packed=True, overwrite=overwrite_f)
        plan_r = Plan(shape=shape, dtype_in=scalar_type(dtype), inverse=False, packed=True, overwrite=overwrite_r)
        _fill_normal(plan_f, dtype)
        symmetrize(plan_f.data_in, packed=True)
        original = np.copy(plan_f.data_in)
        want = np.fft.irfftn(original.astype(np.complex128), s=shape, axes=(0, 1, 2))
        plan_r.data_in[:] = plan_f.execute()
        assert np.allclose(plan_r.data_in, want, atol=1e-6 if dtype == np.complex64 else 1e-13)
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


def test_c2r_reverse(rf):
    """tests/test_transform.py:247-268."""
    from randomfield_b200.transform import Plan, symmetrize
    np.random.seed(seed)
    for overwrite_f, overwrite_r, reuse, dtype in product(TF, TF, TF, complex_types):
        if reuse and not overwrite_f and overwrite_r:
            continue
        plan_f = Plan(shape=shape, dtype_in=dtype, inverse=True, packed=True, overwrite=overwrite_f)
        plan_r = plan_f.create_reverse_plan(reuse_output=reuse, overwrite=overwrite_r)
        _fill_normal(plan_f, dtype)
        symmetrize(plan_f.data_in, packed=True)
        original = np.copy(plan_f.data_in)
        plan_f.execute()
        if not reuse:
            plan_r.data_in[:] = plan_f.data_out
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


def test_r2c_round_trip(rf):
    """tests/test_transform.py:271-284."""
    from randomfield_b200.transform import Plan, scalar_type
    np.random.seed(seed)
    for overwrite_f, overwrite_r, dtype in product(TF, TF, complex_types):
        plan_f = Plan(shape=shape, dtype_in=scalar_type(dtype), inverse=False, packed=True, overwrite=overwrite_f)
        plan_r = Plan(shape=shape, dtype_in=dtype, inverse=True, packed=True, overwrite=overwrite_r)
        plan_f.data_in[:] = np.random.normal(size=shape)
        original = np.copy(plan_f.data_in)
        plan_r.data_in[:] = plan_f.execute()
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


def test_r2c_reverse(rf):
    """tests/test_transform.py:287-298."""
    from randomfield_b200.transform import Plan, scalar_type
    np.random.seed(seed)
    for overwrite_f, overwrite_r, reuse, dtype in product(TF, TF, TF, complex_types):
        plan_f = Plan(shape=shape, dtype_in=scalar_type(dtype), inverse=False, packed=True, overwrite=overwrite_f)
        plan_r = plan_f.create_reverse_plan(reuse_output=reuse, overwrite=overwrite_r)
        plan_f.data_in[:] = np.random.normal(size=shape)
        original = np.copy(plan_f.data_in)
        plan_f.execute()
        if not reuse:
            plan_r.data_in[:] = plan_f.data_out
        result = plan_r.execute()
        assert np.allclose(original, result, atol=1e-6)


@pytest.mark.parametrize("big", [(64, 32, 128), (30, 20, 12)])
def test_plan_execute_matches_numpy_on_larger_grids(rf, big):
    """tests/test_transform.py:301-328 ("methods agree") with numpy as the second backend."""
    from randomfield_b200.transform import Plan, symmetrize
    np.random.seed(seed)
    for dtype in complex_types:
        plan = Plan(shape=big, dtype_in=dtype, inverse=True, packed=True, overwrite=False)
        _fill_normal(plan, dtype)
        symmetrize(plan.data_in, packed=True)
        want = np.fft.irfftn(plan.data_in.astype(np.complex128), s=big, axes=(0, 1, 2))
        got = plan.execute()
        tol = 1e-6 if dtype == np.complex64 else 1e-13
        assert np.sqrt(np.mean((got - want) ** 2)) <= tol * want.std()


# ------------------------------------------------------------------------------------------
# powertools.py
# ------------------------------------------------------------------------------------------
def test_fill(rf):
    """tests/test_powertools.py:15-35, vectorised; plus bit equality with the reference's rounding
    sequence (powertools.py:46-60) up to the last bit of log10."""
    from randomfield_b200.powertools import fill_with_log10k
    from randomfield_b200.transform import Plan
    nx, ny, nz = shape
    k0 = [2 * np.pi / (spacing * n) for n in shape]
    for packed, ctype in product(TF, complex_types):
        plan = Plan(shape=shape, dtype_in=ctype, packed=packed)
        fill_with_log10k(plan.data_in, spacing=spacing, packed=packed)
        j = [np.where(np.arange(n) <= n // 2, np.arange(n), np.arange(n) - n) for n in shape]
        jz = j[2][:nz // 2 + 1] if packed else j[2]
        ksq = ((j[0] * k0[0]) ** 2)[:, None, None] + ((j[1] * k0[1]) ** 2)[None, :, None] + \
              ((jz * k0[2]) ** 2)[None, None, :]
        assert np.isinf(plan.data_in[0, 0, 0].real) and plan.data_in[0, 0, 0].real < 0
        ksq[0, 0, 0] = 1.0
        got = plan.data_in.real.copy()
        got[0, 0, 0] = 0.0
        assert np.all(np.abs(got - np.log10(np.sqrt(ksq))) < 1e-6)
        assert np.all(plan.data_in.imag == 0)
        ref = oracle.fill_with_log10k(np.empty_like(plan.data_in), spacing, packed=packed)
        ulp = np.spacing(np.abs(ref.real[1:, 1:, 1:]).astype(ref.real.dtype))
        assert np.all(np.abs(plan.data_in.real[1:, 1:, 1:] - ref.real[1:, 1:, 1:]) <= 2 * ulp)


def test_bounds(rf):
    """tests/test_powertools.py:38-50."""
    from randomfield_b200.powertools import fill_with_log10k, get_k_bounds
    from randomfield_b200.transform import Plan
    plan = Plan(shape=shape, dtype_in=np.complex64)
    fill_with_log10k(plan.data_in, spacing=spacing)
    kmax1 = 10 ** np.max(plan.data_in.real)
    assert np.isinf(plan.data_in[0, 0, 0].real)
    plan.data_in[0, 0, 0] = np.log10(kmax1)
    kmin1 = 10 ** np.min(plan.data_in.real)
    kmin2, kmax2 = get_k_bounds(plan.data_in, spacing=spacing)
    assert abs((kmin1 - kmin2) / kmin1) < 1e-7
    assert abs((kmax1 - kmax2) / kmax1) < 1e-7


def test_tabulate_sigmas(rf):
    """tests/test_powertools.py:99-128 (equation 60 of astro-ph/0506540), vectorised; plus equality
    with the reference's np.interp evaluation (powertools.py:155-163) to float rounding."""
    import scipy.interpolate
    from randomfield_b200.powertools import fill_with_log10k, tabulate_sigmas, load_default_power
    from randomfield_b200.transform import Plan
    nx, ny, nz = shape
    N3 = nx * ny * nz
    Vbox = N3 * spacing ** 3
    power = load_default_power()
    Pk = scipy.interpolate.interp1d(power['k'], power['Pk'])
    k0 = [2 * np.pi / (spacing * n) for n in shape]
    for packed, ctype in product(TF, complex_types):
        plan = Plan(shape=shape, dtype_in=ctype, packed=packed)
        fill_with_log10k(plan.data_in, spacing=spacing, packed=packed)
        tabulate_sigmas(plan.data_in, power, spacing, packed=packed)
        assert plan.data_in[0, 0, 0] == 0
        j = [np.where(np.arange(n) <= n // 2, np.arange(n), np.arange(n) - n) for n in shape]
        jz = j[2][:nz // 2 + 1] if packed else j[2]
        ksq = ((j[0] * k0[0]) ** 2)[:, None, None] + ((j[1] * k0[1]) ** 2)[None, :, None] + \
              ((jz * k0[2]) ** 2)[None, None, :]
        ksq[0, 0, 0] = k0[0] ** 2
        sigma = N3 * np.sqrt(Pk(np.sqrt(ksq)) / (2 * Vbox))
        sigma[0, 0, 0] = 0
        assert np.all(np.abs(plan.data_in.real - sigma) <= 1e-3 * sigma)
        ref = np.empty_like(plan.data_in)
        oracle.fill_with_log10k(ref, spacing, packed=packed)
        oracle.tabulate_sigmas(ref, power, spacing, packed=packed)
        tol = 3e-6 if ctype == np.complex64 else 1e-12     # f32: log10 differs in the last bit
        assert np.all(np.abs(plan.data_in.real - ref.real) <= tol * np.abs(ref.real))
    with pytest.raises(ValueError):      # table does not cover the grid's k range (powertools.py:147-150)
        small = power[100:200]
        tabulate_sigmas(np.zeros((4, 6, 5), np.complex64), small, spacing)


# ------------------------------------------------------------------------------------------
# random.py
# ------------------------------------------------------------------------------------------
def test_randomize(rf):
    """tests/test_random.py:12-22."""
    from randomfield_b200.random import randomize
    nx, ny, nz = 40, 60, 80
    sigma = 1 + np.arange(nx * ny * nz).reshape(nx, ny, nz)
    for ctype in complex_types:
        data = np.empty((nx, ny, nz), dtype=ctype)
        data.real = sigma
        randomize(data, seed)
        data /= sigma
        assert abs(np.mean(data.real)) < 5e-3
        assert abs(np.mean(data.imag)) < 5e-3
        assert abs(np.std(data.real) - 1) < 5e-3
        assert abs(np.std(data.imag) - 1) < 5e-3
        assert abs(np.mean(data.real * data.imag)) < 5e-3      # re and im are independent


def test_reproducible(rf):
    """tests/test_random.py:25-31."""
    from randomfield_b200.random import randomize
    nx, ny, nz = 4, 6, 8
    data1 = np.ones((nx, ny, nz), dtype=np.complex64)
    randomize(data1, seed)
    data2 = np.ones((nx, ny, nz), dtype=np.complex64)
    randomize(data2, seed)
    assert np.array_equal(data1, data2)
    data3 = np.ones((nx, ny, nz), dtype=np.complex64)
    randomize(data3, seed + 1)
    assert not np.array_equal(data1, data3)


def test_deviate_tails(rf):
    """The Philox + Box-Muller deviates are standard normal far into the tails (the radius
    uses all 32 bits of its word, so |n| reaches 6.76 instead of 5.77): counts beyond 3, 4 and 5 sigma
    within 5 Poisson sigma of the Gaussian expectation, a Kolmogorov-Smirnov distance at the
    sampling level, and the float32 deviates against the float64 restatement of the same words."""
    from scipy.special import erfc, erf
    from randomfield_b200.random import randomize
    n = 1 << 24
    data = np.ones((64, 512, 512), dtype=np.complex64)
    randomize(data, 2027)
    x = np.concatenate([data.real.ravel(), data.imag.ravel()]).astype(np.float64)
    assert x.size == 2 * n
    for t in (3.0, 4.0, 5.0):
        expect = x.size * erfc(t / np.sqrt(2))
        got = np.count_nonzero(np.abs(x) > t)
        assert abs(got - expect) <= 5 * np.sqrt(expect) + 1, (t, got, expect)
    assert np.abs(x).max() < 6.8
    xs = np.sort(x[::16])
    cdf = 0.5 * (1 + erf(xs / np.sqrt(2)))
    ks = np.abs(cdf - (np.arange(xs.size) + 0.5) / xs.size).max()
    assert ks < 2.0 / np.sqrt(xs.size)
    # float64 mode draws the same words: the two precisions agree to float32 rounding
    d64 = np.ones((8, 64, 64), dtype=np.complex128)
    d32 = np.ones((8, 64, 64), dtype=np.complex64)
    randomize(d64, 5)
    randomize(d32, 5)
    assert np.abs(d32 - d64).max() < 5e-6


def test_stage_pipeline_equals_generator(rf):
    """The four stage functions chained as generate.py:191-199 chains them, then Plan.execute, give
    a real field with the reference's statistics (the deviates differ from the fused kernel's
    mapping, the spectrum does not): variance within 5 sigma of the fused Generator's."""
    from randomfield_b200 import powertools, random, transform
    n = 32
    plan = transform.Plan(shape=(n, n, n), dtype_in=np.complex64, packed=True, overwrite=True, inverse=True)
    power = powertools.load_default_power()
    var = []
    for s in range(4):
        powertools.fill_with_log10k(plan.data_in, spacing=2.0, packed=True)
        powertools.tabulate_sigmas(plan.data_in, power, 2.0, packed=True)
        random.randomize(plan.data_in, seed=seed + s)
        transform.symmetrize(plan.data_in, packed=True)
        assert transform.is_hermitian(plan.data_in, packed=True)
        delta = plan.execute()
        assert delta.shape == (n, n, n) and delta.dtype == np.float32
        var.append(np.var(delta.astype(np.float64)))
    gen = rf.Generator(n, n, n, 2.0)
    var2 = [np.var(gen.generate_delta_field(seed=seed + s).astype(np.float64)) for s in range(16)]
    assert abs(np.mean(var) - np.mean(var2)) < 5 * np.std(var2) / np.sqrt(4)


# ------------------------------------------------------------------------------------------
# cosmotools.apply_lognormal_transform / convert_delta_to_density
# ------------------------------------------------------------------------------------------
def test_lognormal(rf):
    """tests/test_cosmotools.py:86-99 on rf_stage_density."""
    from randomfield_b200.cosmotools import apply_lognormal_transform
    growth = 0.3
    sigma = 2.5
    np.random.seed(123)
    delta = np.empty((64, 64, 128), dtype=np.float32)
    delta[:] = sigma * np.random.normal(size=delta.shape)
    assert abs(np.mean(delta)) < 1e-2
    assert abs(np.std(delta) - sigma) < 1e-2 * sigma
    want = oracle.apply_lognormal_transform(delta.copy(), growth, sigma=2.5)
    rho = apply_lognormal_transform(delta, growth, sigma=2.5)
    assert rho.shape == delta.shape
    assert rho is delta or rho.base is delta.base or rho.base is delta
    assert np.all(rho > 0)
    assert abs(np.mean(rho) - 1.) < 1e-3
    assert abs(np.std(rho) - growth * sigma) < 1e-2 * sigma
    assert np.all(np.abs(rho - want) <= 3e-6 * want)


@pytest.mark.parametrize("dtype", float_types)
def test_lognormal_default_sigma_and_axis_tables(rf, dtype):
    """sigma=None means np.std(delta) (cosmotools.py:212-213), computed on the device; growth may be
    a per-z table broadcast on the last axis (generate.py:268-269)."""
    from randomfield_b200.cosmotools import apply_lognormal_transform
    rng = np.random.RandomState(5)
    delta = (1.7 * rng.normal(size=(12, 10, 14))).astype(dtype)
    growth = np.linspace(1.0, 0.5, 14)
    want = oracle.apply_lognormal_transform(delta.astype(np.float64), growth, sigma=None)
    rho = apply_lognormal_transform(delta, growth)
    assert np.all(np.abs(rho - want) <= (2e-5 if dtype == np.float32 else 1e-12) * want)


@pytest.mark.parametrize("shape3,dtype,lognormal", [((16, 12, 64), np.float32, True), ((16, 12, 64), np.float32, False),
                                                     ((12, 10, 6), np.float32, True), ((12, 10, 6), np.float64, True),
                                                     ((12, 10, 6), np.float64, False), ((64, 64, 128), np.float64, True),
                                                     ((128, 64, 256), np.float32, True)])
def test_density_matches_oracle_on_same_delta(rf, shape3, dtype, lognormal):
    """Generator.convert_delta_to_density (rf_density: density_kernel_v4 when nz % 4 == 0, else
    density_kernel) against generate.py:266-273 applied by the oracle to the SAME delta field."""
    gen = rf.Generator(*shape3, 2.0, dtype=dtype, num_plot_sections=2)
    delta = gen.generate_delta_field(seed=31).copy()
    sigma = gen.delta_field_rms
    assert abs(float(sigma) - delta.std(dtype=np.float64)) <= 1e-5 * float(sigma)
    want = oracle.convert_delta_to_density(delta.astype(np.float64), gen.growth_function, gen.mean_matter_density,
                                           float(sigma), lognormal=lognormal)
    rho = gen.convert_delta_to_density(apply_lognormal_transform=lognormal)
    assert rho.shape == shape3 and rho.dtype == dtype
    tol = 5e-6 if dtype == np.float32 else 1e-12
    assert np.all(np.abs(rho - want) <= tol * np.abs(want) + (1e-36 if not lognormal else 0))
    if lognormal:
        assert np.all(rho > 0)


# ------------------------------------------------------------------------------------------
# calculate_newtonian_potential
# ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("shape3,dtype", [((64, 64, 64), np.float32), ((64, 64, 64), np.float64),
                                          ((128, 64, 256), np.float32), ((12, 10, 8), np.float32),
                                          ((256, 256, 128), np.float32)])
@pytest.mark.parametrize("light_cone", [True, False])
def test_newtonian_potential_matches_oracle(rf, shape3, dtype, light_cone):
    """Phi(x) against generate.py:331-343 evaluated by the oracle in float64 on the float64
    restatement of the same Philox k-field: relative to max|Phi|; the 4.8e-36 scale must not flush
    anything to zero (SURVEY section 7 'dynamic range')."""
    gen = rf.Generator(*shape3, 2.0, dtype=dtype)
    gen.generate_delta_field(seed=17, smoothing_length_Mpc_h=1.0, save_potential=True)
    phi = gen.calculate_newtonian_potential(light_cone=light_cone).copy()
    kf, _ = oracle.philox_kfield(shape3, 2.0, oracle.load_default_power(), 17, smoothing=1.0)
    pot = oracle.potential_from_kfield(kf, 2.0)
    want = oracle.newtonian_potential(pot, shape3, gen.cosmology.Om0,
                                      growth=gen.growth_function if light_cone else None,
                                      redshifts=gen.redshifts if light_cone else None)
    assert phi.shape == shape3 and phi.dtype == dtype
    scale = np.abs(want).max()
    assert 1e-36 < scale < 1e-28            # s^-2: tiny but normal numbers in float32
    err = np.sqrt(np.mean((phi.astype(np.float64) - want) ** 2)) / want.std()
    assert err <= (1e-5 if dtype == np.float32 else 1e-6), err
    assert np.abs(phi.astype(np.float64) - want).max() <= (2e-5 if dtype == np.float32 else 1e-5) * scale
    assert np.count_nonzero((phi == 0) & (np.abs(want) > 1e-5 * scale)) == 0     # nothing flushed to zero
    # the reference keeps the saved potential (generate.py:333): a second call gives the same field
    again = gen.calculate_newtonian_potential(light_cone=light_cone)
    assert np.array_equal(again, phi)


def test_newtonian_potential_needs_saved_potential(rf):
    gen = rf.Generator(64, 64, 64, 2.0)
    gen.generate_delta_field(seed=1, save_potential=False)
    with pytest.raises(RuntimeError):
        gen.calculate_newtonian_potential()


def test_plot_requests_fail_before_any_work(rf):
    """plot_slice is outside the path: show_plot / save_plot_name raise before a field is generated."""
    gen = rf.Generator(64, 64, 64, 2.0)
    with pytest.raises(NotImplementedError):
        gen.generate_delta_field(seed=1, show_plot=True)
    assert gen._last is None
    gen.generate_delta_field(seed=1)
    with pytest.raises(NotImplementedError):
        gen.convert_delta_to_density(save_plot_name='x.png')
    assert gen._field_kind == 'delta'


def test_potential_attribute_has_no_side_effect(rf):
    """reading generator.potential / plan_c2r.data_in must not destroy the delta field."""
    gen = rf.Generator(64, 64, 64, 2.0)
    delta = gen.generate_delta_field(seed=3).copy()
    pot = gen.potential
    assert pot.shape == (64, 64, 33) and pot.dtype == np.complex64
    kf = gen.plan_c2r.data_in
    assert np.allclose(pot[1:, 1:, 1:] * 0 + 1, 1)
    want = oracle.potential_from_kfield(kf, 2.0)
    assert np.sqrt(np.mean(np.abs(pot - want) ** 2)) <= 1e-6 * np.sqrt(np.mean(np.abs(want) ** 2))
    assert np.array_equal(gen.device_field().cpu().numpy(), delta)
    rho = gen.convert_delta_to_density()
    assert np.all(rho > 0)


def test_lensing_leaves_the_potential_intact(rf):
    """the reference allocates a separate psi (generate.py:398-399) and leaves dPhi alone;
    a second call with another i_min integrates dPhi again, not psi."""
    gen = rf.Generator(16, 12, 64, 2.0)
    gen.generate_delta_field(seed=11, save_potential=True)
    dphi = gen.calculate_newtonian_potential(light_cone=True)
    keep = dphi.copy()
    psi1 = gen.calculate_lensing_potential(i_min=2)
    assert np.array_equal(dphi, keep)                # the caller's dPhi array did not turn into psi
    assert psi1 is not dphi
    p1 = psi1.copy()
    psi2 = gen.calculate_lensing_potential(i_min=5).copy()
    ref2 = oracle.lensing_potential(keep, gen.DC, gen.DA, Ok0=gen.cosmology.Ok0, i_min=5)
    scale = np.abs(ref2.astype(np.float64)).max()
    assert np.abs(psi2.astype(np.float64) - ref2).max() <= 2e-6 * scale
    ref1 = oracle.lensing_potential(keep, gen.DC, gen.DA, Ok0=gen.cosmology.Ok0, i_min=2)
    assert np.abs(p1.astype(np.float64) - ref1).max() <= 2e-6 * np.abs(ref1.astype(np.float64)).max()
    gen2 = rf.Generator(16, 12, 64, 2.0)
    gen2.generate_delta_field(seed=11)
    with pytest.raises(RuntimeError):
        gen2.calculate_lensing_potential()            # a delta field is not a potential
