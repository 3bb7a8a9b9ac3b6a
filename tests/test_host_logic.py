"""CPU tests of the host-side mirror of the reference interface (table-sized logic that stays
in Python) and of the slab-decomposition bookkeeping, incl. a world_size-2 gloo exchange."""
import os

import numpy as np
import pytest

import oracle
from randomfield_b200 import powertools, slab, transform


# ---- reference tests/test_powertools.py:53-96 re-hosted ---------------------------------------
def _power(n=10):
    p = np.empty(n, dtype=[('k', float), ('Pk', float)])
    p['k'] = np.linspace(0.01, 10., n)
    p['Pk'] = 1.0
    return p


def test_validate_power_accepts_default():
    assert len(powertools.validate_power(powertools.load_default_power())) == 500


@pytest.mark.parametrize("mutate", [
    lambda p: p.tolist(),
    lambda p: p[['k']],
    lambda p: _set(p, 'k', 3, np.nan),
    lambda p: _set(p, 'Pk', 3, np.inf),
    lambda p: _set(p, 'k', 3, p['k'][2]),
    lambda p: _set(p, 'k', 0, 0.0),
    lambda p: _set(p, 'Pk', 4, -1.0),
])
def test_validate_power_rejects(mutate):
    with pytest.raises(ValueError):
        powertools.validate_power(mutate(_power()))


def _set(p, field, i, v):
    p = p.copy()
    p[field][i] = v
    return p


def test_filter_power_matches_oracle_and_rejects_negative():
    p = powertools.load_default_power()
    a, b = powertools.filter_power(p, 3.0), oracle.filter_power(oracle.load_default_power(), 3.0)
    assert np.array_equal(a['Pk'], b['Pk'])
    assert np.array_equal(powertools.filter_power(p, 0.0)['Pk'], p['Pk'])
    with pytest.raises(ValueError):
        powertools.filter_power(p, -1.0)


def test_k_bounds_and_sigma_table():
    data = np.empty((4, 6, 5), dtype=np.complex64)
    assert powertools.get_k_bounds(data, 2.5) == pytest.approx(oracle.get_k_bounds((4, 6, 8), 2.5))
    k, sig = powertools.sigma_table(powertools.load_default_power(), (64, 64, 64), 2.0)
    lk, osig = oracle.sigma_table(oracle.load_default_power(), (64, 64, 64), 2.0)
    assert np.array_equal(sig, osig) and np.allclose(np.log10(k), lk)
    with pytest.raises(ValueError):     # table does not cover the grid's k range: powertools.py:147-150
        powertools.sigma_table(_power(), (64, 64, 64), 0.01)


def test_transform_type_helpers():
    assert transform.scalar_type(np.complex64) == np.float32
    assert transform.complex_type(np.float64) == np.complex128
    with pytest.raises(ValueError):
        transform.scalar_type(np.float32)
    with pytest.raises(ValueError):
        transform.expanded_shape(np.empty((3, 4, 4)))
    with pytest.raises(ValueError):
        transform.expanded_shape(np.empty((4, 4, 4)), packed=True)
    assert transform.expanded_shape(np.empty((4, 6, 5)), packed=True) == (4, 6, 8)


# ---- slab decomposition ---------------------------------------------------------------------------
def test_slab_ranges():
    assert slab.slab_ranges(64, 4, 1) == (16, 32)
    with pytest.raises(ValueError):
        slab.slab_ranges(10, 4, 0)
    assert slab.block_elems(64, 32, 16, 4) == 16 * 8 * 8


def _slab_reference(shape, world):
    """numpy emulation of the slab pipeline in the internal packed-Nyquist layout: returns the
    per-rank x-slabs and the single-rank result they must equal bit for bit."""
    nx, ny, nz = shape
    rng = np.random.RandomState(4)
    kf = (rng.normal(size=(nx, ny, nz // 2)) + 1j * rng.normal(size=(nx, ny, nz // 2))).astype(np.complex64)
    full = np.fft.ifft(np.fft.ifft(kf, axis=0), axis=1)
    nyl, nxl = ny // world, nx // world
    sends = [np.ascontiguousarray(np.fft.ifft(kf[:, r * nyl:(r + 1) * nyl], axis=0)) for r in range(world)]
    return kf, full, sends, nxl, nyl


def test_recv_layout_single_process():
    shape, world = (8, 12, 8), 4
    _, full, sends, nxl, nyl = _slab_reference(shape, world)
    nzc = shape[2] // 2
    for d in range(world):
        recv = np.stack([s.reshape(world, nxl, nyl, nzc)[d] for s in sends])
        rows = slab.recv_to_rows(recv, *shape, world)
        assert np.array_equal(np.fft.ifft(rows, axis=1), full[d * nxl:(d + 1) * nxl])


def _gloo_worker(rank, world, port, shape, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        _, full, sends, nxl, nyl = _slab_reference(shape, world)
        send = torch.from_numpy(sends[rank].view(np.float32).copy())
        recv = torch.empty_like(send)
        slab.all_to_all(send, recv)
        rows = slab.recv_to_rows(recv.numpy().view(np.complex64).reshape(-1), *shape, world)
        ok = np.array_equal(np.fft.ifft(rows, axis=1), full[rank * nxl:(rank + 1) * nxl])
        # global statistics: the two-double all-reduce of SlabEngine.global_stats
        t = torch.tensor([float(rank + 1), float((rank + 1) ** 2), 1.0], dtype=torch.float64)
        dist.all_reduce(t)
        ok = ok and t.tolist() == [3.0, 5.0, 2.0]
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_all_to_all_gloo_world2():
    import torch.multiprocessing as mp
    world, port = 2, 29541 + os.getpid() % 200
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_gloo_worker, args=(world, port, (8, 12, 8), out), nprocs=world, join=True)
        assert dict(out) == {0: True, 1: True}


# ------------------------------------------------------------------------------------------
# lensing potential: oracle restatement and the O(nz) scan form
# ------------------------------------------------------------------------------------------
def test_simps_old_matches_simpson_for_odd_counts_and_avg_rule_for_even():
    import scipy.integrate
    rng = np.random.RandomState(3)
    x = np.cumsum(rng.uniform(0.5, 2.0, size=21))
    y = rng.normal(size=(3, 21))
    # odd sample count: plain composite Simpson on non-uniform abscissae (any scipy)
    assert np.allclose(oracle.simps_old(y, x), scipy.integrate.simpson(y, x=x), rtol=0, atol=1e-13)
    # even count: average of (Simpson on first N-1 + last trapezoid) and (first trapezoid + Simpson on last N-1)
    n = 10
    first = oracle.simps_old(y[..., :n - 1], x[:n - 1]) + 0.5 * (x[n - 1] - x[n - 2]) * (y[..., n - 1] + y[..., n - 2])
    last = oracle.simps_old(y[..., 1:n], x[1:n]) + 0.5 * (x[1] - x[0]) * (y[..., 0] + y[..., 1])
    assert np.allclose(oracle.simps_old(y[..., :n], x[:n]), 0.5 * (first + last), rtol=0, atol=1e-14)
    # exact for cubics on uniform abscissae with an odd count
    xu = np.linspace(0.0, 2.0, 9)
    assert abs(oracle.simps_old(xu ** 3, xu) - 4.0) < 1e-13


@pytest.mark.parametrize("nz,i_min,ok0", [(32, 1, 0.0), (32, 0, 0.0), (33, 2, 0.05), (64, None, -0.02), (10, 3, 0.0)])
def test_lensing_scan_form_equals_reference_loop(nz, i_min, ok0):
    """The parity-split prefix-sum form the device kernel uses (generate.lensing_scan_host) equals
    the reference's one-integral-per-slice loop (oracle.lensing_potential, generate.py:374-410)."""
    from randomfield_b200.generate import cot_k, lensing_scan_host, lensing_tables
    rng = np.random.RandomState(nz)
    DC = np.arange(nz) * 2.0 if nz != 33 else np.cumsum(rng.uniform(1.0, 3.0, size=nz))
    DA = DC * (1 + 0.01 * np.arange(nz))
    dphi = rng.normal(size=(4, 3, nz))
    ref = oracle.lensing_potential(dphi, DC, DA, Ok0=ok0, i_min=i_min)
    im = nz // 32 if i_min is None else i_min
    got = lensing_scan_host(dphi, lensing_tables(DC, cot_k(DC, DA, ok0)), im)
    assert np.abs(got - ref).max() <= 1e-13 * np.abs(ref).max()
    assert np.all(ref[..., :im] == 0)
    with pytest.raises(ValueError):
        oracle.lensing_potential(dphi, DC, DA, i_min=nz)


def test_split_plan_covers_the_slab():
    from randomfield_b200 import slab
    for nxl in (1, 2, 16, 512):
        for frac in (0.1, 0.45, 0.9, 1.0):
            for nsub in (1, 4, 1000):
                for align in (1, 2, 4):
                    xa, subs = slab.split_plan(nxl, 8, frac, nsub, align)
                    edges = [0, xa] + [b for _, b in subs]
                    assert edges[-1] == nxl and all(b > a for a, b in zip(edges, edges[1:]))
                    assert all(a == pb for (a, _), pb in zip(subs, edges[1:]))
                    assert all(e % align == 0 for e in edges) or not subs


def test_split_plan_defaults(monkeypatch):
    """Without RF_SPLIT_A / RF_SPLIT_NSUB: plain chunked schedule below 8 ranks, (0.7, 2) from 8 ranks
    (4096^3 on 8 GPUs: rows [0, 358) with pass 1, then two ranges); the environment overrides both."""
    from randomfield_b200 import slab
    monkeypatch.delenv("RF_SPLIT_A", raising=False)
    monkeypatch.delenv("RF_SPLIT_NSUB", raising=False)
    for world in (2, 4):
        assert slab.split_plan(2048 // world, world) == (2048 // world, [])
    xa, subs = slab.split_plan(512, 8, align=slab.pass3_row_align(4096, 4096))
    assert xa == 358 and subs == [(358, 435), (435, 512)]
    monkeypatch.setenv("RF_SPLIT_A", "1")
    assert slab.split_plan(512, 8) == (512, [])
    monkeypatch.setenv("RF_SPLIT_A", "0.5")
    monkeypatch.setenv("RF_SPLIT_NSUB", "4")
    xa, subs = slab.split_plan(512, 4)
    assert xa == 256 and len(subs) == 4 and subs[-1][1] == 512


def test_pass3_row_align():
    from randomfield_b200 import slab
    # 64 KB pass-3 tiles: nz = 64 -> 256 lines per tile, ny = 64 lines per x row -> 4 rows per tile
    assert slab.pass3_row_align(64, 64) == 4
    assert slab.pass3_row_align(4096, 4096) == 1 and slab.pass3_row_align(256, 128) == 1
