"""Slab decomposition on the GPU (gate 3 of BASELINE.json north_star): the Philox field must be
bit-identical at 1, 2, 4 and 8 ranks.  On a one-GPU box the ranks are driven from one process
(slab.run_virtual_slabs: same kernels, same buffers, device copies for the all-to-all); with
two or more GPUs the real torch.distributed/NCCL path is run under torchrun as well."""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("shape", [(64, 64, 64), (128, 256, 64), (256, 128, 128), (4096, 64, 64), (64, 2048, 64)])
def test_field_bit_identical_across_decompositions(shape):
    from randomfield_b200 import slab
    power = oracle.load_default_power()
    ref, mean1, std1 = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=1)
    assert np.isfinite(ref).all() and abs(std1 - ref.std(dtype=np.float64)) < 1e-5 * std1
    for world in (2, 4, 8):
        got, mean, std = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=world)
        assert np.array_equal(got, ref), "world=%d" % world
        assert abs(std - std1) <= 1e-9 * std1


@pytest.mark.parametrize("shape", [(64, 64, 64), (256, 128, 128), (2048, 128, 64), (128, 4096, 64)])
def test_direct_exchange_bit_identical(shape):
    """Pass 1 storing straight into the receive buffers of the owning ranks (peer memory; here
    all ranks live on one GPU) gives the same bits as the all-to-all path."""
    from randomfield_b200 import slab
    power = oracle.load_default_power()
    ref, _, _ = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=1)
    for world in (2, 4, 8):
        got, _, _ = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=world, direct=True)
        assert np.array_equal(got, ref), "world=%d" % world


@pytest.mark.parametrize("shape,chunks", [((256, 256, 128), 4), ((128, 256, 128), 2), ((2048, 128, 128), 2),
                                          ((128, 4096, 128), 2)])
def test_chunked_exchange_bit_identical(shape, chunks):
    """Pass 1 / copy-engine exchange / pass 2 pipelined over chunks of z columns (rf_chunk_*)
    gives the same bits as the unchunked single-rank run, for every decomposition."""
    from randomfield_b200 import slab
    power = oracle.load_default_power()
    ref, _, std1 = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=1)
    for world in (1, 2, 8):
        got, _, std = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=world, chunks=chunks)
        assert np.array_equal(got, ref), "world=%d" % world
        assert abs(std - std1) <= 1e-9 * std1


@pytest.mark.parametrize("shape,chunks,split", [((256, 256, 128), 4, (0.4, 3)), ((128, 256, 128), 2, (0.5, 64)),
                                                ((2048, 128, 128), 2, (0.3, 2)), ((128, 2048, 128), 2, (0.7, 1))])
def test_split_schedule_bit_identical(shape, chunks, split):
    """The exchange also cut along the destination's x axis (rf_chunk_exchange_x / pass2_x /
    pass3_x: rows [0, xa) travel with pass 1, the rest afterwards in x ranges): same bits, and
    the statistics summed over the ranges' pass-3 launches agree."""
    from randomfield_b200 import slab
    power = oracle.load_default_power()
    zs = np.linspace(0.5, 1.5, shape[2])
    ref, _, std1 = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=1, zscale=zs)
    for world in (1, 2, 4, 8):
        got, _, std = slab.run_virtual_slabs(*shape, 2.0, power, seed=99, world=world, chunks=chunks,
                                             split=split, zscale=zs)
        assert np.array_equal(got, ref), "world=%d" % world
        assert abs(std - std1) <= 1e-9 * std1


def test_slab_field_matches_oracle():
    """and the decomposed field is the right field: Philox restatement + numpy irfftn."""
    from randomfield_b200 import slab
    shape, spacing, seed = (64, 128, 64), 2.0, 7
    power = oracle.load_default_power()
    kf, _ = oracle.philox_kfield(shape, spacing, power, seed)
    truth = np.fft.irfftn(kf, s=shape, axes=(0, 1, 2))
    got, _, _ = slab.run_virtual_slabs(*shape, spacing, power, seed=seed, world=4)
    err = np.sqrt(np.mean((got - truth) ** 2)) / truth.std()
    assert err <= 1e-5


def test_potential_and_zscale_across_decompositions():
    from randomfield_b200 import slab
    shape = (64, 64, 128)
    power = oracle.load_default_power()
    zs = np.linspace(0.5, 1.5, shape[2])
    ref, _, _ = slab.run_virtual_slabs(*shape, 2.0, power, seed=3, world=1, potential=True, zscale=zs)
    got, _, _ = slab.run_virtual_slabs(*shape, 2.0, power, seed=3, world=2, potential=True, zscale=zs)
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("mode", ["2"])
@pytest.mark.parametrize("shape", [(64, 256, 256), (128, 1024, 128), (64, 64, 2048), (256, 512, 1024)])
def test_fused_yz_kernels_bit_identical(monkeypatch, mode, shape):
    """Fused pass-2 + pass-3 kernel (RF_FUSE=2: k23q, in-place tiles in a three-deep ring) wherever
    it fits: same arithmetic in the same order as k2p + k3q."""
    from randomfield_b200 import slab
    power = oracle.load_default_power()
    zs = np.linspace(0.5, 1.5, shape[2])
    monkeypatch.setenv("RF_FUSE", "0")
    ref, _, std1 = slab.run_virtual_slabs(*shape, 2.0, power, seed=5, world=1, zscale=zs)
    monkeypatch.setenv("RF_FUSE", mode)
    for world in (1, 2):
        got, _, std = slab.run_virtual_slabs(*shape, 2.0, power, seed=5, world=world, zscale=zs)
        assert np.array_equal(got, ref), "world=%d" % world
        assert abs(std - std1) <= 1e-9 * std1


def test_fused_yz_is_the_default_on_large_grids(monkeypatch):
    """2^26 cells with ny, nz <= 1024: rf_finish_yz runs k23q_fused_yz unless RF_FUSE=0 -- same bits."""
    from randomfield_b200 import _lib, slab
    shape = (256, 512, 512)
    power = oracle.load_default_power()
    monkeypatch.delenv("RF_FUSE", raising=False)
    e = _lib.Engine(*shape, dtype=np.float32)
    assert e.fused_yz == 2
    e.close()
    small = _lib.Engine(128, 128, 128, dtype=np.float32)
    assert small.fused_yz == 0
    small.close()
    got, _, std = slab.run_virtual_slabs(*shape, 2.0, power, seed=8, world=1)
    monkeypatch.setenv("RF_FUSE", "0")
    ref, _, std0 = slab.run_virtual_slabs(*shape, 2.0, power, seed=8, world=1)
    assert np.array_equal(got, ref) and abs(std - std0) <= 1e-9 * std0


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("direct", ["1", "0", "chunked", "split"])
def test_nccl_two_ranks_matches_single(tmp_path, direct):
    """Real one-process-per-GPU run (direct peer stores, the NCCL all-to-all, and the chunked
    copy-engine exchange): rank slabs written to disk, compared with world=1."""
    script = os.path.join(ROOT, "tests", "multi_worker.py")
    out = str(tmp_path)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port",
           {"1": "29533", "0": "29535", "split": "29539"}.get(direct, "29537"), script, out]
    env = dict(os.environ, RF_DIRECT=direct)
    if direct in ("chunked", "split"):
        env = dict(os.environ, RF_EXCHANGE="chunked", RF_CHUNKS="2")
    if direct == "split":   # rows [0, 64) with pass 1, then two ranges of 32 rows
        env.update(RF_SPLIT_A="0.5", RF_SPLIT_NSUB="2")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    from randomfield_b200 import slab
    assert open(os.path.join(out, "mode.txt")).read() == {"1": "direct", "0": "nccl", "split": "chunked"}.get(direct, direct)
    ref, _, std1 = slab.run_virtual_slabs(256, 256, 128, 2.0, oracle.load_default_power(), seed=42, world=1)
    got = np.concatenate([np.load(os.path.join(out, "rank%d.npy" % r)) for r in range(2)], axis=0)
    assert np.array_equal(got, ref)
    std = float(open(os.path.join(out, "std.txt")).read())
    assert abs(std - std1) <= 1e-9 * std1


@pytest.mark.skipif(_ngpu() < 2, reason="needs two GPUs")
def test_generator_over_two_ranks(tmp_path):
    """randomfield_b200.Generator(group=True) under torchrun (generate.py:74-75,144-146,232-233,
    282-283,353-354 slab-decomposed): the ranks' delta, rho, Phi and psi slabs concatenated along
    axis 0 equal the single-GPU Generator's arrays bit for bit, and delta_field_rms agrees."""
    import randomfield_b200 as rf
    shape = (256, 256, 128)
    script = os.path.join(ROOT, "tests", "multi_worker_gen.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29541", script, str(tmp_path),
           ",".join(map(str, shape))]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    parts = [np.load(os.path.join(str(tmp_path), "gen_rank%d.npz" % k)) for k in range(2)]
    gen = rf.Generator(*shape, 2.0)
    delta = gen.generate_delta_field(seed=77, smoothing_length_Mpc_h=1.0).copy()
    assert np.array_equal(np.concatenate([p["delta"] for p in parts], axis=0), delta)
    assert abs(float(parts[0]["rms"]) - float(gen.delta_field_rms)) <= 1e-6 * float(gen.delta_field_rms)
    assert float(parts[0]["rms"]) == float(parts[1]["rms"])
    # the density uses the rms: equal to float32 rounding of the table built from it
    rho = gen.convert_delta_to_density().copy()
    got = np.concatenate([p["rho"] for p in parts], axis=0)
    assert np.all(np.abs(got - rho) <= 2e-6 * rho)
    phi = gen.calculate_newtonian_potential(light_cone=True).copy()
    assert np.array_equal(np.concatenate([p["phi"] for p in parts], axis=0), phi)
    assert np.array_equal(parts[0]["phi_gathered"], phi)
    psi = gen.calculate_lensing_potential(i_min=3).copy()
    assert np.array_equal(np.concatenate([p["psi"] for p in parts], axis=0), psi)
    assert np.array_equal(np.concatenate([p["phi_after"] for p in parts], axis=0), phi)
    assert [tuple(p["x_slab"]) for p in parts] == [(0, 128), (128, 256)]
