"""Slab decomposition of the seed -> delta(x) path across the GPUs of one box.

One process per GPU (torch.distributed, backend "nccl").  Rank g generates the k-space
y-slab ``iy in [g*ny/G, (g+1)*ny/G)`` for every x (pass 1 is local because every x line is
complete), the ranks exchange one all-to-all of the half-complex field, and rank g finishes
the x-slab ``ix in [g*nx/G, (g+1)*nx/G)`` (passes 2 and 3 are local).  Concatenating the ranks'
results along axis 0 gives the reference's C-order (nx, ny, nz) array.  Mode generation is
counter-based, so the field is bit-identical for every G.

Buffer layouts (complex elements, nzc = nz/2):
  send  [nx][ny/G][nzc]            == [G dest][nx/G][ny/G][nzc]  (dest blocks are contiguous)
  recv  [G src][nx/G][ny/G][nzc]   row y of pass 2 lives in block y // (ny/G)
"""
from __future__ import annotations

import os

import numpy as np

from . import _lib
from . import powertools


def slab_ranges(n, world, rank):
    """[lo, hi) owned by `rank` when an axis of length n is split over `world` ranks."""
    if n % world:
        raise ValueError("axis length {0} is not divisible by {1} ranks".format(n, world))
    step = n // world
    return rank * step, (rank + 1) * step


def block_elems(nx, ny, nz, world):
    """Complex elements in one all-to-all block: (nx/G) * (ny/G) * (nz/2)."""
    return (nx // world) * (ny // world) * (nz // 2)


def recv_to_rows(recv, nx, ny, nz, world):
    """Reorder a receive buffer [G src][nx/G][ny/G][nzc] into [nx/G][ny][nzc] (what pass 2
    reads through its row map).  Host-side restatement used by the CPU tests."""
    nxl, nyl, nzc = nx // world, ny // world, nz // 2
    r = np.asarray(recv).reshape(world, nxl, nyl, nzc)
    return np.ascontiguousarray(np.transpose(r, (1, 0, 2, 3)).reshape(nxl, ny, nzc))


def split_plan(nxl, world, frac=None, nsub=None, align=1):
    """x ranges of the split schedule (include/rf_b200.h, rf_chunk_exchange_x): rows [0, xa) of
    the local x-slab cross NVLink while pass 1 runs, the rest afterwards in `nsub` ranges whose
    pass 2 + pass 3 overlap the next range's transfer.  Returns (xa, [(x0, x1), ...]);
    no ranges = the plain chunked schedule.  Range edges are multiples of `align` rows (pass 3
    works on whole 64 KB tiles of lines).  RF_SPLIT_A = fraction of rows sent with pass 1,
    RF_SPLIT_NSUB = later ranges.  Default: on for 8 ranks and more (0.7, 2), off below.  Measured
    on B200: 4096^3 on 8 GPUs 57.3 ms plain, 53.6 / 55.1 / 53.6 ms with
    (0.6, 2) / (0.7, 2) / (0.8, 2), 56.5 ms with (0.45, 4) -- pass 3 of the early rows hides behind the
    tail of the transfer; on 4 GPUs the kernels bound the step and the split loses 1 %."""
    if frac is None:
        frac = float(os.environ.get("RF_SPLIT_A", "0.7" if world >= 8 else "1"))
    if nsub is None:
        nsub = int(os.environ.get("RF_SPLIT_NSUB", "2" if world >= 8 else "4"))
    units = nxl // align if align > 0 and nxl % align == 0 else 0
    if not 0.0 < frac < 1.0 or nsub < 1 or units < 2:
        return nxl, []
    ua = min(units - 1, max(1, int(round(frac * units))))
    nsub = min(nsub, units - ua)
    edges = [align * (ua + ((units - ua) * i) // nsub) for i in range(nsub + 1)]
    return align * ua, [(edges[i], edges[i + 1]) for i in range(nsub) if edges[i + 1] > edges[i]]


def pass3_row_align(ny, nz):
    """x rows per whole number of pass-3 tiles (64 KB of packed half-complex lines)."""
    lines_per_tile = max(1, 65536 // (nz // 2 * 8))
    return lines_per_tile // np.gcd(lines_per_tile, ny)


def all_to_all(send, recv, group=None):
    """One all-to-all of `world` equal contiguous blocks (torch tensors, any backend)."""
    import torch.distributed as dist
    dist.all_to_all_single(recv, send, group=group)


class SlabEngine(object):
    """The device side of a slab-decomposed Generator on this rank."""

    def __init__(self, nx, ny, nz, spacing, power, smoothing=0.0, dtype=np.float32, group=None,
                 direct=None, chunks=None):
        import torch
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.shape = (nx, ny, nz)
        self.engine = _lib.Engine(nx, ny, nz, dtype=dtype, rank=self.rank, world=self.world)
        k, sig = powertools.sigma_table(powertools.filter_power(power, smoothing), self.shape, spacing)
        self.engine.set_power(k, sig, spacing)
        self.stream = torch.cuda.Stream()
        self.engine.set_stream(self.stream.cuda_stream)
        itemsize = self.engine.complex_dtype.itemsize
        self._xfer_bytes = nx * (ny // self.world) * (nz // 2) * itemsize
        if self.world > 1:
            self._send = self.engine.torch_view(_lib.RF_BUF_FIELD, self._xfer_bytes)
            self._recv = self.engine.torch_view(_lib.RF_BUF_RECV, self._xfer_bytes)
        # Exchange modes (env RF_EXCHANGE / arguments):
        #   "chunked": pass 1, copy-engine pushes into the peers' receive buffers and pass 2 are
        #              pipelined over chunks of z columns (default; rf_chunk_* in the C ABI);
        #   "direct" : pass 1 stores into the peers' receive buffers over NVLink, no collective;
        #   "nccl"   : one NCCL all-to-all between pass 1 and pass 2.
        # "chunked" and "direct" map the peers' receive buffers through CUDA IPC and fence them
        # with 1-element all-reduces.
        mode = os.environ.get("RF_EXCHANGE", "")
        if direct is not None:
            mode = "direct" if direct else (mode if mode in ("nccl", "chunked") else "nccl")
        elif not mode:
            mode = "direct" if os.environ.get("RF_DIRECT") == "1" else (
                "nccl" if os.environ.get("RF_DIRECT") == "0" else "chunked")
        if not 2 <= self.world <= 8:
            mode = "single"
        self.nchunks = 0
        if mode == "chunked":
            want = int(os.environ.get("RF_CHUNKS", "0")) or (chunks or 8)
            while want > 1:
                try:
                    self.engine.chunk_config(want)
                    self.nchunks = want
                    break
                except (RuntimeError, ValueError):
                    want //= 2
            if not self.nchunks:
                mode = "nccl"
        self._xa, self._subs = (split_plan(nx // self.world, self.world, align=pass3_row_align(ny, nz))
                                if mode == "chunked" else (nx // self.world, []))
        self.mode = mode
        self.direct = mode == "direct"
        self._peer_ptrs = []
        if mode in ("direct", "chunked"):
            handles = [None] * self.world
            dist.all_gather_object(handles, self.engine.ipc_export(_lib.RF_BUF_RECV), group=group)
            for r, h in enumerate(handles):
                if r == self.rank:
                    continue
                ptr = _lib.Engine.ipc_open(h)
                self._peer_ptrs.append(ptr)
                self.engine.set_peer_recv(r, ptr)
            if self.direct:
                self.engine.set_direct(True)
            self._fence = torch.zeros(1, device="cuda")
            if mode == "chunked":
                self._comm = torch.cuda.ExternalStream(self.engine.comm_stream)

    @property
    def exchange_bytes(self):
        """Bytes this rank sends to other ranks per step (NVLink, either mode)."""
        return self._xfer_bytes * (self.world - 1) // self.world

    def generate(self, seed, potential=False, zscale=None, want_stats=True):
        """seed -> this rank's real-space x-slab, left resident on the device."""
        import torch
        import torch.distributed as dist
        with torch.cuda.stream(self.stream):
            if self.world == 1:
                self.engine.generate_delta(seed, potential, zscale, want_stats)
            else:
                if self.mode == "chunked":
                    eng, q = self.engine, self._comm
                    nc, xa, subs = self.nchunks, self._xa, self._subs

                    def fence():
                        with torch.cuda.stream(q):
                            dist.all_reduce(self._fence, group=self.group)

                    eng.chunk_begin_step()
                    fence()      # peers have finished reading their receive buffers (previous step)
                    for c in range(nc):
                        eng.chunk_xpass(seed, potential, c)
                        eng.chunk_exchange_x(c, 0, xa, True)
                        fence()  # every rank's copies of chunk c have landed
                        eng.chunk_arrived(c)
                        if c >= 1:
                            eng.chunk_pass2_x(c - 1, 0, xa, c - 1)
                    eng.chunk_pass2_x(nc - 1, 0, xa, nc - 1)
                    for s, (x0, x1) in enumerate(subs):
                        eng.chunk_exchange_x(-1, x0, x1, False)
                        fence()
                        eng.chunk_arrived(nc + s)
                    eng.chunk_pass3_x(0, xa, zscale, want_stats, last=not subs)
                    for s, (x0, x1) in enumerate(subs):
                        eng.chunk_pass2_x(-1, x0, x1, nc + s)
                        eng.chunk_pass3_x(x0, x1, zscale, want_stats, last=s == len(subs) - 1)
                    return
                elif self.direct:
                    # peers have finished reading their receive buffers (previous step) ...
                    dist.all_reduce(self._fence, group=self.group)
                    self.engine.generate_xpass(seed, potential)
                    # ... and every rank's pass-1 stores have landed before anyone reads
                    dist.all_reduce(self._fence, group=self.group)
                else:
                    self.engine.generate_xpass(seed, potential)
                    all_to_all(self._send, self._recv, self.group)
                self.engine.finish_yz(zscale, want_stats)

    def global_stats(self):
        """(mean, std) of the whole field: local sums combined across ranks."""
        import torch
        import torch.distributed as dist
        s1, s2, n = self.engine.stats()
        if self.world > 1:
            t = torch.tensor([s1, s2, float(n)], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, group=self.group)
            s1, s2, n = (float(v) for v in t.tolist())
        mean = s1 / n
        return mean, float(np.sqrt(max(s2 / n - mean * mean, 0.0)))

    def local_field(self):
        return self.engine.copy_field()

    def iter_field_rows(self, rows=None):
        """This rank's x-slab streamed to the host chunk by chunk (Engine.iter_rows): yields
        (x0 within the slab, (n, ny, nz) array in page-locked staging memory)."""
        return self.engine.iter_rows(rows)


def run_virtual_slabs(nx, ny, nz, spacing, power, seed, world, smoothing=0.0, dtype=np.float32,
                      potential=False, zscale=None, direct=False, chunks=0, split=None):
    """All `world` ranks of the slab pipeline driven from ONE process on ONE GPU: each rank's
    engine generates its y-slab, the all-to-all is done with device copies between the
    engines' buffers, each rank finishes its x-slab.  Returns (field (nx, ny, nz), mean, std).
    Used to check that the field is bit-identical for every decomposition."""
    import torch
    shape = (nx, ny, nz)
    k, sig = powertools.sigma_table(powertools.filter_power(power, smoothing), shape, spacing)
    engines = [_lib.Engine(nx, ny, nz, dtype=dtype, rank=r, world=world) for r in range(world)]
    itemsize = engines[0].complex_dtype.itemsize
    xfer = nx * (ny // world) * (nz // 2) * itemsize
    if direct and world > 1:
        for e in engines:
            for r, peer in enumerate(engines):
                e.set_peer_recv(r, peer.device_ptr(_lib.RF_BUF_RECV))
            e.set_direct(True)
    if chunks:
        # chunked exchange: every rank's pass 1 + copy-engine pushes, then every rank's pass 2
        for e in engines:
            e.set_power(k, sig, spacing)
            for r, peer in enumerate(engines):
                e.set_peer_recv(r, peer.device_ptr(_lib.RF_BUF_RECV))
            e.chunk_config(chunks)
        xa, subs = (split_plan(nx // world, world, *split, align=pass3_row_align(ny, nz)) if split
                    else (nx // world, []))
        for e in engines:
            e.chunk_begin_step()
            for c in range(chunks):
                e.chunk_xpass(seed, potential, c)
                e.chunk_exchange_x(c, 0, xa, True)
            for x0, x1 in subs:
                e.chunk_exchange_x(-1, x0, x1, False)
        torch.cuda.synchronize()
        for e in engines:
            for c in range(chunks):
                e.chunk_arrived(c)
                e.chunk_pass2_x(c, 0, xa, c)
            e.chunk_pass3_x(0, xa, zscale, True, last=not subs)
            for s, (x0, x1) in enumerate(subs):
                e.chunk_arrived(chunks + s)
                e.chunk_pass2_x(-1, x0, x1, chunks + s)
                e.chunk_pass3_x(x0, x1, zscale, True, last=s == len(subs) - 1)
    for e in engines:
        if chunks:
            break
        e.set_power(k, sig, spacing)
        e.generate_xpass(seed, potential)
        e.sync()
    if world > 1 and not direct and not chunks:
        sends = [e.torch_view(_lib.RF_BUF_FIELD, xfer).view(world, -1) for e in engines]
        recvs = [e.torch_view(_lib.RF_BUF_RECV, xfer).view(world, -1) for e in engines]
        for d in range(world):
            for s in range(world):
                recvs[d][s].copy_(sends[s][d])
        torch.cuda.synchronize()
    s1 = s2 = 0.0
    n = 0
    parts = []
    for e in engines:
        if not chunks:
            e.finish_yz(zscale, True)
        a, b, c = e.stats()
        s1, s2, n = s1 + a, s2 + b, n + c
        parts.append(e.copy_field())
        e.close()
    mean = s1 / n
    return np.concatenate(parts, axis=0), mean, float(np.sqrt(max(s2 / n - mean * mean, 0.0)))
