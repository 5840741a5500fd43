"""Multi-GPU sharding of the tracer path: one process per GPU, torch.distributed for the plumbing.

Every (pixel, sample) is independent given the counter-based RNG, so the path shards
with no data-path collective until the frame is assembled:

  sample-split   rank g traces samples sample_range(g) of every pixel into an f32 float4 sum buffer;
                 the exchange adds the buffers in rank order and tone-maps (tracer.rs:177-184).

  tile-split     rank g traces every sample of the 32x4-pixel tiles t with t % world == g
                 (forma_rt_set_partition) and stores RGB8 straight into rank 0's frame through a peer
                 pointer: no exchange kernel at all, and bit-exact against one GPU (same sample order).

The reference's only notion of parallelism is render_mt's 4 row bands (tracer.rs:83-134).
"""
from __future__ import annotations


def sample_range(rank: int, world: int, spp_total: int) -> tuple[int, int]:
    """(sample_offset, spp) of `rank`: contiguous, disjoint, covering [0, spp_total)."""
    lo = spp_total * rank // world
    hi = spp_total * (rank + 1) // world
    return lo, hi - lo


def pixel_stripe(rank: int, world: int, npix: int) -> tuple[int, int]:
    """[begin, end) of the pixels `rank` reduces and tone-maps in the fused exchange."""
    return npix * rank // world, npix * (rank + 1) // world


TILE_W, TILE_H = 32, 4  # the partition unit of forma_rt_set_partition (= render_kernel's CTA tile)


def tile_pitch(width: int, world: int) -> int:
    """Smallest pitch >= tiles per row that is coprime with `world`: the deal rotates from row to row."""
    import math
    p = (width + TILE_W - 1) // TILE_W
    while math.gcd(p, world) != 1:
        p += 1
    return p


def tile_owner(x: int, y: int, width: int, world: int) -> int:
    """Rank that traces pixel (x, y) under the tile-interleaved split."""
    return ((y // TILE_H) * tile_pitch(width, world) + x // TILE_W) % world


def owner_map(width: int, height: int, world: int):
    """[height, width] array of owning ranks (numpy; test and tooling aid)."""
    import numpy as np
    ys, xs = np.mgrid[0:height, 0:width]
    return ((ys // TILE_H) * tile_pitch(width, world) + xs // TILE_W) % world


def gather_bytes(dist, payload: bytes, device) -> list[bytes]:
    """all_gather of one fixed-size byte string per rank (CUDA IPC handles travel this way)."""
    import torch
    mine = torch.tensor(list(payload), dtype=torch.uint8, device=device)
    out = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


class FusedExchange:
    """Sample-split exchange over peer memory, with no host-side barrier in the frame.

    Every rank exports two float4 sum buffers (double-buffered) and two flags through CUDA IPC, rank 0 its RGB8 frame
    as well.  Frame e (1, 2, ...) on every rank, all stream-ordered:
        wait   peers' done  >= e - 2      their reduce of frame e - 2 has finished reading my sum[e & 1]
        render my samples -> sum[e & 1]
        set    my ready = e
        wait   peers' ready >= e          their partial sums of frame e are complete in their HBM
        reduce_tonemap on my pixel stripe: peer loads over NVLink, rank-order sum, tone-map, RGB8 into rank 0's frame
        set    my done = e
        rank 0: wait peers' done >= e     the frame is complete in rank 0's memory
    A rank never waits for more than the data it needs; nothing returns to the host until the caller synchronises."""

    mode = "fused peer-memory reduce+tonemap, flag-ordered (CUDA IPC over NVLink, no barrier per frame)"

    def __init__(self, model, dist, device):
        import ctypes
        from . import _lib
        self.m, self.dist = model, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.npix = model.width * model.height

        def alloc(nbytes):
            p = ctypes.c_void_p()
            _lib.check(_lib.lib().forma_rt_device_alloc(model._h, nbytes, ctypes.byref(p)))  # zero-filled
            return p.value

        self.sums = [alloc(self.npix * 16), alloc(self.npix * 16)]
        self.rgb = alloc(self.npix * 3)
        self.flags = alloc(256)  # ready at +0, done at +128 (separate lines)
        payload = b"".join(model.ipc_export(p) for p in (self.sums[0], self.sums[1], self.rgb, self.flags))
        handles = gather_bytes(dist, payload, device)  # a collective: every rank's flags are zero before anyone waits

        def peer(r, k, own):
            return own if r == self.rank else model.ipc_open(handles[r][64 * k:64 * (k + 1)])

        self.peer_sums = [[peer(r, b, self.sums[b]) for r in range(self.world)] for b in (0, 1)]
        self.rgb0 = peer(0, 2, self.rgb)
        peer_flags = [peer(r, 3, self.flags) for r in range(self.world)]
        self.ready = [f for f in peer_flags]
        self.done = [f + 128 for f in peer_flags]
        self.stripe = pixel_stripe(self.rank, self.world, self.npix)
        self.epoch = 0
        self._opened = [p for r in range(self.world) if r != self.rank
                        for p in (self.peer_sums[0][r], self.peer_sums[1][r], peer_flags[r])]
        if self.rank != 0:
            self._opened.append(self.rgb0)

    def close(self):
        """Unmap the peers' buffers (forma_rt_ipc_close).  Call on every rank after a final synchronise + barrier: a
        peer may still be reading this rank's sums until then.  `world` must not exceed the number of GPUs: two
        ranks on one GPU would spin on each other's flags (the wait gives up after FORMA_RT_FLAG_TIMEOUT_MS)."""
        for p in self._opened:
            self.m.ipc_close(p)
        self._opened = []

    def frame(self, spp_per_rank: int, max_depth: int, stream: int) -> int:
        """One frame of spp_per_rank * world samples per pixel; returns kernels launched on this rank."""
        m = self.m
        self.epoch += 1
        e, b, n = self.epoch, self.epoch & 1, 5
        if e > 2:
            m.flag_wait_device(self.done, e - 2, stream)
            n += 1
        m.render_device(spp_per_rank, self.rank * spp_per_rank, max_depth, None, self.sums[b], stream)
        m.flag_set_device(self.ready[self.rank], e, stream)
        m.flag_wait_device(self.ready, e, stream)
        m.reduce_tonemap_device(self.peer_sums[b], spp_per_rank * self.world, self.stripe[0], self.stripe[1],
                                self.rgb0, None, stream)
        m.flag_set_device(self.done[self.rank], e, stream)
        if self.rank == 0:
            m.flag_wait_device(self.done, e, stream)
            n += 1
        return n


class TileExchange:
    """Tile-interleaved split over peer memory: rank 0 exports its RGB8 frame through CUDA IPC, every rank
    renders its own tiles (all samples) straight into it.  One trace launch per rank and frame, no reduce, and no
    host-side barrier: frame e is ordered by two flags in peer-visible memory,
        rank 0: set go = e            everything rank 0 queued before (the caller's copy of frame e - 1) is complete
        all:    wait rank 0's go >= e
        render own tiles -> rank 0's frame
        set my done = e
        rank 0: wait peers' done >= e the frame is complete in rank 0's memory."""

    mode = "tile-interleaved (32x4 tiles), RGB8 stored into rank 0's frame over NVLink (CUDA IPC), flag-ordered"

    def __init__(self, model, dist, device):
        import ctypes
        from . import _lib
        self.m, self.dist = model, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.npix = model.width * model.height

        def alloc(nbytes):
            q = ctypes.c_void_p()
            _lib.check(_lib.lib().forma_rt_device_alloc(model._h, nbytes, ctypes.byref(q)))  # zero-filled
            return q.value

        self.rgb = alloc(self.npix * 3)
        self.flags = alloc(256)  # done at +0, go (rank 0's) at +128
        handles = gather_bytes(dist, model.ipc_export(self.rgb) + model.ipc_export(self.flags), device)
        self.rgb0 = self.rgb if self.rank == 0 else model.ipc_open(handles[0][:64])
        peer_flags = [self.flags if r == self.rank else model.ipc_open(handles[r][64:]) for r in range(self.world)]
        self.done = peer_flags
        self.go = peer_flags[0] + 128
        self.epoch = 0
        self._opened = [p for r, p in enumerate(peer_flags) if r != self.rank] + ([self.rgb0] if self.rank != 0 else [])
        model.set_partition(self.rank, self.world)

    def close(self):
        """Unmap the peers' buffers and give the handle its whole frame back (partition (0, 1)).  Call on every rank
        after a final synchronise + barrier."""
        for p in self._opened:
            self.m.ipc_close(p)
        self._opened = []
        self.m.set_partition(0, 1)

    def frame(self, spp: int, max_depth: int, stream: int) -> int:
        m = self.m
        self.epoch += 1
        e = self.epoch
        if self.rank == 0:
            m.flag_set_device(self.go, e, stream)
        m.flag_wait_device([self.go], e, stream)
        m.render_device(spp, 0, max_depth, self.rgb0, None, stream)
        m.flag_set_device(self.done[self.rank], e, stream)
        if self.rank == 0:
            m.flag_wait_device(self.done, e, stream)
        return 1
