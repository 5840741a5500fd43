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
    """Sample-split exchange over peer memory: every rank exports its float4 sum buffer (and rank 0
    its RGB8 frame) through CUDA IPC; after a barrier each rank runs reduce_tonemap_kernel on its
    pixel stripe — peer loads over NVLink, rank-order sum, tone-map, RGB8 stored into rank 0's frame."""

    mode = "fused peer-memory reduce+tonemap (CUDA IPC over NVLink)"

    def __init__(self, model, dist, device):
        import ctypes
        from . import _lib
        self.m, self.dist = model, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.npix = model.width * model.height
        p, q = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(_lib.lib().forma_rt_device_alloc(model._h, self.npix * 16, ctypes.byref(p)))
        _lib.check(_lib.lib().forma_rt_device_alloc(model._h, self.npix * 3, ctypes.byref(q)))
        self.sum, self.rgb = p.value, q.value
        handles = gather_bytes(dist, model.ipc_export(self.sum) + model.ipc_export(self.rgb), device)
        self.peer_sums = [self.sum if r == self.rank else model.ipc_open(h[:64]) for r, h in enumerate(handles)]
        self.rgb0 = self.rgb if self.rank == 0 else model.ipc_open(handles[0][64:])
        self.stripe = pixel_stripe(self.rank, self.world, self.npix)

    def frame(self, spp_per_rank: int, max_depth: int, stream: int) -> int:
        """One frame of spp_per_rank * world samples per pixel; returns kernels launched on this rank."""
        m = self.m
        m.render_device(spp_per_rank, self.rank * spp_per_rank, max_depth, None, self.sum, stream)
        self.dist.barrier()  # every rank's partial sums are complete in its HBM
        m.reduce_tonemap_device(self.peer_sums, spp_per_rank * self.world, self.stripe[0], self.stripe[1],
                                self.rgb0, None, stream)
        self.dist.barrier()  # rank 0's frame is complete
        return 2


class TileExchange:
    """Tile-interleaved split over peer memory: rank 0 exports its RGB8 frame through CUDA IPC, every rank
    renders its own tiles (all samples) straight into it.  One trace launch per rank and frame, no reduce."""

    mode = "tile-interleaved (32x4 tiles), RGB8 stored into rank 0's frame over NVLink (CUDA IPC)"

    def __init__(self, model, dist, device):
        import ctypes
        from . import _lib
        self.m, self.dist = model, dist
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.npix = model.width * model.height
        q = ctypes.c_void_p()
        _lib.check(_lib.lib().forma_rt_device_alloc(model._h, self.npix * 3, ctypes.byref(q)))
        self.rgb = q.value
        handles = gather_bytes(dist, model.ipc_export(self.rgb), device)
        self.rgb0 = self.rgb if self.rank == 0 else model.ipc_open(handles[0])
        model.set_partition(self.rank, self.world)

    def frame(self, spp: int, max_depth: int, stream: int) -> int:
        self.m.render_device(spp, 0, max_depth, self.rgb0, None, stream)
        self.dist.barrier()  # stream-ordered (NCCL on the launch stream): rank 0's frame is complete
        return 1

