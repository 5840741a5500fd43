"""Multi-GPU sharding of the tracer path: one process per GPU, torch.distributed for the plumbing.

Every (pixel, sample) is independent given the counter-based RNG, so the path shards
with no data-path collective until the frame is assembled:

  sample-split   rank g traces samples sample_range(g) of every pixel into an f32 float4 sum buffer;
                 the exchange adds the buffers in rank order and tone-maps (tracer.rs:177-184).

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
