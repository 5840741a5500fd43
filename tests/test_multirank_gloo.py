"""world_size-2 `gloo` test (CPU): the host-side sharding logic of the N > 1 path.  Each rank's
partial sums come from the oracle (standing in for the device kernel, which needs a GPU); what is
tested is the partition arithmetic, the byte all_gather used for the IPC handles, and that
rank-order reduction + tone-map of the parts reproduces the single-rank frame."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from forma_b200 import distributed as fd
    from oracle import binding as ob
    w, h, spp_total = 48, 32, 6
    objs, cam = ob.builtin_scene(2), ob.camera_new(w, h)
    off, spp = fd.sample_range(rank, world, spp_total)
    _, part, _ = ob.render(objs, cam, w, h, spp, sample_offset=off, seed=5, want_rgb=False)
    # the byte all_gather that carries the CUDA IPC handles
    got = fd.gather_bytes(dist, bytes([rank]) * 128, torch.device("cpu"))
    assert got == [bytes([r]) * 128 for r in range(world)]
    # rank-order sum == what reduce_tonemap_kernel computes on its stripe
    t = torch.from_numpy(part.copy())
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    acc = parts[0].clone()
    for p in parts[1:]:
        acc += p
    lo, hi = fd.pixel_stripe(rank, world, w * h)
    stripe = ob.tonemap(acc.numpy().reshape(-1, 3)[lo:hi], spp_total)
    np.save(os.path.join(out_dir, f"stripe{rank}.npy"), stripe)
    np.save(os.path.join(out_dir, f"acc{rank}.npy"), acc.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_sample_split_two_ranks(tmp_path):
    from oracle import binding as ob
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    w, h, spp = 48, 32, 6
    rgb, lin, _ = ob.render(ob.builtin_scene(2), ob.camera_new(w, h), w, h, spp, seed=5)
    acc = np.load(tmp_path / "acc0.npy")
    assert np.array_equal(acc, np.load(tmp_path / "acc1.npy"))
    np.testing.assert_allclose(acc, lin, rtol=2e-6, atol=1e-6)     # f32 re-association only
    frame = np.concatenate([np.load(tmp_path / "stripe0.npy"), np.load(tmp_path / "stripe1.npy")]).reshape(h, w, 3)
    d = np.abs(frame.astype(int) - rgb.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01                   # at most 1 LSB on rare pixels


def test_partitions_cover_exactly():
    from forma_b200 import distributed as fd
    for world in (1, 2, 3, 4, 8):
        for total in (1, 7, 256, 4096):
            r = [fd.sample_range(g, world, total) for g in range(world)]
            assert r[0][0] == 0 and sum(n for _, n in r) == total
            assert all(r[g][0] + r[g][1] == r[g + 1][0] for g in range(world - 1))
        s = [fd.pixel_stripe(g, world, 1920 * 1080) for g in range(world)]
        assert s[0][0] == 0 and s[-1][1] == 1920 * 1080 and all(s[g][1] == s[g + 1][0] for g in range(world - 1))


def test_tile_interleave_covers_every_pixel_once():
    """forma_rt_set_partition's ownership rule (32x4 tiles dealt round-robin): a partition of the frame,
    balanced to within one tile, for ragged sizes too."""
    from forma_b200 import distributed as fd
    for w, h in ((1920, 1080), (600, 600), (67, 45), (31, 3), (1, 1)):
        for world in (1, 2, 3, 4, 8):
            own = fd.owner_map(w, h, world)
            assert own.shape == (h, w) and own.min() >= 0 and own.max() < world
            assert own[0, 0] == fd.tile_owner(0, 0, w, world) and own[h - 1, w - 1] == fd.tile_owner(w - 1, h - 1, w, world)
            counts = np.bincount(own.ravel(), minlength=world)
            assert counts.sum() == w * h
            tiles = ((w + 31) // 32) * ((h + 3) // 4)
            if tiles >= world:
                assert counts.max() - counts.min() <= 2 * 32 * 4
