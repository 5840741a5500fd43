"""The flag-ordered sample-split exchange against a single-GPU composition of the same partial sums.
torchrun --nproc-per-node N tools/exchange_check.py [frames]  (N GPUs of one box)

Every frame, rank 0's RGB8 frame must equal what ONE GPU gets from the same per-rank sample ranges summed in rank
order (forma_rt_reduce_tonemap_device on local buffers) - byte for byte, for every frame of a run whose per-frame work
varies from rank to rank and from frame to frame (so that ranks run ahead of and behind each other)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import forma_b200  # noqa: E402
from forma_b200.distributed import FusedExchange  # noqa: E402

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 24
world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ts = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(ts)
w, h = 640, 360
m = forma_b200.create_model(w, h, device=local, seed=3)
m.scene.load_builtin(2)
ex = FusedExchange(m, dist, dev)
npix = w * h
bad = 0
host = np.empty((h, w, 3), np.uint8)
refs = []
if rank == 0:  # reference: the same partial sums on one GPU, summed in rank order
    ref_m = forma_b200.create_model(w, h, device=local, seed=3)
    ref_m.scene.load_builtin(2)
    parts = [torch.zeros((npix, 4), dtype=torch.float32, device=dev) for _ in range(world)]
    ref_rgb = torch.zeros(npix * 3, dtype=torch.uint8, device=dev)
for f in range(frames):
    spp = 1 + (f * 7) % 5  # per-frame work varies
    # skew: some ranks get extra untimed work in front of some frames
    if (f + rank) % 3 == 0:
        torch.empty(64 << 20, dtype=torch.uint8, device=dev).fill_(f & 255)
    ex.frame(spp, 50, ts.cuda_stream)
    if rank == 0:
        ts.synchronize()  # rank 0's stream ends with the wait for every peer's done flag
        from cuda.bindings import runtime as rt
        rt.cudaMemcpy(host.ctypes.data, ex.rgb0, npix * 3, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        for r in range(world):
            ref_m.render_device(spp, r * spp, 50, None, parts[r].data_ptr(), ts.cuda_stream)
        ref_m.reduce_tonemap_device([p.data_ptr() for p in parts], spp * world, 0, npix, ref_rgb.data_ptr(), None, ts.cuda_stream)
        ts.synchronize()
        want = ref_rgb.cpu().numpy().reshape(h, w, 3)
        if not np.array_equal(host, want):
            bad += 1
            print(f"frame {f}: {int((host != want).any(-1).sum())} pixels differ", flush=True)
torch.cuda.synchronize()
dist.barrier()
ex.close()  # unmap the peers' buffers (every rank is past its last frame: barrier above)
if rank == 0:
    print(f"exchange_check: {frames} frames on {world} GPUs, {bad} bad", flush=True)
# ---- the tile-interleaved split: rank 0's frame must equal ONE GPU's render of the same frame, bit for bit
from forma_b200.distributed import TileExchange  # noqa: E402
from forma_b200 import _lib  # noqa: E402
tm = forma_b200.create_model(w, h, device=local, seed=5, accel=_lib.ACCEL_LBVH)
tm.scene.load_builtin(2)
tx = TileExchange(tm, dist, dev)
bad_t = 0
if rank == 0:
    one = forma_b200.create_model(w, h, device=local, seed=5, accel=_lib.ACCEL_LBVH)
    one.scene.load_builtin(2)
for f in range(frames):
    spp = 1 + (f * 3) % 4
    if (f + rank) % 3 == 1:
        torch.empty(64 << 20, dtype=torch.uint8, device=dev).fill_(f & 255)
    tx.frame(spp, 50, ts.cuda_stream)
    if rank == 0:
        ts.synchronize()
        from cuda.bindings import runtime as rt
        rt.cudaMemcpy(host.ctypes.data, tx.rgb0, npix * 3, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
        want, _ = one.render(spp)
        if not np.array_equal(host, want):
            bad_t += 1
            print(f"tile frame {f}: {int((host != want).any(-1).sum())} pixels differ", flush=True)
torch.cuda.synchronize()
dist.barrier()
tx.close()  # unmap, and give the handle its whole frame back
if rank == 0:
    print(f"exchange_check (tiles): {frames} frames on {world} GPUs, {bad_t} bad", flush=True)
bad += bad_t
dist.destroy_process_group()
sys.exit(1 if bad else 0)
