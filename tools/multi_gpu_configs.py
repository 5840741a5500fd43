"""Configs D (tile-interleaved split, strong scaling) and E (4K sample-split) on N GPUs of one box (development aid;
bench.py is the contract).  torchrun --nproc-per-node N tools/multi_gpu_configs.py [--spp-e 512] [--spp-d 64]"""
import argparse, importlib.util, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
import forma_b200
from forma_b200.distributed import FusedExchange, TileExchange

ap = argparse.ArgumentParser()
ap.add_argument("--spp-e", type=int, default=512)
ap.add_argument("--spp-d", type=int, default=64)
ap.add_argument("--reps", type=int, default=3)
args = ap.parse_args()
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ts = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(ts)


def timed(fn, reps):
    fn(); dist.barrier(); torch.cuda.synchronize()
    best = None
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier(); torch.cuda.synchronize()
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        t = torch.tensor([a.elapsed_time(b)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = t.item() if best is None else min(best, t.item())
    return best


def total_rays(m):
    r = torch.tensor([float(m.stats().rays)], dtype=torch.float64, device=dev)
    dist.all_reduce(r)
    return r.item()


S = os.path.join(ROOT, "tests", "golden", "scenes")
# ---- E: scene_09 at 3840x2160, spp_e samples per GPU, fused peer-memory reduce + tonemap
m = forma_b200.create_model(3840, 2160, device=local, seed=0)
m.scene.load_json(open(os.path.join(S, "scene_09.json")).read()); m.scene.camera.use_json()
ex = FusedExchange(m, dist, dev)
ms = timed(lambda: ex.frame(args.spp_e, 50, ts.cuda_stream), args.reps)
rays = total_rays(m)
if rank == 0:
    print(json.dumps({"case": f"E scene_09 3840x2160 {args.spp_e} spp/GPU sample-split", "n_gpus": world, "ms_per_frame": round(ms, 3),
                      "rays": int(rays), "mrays_s": round(rays / ms / 1e3, 1), "exchange": ex.mode}), flush=True)
m.close()
# ---- D: 100k spheres at 1080p, spp_d samples, tile-interleaved (strong scaling: the frame is fixed)
spec = importlib.util.spec_from_file_location("gen", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
gen = importlib.util.module_from_spec(spec); spec.loader.exec_module(gen)
m = forma_b200.create_model(1920, 1080, device=local, seed=0)
m.scene.load_json(json.dumps(gen.generate(100_000))); m.scene.camera.use_json()
tx = TileExchange(m, dist, dev)
ms = timed(lambda: tx.frame(args.spp_d, 50, ts.cuda_stream), args.reps)
st = m.stats()
per_rank = torch.zeros(world, 2, dtype=torch.float64, device=dev)
per_rank[rank, 0], per_rank[rank, 1] = float(st.kernel_ms), float(st.rays)
dist.all_reduce(per_rank)
rays = total_rays(m)
if rank == 0:
    print(json.dumps({"case": f"D 100k spheres 1920x1080 {args.spp_d} spp tile-interleaved", "n_gpus": world, "ms_per_frame": round(ms, 3),
                      "rays": int(rays), "mrays_s": round(rays / ms / 1e3, 1), "exchange": tx.mode,
                      "kernel_ms_per_rank": [round(x, 2) for x in per_rank[:, 0].tolist()],
                      "mrays_per_rank": [round(x / 1e6, 1) for x in per_rank[:, 1].tolist()]}), flush=True)
dist.barrier()
dist.destroy_process_group()
