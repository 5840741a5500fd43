#!/usr/bin/env python
"""bench.py — Mrays/s and ms/frame of the tracer hot path on BASELINE.json's headline config
(config C: scenes/scene_09.json, 1920x1080, 256 spp, max depth 50), N GPUs of one node.

    python bench.py --gpus 1 --steps K --warmup W                    (our arm)
    python -m torch.distributed.run ... bench.py --gpus N ...         (N > 1, one rank per GPU)
    python bench.py --impl reference ...                              (the reference's CPU path)

A "step" is one frame: every (pixel, sample) path of the workload traced and written out as RGB8.
`value` = closest-hit queries (get_color calls, tracer.rs:189) per second with the scene resident in
HBM, timed with CUDA events on the launch stream.  `e2e` = the same frame through the host-facing
C ABI call (scene JSON parsed, flattened and copied H2D, frame copied D2H into pinned host memory,
all inside the timed region).  At N > 1 every rank traces its own 256 samples of every pixel
(sample-split, weak scaling: the job is the N*256-spp frame) and the partial f32 sums are combined by
the fused peer-memory reduce + tone-map kernel (or NCCL reduce as a fallback).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, SPP, DEPTH, SEED = 1920, 1080, 256, 50, 0
SCENE = os.path.join(ROOT, "tests", "golden", "scenes", "scene_09.json")
METRIC, UNIT = "Mrays/s at 1080p scene_09 (256 spp, depth 50)", "Mrays/s"


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def flops_model(rays, paths, n_sph, n_tri, n_plane):
    """algorithmic FP32 work of the brute-force path."""
    hits = max(rays - paths, 0)
    return rays * (23 * n_sph + 45 * n_tri + 20 * n_plane) + 60 * paths + 40 * hits


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index, enabled=True):
        self.rows, self.proc, self.index, self.enabled = [], None, index, enabled

    def start(self):
        """Launch nvidia-smi and wait for its first row (it takes seconds to come up on an 8-GPU box, longer than a
        short timed region); call before the last warm-up steps."""
        if not self.enabled:
            return self
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
            t0 = time.perf_counter()
            while not self.rows and time.perf_counter() - t0 < 10.0:
                time.sleep(0.01)
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line)

    def __enter__(self):
        if self.enabled and self.proc is None:
            self.start()
        self.rows.clear()  # keep only the rows sampled inside the timed region
        return self

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.05)
            self.proc.terminate()
            self.t.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0, set()
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6 or not f[0].isdigit():
                continue
            sm.append(int(f[0])); mx = max(mx, int(f[1]))
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_scene():
    from oracle import binding as ob
    from oracle import scene_oracle as so
    d = so.load(SCENE)
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    return ob, objs, ob.camera_look(fr, at, up, fov, W / H)


def cpu_sample(nthreads, spp, rows=None):
    """Times the oracle (the CPU restatement of the reference tracer) on a bounded sample of config C:
    `spp` samples per pixel over rows [0, rows).  Returns (Mrays/s, seconds, rays)."""
    ob, objs, cam = oracle_scene()
    rows = H if rows is None else rows
    t0 = time.perf_counter()
    _, _, rays = ob.render(objs, cam, W, H, spp, max_depth=DEPTH, seed=SEED, y1=rows, nthreads=nthreads,
                           want_linear=False)
    dt = time.perf_counter() - t0
    return rays / dt / 1e6, dt, rays


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be
    built here (no cargo/rustc, un-vendored ../kopek), so this is the oracle port, all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    spp = 4  # bounded sample: 4 of the 256 spp of the full 1080p frame per step (throughput is spp-independent)
    for _ in range(args.warmup):
        cpu_sample(cores, 1)
    t_tot, r_tot = 0.0, 0
    for _ in range(args.steps):
        _, dt, rays = cpu_sample(cores, spp)
        t_tot += dt; r_tot += rays
    v = r_tot / t_tot / 1e6
    sample = f"{spp} of {SPP} spp of the full {W}x{H} frame per step, {cores} threads (row-interleaved)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * t_tot / args.steps * SPP / spp, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "scenes/scene_09.json 1920x1080 256 spp depth 50 (config C)", "seed": SEED,
                   "note": "ms_per_step is extrapolated to 256 spp from the bounded sample"},
        "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def run_ours(args):
    import torch
    import forma_b200

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    npix = W * H
    scene_txt = open(SCENE).read()

    m = forma_b200.create_model(W, H, device=local, seed=SEED)
    m.scene.load_json(scene_txt)
    m.scene.camera.use_json()
    ts = torch.cuda.Stream(device=dev)  # launch stream; the CUDA events below are recorded on it
    torch.cuda.set_stream(ts)
    stream = ts.cuda_stream
    d_rgb = torch.empty(npix * 3, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2, written between timed steps

    ex = None
    d_sum = None
    if world > 1:
        try:
            from forma_b200.distributed import FusedExchange
            ex = FusedExchange(m, dist, dev)
        except Exception as e:  # noqa: BLE001 — fall back to NCCL reduce, and say so in the JSON line
            sys.stderr.write(f"[bench] rank {rank}: IPC exchange unavailable ({e}); using NCCL reduce\n")
            ex = None
        ok = torch.tensor([1 if ex else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if not ok.item():
            ex = None
            d_sum = torch.empty((npix, 4), dtype=torch.float32, device=dev)

    def step():
        """One frame on this rank; returns kernels launched."""
        if world == 1:
            m.render_device(SPP, 0, DEPTH, d_rgb.data_ptr(), None, stream)
            return 1
        if ex:
            return ex.frame(SPP, DEPTH, stream)
        m.render_device(SPP, rank * SPP, DEPTH, None, d_sum.data_ptr(), stream)
        dist.reduce(d_sum, dst=0)
        if rank == 0:
            m.tonemap_device(d_sum.data_ptr(), SPP * world, d_rgb.data_ptr(), 0, npix, stream)
        return 2 if rank == 0 else 1

    def sync():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    clk = ClockSampler(local, enabled=rank == 0).start()  # rank 0 samples its own GPU
    for _ in range(max(args.warmup, 3)):
        step()
    sync()
    rays_step = None
    launches = 0
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kernel_ms = []
    with clk:
        sync()
        t_wall = time.perf_counter()
        for a, b in ev:
            flush.fill_(1)  # L2 flush between timed iterations (untimed)
            a.record()
            launches += step()
            b.record()
        sync()
        t_wall = time.perf_counter() - t_wall
    step_ms = [a.elapsed_time(b) for a, b in ev]
    # trace-kernel duration of the last step, from the library's own events on the same stream
    st = m.stats()  # finalises the last device-resident render: ray counter + kernel events
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    rays = torch.tensor([float(st.rays)], dtype=torch.float64, device=dev)
    if dist:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    total_ms, rays_step = float(total_ms.item()), float(rays.item())
    ms_per_step = total_ms / args.steps
    value = rays_step / (ms_per_step * 1e-3) / 1e6
    kern_ms = float(st.kernel_ms)

    # ---- e2e: the host-facing call, host buffers, copies inside the timed region ----
    e2e = None
    if True:
        host_rgb = torch.empty(npix * 3, dtype=torch.uint8).pin_memory()
        hp = host_rgb.numpy()
        def e2e_step():
            m.scene.load_json(scene_txt)      # parse + flatten + H2D of the packed scene
            m.scene.camera.use_json()
            if world == 1:
                m.render(SPP, max_depth=DEPTH, rgb_out=hp.reshape(H, W, 3))
            else:
                step()
                if rank == 0:
                    # rank 0's stream ends with the wait for every peer's part of the frame; the copy below runs on
                    # the legacy default stream, which does not order itself behind torch's non-blocking streams
                    ts.synchronize()
                    src = ex.rgb0 if ex else d_rgb.data_ptr()
                    _d2h(hp, src, npix * 3)
        for _ in range(2):
            e2e_step()
        sync()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step()
        sync()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if dist:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        launches += args.steps * (1 if world == 1 else 2)
        stx = m.stats()
        h2d = 16 * (stx.n_spheres + 3 * stx.n_triangles + 3 * stx.n_planes + stx.n_objects) + 12 * stx.n_objects + 20
        e2e = {"value": round(rays_step * args.steps / float(dt.item()) / 1e6, 1), "unit": UNIT,
               "ms_per_step": round(1e3 * float(dt.item()) / args.steps, 3),
               "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": npix * 3 + 8,
               "api": "forma_rt_load_json + forma_rt_camera_json + forma_rt_render (host RGB8 buffer, pinned)" if world == 1 else
                      "forma_rt_load_json + forma_rt_camera_json + forma_rt_render_device + flag-ordered exchange on every "
                      "rank, then rank 0 waits for the frame and copies it to a pinned host RGB8 buffer"}

    others = None
    if world == 1 and not args.no_other_configs:
        others = other_configs(local)
    if rank == 0:
        emit(args, world, value, ms_per_step, rays_step, kern_ms, st, clk.summary(), e2e, launches, t_wall,
             ex.mode if ex else ("NCCL reduce + tonemap" if world > 1 else None), others)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


def other_configs(device):
    """BASELINE.json's other single-GPU configs through the host-facing API (untimed part of the run; parity
    for each is in tests/): kernel time from the library's CUDA events, total = kernel + D2H of the result."""
    import importlib.util
    import forma_b200
    scenes = os.path.dirname(SCENE)
    out = {}

    def best(m, fn, reps=3):
        b = None
        for _ in range(reps):
            fn()
            st = m.stats()
            if b is None or st.kernel_ms < b.kernel_ms:
                b = st
        return b

    m = forma_b200.create_model(600, 600, device=device, seed=SEED)
    m.scene.load_json(open(os.path.join(scenes, "scene_01.json")).read()); m.scene.camera.use_json()
    st = best(m, lambda: m.render(50, max_depth=DEPTH))
    out["A scene_01 600x600 50 spp (JSON camera)"] = {"kernel_ms": round(st.kernel_ms, 3), "total_ms": round(st.total_ms, 3),
                                                     "rays": int(st.rays), "mrays_s": round(st.rays / st.kernel_ms / 1e3, 1)}
    m.close()
    m = forma_b200.create_model(W, H, device=device, seed=SEED)
    m.scene.load_json(open(os.path.join(scenes, "scene_05.json")).read()); m.scene.camera.use_json()
    st = best(m, lambda: m.primary_hits())
    out["B scene_05 1920x1080 primary hit id/t dump"] = {"kernel_ms": round(st.kernel_ms, 3), "total_ms": round(st.total_ms, 3),
                                                        "rays": int(st.rays), "mrays_s": round(st.rays / st.kernel_ms / 1e3, 1)}
    spec = importlib.util.spec_from_file_location("gen", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    gen = importlib.util.module_from_spec(spec); spec.loader.exec_module(gen)
    m.scene.load_json(json.dumps(gen.generate(100_000))); m.scene.camera.use_json()
    st = best(m, lambda: m.render(64, max_depth=DEPTH), reps=2)
    out["D 100k spheres 1920x1080 64 spp (LBVH)"] = {"kernel_ms": round(st.kernel_ms, 3), "total_ms": round(st.total_ms, 3),
                                                    "rays": int(st.rays), "mrays_s": round(st.rays / st.kernel_ms / 1e3, 1),
                                                    "bvh_nodes": int(st.bvh_nodes)}
    # config D's roofline is L2 traffic - counted node visits x 32 B + leaf tests x 16 B from an
    # instrumented (untimed) run of the same traversal, over the timed kernel, against the streaming-from-L2 read
    # bandwidth measured on this box by the same call
    ws = m.walk_stats(64, max_depth=DEPTH)
    l2_bytes = ws.node_visits * ws.node_bytes + ws.leaf_tests * ws.leaf_bytes
    gbs = l2_bytes / (st.kernel_ms * 1e-3) / 1e9
    out["D 100k spheres 1920x1080 64 spp (LBVH)"]["l2_roofline"] = {
        "node_visits": int(ws.node_visits), "leaf_tests": int(ws.leaf_tests), "rays_instrumented": int(ws.rays),
        "visits_per_ray": round(ws.node_visits / max(1, ws.rays), 2), "leaf_tests_per_ray": round(ws.leaf_tests / max(1, ws.rays), 2),
        "bytes": int(l2_bytes), "achieved_gbs": round(gbs, 1), "peak_gbs": round(float(ws.l2_read_gbs), 1),
        "frac": round(gbs / ws.l2_read_gbs, 4) if ws.l2_read_gbs > 0 else None,
        "peak_source": "forma_rt_walk_stats_get: 32 MiB buffer streamed out of L2, 8 passes, CUDA events (measured here)"}
    m.close()
    return out


def _d2h(host_np, d_ptr, nbytes):
    from cuda.bindings import runtime as rt  # cuda-python
    rt.cudaMemcpy(host_np.ctypes.data, d_ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)


def emit(args, world, value, ms_per_step, rays_step, kern_ms, st, clocks, e2e, launches, t_wall, exchange, others=None):
    peaks, src = measured_peaks()
    paths_rank = W * H * SPP
    rays_rank = rays_step / world
    flops = flops_model(rays_rank, paths_rank, st.n_spheres, st.n_triangles, st.n_planes)
    sm_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = st.sm_count * 128 * 2 * sm_mhz * 1e6 / 1e12  # TFLOP/s with FMA at the max SM clock
    achieved = flops / (kern_ms * 1e-3) / 1e12 if kern_ms > 0 else None
    hbm_bytes = 16 * (st.n_spheres + 3 * st.n_triangles + 3 * st.n_planes) + W * H * 3
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "scenes/scene_09.json 1920x1080 256 spp depth 50 (config C)", "seed": SEED,
                   "spp_per_gpu": SPP, "spp_total": SPP * world, "parallelism": f"sample-split x{world}",
                   "exchange": exchange, "l2": "256 MiB buffer written between timed steps (L2 flush)",
                   "build": "exact (-fmad=false, IEEE div/sqrt): bit-identical to the CPU oracle",
                   "rays_per_step": int(rays_step), "wall_s_timed_region": round(t_wall, 4)},
        "roofline": {
            "bound": "fp32-issue",  # neither HBM nor tensor: SURVEY §8(d); see DESIGN.md "Roofline"
            "achieved": round(achieved, 3) if achieved else None, "peak": round(fp32_peak, 2), "unit": "TFLOP/s",
            "frac": round(achieved / fp32_peak, 4) if achieved else None,
            # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full (capture r02f, profiles/r01j_render_kernel_C_ncu.md):
            # the 6.22 MB frame is still in L2 when the kernel ends, the scene is 192 B
            "traffic": 50688 if world == 1 else None,
            "issue_slots": {"active_pct": 78.5, "warp_instr_per_path": 7.5, "source": "ncu smsp__issue_active / smsp__inst_executed of this kernel "
                            "on this config (capture r02f, profiles/r01j_render_kernel_C_ncu.md): the kernel is issue-bound; 40 % of its "
                            "instructions are the counter RNG + rejection sampling, which the flop model does not count"},
            "peak_source": f"{st.sm_count} SMs x 128 lanes x 2 x {sm_mhz:.0f} MHz (sm_max_mhz of MEASURED_PEAKS.json, {src}); "
                           "no FMA in the exact build, so its own ceiling is half of this",
            "kernel": "render_kernel", "kernel_ms": round(kern_ms, 4),
            "algorithmic_flops_per_launch": int(flops),
            "hbm": {"bytes_per_launch": int(hbm_bytes),
                    "achieved_gbs": round(hbm_bytes / (kern_ms * 1e-3) / 1e9, 3) if kern_ms > 0 else None,
                    "peak_gbs": peaks.get("hbm_gbs"), "frac": round(hbm_bytes / (kern_ms * 1e-3) / 1e9 / peaks.get("hbm_gbs", 6650.0), 6) if kern_ms > 0 else None}},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
    }
    if others:
        line["other_configs"] = others
    if world == 1 and not args.no_cpu_baseline:
        cores = host_cores()
        v1, dt1, _ = cpu_sample(1, 8)           # faithful: render/save_image are single-threaded (tracer.rs:57-81,160-187)
        vn, dtn, _ = cpu_sample(cores, 32)      # generous: all host cores
        line["cpu_baseline"] = {"value": round(vn, 3), "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"32 of 256 spp of the full 1920x1080 frame, {cores} threads, {dtn:.1f} s",
                                "value_1thread": round(v1, 3),
                                "sample_1thread": f"8 of 256 spp of the full frame, 1 thread, {dt1:.1f} s"}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the untimed A / B / D measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
