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
all inside the timed region).  Config C is an all-sky frame: `culled_paths` of its `rays` run no
intersection test at all (camera ray + RNG + sky + write-out), so the traversal configs (A, B, D and
the reference's built-in scenes) are reported beside it in `other_configs`, each with a CPU figure.

At N > 1 every rank traces its own 256 samples of every pixel (sample-split, WEAK scaling: the job is the
N*256-spp frame) and the partial f32 sums are combined by the fused peer-memory reduce + tone-map kernel
(or NCCL reduce as a fallback); rank 0's frame is then checked against one GPU's composition of the same
partial sums (`exchange_parity`).  `other_configs` at N > 1 carries the north-star multi-GPU configs: C strong
(256 spp total), D tile-split strong (100k spheres, 64 spp) and E sample-split (4K, 512 spp per GPU), each
with its 1-GPU time measured on rank 0 in the same run, the efficiency against it and the exchange's share.
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
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
SCENE = os.path.join(SCENES, "scene_09.json")
METRIC, UNIT = "Mrays/s at 1080p scene_09 (256 spp per GPU, depth 50)", "Mrays/s"
WORKLOAD = "scenes/scene_09.json 1920x1080 256 spp depth 50 (config C)"
CAPTURE = os.path.join(ROOT, "profiles", "r05_render_kernel_C_capture.json")


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


def config_dict(world, exchange, rays_step, t_wall, cores=None):
    """The same keys in both arms (the driver compares the key sets)."""
    return {"workload": WORKLOAD, "seed": SEED, "spp_per_gpu": SPP, "spp_total": SPP * world,
            "parallelism": f"sample-split x{world}" if cores is None else f"{cores} host threads, rows interleaved",
            "exchange": exchange,
            "l2": "256 MiB buffer written between timed steps (L2 flush)" if cores is None else "n/a (CPU arm)",
            "build": "exact (-fmad=false, IEEE div/sqrt): bit-identical to the CPU oracle" if cores is None
                     else "oracle port of the reference tracer, g++ -O3 -ffp-contract=off",
            "rays_per_step": int(rays_step), "wall_s_timed_region": round(t_wall, 4),
            # filled in by the arm they apply to; present in both so that the two lines carry the same keys
            "what_it_measures": None, "culled_paths_per_step": None, "exchange_ms": None, "exchange_parity": None, "note": None}
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




def sphere_generator():
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen", os.path.join(ROOT, "python-extras", "sphere_scene_generator.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    return gen


def oracle_scene(path=SCENE, w=W, h=H, doc=None):
    from oracle import binding as ob
    from oracle import scene_oracle as so
    d = doc if doc is not None else so.load(path)
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    return ob, objs, ob.camera_look(fr, at, up, fov, w / h)


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


def cpu_case(objs, cam, w, h, spp, nthreads, y0=0, y1=None, primary=False):
    """One bounded CPU run of the reference's algorithm (linear scan, tracer.rs:195-200) on any scene."""
    from oracle import binding as ob
    t0 = time.perf_counter()
    if primary:  # the un-jittered pinhole primary ray only: depth 0 stops every path at its first hit (tracer.rs:205)
        _, _, rays = ob.render(objs, cam, w, h, 1, max_depth=0, seed=SEED, y0=y0, y1=y1, nthreads=nthreads,
                               flags=ob.FLAG_NO_JITTER | ob.FLAG_PINHOLE, want_linear=False)
    else:
        _, _, rays = ob.render(objs, cam, w, h, spp, max_depth=DEPTH, seed=SEED, y0=y0, y1=y1, nthreads=nthreads,
                               want_linear=False)
    dt = time.perf_counter() - t0
    return {"mrays_s": round(rays / dt / 1e6, 3), "seconds": round(dt, 2), "rays": int(rays)}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot be
    built here (no cargo/rustc on either machine, un-vendored ../kopek), so this is the oracle port, all host
    threads.  A step = 32 of the 256 spp of the full frame (>= 0.3 s of work per step, so thread start-up does not
    show; throughput is spp-independent)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = host_cores()
    spp = 32
    for _ in range(max(1, min(args.warmup, 2))):
        cpu_sample(cores, 4)
    t_tot, r_tot = 0.0, 0
    t_wall = time.perf_counter()
    for _ in range(args.steps):
        _, dt, rays = cpu_sample(cores, spp)
        t_tot += dt; r_tot += rays
    t_wall = time.perf_counter() - t_wall
    v = r_tot / t_tot / 1e6
    v1, dt1, _ = cpu_sample(1, 4)  # the reference's render / save_image are single-threaded (tracer.rs:57-81,160-187)
    sample = f"{spp} of {SPP} spp of the full {W}x{H} frame per step, {cores} threads (row-interleaved)"
    cfg = config_dict(max(1, args.gpus), None, r_tot / args.steps * SPP / spp, t_wall, cores=cores)
    cfg["note"] = "ms_per_step and rays_per_step are extrapolated to 256 spp from the bounded sample"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(v, 3), "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * t_tot / args.steps * SPP / spp, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": round(v, 3), "unit": UNIT, "cores": cores, "kind": "port", "sample": sample,
                         "value_1thread": round(v1, 3), "sample_1thread": f"4 of {SPP} spp of the full frame, 1 thread, {dt1:.1f} s"},
        "e2e": {"value": round(v, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def _d2h(host_np, d_ptr, nbytes):
    from cuda.bindings import runtime as rt  # cuda-python
    rt.cudaMemcpy(host_np.ctypes.data, d_ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)


def frame_bytes(d_ptr, nbytes):
    out = np.empty(nbytes, np.uint8)
    _d2h(out, d_ptr, nbytes)
    return out


class Rig:
    """What every measurement of a rank shares: device, launch stream, L2-flush buffer, process group."""

    def __init__(self):
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=self.dev)
            self.dist = dist
        self.ts = torch.cuda.Stream(device=self.dev)  # launch stream; the CUDA events below are recorded on it
        torch.cuda.set_stream(self.ts)
        self.stream = self.ts.cuda_stream
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2, written between timed steps

    def sync(self):
        if self.dist:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x):
        t = self.torch.tensor([float(x)], dtype=self.torch.float64, device=self.dev)
        if self.dist:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def time_steps(self, fn, steps, collective=True):
        """ms per step: CUDA events on the launch stream around each step, L2 flushed (untimed) in between; the max
        over ranks when `collective`.  Returns (ms_per_step, wall seconds of the region)."""
        torch = self.torch
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        self.sync() if collective else torch.cuda.synchronize()
        t_wall = time.perf_counter()
        for a, b in ev:
            self.flush.fill_(1)
            a.record()
            fn()
            b.record()
        self.sync() if collective else torch.cuda.synchronize()
        t_wall = time.perf_counter() - t_wall
        ms = sum(a.elapsed_time(b) for a, b in ev) / steps
        return (self.max_over_ranks(ms) if collective else ms), t_wall


def make_exchange(rig, model, kind):
    """FusedExchange / TileExchange over CUDA IPC on every rank, or None on all ranks (and why on stderr)."""
    ex = None
    try:
        from forma_b200 import distributed as fd
        ex = (fd.FusedExchange if kind == "sample" else fd.TileExchange)(model, rig.dist, rig.dev)
    except Exception as e:  # noqa: BLE001 — reported, and the caller falls back or skips
        sys.stderr.write(f"[bench] rank {rig.rank}: IPC exchange unavailable ({e})\n")
    ok = rig.torch.tensor([1 if ex else 0], device=rig.dev)
    rig.dist.all_reduce(ok, op=rig.dist.ReduceOp.MIN)
    return ex if ok.item() else None


def sample_split_parity(rig, ref_model, ex_rgb_ptr, spp_rank, w, h):
    """Rank 0: its frame (after the exchange) against ONE GPU's composition of the same per-rank sample ranges summed
    in rank order.  The composition is bit-exact by construction, so anything but byte equality is 'bad'."""
    torch = rig.torch
    npix = w * h
    got = frame_bytes(ex_rgb_ptr, npix * 3)
    parts = [torch.empty((npix, 4), dtype=torch.float32, device=rig.dev) for _ in range(rig.world)]
    ref = torch.empty(npix * 3, dtype=torch.uint8, device=rig.dev)
    for r, p in enumerate(parts):
        ref_model.render_device(spp_rank, r * spp_rank, DEPTH, None, p.data_ptr(), rig.stream)
    ref_model.reduce_tonemap_device([p.data_ptr() for p in parts], spp_rank * rig.world, 0, npix, ref.data_ptr(), None, rig.stream)
    rig.ts.synchronize()
    return "ok" if np.array_equal(got, ref.cpu().numpy()) else "bad"


def run_ours(args):
    import forma_b200

    rig = Rig()
    torch, world, rank, local, dev = rig.torch, rig.world, rig.rank, rig.local, rig.dev
    npix = W * H
    scene_txt = open(SCENE).read()
    steps, warm = args.steps, max(args.warmup, 3)

    m = forma_b200.create_model(W, H, device=local, seed=SEED)
    m.scene.load_json(scene_txt)
    m.scene.camera.use_json()
    d_rgb = torch.empty(npix * 3, dtype=torch.uint8, device=dev)

    ex = make_exchange(rig, m, "sample") if world > 1 else None
    d_sum = torch.empty((npix, 4), dtype=torch.float32, device=dev) if world > 1 and not ex else None
    launches = [0]

    def step():
        """One frame on this rank."""
        if world == 1:
            m.render_device(SPP, 0, DEPTH, d_rgb.data_ptr(), None, rig.stream)
            launches[0] += 1
        elif ex:
            launches[0] += ex.frame(SPP, DEPTH, rig.stream)
        else:
            m.render_device(SPP, rank * SPP, DEPTH, None, d_sum.data_ptr(), rig.stream)
            rig.dist.reduce(d_sum, dst=0)
            if rank == 0:
                m.tonemap_device(d_sum.data_ptr(), SPP * world, d_rgb.data_ptr(), 0, npix, rig.stream)
            launches[0] += 2 if rank == 0 else 1

    clk = ClockSampler(local, enabled=rank == 0).start()  # rank 0 samples its own GPU
    for _ in range(warm):
        step()
    rig.sync()
    launches[0] = 0
    with clk:
        ms_per_step, t_wall = rig.time_steps(step, steps)
    st = m.stats()  # finalises the last device-resident render: ray counter + kernel events (this rank's trace kernel)
    rays_step = rig.sum_over_ranks(st.rays)
    culled_step = rig.sum_over_ranks(st.culled_paths)
    kern_ms = float(st.kernel_ms)
    value = rays_step / (ms_per_step * 1e-3) / 1e6
    exchange_ms = rig.max_over_ranks(ms_per_step - kern_ms) if world > 1 else None

    # ---- e2e: the host-facing call, host buffers, copies inside the timed region ----
    hp = m.host_alloc(npix * 3)  # pinned, through the C ABI (forma_rt_host_alloc): what a Rust caller's `pixels` would be

    def e2e_step():
        m.scene.load_json(scene_txt)      # parse + flatten + H2D of the packed scene
        m.scene.camera.use_json()
        if world == 1:
            m.render(SPP, max_depth=DEPTH, rgb_out=hp.reshape(H, W, 3))
            launches[0] += 1
        else:
            step()
            if rank == 0:
                # rank 0's stream ends with the wait for every peer's part of the frame; the copy below runs on
                # the legacy default stream, which does not order itself behind torch's non-blocking streams
                rig.ts.synchronize()
                _d2h(hp, ex.rgb0 if ex else d_rgb.data_ptr(), npix * 3)

    for _ in range(2):
        e2e_step()
    rig.sync()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_step()
    rig.sync()
    dt = rig.max_over_ranks(time.perf_counter() - t0)
    stx = m.stats()
    h2d = 16 * (stx.n_spheres + 3 * stx.n_triangles + 3 * stx.n_planes + stx.n_objects) + 12 * stx.n_objects + 20
    e2e = {"value": round(rays_step * steps / dt / 1e6, 1), "unit": UNIT, "ms_per_step": round(1e3 * dt / steps, 3),
           "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": npix * 3 + 16,
           "api": "forma_rt_load_json + forma_rt_camera_json + forma_rt_render into a forma_rt_host_alloc'ed (pinned) RGB8 buffer"
                  if world == 1 else
                  "forma_rt_load_json + forma_rt_camera_json + forma_rt_render_device + flag-ordered exchange on every "
                  "rank, then rank 0 waits for the frame and copies it to a pinned host RGB8 buffer"}

    parity = None
    if world > 1:
        step()
        rig.sync()
        if rank == 0:
            if ex:
                parity = sample_split_parity(rig, m, ex.rgb0, SPP, W, H)
            else:
                parity = "unchecked (NCCL fallback)"
        rig.sync()

    others = None
    if not args.no_other_configs:
        others = other_configs(rig) if world == 1 else multi_gpu_configs(rig, m, ex, d_rgb, max(3, min(steps, 10)))
    if rank == 0:
        emit(args, world, value, ms_per_step, rays_step, culled_step, kern_ms, st, clk.summary(), e2e, launches[0], t_wall,
             ex.mode if ex else ("NCCL reduce + tonemap" if world > 1 else None), others, exchange_ms, parity)
    if rig.dist:
        rig.dist.barrier()
        rig.dist.destroy_process_group()


def other_configs(rig):
    """BASELINE.json's other single-GPU configs and the reference's own scenes through the host-facing API (outside
    the headline's timed region; parity for each is in tests/).  Per case: kernel time from the library's CUDA events
    (best of 3), e2e = wall time of the host-facing render call into a host buffer (D2H inside), the CPU restatement of
    the reference's algorithm on a bounded sample (what the sample was is stated), and a roofline fraction: SURVEY
    §8(d)'s brute-force flop model where the GPU scans, the L2 traffic of the walk where it uses the LBVH."""
    import forma_b200
    from oracle import binding as ob
    from oracle import scene_oracle as so
    peaks, _ = measured_peaks()
    cores = host_cores()
    out = {}

    def best(m, fn, reps=3):
        b, wall = None, None
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            dt = (time.perf_counter() - t0) * 1e3
            st = m.stats()
            if b is None or st.kernel_ms < b.kernel_ms:
                b = st
            wall = dt if wall is None else min(wall, dt)
        return b, wall

    def gpu_fields(st, wall):
        return {"kernel_ms": round(st.kernel_ms, 3), "e2e_ms": round(wall, 3), "rays": int(st.rays),
                "mrays_s": round(st.rays / st.kernel_ms / 1e3, 1), "e2e_mrays_s": round(st.rays / wall / 1e3, 1),
                "gpu_launches": int(st.gpu_launches)}

    def scan_roofline(st):
        fl = flops_model(st.rays, st.paths, st.n_spheres, st.n_triangles, st.n_planes)
        peak = st.sm_count * 128 * 2 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6 / 1e12
        ach = fl / (st.kernel_ms * 1e-3) / 1e12
        return {"bound": "fp32-issue", "achieved": round(ach, 3), "peak": round(peak, 2), "unit": "TFLOP/s",
                "frac": round(ach / peak, 4), "model": "SURVEY 8(d): 23/sphere + 45/triangle + 20/plane per ray, 60/path, 40/hit"}

    def walk_roofline(m, st, spp):
        ws = m.walk_stats(spp, max_depth=DEPTH)
        l2_bytes = ws.node_visits * ws.node_bytes + ws.leaf_tests * ws.leaf_bytes
        gbs = l2_bytes / (st.kernel_ms * 1e-3) / 1e9
        return {"bound": "l2", "node_visits": int(ws.node_visits), "leaf_tests": int(ws.leaf_tests),
                "visits_per_ray": round(ws.node_visits / max(1, ws.rays), 2),
                "leaf_tests_per_ray": round(ws.leaf_tests / max(1, ws.rays), 2), "bytes": int(l2_bytes),
                "achieved_gbs": round(gbs, 1), "peak_gbs": round(float(ws.l2_read_gbs), 1),
                "frac": round(gbs / ws.l2_read_gbs, 4) if ws.l2_read_gbs > 0 else None,
                "peak_source": "forma_rt_walk_stats_get: 32 MiB streamed out of L2, best of 5 warmed rounds of 8 passes, CUDA events"}

    # ---- the reference's own workload: create_model + save_image(model, 50) at 600x600 (frontend/macroquad.rs:5-7,60)
    # on get_simple_scene / get_objects (scenes.rs:6-39,115-156), after the frontend's first update() (eye at (5,0,0))
    for which, name in ((0, "get_simple_scene"), (2, "get_objects")):
        for (w, h, spp) in ((600, 600, 50), (1920, 1080, 64)):
            m = forma_b200.create_model(w, h, device=rig.local, seed=SEED)
            m.scene.load_builtin(which)
            if (w, h) == (600, 600):
                m.scene.camera.orbit((0.0, 0.0, 0.0))
            st, wall = best(m, lambda: m.render(spp, max_depth=DEPTH))
            key = f"{name} {w}x{h} {spp} spp" + (" (the reference workload: save_image(model, 50) after update())" if w == 600 else " (Camera::new)")
            out[key] = {**gpu_fields(st, wall), "roofline": scan_roofline(st)}
            if w == 600:
                cam = ob.camera_new(w, h)
                ob.camera_orbit(cam, (0.0, 0.0, 0.0))
                objs = ob.builtin_scene(which)
                out[key]["cpu_baseline"] = {"all_cores": cpu_case(objs, cam, w, h, spp, cores), "one_thread": cpu_case(objs, cam, w, h, 5, 1),
                                            "cores": cores, "kind": "port", "sample": "the full frame at 50 spp on all cores; 5 of 50 spp on one thread"}
                # tracer::update: one 1-spp frame into the caller's buffer (f-1)
                lat = []
                for _ in range(30):
                    t0 = time.perf_counter()
                    forma_b200.update(m, 0, 0.016)
                    lat.append((time.perf_counter() - t0) * 1e3)
                out[key]["update_1spp"] = {"wall_ms": round(float(np.median(lat[5:])), 3), "kernel_ms": round(m.stats().kernel_ms, 3)}
            m.close()

    # ---- A: scenes/scene_01.json at the reference's default size (BASELINE.json config 1)
    d = so.load(os.path.join(SCENES, "scene_01.json"))
    m = forma_b200.create_model(600, 600, device=rig.local, seed=SEED)
    m.scene.load_json(json.dumps(d)); m.scene.camera.use_json()
    st, wall = best(m, lambda: m.render(50, max_depth=DEPTH))
    _, objs, cam = oracle_scene(w=600, h=600, doc=d)
    out["A scene_01 600x600 50 spp (JSON camera)"] = {
        **gpu_fields(st, wall), "bvh_nodes": int(st.bvh_nodes), "roofline": walk_roofline(m, st, 50),
        "cpu_baseline": {"all_cores": cpu_case(objs, cam, 600, 600, 4, cores), "cores": cores, "kind": "port",
                         "sample": "4 of 50 spp of the full 600x600 frame, linear scan over 363 primitives"}}
    m.close()

    # ---- B: scenes/scene_05.json primary hit id / t dump
    d = so.load(os.path.join(SCENES, "scene_05.json"))
    m = forma_b200.create_model(W, H, device=rig.local, seed=SEED)
    m.scene.load_json(json.dumps(d)); m.scene.camera.use_json()
    st, wall = best(m, lambda: m.primary_hits())
    _, objs, cam = oracle_scene(doc=d)
    out["B scene_05 1920x1080 primary hit id/t dump"] = {
        **gpu_fields(st, wall), "bvh_nodes": int(st.bvh_nodes),
        "cpu_baseline": {"all_cores": cpu_case(objs, cam, W, H, 1, cores, primary=True), "cores": cores, "kind": "port",
                         "sample": "the full 1920x1080 set of primary rays (closest hit, no shading), linear scan over 827 primitives"}}

    # ---- D: 100 000 spheres + ground, LBVH
    d = sphere_generator().generate(100_000)
    txt = json.dumps(d).encode("utf-8")  # 37 MB of JSON; producing it is the generator's business, not the loader's
    t0 = time.perf_counter()
    m.scene.load_json(txt); m.scene.camera.use_json()   # parse + JSON -> RT mapping (host)
    t1 = time.perf_counter()
    m.render(1, max_depth=DEPTH)                         # pack + H2D + on-GPU LBVH build + a 1-spp frame
    load_ms, first_ms = (t1 - t0) * 1e3, (time.perf_counter() - t1) * 1e3
    st, wall = best(m, lambda: m.render(64, max_depth=DEPTH), reps=2)
    _, objs, cam = oracle_scene(w=480, h=270, doc=d)
    out["D 100k spheres 1920x1080 64 spp (LBVH)"] = {
        **gpu_fields(st, wall), "bvh_nodes": int(st.bvh_nodes), "scene_json_parse_and_mapping_ms": round(load_ms, 1), "scene_bytes": len(txt),
        "first_frame_pack_upload_lbvh_build_1spp_ms": round(first_ms, 1),
        "l2_roofline": walk_roofline(m, st, 64),
        "cpu_baseline": {"all_cores": cpu_case(objs, cam, 480, 270, 1, cores, y0=160, y1=162), "cores": cores, "kind": "port",
                         "sample": "reduced resolution and spp: rows 160-161 of a 480x270 frame of the same camera at 1 spp, "
                                   "the reference's linear scan over 100 001 spheres (O(n) per ray)"}}
    m.close()
    return out


def multi_gpu_configs(rig, m_c, ex_c, d_rgb_c, steps):
    """north_star's multi-GPU configs at this N (reported, not the headline): per config the N-GPU frame time (CUDA
    events, max over ranks), the SAME job on one GPU (rank 0 alone, the others idle at a barrier), the efficiency
    between the two, `exchange_ms` = frame time minus the slowest rank's trace kernel (flag waits + the fused
    reduce / the remote stores' completion), and a parity verdict of rank 0's frame."""
    import forma_b200
    torch, world, rank = rig.torch, rig.world, rig.rank
    out = {}

    def solo(fn):
        """ms of fn on rank 0 alone."""
        ms = 0.0
        if rank == 0:
            for _ in range(2):
                fn()
            ms, _ = rig.time_steps(fn, steps, collective=False)
        rig.sync()
        return rig.max_over_ranks(ms)

    def record(name, kind, ms_n, ms_1, rays, model, parity, note):
        kern = rig.max_over_ranks(model.stats().kernel_ms)
        eff = ms_1 / (world * ms_n) if kind == "strong" else ms_1 / ms_n
        out[name] = {"scaling": kind, "n_gpus": world, "ms": round(ms_n, 4), "ms_1gpu_same_job" if kind == "strong" else "ms_1gpu_per_gpu_share": round(ms_1, 4),
                     "efficiency": round(eff, 4), "mrays_s": round(rays / ms_n / 1e3, 1), "rays": int(rays),
                     "trace_kernel_ms_slowest_rank": round(kern, 4), "exchange_ms": round(ms_n - kern, 4),
                     "exchange_parity": parity, "note": note}

    # ---- C strong: the 256-spp frame split over the ranks' sample ranges (latency-bound: 0.55 ms of trace at N = 8)
    if ex_c and SPP % world == 0:
        spr = SPP // world
        for _ in range(3):
            ex_c.frame(spr, DEPTH, rig.stream)
        ms_n, _ = rig.time_steps(lambda: ex_c.frame(spr, DEPTH, rig.stream), steps)
        rays = rig.sum_over_ranks(m_c.stats().rays)
        parity = None
        if rank == 0:
            rig.ts.synchronize()
            parity = sample_split_parity(rig, m_c, ex_c.rgb0, spr, W, H)
        rig.sync()
        ms_1 = solo(lambda: m_c.render_device(SPP, 0, DEPTH, d_rgb_c.data_ptr(), None, rig.stream))
        # stats() now holds rank 0's solo render; the N-rank kernel time is taken from one more split frame
        ex_c.frame(spr, DEPTH, rig.stream); rig.sync()
        record("C strong: scene_09 1080p, 256 spp TOTAL, sample-split", "strong", ms_n, ms_1, rays, m_c, parity,
               f"{spr} spp per GPU; fused peer-memory reduce + tone-map, flag-ordered")

    # ---- C strong, tile-interleaved: the same 256-spp frame with the OTHER partition north_star names - every rank traces
    # all samples of its own 32x4 tiles and stores RGB8 into rank 0's frame; no reduce, bit-exact against one GPU
    m = forma_b200.create_model(W, H, device=rig.local, seed=SEED)
    m.scene.load_json(open(SCENE).read()); m.scene.camera.use_json()
    tx = make_exchange(rig, m, "tile")
    if tx:
        for _ in range(3):
            tx.frame(SPP, DEPTH, rig.stream)
        ms_n, _ = rig.time_steps(lambda: tx.frame(SPP, DEPTH, rig.stream), steps)
        rays = rig.sum_over_ranks(m.stats().rays)
        parity = None
        if rank == 0:
            rig.ts.synchronize()
            got = frame_bytes(tx.rgb0, W * H * 3)
            m_c.render_device(SPP, 0, DEPTH, d_rgb_c.data_ptr(), None, rig.stream)
            rig.ts.synchronize()
            parity = "ok" if np.array_equal(got, d_rgb_c.cpu().numpy()) else "bad"
        rig.sync()
        ms_1 = solo(lambda: m_c.render_device(SPP, 0, DEPTH, d_rgb_c.data_ptr(), None, rig.stream))
        record("C strong: scene_09 1080p, 256 spp TOTAL, tile-interleaved", "strong", ms_n, ms_1, rays, m, parity,
               "every rank traces all 256 samples of its 32x4 tiles and stores RGB8 into rank 0's frame over NVLink; no reduce")
        rig.sync()
        tx.close()
    m.close()

    # ---- E: scene_09 at 3840x2160, 512 spp per GPU, sample-split (N = 8 is BASELINE.json's 4096-spp config E)
    w, h, spp = 3840, 2160, 512
    m = forma_b200.create_model(w, h, device=rig.local, seed=SEED)
    m.scene.load_json(open(SCENE).read()); m.scene.camera.use_json()
    ex = make_exchange(rig, m, "sample")
    if ex:
        for _ in range(2):
            ex.frame(spp, DEPTH, rig.stream)
        ms_n, _ = rig.time_steps(lambda: ex.frame(spp, DEPTH, rig.stream), steps)
        rays = rig.sum_over_ranks(m.stats().rays)
        parity = None
        if rank == 0:
            rig.ts.synchronize()
            parity = sample_split_parity(rig, m, ex.rgb0, spp, w, h)
        rig.sync()
        d4k = torch.empty(w * h * 3, dtype=torch.uint8, device=rig.dev)
        ms_1 = solo(lambda: m.render_device(spp, 0, DEPTH, d4k.data_ptr(), None, rig.stream))
        ex.frame(spp, DEPTH, rig.stream); rig.sync()
        record(f"E: scene_09 3840x2160, {spp * world} spp total = 512 per GPU, sample-split", "weak", ms_n, ms_1, rays, m, parity,
               "BASELINE.json config E is N = 8 (4096 spp); at N < 8 the per-GPU share is kept at 512 spp")
        del d4k
        rig.sync()
        ex.close()
    m.close()

    # ---- D: 100 000 spheres + ground, 1080p, 64 spp, tile-interleaved (every rank builds its own LBVH)
    txt = json.dumps(sphere_generator().generate(100_000))
    m = forma_b200.create_model(W, H, device=rig.local, seed=SEED)
    m.scene.load_json(txt); m.scene.camera.use_json()
    one = None
    if rank == 0:  # the 1-GPU frame: the same job without a partition
        one = forma_b200.create_model(W, H, device=rig.local, seed=SEED)
        one.scene.load_json(txt); one.scene.camera.use_json()
    tx = make_exchange(rig, m, "tile")
    if tx:
        for _ in range(2):
            tx.frame(64, DEPTH, rig.stream)
        ms_n, _ = rig.time_steps(lambda: tx.frame(64, DEPTH, rig.stream), max(2, steps // 2))
        rays = rig.sum_over_ranks(m.stats().rays)
        d1 = torch.empty(W * H * 3, dtype=torch.uint8, device=rig.dev)
        parity = None
        if rank == 0:
            rig.ts.synchronize()
            got = frame_bytes(tx.rgb0, W * H * 3)
            one.render_device(64, 0, DEPTH, d1.data_ptr(), None, rig.stream)
            rig.ts.synchronize()
            parity = "ok" if np.array_equal(got, d1.cpu().numpy()) else "bad"  # bit-exact by design: same sample order per pixel
        rig.sync()
        ms_1 = solo(lambda: one.render_device(64, 0, DEPTH, d1.data_ptr(), None, rig.stream))
        record("D strong: 100k spheres 1080p 64 spp, tile-interleaved", "strong", ms_n, ms_1, rays, m, parity,
               "32x4-pixel tiles dealt round-robin; every rank stores RGB8 into rank 0's frame over NVLink; no reduce")
        rig.sync()
        tx.close()
    if one:
        one.close()
    m.close()
    return out


def static_capture(kern_ms):
    """Counters that only a profiler can give (DRAM traffic, executed warp instructions) come from ONE committed ncu
    capture of this kernel on this config (profiles/r05_render_kernel_C_capture.json, written by
    tools/make_capture_json.py with the commit it was taken at).  They are reported only while the live kernel time is
    within 5 % of the capture's; otherwise null: a stale capture must not pose as a measurement."""
    try:
        with open(CAPTURE) as f:
            cap = json.load(f)
    except (OSError, ValueError):
        return None, "no static capture committed"
    src = f"static capture {cap.get('commit', '?')} ({os.path.basename(CAPTURE)}: ncu --set full, {cap.get('kernel_ms')} ms)"
    if not kern_ms or abs(kern_ms - cap["kernel_ms"]) > 0.05 * cap["kernel_ms"]:
        return None, src + f" - STALE: live kernel {kern_ms:.4f} ms differs by more than 5 %"
    return cap, src


def emit(args, world, value, ms_per_step, rays_step, culled_step, kern_ms, st, clocks, e2e, launches, t_wall, exchange,
         others=None, exchange_ms=None, parity=None):
    peaks, src = measured_peaks()
    paths_rank = W * H * SPP
    rays_rank = rays_step / world
    flops = flops_model(rays_rank, paths_rank, st.n_spheres, st.n_triangles, st.n_planes)
    # what the kernel executes after the beam cull: culled paths run no primitive test (60 flop: jitter, lens, camera, sky, add)
    culled_rank = culled_step / world
    flops_exec = flops_model(rays_rank - culled_rank, paths_rank - culled_rank, st.n_spheres, st.n_triangles, st.n_planes) + 60 * culled_rank
    sm_mhz = float(peaks.get("sm_max_mhz", 1965.0))
    fp32_peak = st.sm_count * 128 * 2 * sm_mhz * 1e6 / 1e12  # TFLOP/s with FMA at the max SM clock
    achieved = flops / (kern_ms * 1e-3) / 1e12 if kern_ms > 0 else None
    hbm_bytes = 16 * (st.n_spheres + 3 * st.n_triangles + 3 * st.n_planes) + W * H * 3
    cap, cap_src = static_capture(kern_ms)
    issue_frac = None
    if cap:  # warp instructions of one launch / issue slots of the launch: SMs x 4 schedulers x clock x time
        issue_frac = round(cap["inst_executed"] / (st.sm_count * 4 * sm_mhz * 1e6 * kern_ms * 1e-3), 4)
    cfg = config_dict(world, exchange, rays_step, t_wall)
    cfg["what_it_measures"] = ("config C is a ray-generation / sky-bound frame: culled_paths of rays_per_step are paths of pixels whose "
                               "whole primary beam misses the scene's bounding sphere and run no intersection test; the traversal "
                               "numbers are in other_configs")
    cfg["culled_paths_per_step"] = int(culled_step)
    cfg["exchange_ms"] = round(exchange_ms, 4) if exchange_ms is not None else None
    cfg["exchange_parity"] = parity
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg,
        "roofline": {
            "bound": "fp32-issue",  # neither HBM nor tensor: SURVEY §8(d); see DESIGN.md "Roofline"
            "achieved": round(achieved, 3) if achieved else None, "peak": round(fp32_peak, 2), "unit": "TFLOP/s",
            "frac": round(achieved / fp32_peak, 4) if achieved else None,
            "frac_executed": round(flops_exec / (kern_ms * 1e-3) / 1e12 / fp32_peak, 4) if kern_ms > 0 else None,
            "issue_frac": issue_frac,
            "traffic": cap["dram_bytes"] if cap and world == 1 else None,
            "issue_slots": {"active_pct": cap["issue_active_pct"], "threads_per_instruction": cap["threads_per_inst"],
                            "warp_instr_per_launch": cap["inst_executed"]} if cap else None,
            "source": cap_src,
            "peak_source": f"{st.sm_count} SMs x 128 lanes x 2 x {sm_mhz:.0f} MHz (sm_max_mhz of MEASURED_PEAKS.json, {src}); "
                           "no FMA in the exact build, so its own ceiling is half of this",
            "how_to_read": "frac = SURVEY 8(d)'s flop model (counts 4 triangle tests per ray the beam cull never runs); frac_executed = "
                           "the flops of what is executed; issue_frac = warp instructions / issue slots, the binding figure of this "
                           "issue-bound kernel (the counter RNG and the rejection sampling are instructions, not model flops)",
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
    ap.add_argument("--no-other-configs", action="store_true", help="skip the A / B / D / built-in-scene (N = 1) or C-strong / D / E (N > 1) measurements")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
