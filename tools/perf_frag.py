"""Times the fragment-shader ray tracers (frag.cu) on the GPU next to the CPU restatement on the host cores.

python tools/perf_frag.py [--steps 20] [--cpu]  ->  one JSON line per (shader, frame size).
Device time = CUDA events around the kernel (forma_frag_stats.kernel_ms); e2e = forma_frag_render into a host RGBA8
frame (kernel + D2H).  An L2 flush (256 MiB memset) runs between timed frames.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--cpu", action="store_true", help="also time the CPU restatement (all host threads)")
    ap.add_argument("--only", type=int, default=0)
    a = ap.parse_args()
    import torch

    from forma_b200.frag import Shader, State

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    for shader in (Shader.RT, Shader.RT2):
        if a.only and a.only != int(shader):
            continue
        for (w, h) in ((1600, 1200), (3840, 2160)):
            res = None if (w, h) == (1600, 1200) else (w, h)
            d_u8 = torch.empty((h, w, 4), dtype=torch.uint8, device="cuda:0")
            host = np.empty((h, w, 4), np.uint8)
            with State(w, h, shader=shader, srgb_surface=True) as st:
                kms, ems = [], []
                for i in range(a.warmup + a.steps):
                    flush.zero_()
                    torch.cuda.synchronize()
                    st.render_device(d_u8.data_ptr(), None, 3.0, resolution=res)
                    st.synchronize()
                    s = st.stats()
                    if i >= a.warmup:
                        kms.append(s.kernel_ms)
                for i in range(a.warmup + a.steps):
                    t0 = time.perf_counter()
                    st.render(3.0, resolution=res)
                    if i >= a.warmup:
                        ems.append((time.perf_counter() - t0) * 1e3)
                rays = st.stats().rays
            out = {"shader": shader.name.lower() + ".wgsl", "w": w, "h": h, "rays": rays,
                   "kernel_ms_median": float(np.median(kms)), "kernel_ms_min": float(np.min(kms)),
                   "mrays_s": rays / np.median(kms) / 1e3, "e2e_ms_median": float(np.median(ems)),
                   "e2e_note": "forma_frag_render into a pageable host RGBA8 frame, wall clock"}
            if a.cpu:
                from oracle import wgsl_rt_binding as wb
                n = os.cpu_count() or 1
                t0 = time.perf_counter()
                wb.render(int(shader), w, h, time=3.0, srgb=True, resolution=res, threads=n)
                dt = time.perf_counter() - t0
                out["cpu_ms"] = dt * 1e3
                out["cpu_threads"] = n
                out["cpu_mrays_s"] = rays / dt / 1e6
            print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
