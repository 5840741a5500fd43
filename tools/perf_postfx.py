"""Post-fx passes on the GPU: device time per pass (CUDA events inside the library, on the handle's stream, device
frames), algorithmic bytes (8 B per texel and pass) against the measured HBM peak, and the CPU restatement timed beside
it.  Frames: 1080p / 4K stay in the 126 MB L2; 8K (132.7 MB per frame) does not.

    python tools/perf_postfx.py [--sizes 1920x1080,3840x2160,7680x4320] [--reps 20] [--no-cpu]
Prints one JSON line per (size, case)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from forma_b200.postfx import Effect, PostProcessor  # noqa: E402

CHAIN = [Effect.PIXELATE, Effect.GRAYSCALE, Effect.NOISE, Effect.INVERT_COLOR, Effect.INTERLACE]


def hbm_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        for k in ("hbm_gbs", "hbm_gbps"):
            if k in d:
                return float(d[k]), "MEASURED_PEAKS.json:" + k
        for k, v in d.items():
            if "hbm" in k.lower() and isinstance(v, (int, float)):
                return float(v), "MEASURED_PEAKS.json:" + k
    except Exception:
        pass
    return 6452.8, "B200_PROFILING.md fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1920x1080,3840x2160,7680x4320")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    peak, peak_src = hbm_peak()
    cases = [(e.name.lower(), [e]) for e in Effect] + [("fused5", CHAIN), ("unfused5", CHAIN)]
    if a.only:
        cases = [c for c in cases if c[0] in a.only.split(",")]
    for size in a.sizes.split(","):
        w, h = (int(x) for x in size.split("x"))
        f = np.random.default_rng(1).integers(0, 256, (h, w, 4), dtype=np.uint8)
        src = torch.from_numpy(f).cuda()
        dst = torch.empty_like(src)
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
        clean = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
        for name, ch in cases:
            with PostProcessor(w, h, unfused=name.startswith("unfused")) as pp:
                pp.update_effects(ch)
                ms = []
                for i in range(a.reps + 3):
                    flush.fill_(i & 255)  # write 256 MiB: evicts the frames from L2 between timed passes ...
                    clean.view(torch.int64).sum()  # ... then read 256 MiB, so that L2 holds clean lines and the timed
                    # pass does not pay for writing the flush buffer's dirty lines back
                    torch.cuda.synchronize()
                    pp.run_device(src.data_ptr(), dst.data_ptr(), 1.5)
                    pp.synchronize()
                    st = pp.stats()
                    if i >= 3:
                        ms.append(st.kernel_ms)
                t = float(np.median(ms))
                rec = {"size": size, "case": name, "effects": [e.name.lower() for e in ch], "passes": st.n_passes,
                       "kernel_ms": round(t, 4), "bytes": st.bytes_moved, "gbps": round(st.bytes_moved / t / 1e6, 1),
                       "frac_of_hbm_peak": round(st.bytes_moved / t / 1e6 / peak, 3), "peak_gbps": peak,
                       "peak_source": peak_src, "l2_flushed": True,
                       "gtexel_s": round(w * h * len(ch) / t / 1e6, 2)}
                if not a.no_cpu and size == "1920x1080":
                    from oracle import postfx_binding as fo
                    cores = os.cpu_count() or 1
                    t0 = time.perf_counter()
                    fo.chain(ch, f, 1.5, threads=cores)
                    dt = time.perf_counter() - t0
                    rec["cpu_ms"] = round(dt * 1e3, 2)
                    rec["cpu_cores"] = cores
                print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
