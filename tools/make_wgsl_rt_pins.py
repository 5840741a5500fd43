"""Writes tests/golden/wgsl_rt_pins.json: sha256 of the fragment-shader checker's frames (regression pins of
oracle/wgsl_rt_oracle.cpp — the reference holds no golden image of rt.wgsl / rt2.wgsl)."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import wgsl_rt_binding as wb  # noqa: E402

cases = [
    {"shader": 1, "w": 1600, "h": 1200, "time": 0.0, "srgb": True, "resolution": None},
    {"shader": 1, "w": 320, "h": 240, "time": 0.0, "srgb": False, "resolution": [320, 240]},
    {"shader": 1, "w": 333, "h": 77, "time": 0.0, "srgb": True, "resolution": [640, 480]},
    {"shader": 1, "w": 1, "h": 1, "time": 0.0, "srgb": False, "resolution": [1, 1]},
    {"shader": 2, "w": 1600, "h": 1200, "time": 0.0, "srgb": True, "resolution": None},
    {"shader": 2, "w": 1600, "h": 1200, "time": 12.5, "srgb": False, "resolution": None},
    {"shader": 2, "w": 320, "h": 240, "time": 3.0, "srgb": False, "resolution": [320, 240]},
    {"shader": 2, "w": 123, "h": 45, "time": -40.0, "srgb": True, "resolution": [200, 100]},
    {"shader": 2, "w": 1, "h": 1, "time": 1.0, "srgb": False, "resolution": [1, 1]},
]
for c in cases:
    u8, lin, rays = wb.render(c["shader"], c["w"], c["h"], time=c["time"], srgb=c["srgb"],
                              resolution=tuple(c["resolution"]) if c["resolution"] else None, threads=8)
    c["sha256_linear"] = hashlib.sha256(lin.tobytes()).hexdigest()
    c["sha256_rgba8"] = hashlib.sha256(u8.tobytes()).hexdigest()
    c["rays"] = rays
with open(os.path.join(ROOT, "tests", "golden", "wgsl_rt_pins.json"), "w") as fh:
    json.dump({"generator": "tools/make_wgsl_rt_pins.py", "cases": cases}, fh, indent=1)
print(len(cases), "pins")
