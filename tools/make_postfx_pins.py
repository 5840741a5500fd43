"""Writes tests/golden/postfx_pins.json: sha256 of the post-fx checker's output on seeded frames (regression pins of
oracle/postfx_oracle.cpp — the reference holds no golden data for its compute filters)."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import postfx_binding as fo  # noqa: E402

cases = []
for e in range(12):
    cases.append({"w": 96, "h": 64, "seed": 100 + e, "effects": [e], "time": 1.5})
cases.append({"w": 50, "h": 20, "seed": 200, "effects": [6], "time": 0.0})
cases.append({"w": 133, "h": 71, "seed": 201, "effects": [1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 0], "time": 7.75})
cases.append({"w": 64, "h": 64, "seed": 202, "effects": [11, 9, 6, 1, 8], "time": 0.5})
for c in cases:
    f = np.random.default_rng(c["seed"]).integers(0, 256, (c["h"], c["w"], 4), dtype=np.uint8)
    c["sha256"] = hashlib.sha256(fo.chain(c["effects"], f, c["time"]).tobytes()).hexdigest()
with open(os.path.join(ROOT, "tests", "golden", "postfx_pins.json"), "w") as fh:
    json.dump({"generator": "tools/make_postfx_pins.py", "cases": cases}, fh, indent=1)
print(len(cases), "pins")
