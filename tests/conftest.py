import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(GOLDEN, "scenes")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the checker (oracle) and the product library once per session (both compile on CPU)."""
    from oracle import binding

    binding.build()
    lib = os.path.join(ROOT, "fo-rma_b200", "libforma_rt.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "fo-rma_b200", "csrc")], stdout=subprocess.DEVNULL)
    yield


def scene_path(name):
    return os.path.join(SCENES, name + ".json")


def load_scene_pair(name, w, h):
    """(json text, oracle object records, oracle camera) for a fixture scene with its JSON camera."""
    from oracle import binding as ob
    from oracle import scene_oracle as so

    path = scene_path(name)
    with open(path) as f:
        txt = f.read()
    d = so.load(path)
    objs = so.flatten(d)
    fr, at, up, fov = so.json_camera(d)
    cam = ob.camera_look(fr, at, up, fov, w / h)
    return txt, objs, cam
