"""The compiled-code consumers of the C ABI (examples/): include/forma_rt.hpp, the C++17 mirror of the reference's
tracer module, driven by a headless copy of frontend/macroquad.rs's loop, and a plain-C11 check of the three public
headers.  CPU: both build, the C program's read-backs equal the Python mirror's, the front-end fails loudly without a
device.  GPU: the front-end's frames and PNG equal what the Python mirror (and, through it, the oracle) produces."""
import os
import re
import subprocess

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from forma_b200 import frag

from conftest import ROOT, SCENES

EX = os.path.join(ROOT, "examples")


@pytest.fixture(scope="module")
def built():
    subprocess.check_call(["make", "-C", EX, "-s"])
    return os.path.join(EX, "build")


def fnv1a(b: bytes) -> int:
    h = 1469598103934665603
    for x in b:
        h = ((h ^ x) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def test_plain_c_program_reads_back_what_the_python_mirror_reads(built):
    out = subprocess.run([os.path.join(built, "abi_check")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    lines = out.stdout.splitlines()
    assert lines[0] == "abi 3"
    m = forma_b200.create_model(64, 48, flags=_lib.HOST_ONLY)
    for which in range(3):
        m.scene.load_builtin(which)
        assert lines[1 + which] == f"builtin {which}: {len(m.scene.objects)} objects"
    m.scene.camera.orbit((0.0, 0.0, 0.016))
    cam = m.scene.camera.get()
    got = [np.float32(x) for x in re.findall(r"[-+0-9.e]+", lines[4].replace("camera", ""))]
    assert got[:3] == [np.float32(v) for v in cam.position] and got[3] == np.float32(cam.radius)
    assert "FORMA_RT_ECUDA" in lines[5] and lines[6] == "effect 6 = flipaxis"
    light = [np.float32(x) for x in lines[7].split()[1:]]
    assert light == list(frag.light_direction(3.0))


def test_headers_are_self_contained_c_and_cxx(built, tmp_path):
    for hdr in ("forma_rt.h", "forma_fx.h", "forma_frag.h"):
        src = tmp_path / "one.c"
        src.write_text(f'#include "{hdr}"\nint main(void) {{ return 0; }}\n')
        subprocess.check_call(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"),
                               "-fsyntax-only", str(src)])
    src = tmp_path / "one.cpp"
    src.write_text('#include "forma_rt.hpp"\n#include "forma_fx.h"\n#include "forma_frag.h"\nint main() { return 0; }\n')
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"),
                           "-fsyntax-only", str(src)])


def test_front_end_fails_loudly_without_a_device(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    out = subprocess.run([os.path.join(built, "frontend_headless"), "32", "24", "1"], capture_output=True, text=True)
    assert out.returncode == 1 and "no CPU fallback" in out.stderr and "frame" not in out.stdout


@pytest.mark.gpu
def test_front_end_frames_equal_the_python_mirror(built, tmp_path):
    from PIL import Image
    w, h, frames, keys, dt, spp, seed = 160, 120, 3, 0x0A, 0.05, 4, 5
    png = tmp_path / "basic.png"
    out = subprocess.run([os.path.join(built, "frontend_headless"), str(w), str(h), str(frames), "--keys", hex(keys),
                          "--dt", str(dt), "--png", str(png), "--spp", str(spp), "--seed", str(seed)],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    hashes = [int(x, 16) for x in re.findall(r"fnv1a ([0-9a-f]{16})", out.stdout)]
    m = forma_b200.create_model(w, h, seed=seed)
    for f in range(frames):
        m.scene.objects[0].translate((-1.0, 0.0, 0.0))
        m.scene.objects[0].rotate((-1.0, 0.25 * f, 0.0))
        forma_b200.update(m, keys, dt)
        assert fnv1a(m.pixels.tobytes()) == hashes[f], f
    rgb, _ = m.render(spp)
    assert np.array_equal(np.asarray(Image.open(png).convert("RGB")), rgb)


@pytest.mark.gpu
def test_front_end_on_a_json_scene(built, tmp_path):
    from PIL import Image
    w, h, spp = 192, 108, 2
    scene = os.path.join(SCENES, "scene_01.json")
    png = tmp_path / "s.png"
    out = subprocess.run([os.path.join(built, "frontend_headless"), str(w), str(h), "1", "--scene", scene, "--png",
                          str(png), "--spp", str(spp)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    m = forma_b200.create_model(w, h, seed=0)
    m.scene.load_json(open(scene).read())
    m.scene.camera.use_json()
    forma_b200.update(m, 0, 0.016)
    assert fnv1a(m.pixels.tobytes()) == int(re.search(r"fnv1a ([0-9a-f]{16})", out.stdout).group(1), 16)
    rgb, _ = m.render(spp)
    assert np.array_equal(np.asarray(Image.open(png).convert("RGB")), rgb)
    # a JSON the loader rejects: the reference's unwrap panics, the mirror throws and the program exits 1
    bad = tmp_path / "bad.json"
    bad.write_text("{ not json")
    out = subprocess.run([os.path.join(built, "frontend_headless"), "8", "8", "1", "--scene", str(bad)],
                         capture_output=True, text=True)
    assert out.returncode == 1 and "forma_rt error -2" in out.stderr
