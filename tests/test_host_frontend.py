"""Host-side logic of the product (no GPU): scene front-end, camera, ABI surface — checked against
the oracle's independent Python/C++ restatement.  Uses FORMA_RT_HOST_ONLY handles."""
import glob
import os
import re

import numpy as np
import pytest

import forma_b200
from forma_b200 import _lib
from forma_b200.tracer import FormaError
from oracle import binding as ob
from oracle import scene_oracle as so

from conftest import ROOT, SCENES, load_scene_pair

HOST = dict(flags=_lib.HOST_ONLY)


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "forma_rt.h")).read()
    declared = set(re.findall(r"FORMA_RT_API[^;]*?\b(forma_rt_\w+)\s*\(", hdr))
    assert len(declared) >= 30
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.forma_rt_abi_version() == 3


def test_no_cpu_fallback_render_fails_loudly_without_device():
    m = forma_b200.create_model(8, 8, **HOST)
    for call in (lambda: m.render(1), lambda: forma_b200.update(m, 0, 0.1), lambda: m.primary_hits(),
                 lambda: forma_b200.save_image(m, 1, "/tmp/x.png")):
        with pytest.raises(FormaError) as e:
            call()
        assert e.value.code == _lib.ECUDA and "no CPU fallback" in str(e.value)


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under fo-rma_b200/ may import, link or load it."""
    pat = re.compile(r"import\s+oracle|from\s+oracle|liboracle|oracle/|oracle\.|-loracle")
    for path in glob.glob(os.path.join(ROOT, "fo-rma_b200", "**", "*"), recursive=True):
        if os.path.isfile(path) and not path.endswith((".so", ".o", ".pyc", ".log")):
            assert not pat.search(open(path, errors="replace").read()), path


def test_create_model_matches_scene_new():  # tracer.rs:19-28, scene.rs:10-15
    m = forma_b200.create_model(600, 600, **HOST)
    assert (m.width, m.height) == (600, 600) and m.pixels.shape == (600 * 600 * 3,) and not m.pixels.any()
    assert np.array_equal(m.scene.camera.as_array(), ob.camera_new(600, 600))
    assert m.scene.object_records().tobytes() == ob.builtin_scene(0).tobytes()


@pytest.mark.parametrize("which", [0, 1, 2])
def test_builtin_scenes(which):  # scenes.rs
    m = forma_b200.create_model(64, 48, **HOST)
    m.scene.load_builtin(which)
    assert m.scene.object_records().tobytes() == ob.builtin_scene(which).tobytes()
    with pytest.raises(FormaError):
        m.scene.load_builtin(7)


@pytest.mark.parametrize("name", sorted(os.path.basename(p)[:-5] for p in glob.glob(os.path.join(SCENES, "*.json"))))
def test_json_mapping_bit_exact(name):
    w, h = 1920, 1080
    txt, objs, cam = load_scene_pair(name, w, h)
    m = forma_b200.create_model(w, h, **HOST)
    m.scene.load_json(txt)
    m.scene.camera.use_json()
    assert m.scene.object_records().tobytes() == objs.tobytes()
    assert np.array_equal(m.scene.camera.as_array().view(np.uint32), cam.view(np.uint32))


def test_scene_counts_match_survey():
    # A = 13 spheres + 350 triangles, B = 826 triangles + 1 sphere, C = 4 triangles
    for name, (ns, nt) in {"scene_01": (13, 350), "scene_05": (1, 826), "scene_09": (0, 4), "scene_07": (0, 0)}.items():
        objs = so.flatten(so.load(os.path.join(SCENES, name + ".json")))
        assert ((objs["kind"] == 0).sum(), (objs["kind"] == 2).sum()) == (ns, nt)


def test_rt_block_and_bad_json():
    base = so.load(os.path.join(SCENES, "scene_09.json"))
    base["objects"].append({"mesh": "sphere", "material": "x", "position": {"x": 1, "y": 2, "z": 3},
                            "rotation": {"x": 0, "y": 0, "z": 0, "w": 1}, "scale": {"x": 3, "y": 1, "z": 1},
                            "rt": {"material": "metal", "color": [0.9, 0.8, 0.7], "fuzz": 0.25}})
    base["objects"].append({"mesh": "no-such-mesh", "material": "x", "position": {"x": 0, "y": 0, "z": 0},
                            "rotation": {"x": 0.1, "y": 0.2, "z": 0.3, "w": 0.9}, "scale": {"x": 2, "y": 2, "z": 2},
                            "rt": {"material": "dielectric"}, "extra_key": [1, {"a": None}]})
    import json
    txt = json.dumps(base)
    m = forma_b200.create_model(32, 32, **HOST)
    m.scene.load_json(txt)
    rec = m.scene.object_records()
    assert rec.tobytes() == so.flatten(base).tobytes()
    assert rec[4]["kind"] == 0 and rec[4]["material"] == 1 and rec[4]["b"][0] == 1.5 and rec[4]["fuzz"] == 0.25
    assert (rec[5:]["kind"] == 2).all() and len(rec) == 4 + 1 + 2 and rec[5]["material"] == 2  # quad fallback
    for bad in ("{", "[]", '{"camera": {}}', txt.replace('"lights"', '"lamps"'), txt.replace('"fov"', '"fow"')):
        with pytest.raises(FormaError) as e:   # the reference panics (scene_loader.rs:4)
            m.scene.load_json(bad)
        assert e.value.code == _lib.EJSON
    assert m.scene.object_records().tobytes() == rec.tobytes()  # a failed load leaves the scene untouched


def test_camera_orbit_translate_update_keys():  # camera.rs:74-122, tracer.rs:31-52
    m = forma_b200.create_model(600, 400, **HOST)
    cam = ob.camera_new(600, 400)
    m.scene.camera.orbit((0.25, -0.5, 1.5)); ob.camera_orbit(cam, (0.25, -0.5, 1.5))
    assert np.array_equal(m.scene.camera.as_array().view(np.uint32), cam.view(np.uint32))
    m.scene.camera.translate((0.5, 0.25, -2.0)); ob.camera_translate(cam, (0.5, 0.25, -2.0))
    assert np.array_equal(m.scene.camera.as_array().view(np.uint32), cam.view(np.uint32))
    m.scene.camera.reset()
    assert np.array_equal(m.scene.camera.as_array(), ob.camera_new(600, 400))


def test_set_object_translate_rotate_semantics():  # plane.rs:62-68 assign; hitable.rs:12-13 no-op
    m = forma_b200.create_model(16, 16, **HOST)
    before = m.scene.object_records()
    m.scene.objects[0].translate((-1.0, 0.5, 0.0))
    m.scene.objects[0].rotate((-1.0, 0.0, 0.0))
    m.scene.objects[1].translate((9.0, 9.0, 9.0))   # a sphere: ignored
    m.scene.objects[1].rotate((9.0, 9.0, 9.0))
    after = m.scene.object_records()
    assert after[0]["a"].tolist() == [-1.0, 0.5, 0.0] and after[0]["b"].tolist() == [-1.0, 0.0, 0.0]
    assert after[1:].tobytes() == before[1:].tobytes()
    with pytest.raises(IndexError):
        m.scene.objects[4]


def test_write_png_roundtrip(tmp_path):
    from PIL import Image
    rgb = np.random.default_rng(0).integers(0, 256, (37, 53, 3), dtype=np.uint8)
    p = str(tmp_path / "a.png")
    _lib.check(_lib.lib().forma_rt_write_png(p.encode(), rgb.ctypes.data, 53, 37))
    assert np.array_equal(np.asarray(Image.open(p).convert("RGB")), rgb)
    rc = _lib.lib().forma_rt_write_png(str(tmp_path / "missing_dir" / "a.png").encode(), rgb.ctypes.data, 53, 37)
    assert rc == _lib.EIO  # the reference panics when ./out is missing (tracer.rs:186)


def test_srgb_encode_table_matches_color_utils():
    """color_utils::srgb_encode (color_utils.rs:49-66) is a u8 -> u8 map; the library's table against an f64
    evaluation of the same formula.  The f32 evaluation errs by < 5e-5 grey levels (a few ulp of a value <= 1, x 255)
    and no entry sits within 2e-4 of a rounding boundary, so the last-ulp freedom of an f32 powf (Rust's, glibc's)
    cannot move any of them."""
    t = forma_b200.srgb_table()
    lin = np.arange(256, dtype=np.float64) / 255.0
    srgb = np.where(lin <= 0.0031308, lin * 12.92, 1.055 * lin ** (1.0 / 2.4) - 0.055) * 255.0
    assert np.abs(srgb - np.floor(srgb) - 0.5).min() > 2e-4
    assert np.array_equal(t, np.floor(srgb + 0.5).astype(np.uint8))
    assert t[0] == 0 and t[255] == 255 and (np.diff(t.astype(int)) >= 0).all()
