"""Host library + C-ABI surface, no GPU needed."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle_util as ou
import raytracing_course_hw_b200 as rt

S = ou.SCENES
ROOT = ou.ROOT


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    return sorted(set(re.findall(r"\b(b200rt_[a-z_0-9]+)\s*\(", text)))


def test_device_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(rt.LIB_DIR, "libb200rt.so"))
    names = _declared("b200rt.h")
    assert set(rt.DEVICE_SYMBOLS) <= set(names)
    for n in names:
        assert hasattr(lib, n), n


def test_host_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(rt.LIB_DIR, "libb200rt_host.so"))
    names = [n for n in _declared("b200rt_host.h") if n.startswith("b200rt_host_")]
    assert set(rt.HOST_SYMBOLS) <= set(names)
    for n in names:
        assert hasattr(lib, n), n


def test_no_cpu_fallback_without_device():
    lib = rt.device_lib()
    if lib.b200rt_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(rt.B200rtError) as e:
        rt.Renderer()
    assert e.value.code == -2


def test_text_parser_modes_and_quirks(tmp_path, capfd):
    sc = rt.load_scene(os.path.join(S, "hw1/sample.txt"))
    assert "UNKNOWN COMMAND: YA" in capfd.readouterr().err   # hw1/sample.txt:25 junk line
    assert sc.mode == rt.MODE_HW1
    assert rt.load_scene(os.path.join(S, "hw2/sample.txt")).mode == rt.MODE_HW2
    sc4 = rt.load_scene(os.path.join(S, "hw4/practice3_3.txt"))
    assert (sc4.mode, sc4.samples, sc4.ray_depth, sc4.width) == (rt.MODE_HW4, 64, 6, 512)
    # TRIANGLE figures make it hw5: non-planes first in BVH order, planes last, FiguresMix light list
    sc5 = rt.load_scene(os.path.join(S, "hw5/mixed.txt"))
    d5 = sc5.desc
    assert (sc5.mode, d5.n_prims, d5.n_bvh_prims, d5.n_light_prims) == (rt.MODE_HW5, 36, 31, 4)
    assert d5.n_nodes >= 2 * 31 - 1 - 31 and d5.n_light_nodes >= 1
    assert rt.load_scene(os.path.join(S, "hw4/practice3_3.txt"), mode=rt.MODE_HW3).mode == rt.MODE_HW3
    # malformed ROTATION (comma) leaves the remaining fields at their defaults
    p = tmp_path / "q.txt"
    p.write_text("DIMENSIONS 4 4\nCAMERA_FOV_X 1\nNEW_PRIMITIVE\nBOX 1 1 1\nROTATION 0, 0.38 0 0.92\nCOLOR 1 0 0\nBOGUS 1\n")
    sc = rt.load_scene(str(p))
    assert "UNKNOWN COMMAND: BOGUS" in capfd.readouterr().err
    assert sc.mode == rt.MODE_HW1


def test_missing_files_fail_cleanly(tmp_path):
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(tmp_path / "nope.txt"))
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(tmp_path / "nope.gltf"), width=4, height=4, samples=1)
    bad = tmp_path / "bad.gltf"
    bad.write_text("{ not json")
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(bad), width=4, height=4, samples=1)


def test_ppm_writer_format(tmp_path):
    img = (np.arange(5 * 7 * 3) % 256).astype(np.uint8).reshape(5, 7, 3)
    out = tmp_path / "o.ppm"
    rt.write_ppm(str(out), img)
    data = out.read_bytes()
    assert data.startswith(b"P6\n7 5\n255\n")          # hw8/src/sceneio.cpp:383-385
    assert np.array_equal(ou.parse_ppm(data), img)


def test_png_decoder_matches_pil():
    PIL = pytest.importorskip("PIL.Image")
    path = os.path.join(S, "hw8/sphere_emission.png")
    ref = np.asarray(PIL.open(path).convert("RGB"))
    sc = rt.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=8, height=8, samples=1)
    d = sc.desc
    assert d.n_images == 1 and d.n_tris == 970 and d.n_light_tris == 960
    img = d.images[0]
    mine = np.ctypeslib.as_array(img.rgb, shape=(img.height, img.width, 3))
    assert (img.height, img.width) == ref.shape[:2]
    assert np.array_equal(mine, ref)
