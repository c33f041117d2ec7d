"""The drop-in process surface (run.sh + CLI) on a real GPU: same argv, same PPM bytes."""
import os
import subprocess

import numpy as np
import pytest

import oracle_util as ou
import raytracing_course_hw_b200 as rt

pytestmark = pytest.mark.gpu
RUN = os.path.join(ou.ROOT, "run.sh")


@pytest.mark.parametrize("name,path", [("hw1_sample", "hw1/sample.txt"), ("hw2_sample", "hw2/sample.txt")])
def test_cli_text_scene_bytes_equal_reference(tmp_path, name, path):
    out = tmp_path / "o.ppm"
    r = subprocess.run([RUN, os.path.join(ou.SCENES, path), str(out)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    data = out.read_bytes()
    ref = ou.golden(name)
    assert data.startswith(b"P6\n%d %d\n255\n" % (ref.shape[1], ref.shape[0]))
    cmp = ou.compare_u8(ou.parse_ppm(data), ref)
    assert cmp["within1"] >= 0.999 and cmp["exact"] >= 0.999, cmp


def test_cli_gltf_scene(tmp_path):
    out = tmp_path / "o.ppm"
    r = subprocess.run([RUN, os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), "128", "96", "256", str(out)],
                       capture_output=True, text=True, env=dict(os.environ, B200RT_STATS="1"))
    assert r.returncode == 0, r.stderr
    assert "Launching version without environment map" in r.stdout      # hw8/run.sh:7
    assert "FINISH" in r.stderr                                          # hw8/src/main.cpp:16
    img = ou.inv_display(ou.parse_ppm(out.read_bytes()))
    truth = ou.inv_display(ou.golden("hw8_sphere_emissive_128x96_spp4096"))
    assert abs(img.mean() - truth.mean()) <= 0.02 * truth.mean()


def test_cli_hw5_and_hw3_text_scenes(tmp_path):
    """Text scenes with triangles are hw5 by content; hw3 shares hw4's grammar and is picked with B200RT_MODE."""
    out = tmp_path / "o.ppm"
    src = open(os.path.join(ou.SCENES, "hw5/mixed.txt")).read().replace("SAMPLES 4", "SAMPLES 512")
    scene = tmp_path / "hw5.txt"
    scene.write_text(src)
    r = subprocess.run([RUN, str(scene), str(out)], capture_output=True, text=True, env=dict(os.environ, B200RT_STATS="1"))
    assert r.returncode == 0, r.stderr
    assert '"mode": 5' in r.stderr
    img = ou.inv_display(ou.parse_ppm(out.read_bytes()))
    truth = ou.inv_display(ou.golden("hw5_mixed_spp2048"))
    assert abs(img.mean() - truth.mean()) <= 0.02 * truth.mean()
    src = open(os.path.join(ou.GOLDEN, "hw4_p3_1_spp4096.txt")).read().replace("SAMPLES 4096", "SAMPLES 512")
    scene.write_text(src)
    r = subprocess.run([RUN, str(scene), str(out)], capture_output=True, text=True, env=dict(os.environ, B200RT_STATS="1", B200RT_MODE="hw3"))
    assert r.returncode == 0, r.stderr
    assert '"mode": 3' in r.stderr
    img = ou.inv_display(ou.parse_ppm(out.read_bytes()))
    truth = ou.inv_display(ou.golden("hw3_p3_1_spp4096"))
    assert abs(img.mean() - truth.mean()) <= 0.02 * truth.mean()


def test_cli_multi_gpu_sample_split(tmp_path):
    """B200RT_GPUS=2: two contexts behind b200rt_group (on two GPUs when the box has them, else sharing device 0 --
    the same code path: replication, sample split, fused peer reduce + resolve)."""
    a, b = tmp_path / "a.ppm", tmp_path / "b.ppm"
    args = [os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), "128", "96", "64"]
    subprocess.run([RUN] + args + [str(a)], check=True, capture_output=True)
    subprocess.run([RUN] + args + [str(b)], check=True, capture_output=True, env=dict(os.environ, B200RT_GPUS="2"))
    cmp = ou.compare_u8(ou.parse_ppm(a.read_bytes()), ou.parse_ppm(b.read_bytes()))
    assert cmp["within1"] >= 0.999, cmp     # same samples, different float summation order


def test_cli_usage_and_missing_file(tmp_path):
    assert subprocess.run([RUN]).returncode == 2
    assert subprocess.run([RUN, str(tmp_path / "nope.txt"), str(tmp_path / "o.ppm")], capture_output=True).returncode == 1


def test_cli_png_and_pfm_outputs(tmp_path):
    """Output container by file name: .png decodes (own decoder) to the PPM's bytes, .pfm holds sum / spp as floats."""
    import ctypes
    args = [os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), "64", "48", "8"]
    ppm, png, pfm = tmp_path / "o.ppm", tmp_path / "o.png", tmp_path / "o.pfm"
    for o in (ppm, png, pfm):
        subprocess.run([RUN] + args + [str(o)], check=True, capture_output=True)
    ref = ou.parse_ppm(ppm.read_bytes())
    sc = rt.load_scene(args[0], width=64, height=48, samples=8)     # only to reach the host library's PNG decoder
    sc.set_env_map(str(png))
    env = sc.desc.env_map.contents
    assert (env.width, env.height) == (64, 48)
    got = np.ctypeslib.as_array(env.rgb, shape=(48, 64, 3))
    assert np.array_equal(got, ref)
    raw = pfm.read_bytes()
    head, body = raw.split(b"\n", 3)[:3], raw.split(b"\n", 3)[3]
    assert head == [b"PF", b"64 48", b"-1.0"]
    mean = np.frombuffer(body, "<f4").reshape(48, 64, 3)[::-1]
    assert np.array_equal(ou.port_resolve(mean * 8, 8), ref) or ou.compare_u8(ou.port_resolve(mean * 8, 8), ref)["within1"] >= 0.999
