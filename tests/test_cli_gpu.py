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
    if rt.device_lib().b200rt_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    a, b = tmp_path / "a.ppm", tmp_path / "b.ppm"
    args = [os.path.join(ou.SCENES, "hw8/sphere_emissive.gltf"), "128", "96", "64"]
    subprocess.run([RUN] + args + [str(a)], check=True, capture_output=True)
    subprocess.run([RUN] + args + [str(b)], check=True, capture_output=True, env=dict(os.environ, B200RT_GPUS="2"))
    cmp = ou.compare_u8(ou.parse_ppm(a.read_bytes()), ou.parse_ppm(b.read_bytes()))
    assert cmp["within1"] >= 0.999, cmp     # same samples, different float summation order


def test_cli_usage_and_missing_file(tmp_path):
    assert subprocess.run([RUN]).returncode == 2
    assert subprocess.run([RUN, str(tmp_path / "nope.txt"), str(tmp_path / "o.ppm")], capture_output=True).returncode == 1
