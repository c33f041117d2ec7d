"""Host-only check of the binned-SAH builder that runs behind the C ABI (csrc/cuda/bvh_rebuild.h):
tests/native/bvh_rebuild_check.cpp builds over seeded triangle soups -- serial path, parallel path, clusters of
identical centroids -- and verifies the invariants the flattener and the kernels rely on (permutation, every
triangle in exactly one leaf of at most 15, boxes nested, children after parents, unreferenced slots empty)."""
import os, shutil, subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_bvh_rebuild_invariants(tmp_path):
    gxx = os.environ.get("B200RT_CXX", "/usr/bin/g++")
    if not os.path.exists(gxx):
        gxx = shutil.which("g++")
    exe = str(tmp_path / "bvh_check")
    subprocess.run([gxx, "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "include"),
                    "-I", os.path.join(ROOT, "raytracing_course_hw_b200", "csrc", "cuda"),
                    os.path.join(ROOT, "tests", "native", "bvh_rebuild_check.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0, r.stdout
    assert r.stdout.count(" ok") == 7      # six soups + the depth-capped comb


def test_host_pool_stress(tmp_path):
    """The persistent worker pool every parallel section of the upload runs on: 8000 short jobs from two caller
    threads, each task exactly once (a lost wake-up here would hang an upload)."""
    gxx = os.environ.get("B200RT_CXX", "/usr/bin/g++")
    if not os.path.exists(gxx):
        gxx = shutil.which("g++")
    exe = str(tmp_path / "pool_check")
    subprocess.run([gxx, "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "include"),
                    "-I", os.path.join(ROOT, "raytracing_course_hw_b200", "csrc", "cuda"),
                    os.path.join(ROOT, "tests", "native", "host_pool_check.cpp"), "-o", exe], check=True)
    r = subprocess.run([exe], stdout=subprocess.PIPE, text=True, timeout=120)
    assert r.returncode == 0 and "host pool ok" in r.stdout, r.stdout
