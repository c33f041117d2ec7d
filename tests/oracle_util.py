"""Test-side access to the oracle (oracle/): the CPU port library and the committed golden images.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use this."""
import ctypes
import gzip
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")
SCENES = os.path.join(ROOT, "tests", "scenes")

_port = None


def port_lib():
    global _port
    if _port is None:
        path = os.path.join(REF_DIR, "liboracle_port.so")
        if not os.path.exists(path):
            from raytracing_course_hw_b200 import build
            build.build_oracle()
        _port = ctypes.CDLL(path)
        _port.oracle_render.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_uint32, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _port.oracle_resolve.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
    return _port


def load_scene(path, **kw):
    """rt.load_scene with the reference's own SAH build for glTF files (triangles and light set in the reference's
    order + both trees): what the byte-exact goldens and the same-stream comparisons were pinned with.  Tests of the
    product's default load (no trees in the description) call rt.load_scene directly."""
    import raytracing_course_hw_b200 as rt
    if str(path).lower().endswith(".gltf"):
        kw.setdefault("reference_bvh", True)
    return rt.load_scene(path, **kw)


def port_render(scene, rng_kind, seed=0, spp_begin=0, spp_end=None, want_counters=False):
    """rng_kind 0 = reference minstd stream, 1 = Philox stream of the CUDA kernels."""
    h, w = scene.height, scene.width
    if spp_end is None:
        spp_end = scene.samples
    accum = np.zeros((h, w, 3), np.float32)
    rgb = np.zeros((h, w, 3), np.uint8)
    ids = np.full((h, w), -2, np.int32)
    counters = np.zeros(6, np.uint64)
    rc = port_lib().oracle_render(scene.desc_ptr, rng_kind, seed, spp_begin, spp_end,
                                  accum.ctypes.data_as(ctypes.c_void_p), rgb.ctypes.data_as(ctypes.c_void_p),
                                  ids.ctypes.data_as(ctypes.c_void_p),
                                  counters.ctypes.data_as(ctypes.c_void_p) if want_counters else None)
    assert rc == 0, f"oracle_render rc={rc}"
    return {"accum": accum, "rgb": rgb, "ids": ids, "counters": counters}


def port_resolve(accum_sum, total_spp):
    a = np.ascontiguousarray(accum_sum, np.float32)
    out = np.zeros(a.shape, np.uint8)
    rc = port_lib().oracle_resolve(a.ctypes.data_as(ctypes.c_void_p), a.shape[0] * a.shape[1], total_spp,
                                   out.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0
    return out


def parse_ppm(data):
    parts = data.split(b"\n", 3)
    assert parts[0] == b"P6" and parts[2] == b"255"
    w, h = map(int, parts[1].split())
    return np.frombuffer(parts[3], np.uint8).reshape(h, w, 3)


def golden(name):
    with gzip.open(os.path.join(GOLDEN, name + ".ppm.gz"), "rb") as f:
        return parse_ppm(f.read())


def has_golden(name):
    return os.path.exists(os.path.join(GOLDEN, name + ".ppm.gz"))


def run_reference(binary, args, out_path):
    """Run an unmodified reference program from oracle/_ref (raises the stack limit)."""
    exe = os.path.join(REF_DIR, binary)
    cmd = ["bash", "-c", 'ulimit -s unlimited 2>/dev/null; exec "$0" "$@"', exe] + [str(a) for a in args]
    subprocess.run(cmd, check=True, stderr=subprocess.DEVNULL, stdout=subprocess.DEVNULL)
    with open(out_path, "rb") as f:
        return parse_ppm(f.read())


def inv_display(rgb8):
    """8-bit display value -> linear radiance (inverse gamma then inverse ACES fit); used to compare
    MC images in linear space when only the reference's PPM is available."""
    y = (rgb8.astype(np.float64) / 255.0) ** 2.2
    a, b, c, d, e = 2.51, 0.03, 2.43, 0.59, 0.14
    # y = (x(ax+b))/(x(cx+d)+e)  =>  (a - c y) x^2 + (b - d y) x - e y = 0
    A, B, C = a - c * y, b - d * y, -e * y
    disc = np.maximum(B * B - 4 * A * C, 0)
    return (-B + np.sqrt(disc)) / (2 * A)


def compare_u8(a, b):
    d = np.abs(a.astype(np.int32) - b.astype(np.int32)).max(axis=2)
    return {"exact": float((d == 0).mean()), "within1": float((d <= 1).mean()), "max": int(d.max())}
