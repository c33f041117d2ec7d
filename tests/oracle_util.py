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


def room_scene(tris=20000):
    """Textured, normal-mapped, emissive synthetic room of the bench scene's family (tools/make_scene.py, seed 1):
    generated on demand (deterministic), never committed."""
    import sys
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import make_scene
    d = os.path.join(tempfile.gettempdir(), f"b200rt_test_room_{tris}_{os.getuid()}")
    if not os.path.exists(os.path.join(d, "b200room.gltf")):
        make_scene.build(d, tris, 1)
    return os.path.join(d, "b200room.gltf")


def room_digest(gltf_path):
    """sha256 over the generated scene's geometry + the DECODED texels of its images (PNG compression may differ
    between zlib builds; the texels may not)."""
    import hashlib
    import raytracing_course_hw_b200 as rt
    h = hashlib.sha256()
    d = os.path.dirname(gltf_path)
    h.update(open(os.path.join(d, "b200room.bin"), "rb").read())
    sc = rt.load_scene(gltf_path, width=4, height=4, samples=1)
    desc = sc.desc
    for i in range(desc.n_images):
        im = desc.images[i]
        h.update(ctypes.string_at(im.rgb, im.width * im.height * 3))
    return h.hexdigest()


def room_matches_golden(gltf_path):
    p = os.path.join(GOLDEN, "hw8_room20k.sha")
    return os.path.exists(p) and open(p).read().strip() == room_digest(gltf_path)


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


_pairs = None


def pairs_traverse(layout, records, tris, rays):
    """CPU walk of the flattened device BVH (oracle/pairs_traverse.c).  Returns (hits (n, 4), per-ray (n, 2) uint32
    {boxes, triangles tested}, total boxes, total triangles)."""
    global _pairs
    if _pairs is None:
        path = os.path.join(REF_DIR, "liboracle_pairs.so")
        if not os.path.exists(path):
            from raytracing_course_hw_b200 import build
            build.build_oracle()
        _pairs = ctypes.CDLL(path)
        _pairs.pairs_traverse.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int] + [ctypes.c_void_p] * 3
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    hits = np.empty((n, 4), np.float32)
    per = np.zeros((n, 2), np.uint32)
    counts = np.zeros(2, np.uint64)
    rc = _pairs.pairs_traverse(ctypes.byref(layout), records.ctypes.data_as(ctypes.c_void_p), tris.ctypes.data_as(ctypes.c_void_p),
                               rays.ctypes.data_as(ctypes.c_void_p), n, hits.ctypes.data_as(ctypes.c_void_p),
                               per.ctypes.data_as(ctypes.c_void_p), counts.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0, f"pairs_traverse rc={rc}"
    return hits, per, int(counts[0]), int(counts[1])


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
