"""GPU parity tests proper: the CUDA path (through the C ABI) against the oracle.

 * deterministic configs (hw1, hw2): primary-hit primitive ids bit-exact, 8-bit pixels vs the
   reference's golden PPM (north_star bar: >= 99.9 % within 1 LSB; we also report exact);
 * Monte-Carlo configs (hw4, hw7, hw8):
     - sample for sample against the CPU port run on the SAME Philox stream (float tolerance), and
     - north_star's statistical bar against the reference's 4096-spp render: RMSE no worse than the
       reference's own at equal spp, mean luminance within 1 %  (display-space PPM inverted to linear,
       SURVEY 8c);
 * K5 resolve bytes against the port's resolve of the same float buffer;
 * size-independent properties at large sizes: sample ranges add up, determinism, seeds differ.
"""
import ctypes
import os

import numpy as np
import pytest

import oracle_util as ou
import raytracing_course_hw_b200 as rt

pytestmark = pytest.mark.gpu
S, G = ou.SCENES, ou.GOLDEN


@pytest.fixture(scope="module", params=["wavefront", "wavefront-counting", "megakernel"])
def renderer(request):
    """The product build (K4w wavefront, no flags), its counting instantiation (FLAG_COUNTERS: what bench.py's
    roofline bytes come from) and the K4 megakernel run the whole suite; the analytic kernels K1-K3 are the
    same under every flag."""
    flags = {"wavefront": 0, "wavefront-counting": rt.FLAG_COUNTERS,
             "megakernel": rt.FLAG_COUNTERS | rt.FLAG_MEGAKERNEL}[request.param]
    r = rt.Renderer(flags=flags)
    r.megakernel = request.param == "megakernel"
    yield r
    r.close()


BIG = os.path.join(ou.REF_DIR, "scenes")        # the reference's large example scenes (copied by oracle/build_ref.sh)


def _scene_file(name):
    """Fixture name -> path; None when an optional large scene is not on this machine."""
    if name == "room20k":
        return ou.room_scene(20000)
    if name == "bench268k":
        import sys
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        import bench
        return bench.scene_path(270000)
    if name.startswith("big:"):
        p = os.path.join(BIG, name[4:])
        return p if os.path.exists(p) else None
    return os.path.join(S, name)


def _primary_material(sc):
    """(H, W) material id of the primary hit through each pixel centre (-1 = miss), from the production traversal
    kernel (b200rt_trace_rays) and the material ids stored in the device's triangle records."""
    d, H, W = sc.desc, sc.height, sc.width
    pos, right, up, fwd = (np.array(list(v), np.float32) for v in (d.cam_pos, d.cam_right, d.cam_up, d.cam_forward))
    ty = np.tan(d.cam_fov / 2)
    tx = ty * W / H
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    nx, ny = tx * (2 * (xs + 0.5) / W - 1), ty * (2 * (ys + 0.5) / H - 1)
    dirs = nx[..., None] * right - ny[..., None] * up + fwd
    dirs /= np.linalg.norm(dirs, axis=2, keepdims=True)
    rays = np.concatenate([np.broadcast_to(pos, dirs.shape), dirs], axis=2).reshape(-1, 6)
    r = rt.Renderer()
    r.upload(sc)
    hits, _, _ = r.trace_rays(rays)
    _, _, tri = r.scene_bvh()
    r.close()
    bits = hits[:, 3].view(np.uint32)
    miss = bits == 0x7FFFFFFF
    mat = tri[:, 3].view(np.int32)[np.where(miss, 0, bits & 0x7FFFFFFF)]
    return np.where(miss, -1, mat).reshape(H, W)


def _class_means_close(acc, port, mat, rel=0.02):
    """Same Philox stream => the two images differ only where a path diverged at a tie (triangle edge, box plane);
    those differences are unbiased noise.  A SYSTEMATIC error in the shading of one material class would shift that
    class's mean: for every class covering >= 2 % of the frame the means agree within `rel` (+ a small floor)."""
    bad = []
    for m in np.unique(mat):
        sel = mat == m
        if sel.mean() < 0.02:
            continue
        a, b = float(acc[sel].mean()), float(port[sel].mean())
        if abs(a - b) > rel * b + 1e-3 * float(port.mean()):
            bad.append((int(m), float(sel.mean()), a, b))
    return bad


def _render(r, sc, b=0, e=None):
    r.upload(sc)
    r.render(b, e if e is not None else sc.samples)
    return r.download_accum()


# ---------------------------------------------------------------- deterministic configs
@pytest.mark.parametrize("name,path", [("hw1_sample", "hw1/sample.txt"), ("hw2_sample", "hw2/sample.txt")])
def test_deterministic_ids_and_pixels(renderer, name, path):
    sc = ou.load_scene(os.path.join(S, path))
    acc = _render(renderer, sc)
    img = renderer.resolve(1)
    ids = renderer.primary_ids()
    port = ou.port_render(sc, 0)
    assert np.array_equal(ids, port["ids"]), "primary-hit primitive ids must be bit-exact"
    cmp = ou.compare_u8(img, ou.golden(name))
    assert cmp["within1"] >= 0.999, cmp
    assert cmp["exact"] >= 0.999, cmp
    # radiance before tonemap: same float operation order => bit-identical on >= 99.9 % of pixels
    same = (acc.view(np.uint32) == port["accum"].view(np.uint32)).all(axis=2).mean()
    assert same >= 0.999, same


def test_deterministic_upres_ids(renderer):
    """Synthetic up-res copy: 1920x1080 ids still bit-exact against the port."""
    sc = ou.load_scene(os.path.join(S, "hw2/sample.txt"), width=960, height=540)
    _render(renderer, sc)
    assert np.array_equal(renderer.primary_ids(), ou.port_render(sc, 0)["ids"])


# ---------------------------------------------------------------- MC: same stream vs the port
def _mean_close(gpu_sum, port_sum, spp, frac_needed, rel=2e-2):
    g, p = gpu_sum / spp, port_sum / spp
    d = np.abs(g - p).max(axis=2)
    ok = d <= rel * (np.abs(p).max(axis=2) + 1e-2)
    return float(ok.mean()) >= frac_needed, float(ok.mean())


@pytest.mark.parametrize("k", [1, 3, 4])
def test_hw4_matches_port_on_same_stream(renderer, k):
    sc = ou.load_scene(os.path.join(S, f"hw4/practice3_{k}.txt"), width=160, height=120, samples=16)
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1)
    ok, frac = _mean_close(acc, port["accum"], 16, 0.99)
    assert ok, frac
    assert abs(acc.mean() - port["accum"].mean()) <= 2e-3 * port["accum"].mean()


@pytest.mark.parametrize("k", [1, 3, 4])
def test_hw3_matches_port_on_same_stream(renderer, k):
    """hw3 = the uniform-hemisphere integrator (hw3/src/scene.cpp:31-106); hw3 ships hw4's scenes, the mode is
    explicit (text scenes with SAMPLES are inferred as hw4)."""
    sc = ou.load_scene(os.path.join(S, f"hw4/practice3_{k}.txt"), width=160, height=120, samples=16, mode=rt.MODE_HW3)
    assert sc.mode == rt.MODE_HW3
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1)
    ok, frac = _mean_close(acc, port["accum"], 16, 0.99)
    assert ok, frac
    assert abs(acc.mean() - port["accum"].mean()) <= 2e-3 * port["accum"].mean()
    # not the hw4 estimator: same scene, same stream, different image
    sc4 = ou.load_scene(os.path.join(S, f"hw4/practice3_{k}.txt"), width=160, height=120, samples=16)
    assert not np.allclose(_render(renderer, sc4), acc)


def test_hw5_matches_port_on_same_stream(renderer):
    """hw5: analytic figures scanned in constant memory + triangles through the library's own BVH (K7) against the
    port, which walks the reference's mixed-figure BVH: same Philox stream, same image up to float rounding."""
    sc = ou.load_scene(os.path.join(S, "hw5/mixed.txt"), samples=16)
    assert sc.mode == rt.MODE_HW5
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1, want_counters=True)
    ok, frac = _mean_close(acc, port["accum"], 16, 0.97, rel=5e-2)
    assert ok, frac
    assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()
    st = renderer.stats()
    assert abs(int(st["rays"]) - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])
    assert abs(int(st["light_queries"]) - int(port["counters"][3])) <= 2e-3 * int(port["counters"][3])


def test_hw5_without_triangles_matches_port_on_same_stream(renderer):
    """hw5 rules on a scene with analytic figures only (planes, rotated box light, ellipsoids, box): the triangle side
    of K7 is empty, the mixed-figure BVH of the description holds boxes and ellipsoids."""
    sc = ou.load_scene(os.path.join(G, "hw4_boxlight_spp4.txt"), samples=16, mode=rt.MODE_HW5)
    assert sc.mode == rt.MODE_HW5 and sc.desc.n_light_prims == 1
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1)
    ok, frac = _mean_close(acc, port["accum"], 16, 0.98, rel=5e-2)
    assert ok, frac
    assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()


def test_hw4_box_light_matches_port_on_same_stream(renderer):
    sc = ou.load_scene(os.path.join(G, "hw4_boxlight_spp4.txt"), samples=16)
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1)
    ok, frac = _mean_close(acc, port["accum"], 16, 0.99)
    assert ok, frac


@pytest.mark.parametrize("path,frac", [("hw8/sphere_emissive.gltf", 0.99), ("hw7/practice7_1.gltf", 0.85),
                                       ("hw7/practice7_4.gltf", 0.90), ("room20k", 0.85)])
def test_gltf_matches_port_on_same_stream(renderer, path, frac):
    """Pixels of the two images are equal unless one of the pixel's ~80 ray queries diverged at a tie (the device hits
    triangles with Moeller-Trumbore, the reference with its plane + projected 2x2 solve: different rounding at
    edges), hence the pixel budget `frac`; what the budget could hide -- a systematic error on one material -- is
    bounded per material class of the primary hit (room20k: all four texture kinds of hw8)."""
    sc = ou.load_scene(_scene_file(path), width=128, height=96, samples=16)
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1, want_counters=True)
    ok, got = _mean_close(acc, port["accum"], 16, frac, rel=5e-2)
    assert ok, got
    assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()
    bad = _class_means_close(acc, port["accum"], _primary_material(sc))
    assert not bad, bad
    st = renderer.stats()
    # same paths => (almost) the same number of closest-hit queries and light-pdf queries
    rays, rays_port = int(st["rays"]), int(port["counters"][0])
    if renderer.megakernel:
        assert abs(rays - rays_port) <= 2e-3 * rays_port
    else:   # the wavefront does not trace last-bounce rays that cannot reach an emitter (test_dark_rays_are_dropped...)
        assert 0.6 * rays_port <= rays <= 1.002 * rays_port
    lq, lq_port = int(st["light_queries"]), int(port["counters"][3])
    # every scheduling stops the last bounce at the emission term: its (unused) light-pdf query is not made
    assert 0.7 * lq_port <= lq <= lq_port
    assert st["paths"] == 128 * 96 * 16


def test_hw8_environment_map_matches_port_on_same_stream(renderer):
    """Misses sample the environment map (equirect lookup + sRGB decode on the device); with a non-black background
    the dark-ray rule must stay off: the ray count equals the port's in every scheduling."""
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=128, height=96, samples=16,
                       env_map=os.path.join(S, "hw8/sphere_emission.png"))
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1, want_counters=True)
    ok, got = _mean_close(acc, port["accum"], 16, 0.99, rel=5e-2)
    assert ok, got
    assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()
    assert abs(int(renderer.stats()["rays"]) - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])
    spp = 64
    renderer.clear()
    renderer.render(0, spp)
    gpu = _linear(renderer.resolve(spp))
    truth = _linear(ou.golden("hw8_sphere_emissive_env_128x96_spp4096"))
    sc.set_frame(samples=spp)
    ref = _linear(ou.port_render(sc, 0)["rgb"])
    shared = np.sqrt((1 + spp / 4096) / (1 - spp / 4096))       # see test_rmse_and_luminance_against_4096spp_reference
    assert np.sqrt(((gpu - truth) ** 2).mean()) <= 1.05 * shared * np.sqrt(((ref - truth) ** 2).mean()) + 1e-4


@pytest.mark.parametrize("path,w,h", [("hw8/sphere_emissive.gltf", 101, 67), ("hw7/practice7_4.gltf", 37, 1), ("hw7/practice7_1.gltf", 1, 1)])
def test_ragged_frames_match_port(renderer, path, w, h):
    """Frames that are not multiples of the 8x4 pixel tile (partial tiles leave dead slots in a wave; the first queue
    is then compacted with atomics instead of being the identity), down to a single pixel."""
    sc = ou.load_scene(os.path.join(S, path), width=w, height=h, samples=16)
    acc = _render(renderer, sc)
    assert acc.shape == (h, w, 3)
    port = ou.port_render(sc, 1)
    assert np.isfinite(acc).all()
    assert renderer.stats()["paths"] == w * h * 16
    if w * h > 64:
        ok, got = _mean_close(acc, port["accum"], 16, 0.95, rel=5e-2)
        assert ok, got
        assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()
    else:   # a handful of pixels: one path that rounds to another triangle moves the mean, so only a loose bound
        assert abs(acc.mean() - port["accum"].mean()) <= 0.25 * port["accum"].mean() + 1e-3


@pytest.mark.parametrize("path", ["hw8/sphere_emissive.gltf", "hw7/practice7_4.gltf"])
def test_dark_rays_are_dropped_without_changing_the_image(path, monkeypatch):
    """With a black background the last bounce only adds the emission of what it hits, so the wavefront ends a path one
    bounce early when its last ray pierces no emissive triangle (light stage: pdf sum == 0; shade: misses the emitters'
    boxes).  Those paths would have added exactly zero: the image is bit-identical with the rule off, the ray count
    with the rule off equals the port's, and fewer rays are traced with it on."""
    sc = ou.load_scene(os.path.join(S, path), width=128, height=96, samples=16)
    port = ou.port_render(sc, 1, want_counters=True)
    out = {}
    for drop in ("1", "0"):
        r = rt.Renderer(flags=0 if drop == "1" else rt.FLAG_KEEP_DARK_RAYS)
        out[drop] = (_render(r, sc), int(r.stats()["rays"]))
        r.close()
    assert np.array_equal(out["1"][0], out["0"][0])
    assert abs(out["0"][1] - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])
    assert out["1"][1] < out["0"][1]


@pytest.mark.parametrize("w,h", [(128, 96), (101, 67)])
def test_implicit_camera_state_is_the_stored_one(w, h, monkeypatch):
    """The first extend / shade of a wave derive a camera path's state from its slot (origin = camera, throughput 1,
    radiance 0, Philox key / counter) instead of reading what wf_generate would have stored: bit-identical image."""
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=w, height=h, samples=8)
    out = {}
    for v in ("1", "0"):
        r = rt.Renderer(flags=0 if v == "1" else rt.FLAG_EXPLICIT_CAMERA_STATE)
        out[v] = (_render(r, sc), r.stats()["rays"])
        r.close()
    assert np.array_equal(out["1"][0], out["0"][0]) and out["1"][1] == out["0"][1]


def test_exported_scene_renders_the_same_image_in_another_context():
    """b200rt_export_scene / b200rt_import_scene_*: the device blob of one context, copied device to device into
    another context's arena and rebased (textures, materials, both trees), renders bit-identical frames -- the path
    raytracing_course_hw_b200.dist.upload_replicated takes over NCCL on several GPUs."""
    torch = pytest.importorskip("torch")
    from raytracing_course_hw_b200 import dist as rtdist
    dev = torch.device("cuda", 0)
    sc = ou.load_scene(_room(), width=160, height=96, samples=4)
    a, b = rt.Renderer(), rt.Renderer()
    a.upload(sc)
    hdr, src, n = a.export_scene()
    dst = b.import_scene_begin(hdr)
    assert dst != src and n > 0
    rtdist.device_bytes(dst, n, dev).copy_(rtdist.device_bytes(src, n, dev))
    torch.cuda.synchronize()
    b.import_scene_end(hdr, sc)
    imgs = []
    for r in (a, b):
        r.render(0, 4)
        imgs.append(r.download_accum())
    assert np.array_equal(imgs[0], imgs[1]) and imgs[0].max() > 0
    c = rt.Renderer()
    c.replicate_from(a, sc)             # the same three steps in one call (what the CLI does for B200RT_GPUS > 1)
    c.render(0, 4)
    assert np.array_equal(c.download_accum(), imgs[0])
    c.close()
    # text scenes are not exported (they are a few hundred bytes of constant memory)
    a.upload(ou.load_scene(os.path.join(S, "hw4/practice3_1.txt")))
    with pytest.raises(rt.B200rtError):
        a.export_scene()
    with pytest.raises(rt.B200rtError):
        b.import_scene_begin(b"\0" * 64)
    a.close(); b.close()


def test_reference_tree_and_rebuilt_tree_agree():
    """FLAG_REFERENCE_BVH traverses the description's SAH tree, the default the library's own binned-SAH
    tree over the same triangles: same closest hits (ties aside) => same image on the same stream."""
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=128, height=96, samples=8)
    imgs, rays = [], []
    for flags in (0, rt.FLAG_REFERENCE_BVH):
        r = rt.Renderer(flags=flags)
        r.upload(sc)
        r.render(0, 8)
        imgs.append(r.download_accum().astype(np.float64))
        rays.append(int(r.stats()["rays"]))
        r.close()
    assert abs(rays[0] - rays[1]) <= 1e-3 * rays[1]
    close = np.abs(imgs[0] - imgs[1]).max(axis=2) <= 1e-3 * np.maximum(1.0, imgs[1].max(axis=2))
    assert close.mean() >= 0.995, close.mean()


@pytest.mark.parametrize("path,frac", [("hw7/practice7_1.gltf", 0.85), ("hw6/practice7_1_glass.gltf", 0.85)])
def test_hw6_matches_port_on_same_stream(renderer, path, frac):
    """hw6 mode (hw5 integrator on triangles; the device traverses its own binned-SAH tree, the port the
    reference's): same Philox stream => same paths up to float rounding of the hit points."""
    sc = ou.load_scene(os.path.join(S, path), width=128, height=96, samples=16, mode=6)
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1, want_counters=True)
    ok, got = _mean_close(acc, port["accum"], 16, frac, rel=5e-2)
    assert ok, got
    assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean()
    st = renderer.stats()
    assert abs(int(st["rays"]) - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])
    assert abs(int(st["light_queries"]) - int(port["counters"][3])) <= 2e-3 * int(port["counters"][3])
    assert st["paths"] == 128 * 96 * 16


@pytest.mark.parametrize("path", ["hw6/practice7_1_glass.gltf", "hw7/practice7_1.gltf"])
def test_hw6_wavefront_and_per_pixel_kernel_trace_the_same_paths(path):
    """hw6 runs through the wavefront pipeline (hw6_surface / hw6_finish as its shade stage; the default above
    B200RT_HW6_PIXEL_KERNEL_MAX_TRIS triangles) and through the one-thread-per-pixel kernel K6 (small scenes,
    FLAG_MEGAKERNEL): the same two device functions on the same Philox stream.  The
    wavefront drops rays that provably add nothing (FLAG_KEEP_DARK_RAYS turns that off), so with the flag the ray
    counts are EQUAL; the images agree to float rounding (different kernels, different FMA contraction) on
    all but the few pixels where a rounding flips a hit at a triangle edge."""
    sc = ou.load_scene(os.path.join(S, path), width=160, height=120, samples=8, mode=6)
    out = {}
    W = rt.FLAG_WAVEFRONT       # (these scenes are small enough for the per-pixel kernel to be the default)
    for name, flags in (("wave", W | rt.FLAG_KEEP_DARK_RAYS), ("wave-default", W), ("wave-counting", W | rt.FLAG_KEEP_DARK_RAYS | rt.FLAG_COUNTERS),
                        ("pixel", rt.FLAG_MEGAKERNEL)):
        r = rt.Renderer(flags=flags)
        try:
            out[name] = (_render(r, sc, 0, 8), r.stats())
        finally:
            r.close()
    a, b = out["wave"][0], out["pixel"][0]
    assert out["wave"][1]["paths"] == out["pixel"][1]["paths"] == 160 * 120 * 8
    assert abs(int(out["wave"][1]["rays"]) - int(out["pixel"][1]["rays"])) <= 1e-3 * int(out["pixel"][1]["rays"])
    # (the per-pixel kernel also makes the last bounce's light-pdf query, whose weight multiplies a ray that is never traced)
    assert 0.7 * int(out["pixel"][1]["light_queries"]) <= int(out["wave"][1]["light_queries"]) <= 1.001 * int(out["pixel"][1]["light_queries"])
    close = np.abs(a - b).max(axis=2) <= 1e-3 * (np.abs(b).max(axis=2) + 1e-2)
    assert close.mean() >= 0.99, close.mean()
    assert abs(a.mean() - b.mean()) <= 2e-3 * b.mean()
    # dropping dark rays and counting change nothing in the image
    assert np.array_equal(out["wave-default"][0], a) and np.array_equal(out["wave-counting"][0], a)
    assert out["wave-default"][1]["rays"] <= out["wave"][1]["rays"]


@pytest.mark.parametrize("name,mode,frac", [("big:hw7/practice7_2.gltf", 0, 0.99), ("big:hw7/practice7_3.gltf", 0, 0.99),
                                            ("big:hw6/practice6_1.gltf", 6, 0.97), ("big:hw6/practice6_2.gltf", 6, 0.97),
                                            ("bench268k", 0, 0.98)])
def test_large_scenes_match_port_on_same_stream(name, mode, frac):
    """The sizes the headline is quoted on (the reference's large example scenes, copied next to its binaries by
    oracle/build_ref.sh, and the 268 k-triangle bench scene) at a small frame: the product's default scheduling for the
    scene against the port on the same Philox stream -- image, mean, and with FLAG_KEEP_DARK_RAYS the ray and
    light-query counts.  hw6 scenes of this size run the wavefront pipeline."""
    path = _scene_file(name)
    if path is None:
        pytest.skip("the reference's example scenes are not on this machine (oracle/build_ref.sh copies them)")
    sc = ou.load_scene(path, width=96, height=64, samples=8, **({"mode": mode} if mode else {}))
    r = rt.Renderer(flags=rt.FLAG_KEEP_DARK_RAYS)
    try:
        acc = _render(r, sc)
        st = r.stats()
    finally:
        r.close()
    port = ou.port_render(sc, 1, want_counters=True)
    ok, got = _mean_close(acc, port["accum"], 8, frac, rel=5e-2)
    assert ok, got
    assert abs(acc.mean() - port["accum"].mean()) <= 2e-2 * port["accum"].mean()
    rays, rays_port = int(st["rays"]), int(port["counters"][0])
    assert abs(rays - rays_port) <= 1e-3 * rays_port, (rays, rays_port)
    lq, lq_port = int(st["light_queries"]), int(port["counters"][3])
    assert 0.7 * lq_port <= lq <= 1.005 * lq_port, (lq, lq_port)
    assert st["paths"] == 96 * 64 * 8
    print(f"{name}: pixels within 5 % {got:.4f}, rays {rays} vs {rays_port}, light queries {lq} vs {lq_port}")


def test_hw6_without_light_sampling_in_both_schedulings():
    """hw6 rules on a description whose emitter list is empty (Mix = Cosine alone, k = 1) over a bright background: the
    branch of the wavefront shade stage that finishes diffuse bounces itself and sends every ray -- diffuse, mirror,
    dielectric -- straight to the next bounce, against the per-pixel kernel and the port on the same stream."""
    sc = ou.load_scene(os.path.join(S, "hw6/practice7_1_glass.gltf"), width=128, height=96, samples=8, mode=6)
    sc.desc.n_light_tris = 0
    sc.desc.n_light_nodes = 0
    sc.desc.bg_color[0], sc.desc.bg_color[1], sc.desc.bg_color[2] = 0.6, 0.7, 0.9
    out = {}
    for name, flags in (("wave", rt.FLAG_WAVEFRONT), ("pixel", rt.FLAG_MEGAKERNEL)):
        r = rt.Renderer(flags=flags)
        try:
            out[name] = (_render(r, sc, 0, 8), r.stats())
        finally:
            r.close()
    port = ou.port_render(sc, 1, want_counters=True)
    for name, (acc, st) in out.items():
        ok, got = _mean_close(acc, port["accum"], 8, 0.97, rel=5e-2)
        assert ok, (name, got)
        assert abs(acc.mean() - port["accum"].mean()) <= 1e-2 * port["accum"].mean(), name
        assert abs(int(st["rays"]) - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0]), name
        assert int(st["light_queries"]) == 0, name
    assert acc.mean() > 0.3      # the background is seen


# ---------------------------------------------------------------- MC: north_star statistical bar
def _linear(img8):
    return ou.inv_display(img8)


@pytest.mark.parametrize("name,path,kind", [
    ("hw8_sphere_emissive_128x96", "hw8/sphere_emissive.gltf", "gltf"),
    ("hw7_practice7_1_128x96", "hw7/practice7_1.gltf", "gltf"),
    ("hw7_practice7_4_128x96", "hw7/practice7_4.gltf", "gltf"),
    ("hw8_room20k_128x96", "room20k", "room"),
    ("hw6_practice7_1_128x96", "hw7/practice7_1.gltf", "gltf6"),
    ("hw6_practice7_1_glass_128x96", "hw6/practice7_1_glass.gltf", "gltf6"),
    ("hw4_p3_1", "hw4_p3_1_spp4096.txt", "text"),
    ("hw4_p3_3", "hw4_p3_3_spp4096.txt", "text"),
    ("hw4_p3_4", "hw4_p3_4_spp4096.txt", "text"),
    ("hw3_p3_1", "hw4_p3_1_spp4096.txt", "text3"),
    ("hw3_p3_3", "hw4_p3_3_spp4096.txt", "text3"),
    ("hw3_p3_4", "hw4_p3_4_spp4096.txt", "text3"),
    ("hw5_mixed", "hw5_mixed_spp2048.txt", "text"),
])
def test_rmse_and_luminance_against_4096spp_reference(renderer, name, path, kind):
    spp = 64
    if kind == "room":      # textured hw8 room: baseColor / metallicRoughness / normal-map / emissive textures
        if not ou.has_golden(name + "_spp4096") or not ou.room_matches_golden(_scene_file(path)):
            pytest.skip("room golden absent or rendered from a different generated scene")
        sc = ou.load_scene(_scene_file(path), width=128, height=96, samples=spp)
    elif kind == "gltf":
        sc = ou.load_scene(os.path.join(S, path), width=128, height=96, samples=spp)
    elif kind == "gltf6":
        sc = ou.load_scene(os.path.join(S, path), width=128, height=96, samples=spp, mode=6)
    elif kind == "text3":
        sc = ou.load_scene(os.path.join(G, path), samples=spp, mode=rt.MODE_HW3)
    else:
        sc = ou.load_scene(os.path.join(G, path), samples=spp)
    # hw5: 2048 spp is as far as the reference binary gets on the fixture (tools/make_golden.py)
    truth = _linear(ou.golden(name + ("_spp2048" if name.startswith("hw5") else "_spp4096")))
    renderer.upload(sc)
    renderer.render(0, spp)
    gpu = _linear(renderer.resolve(spp))
    ref_equal_spp = _linear(ou.port_render(sc, 0)["rgb"])     # == the reference binary's image (test_oracle)
    rmse_gpu = np.sqrt(((gpu - truth) ** 2).mean())
    rmse_ref = np.sqrt(((ref_equal_spp - truth) ** 2).mean())
    # The reference's equal-spp image is not independent of the ground truth: same per-pixel seed, so its `spp`
    # samples are the first `spp` of the truth's S samples.  Its error against the truth is therefore smaller by
    # sqrt(1 - spp/S) than against the true mean, while an independent estimator's is larger by sqrt(1 + spp/S)
    # (the truth's own noise).  The bar "no worse than the reference at equal spp" is applied after that
    # correction, with 5 % for the run-to-run spread of the RMSE itself.
    truth_spp = 2048 if name.startswith("hw5") else 4096
    shared = np.sqrt((1 + spp / truth_spp) / (1 - spp / truth_spp))
    assert rmse_gpu <= 1.05 * shared * rmse_ref + 1e-4, (rmse_gpu, rmse_ref, shared)
    # mean luminance within 1 % of the 4096-spp reference (use 1024 spp to push MC noise well below 1 %)
    renderer.clear()
    renderer.render(0, 1024)
    lum = _linear(renderer.resolve(1024)).mean()
    assert abs(lum - truth.mean()) <= 0.01 * truth.mean(), (lum, truth.mean())


@pytest.mark.parametrize("name,path", [("hw6_practice7_1_128x96", "hw7/practice7_1.gltf"), ("hw6_practice7_1_glass_128x96", "hw6/practice7_1_glass.gltf")])
def test_hw6_wavefront_rmse_and_luminance_against_4096spp_reference(name, path):
    """The north_star statistical bar for hw6 through the WAVEFRONT pipeline (the fixtures are small enough for the
    per-pixel kernel to be the default, which the parametrised test above covers): RMSE against the hw6 binary's
    4096-spp image no worse than the binary's own at equal spp, mean luminance within 1 %."""
    spp = 64
    sc = ou.load_scene(os.path.join(S, path), width=128, height=96, samples=spp, mode=6)
    truth = _linear(ou.golden(name + "_spp4096"))
    r = rt.Renderer(flags=rt.FLAG_WAVEFRONT)
    try:
        r.upload(sc)
        r.render(0, spp)
        gpu = _linear(r.resolve(spp))
        r.clear()
        r.render(0, 1024)
        lum = _linear(r.resolve(1024)).mean()
    finally:
        r.close()
    ref_equal_spp = _linear(ou.port_render(sc, 0)["rgb"])
    rmse_gpu = np.sqrt(((gpu - truth) ** 2).mean())
    rmse_ref = np.sqrt(((ref_equal_spp - truth) ** 2).mean())
    shared = np.sqrt((1 + spp / 4096) / (1 - spp / 4096))      # (see the test above)
    assert rmse_gpu <= 1.05 * shared * rmse_ref + 1e-4, (rmse_gpu, rmse_ref, shared)
    assert abs(lum - truth.mean()) <= 0.01 * truth.mean(), (lum, truth.mean())


# ---------------------------------------------------------------- K5 resolve
@pytest.mark.parametrize("path", ["hw8/sphere_emissive.gltf", "hw7/practice7_4.gltf"])
def test_resolve_bytes_match_port(renderer, path):
    sc = ou.load_scene(os.path.join(S, path), width=256, height=192, samples=8)
    acc = _render(renderer, sc)
    img = renderer.resolve(8)
    cmp = ou.compare_u8(img, ou.port_resolve(acc, 8))
    assert cmp["within1"] == 1.0 and cmp["exact"] >= 0.999, cmp


def test_resolve_edge_values(renderer):
    """Zeros, huge values, NaN-free ramps through upload_accum -> K5."""
    sc = ou.load_scene(os.path.join(S, "hw7/practice7_1.gltf"), width=64, height=64, samples=1)
    _render(renderer, sc)
    ramp = np.zeros((64, 64, 3), np.float32)
    ramp[..., 0] = np.linspace(0, 4, 64 * 64, dtype=np.float32).reshape(64, 64)
    ramp[..., 1] = 1e6
    renderer.upload_accum(ramp)
    assert np.array_equal(renderer.resolve(1), ou.port_resolve(ramp, 1))


def _room(tris=20000):
    return ou.room_scene(tris)


def test_every_build_renders_the_same_image():
    """Plain wavefront, counting wavefront, one wave in flight, and the megakernel evaluate the same
    arithmetic on the same Philox stream: bit-identical radiance, and a second process-independent context
    reproduces it (regression: a register-starved counting build of wf_shade once shaded ~0.1 % of the hits
    of a bounce differently).  The port on the same stream bounds the absolute error."""
    sc = ou.load_scene(_room(), width=320, height=180, samples=2)
    imgs = {}
    for name, flags in (("plain", 0), ("plain-again", 0), ("counting", rt.FLAG_COUNTERS), ("serial", rt.FLAG_SERIAL_WAVES),
                        ("megakernel", rt.FLAG_MEGAKERNEL), ("megakernel-counting", rt.FLAG_MEGAKERNEL | rt.FLAG_COUNTERS)):
        r = rt.Renderer(flags=flags)
        try:
            imgs[name] = _render(r, sc, 0, 2)
        finally:
            r.close()
    for name, img in imgs.items():
        assert np.array_equal(img, imgs["plain"]), (name, int((img != imgs["plain"]).any(axis=2).sum()))
    port = ou.port_render(sc, 1)["accum"]
    ok, frac = _mean_close(imgs["plain"], port, 2, 0.97, rel=5e-2)
    assert ok, frac
    assert abs(imgs["plain"].mean() - port.mean()) <= 1e-2 * port.mean()


# ---------------------------------------------------------------- size-independent properties
def test_sample_ranges_add_up_and_are_deterministic(renderer):
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=640, height=360, samples=32)
    full = _render(renderer, sc, 0, 32)
    again = _render(renderer, sc, 0, 32)
    assert np.array_equal(full, again), "no float atomics: the same launch is bit-reproducible"
    renderer.clear()
    for b in range(0, 32, 8):                      # what 4 GPUs would each render
        renderer.render(b, b + 8)
    parts = renderer.download_accum()
    assert np.allclose(full, parts, rtol=1e-4, atol=1e-5)
    # a different seed gives a different (but statistically equal) image
    r2 = rt.Renderer(seed=7)
    try:
        other = _render(r2, sc, 0, 32)
    finally:
        r2.close()
    assert not np.array_equal(other, full)
    assert abs(other.mean() - full.mean()) <= 0.02 * full.mean()


def test_full_size_frame_properties(renderer):
    """BASELINE config size (4K): every pixel written, finite, non-negative, stats consistent."""
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=3840, height=2160, samples=2)
    acc = _render(renderer, sc)
    st = renderer.stats()
    assert st["paths"] == 3840 * 2160 * 2
    assert st["paths"] <= st["rays"] <= 6 * st["paths"]
    assert np.isfinite(acc).all() and (acc >= 0).all()
    assert acc.reshape(-1, 3).max(axis=1).astype(bool).mean() > 0.3   # the sphere + light cover the frame
    img = renderer.resolve(2)
    assert img.shape == (2160, 3840, 3)


@pytest.mark.parametrize("height", [8192, 8200])
def test_wave_of_two_to_the_26_slots_and_more(height):
    """8192 x 8192 = 2^26 pixels is the first wave whose slot index leaves no room for the light-cut mask in a queue
    entry (the light stage then tests the cut boxes itself); 8192 x 8200 has slots ABOVE 2^26 (its last eight rows) --
    round 1 masked the queue entry to 26 bits there: those paths were dropped and the low slots they aliased onto
    (the first eight rows) finished the wrong bounce.  Closed, lit room (every pixel carries indirect light): same
    camera => same mean as a small frame, and the first / last eight rows look like the eight rows next to them."""
    path = _room()
    small = ou.load_scene(path, width=256, height=256, samples=64)
    big = ou.load_scene(path, width=8192, height=height, samples=1)
    r = rt.Renderer()
    ref = (_render(r, small) / 64).mean()
    acc = _render(r, big)
    st = r.stats()
    r.close()
    assert st["paths"] == 8192 * height and np.isfinite(acc).all()
    assert abs(acc.mean() - ref) <= 0.02 * ref, (acc.mean(), ref)
    for a, b in ((acc[:8], acc[8:16]), (acc[-8:], acc[-16:-8])):
        assert a.mean() > 0 and abs(a.mean() - b.mean()) <= 0.2 * b.mean(), (a.mean(), b.mean())


@pytest.mark.parametrize("mode_lines,mode", [("", rt.MODE_HW1), ("AMBIENT_LIGHT 0.2 0.2 0.2\nRAY_DEPTH 3\n", rt.MODE_HW2),
                                             ("SAMPLES 4\nRAY_DEPTH 3\n", rt.MODE_HW4), ("SAMPLES 4\nRAY_DEPTH 3\n", rt.MODE_HW3)])
def test_empty_text_scene_is_background(renderer, tmp_path, mode_lines, mode):
    """No primitive at all: every ray misses, the frame is the background colour in every text mode (the reference
    loops over an empty figure list, hw1/src/scene.cpp:9-19, hw4/src/scene.cpp:30-48)."""
    p = tmp_path / "empty.txt"
    p.write_text("DIMENSIONS 33 17\nBG_COLOR 0.25 0.5 0.75\nCAMERA_POSITION 0 0 0\nCAMERA_RIGHT 1 0 0\nCAMERA_UP 0 1 0\n"
                 "CAMERA_FORWARD 0 0 -1\nCAMERA_FOV_X 1.0\n" + mode_lines)
    sc = ou.load_scene(str(p), mode=mode if mode == rt.MODE_HW3 else 0)
    assert sc.mode == mode and sc.desc.n_prims == 0
    acc = _render(renderer, sc)
    n = 1 if mode <= rt.MODE_HW2 else 4
    assert np.allclose(acc / n, np.array([0.25, 0.5, 0.75], np.float32), rtol=1e-6)
    port = ou.port_render(sc, 1 if mode > rt.MODE_HW2 else 0)
    assert np.array_equal(renderer.resolve(n), port["rgb"])


def test_gltf_scene_without_emitters_under_an_environment_map(renderer, tmp_path):
    """No emissive triangle (FiguresMix is empty, Mix has two components, hw8/src/scene.cpp:64-74): the wavefront's
    shade stage finishes every bounce itself and there is no light stage; all light comes from the environment map."""
    import re
    import shutil
    for f in ("sphere_emissive.bin", "sphere_emission.png"):
        shutil.copy(os.path.join(S, "hw8", f), tmp_path / f)
    src = open(os.path.join(S, "hw8/sphere_emissive.gltf")).read()
    dark, n = re.subn(r'"emissiveFactor"\s*:\s*\[[^\]]*\]', '"emissiveFactor": [0, 0, 0]', src)
    assert n >= 1
    (tmp_path / "dark.gltf").write_text(dark)
    sc = ou.load_scene(str(tmp_path / "dark.gltf"), width=96, height=64, samples=16, env_map=os.path.join(S, "hw8/sphere_emission.png"))
    assert sc.desc.n_light_tris == 0 and sc.desc.n_tris == 970
    acc = _render(renderer, sc)
    port = ou.port_render(sc, 1, want_counters=True)
    assert port["accum"].mean() > 0
    ok, got = _mean_close(acc, port["accum"], 16, 0.99, rel=5e-2)
    assert ok, got
    assert abs(int(renderer.stats()["rays"]) - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])
    assert int(renderer.stats()["light_queries"]) == 0


def test_errors_are_reported_not_thrown(renderer):
    r = rt.Renderer()
    try:
        with pytest.raises(rt.B200rtError) as e:
            r.render(0, 1)
        assert e.value.code == -4
    finally:
        r.close()
    sc = ou.load_scene(os.path.join(S, "hw7/practice7_1.gltf"), width=16, height=16, samples=1)
    renderer.upload(sc)
    with pytest.raises(rt.B200rtError):
        renderer.render(4, 4)          # empty sample range


# ---- several GPUs behind the library (b200rt_group_*, fused peer reduce + resolve) -----------------------------
@pytest.mark.parametrize("w,h,members", [(128, 96, 2), (101, 67, 3)])
def test_group_on_one_device_equals_one_context(w, h, members):
    """Members sharing device 0 take the multi-GPU path (replication, sample split by global index, ONE kernel that
    sums the members' buffers in member order and resolves): the frame sum is EXACTLY own + peer1 + ... in float32,
    its bytes are exactly the single-context resolve of that sum, and they agree with a one-context render of the same
    samples up to float summation order.  101 x 67 x 3 floats is not a multiple of four: the scalar tail runs."""
    sc = ou.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=w, height=h, samples=12)
    g = rt.Group([0] * members)
    assert len(g) == members
    g.upload(sc)
    g.render(0, 12)
    total = g.download_accum()
    parts = [g.member_accum(i) for i in range(members)]
    expect = parts[0].copy()
    for p_ in parts[1:]:
        expect += p_
    assert np.array_equal(total, expect) and all(p_.max() > 0 for p_ in parts)
    img = g.resolve(12)
    one = rt.Renderer()
    one.upload(sc)
    one.render(0, 12)
    single = one.download_accum()
    assert np.allclose(total, single, rtol=1e-5, atol=1e-6)
    one.upload_accum(total)
    assert np.array_equal(one.resolve(12), img)
    st = g.stats()
    assert st["paths"] == w * h * 12
    one.close(); g.close()


def test_group_text_scene_and_deterministic_modes():
    """hw4 text scenes upload per member and split samples; hw2 (one sample per pixel) renders on member 0 alone and
    the others contribute zero buffers: bytes equal the single-context frame exactly."""
    sc = ou.load_scene(os.path.join(S, "hw4/practice3_1.txt"), width=96, height=64, samples=8)
    g = rt.Group([0, 0])
    g.upload(sc); g.render(0, 8)
    one = rt.Renderer(); one.upload(sc); one.render(0, 8)
    assert np.allclose(g.download_accum(), one.download_accum(), rtol=1e-5, atol=1e-6)
    sc2 = ou.load_scene(os.path.join(S, "hw2/sample.txt"))
    g.upload(sc2); g.render()
    one.upload(sc2); one.render()
    assert np.array_equal(g.resolve(1), one.resolve(1))
    one.close(); g.close()


def test_group_hw6_large_scene_takes_the_wavefront_on_every_member():
    """An hw6 scene above the per-pixel kernel's size limit, replicated to a second member: the replica keeps the
    triangle count (the scheduling rule reads it), so both members run the wavefront pipeline and the frame equals the
    one-context frame up to float summation order -- and is bit-identical to a forced-wavefront single context when the
    group has one member."""
    path = _scene_file("big:hw6/practice6_1.gltf")
    if path is None:
        pytest.skip("the reference's example scenes are not on this machine (oracle/build_ref.sh copies them)")
    sc = ou.load_scene(path, width=96, height=64, samples=8, mode=6)
    one = rt.Renderer(flags=rt.FLAG_WAVEFRONT)
    one.upload(sc); one.render(0, 8)
    single = one.download_accum()
    for members in (1, 2):
        g = rt.Group([0] * members)
        g.upload(sc); g.render(0, 8)
        total = g.download_accum()
        if members == 1:
            assert np.array_equal(total, single)
        else:
            assert np.allclose(total, single, rtol=1e-5, atol=1e-6)
            assert all(g.member_accum(i).max() > 0 for i in range(members))
        assert g.stats()["paths"] == 96 * 64 * 8
        g.close()
    one.close()


# ---- the product's default load: no trees in the description ----------------------------------------------------
@pytest.mark.parametrize("path,mode", [("hw7/practice7_4.gltf", 0), ("hw8/sphere_emissive.gltf", 0), ("hw7/practice7_1.gltf", 6)])
def test_default_load_without_trees_matches_port_on_same_stream(path, mode):
    """b200rt_host_load_gltf delivers triangles in file order and no trees; the library builds its own, the port
    traverses private ones (oracle/port_mc.cpp) and samples lights in the description's order: same Philox stream,
    same image up to edge ties; ray counts equal within 0.2 %."""
    sc = rt.load_scene(os.path.join(S, path), width=128, height=96, samples=16, mode=mode)
    assert sc.desc.n_nodes == 0 and not sc.desc.nodes and not sc.desc.light_nodes
    # KEEP_DARK_RAYS: the wavefront otherwise ends paths whose last ray cannot reach an emitter one bounce early
    # (bit-identical image, fewer rays: test_dark_rays_are_dropped...), and the port traces them all
    r = rt.Renderer(flags=rt.FLAG_KEEP_DARK_RAYS)
    acc = _render(r, sc)
    rays = int(r.stats()["rays"])
    r.close()
    port = ou.port_render(sc, 1, want_counters=True)
    p = port["accum"]
    d = np.abs(acc - p).max(axis=2) / 16
    ok = (d <= 5e-2 * (np.abs(p).max(axis=2) / 16 + 1e-2)).mean()
    assert ok >= 0.9, ok
    assert abs(acc.mean() - p.mean()) <= 2e-2 * p.mean()
    assert abs(rays - int(port["counters"][0])) <= 2e-3 * int(port["counters"][0])


def test_reference_bvh_flag_needs_the_description_trees():
    sc = rt.load_scene(os.path.join(S, "hw7/practice7_1.gltf"), width=16, height=16, samples=1)
    r = rt.Renderer(flags=rt.FLAG_REFERENCE_BVH)
    with pytest.raises(rt.B200rtError):
        r.upload(sc)
    r.close()


def test_hw7_file_loads_under_hw8_rules_with_generated_attributes():
    """hw8/src/sceneio.cpp:88-93 reads TANGENT and TEXCOORD_0 unconditionally; practice7_1.gltf has neither.  The loader
    generates them (uv = 0, tangent in the normal's frame): the file renders under hw8 rules like the port."""
    sc = rt.load_scene(os.path.join(S, "hw7/practice7_1.gltf"), width=96, height=64, samples=16, mode=rt.MODE_HW8)
    assert sc.mode == rt.MODE_HW8
    r = rt.Renderer()
    acc = _render(r, sc)
    r.close()
    p = ou.port_render(sc, 1)["accum"]
    d = np.abs(acc - p).max(axis=2) / 16
    assert (d <= 5e-2 * (np.abs(p).max(axis=2) / 16 + 1e-2)).mean() >= 0.9
    assert abs(acc.mean() - p.mean()) <= 2e-2 * p.mean()


# ---- the counters the roofline is built from, checked against a CPU walk of the SAME flattened tree ----
def _fixed_rays(sc, n_cam, n_rand, seed):
    """Camera rays of the scene (coherent) + seeded random segments between points of the scene's box (incoherent)."""
    d = sc.desc
    pos, right, up, fwd = (np.array(list(v), np.float32) for v in (d.cam_pos, d.cam_right, d.cam_up, d.cam_forward))
    rng = np.random.default_rng(seed)
    ty = np.tan(d.cam_fov / 2)
    tx = ty * sc.width / sc.height
    uv = rng.random((n_cam, 2), dtype=np.float32) * 2 - 1
    dirs = uv[:, :1] * tx * right - uv[:, 1:] * ty * up + fwd
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    cam = np.concatenate([np.broadcast_to(pos, dirs.shape), dirs], axis=1)
    tris = np.frombuffer(ctypes.string_at(d.tris, d.n_tris * 148), np.uint8).reshape(d.n_tris, 148)[:, :36].copy().view(np.float32).reshape(-1, 3)
    lo, hi = tris.min(axis=0), tris.max(axis=0)
    a = lo + rng.random((n_rand, 3), dtype=np.float32) * (hi - lo)
    b = lo + rng.random((n_rand, 3), dtype=np.float32) * (hi - lo)
    dd = b - a
    dd /= np.maximum(np.linalg.norm(dd, axis=1, keepdims=True), 1e-6)
    return np.concatenate([cam, np.concatenate([a, dd], axis=1)]).astype(np.float32)


@pytest.mark.parametrize("which", ["practice7_4", "room20k", "sphere"])
def test_device_counters_equal_a_cpu_walk_of_the_same_flattened_bvh(which):
    """b200rt_trace_rays runs the production wf_extend (counting build) on a fixed ray set; oracle/pairs_traverse.c walks
    the records downloaded from the device with the same ordered traversal.  Box and triangle arithmetic are explicitly
    rounded / fused on the device and restated with the same operations on the CPU: the hits (t, u, v, triangle, side)
    are bit-identical and the counters EQUAL -- these are the counts bench.py's algorithmic bytes are built from."""
    path = {"practice7_4": os.path.join(S, "hw7/practice7_4.gltf"), "sphere": os.path.join(S, "hw8/sphere_emissive.gltf"), "room20k": _room()}[which]
    sc = rt.load_scene(path, width=160, height=96, samples=1)
    r = rt.Renderer()
    r.upload(sc)
    rays = _fixed_rays(sc, 60000, 60000, 7)
    hits, nb, nt = r.trace_rays(rays)
    lay, rec, tri = r.scene_bvh()
    r.close()
    assert lay.n_tris == sc.desc.n_tris and lay.quant_bits == 15
    chits, per, cb, ct = ou.pairs_traverse(lay, rec, tri, rays)
    assert cb > 0 and ct > 0 and nb % 2 == 0
    bits, cbits = hits[:, 3].view(np.uint32), chits[:, 3].view(np.uint32)
    assert np.array_equal(bits, cbits), int((bits != cbits).sum())
    hit = bits != 0x7FFFFFFF
    assert hit.mean() > 0.3          # (a closed room has no misses at all)
    assert np.array_equal(hits[hit, :3].view(np.uint32), chits[hit, :3].view(np.uint32))
    assert (nb, nt) == (cb, ct), (nb, cb, nt, ct)
    assert int(per[:, 0].sum()) == nb and per[:, 0].max() < 2000


def test_no_firefly_clamp_mode_only_adds_energy():
    """B200RT_FLAG_NO_FIREFLY_CLAMP: the reference ends a path when a bounce weight exceeds 6 in some
    channel (DIRTY_DENOISE_HACK_BUBEN, hw8/src/scene.cpp:161 -- biased).  Without the clamp the same paths continue:
    every pixel's sum is >= the clamped one's (same Philox stream, radiance is non-negative), a visible share of pixels
    gains energy, nothing is NaN / inf, and the flag changes nothing else (the default stays the reference's image)."""
    sc = ou.load_scene(os.path.join(S, "hw7/practice7_4.gltf"), width=128, height=96, samples=32)
    a, b = rt.Renderer(), rt.Renderer(flags=rt.FLAG_NO_FIREFLY_CLAMP)
    clamped, free = _render(a, sc), _render(b, sc)
    m = rt.Renderer(flags=rt.FLAG_NO_FIREFLY_CLAMP | rt.FLAG_MEGAKERNEL)
    assert np.array_equal(_render(m, sc), free)                      # every scheduling honours the flag
    for r in (a, b, m):
        r.close()
    assert np.isfinite(free).all()
    assert (free >= clamped - 1e-6 * np.abs(clamped)).all()
    gained = (free > clamped * 1.0001).any(axis=2).mean()
    assert 0.001 < gained < 0.9, gained
    assert free.mean() > clamped.mean()
