#!/usr/bin/env python3
"""Small renders of every mode for `compute-sanitizer --tool memcheck|racecheck|initcheck python tools/sanitize_probe.py`."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracing_course_hw_b200 as rt
S = os.path.join(ROOT, "tests", "scenes")
cases = [("hw1/sample.txt", {}, 0), ("hw2/sample.txt", {}, 0), ("hw4/practice3_3.txt", dict(width=64, height=48, samples=4), 0),
         ("hw4/practice3_3.txt", dict(width=64, height=48, samples=4, mode=3), 0), ("hw5/mixed.txt", dict(width=64, height=48, samples=4), 0),
         ("hw7/practice7_1.gltf", dict(width=64, height=48, samples=4, mode=6), 0), ("hw7/practice7_4.gltf", dict(width=61, height=47, samples=4), 0),
         ("hw8/sphere_emissive.gltf", dict(width=64, height=48, samples=4), 0), ("hw8/sphere_emissive.gltf", dict(width=64, height=48, samples=4), rt.FLAG_MEGAKERNEL),
         ("hw8/sphere_emissive.gltf", dict(width=64, height=48, samples=4), rt.FLAG_COUNTERS)]
for path, kw, flags in cases:
    r = rt.Renderer(flags=flags)
    sc = rt.load_scene(os.path.join(S, path), **kw)
    r.upload(sc)
    r.render()
    img = r.resolve(sc.samples)
    print(path, kw, flags, "mean", float(img.mean()), flush=True)
    r.close()

# round 2: several contexts behind one group (fused peer reduce + resolve; 61 x 47 x 3 floats is not a multiple of
# four: the scalar tail runs), the ray-set query of the production traversal kernel, the allocation-free resolve
import numpy as np
sc = rt.load_scene(os.path.join(S, "hw7/practice7_4.gltf"), width=61, height=47, samples=6)
g = rt.Group([0, 0, 0])
g.upload(sc)
g.render(0, 6)
print("group", float(g.resolve(6).mean()), float(g.download_accum().mean()), flush=True)
g.close()
r = rt.Renderer()
r.upload(sc)
rays = np.random.default_rng(1).standard_normal((5000, 6)).astype(np.float32)
hits, nb, nt = r.trace_rays(rays)
lay, rec, tri = r.scene_bvh()
print("trace_rays", nb, nt, rec.shape, tri.shape, flush=True)
r.render(0, 6)
r.synchronize()
ptr, _ = r.accum_device_ptr()
print("noalloc", float(rt.resolve_noalloc(0, ptr, 61, 47, 6, sc.mode, stream=r.stream()).mean()), flush=True)
r.close()
