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
