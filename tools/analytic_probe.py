#!/usr/bin/env python3
"""Short, ncu-friendly run of the analytic / per-path kernels: K3 (hw4 practice3_3), K7 (hw5 mixed), K6 (hw6 practice7_1),
K2 (hw2 sample).  One launch each after a warm-up launch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracing_course_hw_b200 as rt
S = os.path.join(ROOT, "tests", "scenes")
r = rt.Renderer()
for path, mode, kw in (("hw4/practice3_3.txt", 0, dict(width=960, height=540, samples=32)), ("hw5/mixed.txt", 0, dict(width=960, height=540, samples=32)),
                       ("hw7/practice7_1.gltf", 6, dict(width=960, height=540, samples=32)), ("hw2/sample.txt", 0, dict(width=1920, height=1080))):
    sc = rt.load_scene(os.path.join(S, path), mode=mode, **kw)
    r.upload(sc)
    for _ in range(2):
        r.render(0, sc.samples)
    print(path, r.stats()["kernel_ms"], "ms", flush=True)
r.close()
