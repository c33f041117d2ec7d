#!/usr/bin/env python3
"""Times the end-to-end call sequence of bench.py's e2e leg piece by piece (host wall clock)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
import raytracing_course_hw_b200 as rt
W, H, SPP = 3840, 2160, 16
sc = rt.load_scene(bench.scene_path(270000), width=W, height=H, samples=SPP)
r = rt.Renderer()
out = np.empty((H, W, 3), np.uint8)
for i in range(3):
    t0 = time.perf_counter(); r.upload(sc)
    t1 = time.perf_counter(); r.render(i * SPP, (i + 1) * SPP); r.synchronize()
    t2 = time.perf_counter(); r.resolve(SPP, out)
    t3 = time.perf_counter()
    print(f"iter {i}: upload {1e3 * (t1 - t0):.1f} ms  render {1e3 * (t2 - t1):.1f} ms  resolve {1e3 * (t3 - t2):.1f} ms", flush=True)
r.close()
