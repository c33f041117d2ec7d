#!/usr/bin/env python3
"""Throughput of the analytic-scene kernels (K3 hw3 / hw4, K7 hw5, K6 hw6, K2 hw2) on their BASELINE scenes: kernel time from the
library's CUDA events.  Development / measurement tool (profiles/rNN_analytic_speed.log is a copy of its output)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracing_course_hw_b200 as rt
S = os.path.join(ROOT, "tests", "scenes")
CASES = [("hw4 practice3_1", "hw4/practice3_1.txt", 0, 1920, 1080, 256), ("hw4 practice3_3", "hw4/practice3_3.txt", 0, 1920, 1080, 256),
         ("hw4 practice3_4", "hw4/practice3_4.txt", 0, 1920, 1080, 256), ("hw4 practice3_5", "hw4/practice3_5.txt", 0, 1920, 1080, 256),
         ("hw3 practice3_3", "hw4/practice3_3.txt", rt.MODE_HW3, 1920, 1080, 256), ("hw5 mixed", "hw5/mixed.txt", 0, 1920, 1080, 256),
         ("hw6 glass", "hw6/practice7_1_glass.gltf", rt.MODE_HW6, 1920, 1080, 64),
         ("hw2 sample 4K", "hw2/sample.txt", 0, 3840, 2160, 1)]
only = sys.argv[1:]      # optional: substrings of the case names to run
r = rt.Renderer()
for name, path, mode, w, h, spp in CASES:
    if only and not any(o in name for o in only):
        continue
    sc = rt.load_scene(os.path.join(S, path), width=w, height=h, samples=spp, mode=mode)
    r.upload(sc)
    r.render(0, spp)
    r.synchronize()
    best = None
    for _ in range(2):
        r.clear()
        r.render(0, spp)
        st = r.stats()
        best = st if best is None or st["kernel_ms"] < best["kernel_ms"] else best
    print(f"{name:18s} {w}x{h}x{spp}: {best['kernel_ms']:9.3f} ms  {best['paths'] / best['kernel_ms'] / 1e3:9.1f} Mpaths/s  {best['rays'] / best['kernel_ms'] / 1e3:9.1f} Mrays/s  rays/path {best['rays'] / max(best['paths'], 1):.2f}", flush=True)
r.close()
