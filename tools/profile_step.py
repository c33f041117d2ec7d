#!/usr/bin/env python3
"""Short, ncu-friendly run of the headline step (same scene and kernels as bench.py, smaller frame)."""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench
import raytracing_course_hw_b200 as rt

ap = argparse.ArgumentParser()
ap.add_argument("--width", type=int, default=1920)
ap.add_argument("--height", type=int, default=1080)
ap.add_argument("--spp", type=int, default=4)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--tris", type=int, default=270000)
ap.add_argument("--scene", default=None)
ap.add_argument("--counters", action="store_true")
ap.add_argument("--megakernel", action="store_true")
ap.add_argument("--serial", action="store_true", help="one wave in flight: the stage times are exclusive kernel times")
a = ap.parse_args()
path = a.scene or bench.scene_path(a.tris)
sc = rt.load_scene(path, width=a.width, height=a.height, samples=a.spp)
r = rt.Renderer(flags=(rt.FLAG_COUNTERS if a.counters else 0) | (rt.FLAG_MEGAKERNEL if a.megakernel else 0) | (rt.FLAG_SERIAL_WAVES if a.serial else 0))
r.upload(sc)
for i in range(a.steps):
    r.render(i * a.spp, (i + 1) * a.spp)
    st = r.stats()
    print(f"step {i}: {st['kernel_ms']:.3f} ms  {st['paths'] / st['kernel_ms'] / 1e3:.1f} Mpaths/s  {st['rays'] / st['kernel_ms'] / 1e3:.1f} Mrays/s"
          f"  | extend {st['extend_ms']:.2f} shade {st['shade_ms']:.2f} light {st['light_ms']:.2f} ms (bracketed per launch)", flush=True)
if a.counters:
    print({k: v for k, v in st.items()})
img = r.resolve(a.spp * a.steps)
print("mean pixel", float(img.mean()))
r.close()
