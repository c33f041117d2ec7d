#!/usr/bin/env python3
"""Ad-hoc GPU-vs-oracle comparison used while bringing kernels up (prints, asserts nothing)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import raytracing_course_hw_b200 as rt
import oracle_util as ou

S = ou.SCENES


def show(name, gpu_acc, port_acc, spp):
    g, p = gpu_acc / spp, port_acc / spp
    d = np.abs(g - p).max(axis=2)
    rel = d / (np.abs(p).max(axis=2) + 1e-3)
    print(f"{name}: mean gpu {g.mean():.6f} port {p.mean():.6f}  px rel<1e-3 {(rel < 1e-3).mean():.4f} rel<5e-2 {(rel < 5e-2).mean():.4f} max abs {d.max():.4f}", flush=True)


def main():
    r = rt.Renderer(flags=rt.FLAG_COUNTERS)
    for name, path in (("hw1", "hw1/sample.txt"), ("hw2", "hw2/sample.txt")):
        sc = rt.load_scene(os.path.join(S, path))
        r.upload(sc); r.render(); img = r.resolve(1); ids = r.primary_ids(); st = r.stats()
        o = ou.port_render(sc, 0)
        print(name, "rgb", ou.compare_u8(img, o["rgb"]), "ids equal", float((ids == o["ids"]).mean()), "ms", st["kernel_ms"], flush=True)
        acc = r.download_accum()
        print("   accum bit-equal frac", float((acc.view(np.uint32) == o["accum"].view(np.uint32)).all(axis=2).mean()))
    for k in (1, 3, 4):
        sc = rt.load_scene(os.path.join(S, f"hw4/practice3_{k}.txt"), width=160, height=120, samples=16)
        r.upload(sc); r.render(); acc = r.download_accum(); st = r.stats()
        o = ou.port_render(sc, 1)
        show(f"hw4_{k}", acc, o["accum"], 16)
        print("   ", {k2: st[k2] for k2 in ("kernel_ms", "paths", "rays")})
    for name, path in (("hw7_1", "hw7/practice7_1.gltf"), ("hw7_4", "hw7/practice7_4.gltf"), ("hw8", "hw8/sphere_emissive.gltf")):
        sc = rt.load_scene(os.path.join(S, path), width=128, height=96, samples=16)
        r.upload(sc); r.render(); acc = r.download_accum(); st = r.stats()
        o = ou.port_render(sc, 1, want_counters=True)
        show(name, acc, o["accum"], 16)
        print("    gpu", st)
        print("    port counters", o["counters"].tolist())
        img = r.resolve(16)
        print("    resolve vs port_resolve", ou.compare_u8(img, ou.port_resolve(acc, 16)))
    # throughput probe
    sc = rt.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=1920, height=1080, samples=16)
    r2 = rt.Renderer()
    r2.upload(sc)
    for _ in range(3):
        r2.clear(); r2.render(); st = r2.stats()
        print("hw8 1080p x16:", st["kernel_ms"], "ms ->", st["paths"] / st["kernel_ms"] / 1e3, "Mpaths/s", st["rays"] / st["kernel_ms"] / 1e3, "Mrays/s", flush=True)


if __name__ == "__main__":
    main()
