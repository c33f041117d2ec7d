#!/usr/bin/env python3
"""Throughput of every BASELINE config (C1..C5) on one GPU, with the unmodified reference binary timed
next to each on the box's host cores (bounded sample, load time subtracted).  Development/measurement tool: writes
one JSON document (profiles/rNN_config_sweep.json is a copy of its output).

  python tools/config_sweep.py [--quick] [--out gpurun_out/config_sweep.json]

GPU numbers: wall clock around b200rt_render + b200rt_resolve with the scene resident (same definition as bench.py's
`value`), kernel time from the library's own CUDA-event brackets.  CPU numbers: oracle/_ref/hwN (TEST
INFRASTRUCTURE, used here only as the reported baseline) at a reduced frame / sample count, linear in both.
"""
import argparse, json, os, re, subprocess, sys, tempfile, time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench
import raytracing_course_hw_b200 as rt

REF = os.path.join(ROOT, "oracle", "_ref")
SC = os.path.join(ROOT, "tests", "scenes")
BIG = os.path.join(REF, "scenes")


def text_copy(path, width, height, samples):
    """Copy of a text scene with DIMENSIONS / SAMPLES replaced (what SURVEY 8d calls a synthetic up-res copy)."""
    s = open(path).read()
    if width:
        s = re.sub(r"DIMENSIONS\s+\d+\s+\d+", f"DIMENSIONS {width} {height}", s)
    if samples:
        s = re.sub(r"SAMPLES\s+\d+", f"SAMPLES {samples}", s)
    f = tempfile.NamedTemporaryFile("w", suffix=".txt", delete=False)
    f.write(s)
    f.close()
    return f.name


def run_ref(hw, scene, width, height, samples, is_text):
    exe = os.path.join(REF, f"hw{hw}")
    if not os.path.exists(exe):
        return None
    env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1))
    scene = os.path.abspath(scene)
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "o.ppm")
        if is_text:
            p = text_copy(scene, width, height, samples)
            argv = [exe, p, out]
        else:
            argv = [exe, scene, str(width), str(height), str(samples), out]
        t = time.perf_counter()
        subprocess.run(["bash", "-c", 'ulimit -s unlimited 2>/dev/null; exec "$0" "$@"'] + argv, check=True, env=env,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, cwd=os.path.dirname(os.path.abspath(scene)),
                       timeout=600)      # the reference's rejection loops are unbounded (hw5 hangs on its fixture beyond 2048 spp)
        dt = time.perf_counter() - t
        if not os.path.exists(out) or os.path.getsize(out) < width * height * 3:
            raise RuntimeError(f"reference hw{hw} wrote no image for {scene}")
        if is_text:
            os.unlink(p)
    return dt


def cpu_rate(hw, scene, is_text, w, h, spp, budget_s):
    """Mpaths/s of the reference: grow the sample until it runs for ~budget_s, subtract the 1x1x1 run."""
    load = run_ref(hw, scene, 1, 1, 1, is_text)
    if load is None:
        return None
    dt = None
    for _ in range(6):
        dt = run_ref(hw, scene, w, h, spp, is_text) - load
        if dt >= 0.3 * budget_s or w * h * spp >= 1 << 28:
            break
        grow = min(8.0, budget_s / max(dt, 1e-3))
        if grow >= 2 and spp > 1 and hw >= 3:
            spp = int(spp * min(grow, 4))
        else:
            f = min(grow, 4) ** 0.5
            w, h = int(w * f), int(h * f)
    return {"mpaths_s": w * h * spp / max(dt, 1e-6) / 1e6, "sample": f"{w}x{h}x{spp}", "seconds": dt, "load_s": load,
            "threads": 1 if hw <= 4 else (os.cpu_count() or 1)}


def gpu_rate(r, scene, w, h, spp, mode, reps):
    sc = rt.load_scene(scene, width=w or None, height=h or None, samples=spp or None, mode=mode)
    t0 = time.perf_counter()
    r.upload(sc)
    r.synchronize()
    upload_s = time.perf_counter() - t0
    W, H, S = sc.width, sc.height, sc.samples
    chunk = max(1, min(S, (1 << 31) // (W * H * 2) if W * H * S > (1 << 30) else S))
    def frame():
        r.clear()
        t = time.perf_counter()
        kms, rays = 0.0, 0
        for b in range(0, S, chunk):
            r.render(b, min(S, b + chunk))
            st = r.stats()
            kms += st["kernel_ms"]; rays += int(st["rays"])
        r.resolve(S)
        return time.perf_counter() - t, kms, rays
    frame()
    best = min((frame() for _ in range(reps)), key=lambda x: x[0])
    paths = W * H * S
    sc.close()
    return {"frame": f"{W}x{H}x{S}", "paths": paths, "seconds": best[0], "kernel_ms": best[1], "mpaths_s": paths / best[0] / 1e6,
            "mrays_s": best[2] / best[0] / 1e6, "rays_per_path": best[2] / paths, "upload_s": upload_s}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true", help="small frames (smoke run of the tool itself)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--cpu-budget", type=float, default=4.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--grid", action="store_true",
                    help="the resolution x spp grids of SURVEY 8d instead of the config list (GPU only: C3 = the five hw4 scenes at "
                         "512^2 / 1080p x 64 .. 4096 spp, C5 = sphere_emissive and the bench scene at 1080p / 4K x 64 .. 1024 spp)")
    a = ap.parse_args()
    q = a.quick
    room = bench.scene_path(270000)
    big = lambda hw, n: os.path.join(BIG, f"hw{hw}", n) if os.path.exists(os.path.join(BIG, f"hw{hw}", n)) else None
    rows = []
    # (config, label, hw (reference binary), scene, is_text, mode, w, h, spp, cpu sample w,h,spp)
    C = []
    for w, h in ((0, 0), (1920, 1080), (3840, 2160), (7680, 4320)):
        C.append(("C1", "hw1/sample.txt", 1, os.path.join(SC, "hw1/sample.txt"), True, 0, w, h, 0, (640, 480, 1)))
    for w, h in ((0, 0), (1920, 1080), (3840, 2160), (7680, 4320)):
        C.append(("C2", "hw2/sample.txt", 2, os.path.join(SC, "hw2/sample.txt"), True, 0, w, h, 0, (332, 255, 1)))
    for k in (1, 2, 3, 4, 5):
        C.append(("C3", f"hw4/practice3_{k}.txt", 4, os.path.join(SC, f"hw4/practice3_{k}.txt"), True, 0, 0, 0, 0, (160, 120, 4)))
    for k in (1, 3):
        C.append(("C3", f"hw4/practice3_{k}.txt", 4, os.path.join(SC, f"hw4/practice3_{k}.txt"), True, 0, 1920, 1080, 256, (160, 120, 4)))
    # hw3 ships hw4's scenes (mode picks the integrator); hw5 has no shipped scene: the test fixture, up-res
    for k in (1, 3):
        C.append(("C3/hw3", f"hw4/practice3_{k}.txt as hw3", 3, os.path.join(SC, f"hw4/practice3_{k}.txt"), True, 3, 1920, 1080, 256, (160, 120, 4)))
    C.append(("C3/hw5", "hw5/mixed.txt", 5, os.path.join(SC, "hw5/mixed.txt"), True, 5, 1920, 1080, 256, (160, 120, 4)))
    for n in ("practice6_1.gltf", "practice6_2.gltf"):
        if big(6, n):
            # the reference's hw6 tree is degenerate (all sort keys equal): practice6_2 takes minutes per 64 k paths on the CPU
            C.append(("C4", "hw6/" + n, 6, big(6, n), False, 6, 1920, 1080, 256, (64, 36, 2) if "6_2" in n else (160, 90, 4)))
    for n in ("practice7_1.gltf", "practice7_2.gltf", "practice7_3.gltf", "practice7_4.gltf"):
        p = big(7, n) or (os.path.join(SC, "hw7", n) if os.path.exists(os.path.join(SC, "hw7", n)) else None)
        if p:
            C.append(("C4", "hw7/" + n, 7, p, False, 7, 1920, 1080, 256, (320, 180, 4)))
    C.append(("C5", "hw8/sphere_emissive.gltf", 8, os.path.join(SC, "hw8/sphere_emissive.gltf"), False, 8, 3840, 2160, 64 if not q else 4, (320, 180, 4)))
    C.append(("C5", "b200room (tools/make_scene.py --tris 270000)", 8, room, False, 8, 3840, 2160, 64 if not q else 4, (320, 180, 4)))
    C.append(("C5", "b200room (tools/make_scene.py --tris 270000)", 8, room, False, 8, 1920, 1080, 256 if not q else 4, None))

    if a.grid:
        C = []
        for k in (1, 2, 3, 4, 5):
            for w, h in ((512, 512), (1920, 1080)):
                for spp in (64, 256, 1024, 4096):
                    C.append(("C3", f"hw4/practice3_{k}.txt", 4, os.path.join(SC, f"hw4/practice3_{k}.txt"), True, 0, w, h, spp, None))
        for label, scene in (("hw8/sphere_emissive.gltf", os.path.join(SC, "hw8/sphere_emissive.gltf")), ("b200room (tools/make_scene.py --tris 270000)", room)):
            for w, h in ((1920, 1080), (3840, 2160)):
                for spp in (64, 256, 1024):
                    C.append(("C5", label, 8, scene, False, 8, w, h, spp, None))
    r = rt.Renderer()
    cpu_cache = {}
    for cfg, label, hw, scene, is_text, mode, w, h, spp, cpu in C:
        if q and w:
            w, h = w // 4, h // 4
            spp = min(spp, 4) if spp else spp
        try:
            g = gpu_rate(r, scene, w, h, spp, mode, 1 if a.grid else 2)
        except Exception as e:
            g = {"error": str(e)}
        row = {"config": cfg, "scene": label, "reference_binary": f"hw{hw}", "gpu": g}
        if not a.no_cpu and cpu is not None:
            key = (hw, scene)
            if key not in cpu_cache:
                try:
                    cpu_cache[key] = cpu_rate(hw, scene, is_text, cpu[0], cpu[1], cpu[2], a.cpu_budget if not q else 0.5)
                except Exception as e:
                    cpu_cache[key] = {"error": str(e)}
            row["cpu_reference"] = cpu_cache[key]
            c = cpu_cache[key]
            if c and "mpaths_s" in c and "mpaths_s" in g:
                row["gpu_over_cpu"] = g["mpaths_s"] / c["mpaths_s"]
        rows.append(row)
        print(json.dumps(row), flush=True)
    r.close()
    doc = {"host_cores": os.cpu_count(), "rows": rows,
           "note": "GPU: wall clock of render + resolve, scene resident, best of 2 frames after a warm-up frame. CPU: unmodified "
                   "reference binary, sample grown to ~cpu-budget seconds, 1x1x1 run subtracted; hw1-hw4 are single-threaded programs."}
    if a.out:
        json.dump(doc, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
