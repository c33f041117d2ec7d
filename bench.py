#!/usr/bin/env python3
"""Headline benchmark: Mpaths/s of the hw8 glTF path tracer on B200 (BASELINE.json metric).

Workload (config.workload): "hw8 largest glTF example: textured metallic-roughness, 4K, 1024 spp".
The reference's own largest hw8 scenes (room, sibenik) ship without geometry (.MISSING_LARGE_BLOBS),
so the scene is the procedural stand-in of the same size class written by tools/make_scene.py
(268 502 triangles, 5 textures, 224 emissive triangles) -- loadable by the unmodified reference too.

A STEP is one pass of the hot path over one batch: the full 3840x2160 frame x `--spp-per-step`
global sample indices (b200rt_render of that range, accumulated on device).  K steps therefore
render K * spp_per_step samples per pixel of the 1024-spp job; Mpaths/s does not depend on the batch.

  value  : kernel-side throughput, scene already resident in HBM, CUDA events on the launch stream
  e2e    : same batch through the C ABI from HOST buffers: b200rt_upload_scene (flatten + H2D) +
           b200rt_render + b200rt_resolve (tonemap + D2H of the RGB8 frame) inside the timed region
  N > 1  : one process per GPU (torchrun), scene replicated, rank r renders its own sample indices
           (weak scaling: per-GPU batch fixed), ONE NCCL reduce of the radiance buffers to rank 0
           after the last step, inside the timed region; time = max over ranks.

--impl reference : the UNMODIFIED reference hw8 binary (oracle/_ref/hw8, built from /root/reference by
oracle/build_ref.sh) on the box's host cores, same scene, bounded sample per step.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

METRIC = "Mpaths/s on hw8 glTF BVH scene (4K, textured metallic-roughness)"
WORKLOAD = "hw8 largest glTF example: textured metallic-roughness, 4K, 1024 spp"


def scene_path(tris):
    import make_scene
    d = os.path.join(tempfile.gettempdir(), f"b200rt_bench_scene_{tris}_{os.getuid()}")
    p = os.path.join(d, "b200room.gltf")
    if not os.path.exists(p):
        tmp = d + f".tmp{os.getpid()}"
        make_scene.build(tmp, tris, 1)
        try:
            os.rename(tmp, d)
        except OSError:
            pass  # another rank won the race
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows, self.proc, self.device = [], None, device

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.device), "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 9 for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def algorithmic_bytes(st):
    """32 B per node box tested, 48 B per triangle record tested, 100 B shading record
    per shaded hit, 16 B per bilinear texture lookup; same for the light-BVH queries."""
    return (32 * (st["node_tests"] + st["light_node_tests"]) + 48 * (st["tri_tests"] + st["light_tri_tests"]) +
            100 * st["shaded_hits"] + 16 * st["tex_lookups"])


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p)).get("dram_bytes_per_launch")
        except Exception:
            return None
    return None


def reference_arm(args, rank):
    """Times the reference's own CPU implementation (hw8 binary) on the host cores."""
    if rank != 0:
        return
    ncores = os.cpu_count() or 1
    gltf = scene_path(args.tris)
    exe = os.path.join(ROOT, "oracle", "_ref", "hw8")
    w, h, spp = args.ref_width, args.ref_height, args.ref_spp
    env = dict(os.environ, OMP_NUM_THREADS=str(ncores))

    if os.path.exists(exe):
        kind = "reference"

        def run(ww, hh, ss):
            with tempfile.TemporaryDirectory() as td:
                t = time.perf_counter()
                subprocess.run(["bash", "-c", 'ulimit -s unlimited 2>/dev/null; exec "$0" "$@"', exe, gltf, str(ww), str(hh), str(ss),
                                os.path.join(td, "o.ppm")], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                return time.perf_counter() - t
        load = min(run(1, 1, 1) for _ in range(2))      # load + BVH build, subtracted
        step = lambda: max(run(w, h, spp) - load, 1e-6)
    else:
        # the oracle port with the reference's minstd stream (byte-identical images, tests/test_oracle.py)
        kind = "port"
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_util as ou
        import raytracing_course_hw_b200 as rt
        sc = rt.load_scene(gltf, width=w, height=h, samples=spp)

        def step():
            t = time.perf_counter()
            ou.port_render(sc, 0)
            return time.perf_counter() - t
    for _ in range(args.warmup):
        step()
    times = [step() for _ in range(args.steps)]
    total = sum(times)
    paths = w * h * spp * args.steps
    v = paths / total / 1e6
    sample = f"{w}x{h}x{spp}spp per step on the same scene; load+BVH time (1x1x1 run) subtracted"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "scene": f"tools/make_scene.py --tris {args.tris}", "sample": sample},
        "cpu_baseline": {"value": v, "unit": "Mpaths/s", "cores": ncores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def cpu_baseline(args):
    """Bounded reference run on the host cores, reported next to the GPU number (rank 0, N = 1)."""
    ncores = os.cpu_count() or 1
    gltf = scene_path(args.tris)
    exe = os.path.join(ROOT, "oracle", "_ref", "hw8")
    w, h, spp = args.ref_width, args.ref_height, args.ref_spp
    sample = f"{w}x{h}x{spp}spp of the same scene, load+BVH (1x1x1 run) subtracted"
    env = dict(os.environ, OMP_NUM_THREADS=str(ncores))
    try:
        if os.path.exists(exe):
            def run(ww, hh, ss):
                with tempfile.TemporaryDirectory() as td:
                    t = time.perf_counter()
                    subprocess.run(["bash", "-c", 'ulimit -s unlimited 2>/dev/null; exec "$0" "$@"', exe, gltf, str(ww), str(hh), str(ss),
                                    os.path.join(td, "o.ppm")], check=True, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
                    return time.perf_counter() - t
            load = run(1, 1, 1)
            t = max(run(w, h, spp) - load, 1e-6)
            kind = "reference"
        else:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_util as ou
            import raytracing_course_hw_b200 as rt
            sc = rt.load_scene(gltf, width=w, height=h, samples=spp)
            t0 = time.perf_counter()
            ou.port_render(sc, 0)
            t = time.perf_counter() - t0
            kind = "port"
        return {"value": w * h * spp / t / 1e6, "unit": "Mpaths/s", "cores": ncores, "kind": kind, "sample": sample}
    except Exception as e:  # the baseline is a reported figure; never fail the GPU bench over it
        return {"value": None, "unit": "Mpaths/s", "cores": ncores, "kind": "unavailable", "sample": f"failed: {e}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=3840)
    ap.add_argument("--height", type=int, default=2160)
    ap.add_argument("--spp-per-step", type=int, default=16)
    ap.add_argument("--tris", type=int, default=270000)
    ap.add_argument("--ref-width", type=int, default=640)
    ap.add_argument("--ref-height", type=int, default=360)
    ap.add_argument("--ref-spp", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import numpy as np
    import torch
    import raytracing_course_hw_b200 as rt
    from raytracing_course_hw_b200 import dist as rtdist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback)")
    if world > 1:   # one process per GPU on one host: share the cores for the host side of the upload
        os.environ.setdefault("B200RT_HOST_THREADS", str(max(2, min(16, (os.cpu_count() or 16) // world))))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL writes its banner ("NCCL version ...", NCCL_DEBUG) to fd 1 when the communicator comes up:
        # route fd 1 to stderr until then so that stdout carries exactly one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            warm = torch.zeros(1, device="cuda")
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    if rank == 0:
        gltf = scene_path(args.tris)
    if dist:
        dist.barrier()
    gltf = scene_path(args.tris)
    W, H, SPP = args.width, args.height, args.spp_per_step
    scene = rt.load_scene(gltf, width=W, height=H, samples=SPP)
    d = scene.desc

    # counters from one counted (untimed) launch of the same step: counts are deterministic
    rc = rt.Renderer(device=local, flags=rt.FLAG_COUNTERS)
    rc.upload(scene)
    rc.render(rank * SPP, (rank + 1) * SPP)
    counted = rc.stats()
    rc.close()

    r = rt.Renderer(device=local)
    r.upload(scene)
    ext = torch.cuda.ExternalStream(r.stream(), device=torch.device("cuda", local))

    from raytracing_course_hw_b200.dist import accum_tensor, reduce_to_rank0

    def step(i):
        # global sample indices: step i, rank `rank` (disjoint across ranks and steps)
        b = (i * world + rank) * SPP
        r.render(b, b + SPP)

    launches = 0
    r.render(rank * SPP, (rank + 1) * SPP)
    r_launches = r.stats()["launches"]      # kernels per step (generate + per-bounce stages + accumulate)
    for i in range(args.warmup):
        step(i)
    r.synchronize()
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    r.clear()
    r.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if dist:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record(ext)
    for i in range(args.steps):
        step(i)
        launches += r_launches
    if dist:
        # ONE reduce of the radiance buffers per frame (north_star), on the render stream
        with torch.cuda.stream(ext):
            reduce_to_rank0(accum_tensor(r, torch.device("cuda", local)))
    e1.record(ext)
    r.synchronize()
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    last = r.stats()               # CUDA-event time of the dominant kernel in the last step
    if dist:
        tt = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms = float(tt.item())
    paths_per_step = W * H * SPP * world
    value = paths_per_step * args.steps / (ms * 1e-3) / 1e6

    # Dominant kernel = wf_extend (closest-hit traversal; largest share in the ncu launch list, profiles/).
    # The timed region keeps two waves in flight on two streams, so CUDA-event brackets around a launch
    # there include time shared with the other wave's kernels.  The kernel's own duration is therefore
    # measured right after the timed region on the same step with ONE wave in flight
    # (FLAG_SERIAL_WAVES: same kernels, same launches, events on the launching stream).
    r.close()
    rs = rt.Renderer(device=local, flags=rt.FLAG_SERIAL_WAVES)
    rs.upload(scene)
    rs.render(rank * SPP, (rank + 1) * SPP)
    kt, et, st_, lt, el = [], [], [], [], 0
    for i in range(min(args.steps, 4)):
        b = ((args.steps + i) * world + rank) * SPP
        rs.render(b, b + SPP)
        s_i = rs.stats()
        kt.append(s_i["kernel_ms"]); et.append(s_i["extend_ms"]); st_.append(s_i["shade_ms"]); lt.append(s_i["light_ms"])
        el = s_i["extend_launches"]
    rs.close()
    k_ms = sum(kt) / len(kt)
    ext_ms = sum(et) / len(et)
    peak, peak_src = measured_peak()
    # 32 B per node box tested + 48 B per triangle record tested, summed over the
    # extend launches of one step (counts from the counted run of the same step)
    ext_bytes = 32 * counted["node_tests"] + 48 * counted["tri_tests"]
    abytes = algorithmic_bytes(counted)
    achieved = ext_bytes / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic(), "kernel": "wf_extend (persistent closest-hit traversal)",
                "kernel_ms_per_launch": ext_ms / max(el, 1), "launches_per_step": el, "kernel_ms_per_step": ext_ms,
                "algorithmic_bytes_per_step": ext_bytes, "algorithmic_bytes_per_launch": ext_bytes / max(el, 1),
                "share_of_step": ext_ms / k_ms if k_ms else None,
                "serial_step": {"ms": k_ms, "extend_ms": ext_ms, "shade_ms": sum(st_) / len(st_), "light_ms": sum(lt) / len(lt),
                                "algorithmic_bytes": abytes, "achieved_gbs": abytes / (k_ms * 1e-3) / 1e9},
                "timed_region_step": {"ms": ms / args.steps, "extend_ms_bracketed": last["extend_ms"],
                                      "shade_ms_bracketed": last["shade_ms"], "light_ms_bracketed": last["light_ms"],
                                      "waves_in_flight": 2},
                "peak_source": peak_src,
                "note": "kernel durations: CUDA events around each launch, one wave in flight (serial_step); the timed region overlaps "
                        "two waves, its brackets include time-sharing. Node and triangle gathers are served by L1/L2 (scene working "
                        "set 47 MB < 126 MB L2), so algorithmic bytes / time can exceed the HBM copy peak: DRAM bytes per launch are in "
                        "`traffic`; ncu shows the kernel issue-bound (profiles/)"}

    # e2e: host scene description -> upload (flatten + H2D) -> render -> resolve -> RGB8 on the host
    r2 = rt.Renderer(device=local)
    out = np.empty((H, W, 3), np.uint8)
    h2d = (d.n_nodes * 32 + d.n_tris * (48 + 96) + d.n_light_tris * (48 + 48) + d.n_light_nodes * 32 +
           sum(d.images[i].width * d.images[i].height * 4 for i in range(d.n_images)))
    d2h = W * H * 3

    if dist and rank == 0:      # the other ranks import rank 0's upload: its host side may use every core now
        os.environ["B200RT_HOST_THREADS"] = str(min(16, os.cpu_count() or 16))

    def e2e_step(i):
        # N > 1: ONE host-side upload per step (rank 0), replicated device to device by an NCCL broadcast
        if dist:
            rtdist.upload_replicated(r2, scene, rank, world, device=torch.device("cuda", local))
        else:
            r2.upload(scene)
        b = (i * world + rank) * SPP
        r2.render(b, b + SPP)
        r2.resolve(SPP, out)

    e2e_step(0)
    if dist:
        dist.barrier()
    t0 = time.perf_counter()
    n_e2e = max(2, min(args.steps, 4))
    for i in range(n_e2e):
        e2e_step(i)
    e2e_s = time.perf_counter() - t0
    if dist:
        tt = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e_s = float(tt.item())
    e2e_value = paths_per_step * n_e2e / e2e_s / 1e6
    r2.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "scene": f"tools/make_scene.py --tris {args.tris} ({d.n_tris} triangles, {d.n_light_tris} emissive, {d.n_images} textures)",
                       "frame": f"{W}x{H}", "spp_per_step": SPP, "ray_depth": 6,
                       "parallelism": f"sample-split x{world}, one NCCL reduce per frame; e2e: rank 0 uploads the scene, NCCL broadcast replicates it" if world > 1 else "single GPU",
                       "l2": "accumulation slab + frame written per step = 199 MB > 126 MB L2; scene reads are L2-resident by design"},
            "mrays_per_s": value * counted["rays"] / max(counted["paths"], 1),
            "rays_per_path": counted["rays"] / max(counted["paths"], 1),
            "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "steps": n_e2e},
            "gpu_launches": launches,
            "clocks": clocks, "roofline": roofline,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
        print(json.dumps(line), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
