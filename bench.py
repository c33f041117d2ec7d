#!/usr/bin/env python3
"""Headline benchmark: Mpaths/s of the hw8 glTF path tracer on B200 (BASELINE.json metric).

Workload (config.workload): "hw8 largest glTF example: textured metallic-roughness, 4K, 1024 spp".
The reference's own largest hw8 scenes (room, sibenik) ship without geometry (.MISSING_LARGE_BLOBS),
so the scene is the procedural stand-in of the same size class written by tools/make_scene.py
(268 502 triangles, 5 textures, 224 emissive triangles) -- loadable by the unmodified reference too.

ONE frame is rendered by ALL GPUs (strong scaling, north_star's sample split): a STEP is one batch of
`--spp-per-step` global sample indices of the full 3840x2160 frame; rank g renders the contiguous slice
[g*B/G, (g+1)*B/G) of every batch into its own float buffer.  K steps therefore add K * spp_per_step
samples per pixel to one frame (16 steps of 64 = the named 1024-spp frame).

  value  : scene already resident in HBM; CUDA events on the launch stream around the K steps AND the
           end of the frame: the fused peer-memory reduce + resolve kernel on rank 0 (it reads the other
           ranks' buffers over NVLink through CUDA IPC mappings) and the D2H copy of the RGB8 frame.
  e2e    : ONE whole `--frame-spp` (1024) frame through the public API starting from the .gltf FILE:
           load (parse, PNG decode, triangle assembly) -> b200rt_upload_scene (BVH build, flatten, H2D) on
           rank 0 -> NCCL broadcast of the device-resident scene -> render slices -> fused reduce +
           resolve -> RGB8 on the host.  Wall clock, max over ranks.

--impl reference : the UNMODIFIED reference hw8 binary (oracle/_ref/hw8, built from /root/reference by
oracle/build_ref.sh) on the box's host cores, same scene, a bounded sample per step (640x360x4 spp: cost
is linear in pixels x samples, so Mpaths/s carries over -- an extrapolation, stated in `sample`).
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
    """SURVEY 8(d) with the bytes the device layout defines: one 32-byte DPair record holds BOTH child boxes, and
    the kernels count boxes (node_tests += 2 per record) => 16 B per box tested; 48 B per triangle record tested;
    100 B shading record per shaded hit; 16 B per bilinear texture lookup; same for the light-BVH queries."""
    return (16 * (st["node_tests"] + st["light_node_tests"]) + 48 * (st["tri_tests"] + st["light_tri_tests"]) +
            100 * st["shaded_hits"] + 16 * st["tex_lookups"])


KERNEL_SOURCES = ("wavefront.cuh", "gltf_kernel.cuh", "device_scene.h", "wave_state.h")


def kernel_source_hash():
    """Identity of the traversal kernels' source: profiles/traffic.json stores it at capture time, so a capture of
    an older kernel is flagged instead of being silently reused."""
    import hashlib
    h = hashlib.sha256()
    for f in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "raytracing_course_hw_b200", "csrc", "cuda", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_summary():
    """profiles/traffic.json: the `ncu --set full` capture of wf_extend inside this bench command (tools/ncu_traffic.py)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p))
    except Exception:
        return {}


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
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong",
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
    ap.add_argument("--spp-per-step", type=int, default=64, help="global sample indices per step, split over the GPUs")
    ap.add_argument("--frame-spp", type=int, default=1024, help="samples per pixel of the e2e frame (0 = skip e2e)")
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
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # NCCL writes its banner ("NCCL version ...", NCCL_DEBUG) to fd 1 when the communicator comes up:
        # route fd 1 to stderr until then so that stdout carries exactly one JSON line
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device="cuda")
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)

    def barrier():
        if dist:
            dist.barrier()

    def max_over_ranks(x):
        if not dist:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(d, keys):
        if not dist:
            return d
        t = torch.tensor([float(d[k]) for k in keys], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return dict(d, **{k: int(v) for k, v in zip(keys, t.tolist())})

    gltf = scene_path(args.tris) if rank == 0 else None
    barrier()
    gltf = scene_path(args.tris)
    W, H, B = args.width, args.height, args.spp_per_step
    if B < world:
        raise SystemExit("--spp-per-step must be at least the number of GPUs")
    MODE = rt.MODE_HW8

    def my_slice(first, count):
        """Contiguous slice of the global sample indices [first, first + count) of this rank."""
        return first + count * rank // world, first + count * (rank + 1) // world

    # ---- scene: ONE host-side load + upload (rank 0), replicated device to device ------------------------------
    scene = rt.load_scene(gltf, width=W, height=H, samples=B) if rank == 0 else None
    n_tris, n_lights, n_images = (scene.desc.n_tris, scene.desc.n_light_tris, scene.desc.n_images) if rank == 0 else (0, 0, 0)
    COUNT_KEYS = ["paths", "rays", "light_queries", "node_tests", "tri_tests", "light_node_tests", "light_tri_tests", "shaded_hits", "tex_lookups"]

    # counters of step 0 from one counted (untimed) launch of each rank's slice: counts are deterministic
    rc = rt.Renderer(device=local, flags=rt.FLAG_COUNTERS)
    rtdist.upload_replicated(rc, scene, rank, world, device=dev, frame=(W, H, B, MODE))
    rc.render(*my_slice(0, B))
    counted = sum_over_ranks(rc.stats(), COUNT_KEYS)
    rc.close()

    r = rt.Renderer(device=local)
    scene_bytes = rtdist.upload_replicated(r, scene, rank, world, device=dev, frame=(W, H, B, MODE))
    ext = torch.cuda.ExternalStream(r.stream(), device=dev)
    out = np.empty((H, W, 3), np.uint8)

    def step(i):
        r.render(*my_slice(i * B, B))

    step(0)                                     # allocates the frame buffers the IPC handles point at
    launches_per_step = r.stats()["launches"]
    peers = rtdist.PeerFrame(r, rank, world)    # rank 0 maps the other ranks' accumulation buffers (once)
    for i in range(args.warmup):
        step(i)
    peers.resolve(B * args.warmup, out)
    r.clear()
    r.synchronize()
    barrier()
    torch.cuda.synchronize()

    # ---- timed region: K steps of ONE frame + its end (fused peer reduce + resolve + D2H on rank 0) ------------
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    e0.record(ext)
    for i in range(args.steps):
        step(i)
    peers.resolve(B * args.steps, out)
    e1.record(ext)
    r.synchronize()
    torch.cuda.synchronize()
    barrier()
    clocks = sampler.stop()
    ms = max_over_ranks(e0.elapsed_time(e1))
    last = r.stats()
    launches = launches_per_step * args.steps + (1 if rank == 0 else 0)
    paths_per_step = W * H * B
    value = paths_per_step * args.steps / (ms * 1e-3) / 1e6

    # ---- dominant kernel: wf_extend (closest-hit traversal; largest share in the ncu launch list, profiles/) ----
    # The timed region keeps two waves in flight on two streams, so event brackets around a launch there include time
    # shared with the other wave's kernels.  The kernel's own duration is measured right after the timed region on
    # the same slices with ONE wave in flight (FLAG_SERIAL_WAVES: same kernels, same launches, CUDA events on the
    # launching stream).
    rs = rt.Renderer(device=local, flags=rt.FLAG_SERIAL_WAVES)
    rs.replicate_from(r, None, frame=(W, H, B, MODE))
    rs.render(*my_slice(0, B))
    kt, et, st_, lt, el = [], [], [], [], 0
    for i in range(min(args.steps, 3)):
        rs.render(*my_slice((args.steps + i) * B, B))
        s_i = rs.stats()
        kt.append(s_i["kernel_ms"]); et.append(s_i["extend_ms"]); st_.append(s_i["shade_ms"]); lt.append(s_i["light_ms"])
        el = s_i["extend_launches"]
    rs.close()
    k_ms, ext_ms = sum(kt) / len(kt), sum(et) / len(et)
    peak, peak_src = measured_peak()
    ncu = ncu_summary()
    m = ncu.get("metrics_mean", {})
    stale = bool(ncu) and ncu.get("kernel_source_hash") != kernel_source_hash()
    # algorithmic bytes of this rank's extend launches in one step (counts of the counted run, scaled to the rank's share)
    ext_bytes = (16 * counted["node_tests"] + 48 * counted["tri_tests"]) / world
    ms_per_launch = ext_ms / max(el, 1)
    traffic = ncu.get("dram_bytes_per_launch")
    # the capture profiles launches of the N = 1 step; DRAM bytes scale with the rays of a launch
    traffic_here = traffic * (ext_bytes / max(el, 1)) / ncu["algorithmic_bytes_per_launch"] if traffic and ncu.get("algorithmic_bytes_per_launch") else traffic
    dram_gbs = traffic_here / (ms_per_launch * 1e-3) / 1e9 if traffic_here and ms_per_launch > 0 else None
    issue = m.get("smsp__issue_active.avg.pct_of_peak_sustained_active")
    lanes = m.get("smsp__thread_inst_executed_per_inst_executed.ratio")
    roofline = {
        # SURVEY 8(d) names memory bandwidth as the bounding roofline of this gather-heavy path, so the fraction
        # reported against the HBM peak is the kernel's MEASURED DRAM traffic (ncu) over its live duration.  It is
        # low by design: the scene (57 MB) is L2-resident, DRAM only streams the wave state.  What actually limits
        # the kernel is instruction issue at partial lane occupancy: see `limiter`.
        "bound": "hbm", "achieved": dram_gbs, "peak": peak, "unit": "GB/s", "frac": dram_gbs / peak if dram_gbs else None,
        "traffic": traffic_here, "peak_source": peak_src,
        "kernel": "wf_extend (persistent closest-hit traversal)", "kernel_ms_per_launch": ms_per_launch,
        "launches_per_step": el, "kernel_ms_per_step": ext_ms, "share_of_step": ext_ms / k_ms if k_ms else None,
        "limiter": {"name": "instruction issue x lane occupancy (ncu)", "issue_slots_busy_frac": issue / 100 if issue else None,
                    "active_lanes_of_32": lanes, "useful_lane_issue_frac": issue / 100 * lanes / 32 if issue and lanes else None,
                    "l2_throughput_frac": (m.get("lts__throughput.avg.pct_of_peak_sustained_elapsed") or 0) / 100 or None,
                    "l1tex_throughput_frac": (m.get("l1tex__throughput.avg.pct_of_peak_sustained_elapsed") or 0) / 100 or None,
                    "l1_hit": m.get("l1tex__t_sector_hit_rate.pct"), "l2_hit": m.get("lts__t_sector_hit_rate.pct"),
                    "registers": m.get("launch__registers_per_thread"), "ncu_capture_stale": stale,
                    "source": "profiles/traffic.json (ncu --set full of this command; static, re-captured when the kernels change)"},
        "algorithmic": {"bytes_per_launch": ext_bytes / max(el, 1), "bytes_per_step": ext_bytes,
                        "gbs": ext_bytes / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None,
                        "definition": "32 B per DPair record (two child boxes) + 48 B per triangle record tested; served mostly by L1/L2, NOT a DRAM rate"},
        "serial_step": {"ms": k_ms, "extend_ms": ext_ms, "shade_ms": sum(st_) / len(st_), "light_ms": sum(lt) / len(lt),
                        "algorithmic_bytes": algorithmic_bytes(counted) / world},
        "timed_region_step": {"ms": ms / args.steps, "extend_ms_bracketed": last["extend_ms"], "shade_ms_bracketed": last["shade_ms"],
                              "light_ms_bracketed": last["light_ms"], "waves_in_flight": 2},
    }

    # ---- e2e: ONE whole frame from the .gltf FILE to RGB8 on the host -------------------------------------------
    e2e = None
    F = args.frame_spp
    if F >= world:
        r2 = rt.Renderer(device=local)
        rtdist.upload_replicated(r2, scene, rank, world, device=dev, frame=(W, H, F, MODE))
        r2.render(*my_slice(0, 1 * world))                     # warm: buffers, IPC mappings, resolve path
        peers2 = rtdist.PeerFrame(r2, rank, world)
        peers2.resolve(world, out)
        stages = {}

        def lap(name, t_prev):
            t = time.perf_counter()
            stages[name] = stages.get(name, 0.0) + (t - t_prev) * 1e3
            return t

        barrier()
        t0 = t = time.perf_counter()
        sc2 = rt.load_scene(gltf, width=W, height=H, samples=F) if rank == 0 else None
        t = lap("load_ms", t)
        h2d = rtdist.upload_replicated(r2, sc2, rank, world, device=dev, frame=(W, H, F, MODE))
        r2.synchronize()
        t = lap("upload_and_replicate_ms", t)
        r2.render(*my_slice(0, F))
        r2.synchronize()
        t = lap("render_ms", t)
        peers2.resolve(F, out)
        t = lap("reduce_resolve_d2h_ms", t)
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        peers2.close()
        r2.close()
        e2e = {"value": W * H * F / e2e_s / 1e6, "unit": "Mpaths/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(W * H * 3),
               "steps": 1, "frame": f"{W}x{H}x{F}spp", "seconds": e2e_s, "stages_rank0": stages,
               "from": "the .gltf file on disk (parse + PNG decode + triangle assembly) to RGB8 in host memory"}
    peers.close()
    r.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "scene": f"tools/make_scene.py --tris {args.tris} ({n_tris} triangles, {n_lights} emissive, {n_images} textures)",
                       "frame": f"{W}x{H}", "spp_per_step": B, "frame_spp_timed": B * args.steps, "ray_depth": 6,
                       "parallelism": (f"sample-split x{world} of ONE frame: rank g renders [g*B/G,(g+1)*B/G) of every {B}-sample batch; one fused peer-memory "
                                       "reduce + resolve kernel on rank 0 (CUDA IPC mappings over NVLink) per frame, inside the timed region; "
                                       "scene replicated by one NCCL broadcast of rank 0's device blob") if world > 1 else "single GPU",
                       "l2": "per step every wave streams > 1 GB of path state; accumulation + frame = 124 MB; scene reads are L2-resident by design"},
            "mrays_per_s": value * counted["rays"] / max(counted["paths"], 1),
            "rays_per_path": counted["rays"] / max(counted["paths"], 1),
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
        }
        if e2e:
            line["e2e"] = e2e
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args)
            if line["cpu_baseline"].get("value"):
                line["vs_cpu_baseline"] = {"value_ratio": value / line["cpu_baseline"]["value"],
                                           "e2e_ratio": e2e["value"] / line["cpu_baseline"]["value"] if e2e else None,
                                           "note": "CPU arm = bounded sample of the same scene, extrapolated linearly in pixels x samples"}
        print(json.dumps(line), flush=True)
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
