"""Build recipes: explicit nvcc / g++ invocations, outputs IN-TREE under raytracing_course_hw_b200/lib
(the .so files are git-ignored but travel to the GPU box with the snapshot).

  libb200rt.so       CUDA render core, sm_100a only (C ABI: include/b200rt.h)
  libb200rt_host.so  host scene library: parsers, SAH BVH build, PPM (C ABI: include/b200rt_host.h)
  b200rt             CLI with the reference's argv conventions (run.sh)
  oracle/_ref/*      TEST INFRASTRUCTURE: unmodified reference binaries + the CPU port
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "lib")
INC = os.path.join(ROOT, "include")
ORACLE = os.path.join(ROOT, "oracle")
# NOT $CXX: this image exports CXX=/opt/gcc/bin/g++, which lacks libgomp.spec
GXX = os.environ.get("B200RT_CXX", "/usr/bin/g++")
NVCC = os.environ.get("B200RT_NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
STB_DIR = os.environ.get("B200RT_STB_DIR", "/root/reference/stb")


def _newer(srcs, out):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(s) > t for s in srcs if os.path.exists(s))


def _run(cmd, **kw):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, **kw)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise RuntimeError("build step failed: " + cmd[0])
    return r.stdout


def _all_under(d, exts):
    out = []
    for base, _, files in os.walk(d):
        out += [os.path.join(base, f) for f in files if f.endswith(exts)]
    return out


def build_cuda(force=False, verbose=False):
    """libb200rt.so: every CUDA kernel, compiled for sm_100a with -lineinfo."""
    os.makedirs(LIB, exist_ok=True)
    cu = os.path.join(CSRC, "cuda")
    out = os.path.join(LIB, "libb200rt.so")
    deps = _all_under(cu, (".cu", ".cuh", ".h")) + _all_under(os.path.join(CSRC, "common"), (".h",)) + _all_under(INC, (".h",))
    if not force and not _newer(deps, out):
        return out
    common = [NVCC] + ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-I", INC, "-I", cu,
                              "-ccbin", GXX, "-diag-suppress", "550"]
    common += os.environ.get("B200RT_NVCC_DEFS", "").split()      # development sweeps only
    if verbose:
        common += ["-Xptxas", "-v"]
    objs = []
    # analytic kernels: hw1 / hw2 without FMA contraction (bit-exact), the Monte-Carlo modes of the same file as a
    # second unit with the default contraction; path tracer: default FMA
    units = (("b200rt.cu", [], None), ("upload.cu", [], None), ("scheduler.cu", [], None), ("multi_gpu.cu", [], None),
             ("noalloc.cu", [], None), ("analytic_kernels.cu", ["-fmad=false"], None),
             ("analytic_kernels.cu", ["-DB200RT_ANALYTIC_MC_UNIT"], "analytic_mc.o"))
    procs = []
    for src, extra, obj_name in units:        # the units compile concurrently (scheduler.cu holds the big kernels)
        obj = os.path.join(LIB, obj_name or src.replace(".cu", ".o"))
        cmd = common + extra + ["-c", os.path.join(cu, src), "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for cmd, pr in procs:
        log = pr.communicate()[0]
        if pr.returncode != 0:
            sys.stderr.write(" ".join(cmd) + "\n" + log + "\n")
            raise RuntimeError("build step failed: " + cmd[0])
        if verbose:
            print(log)
    _run([NVCC] + ARCH + ["-shared", "-ccbin", GXX, "-o", out] + objs)
    return out


def build_host(force=False):
    """libb200rt_host.so: text/glTF loaders + SAH BVH builder (no CUDA)."""
    os.makedirs(LIB, exist_ok=True)
    hd = os.path.join(CSRC, "host")
    out = os.path.join(LIB, "libb200rt_host.so")
    srcs = [os.path.join(hd, f) for f in ("text_scene.cpp", "gltf_scene.cpp", "host_api.cpp", "png_decode.cpp")]
    if not force and not _newer(_all_under(hd, (".cpp", ".h")) + _all_under(os.path.join(CSRC, "common"), (".h",)) + _all_under(INC, (".h",)), out):
        return out
    flags = ["-O2", "-std=c++17", "-Wall", "-Wextra", "-fPIC", "-shared", "-I", INC]
    if os.path.exists(os.path.join(STB_DIR, "stb_image.h")):
        # third-party decoder for JPEG & co (PNG is decoded by the built-in decoder either way)
        flags += ["-DB200RT_HAVE_STB", "-I", STB_DIR]
    _run([GXX] + flags + srcs + ["-o", out])
    return out


def build_cli(force=False):
    os.makedirs(LIB, exist_ok=True)
    src = os.path.join(CSRC, "cli", "main.cpp")
    out = os.path.join(LIB, "b200rt")
    if not os.path.exists(src):
        return None
    if not force and not _newer([src] + _all_under(INC, (".h",)), out) and \
            not _newer([os.path.join(LIB, "libb200rt.so"), os.path.join(LIB, "libb200rt_host.so")], out):
        return out
    _run([GXX, "-O2", "-std=c++17", "-Wall", "-Wextra", "-I", INC, src, "-L", LIB, "-lb200rt", "-lb200rt_host",
          "-Wl,-rpath,$ORIGIN", "-o", out])
    return out


def build_oracle(force=False):
    """TEST INFRASTRUCTURE: the CPU port (always) and the unmodified reference (when /root/reference exists)."""
    ref_out = os.path.join(ORACLE, "_ref")
    os.makedirs(ref_out, exist_ok=True)
    out = os.path.join(ref_out, "liboracle_port.so")
    srcs = [os.path.join(ORACLE, f) for f in ("port_whitted.cpp", "port_mc.cpp", "port_api.cpp")]
    deps = srcs + [os.path.join(ORACLE, "port_common.h"), os.path.join(ORACLE, "port_figures.inl")] + _all_under(INC, (".h",))
    if force or _newer(deps, out):
        _run([GXX, "-O3", "-DNDEBUG", "-std=c++17", "-fopenmp", "-fPIC", "-shared", "-Wall", "-Wextra"] + srcs + ["-o", out])
    # CPU walk of the flattened device BVH (SURVEY 8d counter cross-check): plain C, no FMA contraction
    pt_src, pt_out = os.path.join(ORACLE, "pairs_traverse.c"), os.path.join(ref_out, "liboracle_pairs.so")
    if force or _newer([pt_src], pt_out):
        _run([os.environ.get("B200RT_CC", "/usr/bin/gcc"), "-O2", "-ffp-contract=off", "-std=c11", "-fPIC", "-shared", "-Wall", "-Wextra",
              pt_src, "-lm", "-o", pt_out])
    _run(["bash", os.path.join(ORACLE, "build_ref.sh")])
    return out


def build_all(force=False, verbose=False):
    build_cuda(force, verbose)
    build_host(force)
    build_cli(force)
    build_oracle(force)


if __name__ == "__main__":
    build_all(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print("built into", LIB)
