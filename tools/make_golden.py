#!/usr/bin/env python3
"""Generate tests/golden/* by running the UNMODIFIED reference binaries (oracle/_ref/hwN, built by
oracle/build_ref.sh from /root/reference).  Run in the build container only; the fixtures are
committed because neither /root/reference nor (necessarily) oracle/_ref exist on the GPU box.

Fixtures:
  hw1_sample.ppm.gz, hw2_sample.ppm.gz            deterministic configs, full size
  hw4_p3_{k}.txt + hw4_p3_{k}_spp4.ppm.gz         hw4 scenes shrunk to 160x120, SAMPLES 4 (minstd stream)
  hw4_p3_{k}_spp4096.ppm.gz                       ground truth for the RMSE tests (160x120, SAMPLES 4096)
  hw4_boxlight_spp4 / hw3_boxlight_spp4           planes + rotated BOX light + analytic bodies (subset of the hw5 fixture)
  hw8_sphere_emissive_env_128x96_spp{4,4096}      hw8 with sphere_emission.png as the environment map
  hw5_mixed_spp{4,2048}.ppm.gz                    the hw5 binary on tests/scenes/hw5/mixed.txt (tools/make_hw5_fixture.py)
  hw3_p3_{k}_spp{4,4096}.ppm.gz                   the hw3 binary on the same shrunk inputs (hw3 ships the same scenes)
  hw7_{name}_128x96_spp4.ppm.gz / _spp4096        hw7 binaries on practice7_1, practice7_4
  hw8_sphere_emissive_128x96_spp4.ppm.gz / _spp4096
  hw8_room20k_128x96_spp{4,4096}.ppm.gz + .sha    hw8 binary on the 20 k-triangle room of tools/make_scene.py (seed 1): the only
                                                  fixture with baseColor, metallicRoughness and normal-map textures; the
                                                  .sha file pins the generated scene files (regenerated at test time)
  hw6_{name}_128x96_spp4.ppm.gz / _spp4096        hw6 binary on practice7_1 (mirrors + diffuse) and on
                                                  hw6/practice7_1_glass (same geometry, cube alpha 0.5 => dielectric)
"""
import gzip, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref")
GOLD = os.path.join(ROOT, "tests", "golden")
SC = os.path.join(ROOT, "tests", "scenes")


def run(binary, args, out_name):
    dst = os.path.join(GOLD, out_name + ".ppm.gz")
    if os.path.exists(dst) and "--force" not in sys.argv:
        return
    with tempfile.TemporaryDirectory() as td:
        tmp = os.path.join(td, "o.ppm")
        cmd = ["bash", "-c", 'ulimit -s unlimited; exec "$0" "$@"', os.path.join(REF, binary)] + [a if a != "@OUT" else tmp for a in args]
        subprocess.run(cmd, check=True, stderr=subprocess.DEVNULL)
        with open(tmp, "rb") as f, gzip.GzipFile(dst, "wb", mtime=0) as g:
            g.write(f.read())
    print("wrote", dst, flush=True)


def main():
    os.makedirs(GOLD, exist_ok=True)
    run("hw1", [os.path.join(SC, "hw1/sample.txt"), "@OUT"], "hw1_sample")
    run("hw2", [os.path.join(SC, "hw2/sample.txt"), "@OUT"], "hw2_sample")
    for k in (1, 3, 4):
        src = open(os.path.join(SC, f"hw4/practice3_{k}.txt")).read()
        src = re.sub(r"DIMENSIONS \d+ \d+", "DIMENSIONS 160 120", src)
        for spp in (4, 4096):
            txt = os.path.join(GOLD, f"hw4_p3_{k}_spp{spp}.txt")
            open(txt, "w").write(re.sub(r"SAMPLES \d+", f"SAMPLES {spp}", src))
        run("hw4", [os.path.join(GOLD, f"hw4_p3_{k}_spp4.txt"), "@OUT"], f"hw4_p3_{k}_spp4")
        # hw3 ships the same scenes (only SAMPLES differs): the shrunk hw4 inputs feed the hw3 binary too
        run("hw3", [os.path.join(GOLD, f"hw4_p3_{k}_spp4.txt"), "@OUT"], f"hw3_p3_{k}_spp4")
    # box-light subset of the hw5 fixture under hw4 and hw3 rules; the hw5 fixture itself (tools/make_hw5_fixture.py)
    run("hw4", [os.path.join(GOLD, "hw4_boxlight_spp4.txt"), "@OUT"], "hw4_boxlight_spp4")
    run("hw3", [os.path.join(GOLD, "hw4_boxlight_spp4.txt"), "@OUT"], "hw3_boxlight_spp4")
    run("hw5", [os.path.join(GOLD, "hw4_boxlight_spp4.txt"), "@OUT"], "hw5_boxlight_spp4")      # hw5 rules, no triangle at all
    hw5 = open(os.path.join(SC, "hw5/mixed.txt")).read()
    # 2048, not 4096: between 2048 and 3072 samples some pixel's generator reaches a state from which the reference's
    # unbounded rejection loop (BoxLight / EllipsoidLight::sample, hw5/src/include/distributions.h:88-104) never
    # returns -- the hw5 binary then spins forever on this scene (observed; the port caps the loop for the counter RNG)
    open(os.path.join(GOLD, "hw5_mixed_spp2048.txt"), "w").write(re.sub(r"SAMPLES \d+", "SAMPLES 2048", hw5))
    run("hw5", [os.path.join(SC, "hw5/mixed.txt"), "@OUT"], "hw5_mixed_spp4")
    run("hw5", [os.path.join(GOLD, "hw5_mixed_spp2048.txt"), "@OUT"], "hw5_mixed_spp2048")
    for name, path, binary in (("hw7_practice7_1", "hw7/practice7_1.gltf", "hw7"),
                               ("hw7_practice7_4", "hw7/practice7_4.gltf", "hw7"),
                               ("hw8_sphere_emissive", "hw8/sphere_emissive.gltf", "hw8"),
                               ("hw6_practice7_1", "hw7/practice7_1.gltf", "hw6"),
                               ("hw6_practice7_1_glass", "hw6/practice7_1_glass.gltf", "hw6")):
        for spp in (4, 4096):
            run(binary, [os.path.join(SC, path), "128", "96", str(spp), "@OUT"], f"{name}_128x96_spp{spp}")
    # hw8 with an environment map (argv[6], hw8/src/main.cpp:12-14): the fixture's emission texture doubles as the
    # equirectangular image -- misses then return texels instead of black
    env = os.path.join(SC, "hw8/sphere_emission.png")
    for spp in (4, 4096):
        run("hw8", [os.path.join(SC, "hw8/sphere_emissive.gltf"), "128", "96", str(spp), "@OUT", env], f"hw8_sphere_emissive_env_128x96_spp{spp}")
    # textured room (tools/make_scene.py is deterministic: the scene is regenerated wherever the tests run)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_util as ou
    room = ou.room_scene(20000)
    open(os.path.join(GOLD, "hw8_room20k.sha"), "w").write(ou.room_digest(room) + "\n")
    for spp in (4, 4096):
        run("hw8", [room, "128", "96", str(spp), "@OUT"], f"hw8_room20k_128x96_spp{spp}")
    for k in (1, 3, 4):  # slow (single-threaded reference): last
        run("hw4", [os.path.join(GOLD, f"hw4_p3_{k}_spp4096.txt"), "@OUT"], f"hw4_p3_{k}_spp4096")
        run("hw3", [os.path.join(GOLD, f"hw4_p3_{k}_spp4096.txt"), "@OUT"], f"hw3_p3_{k}_spp4096")


if __name__ == "__main__":
    main()
