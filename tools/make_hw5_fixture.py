#!/usr/bin/env python3
"""Writes tests/scenes/hw5/mixed.txt: a synthetic hw5 text scene (the reference ships none).

Every figure type and material of hw5 is present: planes (scanned), rotated boxes and ellipsoids and triangles in the
BVH, emissive BOX / ELLIPSOID / TRIANGLE figures (all three FiguresMix light kinds), a mirror, a dielectric, and one
triangle that carries its own POSITION and ROTATION.  Deterministic (no random numbers)."""
import math
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(ROOT, "tests", "scenes", "hw5", "mixed.txt")
L = []
L += ["DIMENSIONS 160 120", "BG_COLOR 0.02 0.02 0.03", "CAMERA_POSITION 0 0.2 4.6", "CAMERA_RIGHT 1 0 0", "CAMERA_UP 0 1 0",
      "CAMERA_FORWARD 0 0 -1", "CAMERA_FOV_X 1.05", "RAY_DEPTH 6", "SAMPLES 4"]


def prim(shape, *attrs):
    L.append("NEW_PRIMITIVE")
    L.append(shape)
    L.extend(attrs)


# room: five planes
prim("PLANE 0 1 0", "POSITION 0 -1.5 0", "COLOR 0.7 0.7 0.7")
prim("PLANE 0 -1 0", "POSITION 0 2 0", "COLOR 0.7 0.7 0.7")
prim("PLANE 0 0 1", "POSITION 0 0 -2.5", "COLOR 0.7 0.7 0.7")
prim("PLANE 1 0 0", "POSITION -2.2 0 0", "COLOR 0.75 0.2 0.2")
prim("PLANE -1 0 0", "POSITION 2.2 0 0", "COLOR 0.2 0.3 0.75")
# lights of all three kinds
prim("BOX 0.5 0.05 0.35", "POSITION -0.9 1.9 -0.6", "ROTATION 0 0.2588 0 0.9659", "COLOR 0 0 0", "EMISSION 9 8 7")
prim("ELLIPSOID 0.18 0.12 0.18", "POSITION 1.3 1.2 -1.2", "COLOR 0 0 0", "EMISSION 12 10 6")
prim("TRIANGLE 0.6 1.95 0.9 1.6 1.95 0.9 0.6 1.95 0.1", "COLOR 0 0 0", "EMISSION 6 7 9")
prim("TRIANGLE 1.6 1.95 0.1 0.6 1.95 0.1 1.6 1.95 0.9", "COLOR 0 0 0", "EMISSION 6 7 9")
# analytic bodies
prim("ELLIPSOID 0.55 0.55 0.55", "POSITION -1.1 -0.95 -0.9", "METALLIC", "COLOR 0.95 0.9 0.8")
prim("BOX 0.4 0.6 0.4", "POSITION 1.1 -0.9 -1.3", "ROTATION 0 0.3827 0 0.9239", "COLOR 0.3 0.7 0.35")
prim("ELLIPSOID 0.4 0.5 0.4", "POSITION 0.35 -1.0 0.5", "DIELECTRIC", "IOR 1.5", "COLOR 0.9 0.95 1.0")
# an octahedron of diffuse triangles
c = (-0.2, 0.15, -1.0)
r = 0.55
v = [(c[0] + r, c[1], c[2]), (c[0] - r, c[1], c[2]), (c[0], c[1] + r, c[2]), (c[0], c[1] - r, c[2]), (c[0], c[1], c[2] + r), (c[0], c[1], c[2] - r)]
faces = [(0, 2, 4), (2, 1, 4), (1, 3, 4), (3, 0, 4), (2, 0, 5), (1, 2, 5), (3, 1, 5), (0, 3, 5)]
for k, (a, b, d) in enumerate(faces):
    col = (0.3 + 0.08 * k, 0.8 - 0.07 * k, 0.4 + 0.05 * (k % 3))
    prim("TRIANGLE " + " ".join(f"{x:.4f}" for p in (v[a], v[b], v[d]) for x in p), f"COLOR {col[0]:.3f} {col[1]:.3f} {col[2]:.3f}")
# a mirror triangle pair leaning on the back wall, and a glass triangle
prim("TRIANGLE -2.0 -1.4 -2.3 -0.6 -1.4 -2.3 -2.0 1.2 -2.45", "METALLIC", "COLOR 0.9 0.9 0.9")
prim("TRIANGLE -0.6 1.2 -2.45 -2.0 1.2 -2.45 -0.6 -1.4 -2.3", "METALLIC", "COLOR 0.9 0.9 0.9")
prim("TRIANGLE 1.2 -1.45 1.2 2.0 -1.45 0.6 1.6 0.2 0.9", "DIELECTRIC", "IOR 1.33", "COLOR 0.8 1.0 0.9")
# a fan of small floor-level triangles (deeper BVH)
for k in range(12):
    a0, a1 = 2 * math.pi * k / 12, 2 * math.pi * (k + 1) / 12
    p0 = (0.9 * math.cos(a0) - 0.3, -1.49, 0.9 * math.sin(a0) + 1.6)
    p1 = (0.9 * math.cos(a1) - 0.3, -1.49, 0.9 * math.sin(a1) + 1.6)
    pc = (-0.3, -1.3, 1.6)
    prim("TRIANGLE " + " ".join(f"{x:.4f}" for p in (pc, p1, p0) for x in p), f"COLOR {0.4 + 0.04 * k:.3f} 0.5 {0.9 - 0.04 * k:.3f}")
# a triangle with its own frame: vertices in local space, placed by POSITION + ROTATION
prim("TRIANGLE -0.4 0 0 0.4 0 0 0 0.7 0", "POSITION 1.5 0.3 -0.2", "ROTATION 0.1830 0.6830 0.1830 0.6830", "COLOR 0.85 0.75 0.2")
os.makedirs(os.path.dirname(out), exist_ok=True)
open(out, "w").write("\n".join(L) + "\n")
print("wrote", out, len(L), "lines")
# The planes, the rotated BOX light and the three analytic bodies are valid hw3 / hw4 grammar too: that subset pins the
# BoxLight sampler of the hw4 port (none of the reference's own hw4 scenes has a box light).
blocks = ("\n".join(L) + "\n").split("NEW_PRIMITIVE\n")
sub = blocks[0] + "".join("NEW_PRIMITIVE\n" + b for b in blocks[1:7] + blocks[10:13])
out4 = os.path.join(ROOT, "tests", "golden", "hw4_boxlight_spp4.txt")
open(out4, "w").write(sub)
print("wrote", out4)
