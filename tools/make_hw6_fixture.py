#!/usr/bin/env python3
"""tests/scenes/hw6/practice7_1_glass.gltf: the hw7 Cornell-box fixture re-materialised for hw6 mode.

The two boxes get baseColorFactor alpha 0.5, which the hw6 loader maps to DIELECTRIC (ior 1.5,
hw6/src/sceneio.cpp:179-183); the walls without metallicFactor are mirrors under hw6 rules and the
ones with metallicFactor 0 diffuse, so one scene exercises all three branches of the integrator.
The boxes are lifted 5 cm off the floor: coplanar glass/floor faces would make the winner of the
closest-hit tie depend on the rounding of t (plane equation in the reference, Moeller-Trumbore on the
device), which changes the event sequence of a path although not the radiance statistics.
The geometry buffer is shared with the hw7 fixture (uri ../hw7/practice7_1.bin)."""
import json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
g = json.load(open(os.path.join(ROOT, "tests/scenes/hw7/practice7_1.gltf")))
for m in g["materials"]:
    if m.get("name") == "Cube":
        m["pbrMetallicRoughness"]["baseColorFactor"] = [0.9, 0.95, 1.0, 0.5]
for n in g["nodes"]:
    if n.get("name", "").startswith("Cube"):
        n["translation"][1] += 0.05
for b in g["buffers"]:
    b["uri"] = "../hw7/" + b["uri"]
os.makedirs(os.path.join(ROOT, "tests/scenes/hw6"), exist_ok=True)
json.dump(g, open(os.path.join(ROOT, "tests/scenes/hw6/practice7_1_glass.gltf"), "w"), indent=1)
