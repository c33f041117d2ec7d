#!/usr/bin/env python3
"""Procedural hw8-compatible glTF scene ("b200room") for the headline benchmark.

Why: BASELINE.json names "the largest hw8 glTF example" (room.gltf, 269 966 triangles) but its
geometry blob is absent from the reference checkout (.MISSING_LARGE_BLOBS), and /root/reference does
not exist on the GPU box anyway.  This script writes a deterministic stand-in of the same size class
that BOTH the unmodified reference hw8 binary and this repo's loader accept: every primitive has
POSITION / NORMAL / TEXCOORD_0 / TANGENT / indices / material, every bufferView has byteOffset,
"images" + "textures"(sampler, source) + "samplers" are present.

Content: a closed room (tessellated, textured walls/floor), an emissive ceiling panel and two small
emissive lamps, and a set of displaced spheres / tori with metallic-roughness, base-colour, emissive
and normal textures.  Textures are procedural PNGs (zlib, no third-party code).

usage: make_scene.py OUT_DIR [--tris 270000] [--seed 1]
"""
import argparse
import json
import os
import struct
import zlib

import numpy as np


# ------------------------------------------------------------------------------------ PNG
def write_png(path, rgb):
    h, w, _ = rgb.shape
    raw = b"".join(b"\x00" + rgb[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def value_noise(rng, size, octaves=5):
    out = np.zeros((size, size), np.float64)
    amp, total = 1.0, 0.0
    for o in range(octaves):
        n = 4 << o
        g = rng.random((n, n))
        g = np.concatenate([g, g[:1]], 0)
        g = np.concatenate([g, g[:, :1]], 1)
        t = np.linspace(0, n, size, endpoint=False)
        i = t.astype(int)
        f = t - i
        f = f * f * (3 - 2 * f)
        a = g[i][:, i] * (1 - f)[None, :] + g[i][:, i + 1] * f[None, :]
        b = g[i + 1][:, i] * (1 - f)[None, :] + g[i + 1][:, i + 1] * f[None, :]
        out += amp * (a * (1 - f)[:, None] + b * f[:, None])
        total += amp
        amp *= 0.5
    return out / total


def make_textures(out_dir, rng):
    size = 1024
    n = value_noise(rng, size)
    yy, xx = np.mgrid[0:size, 0:size]
    # 0: tiles (base colour)
    tile = ((xx // 128 + yy // 128) % 2).astype(np.float64)
    grout = ((xx % 128 < 4) | (yy % 128 < 4)).astype(np.float64)
    base = np.stack([0.75 + 0.2 * tile, 0.7 + 0.15 * n, 0.6 + 0.3 * (1 - tile)], -1) * (1 - 0.6 * grout[..., None])
    write_png(os.path.join(out_dir, "tiles.png"), (np.clip(base, 0, 1) * 255).astype(np.uint8))
    # 1: marble (base colour)
    m = 0.5 + 0.5 * np.sin(xx / 40.0 + 8 * n)
    marble = np.stack([0.85 * m + 0.1, 0.8 * m + 0.12, 0.78 * m + 0.2], -1)
    write_png(os.path.join(out_dir, "marble.png"), (np.clip(marble, 0, 1) * 255).astype(np.uint8))
    # 2: metallic-roughness (G = roughness, B = metallic)
    n2 = value_noise(rng, 512)
    mr = np.stack([np.zeros_like(n2), 0.15 + 0.7 * n2, (n2 > 0.5).astype(np.float64)], -1)
    write_png(os.path.join(out_dir, "metalrough.png"), (np.clip(mr, 0, 1) * 255).astype(np.uint8))
    # 3: normal map from a height field
    hgt = value_noise(rng, 512, 6)
    dx = np.roll(hgt, -1, 1) - np.roll(hgt, 1, 1)
    dy = np.roll(hgt, -1, 0) - np.roll(hgt, 1, 0)
    nrm = np.stack([-dx * 24, -dy * 24, np.ones_like(hgt)], -1)
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    write_png(os.path.join(out_dir, "normal.png"), ((nrm * 0.5 + 0.5) * 255).astype(np.uint8))
    # 4: emissive pattern
    e = (np.sin(xx[:512, :512] / 9.0) * np.sin(yy[:512, :512] / 9.0) > 0).astype(np.float64)
    em = np.stack([0.9 + 0.1 * e, 0.8 * e + 0.15, 0.5 * e + 0.1], -1)
    write_png(os.path.join(out_dir, "emission.png"), (np.clip(em, 0, 1) * 255).astype(np.uint8))
    return ["tiles.png", "marble.png", "metalrough.png", "normal.png", "emission.png"]


# ------------------------------------------------------------------------------------ meshes
def grid(origin, du, dv, nu, nv, uv_scale=1.0):
    """Planar grid; normal = du x dv (CCW seen from the normal side)."""
    origin, du, dv = (np.asarray(a, np.float64) for a in (origin, du, dv))
    u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    pos = origin + u[..., None] * du + v[..., None] * dv
    nrm = np.cross(du, dv)
    nrm /= np.linalg.norm(nrm)
    tan = du / np.linalg.norm(du)
    return _finish(pos, np.broadcast_to(nrm, pos.shape), np.stack([u * uv_scale, v * uv_scale], -1),
                   np.broadcast_to(tan, pos.shape), nu, nv)


def _finish(pos, nrm, uv, tan, nu, nv, flip=False):
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nv + 1, nu + 1)
    a, b, c, d = idx[:-1, :-1], idx[:-1, 1:], idx[1:, 1:], idx[1:, :-1]
    tris = np.stack([np.stack([a, b, c], -1), np.stack([a, c, d], -1)], 2).reshape(-1, 3)
    if flip:
        tris = tris[:, ::-1]
    t4 = np.concatenate([tan.reshape(-1, 3), np.ones((tan.size // 3, 1))], 1)
    return (pos.reshape(-1, 3).astype(np.float32), nrm.reshape(-1, 3).astype(np.float32),
            uv.reshape(-1, 2).astype(np.float32), t4.astype(np.float32), tris.astype(np.uint32))


def param_surface(fn, nu, nv, uv_scale=(1.0, 1.0)):
    """Closed parametric surface with finite-difference normals/tangents (smooth shading); the
    orientation is chosen so that normals (and the CCW winding) point away from the centroid."""
    u, v = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    e = 1e-4
    p = fn(u, v)
    pu = (fn(u + e, v) - fn(u - e, v)) / (2 * e)
    pv = (fn(u, v + e) - fn(u, v - e)) / (2 * e)
    n = np.cross(pu, pv)
    ln = np.linalg.norm(n, axis=-1, keepdims=True)
    n = np.where(ln > 1e-9, n / np.maximum(ln, 1e-12), np.array([0.0, 1.0, 0.0]))
    lt = np.linalg.norm(pu, axis=-1, keepdims=True)
    t = np.where(lt > 1e-9, pu / np.maximum(lt, 1e-12), np.array([1.0, 0.0, 0.0]))
    flip = bool(((p - p.reshape(-1, 3).mean(0)) * n).sum(-1).mean() < 0)
    if flip:
        n = -n
    return _finish(p, n, np.stack([u * uv_scale[0], v * uv_scale[1]], -1), t, nu, nv, flip)


def sphere_fn(center, radius, bump=0.0, freq=6):
    c = np.asarray(center, np.float64)

    def fn(u, v):
        theta, phi = 2 * np.pi * u, np.pi * (v * 0.998 + 0.001)
        r = radius * (1 + bump * np.sin(freq * theta) * np.sin(freq * phi))
        # (u, v) -> outward normal pu x pv requires this orientation
        return c + np.stack([r * np.sin(phi) * np.cos(theta), r * np.cos(phi), -r * np.sin(phi) * np.sin(theta)], -1) * np.array([1, 1, 1.0])
    return fn


def torus_fn(center, R, r, twist=0.0):
    c = np.asarray(center, np.float64)

    def fn(u, v):
        a, b = 2 * np.pi * u, 2 * np.pi * v
        rr = r * (1 + twist * np.sin(3 * a + 2 * b))
        return c + np.stack([(R + rr * np.cos(b)) * np.cos(a), rr * np.sin(b), -(R + rr * np.cos(b)) * np.sin(a)], -1)
    return fn


# ------------------------------------------------------------------------------------ glTF
class Builder:
    def __init__(self):
        self.blob = bytearray()
        self.views, self.accessors, self.meshes, self.nodes = [], [], [], []

    def _acc(self, arr, ctype, kind):
        while len(self.blob) % 4:
            self.blob.append(0)
        off = len(self.blob)
        data = np.ascontiguousarray(arr).tobytes()
        self.blob += data
        self.views.append({"buffer": 0, "byteOffset": off, "byteLength": len(data)})
        acc = {"bufferView": len(self.views) - 1, "componentType": ctype, "count": int(arr.shape[0]) if kind != "SCALAR" else int(arr.size), "type": kind}
        if kind == "VEC3" and ctype == 5126:
            acc["min"] = [float(x) for x in arr.min(0)]
            acc["max"] = [float(x) for x in arr.max(0)]
        self.accessors.append(acc)
        return len(self.accessors) - 1

    def add(self, mesh, material, name, translation=None):
        pos, nrm, uv, tan, tris = mesh
        prim = {"attributes": {"POSITION": self._acc(pos, 5126, "VEC3"), "NORMAL": self._acc(nrm, 5126, "VEC3"),
                               "TEXCOORD_0": self._acc(uv, 5126, "VEC2"), "TANGENT": self._acc(tan, 5126, "VEC4")},
                "indices": self._acc(tris.reshape(-1), 5125, "SCALAR"), "material": material}
        self.meshes.append({"name": name, "primitives": [prim]})
        node = {"name": name, "mesh": len(self.meshes) - 1}
        if translation is not None:
            node["translation"] = [float(x) for x in translation]
        self.nodes.append(node)
        return tris.shape[0]


def build(out_dir, target_tris, seed):
    os.makedirs(out_dir, exist_ok=True)
    rng = np.random.default_rng(seed)
    images = make_textures(out_dir, rng)
    T_TILES, T_MARBLE, T_MR, T_NORMAL, T_EMIT = range(5)

    def mat(name, color=(1, 1, 1), metallic=0.0, rough=0.6, base=None, mr=None, normal=None, emissive=None, etex=None, strength=None):
        m = {"name": name, "pbrMetallicRoughness": {"baseColorFactor": [*map(float, color), 1.0], "metallicFactor": float(metallic), "roughnessFactor": float(rough)}}
        if base is not None:
            m["pbrMetallicRoughness"]["baseColorTexture"] = {"index": base}
        if mr is not None:
            m["pbrMetallicRoughness"]["metallicRoughnessTexture"] = {"index": mr}
        if normal is not None:
            m["normalTexture"] = {"index": normal}
        if emissive is not None:
            m["emissiveFactor"] = [*map(float, emissive)]
        if etex is not None:
            m["emissiveTexture"] = {"index": etex}
        if strength is not None:
            m["extensions"] = {"KHR_materials_emissive_strength": {"emissiveStrength": float(strength)}}
        return m

    materials = [
        mat("floor", base=T_TILES, rough=0.5, normal=T_NORMAL),                       # 0
        mat("wall_white", color=(0.8, 0.8, 0.8), rough=0.9),                          # 1
        mat("wall_red", color=(0.75, 0.2, 0.18), rough=0.9),                          # 2
        mat("wall_blue", color=(0.2, 0.3, 0.75), rough=0.9),                          # 3
        mat("marble", base=T_MARBLE, rough=0.35),                                     # 4
        mat("worn_metal", color=(0.95, 0.8, 0.55), metallic=1.0, rough=1.0, mr=T_MR, normal=T_NORMAL),  # 5
        mat("chrome", color=(0.95, 0.95, 0.95), metallic=1.0, rough=0.12),            # 6
        mat("plastic", color=(0.15, 0.65, 0.3), rough=0.3),                           # 7
        mat("panel", color=(0, 0, 0), emissive=(1.0, 0.95, 0.85), strength=14.0),     # 8
        mat("lamp", color=(0, 0, 0), emissive=(1.0, 0.8, 0.6), etex=T_EMIT, strength=20.0),  # 9
    ]

    b = Builder()
    H = 5.0  # room half size
    fixed = 0
    # room: normals point inward
    wall_n = 24
    fixed += b.add(grid((-H, -H, H), (2 * H, 0, 0), (0, 0, -2 * H), wall_n, wall_n, 4.0), 0, "floor")
    fixed += b.add(grid((-H, H, -H), (2 * H, 0, 0), (0, 0, 2 * H), wall_n, wall_n), 1, "ceiling")
    fixed += b.add(grid((-H, -H, -H), (2 * H, 0, 0), (0, 2 * H, 0), wall_n, wall_n), 1, "back")
    fixed += b.add(grid((-H, -H, H), (0, 0, -2 * H), (0, 2 * H, 0), wall_n, wall_n), 2, "left")
    fixed += b.add(grid((H, -H, -H), (0, 0, 2 * H), (0, 2 * H, 0), wall_n, wall_n), 3, "right")
    fixed += b.add(grid((H, -H, H), (-2 * H, 0, 0), (0, 2 * H, 0), wall_n, wall_n), 1, "front")
    # emitters: ceiling panel (facing down) + two lamps  (~ 2*4*4 + 2*2*8*6 = 224 emissive triangles)
    fixed += b.add(grid((-1.5, H - 0.02, -1.5), (3, 0, 0), (0, 0, 3), 4, 4), 8, "panel")
    fixed += b.add(param_surface(sphere_fn((-3.2, -1.0, -3.0), 0.35), 8, 6), 9, "lamp_a")
    fixed += b.add(param_surface(sphere_fn((3.4, 1.5, -2.0), 0.3), 8, 6), 9, "lamp_b")

    # objects share the remaining budget; each (nu x nv x 2) triangles
    objects = [
        ("bumpy_marble", sphere_fn((-2.2, -3.4, -1.0), 1.6, 0.04, 9), 4, 0.30, (3.0, 2.0)),
        ("worn_metal_torus", torus_fn((2.0, -3.9, 0.5), 1.6, 0.55, 0.15), 5, 0.30, (6.0, 2.0)),
        ("chrome_ball", sphere_fn((0.2, -4.0, 2.2), 1.0), 6, 0.12, (1.0, 1.0)),
        ("plastic_torus", torus_fn((-1.0, 0.5, -3.0), 1.3, 0.4, 0.3), 7, 0.16, (1.0, 1.0)),
        ("marble_ball", sphere_fn((3.2, -0.5, -3.2), 1.1, 0.02, 14), 4, 0.12, (2.0, 1.0)),
    ]
    budget = max(target_tris - fixed, 2000)
    total = fixed
    for name, fn, material, share, uvs in objects:
        n = max(int(np.sqrt(budget * share / 2)), 8)
        total += b.add(param_surface(fn, n, n, uvs), material, name)

    # camera: in the front part of the room looking at -z, slightly down
    ang = -0.12
    cam_q = [float(np.sin(ang / 2)), 0.0, 0.0, float(np.cos(ang / 2))]
    b.nodes.append({"name": "camera", "camera": 0, "translation": [0.0, 0.4, 4.6], "rotation": cam_q})

    gltf = {
        "asset": {"version": "2.0", "generator": "b200rt tools/make_scene.py"},
        "extensionsUsed": ["KHR_materials_emissive_strength"],
        "scene": 0, "scenes": [{"nodes": list(range(len(b.nodes)))}],
        "nodes": b.nodes, "meshes": b.meshes, "materials": materials,
        "cameras": [{"type": "perspective", "perspective": {"yfov": 0.9, "znear": 0.1, "aspectRatio": 16.0 / 9.0}}],
        "samplers": [{"magFilter": 9729, "minFilter": 9987, "wrapS": 10497, "wrapT": 10497}],
        "images": [{"uri": u} for u in images],
        "textures": [{"sampler": 0, "source": i} for i in range(len(images))],
        "accessors": b.accessors, "bufferViews": b.views,
        "buffers": [{"uri": "b200room.bin", "byteLength": len(b.blob)}],
    }
    with open(os.path.join(out_dir, "b200room.bin"), "wb") as f:
        f.write(bytes(b.blob))
    path = os.path.join(out_dir, "b200room.gltf")
    with open(path, "w") as f:
        json.dump(gltf, f)
    return path, total


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("out_dir")
    ap.add_argument("--tris", type=int, default=270000)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    path, n = build(a.out_dir, a.tris, a.seed)
    print(f"{path}: {n} triangles")


if __name__ == "__main__":
    main()
