"""Host library + C-ABI surface, no GPU needed."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle_util as ou
import raytracing_course_hw_b200 as rt

S = ou.SCENES
ROOT = ou.ROOT


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    return sorted(set(re.findall(r"\b(b200rt_[a-z_0-9]+)\s*\(", text)))


def test_device_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(rt.LIB_DIR, "libb200rt.so"))
    names = _declared("b200rt.h")
    assert set(rt.DEVICE_SYMBOLS) <= set(names)
    for n in names:
        assert hasattr(lib, n), n


def test_host_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(rt.LIB_DIR, "libb200rt_host.so"))
    names = [n for n in _declared("b200rt_host.h") if n.startswith("b200rt_host_")]
    assert set(rt.HOST_SYMBOLS) <= set(names)
    for n in names:
        assert hasattr(lib, n), n


def test_no_cpu_fallback_without_device():
    lib = rt.device_lib()
    if lib.b200rt_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(rt.B200rtError) as e:
        rt.Renderer()
    assert e.value.code == -2


def test_text_parser_modes_and_quirks(tmp_path, capfd):
    sc = rt.load_scene(os.path.join(S, "hw1/sample.txt"))
    assert "UNKNOWN COMMAND: YA" in capfd.readouterr().err   # hw1/sample.txt:25 junk line
    assert sc.mode == rt.MODE_HW1
    assert rt.load_scene(os.path.join(S, "hw2/sample.txt")).mode == rt.MODE_HW2
    sc4 = rt.load_scene(os.path.join(S, "hw4/practice3_3.txt"))
    assert (sc4.mode, sc4.samples, sc4.ray_depth, sc4.width) == (rt.MODE_HW4, 64, 6, 512)
    # TRIANGLE figures make it hw5: non-planes first in BVH order, planes last, FiguresMix light list
    sc5 = rt.load_scene(os.path.join(S, "hw5/mixed.txt"))
    d5 = sc5.desc
    assert (sc5.mode, d5.n_prims, d5.n_bvh_prims, d5.n_light_prims) == (rt.MODE_HW5, 36, 31, 4)
    assert d5.n_nodes >= 2 * 31 - 1 - 31 and d5.n_light_nodes >= 1
    assert rt.load_scene(os.path.join(S, "hw4/practice3_3.txt"), mode=rt.MODE_HW3).mode == rt.MODE_HW3
    # malformed ROTATION (comma) leaves the remaining fields at their defaults
    p = tmp_path / "q.txt"
    p.write_text("DIMENSIONS 4 4\nCAMERA_FOV_X 1\nNEW_PRIMITIVE\nBOX 1 1 1\nROTATION 0, 0.38 0 0.92\nCOLOR 1 0 0\nBOGUS 1\n")
    sc = rt.load_scene(str(p))
    assert "UNKNOWN COMMAND: BOGUS" in capfd.readouterr().err
    assert sc.mode == rt.MODE_HW1


def test_missing_files_fail_cleanly(tmp_path):
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(tmp_path / "nope.txt"))
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(tmp_path / "nope.gltf"), width=4, height=4, samples=1)
    bad = tmp_path / "bad.gltf"
    bad.write_text("{ not json")
    with pytest.raises(rt.B200rtError):
        rt.load_scene(str(bad), width=4, height=4, samples=1)


def test_ppm_writer_format(tmp_path):
    img = (np.arange(5 * 7 * 3) % 256).astype(np.uint8).reshape(5, 7, 3)
    out = tmp_path / "o.ppm"
    rt.write_ppm(str(out), img)
    data = out.read_bytes()
    assert data.startswith(b"P6\n7 5\n255\n")          # hw8/src/sceneio.cpp:383-385
    assert np.array_equal(ou.parse_ppm(data), img)


def test_png_decoder_matches_pil():
    PIL = pytest.importorskip("PIL.Image")
    path = os.path.join(S, "hw8/sphere_emission.png")
    ref = np.asarray(PIL.open(path).convert("RGB"))
    sc = rt.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=8, height=8, samples=1)
    d = sc.desc
    assert d.n_images == 1 and d.n_tris == 970 and d.n_light_tris == 960
    img = d.images[0]
    mine = np.ctypeslib.as_array(img.rgb, shape=(img.height, img.width, 3))
    assert (img.height, img.width) == ref.shape[:2]
    assert np.array_equal(mine, ref)


def test_default_gltf_load_has_no_trees_and_the_same_triangles():
    """b200rt_host_load_gltf: no reference SAH build -- triangles in file order, light set = the emissive ones in
    file order, nodes / light_nodes NULL.  The reference-ordered load holds the same triangles in another order."""
    p = os.path.join(S, "hw7/practice7_4.gltf")
    a = rt.load_scene(p, width=8, height=8, samples=1)
    b = rt.load_scene(p, width=8, height=8, samples=1, reference_bvh=True)
    da, db = a.desc, b.desc
    assert da.n_nodes == 0 and not da.nodes and da.n_light_nodes == 0 and not da.light_nodes
    assert db.n_nodes > 0 and db.nodes and db.n_light_nodes > 0
    assert da.n_tris == db.n_tris and da.n_light_tris == db.n_light_tris > 0
    tri_bytes = 148

    def rows(ptr, n):
        raw = ctypes.string_at(ptr, n * tri_bytes)
        return sorted(raw[i * tri_bytes:(i + 1) * tri_bytes] for i in range(n))
    assert rows(da.tris, da.n_tris) == rows(db.tris, db.n_tris)
    assert rows(da.light_tris, da.n_light_tris) == rows(db.light_tris, db.n_light_tris)


def test_port_with_private_trees_agrees_with_reference_trees():
    """The oracle port traverses private trees when the description has none: with the light set's ORDER made equal
    (one emitter) the two loads render the same Philox image up to tie / summation-order rounding."""
    import oracle_util as ou
    p = os.path.join(S, "hw7/practice7_1.gltf")
    a = rt.load_scene(p, width=48, height=32, samples=4)
    b = rt.load_scene(p, width=48, height=32, samples=4, reference_bvh=True)
    ia, ib = ou.port_render(a, 1)["accum"], ou.port_render(b, 1)["accum"]
    assert abs(ia.mean() - ib.mean()) <= 0.05 * ib.mean()


def test_generated_tangents_are_unit_and_orthogonal_to_the_normals():
    sc = rt.load_scene(os.path.join(S, "hw7/practice7_1.gltf"), width=8, height=8, samples=1, mode=rt.MODE_HW8)
    d = sc.desc
    raw = np.frombuffer(ctypes.string_at(d.tris, d.n_tris * 148), np.uint8).reshape(d.n_tris, 148)
    f = raw[:, :144].copy().view(np.float32).reshape(d.n_tris, 36)
    n = f[:, 15:24].reshape(-1, 3, 3)
    t = f[:, 24:36].reshape(-1, 3, 4)
    assert np.allclose(np.linalg.norm(t[..., :3], axis=2), 1, atol=1e-4)
    assert np.abs((t[..., :3] * n).sum(axis=2)).max() < 1e-3
    assert np.all(t[..., 3] == 1)


def test_png_writer_round_trips_through_the_decoder_and_rejects_truncation(tmp_path):
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    p = tmp_path / "a.png"
    rt.write_png(str(p), img)
    sc = rt.load_scene(os.path.join(S, "hw8/sphere_emissive.gltf"), width=8, height=8, samples=1)
    sc.set_env_map(str(p))
    env = sc.desc.env_map.contents
    assert (env.width, env.height) == (53, 37)
    assert np.array_equal(np.ctypeslib.as_array(env.rgb, shape=(37, 53, 3)), img)
    # the shipped PNG (deflate with dynamic Huffman blocks) still decodes to the committed texels
    sc.set_env_map(os.path.join(S, "hw8/sphere_emission.png"))
    env = sc.desc.env_map.contents
    tex = np.ctypeslib.as_array(env.rgb, shape=(env.height, env.width, 3))
    import zlib
    assert zlib.crc32(tex.tobytes()) == zlib.crc32(_stb_free_decode(os.path.join(S, "hw8/sphere_emission.png")).tobytes())
    data = p.read_bytes()
    (tmp_path / "cut.png").write_bytes(data[:len(data) // 2])
    with pytest.raises(rt.B200rtError):
        sc.set_env_map(str(tmp_path / "cut.png"))
    # a header announcing a huge image over a tiny stream is rejected before any allocation
    bad = bytearray(data)
    bad[16:24] = (30000).to_bytes(4, "big") + (30000).to_bytes(4, "big")
    (tmp_path / "huge.png").write_bytes(bytes(bad))
    with pytest.raises(rt.B200rtError):
        sc.set_env_map(str(tmp_path / "huge.png"))


def _stb_free_decode(path):
    """Independent PNG decode with zlib + numpy (8-bit RGB / RGBA, non-interlaced) to check the built-in decoder."""
    import struct, zlib
    raw = open(path, "rb").read()
    pos, idat, w = 8, b"", None
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    ch = {2: 3, 6: 4, 0: 1}[ctype]
    assert depth == 8
    data = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, w * ch + 1)
    out = np.zeros((h, w * ch), np.uint8)
    prev = np.zeros(w * ch, np.int32)
    for y in range(h):
        ft, line = int(data[y, 0]), data[y, 1:].astype(np.int32)
        cur = np.zeros(w * ch, np.int32)
        if ft == 0:
            cur = line
        elif ft == 2:
            cur = (line + prev) & 255
        else:
            for i in range(w * ch):
                a = cur[i - ch] if i >= ch else 0
                b = prev[i]
                c = prev[i - ch] if i >= ch else 0
                if ft == 1:
                    pr = a
                elif ft == 3:
                    pr = (a + b) >> 1
                else:
                    pa, pb, pc = abs(b - c), abs(a - c), abs(a + b - 2 * c)
                    pr = a if pa <= pb and pa <= pc else (b if pb <= pc else c)
                cur[i] = (line[i] + pr) & 255
        out[y] = cur
        prev = cur
    out = out.reshape(h, w, ch)
    return np.repeat(out, 3, axis=2) if ch == 1 else out[:, :, :3]


def test_gltf_loader_survives_mutated_files(tmp_path):
    """The reference reads missing JSON members through compiled-out asserts and indexes buffers unchecked
    (hw8/src/sceneio.cpp:28-37, 82-98: undefined behaviour on malformed files); this loader promises an error code.
    300 seeded mutations of practice7_4.gltf (byte flips, deletions, truncations, number edits; a truncated .bin;
    index / count / offset values pushed out of range) are loaded in a SUBPROCESS: it must exit normally, every load
    either fails with a B200rtError or yields a description whose counts are sane."""
    import subprocess, sys, textwrap
    src = os.path.join(S, "hw7/practice7_4.gltf")
    text = open(src, "rb").read()
    blob = open(os.path.join(S, "hw7/practice7_4.bin"), "rb").read()
    rng = np.random.default_rng(11)
    paths = []
    for i in range(300):
        t = bytearray(text)
        kind = i % 6
        if kind == 0:        # flip random bytes
            for _ in range(int(rng.integers(1, 6))):
                t[int(rng.integers(0, len(t)))] = int(rng.integers(32, 127))
        elif kind == 1:      # delete a slice
            a = int(rng.integers(0, len(t) - 1)); del t[a:a + int(rng.integers(1, 200))]
        elif kind == 2:      # truncate
            del t[int(rng.integers(1, len(t))):]
        elif kind == 3:      # blow up some number
            s = t.decode("latin1")
            nums = [m for m in re.finditer(r"(?<=[:\[, ])\d+(?=[,\]\} \n])", s)]
            m = nums[int(rng.integers(0, len(nums)))]
            s = s[:m.start()] + str(int(rng.choice([0, 7, 99999999, 4294967295, 2 ** 40]))) + s[m.end():]
            t = bytearray(s.encode("latin1"))
        elif kind == 4:      # duplicate a slice somewhere else
            a = int(rng.integers(0, len(t) - 50)); b = int(rng.integers(0, len(t)))
            t[b:b] = t[a:a + int(rng.integers(1, 50))]
        d = tmp_path / f"m{i}"
        d.mkdir()
        (d / "practice7_4.gltf").write_bytes(bytes(t))
        (d / "practice7_4.bin").write_bytes(blob[:int(rng.integers(0, len(blob)))] if kind == 5 else blob)
        paths.append(str(d / "practice7_4.gltf"))
    (tmp_path / "list.txt").write_text("\n".join(paths))
    code = textwrap.dedent(f"""
        import sys
        sys.path.insert(0, {ROOT!r})
        import raytracing_course_hw_b200 as rt
        ok = bad = 0
        for p in open({str(tmp_path / 'list.txt')!r}).read().split("\\n"):
            for mode in (0, 8):
                try:
                    sc = rt.load_scene(p, width=8, height=8, samples=1, mode=mode)
                    d = sc.desc
                    assert 0 <= d.n_tris < 1 << 27 and 0 <= d.n_light_tris <= d.n_tris and d.n_materials >= 0
                    ok += 1
                except rt.B200rtError:
                    bad += 1
        print("loaded", ok, "rejected", bad)
    """)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-2000:])
    ok, bad = (int(x) for x in re.findall(r"loaded (\d+) rejected (\d+)", r.stdout)[0])
    assert ok + bad == 600 and bad > 100 and ok > 50, r.stdout       # both outcomes occur: the mutations are neither all fatal nor all harmless


def test_png_decoder_survives_mutated_files(tmp_path):
    """Same for the built-in PNG decoder (dynamic-Huffman deflate stream of the shipped texture): 120 seeded mutations /
    truncations, decoded in a subprocess through b200rt_host_set_env_map; never a crash, never a partial image."""
    import subprocess, sys, textwrap
    data = open(os.path.join(S, "hw8/sphere_emission.png"), "rb").read()
    rng = np.random.default_rng(5)
    names = []
    for i in range(120):
        t = bytearray(data)
        if i % 3 == 0:
            del t[int(rng.integers(8, len(t))):]
        elif i % 3 == 1:
            for _ in range(int(rng.integers(1, 8))):
                t[int(rng.integers(8, len(t)))] = int(rng.integers(0, 256))
        else:
            a = int(rng.integers(33, len(t) - 64)); del t[a:a + int(rng.integers(1, 64))]
        (tmp_path / f"m{i}.png").write_bytes(bytes(t))
        names.append(str(tmp_path / f"m{i}.png"))
    (tmp_path / "list.txt").write_text("\n".join(names))
    code = textwrap.dedent(f"""
        import sys
        sys.path.insert(0, {ROOT!r})
        import raytracing_course_hw_b200 as rt
        sc = rt.load_scene({os.path.join(S, 'hw8/sphere_emissive.gltf')!r}, width=8, height=8, samples=1)
        ok = bad = 0
        for p in open({str(tmp_path / 'list.txt')!r}).read().split("\\n"):
            try:
                sc.set_env_map(p)
                e = sc.desc.env_map.contents
                assert e.width == 2048 and e.height == 1024 or e.width > 0
                ok += 1
            except rt.B200rtError:
                bad += 1
        print("decoded", ok, "rejected", bad)
    """)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-2000:])
    ok, bad = (int(x) for x in re.findall(r"decoded (\d+) rejected (\d+)", r.stdout)[0])
    assert ok + bad == 120 and bad >= 40, r.stdout
