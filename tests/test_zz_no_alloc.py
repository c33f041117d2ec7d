"""b200rt_resolve_noalloc (include/b200rt.h): the resolve through the library's static __device__ scratch, for callers
that must not allocate.  It has no counterpart in the reference, so the
check is against the existing entry point b200rt_resolve on the same radiance buffer: same arithmetic => the SAME
BYTES (tolerance 0, tighter than the 1-LSB bar the other resolve tests use).  Collected last on purpose.
Passes alone in a fresh process and after the other tests (it creates its own contexts)."""
import ctypes
import os

import numpy as np
import pytest

import oracle_util as ou
import raytracing_course_hw_b200 as rt

pytestmark = pytest.mark.gpu
S = ou.SCENES


def _frame(r, path, **kw):
    sc = rt.load_scene(os.path.join(S, path), **kw)
    r.upload(sc)
    r.render(0, sc.samples)
    r.synchronize()
    return sc


def test_no_alloc_resolve_equals_resolve_on_two_calls_in_a_row():
    r = rt.Renderer()
    try:
        # call 1: Monte-Carlo frame (hw8 rules, tonemapped mean of 8 samples)
        sc = _frame(r, "hw8/sphere_emissive.gltf", width=128, height=96, samples=8)
        ptr, n = r.accum_device_ptr()
        assert n == 128 * 96 * 3
        a = rt.resolve_noalloc(0, ptr, 128, 96, 8, sc.mode, stream=r.stream())
        assert np.array_equal(a, r.resolve(8)) and a.max() > 0
        # call 2, right after, on a SMALLER frame of another mode (hw1: raw colour, no tonemap): the scratch still
        # holds call 1's bytes beyond this frame -- none of them may leak into the result
        sc = _frame(r, "hw1/sample.txt", width=80, height=60)
        ptr, n = r.accum_device_ptr()
        assert n == 80 * 60 * 3
        b = rt.resolve_noalloc(0, ptr, 80, 60, 1, sc.mode)          # default stream
        assert np.array_equal(b, r.resolve(1)) and len(np.unique(b)) > 2
        # and once more on the first kind of frame: the result does not depend on what the scratch held before
        sc = _frame(r, "hw8/sphere_emissive.gltf", width=128, height=96, samples=8)
        ptr, _ = r.accum_device_ptr()
        assert np.array_equal(rt.resolve_noalloc(0, ptr, 128, 96, 8, sc.mode, stream=r.stream()), a)
    finally:
        r.close()


def test_no_alloc_resolve_rejects_what_it_cannot_hold():
    r = rt.Renderer()
    try:
        _frame(r, "hw1/sample.txt", width=64, height=48)
        ptr, _ = r.accum_device_ptr()
        w = 2048
        h = rt.NOALLOC_MAX_PIXELS // w + 1            # one row above the static scratch
        out = np.zeros((4, 4, 3), np.uint8)           # never written: the size check comes first
        with pytest.raises(rt.B200rtError) as e:
            rt.resolve_noalloc(0, ptr, w, h, 1, rt.MODE_HW1, out=out)
        assert e.value.code == rt.ERR_TOO_LARGE and "B200RT_NOALLOC_MAX_PIXELS" in str(e.value)
        for bad in (dict(width=0), dict(total_spp=0), dict(mode=0), dict(device=-1), dict(device=4096), dict(accum_device_ptr=0)):
            kw = dict(device=0, accum_device_ptr=ptr, width=64, height=48, total_spp=1, mode=rt.MODE_HW1)
            kw.update(bad)
            with pytest.raises(rt.B200rtError):
                rt.resolve_noalloc(**kw)
    finally:
        r.close()
