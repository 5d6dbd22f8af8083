"""Host side of progressive rendering: checkpoint files and the scene hash.  No GPU needed."""
import ctypes as C
import os

import numpy as np
import pytest

import raytracing_gpu_b200 as rt
from raytracing_gpu_b200.capi import lib, _check
from tests import helpers


def small_scene(radius=1.0):
    sc = rt.Scene()
    m = sc.mat_lambertian(sc.tex_solid(0.5, 0.5, 0.5))
    sc.set_world(sc.sphere(0.0, 0.0, 0.0, radius, m))
    sc.set_camera([0, 0, 5], [0, 0, 0], [0, 1, 0], 40.0, 1.0, 0.0, 5.0, 0.0, 1.0)
    return sc


def test_scene_hash_tracks_content_and_semantics():
    a, b, c = small_scene(), small_scene(), small_scene(1.5)
    assert a.scene_hash() == b.scene_hash() != c.scene_hash()
    h = a.scene_hash()
    a.set_semantics(rt.SEM_SPHERE_FAR_ROOT)
    assert a.scene_hash() != h
    a.set_semantics(rt.SEM_REFERENCE)
    assert a.scene_hash() == h
    with pytest.raises(rt.RtError, match="unknown bits"):
        a.set_semantics(8)
    g = helpers.load_golden("cornell_smoke")          # replayed programs hash like directly built scenes: by content
    s1, s2 = rt.Scene(), rt.Scene()
    s1.replay(g["program"]); s2.replay(g["program"])
    assert s1.scene_hash() == s2.scene_hash()


def test_checkpoint_round_trip_and_damage(tmp_path):
    sc = small_scene()
    pr = rt.ProgressiveRender(sc, 7, 5, spp_total=64, max_depth=12, seed=99)
    rng = np.random.default_rng(4)
    pr.sums[:] = rng.random(pr.sums.shape, dtype=np.float32) * 17
    pr.samples_done = 23
    path = str(tmp_path / "frame.ckpt")
    pr.save(path)
    assert not os.path.exists(path + ".tmp")
    info = rt.ProgressiveRender.peek(path)
    assert (info.width, info.height, info.seed, info.max_depth, info.samples_done) == (7, 5, 99, 12, 23)
    assert info.scene_hash == sc.scene_hash()
    back = rt.ProgressiveRender.resume(sc, path, spp_total=64)
    assert back.samples_done == 23 and back.seed == 99 and back.max_depth == 12
    assert (back.sums == pr.sums).all()
    np.testing.assert_array_equal(back.image(), (np.float32(1.0) / np.float32(23)) * pr.sums)

    with pytest.raises(rt.RtError, match="another scene"):
        rt.ProgressiveRender.resume(small_scene(2.0), path, spp_total=64)
    with pytest.raises(rt.RtError, match="more than spp_total"):
        rt.ProgressiveRender.resume(sc, path, spp_total=16)

    raw = bytearray(open(path, "rb").read())
    assert len(raw) == 64 + 7 * 5 * 3 * 4
    raw[100] ^= 0x40                                   # one flipped payload bit
    open(path, "wb").write(raw)
    with pytest.raises(rt.RtError, match="corrupt"):
        rt.ProgressiveRender.resume(sc, path, spp_total=64)
    open(path, "wb").write(raw[:90])
    with pytest.raises(rt.RtError, match="truncated"):
        rt.ProgressiveRender.resume(sc, path, spp_total=64)
    open(path, "wb").write(b"P3\n1 1\n255\n0 0 0\n" + bytes(80))
    with pytest.raises(rt.RtError, match="not a checkpoint"):
        rt.ProgressiveRender.peek(path)
    with pytest.raises(rt.RtError, match="cannot open"):
        rt.ProgressiveRender.peek(str(tmp_path / "missing.ckpt"))
    info = rt.CheckpointInfo(width=7, height=5)
    small = np.zeros(10, np.float32)
    pr.save(path)
    with pytest.raises(rt.RtError, match="too small"):
        _check(lib().rt_checkpoint_load(os.fsencode(path), C.byref(info), small.ctypes.data, small.size))


def test_progressive_without_samples_has_no_image():
    pr = rt.ProgressiveRender(small_scene(), 4, 4, spp_total=8)
    assert not pr.done
    with pytest.raises(rt.RtError, match="no samples"):
        pr.image()
