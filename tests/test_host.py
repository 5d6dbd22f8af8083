"""CPU tests of the host side of the C-ABI: exported symbols, scene recording, flattening, error behaviour."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import raytracing_gpu_b200 as rt
from tests import helpers

ROOT = helpers.ROOT


def declared_functions():
    names = set()
    for hdr in ("rt_b200.h", "rt_catalog.h", "rt_obj.h", "rt_output.h"):      # every C-ABI header of include/
        path = os.path.join(ROOT, "include", hdr)
        if not os.path.exists(path):
            continue
        src = open(path).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names |= set(re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", src))
    return sorted(names)


def test_library_exports_every_declared_symbol():
    names = declared_functions()
    assert len(names) > 40
    out = subprocess.check_output(["nm", "-D", "--defined-only", rt.lib_path()], text=True)
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    missing = [n for n in names if n not in exported]
    assert not missing, missing
    L = C.CDLL(rt.lib_path())
    for n in names:
        getattr(L, n)


def test_image_height_reproduces_reference_rounding():
    # render.h:33 with aspect = 16.0f/9.0f: one row short
    a = np.float32(16.0) / np.float32(9.0)
    assert rt.image_height(400, a) == 224
    assert rt.image_height(1200, a) == 674
    assert rt.image_height(1920, a) == 1079
    assert rt.image_height(3840, a) == 2159
    assert rt.image_height(800, 1.0) == 800


def test_quantize_frame_matches_color_h():
    rng = np.random.default_rng(0)
    img = (rng.random((5, 7, 3), dtype=np.float32) * 1.3).astype(np.float32)
    q = rt.quantize_frame(img)
    want = (256 * np.clip(np.sqrt(img[::-1]), 0, np.float32(0.999))).astype(np.int32).astype(np.uint8)   # color.h:23-44
    np.testing.assert_array_equal(q, want)


def test_camera_matches_reference_constructor_bits():
    # big_scene1's camera (scenes.h:196) built through rt_set_camera must equal the raw camera the reference built
    g = helpers.load_golden("big1")
    sc = rt.Scene(); sc.replay(g["program"]); raw = helpers.blob_camera(sc.export())
    s2 = rt.Scene()
    m = s2.mat_lambertian(s2.tex_solid(0.5, 0.5, 0.5))
    s2.set_world(s2.list([s2.sphere(0, 0, 0, 1, m)], 1))
    s2.set_camera([13.0, 2.0, -3.0], [0, 0, 0], [0, 1, 0], 20.0, float(np.float32(16.0) / np.float32(9.0)), 0.1, 10.0, 0.0, 1.0)
    mine = helpers.blob_camera(s2.export())
    np.testing.assert_array_equal(mine.view(np.uint32), raw.view(np.uint32))


def test_rotate_and_medium_constructors_match_reference_bits():
    # cornell_smoke: rotate_y(15), rotate_y(-18), constant_medium(0.01): degree/density forms == exported exact forms
    g = helpers.load_golden("cornell_smoke")
    sc = rt.Scene(); sc.replay(g["program"]); ref_blob = sc.export()
    s = rt.Scene()
    tex = lambda r, gg, b: s.tex_solid(r, gg, b)
    red, white, green = (s.mat_lambertian(tex(.65, .05, .05)), s.mat_lambertian(tex(.73, .73, .73)), s.mat_lambertian(tex(.12, .45, .15)))
    light = s.mat_diffuse_light(tex(15, 15, 15))
    objs = [s.yz_rect(0, 555, 0, 555, 555, green), s.yz_rect(0, 555, 0, 555, 0, red), s.xz_rect(113, 443, 127, 432, 554, light),
            s.xz_rect(0, 555, 0, 555, 0, white), s.xz_rect(0, 555, 0, 555, 555, white), s.xy_rect(0, 555, 0, 555, 555, white)]
    b1 = s.translate(s.rotate_y(s.box([0, 0, 0], [165, 330, 165], white), 15.0), 265, 0, 295)
    objs.append(s.constant_medium(b1, 0.01, tex(0, 0, 0)))
    b2 = s.translate(s.rotate_y(s.box([0, 0, 0], [165, 165, 165], white), -18.0), 130, 0, 65)
    objs.append(s.constant_medium(b2, 0.01, tex(1, 1, 1)))
    s.set_world(s.list(objs, len(objs)))
    s.set_background(0, 0, 0)
    s.set_camera([278, 278, -800], [278, 278, 0], [0, 1, 0], 40.0, 1.0, 0.0, 10.0, 0.0, 1.0)
    mine = s.export()
    hm, hr = helpers.blob_header(mine), helpers.blob_header(ref_blob)
    np.testing.assert_array_equal(hm[1:8], hr[1:8])            # leaf ids, prims, bprims, tri, xops, chains, media
    # geometry + transforms + media sections are bit-identical (materials/textures are indexed differently)
    o_m, o_r = helpers.OracleScene(mine), helpers.OracleScene(ref_blob)
    im, tm = o_m.first_hit(g["rays"], xi=g["xi"])
    np.testing.assert_array_equal(im, g["ref_id"])
    np.testing.assert_array_equal(tm.view(np.uint32), g["ref_t"].view(np.uint32))
    del o_r


def test_flatten_box_and_medium_leaf_ids():
    s = rt.Scene()
    m = s.mat_lambertian(s.tex_solid(1, 1, 1))
    box = s.box([0, 0, 0], [1, 1, 1], m)
    med = s.constant_medium(s.sphere(0, 0, 0, 1, m), 0.5, s.tex_solid(1, 1, 1))
    s.set_world(s.list([s.sphere(0, 0, 0, 1, m), box, med, s.sphere(0, 0, 0, 2, m)], 4))
    s.set_camera([0, 0, 5], [0, 0, 0], [0, 1, 0], 40.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    assert s.flatten() == 1 + 6 + 2 + 1
    hdr = helpers.blob_header(s.export())
    assert (hdr[2], hdr[3], hdr[7]) == (8, 1, 1)               # prims, boundary prims, media


def test_error_behaviour_is_codes_not_exit():
    s = rt.Scene()
    with pytest.raises(rt.RtError, match="bad material index"):
        s.sphere(0, 0, 0, 1, 5)
    with pytest.raises(rt.RtError, match="bad texture index"):
        s.mat_lambertian(3)
    with pytest.raises(rt.RtError, match="no world"):
        s.flatten()
    m = s.mat_lambertian(s.tex_solid(1, 1, 1))
    s.set_world(s.sphere(0, 0, 0, 1, m))
    with pytest.raises(rt.RtError, match="no camera"):
        s.flatten()
    s.set_camera([0, 0, 5], [0, 0, 0], [0, 1, 0], 40.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    assert s.flatten() == 1
    with pytest.raises(rt.RtError, match="not committed"):
        s.first_hit(np.zeros((1, 7), np.float32))
    with pytest.raises(rt.RtError, match="not committed"):
        s.render(4, 4, 1)
    with pytest.raises(rt.RtError, match="bad magic"):
        s.replay(np.zeros(8, np.uint32))
    g = helpers.load_golden("basic")
    with pytest.raises(rt.RtError, match="truncated"):
        rt.Scene().replay(g["program"][:20])
    with pytest.raises(rt.RtError):                      # no device in this container / bad split refused before any work
        s.render_multi(8, 8, 1, split="tile", devices=[0])
    # parameters that drive device loops are validated where the scene is recorded
    t = rt.Scene()
    pn = t.perlin_tables(np.ones(768, np.float32), np.zeros(768, np.int32))
    with pytest.raises(rt.RtError, match="turbulence depth"):
        t.tex_turbulent(pn, 1.0, 0x7fffffff)
    with pytest.raises(rt.RtError, match="turbulence depth"):
        t.tex_turbulent(pn, 1.0, -1)
    assert t.tex_turbulent(pn, 1.0, 7) >= 0
    c = t.tex_solid(0, 0, 0)
    for _ in range(8):
        c = t.tex_checker(c, c)
    with pytest.raises(rt.RtError, match="nested more than 8"):
        t.tex_checker(c, c)
    # an idempotent flatten keeps the flattening (and with it a committed device copy) current; an edit invalidates it
    h1 = s.scene_hash()
    assert s.flatten() == 1 and s.scene_hash() == h1
    s.set_background(0.1, 0.2, 0.3)
    assert s.flatten() == 1 and s.scene_hash() != h1


def test_nested_medium_is_rejected():
    s = rt.Scene()
    m = s.mat_lambertian(s.tex_solid(1, 1, 1))
    inner = s.constant_medium(s.sphere(0, 0, 0, 1, m), 0.5, s.tex_solid(1, 1, 1))
    outer = s.constant_medium(inner, 0.5, s.tex_solid(1, 1, 1))
    s.set_world(outer)
    s.set_camera([0, 0, 5], [0, 0, 0], [0, 1, 0], 40.0, 1.0, 0.0, 1.0, 0.0, 1.0)
    with pytest.raises(rt.RtError, match="nested"):
        s.flatten()


def test_mesh_matches_reference_create_meshes_d():
    # rt_mesh must build the same triangles the reference's create_meshes_d built for the golden mesh scene
    g = helpers.load_golden("mesh")
    v, idx, uv, nrm = helpers.procedural_mesh(16, 24)
    img = helpers.procedural_image(64, 64, seed=13)
    s = rt.Scene()
    mat = s.mat_lambertian(s.tex_image(s.image(img, 64, 64, 3)))
    mesh = s.mesh(v, len(v), idx, len(idx) // 3, uv, nrm, mat)
    ground = s.sphere(0, -100, -1, 100, s.mat_lambertian(s.tex_solid(0, 1, 0)))
    s.set_world(s.list([mesh, ground], 2))
    s.set_background(0.7, 0.8, 1.0)
    c = helpers.MESH_CAMERA
    s.set_camera(c[0:3], c[3:6], c[6:9], float(c[9]), float(c[10]), float(c[11]), float(c[12]), float(c[13]), float(c[14]))
    o = helpers.OracleScene(s.export())
    ids, t = o.first_hit(g["rays"], xi=g["xi"])
    np.testing.assert_array_equal(ids, g["ref_id"])
    np.testing.assert_array_equal(t.view(np.uint32), g["ref_t"].view(np.uint32))


def test_commit_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    g = helpers.load_golden("basic")
    s = rt.Scene(); s.replay(g["program"])
    with pytest.raises(rt.RtError, match="CUDA"):
        s.commit(0)


def test_struct_layouts_of_the_binding_match_the_library():
    """rt_struct_size (include/rt_b200.h): every struct the ctypes binding mirrors by hand has the size the library was compiled
    with; an unknown id answers 0.  (The binding performs the same check when it loads the library.)"""
    import ctypes as C
    L = rt.lib()
    assert len(rt.capi.STRUCT_MIRRORS) == 9
    for sid, cls in enumerate(rt.capi.STRUCT_MIRRORS):
        assert L.rt_struct_size(sid) == C.sizeof(cls) > 0, cls.__name__
    assert L.rt_struct_size(9) == 0 and L.rt_struct_size(-1) == 0
