"""CPU tests of the OBJ / MTL loader (include/rt_obj.h), the replacement for the reference's assimp path
(triangle_mesh.h:51-143): checked against an independent Python parse of the same text."""
import os

import numpy as np
import pytest

import raytracing_gpu_b200 as rt

CUBE = """# unit cube: quads, one pentagon cap, negative indices, shared positions with different uvs
mtllib cube.mtl
o Cube
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0 0 1
v 1 0 1
v 1 1 1
v 0 1 1
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vn 0 0 -1
vn 0 0 1
vn 0 -1 0
vn 1 0 0
vn 0 1 0
vn -1 0 0
usemtl Skin
s off
f 1/1/1 4/4/1 3/3/1 2/2/1
f 5/1/2 6/2/2 7/3/2 8/4/2
f 1/1/3 2/2/3 6/3/3 5/4/3
f 2/1/4 3/2/4 7/3/4 6/4/4
f -6/1/5 -5/2/5 -1/3/5 -2/4/5
f 4//6 1//6 5//6 8//6
g cap
v 0.5 1.5 0.5
f 4 3 7 8 9
"""
MTL = """newmtl Other
Kd 1 1 1
newmtl Skin
Kd 0.8 0.8 0.8
map_Kd -s 1 1 1 textures/skin.jpg
map_Kd second.jpg
"""


def python_parse(text):
    """Independent restatement: unified (v, vt, vn) corners, fan triangulation, first use first."""
    pos, tex, nrm, tris = [], [], [], []
    for line in text.splitlines():
        tok = line.split()
        if not tok or tok[0].startswith("#"):
            continue
        if tok[0] == "v":
            pos.append([float(x) for x in tok[1:4]])
        elif tok[0] == "vt":
            tex.append([float(x) for x in (tok[1:3] + ["0"])[:2]])
        elif tok[0] == "vn":
            nrm.append([float(x) for x in tok[1:4]])
        elif tok[0] == "f":
            face = []
            for c in tok[1:]:
                f = (c.split("/") + ["", ""])[:3]
                ix = []
                for val, count in zip(f, (len(pos), len(tex), len(nrm))):
                    i = int(val) if val else 0
                    ix.append(i - 1 if i > 0 else (count + i if i < 0 else -1))
                face.append(tuple(ix))
            for i in range(1, len(face) - 1):
                tris.append((face[0], face[i], face[i + 1]))
    seen, verts, uvs, normals, indices = {}, [], [], [], []
    for tri in tris:
        for c in tri:
            if c not in seen:
                seen[c] = len(verts)
                verts.append(pos[c[0]])
                uvs.append(tex[c[1]] if c[1] >= 0 else [0.0, 0.0])
                normals.append(nrm[c[2]] if c[2] >= 0 else [0.0, 1.0, 0.0])
            indices.append(seen[c])
    return (np.array(verts, np.float32), np.array(uvs, np.float32), np.array(normals, np.float32),
            np.array(indices, np.uint32).reshape(-1, 3))


def test_parse_matches_independent_python_parse():
    m = rt.ObjMesh(text=CUBE, mtl_text=MTL)
    v, uv, n, idx = python_parse(CUBE)
    assert m.n_triangles == 6 * 2 + 3 == len(idx)
    assert not m.normals_generated
    np.testing.assert_array_equal(m.indices, idx)
    np.testing.assert_array_equal(m.vertices, v)
    np.testing.assert_array_equal(m.uvs, uv)
    np.testing.assert_array_equal(m.normals, n)
    assert m.diffuse_map == "textures/skin.jpg"          # first map_Kd, options skipped


def test_load_from_file_reads_mtllib(tmp_path):
    (tmp_path / "cube.obj").write_text(CUBE)
    (tmp_path / "cube.mtl").write_text(MTL)
    m = rt.ObjMesh(path=str(tmp_path / "cube.obj"))
    assert m.n_triangles == 15 and m.diffuse_map == "textures/skin.jpg"
    (tmp_path / "cube.mtl").unlink()                     # a missing .mtl is not an error
    assert rt.ObjMesh(path=str(tmp_path / "cube.obj")).diffuse_map == ""


def test_generated_normals_are_smooth_and_unit():
    # an open pyramid without vn records: the apex normal is the normalised sum of the face normals
    text = "v 0 1 0\nv -1 0 -1\nv 1 0 -1\nv 1 0 1\nv -1 0 1\nf 1 3 2\nf 1 4 3\nf 1 5 4\nf 1 2 5\n"
    m = rt.ObjMesh(text=text)
    assert m.normals_generated and m.n_vertices == 5 and m.n_triangles == 4
    np.testing.assert_allclose(np.linalg.norm(m.normals, axis=1), 1.0, rtol=1e-6)
    np.testing.assert_allclose(m.normals[0], [0, 1, 0], atol=1e-6)
    np.testing.assert_array_equal(m.uvs, np.zeros((5, 2), np.float32))


@pytest.mark.parametrize("text, needle", [("v 0 0 0\nv 1 0 0\nv 0 1 0\n", "no faces"), ("v 0 0 0\nf 1 2 3\n", "out of range"),
                                          ("v 0 0\nf 1 1 1\n", "short v"), ("v 0 0 0\nv 1 0 0\nf 1 2\n", "fewer than 3")])
def test_malformed_input_is_an_error_not_a_crash(text, needle):
    with pytest.raises(rt.RtError, match=needle):
        rt.ObjMesh(text=text)


def test_missing_file_is_an_error():
    with pytest.raises(rt.RtError, match="cannot read"):
        rt.ObjMesh(path="/nonexistent/none.obj")


def test_mesh_goes_into_a_scene_as_one_triangle_mesh():
    m = rt.ObjMesh(text=CUBE, mtl_text=MTL)
    sc = rt.Scene()
    mat = sc.mat_lambertian(sc.tex_solid(0.5, 0.5, 0.5))
    h = m.add_to_scene(sc, mat)
    sc.set_world(sc.list([h]))
    sc.set_camera([0, 0, -3], [0, 0, 0], [0, 1, 0], 40.0, 1.0, 0.0, 10.0, 0.0, 1.0)
    assert sc.flatten() == m.n_triangles                 # one leaf id per triangle


@pytest.mark.skipif(not os.path.exists("/root/reference/assets/door/door.obj"), reason="reference assets not mounted")
def test_reference_door_asset_counts():
    # 3294 positions, 2165 faces = 2155 quads + 8 triangles + 2 12-gons -> 4338 triangles
    m = rt.ObjMesh(path="/root/reference/assets/door/door.obj")
    assert m.n_triangles == 2155 * 2 + 8 + 2 * 10
    assert not m.normals_generated
    assert m.diffuse_map.endswith(".jpg") or m.diffuse_map == ""
    v, uv, n, idx = python_parse(open("/root/reference/assets/door/door.obj").read())
    np.testing.assert_array_equal(m.indices, idx)
    np.testing.assert_array_equal(m.vertices, v)
    np.testing.assert_array_equal(m.uvs, uv)
    np.testing.assert_array_equal(m.normals, n)
