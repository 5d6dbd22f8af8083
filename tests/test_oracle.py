"""CPU tests: the plain-C oracle (oracle/rt_oracle.c) against the reference-generated golden fixtures and, when it
has been built, against the reference's own code (oracle/_ref)."""
import numpy as np
import pytest

import raytracing_gpu_b200 as rt
from tests import helpers

ALL = helpers.NAMED_SCENES + ["earth", "final", "mesh", "media_mix"]   # media_mix: overlapping media, medium listed before a nearer surface


def scene_from_golden(name):
    g = helpers.load_golden(name)
    sc = rt.Scene()
    sc.replay(g["program"])
    return g, sc


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32-10
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        c = np.asarray(ctr, np.uint32)
        k = np.asarray(key, np.uint32)
        out = np.zeros(4, np.uint32)
        oracle.rto_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
        assert tuple(int(x) for x in out) == want


def test_philox_matches_torch_free_reference(oracle):
    # independent pure-Python restatement of the round function
    def py_philox(ctr, key):
        c, k = list(ctr), list(key)
        for _ in range(10):
            p0, p1 = 0xD2511F53 * c[0], 0xCD9E8D57 * c[2]
            c = [((p1 >> 32) ^ c[1] ^ k[0]) & 0xffffffff, p1 & 0xffffffff, ((p0 >> 32) ^ c[3] ^ k[1]) & 0xffffffff, p0 & 0xffffffff]
            k = [(k[0] + 0x9E3779B9) & 0xffffffff, (k[1] + 0xBB67AE85) & 0xffffffff]
        return c
    rng = np.random.default_rng(5)
    for _ in range(50):
        c = rng.integers(0, 2**32, 4, dtype=np.uint64).astype(np.uint32)
        k = rng.integers(0, 2**32, 2, dtype=np.uint64).astype(np.uint32)
        out = np.zeros(4, np.uint32)
        oracle.rto_philox4x32_10(c.ctypes.data, k.ctypes.data, out.ctypes.data)
        assert [int(x) for x in out] == py_philox([int(x) for x in c], [int(x) for x in k])


def test_uniform_mapping_is_curand_range(oracle):
    assert oracle.rto_u01(0) > 0.0                      # (0, 1], as curand_uniform
    assert oracle.rto_u01(0xffffffff) == 1.0
    assert abs(oracle.rto_u01(0x80000000) - 0.5) < 1e-6


@pytest.mark.parametrize("name", ALL)
def test_first_hit_matches_reference_golden(name):
    g, sc = scene_from_golden(name)
    assert sc.flatten() == int(g["num_leaves"])
    o = helpers.OracleScene(sc.export())
    ids, t = o.first_hit(g["rays"], xi=g["xi"])
    np.testing.assert_array_equal(ids, g["ref_id"])
    np.testing.assert_array_equal(t.view(np.uint32), g["ref_t"].view(np.uint32))   # bit-exact


def test_media_are_exercised_by_the_fixture():
    g = helpers.load_golden("cornell_smoke")
    assert np.isin(g["ref_id"], [6, 13]).sum() > 20          # the two constant_medium leaf ids
    g = helpers.load_golden("final")
    # sphere-bounded media are inert under reference semantics
    sc = rt.Scene(); sc.replay(g["program"]); blob = sc.export()
    hdr = helpers.blob_header(blob)
    assert hdr[7] == 2                                       # n_media
    o = helpers.OracleScene(blob)
    ids_with, _ = o.first_hit(g["rays"], xi=g["xi"])
    ids_without, _ = o.first_hit(g["rays"], xi=None)
    np.testing.assert_array_equal(ids_with, ids_without)


@pytest.mark.parametrize("name", ["first", "two_perlin", "cornell", "cornell_smoke", "earth", "triangles", "mesh"])
def test_converged_image_matches_reference_render(name):
    """Statistical image parity: the reference's own render()/color_f() output (XORWOW, racy camera stream) against
    the restatement (Philox).  Bound: RMSE no worse than 1.3x the restatement's own seed-to-seed RMSE + 2e-3, and
    image means within 1 %."""
    g, sc = scene_from_golden(name)
    o = helpers.OracleScene(sc.export())
    ref = g["image"]
    h, w, _ = ref.shape
    spp = int(g["image_spp"])
    a, _ = o.render(w, h, spp, seed=1984)
    b, _ = o.render(w, h, spp, seed=4891)
    noise = helpers.rmse(a, b)
    assert helpers.rmse(a, ref) <= 1.3 * noise + 2e-3
    assert abs(a.mean() - ref.mean()) <= 0.01 * ref.mean() + 1e-4
    assert not np.isnan(a).any()


@pytest.mark.slow
def test_big1_converged_image_matches_reference_render():
    g, sc = scene_from_golden("big1")
    o = helpers.OracleScene(sc.export())
    ref = g["image"]
    h, w, _ = ref.shape
    a, _ = o.render(w, h, 256, seed=1984)
    b, _ = o.render(w, h, 256, seed=4891)
    noise = helpers.rmse(a, b)                 # at 256 spp; the reference image has 1024 spp (half the noise)
    assert helpers.rmse(a, ref) <= 1.1 * noise
    assert abs(a.mean() - ref.mean()) <= 0.01 * ref.mean()


def test_sample_partition_independence():
    """Philox keying: rendering samples [0,8) at once equals rendering [0,3) + [3,8) (up to float addition order)."""
    g, sc = scene_from_golden("cornell_smoke")
    o = helpers.OracleScene(sc.export())
    full, _ = o.render(24, 24, 8)
    a, _ = o.render(24, 24, 3, sample_first=0)
    b, _ = o.render(24, 24, 5, sample_first=3)
    np.testing.assert_allclose(full, (3 * a + 5 * b) / 8, rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["big1", "cornell_smoke", "triangles"])
def test_first_hit_live_against_reference(ref, name):
    """Fresh random rays through the reference's own hit code (incl. its bvh_node for big1)."""
    h = ref.ref_scene_named(name.encode())
    prog = helpers.ref_program(ref, h)
    sc = rt.Scene(); sc.replay(prog); blob = sc.export()
    o = helpers.OracleScene(blob)
    rays, xi = helpers.make_rays(blob, 20000, seed=99)
    ids = np.empty(len(rays), np.int32); t = np.empty(len(rays), np.float32)
    ref.ref_first_hit(h, rays.ctypes.data, len(rays), 0.001, xi.ctypes.data, ids.ctypes.data, t.ctypes.data, 4)
    oi, ot = o.first_hit(rays, xi=xi)
    np.testing.assert_array_equal(oi, ids)
    np.testing.assert_array_equal(ot.view(np.uint32), t.view(np.uint32))
