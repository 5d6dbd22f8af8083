"""Test helpers: loaders for the two oracles (TEST INFRASTRUCTURE under oracle/), scene programs, ray sets."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
ORACLE_DIR = os.path.join(ROOT, "oracle")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

NAMED_SCENES = ["basic", "first", "big1", "two_spheres", "two_perlin", "cornell", "cornell_smoke", "triangle", "triangles"]

_oracle = None
_ref = None


def oracle_lib():
    """oracle/librt_oracle.so — the plain-C restatement; built on demand (gcc only)."""
    global _oracle
    if _oracle is not None:
        return _oracle
    path = os.path.join(ORACLE_DIR, "librt_oracle.so")
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "librt_oracle.so"])
    L = C.CDLL(path)
    vp = C.c_void_p
    L.rto_scene_load.restype = vp
    L.rto_scene_load.argtypes = [vp, C.c_size_t]
    L.rto_scene_free.argtypes = [vp]
    L.rto_first_hit.argtypes = [vp, vp, C.c_int, C.c_float, vp, vp, vp]
    L.rto_philox4x32_10.argtypes = [vp, vp, vp]
    L.rto_u01.restype = C.c_float
    L.rto_u01.argtypes = [C.c_uint32]
    L.rto_render_rows.restype = C.c_uint64
    L.rto_render_rows.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_int, C.c_uint32, vp]
    L.rto_sample_radiance.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_uint32, vp, vp, C.c_int, vp, vp]
    _oracle = L
    return L


def ref_lib():
    """oracle/_ref/libref_oracle.so — the reference's own sources compiled for the host; None if absent."""
    global _ref
    if _ref is not None:
        return _ref
    path = os.path.join(ORACLE_DIR, "_ref", "libref_oracle.so")
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    vp = C.c_void_p
    L.ref_scene_named.argtypes = [C.c_char_p]
    L.ref_scene_earth.argtypes = [vp, C.c_int, C.c_int, C.c_int]
    L.ref_scene_final.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_float]
    L.ref_scene_mesh.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]
    L.ref_program_words.restype = C.c_size_t
    L.ref_aspect.restype = C.c_float
    L.ref_first_hit.argtypes = [C.c_int, vp, C.c_int, C.c_float, vp, vp, vp, C.c_int]
    L.ref_render.argtypes = [C.c_int] * 7 + [vp]
    _ref = L
    return L


def ref_output_pipeline(frames, workdir):
    """The reference's own write_frame_buffer + average_images (color.h) on (n, H, W, 3) linear float frames, run by
    oracle/_ref/ref_output_tool in its own process.  Returns (final uint8 image, directory holding the PPMs) or None
    when the tool was never built."""
    import subprocess
    tool = os.path.join(ORACLE_DIR, "_ref", "ref_output_tool")
    if not os.path.exists(tool):
        return None
    frames = np.ascontiguousarray(frames, np.float32)
    n, h, w, _ = frames.shape
    din = os.path.join(workdir, "in")
    os.mkdir(din)
    raw_in, raw_out = os.path.join(workdir, "frames.f32"), os.path.join(workdir, "out.raw")
    frames.tofile(raw_in)
    subprocess.check_call([tool, str(n), str(w), str(h), raw_in, din, raw_out])
    data = open(raw_out, "rb").read()
    header, _, body = data.partition(b"\n")
    assert header.split() == [b"RAW8", str(w).encode(), str(h).encode()]
    return np.frombuffer(body, np.uint8).reshape(h, w, 3).copy(), din


def ref_program(L, handle):
    n = L.ref_program_words(handle)
    prog = np.zeros(n, np.uint32)
    L.ref_program_copy(handle, prog.ctypes.data)
    return prog


class OracleScene:
    """A flattened-scene blob loaded into the C oracle."""

    def __init__(self, blob):
        self.L = oracle_lib()
        self.blob = np.ascontiguousarray(blob, dtype=np.uint8)
        self.h = self.L.rto_scene_load(self.blob.ctypes.data, self.blob.size)
        assert self.h, "oracle rejected the blob"

    def first_hit(self, rays, t_min=0.001, xi=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 7)
        n = rays.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        xi_a = np.ascontiguousarray(xi, np.float32) if xi is not None else None
        self.L.rto_first_hit(self.h, rays.ctypes.data, n, t_min, xi_a.ctypes.data if xi_a is not None else None,
                             ids.ctypes.data, t.ctypes.data)
        return ids, t

    def render(self, width, height, spp, max_depth=50, seed=1984, sample_first=0, workers=None):
        """Mean image (H, W, 3), row 0 = bottom; rows fanned out over processes."""
        from concurrent.futures import ThreadPoolExecutor
        workers = workers or os.cpu_count() or 1
        sums = np.zeros((height, width, 3), np.float32)
        rows = [(r, min(height, r + 4)) for r in range(0, height, 4)]

        def job(rr):   # ctypes releases the GIL during the call
            return self.L.rto_render_rows(self.h, width, height, rr[0], rr[1], sample_first, spp, max_depth, seed, sums.ctypes.data)
        with ThreadPoolExecutor(workers) as ex:
            rays = sum(ex.map(job, rows))
        return (np.float32(1.0) / np.float32(spp)) * sums, rays

    def sample_radiance(self, width, height, pix, smp, max_depth=50, seed=1984):
        pix = np.ascontiguousarray(pix, np.uint32)
        smp = np.ascontiguousarray(smp, np.uint32)
        out = np.empty((pix.size, 3), np.float32)
        segs = np.empty(pix.size, np.uint32)
        self.L.rto_sample_radiance(self.h, width, height, max_depth, seed, pix.ctypes.data, smp.ctypes.data, pix.size,
                                   out.ctypes.data, segs.ctypes.data)
        return out, segs

    def __del__(self):
        try:
            self.L.rto_scene_free(self.h)
        except Exception:
            pass


def blob_camera(blob):
    """rt_flat_camera (24 floats) from an export blob (rt_flat_header: 80-byte prefix)."""
    return np.frombuffer(np.ascontiguousarray(blob)[80:176].tobytes(), np.float32)


def blob_header(blob):
    return np.frombuffer(np.ascontiguousarray(blob)[:64].tobytes(), np.uint32)


def make_rays(blob, n, seed):
    """Deterministic ray set for a scene: half primary rays through random film points (random shutter time),
    half secondary-like rays (origins jittered around the eye->scene segment, random directions)."""
    cam = blob_camera(blob)
    origin, llc, hor, ver = cam[0:3], cam[3:6], cam[6:9], cam[9:12]
    rng = np.random.default_rng(seed)
    su = rng.random(n, dtype=np.float32)
    sv = rng.random(n, dtype=np.float32)
    d = (llc + su[:, None] * hor + sv[:, None] * ver - origin).astype(np.float32)
    o = np.tile(origin, (n, 1)).astype(np.float32)
    tm = rng.random((n, 1), dtype=np.float32)
    rays = np.concatenate([o, d, tm], axis=1).astype(np.float32)
    h = n // 2
    # secondary-like: start somewhere along the primary ray, head in a random direction
    frac = rng.random((n - h, 1), dtype=np.float32)
    rays[h:, 0:3] = rays[h:, 0:3] + frac * rays[h:, 3:6] * np.float32(0.9)
    rays[h:, 3:6] = rng.random((n - h, 3), dtype=np.float32) - np.float32(0.5)
    xi = (rng.random(n, dtype=np.float32) * np.float32(0.998) + np.float32(0.001)).astype(np.float32)
    return np.ascontiguousarray(rays), xi


def load_golden(name):
    path = os.path.join(GOLDEN, name + ".npz")
    return np.load(path)


def rmse(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)))


from workloads import procedural_image, procedural_mesh, MESH_CAMERA  # noqa: E402,F401  (synthetic inputs shared with bench.py)


def channel_rmse(a, b):
    """Per-channel RMSE of two (H, W, 3) images."""
    d = np.asarray(a, np.float64) - np.asarray(b, np.float64)
    return np.sqrt((d * d).mean(axis=(0, 1)))


def channel_psnr(a, b, peak=1.0):
    """Per-channel PSNR in dB of two linear-radiance images clipped to [0, peak] (what an 8-bit output stage would show)."""
    d = np.clip(np.asarray(a, np.float64), 0, peak) - np.clip(np.asarray(b, np.float64), 0, peak)
    mse = (d * d).mean(axis=(0, 1))
    return 10.0 * np.log10(peak * peak / np.maximum(mse, 1e-20))


def media_leaf_ids(blob):
    """Leaf ids of the constant_medium objects in an export blob (layout: include/rt_flat.h)."""
    h = blob_header(blob)
    pad16 = lambda n: (n + 15) & ~15
    off = 176 + pad16(64 * int(h[2])) + pad16(64 * int(h[3])) + pad16(96 * int(h[4])) + pad16(16 * int(h[5])) + pad16(16 * int(h[6]))
    n = int(h[7])
    m = np.frombuffer(np.ascontiguousarray(blob)[off:off + 32 * n].tobytes(), np.uint32).reshape(n, 8)
    return m[:, 4].astype(np.int32)


def assert_first_hit_parity(blob, ids, t, want_ids, want_t):
    """The parity bar for first-hit queries: ids exact; t bit-identical for surface hits; for constant_medium hits
    (whose t goes through libm logf, constant_medium.h:57) within 1e-6 relative — both inside BASELINE's 1e-5."""
    np.testing.assert_array_equal(ids, want_ids)
    med = np.isin(ids, media_leaf_ids(blob))
    np.testing.assert_array_equal(t[~med].view(np.uint32), want_t[~med].view(np.uint32))
    if med.any():
        np.testing.assert_allclose(t[med], want_t[med], rtol=1e-6, atol=0)


def ref_cuda_probe(scene_name, rays, image=None, exe="ref_cuda_nofma", with_ids=False):
    """First-hit distances from the REFERENCE's own device code: oracle/_ref/<exe> builds the named scene with the
    reference's kernels and runs world->hit (hittable_list.h / bvh.h) on `rays` in one CUDA thread per ray.  Returns
    float32 t (inf = miss) — with_ids: (t, int32 DFS leaf ids read back from the materials the harness tagged on the device, info) —
    or None when the binary was not built (it is compiled from /root/reference by
    oracle/Makefile in the authoring container and travels to the GPU box)."""
    import subprocess
    import tempfile
    path = os.path.join(ROOT, "oracle", "_ref", exe)
    if not os.path.exists(path):
        return None
    with tempfile.TemporaryDirectory() as d:
        rp, op = os.path.join(d, "rays.f32"), os.path.join(d, "out.f32")
        np.ascontiguousarray(rays, np.float32).tofile(rp)
        extra = []
        if image is not None:
            ip = os.path.join(d, "image.rgb8")
            np.ascontiguousarray(image, np.uint8).tofile(ip)
            extra = [ip, str(image.shape[1]), str(image.shape[0])]
        res = subprocess.run([path, scene_name, "probe", rp, op] + extra, check=True, capture_output=True, text=True)
        out = np.fromfile(op, np.float32)
    n = len(rays)
    if not with_ids:
        return out[:n]
    import json
    return out[:n], out[4 * n:5 * n].view(np.int32), json.loads(res.stdout.strip().splitlines()[-1])
