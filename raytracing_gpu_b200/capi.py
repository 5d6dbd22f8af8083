"""ctypes binding of include/rt_b200.h.  Mirrors the C-ABI one to one (same names minus the ``rt_`` prefix)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def lib_path():
    """The product library.  RT_B200_LIB selects another build of the SAME sources (A/B of compile-time variants)."""
    return os.environ.get("RT_B200_LIB") or os.path.join(_HERE, "librt_b200.so")


class RtError(RuntimeError):
    pass


class RenderSettings(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("samples_per_pixel", C.c_uint32),
                ("max_depth", C.c_uint32), ("seed", C.c_uint32), ("sample_first", C.c_uint32),
                ("sample_count", C.c_uint32), ("flags", C.c_uint32), ("pool_paths", C.c_uint32),
                ("tile_index", C.c_uint32), ("tile_count", C.c_uint32), ("extend_defer", C.c_uint32), ("traversal", C.c_uint32),
                ("debug_flags", C.c_uint32), ("reserved", C.c_uint32 * 2)]


class RenderStats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("rays", C.c_uint64), ("iterations", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("ms_total", C.c_float), ("ms_generate", C.c_float),
                ("ms_extend", C.c_float), ("ms_shade", C.c_float), ("ms_finish", C.c_float),
                ("nodes_visited", C.c_uint64), ("prim_tests", C.c_uint64), ("node_slots", C.c_uint64),
                ("pool_paths", C.c_uint64), ("reserved", C.c_uint64 * 2)]


class BuildStats(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("n_nodes", C.c_uint32), ("ms_upload", C.c_float), ("ms_morton", C.c_float),
                ("ms_sort", C.c_float), ("ms_hierarchy", C.c_float), ("ms_refit", C.c_float),
                ("ms_collapse", C.c_float), ("ms_total", C.c_float), ("bvh_depth", C.c_uint32), ("build_rounds", C.c_uint32),
                ("sah_internal", C.c_float), ("ploc_radius", C.c_uint32), ("builder", C.c_uint32),
                ("n_wide_nodes", C.c_uint32), ("ms_host", C.c_float), ("ms_wide", C.c_float), ("wide_depth", C.c_uint32)]


class BuildSettings(C.Structure):
    _fields_ = [("builder", C.c_uint32), ("sah_max_prims", C.c_uint32), ("ploc_radius", C.c_uint32), ("ploc_top", C.c_uint32),
                ("wide_nodes", C.c_uint32), ("flags", C.c_uint32), ("reserved", C.c_uint32 * 10)]


BUILD_AUTO, BUILD_KARRAS, BUILD_PLOC, BUILD_HOST_SAH, BUILD_GPU_SAH = 0, 1, 2, 3, 4
BUILDERS = {"auto": 0, "karras": 1, "ploc": 2, "host_sah": 3, "sah": 4, "gpu_sah": 4}
BUILDER_KARRAS, BUILDER_PLOC, BUILDER_HOST_SAH, BUILDER_GPU_SAH = 0, 1, 2, 3      # rt_build_stats.builder
WIDE_AUTO, WIDE_OFF, WIDE_ON = 0, 1, 2
BUILD_TRACE = 1
DEFER_AUTO, DEFER_OFF = 0, 1
TRAVERSAL_FORCE = 0x100
DEBUG_NO_GRAPH, DEBUG_RASTER_FILM, DEBUG_WARP_SHADE, DEBUG_BLOCK_SHADE = 1, 2, 4, 8


class MultiSettings(C.Structure):
    _fields_ = [("n_devices", C.c_uint32), ("device_ids", C.c_int32 * 16), ("split_mode", C.c_uint32), ("reserved", C.c_uint32 * 6)]


class MultiStats(C.Structure):
    _fields_ = [("n_devices", C.c_uint32), ("peer_reduce", C.c_uint32), ("ms_commit_max", C.c_float), ("ms_render", C.c_float * 16),
                ("ms_reduce", C.c_float), ("ms_wall", C.c_float), ("samples", C.c_uint64), ("rays", C.c_uint64), ("reserved", C.c_uint64 * 4)]


SPLIT_SAMPLE, SPLIT_TILE = 0, 1


class CatalogParams(C.Structure):
    _fields_ = [("image", C.c_void_p), ("image_width", C.c_int), ("image_height", C.c_int), ("image_bytes_per_pixel", C.c_int),
                ("aspect", C.c_float), ("vertices", C.c_void_p), ("n_vertices", C.c_int), ("indices", C.c_void_p),
                ("n_triangles", C.c_int), ("uvs", C.c_void_p), ("normals", C.c_void_p), ("camera15", C.c_void_p), ("host_argument_order", C.c_int)]


class DrawSettings(C.Structure):
    _fields_ = [("image_width", C.c_uint32), ("image_height", C.c_uint32), ("aspect_ratio", C.c_float),
                ("samples_per_pixel_per_fb", C.c_uint32), ("no_fb", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint32), ("device", C.c_uint32), ("n_devices", C.c_uint32), ("reserved", C.c_uint32 * 7)]


class CheckpointInfo(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("seed", C.c_uint32), ("max_depth", C.c_uint32),
                ("samples_done", C.c_uint32), ("reserved0", C.c_uint32), ("scene_hash", C.c_uint64),
                ("reserved", C.c_uint64 * 3)]


RENDER_COUNTERS = 1
RENDER_TIMERS = 2
RENDER_ACCUMULATE = 4
SEM_REFERENCE, SEM_SPHERE_FAR_ROOT, SEM_UNIT_TRIANGLE_NORMALS, SEM_FIXED = 0, 1, 2, 3

_lib = None


def lib():
    """Load librt_b200.so.  Fails loudly when it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RtError(f"{path} is missing: build it with `make -C raytracing_gpu_b200/csrc` "
                      "(there is no CPU fallback)")
    L = C.CDLL(path)
    vp, i32, f32, u32p = C.c_void_p, C.c_int, C.c_float, C.POINTER(C.c_uint32)
    sig = {
        "rt_last_error": (C.c_char_p, []), "rt_version": (C.c_char_p, []), "rt_struct_size": (C.c_size_t, [i32]),
        "rt_scene_create": (i32, [C.POINTER(vp)]), "rt_scene_destroy": (None, [vp]),
        "rt_tex_solid": (i32, [vp, f32, f32, f32]), "rt_tex_checker": (i32, [vp, i32, i32]),
        "rt_perlin_tables": (i32, [vp, vp, vp]), "rt_tex_noise": (i32, [vp, i32, f32]),
        "rt_tex_turbulent": (i32, [vp, i32, f32, i32]), "rt_tex_marble": (i32, [vp, i32, f32]),
        "rt_image": (i32, [vp, vp, i32, i32, i32]), "rt_tex_image": (i32, [vp, i32]),
        "rt_mat_lambertian": (i32, [vp, i32]), "rt_mat_metal": (i32, [vp, i32, f32]),
        "rt_mat_dielectric": (i32, [vp, f32]), "rt_mat_diffuse_light": (i32, [vp, i32]),
        "rt_mat_isotropic": (i32, [vp, i32]),
        "rt_sphere": (i32, [vp, f32, f32, f32, f32, i32]),
        "rt_moving_sphere": (i32, [vp, vp, vp, f32, f32, f32, i32]),
        "rt_xy_rect": (i32, [vp, f32, f32, f32, f32, f32, i32]), "rt_xz_rect": (i32, [vp, f32, f32, f32, f32, f32, i32]),
        "rt_yz_rect": (i32, [vp, f32, f32, f32, f32, f32, i32]), "rt_box": (i32, [vp, vp, vp, i32]),
        "rt_triangle": (i32, [vp, vp, vp, vp, i32]), "rt_mesh": (i32, [vp, vp, i32, vp, i32, vp, vp, i32]),
        "rt_translate": (i32, [vp, i32, f32, f32, f32]), "rt_rotate_y": (i32, [vp, i32, f32]),
        "rt_rotate_y_sincos": (i32, [vp, i32, f32, f32]), "rt_constant_medium": (i32, [vp, i32, f32, i32]),
        "rt_constant_medium_nid": (i32, [vp, i32, f32, i32]), "rt_list": (i32, [vp, vp, i32]),
        "rt_bvh": (i32, [vp, vp, i32]), "rt_set_world": (i32, [vp, i32]),
        "rt_set_background": (i32, [vp, f32, f32, f32]),
        "rt_set_camera": (i32, [vp, vp, vp, vp, f32, f32, f32, f32, f32, f32]), "rt_set_camera_raw": (i32, [vp, vp]),
        "rt_set_semantics": (i32, [vp, C.c_uint32]), "rt_scene_hash": (i32, [vp, C.POINTER(C.c_uint64)]),
        "rt_scene_replay": (i32, [vp, vp, C.c_size_t]), "rt_scene_flatten": (i32, [vp]),
        "rt_scene_export_size": (C.c_size_t, [vp]), "rt_scene_export": (i32, [vp, vp, C.c_size_t]),
        "rt_image_height": (i32, [i32, f32]), "rt_scene_commit": (i32, [vp, i32, C.POINTER(BuildStats)]),
        "rt_scene_commit_ex": (i32, [vp, i32, C.POINTER(BuildSettings), C.POINTER(BuildStats)]),
        "rt_measure_fp32_peak": (i32, [i32, C.POINTER(C.c_double)]), "rt_debug_fail_allocation": (None, [i32]), "rt_scene_bvh_nodes": (i32, [vp, vp, C.c_size_t]),
        "rt_first_hit": (i32, [vp, vp, i32, f32, vp, vp, vp]), "rt_first_hit_static": (i32, [vp, vp, i32, f32, vp, vp, vp]),
        "rt_render": (i32, [vp, C.POINTER(RenderSettings), vp, C.POINTER(RenderStats)]),
        "rt_render_device": (i32, [vp, C.POINTER(RenderSettings), vp, vp, C.POINTER(RenderStats)]),
        "rt_quantize_frame": (i32, [vp, i32, i32, vp]),
        "rt_render_multi": (i32, [vp, C.POINTER(RenderSettings), C.POINTER(MultiSettings), vp, C.POINTER(MultiStats)]),
        "rt_catalog_scene": (i32, [vp, C.c_char_p, C.POINTER(CatalogParams), C.POINTER(C.c_float)]),
        # include/rt_obj.h
        "rt_obj_load": (i32, [C.c_char_p, C.POINTER(vp)]), "rt_obj_parse": (i32, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(vp)]),
        "rt_obj_free": (None, [vp]), "rt_obj_num_vertices": (C.c_uint32, [vp]), "rt_obj_num_triangles": (C.c_uint32, [vp]),
        "rt_obj_vertices": (vp, [vp]), "rt_obj_uvs": (vp, [vp]), "rt_obj_normals": (vp, [vp]), "rt_obj_indices": (vp, [vp]),
        "rt_obj_normals_generated": (i32, [vp]), "rt_obj_diffuse_map": (C.c_char_p, [vp]),
        "rt_obj_add_to_scene": (i32, [vp, vp, i32]),
        # include/rt_output.h
        "rt_average_frames": (i32, [vp, i32, i32, i32, vp]), "rt_write_ppm": (i32, [C.c_char_p, vp, i32, i32]),
        "rt_write_png": (i32, [C.c_char_p, vp, i32, i32]), "rt_write_npy": (i32, [C.c_char_p, vp, i32, i32]),
        "rt_draw": (i32, [vp, C.POINTER(DrawSettings), C.c_char_p, vp, C.POINTER(C.c_uint32), C.POINTER(RenderStats)]),
        "rt_checkpoint_save": (i32, [C.c_char_p, C.POINTER(CheckpointInfo), vp]),
        "rt_checkpoint_load": (i32, [C.c_char_p, C.POINTER(CheckpointInfo), vp, C.c_size_t]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)      # AttributeError here means the .so does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    # the structs above are mirrored by hand: refuse a library whose layout differs (include/rt_b200.h, rt_struct_id)
    for sid, cls in enumerate(STRUCT_MIRRORS):
        if L.rt_struct_size(sid) != C.sizeof(cls):
            raise RtError(f"{path}: sizeof({cls.__name__}) is {L.rt_struct_size(sid)} in the library, {C.sizeof(cls)} in this binding")
    _lib = L
    return L


EXPORTED_SYMBOLS = None  # filled lazily by tests from include/rt_b200.h
# in the order of enum rt_struct_id
STRUCT_MIRRORS = (RenderSettings, RenderStats, BuildSettings, BuildStats, MultiSettings, MultiStats, DrawSettings, CheckpointInfo, CatalogParams)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def image_height(width, aspect):
    return lib().rt_image_height(int(width), float(aspect))


image_height_of = image_height


def measure_fp32_peak(device=0):
    """TFLOP/s of an FMA microbenchmark on `device` (rt_measure_fp32_peak)."""
    v = C.c_double()
    _check(lib().rt_measure_fp32_peak(int(device), C.byref(v)))
    return v.value


def quantize_frame(rgb):
    rgb = _f32(rgb)
    h, w, _ = rgb.shape
    out = np.empty((h, w, 3), np.uint8)
    _check(lib().rt_quantize_frame(rgb.ctypes.data, w, h, out.ctypes.data))
    return out


def average_frames(frames):
    """include/rt_output.h: average_images' arithmetic on (n, H, W, 3) uint8 quantised frames."""
    fr = np.ascontiguousarray(frames, dtype=np.uint8)
    n, h, w, _ = fr.shape
    out = np.empty((h, w, 3), np.uint8)
    _check(lib().rt_average_frames(fr.ctypes.data, n, w, h, out.ctypes.data))
    return out


def write_ppm(path, rgb8):
    px = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _check(lib().rt_write_ppm(os.fsencode(path), px.ctypes.data, px.shape[1], px.shape[0]))


def write_png(path, rgb8):
    px = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _check(lib().rt_write_png(os.fsencode(path), px.ctypes.data, px.shape[1], px.shape[0]))


def write_npy(path, rgb_linear):
    px = np.ascontiguousarray(rgb_linear, dtype=np.float32)
    _check(lib().rt_write_npy(os.fsencode(path), px.ctypes.data, px.shape[1], px.shape[0]))


def _check(rc):
    if rc < 0:
        raise RtError(f"rt_b200 error {rc}: {lib().rt_last_error().decode()}")
    return rc


class Scene:
    """One rt_scene handle.  Method names and argument order follow include/rt_b200.h."""

    def __init__(self):
        self._h = C.c_void_p()
        _check(lib().rt_scene_create(C.byref(self._h)))

    def close(self):
        if self._h:
            lib().rt_scene_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __getattr__(self, name):
        # scalar-argument constructors forward generically: scene.sphere(cx, cy, cz, r, mat) etc.
        fn = getattr(lib(), "rt_" + name, None)
        if fn is None:
            raise AttributeError(name)

        def call(*args):
            for a in args:
                if isinstance(a, (list, tuple, np.ndarray)):
                    raise TypeError(f"rt_{name}: array argument needs the explicit wrapper")
            return _check(fn(self._h, *args))
        return call

    # ---- constructors with array arguments ------------------------------------------------------
    def moving_sphere(self, c0, c1, t0, t1, radius, mat):
        a, b = _f32(c0), _f32(c1)
        return _check(lib().rt_moving_sphere(self._h, a.ctypes.data, b.ctypes.data, t0, t1, radius, mat))

    def box(self, p0, p1, mat):
        a, b = _f32(p0), _f32(p1)
        return _check(lib().rt_box(self._h, a.ctypes.data, b.ctypes.data, mat))

    def triangle(self, v, uv, normals, mat):
        a, b = _f32(v).reshape(9), _f32(uv).reshape(6)
        n = _f32(normals).reshape(9) if normals is not None else None
        return _check(lib().rt_triangle(self._h, a.ctypes.data, b.ctypes.data, n.ctypes.data if n is not None else None, mat))

    def mesh(self, vertices, n_vertices, indices, n_triangles, uvs, normals, mat):
        v, u, n = _f32(vertices), _f32(uvs), _f32(normals)
        ix = np.ascontiguousarray(indices, dtype=np.uint32)
        return _check(lib().rt_mesh(self._h, v.ctypes.data, int(n_vertices), ix.ctypes.data, int(n_triangles), u.ctypes.data,
                                    n.ctypes.data, mat))

    def image(self, pixels, width, height, bytes_per_pixel):
        px = np.ascontiguousarray(pixels, dtype=np.uint8)
        assert px.size == width * height * bytes_per_pixel
        return _check(lib().rt_image(self._h, px.ctypes.data, width, height, bytes_per_pixel))

    def perlin_tables(self, ranvec, perm):
        r = _f32(ranvec).reshape(768)
        p = np.ascontiguousarray(perm, dtype=np.int32).reshape(768)
        return _check(lib().rt_perlin_tables(self._h, r.ctypes.data, p.ctypes.data))

    def list(self, children, n=None):
        c = np.ascontiguousarray(children, dtype=np.int32)
        return _check(lib().rt_list(self._h, c.ctypes.data, c.size if n is None else n))

    def bvh(self, children, n=None):
        c = np.ascontiguousarray(children, dtype=np.int32)
        return _check(lib().rt_bvh(self._h, c.ctypes.data, c.size if n is None else n))

    def set_camera(self, lookfrom, lookat, vup, vfov, aspect, aperture, focus_dist, t0, t1):
        a, b, c = _f32(lookfrom), _f32(lookat), _f32(vup)
        return _check(lib().rt_set_camera(self._h, a.ctypes.data, b.ctypes.data, c.ctypes.data, vfov, aspect, aperture, focus_dist, t0, t1))

    def set_camera_raw(self, cam24):
        c = _f32(cam24).reshape(24)
        return _check(lib().rt_set_camera_raw(self._h, c.ctypes.data))

    @staticmethod
    def from_catalog(name, image=None, aspect=0.0, mesh=None, camera15=None, host_argument_order=False):
        """Library scene (include/rt_catalog.h).  image: (H, W, 3|4) uint8; mesh: (vertices, indices, uvs, normals).
        host_argument_order: RNG draws passed as function arguments in the host compiler's order instead of nvcc's.
        Returns (scene, aspect_ratio)."""
        sc = Scene()
        p = CatalogParams()
        p.host_argument_order = 1 if host_argument_order else 0
        keep = []
        if image is not None:
            img = np.ascontiguousarray(image, dtype=np.uint8)
            keep.append(img)
            p.image, p.image_height, p.image_width, p.image_bytes_per_pixel = img.ctypes.data, img.shape[0], img.shape[1], img.shape[2]
        p.aspect = float(aspect)
        if mesh is not None:
            v, idx, uv, nrm = (_f32(mesh[0]), np.ascontiguousarray(mesh[1], dtype=np.uint32), _f32(mesh[2]), _f32(mesh[3]))
            cam = _f32(camera15)
            keep += [v, idx, uv, nrm, cam]
            p.vertices, p.n_vertices, p.indices, p.n_triangles = v.ctypes.data, v.size // 3, idx.ctypes.data, idx.size // 3
            p.uvs, p.normals, p.camera15 = uv.ctypes.data, nrm.ctypes.data, cam.ctypes.data
        a = C.c_float()
        _check(lib().rt_catalog_scene(sc._h, name.encode(), C.byref(p), C.byref(a)))
        return sc, a.value

    # ---- helpers with array arguments -----------------------------------------------------------
    def replay(self, program):
        program = np.ascontiguousarray(program, dtype=np.uint32)
        return _check(lib().rt_scene_replay(self._h, program.ctypes.data, program.size))

    def flatten(self):
        return _check(lib().rt_scene_flatten(self._h))

    def export(self):
        self.flatten()
        n = lib().rt_scene_export_size(self._h)
        buf = np.zeros(n, np.uint8)
        _check(lib().rt_scene_export(self._h, buf.ctypes.data, n))
        return buf

    def commit(self, device=0, builder="auto", sah_max_prims=0, ploc_radius=0, ploc_top=0, wide_nodes=WIDE_AUTO, trace=False):
        """rt_scene_commit_ex.  builder: "auto" | "sah" (GPU sweep) | "host_sah" | "ploc" | "karras"."""
        bs = BuildSettings(builder=BUILDERS[builder] if isinstance(builder, str) else int(builder), sah_max_prims=sah_max_prims,
                           ploc_radius=ploc_radius, ploc_top=ploc_top, wide_nodes=wide_nodes, flags=BUILD_TRACE if trace else 0)
        st = BuildStats()
        _check(lib().rt_scene_commit_ex(self._h, int(device), C.byref(bs), C.byref(st)))
        return st

    def bvh_nodes(self):
        """The committed scene's binary BVH: (n_nodes, 16) uint32 words (include/rt_b200.h rt_scene_bvh_nodes)."""
        cap = 1 << 16
        while True:
            buf = np.zeros((cap, 16), np.uint32)
            rc = lib().rt_scene_bvh_nodes(self._h, buf.ctypes.data, buf.nbytes)
            if rc >= 0:
                return buf[:rc].copy()
            if "too small" not in lib().rt_last_error().decode() or cap > (1 << 26):
                _check(rc)
            cap *= 8

    def first_hit(self, rays, t_min=0.001, xi=None, static=False):
        """rt_first_hit; static=True: rt_first_hit_static (no allocation, at most RT_FIRST_HIT_STATIC_MAX rays)."""
        rays = _f32(rays).reshape(-1, 7)
        n = rays.shape[0]
        ids = np.empty(n, np.int32)
        t = np.empty(n, np.float32)
        xi_a = _f32(xi) if xi is not None else None
        _check((lib().rt_first_hit_static if static else lib().rt_first_hit)(self._h, rays.ctypes.data, n, float(t_min), xi_a.ctypes.data if xi_a is not None else None,
                                  ids.ctypes.data, t.ctypes.data))
        return ids, t

    def render(self, width, height, spp, max_depth=50, seed=1984, sample_first=0, sample_count=0, flags=0, pool_paths=0,
               tile_index=0, tile_count=0, extend_defer=0, traversal=0, debug_flags=0):
        st = RenderSettings(width=width, height=height, samples_per_pixel=spp, max_depth=max_depth, seed=seed,
                            sample_first=sample_first, sample_count=sample_count, flags=flags, pool_paths=pool_paths,
                            tile_index=tile_index, tile_count=tile_count, extend_defer=extend_defer, traversal=traversal,
                            debug_flags=debug_flags)
        out = np.empty((height, width, 3), np.float32)
        stats = RenderStats()
        _check(lib().rt_render(self._h, C.byref(st), out.ctypes.data, C.byref(stats)))
        return out, stats

    def render_multi(self, width, height, spp, devices=None, split="sample", max_depth=50, seed=1984, pool_paths=0):
        """rt_render_multi: one frame on several GPUs of this node (devices=None: every visible one; devices[0] is the root
        that reduces).  Returns (image, MultiStats)."""
        st = RenderSettings(width=width, height=height, samples_per_pixel=spp, max_depth=max_depth, seed=seed, pool_paths=pool_paths)
        ms = MultiSettings(split_mode={"sample": SPLIT_SAMPLE, "tile": SPLIT_TILE}[split])
        if devices is not None:
            ms.n_devices = len(devices)
            for k, d in enumerate(devices):
                ms.device_ids[k] = int(d)
        out = np.empty((height, width, 3), np.float32)
        stats = MultiStats()
        _check(lib().rt_render_multi(self._h, C.byref(st), C.byref(ms), out.ctypes.data, C.byref(stats)))
        return out, stats

    def scene_hash(self):
        h = C.c_uint64()
        _check(lib().rt_scene_hash(self._h, C.byref(h)))
        return h.value

    def render_into(self, sums, spp_total, sample_first, sample_count, max_depth=50, seed=1984, flags=0, pool_paths=0):
        """Progressive rendering (RT_RENDER_ACCUMULATE): adds sample indices [sample_first, sample_first + sample_count)
        of every pixel to `sums` (H x W x 3 float32, modified in place)."""
        assert sums.dtype == np.float32 and sums.flags["C_CONTIGUOUS"] and sums.ndim == 3 and sums.shape[2] == 3
        st = RenderSettings(width=sums.shape[1], height=sums.shape[0], samples_per_pixel=spp_total, max_depth=max_depth, seed=seed,
                            sample_first=sample_first, sample_count=sample_count, flags=flags | RENDER_ACCUMULATE, pool_paths=pool_paths)
        stats = RenderStats()
        _check(lib().rt_render(self._h, C.byref(st), sums.ctypes.data, C.byref(stats)))
        return stats

    def draw(self, image_width, aspect_ratio, samples_per_pixel_per_fb, no_fb, max_depth=50, seed=1984, png_path=None,
             image_height=0, device=0, n_devices=0):
        """include/rt_output.h rt_draw: the reference's draw() (render.h:118) — batches, 8-bit quantisation, averaging, PNG."""
        st = DrawSettings(image_width=image_width, image_height=image_height, aspect_ratio=aspect_ratio,
                          samples_per_pixel_per_fb=samples_per_pixel_per_fb, no_fb=no_fb, max_depth=max_depth, seed=seed, device=device,
                          n_devices=n_devices)
        h = image_height or (image_height_of(image_width, aspect_ratio) if aspect_ratio > 0 and image_width > 0 else 0)
        out = np.empty((max(h, 1), max(image_width, 1), 3), np.uint8)
        hh = C.c_uint32()
        stats = RenderStats()
        _check(lib().rt_draw(self._h, C.byref(st), os.fsencode(png_path) if png_path else None, out.ctypes.data, C.byref(hh), C.byref(stats)))
        assert hh.value == h
        return out, stats

    def render_device(self, d_accum_ptr, width, height, spp, max_depth=50, seed=1984, sample_first=0, sample_count=0,
                      flags=0, pool_paths=0, stream=None, tile_index=0, tile_count=0, extend_defer=0, traversal=0, debug_flags=0):
        """Accumulate sums into a caller-owned device buffer (e.g. a torch CUDA tensor's data_ptr())."""
        st = RenderSettings(width=width, height=height, samples_per_pixel=spp, max_depth=max_depth, seed=seed,
                            sample_first=sample_first, sample_count=sample_count, flags=flags, pool_paths=pool_paths,
                            tile_index=tile_index, tile_count=tile_count, extend_defer=extend_defer, traversal=traversal,
                            debug_flags=debug_flags)
        stats = RenderStats()
        _check(lib().rt_render_device(self._h, C.byref(st), C.c_void_p(d_accum_ptr), C.c_void_p(stream or 0), C.byref(stats)))
        return stats


class ProgressiveRender:
    """Progressive / resumable rendering over rt_render(RT_RENDER_ACCUMULATE) and rt_checkpoint_save / _load
    (include/rt_output.h).  The frame is the same whatever the slicing: sample k of pixel p draws its
    random numbers from (p, k) alone.

        pr = ProgressiveRender(scene, 640, 360, spp_total=1024)
        while not pr.done:
            pr.step(64); pr.save("frame.ckpt")          # display pr.image() in between
        # after a restart:  pr = ProgressiveRender.resume(scene, "frame.ckpt", spp_total=1024)
    """

    def __init__(self, scene, width, height, spp_total, max_depth=50, seed=1984):
        self.scene, self.spp_total, self.max_depth, self.seed = scene, int(spp_total), int(max_depth), int(seed)
        self.sums = np.zeros((height, width, 3), np.float32)
        self.samples_done = 0

    @property
    def done(self):
        return self.samples_done >= self.spp_total

    def step(self, n_samples):
        n = min(int(n_samples), self.spp_total - self.samples_done)
        if n <= 0:
            return None
        stats = self.scene.render_into(self.sums, self.spp_total, self.samples_done, n, max_depth=self.max_depth, seed=self.seed)
        self.samples_done += n
        return stats

    def image(self):
        """Linear means over the samples rendered so far (row 0 = bottom, as rt_render)."""
        if self.samples_done == 0:
            raise RtError("no samples rendered yet")
        return (np.float32(1.0) / np.float32(self.samples_done)) * self.sums

    def save(self, path):
        info = CheckpointInfo(width=self.sums.shape[1], height=self.sums.shape[0], seed=self.seed, max_depth=self.max_depth,
                              samples_done=self.samples_done, scene_hash=self.scene.scene_hash())
        _check(lib().rt_checkpoint_save(os.fsencode(path), C.byref(info), self.sums.ctypes.data))

    @staticmethod
    def peek(path):
        info = CheckpointInfo()
        _check(lib().rt_checkpoint_load(os.fsencode(path), C.byref(info), None, 0))
        return info

    @classmethod
    def resume(cls, scene, path, spp_total):
        """Continue the render `path` holds.  Refuses a checkpoint of another scene (rt_scene_hash) or one that already
        holds more samples than spp_total."""
        info = cls.peek(path)
        if info.scene_hash != scene.scene_hash():
            raise RtError(f"{path}: checkpoint belongs to another scene (or other semantics)")
        if info.samples_done > spp_total:
            raise RtError(f"{path}: holds {info.samples_done} samples per pixel, more than spp_total={spp_total}")
        pr = cls(scene, info.width, info.height, spp_total, max_depth=info.max_depth, seed=info.seed)
        _check(lib().rt_checkpoint_load(os.fsencode(path), C.byref(info), pr.sums.ctypes.data, pr.sums.size))
        pr.samples_done = info.samples_done
        return pr


class ObjMesh:
    """include/rt_obj.h: a Wavefront OBJ parsed by the library's own loader (replaces the reference's assimp path)."""

    def __init__(self, path=None, text=None, mtl_text=None):
        self._h = C.c_void_p()
        if path is not None:
            _check(lib().rt_obj_load(os.fsencode(path), C.byref(self._h)))
        else:
            t = text.encode() if isinstance(text, str) else bytes(text)
            m = None if mtl_text is None else (mtl_text.encode() if isinstance(mtl_text, str) else bytes(mtl_text))
            _check(lib().rt_obj_parse(t, len(t), m, len(m) if m is not None else 0, C.byref(self._h)))

    def __del__(self):
        try:
            if self._h:
                lib().rt_obj_free(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def _array(self, fn, count, dtype):
        ptr = fn(self._h)
        if not ptr or count == 0:
            return np.zeros(0, dtype)
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float if dtype == np.float32 else C.c_uint32)), shape=(count,)).copy()

    @property
    def n_vertices(self):
        return int(lib().rt_obj_num_vertices(self._h))

    @property
    def n_triangles(self):
        return int(lib().rt_obj_num_triangles(self._h))

    @property
    def vertices(self):
        return self._array(lib().rt_obj_vertices, 3 * self.n_vertices, np.float32).reshape(-1, 3)

    @property
    def uvs(self):
        return self._array(lib().rt_obj_uvs, 2 * self.n_vertices, np.float32).reshape(-1, 2)

    @property
    def normals(self):
        return self._array(lib().rt_obj_normals, 3 * self.n_vertices, np.float32).reshape(-1, 3)

    @property
    def indices(self):
        return self._array(lib().rt_obj_indices, 3 * self.n_triangles, np.uint32).reshape(-1, 3)

    @property
    def normals_generated(self):
        return bool(lib().rt_obj_normals_generated(self._h))

    @property
    def diffuse_map(self):
        return lib().rt_obj_diffuse_map(self._h).decode()

    def add_to_scene(self, scene, mat):
        return _check(lib().rt_obj_add_to_scene(scene._h, self._h, int(mat)))
