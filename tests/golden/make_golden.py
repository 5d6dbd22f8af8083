"""Generate tests/golden/*.npz from the reference itself (oracle/_ref = the reference's own sources compiled for the
host).  Run in the build container, where /root/reference exists:

    make -C oracle ref && python tests/golden/make_golden.py

Each fixture holds: the scene program exported from the reference's object graph, a fixed ray set, the
reference's first-hit (leaf id, t) for it (media draws scripted with `xi`), and for some scenes a converged
low-resolution linear image rendered by the reference's own render()/color_f().
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import helpers  # noqa: E402
import raytracing_gpu_b200 as rt  # noqa: E402

N_RAYS = 2048
# scene -> (image width, spp) for the converged reference image (height follows the reference's own rule)
IMAGES = {"first": (64, 1024), "big1": (80, 1024), "two_perlin": (64, 512), "cornell": (48, 2048), "cornell_smoke": (48, 2048),
          "earth": (64, 512), "final": (64, 1024), "mesh": (64, 512), "triangles": (64, 256), "media_mix": (64, 1024)}


def build(L, name):
    if name == "earth":
        img = helpers.procedural_image(256, 128, seed=11)
        return L.ref_scene_earth(img.ctypes.data, 256, 128, 3)
    if name == "final":
        img = helpers.procedural_image(128, 64, seed=12)
        return L.ref_scene_final(img.ctypes.data, 128, 64, 3, 1.0)
    if name == "mesh":
        v, idx, uv, nrm = helpers.procedural_mesh(16, 24)
        img = helpers.procedural_image(64, 64, seed=13)
        return L.ref_scene_mesh(v.ctypes.data, len(v), idx.ctypes.data, len(idx) // 3, uv.ctypes.data, nrm.ctypes.data,
                                img.ctypes.data, 64, 64, 3, helpers.MESH_CAMERA.ctypes.data)
    return L.ref_scene_named(name.encode())


# second set: converged reference renders at 256 pixels wide for the per-channel RMSE / PSNR test (float16 keeps the
# fixtures small; its 5e-4 relative rounding is far below the Monte-Carlo noise at these sample counts)
WIDE_IMAGES = {"final": (256, 1024), "cornell_smoke": (256, 1024), "big1": (256, 512)}


def main_wide():
    L = helpers.ref_lib()
    assert L is not None, "build oracle/_ref first"
    threads = os.cpu_count() or 1
    for name, (w, spp) in WIDE_IMAGES.items():
        t0 = time.time()
        h = build(L, name)
        hgt = L.ref_image_height(h, w)
        fb = np.zeros((hgt, w, 3), np.float32)
        L.ref_render(h, w, hgt, spp, 50, 0, threads, fb.ctypes.data)
        np.savez_compressed(os.path.join(helpers.GOLDEN, name + "_256.npz"), image=fb.astype(np.float16), image_spp=np.int32(spp))
        print(f"{name}_256: {w}x{hgt}@{spp} mean={fb.mean():.4f} max={fb.max():.2f} ({time.time() - t0:.1f}s)")


def main():
    L = helpers.ref_lib()
    assert L is not None, "build oracle/_ref first"
    threads = os.cpu_count() or 1
    only = [a for a in sys.argv[1:] if not a.startswith("--")]
    for k, name in enumerate(helpers.NAMED_SCENES + ["earth", "final", "mesh", "media_mix"]):
        if only and name not in only:
            continue
        t0 = time.time()
        h = build(L, name)
        assert h >= 0, name
        prog = helpers.ref_program(L, h)
        sc = rt.Scene()
        sc.replay(prog)
        blob = sc.export()
        rays, xi = helpers.make_rays(blob, N_RAYS, seed=1000 + k)
        ids = np.empty(N_RAYS, np.int32)
        t = np.empty(N_RAYS, np.float32)
        L.ref_first_hit(h, rays.ctypes.data, N_RAYS, 0.001, xi.ctypes.data, ids.ctypes.data, t.ctypes.data, threads)
        out = dict(program=prog, rays=rays, xi=xi, ref_id=ids, ref_t=t, num_leaves=np.int32(L.ref_num_leaves(h)))
        if name in IMAGES:
            w, spp = IMAGES[name]
            hgt = L.ref_image_height(h, w)
            fb = np.zeros((hgt, w, 3), np.float32)
            L.ref_render(h, w, hgt, spp, 50, 0, threads, fb.ctypes.data)
            out.update(image=fb, image_spp=np.int32(spp))
        np.savez_compressed(os.path.join(helpers.GOLDEN, name + ".npz"), **out)
        print(f"{name}: leaves={int(out['num_leaves'])} hits={(ids >= 0).sum()}/{N_RAYS} program={prog.size} words "
              f"image={'%dx%d@%d' % (w, hgt, spp) if name in IMAGES else '-'} ({time.time() - t0:.1f}s)")


if __name__ == "__main__":
    if "--wide" in sys.argv:
        main_wide()        # only the 256-pixel images (the first set is not regenerated)
    elif len(sys.argv) > 1:
        main()             # only the scenes named on the command line
    else:
        main()
        main_wide()
