import sys, numpy as np
sys.path.insert(0, '.')
sys.path.insert(0, 'tests')
import raytracing_gpu_b200 as rt
import bench
from workloads import WORKLOADS, image_height
for name in ("final", "cornell_smoke", "big1", "two_perlin", "mesh"):
    scene, width, aspect, _, desc = WORKLOADS[name]
    sc, _ = bench.build_scene(rt, scene, aspect)
    sc.commit(0)
    w, h = 192, image_height(192, aspect)
    for pool in (0, 4096):
        a, sa = sc.render(w, h, 8, seed=3, pool_paths=pool)
        b, sb = sc.render(w, h, 8, seed=3, pool_paths=pool, debug_flags=rt.capi.DEBUG_BLOCK_SHADE)
        ok = np.allclose(a, b, rtol=1e-4, atol=1e-5)
        print(name, pool, "samples", sa.samples, sb.samples, "rays", sa.rays, sb.rays, "iters", sa.iterations, sb.iterations, "match", ok, float(np.abs(a-b).max()), flush=True)
