import sys, numpy as np
sys.path.insert(0, '.')
import raytracing_gpu_b200 as rt
import bench
from workloads import WORKLOADS, image_height
F = rt.capi.TRAVERSAL_FORCE
for name in ("final", "cornell_smoke", "big1", "two_perlin"):
    scene, width, aspect, _, desc = WORKLOADS[name]
    sc, _ = bench.build_scene(rt, scene, aspect)
    sc.commit(0)
    w, h = 192, image_height(192, aspect)
    for pool in (0, 4096):
        a, sa = sc.render(w, h, 8, seed=3, pool_paths=pool, flags=1)
        b, sb = sc.render(w, h, 8, seed=3, pool_paths=pool, flags=1, traversal=F | 16)
        print(name, pool, "rays", sa.rays, sb.rays, "nodes", sa.nodes_visited, sb.nodes_visited, "prims", sa.prim_tests, sb.prim_tests,
              "match", np.allclose(a, b, rtol=1e-4, atol=1e-5), float(np.abs(a - b).max()), flush=True)
