"""Host SAH vs PLOC over mesh sizes: build time, node visits per ray, extend time (GPU box)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import raytracing_gpu_b200 as rt
from tests import helpers
img = np.asarray(helpers.procedural_image(256, 256, seed=13))
for nt in (32, 64, 90, 128, 180, 256):
    mesh = helpers.procedural_mesh(nt, nt, seed=3)
    for which in ("ploc", "sah"):
        os.environ["RT_BVH"] = which
        sc, _ = rt.Scene.from_catalog("mesh", image=img, mesh=mesh, camera15=helpers.MESH_CAMERA)
        sc.commit(0)
        t0 = time.time(); st = sc.commit(0); wall = (time.time() - t0) * 1e3
        _, rs = sc.render(960, 540, 16, flags=1)
        _, rs2 = sc.render(960, 540, 16, flags=2)
        print(f"{which:5s} prims={st.n_prims:7d} commit_wall_ms={wall:7.1f} hier_ms={st.ms_hierarchy + st.ms_refit:7.2f} depth={st.bvh_depth:3d} sah={st.sah_internal:7.2f} "
              f"nodes/ray={rs.nodes_visited / rs.rays:6.2f} extend_ms={rs2.ms_extend:6.2f}", flush=True)
