"""Prints LBVH build statistics and traversal counters for one workload (GPU box; used by tools/ab.sh)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import raytracing_gpu_b200 as rt
name = sys.argv[1]; spp = int(sys.argv[2])
scene_name, width, aspect, _, _ = bench.WORKLOADS[name]
sc, _ = bench.build_scene(rt, scene_name, aspect)
sc.commit(0)
t0 = time.time(); st = sc.commit(0); wall = (time.time() - t0) * 1e3
h = rt.image_height(width, np.float32(aspect))
img, rs = sc.render(width, h, spp, flags=1)
img, rs2 = sc.render(width, h, spp, flags=2)
print(f"{name} bvh={os.environ.get('RT_BVH','auto')}/{('karras','ploc','host-sah')[st.builder]} r={os.environ.get('RT_PLOC_RADIUS','16')} prims={st.n_prims} build_ms={st.ms_total:.1f} (wall {wall:.1f}; upload {st.ms_upload:.1f} sort {st.ms_sort:.1f} hier {st.ms_hierarchy:.2f} "
      f"merge {st.ms_refit:.1f}) radius={st.ploc_radius} rounds={st.build_rounds} depth={st.bvh_depth} sah={st.sah_internal:.2f} nodes/ray={rs.nodes_visited/rs.rays:.2f} "
      f"prims/ray={rs.prim_tests/rs.rays:.2f} extend_ms={rs2.ms_extend:.1f} shade_ms={rs2.ms_shade:.1f}")
