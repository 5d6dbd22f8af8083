"""Where does rt_scene_commit spend its time?  Host wall clock per phase (RT_BUILD_TRACE, stderr) + the build's CUDA-event
stats, for the GPU sweep and for round 1's host sweep.  usage: python tools/commit_probe.py <workload> [reps]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import raytracing_gpu_b200 as rt  # noqa: E402
from workloads import WORKLOADS  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "final"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
scene, width, aspect, _, desc = WORKLOADS[name]
sc, _ = bench.build_scene(rt, scene, aspect)
print(json.dumps({"workload": desc, "host_cores": os.cpu_count()}))
for builder in ("auto", "host_sah", "ploc"):
    for k in range(reps):
        t0 = time.perf_counter()
        st = sc.commit(0, builder=builder, trace=(k == reps - 1))
        wall = (time.perf_counter() - t0) * 1e3
        print(json.dumps({"builder_asked": builder, "builder": st.builder, "rep": k, "wall_ms": round(wall, 3), "host_ms": round(st.ms_host, 3),
                          "gpu_ms": round(st.ms_total, 3), "upload": round(st.ms_upload, 3), "morton": round(st.ms_morton, 3),
                          "sort": round(st.ms_sort, 3), "hierarchy": round(st.ms_hierarchy, 3), "refit": round(st.ms_refit, 3),
                          "permute": round(st.ms_collapse, 3), "prims": st.n_prims, "depth": st.bvh_depth, "sah": round(st.sah_internal, 4)}), flush=True)
