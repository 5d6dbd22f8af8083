"""Where does rt_scene_commit spend its time?  Host wall clock per phase (RT_COMMIT_TRACE) + the build's CUDA-event stats.
    python tools/commit_probe.py [workload]      (bench.py workloads: final, mesh, cornell_smoke, big1)"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import raytracing_gpu_b200 as rt
name = sys.argv[1] if len(sys.argv) > 1 else "mesh"
scene, width, aspect, spp, desc = bench.WORKLOADS[name]
sc, _ = bench.build_scene(rt, scene, aspect)
t0 = time.time(); sc.export(); print(f"flatten + export: {1e3 * (time.time() - t0):.1f} ms", flush=True)
for k in range(4):
    if k == 3: os.environ["RT_COMMIT_TRACE"] = "1"
    t0 = time.time(); st = sc.commit(0); dt = 1e3 * (time.time() - t0)
    print(f"commit {k}: {dt:.1f} ms wall | events: upload {st.ms_upload:.1f} morton {st.ms_morton:.1f} sort {st.ms_sort:.1f} "
          f"hierarchy {st.ms_hierarchy:.1f} refit {st.ms_refit:.1f} collapse {st.ms_collapse:.1f} total {st.ms_total:.1f} | "
          f"builder {st.builder} rounds {st.build_rounds} prims {st.n_prims}", flush=True)
