"""Same-box A/B of the run-time variants of the wavefront (one gpurun call; box-to-box spread on this pool is +-5 %).
usage: python tools/ab_variants.py <workload> <spp> [reps]
Prints one JSON line per variant: ms per frame (CUDA events around the whole frame, CUDA graph on), per-kernel ms from a
separate timers pass, nodes / primitive tests per ray from a counters pass."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import raytracing_gpu_b200 as rt  # noqa: E402
from workloads import SEED, WORKLOADS, image_height  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "final"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 128
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
scene, width, aspect, _, desc = WORKLOADS[name]
height = image_height(width, aspect)
sc, _ = bench.build_scene(rt, scene, aspect)
wide = {"on": rt.capi.WIDE_ON, "off": rt.capi.WIDE_OFF}.get(os.environ.get("AB_WIDE", ""), rt.capi.WIDE_AUTO)
st = sc.commit(0, wide_nodes=wide)
dev = torch.device("cuda", 0)
accum = torch.zeros(height, width, 3, device=dev)
stream = torch.cuda.current_stream().cuda_stream
F = rt.capi.TRAVERSAL_FORCE


def run(label, **kw):
    try:
        sc.render_device(accum.data_ptr(), 8, 8, 1, seed=SEED, stream=stream, **kw)
    except rt.RtError as e:
        if "experiments build" in str(e):
            return
        raise

    def once(flags=0):
        accum.zero_()
        return sc.render_device(accum.data_ptr(), width, height, spp, seed=SEED, flags=flags, stream=stream, **kw)
    once(); once()
    ms = []
    for _ in range(reps):
        torch.cuda.synchronize()
        ms.append(once().ms_total)
    t = once(flags=2)
    c = once(flags=1)
    print(json.dumps({"workload": name, "spp": spp, "variant": label, "ms": round(float(np.median(ms)), 3), "ms_all": [round(m, 2) for m in ms],
                      "msamples_per_s": round(width * height * spp / np.median(ms) / 1e3, 1),
                      "extend_ms": round(t.ms_extend, 2), "shade_ms": round(t.ms_shade, 2), "turnover_ms": round(t.ms_finish, 2),
                      "nodes_per_ray": round(c.nodes_visited / c.rays, 3), "prims_per_ray": round(c.prim_tests / c.rays, 3),
                      "lane_util": round(c.nodes_visited / max(c.node_slots, 1), 3), "rays": int(c.rays), "iterations": int(t.iterations)}), flush=True)


print(json.dumps({"workload": desc, "prims": st.n_prims, "builder": st.builder, "bvh_depth": st.bvh_depth, "sah": st.sah_internal,
                  "build_ms": st.ms_total, "host_ms": st.ms_host, "n_nodes": st.n_nodes, "n_wide_nodes": st.n_wide_nodes, "wide_depth": st.wide_depth,
                  "ms_wide": st.ms_wide}), flush=True)
deferred = st.n_prims >= 65536
tag = os.path.basename(rt.lib_path()) + (" wide" if st.n_wide_nodes else " binary")
run(f"default [{tag}]")
run("block_shade", debug_flags=rt.capi.DEBUG_BLOCK_SHADE)
if not deferred:
    run("pack_leaves", traversal=F | 16)
run("warp_shade", debug_flags=rt.capi.DEBUG_WARP_SHADE)
if os.environ.get("AB_ONLY_DEFAULT"):
    sys.exit(0)
if not deferred:
    for v in range(8):
        run(f"traversal_{v}" + "".join(s for b, s in ((1, "+tcull"), (2, "+smem"), (4, "+minmax")) if v & b), traversal=F | v)
    run("traversal_4 (= round 1 kernels)", traversal=F | 4)
    for below in (4, 8, 12):
        run(f"defer_{below}", extend_defer=below + 1)
else:
    run("defer_off", extend_defer=rt.capi.DEFER_OFF)
    run("defer_off+tcull", extend_defer=rt.capi.DEFER_OFF, traversal=F | 1)
    for below in (4, 12, 16, 20):
        run(f"defer_{below}", extend_defer=below + 1)
    run("defer_off+warp_shade", extend_defer=rt.capi.DEFER_OFF, debug_flags=rt.capi.DEBUG_WARP_SHADE)
