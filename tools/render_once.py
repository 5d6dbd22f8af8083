"""One frame of a workload through rt_render_device (for ncu / sanitizer captures).
usage: python tools/render_once.py <workload> <spp> [debug_flags] [width]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import raytracing_gpu_b200 as rt  # noqa: E402
from workloads import SEED, WORKLOADS, image_height  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "final"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
dbg = int(sys.argv[3]) if len(sys.argv) > 3 else 0
scene, width, aspect, _, desc = WORKLOADS[name]
if len(sys.argv) > 4:
    width = int(sys.argv[4])
height = image_height(width, aspect)
sc, _ = bench.build_scene(rt, scene, aspect)
sc.commit(0)
accum = torch.zeros(height, width, 3, device="cuda:0")
st = sc.render_device(accum.data_ptr(), width, height, spp, seed=SEED, debug_flags=dbg, stream=torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
print(name, width, height, spp, "debug", dbg, "ms", round(st.ms_total, 3), "iterations", st.iterations, "rays", st.rays)
