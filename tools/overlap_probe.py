"""Experiment: two wavefront pipelines sharing one GPU (half-size persistent grids, RT_GRID_DIV=2), each rendering half
of the samples into the same accumulation buffer on its own stream, from two host threads — does extend of one overlap
the latency-bound shade of the other?   python tools/overlap_probe.py [spp] [lag_ms]"""
import os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import raytracing_gpu_b200 as rt

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 128
lag = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
name, width, aspect, _, _ = bench.WORKLOADS[os.environ.get("WORKLOAD", "final")]
height = rt.image_height(width, np.float32(aspect))
dev = torch.device("cuda", 0)
accum = torch.zeros(height, width, 3, device=dev)

def scene():
    sc, _ = bench.build_scene(rt, name, aspect)
    sc.commit(0)
    return sc

def run(scenes, streams, pool):
    """every scene renders its contiguous share of spp on its stream; returns ms (CUDA events on the default stream)"""
    n = len(scenes)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    accum.zero_()
    e0.record()
    for s in streams: s.wait_stream(torch.cuda.current_stream())
    def work(k):
        if k: time.sleep(lag * 1e-3)
        first = k * spp // n; count = (k + 1) * spp // n - first
        scenes[k].render_device(accum.data_ptr(), width, height, spp, seed=1984, sample_first=first, sample_count=count,
                                pool_paths=pool, stream=streams[k].cuda_stream)
    th = [threading.Thread(target=work, args=(k,)) for k in range(n)]
    [t.start() for t in th]; [t.join() for t in th]
    for s in streams: torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)

os.environ.pop("RT_GRID_DIV", None)
one = [scene()]
s1 = [torch.cuda.Stream()]
for _ in range(2): run(one, s1, 1 << 22)
t_one = [run(one, s1, 1 << 22) for _ in range(3)]
ref = (accum / spp).cpu().numpy()
os.environ["RT_GRID_DIV"] = "2"
two = [scene(), scene()]
s2 = [torch.cuda.Stream(), torch.cuda.Stream()]
for pool in (1 << 21, 1 << 22):
    for _ in range(2): run(two, s2, pool)
    t_two = [run(two, s2, pool) for _ in range(3)]
    img = (accum / spp).cpu().numpy()
    ok = np.allclose(img, ref, rtol=1e-3, atol=1e-4)
    npix = width * height
    print(f"spp {spp} lag {lag} ms | one pipeline (4Mi pool): {min(t_one):.1f} ms = {npix * spp / min(t_one) / 1e3:.0f} Msamples/s | "
          f"two pipelines (pool {pool >> 20}Mi each): {min(t_two):.1f} ms = {npix * spp / min(t_two) / 1e3:.0f} Msamples/s | same image: {ok}", flush=True)
