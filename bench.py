#!/usr/bin/env python
"""bench.py — path-tracing throughput (Msamples/s) of the B200-native core on the reference's headline scene.

One "step" = one full pass of the hot path over one frame: clear the accumulation buffer, run the wavefront
pipeline (generate / extend / shade) over every sample of the frame, (N > 1: one NCCL reduce of the float
accumulation buffers,) scale by 1/spp.

Workload (config.workload): BASELINE.json config 3, the Next-Week final scene, 1920 x 1079 (the height the
reference computes for 1920 at 16:9, render.h:33), 1024 samples per pixel PER GPU, max depth 50 — weak scaling:
at N GPUs the frame has 1024*N spp, sample-split, rank r renders samples [1024 r, 1024 (r+1)).
The earth texture is synthetic (reference assets are not redistributed); all other scene data is the scene itself.

  value      whole-job Msamples/s, scene resident in HBM, CUDA events, max over ranks
  e2e        same metric through the C-ABI with host buffers: rt_scene_commit (H2D of the flattened scene + GPU
             LBVH build) + render + reduce + D2H of the W*H*3 float image, every step
  roofline   dominant kernel (extend): algorithmic 48 B per ray segment / its CUDA-event time vs measured HBM copy
             bandwidth; the binding roofline (FP32 issue) is reported next to it as roofline_fp32
  workloads  (N = 1) the other BASELINE configs, same run: config 1, config 2 and the config-4 stand-in — value, e2e, kernel ms
  strong     BASELINE config 5: the 3840 x 2159 frame at 4096 spp TOTAL, sample-split over the N ranks + one NCCL reduce
             (reduce_ms timed separately)
  reference_cuda
             the reference's own CUDA build (oracle/_ref/ref_cuda) on the same GPU, bounded sample, N = 1 only
  cpu_baseline / --impl reference
             the reference's own render()/color_f() (oracle/_ref, its sources compiled for the host) on the box's
             host cores, on a bounded sample of the same frame
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from workloads import MESH_CAMERA, SEED, WORKLOADS, big_mesh, image_height, procedural_image, synthetic_earth  # noqa: E402

# unit utilisation of the dominant kernel from the committed ncu capture of this workload: a profiler reading recorded once per
# kernel version — NOT measured by this run.  `captured_at` names the library version; the block is dropped when it differs.
NCU_EVIDENCE = {}
_ncu_path = os.path.join(ROOT, "profiles", "ncu_evidence.json")
if os.path.exists(_ncu_path):
    NCU_EVIDENCE = json.load(open(_ncu_path))


def build_scene(rt, name, aspect):
    if name == "final":
        return rt.Scene.from_catalog("final", image=synthetic_earth(), aspect=float(np.float32(aspect)))
    if name == "mesh":
        img = np.asarray(procedural_image(2048, 2048, seed=13))
        return rt.Scene.from_catalog("mesh", image=img, mesh=big_mesh(), camera15=MESH_CAMERA)
    return rt.Scene.from_catalog(name)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[k] for r in self.rows if len(r) >= 6 for k in range(4) if r[2 + k].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        return json.load(open(path)), "measured"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


# --------------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own CPU-compiled render() on the host cores
# --------------------------------------------------------------------------------------------------------------
def reference_cpu(workload, steps, warmup, budget_s=12.0):
    """Times oracle/_ref (kind "reference") — or, where it was never built, the C restatement (kind "port") — on a
    bounded sample of the workload: the full-resolution frame at a reduced sample count."""
    from tests import helpers
    scene_name, width, aspect, spp, desc = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    L = helpers.ref_lib()
    height = image_height(width, aspect)
    if L is not None:
        if scene_name == "final":
            img = synthetic_earth()
            h = L.ref_scene_final(img.ctypes.data, img.shape[1], img.shape[0], 3, C.c_float(np.float32(aspect)))
        elif scene_name == "mesh":
            v, idx, uv, nrm = big_mesh(180, 180)   # the reference's one-thread O(n^2) BVH build cannot take 1M triangles
            img = np.asarray(procedural_image(2048, 2048, seed=13))
            h = L.ref_scene_mesh(v.ctypes.data, len(v), idx.ctypes.data, len(idx) // 3, uv.ctypes.data, nrm.ctypes.data,
                                 img.ctypes.data, 2048, 2048, 3, MESH_CAMERA.ctypes.data)
        else:
            h = L.ref_scene_named(scene_name.encode())
        fb = np.zeros((height, width, 3), np.float32)
        times = (C.c_double * 2)()

        def run(ns):
            L.ref_render(h, width, height, ns, 50, 0, cores, fb.ctypes.data)
            L.ref_last_times(times)
            return times[1]          # the render pass only; render_init is excluded like on the GPU side
        kind = "reference"
    else:
        import raytracing_gpu_b200 as rt    # only to export the scene blob the C restatement reads; nothing is rendered with it
        sc, _ = build_scene(rt, scene_name, aspect)
        o = helpers.OracleScene(sc.export())

        def run(ns):
            t0 = time.time()
            o.render(width, height, ns, seed=SEED, workers=cores)
            return time.time() - t0
        kind = "port"
    t1 = run(1)                                            # calibration pass (also the warm-up)
    ns = int(max(1, min(spp, budget_s / max(t1, 1e-3) / max(1, steps + warmup))))
    for _ in range(max(0, warmup - 1)):
        run(ns)
    ts = [run(ns) for _ in range(steps)]
    t = float(np.mean(ts))
    value = width * height * ns / t / 1e6
    sample = f"{width}x{height} full frame, {ns} of {spp} spp, depth 50, render pass only, {cores} host threads"
    return {"value": value, "unit": "Msamples/s", "cores": cores, "kind": kind, "sample": sample, "spp_timed": ns}, t * 1e3


def reference_cuda(workload, budget_s=20.0):
    """The reference's own CUDA build on this GPU (oracle/_ref/ref_cuda: its unmodified headers compiled by nvcc for
    sm_100, harness oracle/ref_cuda_harness.cu) on a bounded sample of the workload: the full-resolution frame at a
    sample count chosen from a 1-spp calibration launch.  Reported next to the product's number; never on its path."""
    import tempfile
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_cuda")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_cuda was not built (needs /root/reference at build time)"}
    scene_name, width, aspect, spp, desc = WORKLOADS[workload]
    if scene_name == "mesh":
        return {"unavailable": "the reference's one-thread O(n^2) BVH build cannot take the 1M-triangle stand-in mesh"}
    with tempfile.TemporaryDirectory() as d:
        extra = []
        if scene_name == "final":
            img = synthetic_earth()
            path = os.path.join(d, "earth.rgb8")
            img.tofile(path)
            extra = [path, str(img.shape[1]), str(img.shape[0])]

        def run(ns, timeout):
            out = subprocess.run([exe, scene_name, str(width), str(ns), "50"] + extra, capture_output=True, text=True, timeout=timeout)
            if out.returncode != 0:
                raise RuntimeError(f"ref_cuda rc={out.returncode}: {out.stderr[-300:]}")
            return json.loads(out.stdout.strip().splitlines()[-1])
        try:
            cal = run(1, 300)
            ns = int(max(1, min(spp, budget_s * 1e3 / max(cal["ms_render"], 1e-3))))
            res = cal if ns == 1 else run(ns, 600)
        except Exception as e:                               # a crash or time-out of the reference is a result, not an error of this bench
            return {"unavailable": f"reference CUDA build failed here: {e}"[:300]}
    return {"value": res["msamples_per_s"], "unit": "Msamples/s", "kind": "reference CUDA build (nvcc sm_100, unmodified headers)",
            "sample": f"{res['width']}x{res['height']} full frame, {res['spp']} of {spp} spp, depth 50, render<<<>>> launch only (CUDA events)",
            "ms_render": res["ms_render"], "ms_scene_build": res["ms_build"], "ms_render_init": res["ms_render_init"],
            "nan_pixels": res["nan_pixels"], "image_mean": res["image_mean"]}


def config_of(workload, world, width, height, spp, partition):
    """The config dict both arms print (the driver compares them key by key)."""
    return {"workload": WORKLOADS[workload][4], "width": width, "height": height, "spp_per_gpu": spp, "spp_total": spp * world,
            "max_depth": 50, "partition": f"{partition}-split x{world} + one NCCL reduce" if world > 1 else "single GPU"}


def print_reference_line(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene_name, width, aspect, spp, desc = WORKLOADS[args.workload]
    width = args.width or width
    spp = args.spp or spp
    world = max(1, args.gpus)
    cb, ms = reference_cpu(args.workload, args.steps, args.warmup, budget_s=60.0)
    cfg = config_of(args.workload, world, width, image_height(width, aspect), spp, args.partition)
    note = (f"reference's own render()/color_f() compiled for the host (oracle/_ref), CPU only; each step renders "
            f"{cb['spp_timed']} of the spp_per_gpu samples of the full frame (cpu_baseline.sample)")
    line = {"impl": "reference", "metric": "Msamples/s", "value": cb["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "note": note, "spp_timed": cb["spp_timed"],
            "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# --------------------------------------------------------------------------------------------------------------
# product arm
# --------------------------------------------------------------------------------------------------------------
class Frame:
    """One workload on this rank's GPU: committed scene, device accumulation buffer, pinned host image."""

    def __init__(self, rt, torch, dist, workload, local, rank, world, partition, pool, width=0, spp=0, strong=False):
        scene_name, w, aspect, s, self.desc = WORKLOADS[workload]
        self.rt, self.torch, self.dist = rt, torch, dist
        self.workload, self.scene_name, self.local, self.rank, self.world, self.partition, self.pool = workload, scene_name, local, rank, world, partition, pool
        self.width = width or w
        self.spp = spp or s                                   # per GPU (weak) or total (strong)
        self.total_spp = self.spp if strong else self.spp * world
        self.height = image_height(self.width, aspect)
        self.npix = self.width * self.height
        self.dev = torch.device("cuda", local)
        self.sc, _ = build_scene(rt, scene_name, aspect)
        self.build = self.sc.commit(local)
        self.blob_bytes = int(len(self.sc.export()))
        self.image_bytes = {"final": 3410 * 1518 * 4, "mesh": 2048 * 2048 * 4}.get(scene_name, 0)     # RGBA8 as uploaded
        self.stream = torch.cuda.current_stream().cuda_stream
        self.accum = torch.zeros(self.height, self.width, 3, device=self.dev, dtype=torch.float32)
        self.out = torch.empty_like(self.accum)
        self.host_img = torch.empty(self.height, self.width, 3, dtype=torch.float32).pin_memory()
        self.reduce_ms = []

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def step(self, flags=0, time_reduce=False, **kw):
        torch, dist = self.torch, self.dist
        self.accum.zero_()
        if self.partition == "tile":
            first, count, ti, tc = 0, self.total_spp, self.rank, self.world
        else:
            base, rem = divmod(self.total_spp, self.world)
            count, first, ti, tc = base + (1 if self.rank < rem else 0), self.rank * base + min(self.rank, rem), 0, 1
        st = self.sc.render_device(self.accum.data_ptr(), self.width, self.height, self.total_spp, max_depth=50, seed=SEED, sample_first=first,
                                   sample_count=count, flags=flags, pool_paths=self.pool, stream=self.stream, tile_index=ti, tile_count=tc, **kw)
        if self.world > 1:
            if time_reduce:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                dist.reduce(self.accum, dst=0, op=dist.ReduceOp.SUM)
                e1.record()
                e1.synchronize()
                self.reduce_ms.append(e0.elapsed_time(e1))
            else:
                dist.reduce(self.accum, dst=0, op=dist.ReduceOp.SUM)
        torch.mul(self.accum, 1.0 / self.total_spp, out=self.out)
        return st

    def e2e_step(self):
        self.sc.commit(self.local)
        self.step()
        if self.rank == 0:
            self.host_img.copy_(self.out, non_blocking=True)

    def timed(self, fn, k):
        torch, dist = self.torch, self.dist
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        res = [fn() for _ in range(k)]
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), res

    def msamples(self, ms_step):
        return self.npix * self.total_spp / (ms_step * 1e-3) / 1e6


KERNEL_MS_NOTE = ("per-kernel CUDA events of a separate pass with direct launches (no CUDA graph); shade = k_bin + k_shade_bins, "
                  "whose last block also runs the turnover (turnover: what is left between shade's end and the next extend)")


def short_record(fr, steps, warmup):
    """value / e2e / kernel ms of one workload (used for the per-scene `workloads` block)."""
    for _ in range(warmup):
        fr.step()
    ms, stats = fr.timed(fr.step, steps)
    ms_step = ms / steps
    _, ts = fr.timed(lambda: fr.step(flags=2), 1)
    fr.e2e_step()
    ms_e2e, _ = fr.timed(fr.e2e_step, steps)
    rays = float(np.mean([s.rays for s in stats]))
    t = ts[0]
    return {"workload": fr.desc, "width": fr.width, "height": fr.height, "spp": fr.total_spp, "steps": steps, "value": fr.msamples(ms_step),
            "unit": "Msamples/s", "ms_per_step": ms_step, "mrays_per_s": rays / (ms_step * 1e-3) / 1e6,
            "e2e": fr.msamples(ms_e2e / steps), "prims": int(fr.build.n_prims), "bvh_builder": int(fr.build.builder),
            "bvh_build_ms": float(fr.build.ms_total), "commit_host_ms": float(fr.build.ms_host),
            "kernel_ms_per_step": {"extend": t.ms_extend, "shade": t.ms_shade, "generate": t.ms_generate, "turnover": t.ms_finish,
                                   "note": KERNEL_MS_NOTE}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="final", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel per GPU")
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--pool", type=int, default=0, help="wavefront pool size in paths (0 = library default)")
    ap.add_argument("--partition", default="sample", choices=["sample", "tile"],
                    help="multi-GPU split: sample ranges (default) or interleaved film blocks; one NCCL reduce either way")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the per-scene `workloads` block and the config-5 `strong` record")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        print_reference_line(args)
        return

    import torch
    import torch.distributed as dist
    import raytracing_gpu_b200 as rt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    fr = Frame(rt, torch, dist, args.workload, local, rank, world, args.partition, args.pool, width=args.width, spp=args.spp)
    build = fr.build
    for _ in range(args.warmup):
        fr.step()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ms_total, stats = fr.timed(fr.step, args.steps)
    clk = clocks.stop() if rank == 0 else None
    ms_step = ms_total / args.steps
    value = fr.msamples(ms_step)
    rays = torch.tensor([float(np.mean([s.rays for s in stats]))], device=fr.dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rays)
    rays_per_step = float(rays.item())
    launches = int(sum(s.kernel_launches for s in stats)) + 2 * args.steps      # + zero + scale kernels per step

    # dominant kernel (extend): separate pass with per-kernel CUDA events and direct launches (not part of the timed region above)
    _, tstats = fr.timed(lambda: fr.step(flags=2), 1)
    ts = tstats[0]
    ext_launches = int(ts.iterations)
    ext_ms = ts.ms_extend / max(ext_launches, 1)
    ext_bytes = 48.0 * ts.rays / max(ext_launches, 1)                 # ray 32 B in + hit 16 B out per segment
    peaks, peak_src = measured_peaks()
    extend_kernel = "k_extend_defer" if int(build.n_prims) >= 65536 else "k_extend"   # wavefront.cu: SUSPEND_MIN_PRIMS
    achieved = ext_bytes / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else 0.0
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath)).get(args.workload, {})
        if tj.get("captured_at") == rt.lib().rt_version().decode() and "extend_dram_bytes_per_ray" in tj:
            # ncu's DRAM bytes of one full-pool launch, per ray, x the rays of this run's average launch (drain launches are smaller)
            traffic = tj["extend_dram_bytes_per_ray"] * ts.rays / max(ext_launches, 1)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": traffic, "kernel": extend_kernel, "peak_source": peak_src, "launches_per_step": ext_launches,
                "avg_launch_ms": ext_ms, "share_of_step": ts.ms_extend / max(ts.ms_total, 1e-9),
                "kernel_ms_per_step": {"extend": ts.ms_extend, "shade": ts.ms_shade, "generate": ts.ms_generate, "turnover": ts.ms_finish,
                                       "total_of_this_pass": ts.ms_total,
                                       "note": "separate pass with direct launches and 3 events per iteration: sums exceed ms_per_step, "
                                               "which is timed with 8 iterations per CUDA graph; shade = k_bin + k_shade_bins (its last block runs the turnover)"}}
    # FP32 roofline (the binding one): algorithmic flops from measured traversal counters (SURVEY §8d figures) against the FMA
    # microbenchmark run on THIS GPU
    _, cstats = fr.timed(lambda: fr.step(flags=1), 1)
    cs = cstats[0]
    flops = 24.0 * 2 * cs.nodes_visited + 45.0 * cs.prim_tests + 55.0 * cs.rays
    fp32_nominal = 148 * 128 * 2 * (peaks.get("sm_max_mhz", 1965.0) * 1e6) / 1e12
    try:
        fp32_peak, fp32_src = rt.capi.measure_fp32_peak(local), "measured: FMA microbenchmark on this GPU (rt_measure_fp32_peak), best of 5"
    except Exception as e:                                            # never lose the bench line over the probe
        fp32_peak, fp32_src = fp32_nominal, f"nominal (probe failed: {e})"
    roofline_fp32 = {"bound": "fp32", "achieved": flops / (ms_step * 1e-3) / 1e12, "peak": fp32_peak, "unit": "TFLOP/s",
                     "frac": flops / (ms_step * 1e-3) / 1e12 / fp32_peak, "peak_source": fp32_src, "peak_nominal": fp32_nominal,
                     "nodes_per_ray": cs.nodes_visited / max(cs.rays, 1), "prim_tests_per_ray": cs.prim_tests / max(cs.rays, 1),
                     "rays_per_sample": cs.rays / max(cs.samples, 1),
                     "warp_lane_utilisation_node_loop": cs.nodes_visited / max(cs.node_slots, 1)}

    # end to end through the C-ABI with host buffers: commit (H2D + BVH build) + render + reduce + D2H
    e2e = None
    if not args.no_e2e:
        fr.e2e_step()
        ms_e2e, _ = fr.timed(fr.e2e_step, args.steps)
        b2 = fr.sc.commit(local)
        e2e = {"value": fr.msamples(ms_e2e / args.steps), "unit": "Msamples/s",
               "h2d_bytes_per_step": fr.blob_bytes + fr.image_bytes, "d2h_bytes_per_step": fr.npix * 12,
               "includes": "rt_scene_commit (scene tables + image texture H2D from pinned staging, BVH build on the GPU) + render + reduce + image D2H",
               "commit_host_ms": float(b2.ms_host), "commit_gpu_ms": float(b2.ms_total), "bvh_builder": int(b2.builder)}

    # per-scene block (the metric is per scene) and BASELINE config 5 (strong scaling)
    extras, strong = None, None
    if not args.no_extras and args.workload == "final":
        if world == 1:
            extras = {}
            for name in ("big1", "cornell_smoke", "mesh"):
                f2 = Frame(rt, torch, dist, name, local, rank, world, args.partition, args.pool)
                extras[name] = short_record(f2, 2, 3)
                del f2
                torch.cuda.empty_cache()
        f5 = Frame(rt, torch, dist, "final_c5", local, rank, world, "sample", args.pool, strong=True)
        f5.sc.render_device(f5.accum.data_ptr(), f5.width, f5.height, 8, max_depth=50, seed=SEED, stream=f5.stream)   # warm-up: pools, graph
        ms5, st5 = f5.timed(lambda: f5.step(time_reduce=True), 1)
        rays5 = torch.tensor([float(st5[0].rays)], device=f5.dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(rays5)
        strong = {"workload": f5.desc, "width": f5.width, "height": f5.height, "spp_total": f5.total_spp, "n_gpus": world, "steps": 1,
                  "scaling": "strong", "value": f5.msamples(ms5), "unit": "Msamples/s", "ms_per_step": ms5,
                  "mrays_per_s": float(rays5.item()) / (ms5 * 1e-3) / 1e6, "reduce_ms": (f5.reduce_ms[-1] if f5.reduce_ms else 0.0),
                  "render_ms_this_rank": float(st5[0].ms_total), "accum_bytes": f5.npix * 12,
                  "partition": f"sample-split x{world} + one NCCL reduce" if world > 1 else "single GPU"}
        del f5

    cb = rc = None
    if rank == 0 and not args.no_cpu_baseline:
        cb, _ = reference_cpu(args.workload, 1, 1)
        if world == 1:
            rc = reference_cuda(args.workload)

    if rank == 0:
        cfg = config_of(args.workload, world, fr.width, fr.height, fr.spp, args.partition)
        extra_cfg = {"l2": "working set (pool + queues) > L2; accumulation buffer re-zeroed each step",
                     "prims": int(build.n_prims), "bvh_nodes": int(build.n_nodes), "bvh_build_ms": float(build.ms_total),
                     "bvh_builder": {0: "karras (GPU)", 1: "PLOC (GPU) + SAH-swept top (GPU)", 2: "host SAH sweep", 3: "SAH sweep (GPU)"}[int(build.builder)]}
        ncu = NCU_EVIDENCE.get(args.workload)
        if ncu and ncu.get("captured_at") != rt.lib().rt_version().decode():
            ncu = None
        line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": cfg, "scene": extra_cfg,
                "mrays_per_s": rays_per_step / (ms_step * 1e-3) / 1e6, "rays_per_sample": rays_per_step / (fr.npix * fr.total_spp),
                "gpu_launches": launches, "clocks": clk, "roofline": roofline, "roofline_fp32": roofline_fp32,
                "ncu_evidence": ncu, "e2e": e2e, "workloads": extras, "strong": strong,
                "cpu_baseline": cb, "reference_cuda": rc,
                "vs_reference_cuda": (value / rc["value"] if rc and rc.get("value") else None),
                "library": rt.lib().rt_version().decode()}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
