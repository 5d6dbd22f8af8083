"""Runs the reference CUDA build's device intersection code (oracle/_ref/ref_cuda[_nofma] probe) on fixed ray sets and
compares hit distances with this core (GPU box)."""
import os, subprocess, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import raytracing_gpu_b200 as rt
from tests import helpers
for exe_name in ("ref_cuda", "ref_cuda_nofma"):
    exe = os.path.join(helpers.ROOT, "oracle", "_ref", exe_name)
    for name in ("cornell", "cornell_smoke", "big1", "final"):
        img = bench.synthetic_earth() if name == "final" else None
        sc, _ = rt.Scene.from_catalog(name, image=img, aspect=float(np.float32(16.0) / np.float32(9.0)) if name == "final" else 0.0)
        sc.commit(0)
        rays, _ = helpers.make_rays(sc.export(), 20000, seed=99)
        ids, t = sc.first_hit(rays)                      # media skipped (stochastic on the reference side)
        with tempfile.TemporaryDirectory() as d:
            rays.tofile(os.path.join(d, "rays.f32"))
            extra = []
            if name == "final":
                img.tofile(os.path.join(d, "earth.rgb8")); extra = [os.path.join(d, "earth.rgb8"), "3410", "1518"]
            subprocess.check_call([exe, name, "probe", os.path.join(d, "rays.f32"), os.path.join(d, "out.f32")] + extra, stdout=subprocess.DEVNULL)
            out = np.fromfile(os.path.join(d, "out.f32"), np.float32)
        n = len(rays); rt_ = out[:n]
        both = np.isfinite(t) & np.isfinite(rt_)
        agree = (np.isfinite(t) == np.isfinite(rt_)).mean()
        rel = np.abs(t[both] - rt_[both]) / np.abs(rt_[both])
        print(f"{exe_name} {name}: rays={n} hit/miss agreement={agree:.4f} both-hit={both.sum()} rel err: max={rel.max():.3e} p99={np.quantile(rel,0.99):.3e} within1e-5={(rel<=1e-5).mean():.4f} bit-identical={(t[both]==rt_[both]).mean():.4f}")
