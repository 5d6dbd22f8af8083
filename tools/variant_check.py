"""GPU check: run-time variants of the wavefront (RT_SUSPEND = deferred extend, RT_LANES = two lanes) trace the same
paths as the default — same sample and ray counts, same image up to the order of the float atomics.
    python tools/variant_check.py ["VAR=val VAR=val" ...]"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import raytracing_gpu_b200 as rt
from tests import helpers
variants = sys.argv[1:] or ["RT_SUSPEND=8", "RT_SUSPEND=16", "RT_LANES=2", "RT_LANES=2 RT_SUSPEND=10"]
KEYS = ("RT_SUSPEND", "RT_LANES", "RT_LANE_GRID")
for name in ["final", "cornell_smoke", "big1", "mesh"]:
    g = helpers.load_golden(name)
    res = {}
    for label in [""] + variants:
        for k in KEYS: os.environ.pop(k, None)
        for kv in label.split(): os.environ[kv.split("=")[0]] = kv.split("=")[1]
        sc = rt.Scene(); sc.replay(g["program"]); sc.commit(0)
        img, st = sc.render(256, 144, 8, seed=7, flags=1, pool_paths=65536)
        res[label or "default"] = (img, st.rays, st.samples, st.nodes_visited, st.node_slots, st.kernel_launches)
    base = res["default"]
    for label, r in res.items():
        ok = r[1] == base[1] and r[2] == base[2] and np.allclose(r[0], base[0], rtol=1e-4, atol=1e-5)
        print(name, f"[{label}]", "rays", r[1], "nodes", r[3], "util", round(r[3] / max(1, r[4]), 3), "launches", r[5], "OK" if ok else "MISMATCH", flush=True)
        assert ok
print("variant_check ok")
