"""Small end-to-end pass for compute-sanitizer (memcheck / racecheck): every kernel family once, at sizes a sanitizer finishes in
a minute — GPU SAH sweep + PLOC + Karras builds, wide collapse, first-hit (both entry points), the wavefront with the block-tile and
the warp-tile shade, the deferred extend with a small pool (continuation records, finished-resumed region, dead slots), the
multi-GPU entry on one device (reduce kernel).  usage: compute-sanitizer --tool memcheck python tools/sanitize_probe.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import raytracing_gpu_b200 as rt  # noqa: E402
from tests import helpers  # noqa: E402

C = rt.capi
for name in ("cornell_smoke", "final", "mesh"):
    g = helpers.load_golden(name)
    sc = rt.Scene()
    sc.replay(g["program"])
    for builder, wide in (("auto", C.WIDE_OFF), ("ploc", C.WIDE_ON), ("karras", C.WIDE_OFF), ("auto", C.WIDE_ON)):
        st = sc.commit(0, builder=builder, wide_nodes=wide)
        ids, t = sc.first_hit(g["rays"][:512], xi=g["xi"][:512])
        ids2, t2 = sc.first_hit(g["rays"][:512], xi=g["xi"][:512], static=True)
        assert (ids == ids2).all() and (ids == g["ref_id"][:512]).all()
    kw = dict(width=48, height=32, spp=3, seed=5, pool_paths=2048)
    a, sa = sc.render(extend_defer=C.DEFER_OFF, **kw)                                   # wide tree, single-pass extend
    b, sb = sc.render(extend_defer=5, **kw)                                             # deferred extend, tiny pool
    try:
        c, s_c = sc.render(extend_defer=C.DEFER_OFF, debug_flags=C.DEBUG_WARP_SHADE, **kw)  # warp-tile shade (experiments build)
    except rt.RtError:
        c, s_c = a, sa
    d, sd = sc.render(extend_defer=C.DEFER_OFF, debug_flags=C.DEBUG_NO_GRAPH, **kw)
    sc.commit(0, wide_nodes=C.WIDE_OFF)
    e, se = sc.render(extend_defer=5, **kw)                                             # binary tree, deferred
    assert sa.rays == sb.rays == s_c.rays == sd.rays == se.rays
    m, ms = sc.render_multi(48, 32, 3, devices=[0], seed=5, pool_paths=2048)
    assert ms.rays == sa.rays
    print(name, "ok: prims", st.n_prims, "rays", sa.rays, flush=True)
print("sanitize probe done")
