"""rt_first_hit_static — the allocation-free variant of rt_first_hit (include/rt_b200.h; no counterpart in the reference).
Collected last on purpose."""
import numpy as np
import pytest

import raytracing_gpu_b200 as rt
from tests import helpers

pytestmark = pytest.mark.gpu


def test_first_hit_static_equals_first_hit_on_two_calls_in_a_row():
    """Through the C-ABI: same (leaf id, t) as rt_first_hit, bit for bit (integer ids; t is the same arithmetic in the same
    kernel body, so the tolerance is zero), on two different ray sets in a row — the second call reuses the shared static
    array that still holds the first call's data — and a request beyond the stated maximum is refused with a clear error."""
    g = helpers.load_golden("final")
    sc = rt.Scene()
    sc.replay(g["program"])
    sc.commit(0)
    blob = sc.export()
    for n, seed, with_xi in ((50000, 7, True), (777, 8, True), (20000, 9, False)):   # shrinking n: stale words stay behind the second call's
        rays, xi = helpers.make_rays(blob, n, seed=seed)
        xi = xi if with_xi else None
        want_id, want_t = sc.first_hit(rays, xi=xi)
        got_id, got_t = sc.first_hit(rays, xi=xi, static=True)
        assert (got_id == want_id).all()
        assert (got_t.view(np.uint32) == want_t.view(np.uint32)).all()
        assert (want_id >= 0).mean() > 0.3          # the comparison is not about misses only
    # a second scene on the same GPU shares the array
    g2 = helpers.load_golden("cornell_smoke")
    s2 = rt.Scene()
    s2.replay(g2["program"])
    s2.commit(0)
    a = s2.first_hit(g2["rays"], xi=g2["xi"])
    b = s2.first_hit(g2["rays"], xi=g2["xi"], static=True)
    assert (a[0] == b[0]).all() and (a[1].view(np.uint32) == b[1].view(np.uint32)).all()
    helpers.assert_first_hit_parity(s2.export(), b[0], b[1], g2["ref_id"], g2["ref_t"])
    too_many = np.zeros((65537, 7), np.float32)
    too_many[:, 5] = 1.0
    with pytest.raises(rt.RtError, match="RT_FIRST_HIT_STATIC_MAX"):
        sc.first_hit(too_many, static=True)
    ids, t = sc.first_hit(np.zeros((0, 7), np.float32), static=True)
    assert ids.size == 0 and t.size == 0
