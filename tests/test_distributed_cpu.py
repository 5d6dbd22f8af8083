"""World-size-2 test of the multi-GPU host logic on CPU (gloo): sample-split + one sum-reduce.  The per-rank render
is done by the C oracle (test infrastructure) because there is no GPU here; on the B200 box the same
render_distributed() drives Scene.render_device (see bench.py)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import raytracing_gpu_b200 as rt
from raytracing_gpu_b200.parallel import film_order, render_distributed, sample_range, tile_mask
from tests import helpers

W, H, SPP = 72, 40, 7     # 2880 pixels = 3 film blocks of 1024: both ranks own some


def test_sample_ranges_partition_exactly():
    for spp in (1, 7, 16, 1024, 4097):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                f, c = sample_range(spp, world, r)
                seen += list(range(f, f + c))
            assert seen == list(range(spp))


def test_film_order_and_tile_masks_partition_the_frame():
    for w, h in [(24, 24), (100, 37), (8, 4), (5, 3), (1920, 1079)]:
        order = film_order(w, h)
        assert sorted(order.tolist()) == list(range(w * h))          # a permutation of the frame
        if w >= 8 and h >= 4:
            first = order[:32]
            assert set(first // w) == {0, 1, 2, 3} and set(first % w) == set(range(8))   # first warp = one 8x4 tile
        for world in (1, 2, 3, 8):
            total = sum(tile_mask(w, h, world, r).astype(int) for r in range(world))
            assert (total == 1).all()


def _worker(rank, world, port, blob, out_path, partition="sample"):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    o = helpers.OracleScene(blob)

    def render_sums(first, count, accum, tile_index, tile_count):
        mean, _ = o.render(W, H, count, sample_first=first, workers=1)
        share = mean * np.float32(count)
        if tile_count > 1:      # the oracle renders whole frames: keep this rank's film blocks only
            share = share * tile_mask(W, H, tile_count, tile_index)[..., None]
        accum += torch.from_numpy(share.astype(np.float32))
    img = render_distributed(render_sums, H, W, SPP, torch.device("cpu"), partition=partition)
    if rank == 0:
        np.save(out_path, img.numpy())
    dist.destroy_process_group()


import pytest


@pytest.mark.parametrize("partition", ["sample", "tile"])
def test_two_rank_split_equals_single_rank(tmp_path, partition):
    g = helpers.load_golden("cornell_smoke")
    sc = rt.Scene(); sc.replay(g["program"])
    blob = sc.export()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "img.npy")
    mp.spawn(_worker, args=(2, port, blob, out, partition), nprocs=2, join=True)
    two = np.load(out)
    one, _ = helpers.OracleScene(blob).render(W, H, SPP)
    np.testing.assert_allclose(two, one, rtol=2e-5, atol=1e-6)
