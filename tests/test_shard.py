"""Host-side multi-GPU logic: partitioning rules, and a world_size-2 gloo run on CPU showing that
frame-pair sharding and tile sharding reproduce the unsharded result with no data-path collective (the per-rank
compute is the CPU oracle here; on the GPU box the same plan drives the CUDA path, tests/test_gpu_parity_dat.py)."""
import os
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from videoframeinterpolation_b200.shard import default_grid, pair_range, tile_plan, tiles_of_rank  # noqa: E402


def test_pair_range_partitions():
    for n in (0, 1, 7, 8, 9, 64):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                b, e = pair_range(r, world, n)
                assert 0 <= b <= e <= n
                seen += list(range(b, e))
            assert seen == list(range(n))
            sizes = [pair_range(r, world, n)[1] - pair_range(r, world, n)[0] for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        pair_range(2, 2, 8)


@pytest.mark.parametrize("H,W,grid,halo", [(2160, 3840, (2, 4), 64), (1088, 1920, (1, 8), 40), (256, 448, (2, 2), 16),
                                           (64, 64, (1, 1), 32)])
def test_tile_plan_covers_frame(H, W, grid, halo):
    tiles = tile_plan(H, W, grid, halo)
    assert len(tiles) == grid[0] * grid[1]
    cover = torch.zeros(H, W, dtype=torch.int32)
    for t in tiles:
        assert t.y0 % 16 == 0 and t.x0 % 16 == 0 and t.y1 % 16 == 0 and t.x1 % 16 == 0      # crops stay /16
        assert t.y0 <= t.iy0 < t.iy1 <= t.y1 and t.x0 <= t.ix0 < t.ix1 <= t.x1
        assert t.iy0 - t.y0 >= min(halo, t.iy0) and t.y1 - t.iy1 >= min(halo, H - t.iy1)    # halo unless at the frame edge
        cover[t.iy0:t.iy1, t.ix0:t.ix1] += 1
    assert int(cover.min()) == 1 and int(cover.max()) == 1                                     # interiors tile the frame exactly
    with pytest.raises(ValueError):
        tile_plan(H + 1, W, grid, halo)


def test_default_grid():
    assert default_grid(8) == (2, 4) and default_grid(4) == (2, 2) and default_grid(2) == (1, 2) and default_grid(1) == (1, 1)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle
    import _det as det
    torch.set_num_threads(1)
    # --- frame-pair sharding: 5 pairs over 2 ranks, warp + splat per pair ---
    n = 5
    img = det.normal("shard.img", (n, 4, 24, 40))
    flow = det.flow_field("shard.flow", n, 24, 40, sigma=2.0)
    metric = det.normal("shard.metric", (n, 1, 24, 40))
    b, e = pair_range(rank, world, n)
    mine = torch.cat((oracle.bwarp(img[b:e], flow[b:e]), oracle.fwarp(img[b:e], flow[b:e], metric[b:e], "soft")), 1)
    # --- tile sharding: bwarp-free pointwise+local op on tiles must equal the whole frame in the interiors ---
    frame = det.normal("shard.frame", (1, 3, 64, 96))
    k = det.normal("shard.k", (3, 3, 5, 5))
    whole = torch.nn.functional.conv2d(frame, k, padding=2)
    part = torch.zeros_like(whole)
    for tl in tiles_of_rank(tile_plan(64, 96, (2, 2), 16), rank, world):
        ys, xs = tl.crop
        pred = torch.nn.functional.conv2d(frame[:, :, ys, xs], k, padding=2)
        iy, ix = tl.interior_in_crop
        oy, ox = tl.interior
        part[:, :, oy, ox] = pred[:, :, iy, ix]
    # reporting only (NOT on the data path): gather the shards to rank 0 and compare with the unsharded result
    shards = [None] * world
    dist.all_gather_object(shards, (b, e, mine, part))
    if rank == 0:
        full = torch.cat((oracle.bwarp(img, flow), oracle.fwarp(img, flow, metric, "soft")), 1)
        got = torch.cat([s[2] for s in sorted(shards, key=lambda s: s[0])], 0)
        tiles_sum = sum(s[3] for s in shards)
        q.put((bool(torch.equal(got, full)), float((tiles_sum - whole).abs().max())))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_sharding_matches_unsharded():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, tile_err = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok, "pair-sharded outputs differ from the unsharded run"
    assert tile_err < 1e-5, f"tile-sharded interiors differ from the whole frame by {tile_err}"
