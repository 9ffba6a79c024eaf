"""N > 1 path on CPU: world_size-2 gloo run of the tile (superblock rows x columns) sharding + gather
(rav1e_b200.shard).  The compute inside each rank is the oracle (no GPU here); what
is under test is the partition and the merge: shards are disjoint, cover every
block, and the gathered array equals the whole-frame result."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

W, H, RNG, LAM = 128, 200, 6, 21


def _planes(r1o):
    rng = np.random.default_rng(11)
    o_ref = r1o.OPlane(W, H, 0, 0, 88, 88, 1)
    o_org = r1o.OPlane(W, H, 0, 0, 88, 88, 1)
    o_ref.visible[...] = rng.integers(0, 256, size=(H, W))
    o_org.visible[...] = np.roll(o_ref.visible, (2, -1), axis=(0, 1))
    o_ref.pad()
    o_org.pad()
    return o_org, o_ref


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import r1o
    from rav1e_b200 import api, shard
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mask = (1 << api.BLOCK["BLOCK_8X8"]) | (1 << api.BLOCK["BLOCK_16X16"]) | (1 << api.BLOCK["BLOCK_64X64"])
    (lo, hi), (clo, chi) = shard.split_tiles(W, H, world)[rank]  # 128 x 200: two tile columns of 4 superblock rows
    slots = shard.shard_slots(W, H, mask, lo, hi, clo, chi)
    o_org, o_ref = _planes(r1o)
    mv, cost = r1o.frame_me(o_org, o_ref, RNG, RNG, LAM, mask, 0, 0, 0, 8, idx=slots)
    local = np.zeros(cost.size, dtype=api.ME_RESULT)
    local["mv"]["row"], local["mv"]["col"], local["cost"] = mv[:, 0], mv[:, 1], cost
    merged = shard.gather_frame_results(local, slots, cost.size, dist)
    if rank == 0:
        q.put((merged["mv"]["row"].copy(), merged["mv"]["col"].copy(), merged["cost"].copy()))
    dist.barrier()
    dist.destroy_process_group()


def test_sb_row_split_is_a_partition():
    sys.path.insert(0, ROOT)
    from rav1e_b200 import api, shard
    for world in (1, 2, 3, 4, 8):
        parts = shard.split_sb_rows(1080, world)
        assert parts[0][0] == 0 and parts[-1][1] == 17
        assert all(parts[i][1] == parts[i + 1][0] for i in range(world - 1))
        sl = [shard.shard_slots(1920, 1080, api.ALL_SIZES_TO_64, lo, hi) for lo, hi in parts]
        allidx = np.sort(np.concatenate(sl))
        p = api.make_frame_me_params(1920, 1080, 8, 8, 1)
        assert np.array_equal(allidx, np.arange(api.frame_me_count(p)))
    assert shard.split_frames(5, 2) == [[0, 2, 4], [1, 3]]


def test_tile_split_is_a_partition_and_balanced():
    sys.path.insert(0, ROOT)
    from rav1e_b200 import api, shard
    for (w, h) in ((3840, 2160), (1920, 1080), (256, 200), (64, 64)):
        p = api.make_frame_me_params(w, h, 8, 8, 1)
        rows, cols = (h + 63) // 64, (w + 63) // 64
        for world in (1, 2, 3, 4, 6, 8):
            tiles = shard.split_tiles(w, h, world)
            assert len(tiles) == world
            sl = [shard.shard_slots(w, h, api.ALL_SIZES_TO_64, lo, hi, clo, chi) for (lo, hi), (clo, chi) in tiles]
            assert np.array_equal(np.sort(np.concatenate(sl)), np.arange(api.frame_me_count(p))), (w, h, world)
            assert sum((hi - lo) * (chi - clo) for (lo, hi), (clo, chi) in tiles) == rows * cols
    # 4K over 8 ranks: 2 x 4 tiles of 17 x 15 superblocks -- every rank the same share (rows alone: 5 of 34 at most)
    t8 = shard.split_tiles(3840, 2160, 8)
    assert {(hi - lo, chi - clo) for (lo, hi), (clo, chi) in t8} == {(17, 15)}


def test_tile_rectangles_cover_what_a_tile_can_touch():
    """shard.tile_plane_rows / tile_plane_cols (what a rank uploads with r1c_plane_upload_rect_async): inside the
    allocation, on 128-byte lines, holding the tile +- (range + 16) pixels, and shrinking with the tile grid"""
    sys.path.insert(0, ROOT)
    from rav1e_b200 import api, capi, shard
    lib = capi.load()
    for px_bytes in (1, 2):
        cfg = capi.PlaneCfg()
        assert lib.r1c_plane_cfg_new(cfg, 3840, 2160, 0, 0, 88, 88, px_bytes) == 0
        for world in (1, 2, 4, 8):
            area = 0
            for (lo, hi), (clo, chi) in shard.split_tiles(3840, 2160, world):
                r0, r1 = shard.tile_plane_rows(cfg, lo, hi, 64)
                c0, c1 = shard.tile_plane_cols(cfg, clo, chi, 64)
                assert 0 <= r0 < r1 <= cfg.alloc_height and 0 <= c0 < c1 <= cfg.stride
                assert c0 % 64 == 0 and (c1 % 64 == 0 or c1 == cfg.stride)
                # the tile's pixels +- (64 + 16), clipped to the allocation
                assert r0 <= max(cfg.yorigin + lo * 64 - 80, 0) and r1 >= min(cfg.yorigin + min(hi * 64, 2160) + 80, cfg.alloc_height)
                assert c0 <= max(cfg.xorigin + clo * 64 - 80, 0) and c1 >= min(cfg.xorigin + min(chi * 64, 3840) + 80, cfg.stride)
                area += (r1 - r0) * (c1 - c0)
            if world == 8:  # 2 x 4 tiles: a rank moves ~ 1/5 of a plane, not all of it
                assert area / 8 < 0.25 * cfg.alloc_height * cfg.stride


@pytest.mark.timeout(300)
def test_two_rank_gloo_gather_equals_whole_frame():
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    from oracle import r1o
    from rav1e_b200 import api
    r1o.build()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    row, col, cost = q.get(timeout=240)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    mask = (1 << api.BLOCK["BLOCK_8X8"]) | (1 << api.BLOCK["BLOCK_16X16"]) | (1 << api.BLOCK["BLOCK_64X64"])
    o_org, o_ref = _planes(r1o)
    mv, want = r1o.frame_me(o_org, o_ref, RNG, RNG, LAM, mask, 0, 0, 0, 8)
    assert np.array_equal(cost, want) and np.array_equal(row, mv[:, 0]) and np.array_equal(col, mv[:, 1])
