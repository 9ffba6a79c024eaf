"""N > 1 path on the GPUs: the CUDA-side twin of tests/test_sharding_gloo.py.  Two ranks, one GPU each, NCCL process
group: every rank runs r1c_frame_me on its own device for its tile (superblock rows x columns) of ONE frame pair
(rav1e_b200.shard.split_tiles, the split the reference makes by tile, src/encoder.rs:2772-2781), the (mv, cost)
shards are gathered to rank 0 through the process group and must equal both the whole-frame search of one GPU and the
oracle.  Skipped on boxes with fewer than two GPUs (the driver's 1-GPU tier); `gpurun --gpus 2` runs it."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, RNG, LAM = 256, 200, 8, 21


def _planes(r1o, pb, bd):
    rng = np.random.default_rng(12)
    o_ref = r1o.OPlane(W, H, 0, 0, 88, 88, pb)
    o_org = r1o.OPlane(W, H, 0, 0, 88, 88, pb)
    o_ref.visible[...] = rng.integers(0, 1 << bd, size=(H, W))
    o_org.visible[...] = np.roll(o_ref.visible, (3, -2), axis=(0, 1))
    o_ref.pad()
    o_org.pad()
    return o_org, o_ref


def _worker(rank, world, port, pb, bd, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import r1o
    import rav1e_b200 as r
    from rav1e_b200 import shard
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ctx = r.Context(rank)
    o_org, o_ref = _planes(r1o, pb, bd)
    org = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(o_org.data)
    ref = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(o_ref.data)
    (lo, hi), (clo, chi) = shard.split_tiles(W, H, world)[rank]
    p = r.make_frame_me_params(W, H, RNG, RNG, LAM, refine=1, use_satd=1, sb_rows=(lo, hi), sb_cols=(clo, chi))
    local = r.frame_me(ctx, org, ref, p, bd)
    slots = shard.shard_slots(W, H, r.ALL_SIZES_TO_64, lo, hi, clo, chi)
    merged = shard.gather_frame_results(local, slots, local.size, dist)
    if rank == 0:
        whole = r.frame_me(ctx, org, ref, r.make_frame_me_params(W, H, RNG, RNG, LAM, refine=1, use_satd=1), bd)
        q.put((merged["mv"]["row"].copy(), merged["mv"]["col"].copy(), merged["cost"].copy(),
               whole["mv"]["row"].copy(), whole["mv"]["col"].copy(), whole["cost"].copy(), ctx.launches))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10)])
def test_two_rank_nccl_tile_shards_equal_whole_frame(pb, bd):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    sys.path.insert(0, ROOT)
    from oracle import r1o
    import rav1e_b200 as r
    r1o.build()
    mpc = mp.get_context("spawn")
    q = mpc.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [mpc.Process(target=_worker, args=(k, 2, port, pb, bd, q)) for k in range(2)]
    for p in procs:
        p.start()
    row, col, cost, wrow, wcol, wcost, launches = q.get(timeout=480)
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert launches > 0
    assert np.array_equal(cost, wcost) and np.array_equal(row, wrow) and np.array_equal(col, wcol)
    o_org, o_ref = _planes(r1o, pb, bd)
    mv, want = r1o.frame_me(o_org, o_ref, RNG, RNG, LAM, r.ALL_SIZES_TO_64, 1, 1, 0, bd)
    assert np.array_equal(cost, want) and np.array_equal(row, mv[:, 0]) and np.array_equal(col, mv[:, 1])
