"""Within-frame sharding of the hot path over the GPUs of one box (north_star: "by AV1 tile"), measured by
bench.py at every N next to the frame-sharded headline:

  * me_strong():  ONE 3840x2160 10-bit frame pair, full-search ME +-64 + SATD sub-pel tail, the frame cut into one tile of
                  whole superblocks per rank (rav1e_b200.shard.split_tiles; the reference splits the same work by tile,
                  src/encoder.rs:2772-2781), both planes resident on every rank, every rank packs the (mv, cost) results of
                  its tile (r1c_frame_me_compact_dev) and copies them to page-locked host memory.  Strong scaling: the
                  frame is fixed, time = max over ranks.
  * mc_tx_sharded(): the 4K 10-bit MC put list (19 block sizes, one mixed list) split by tile column (8 columns of 480 px,
                  src/tiling/tiler.rs:87-109) and the forward-transform blocks split evenly: same rule, no collective.

No collective touches the data path; torch.distributed is used for the barrier and the max-over-ranks of the times."""
import ctypes as C

import numpy as np

import bench_kernels
from rav1e_b200 import shard

W4K, H4K, BD4K, RANGE, LAMBDA = 3840, 2160, 10, 64, 160


def _max_over_ranks(torch, dist, world, v):
    if world == 1:
        return v
    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _barrier(torch, dist, world):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def me_strong(ctx, r, torch, dist, ext, flush, rank, world, steps, warmup=2):
    lib = ctx.lib
    steps0 = steps
    cur, ref = bench_kernels.synth_frame_pair(11, W4K, H4K, BD4K)
    p_org = r.Plane(ctx, W4K, H4K, 0, 0, 88, 88, 2)
    p_ref = r.Plane(ctx, W4K, H4K, 0, 0, 88, 88, 2)
    h_org = torch.from_numpy(bench_kernels.padded(p_org.cfg, cur)).pin_memory()
    h_ref = torch.from_numpy(bench_kernels.padded(p_ref.cfg, ref)).pin_memory()
    p_org.upload(h_org.numpy())
    p_ref.upload(h_ref.numpy())
    tiles = shard.split_tiles(W4K, H4K, world)
    (lo, hi), (clo, chi) = tiles[rank]
    # the rank's tile goes to two contexts (two streams) as two sub-tiles, so that the tail of one sub-tile's kernels is
    # filled by the other's (a rank has 5 waves of work items or fewer per kernel at N = 8)
    if chi - clo >= 2:
        cm = (clo + chi + 1) // 2
        subs = [((lo, hi), (clo, cm)), ((lo, hi), (cm, chi))]
    elif hi - lo >= 2:
        rm = (lo + hi + 1) // 2
        subs = [((lo, rm), (clo, chi)), ((rm, hi), (clo, chi))]
    else:
        subs = [((lo, hi), (clo, chi))]
    ctxs = [ctx] + [r.Context(torch.cuda.current_device()) for _ in subs[1:]]
    exts = [ext] + [torch.cuda.ExternalStream(c.stream, device=torch.cuda.current_device()) for c in ctxs[1:]]
    n_res = r.frame_me_count(r.make_frame_me_params(W4K, H4K, RANGE, RANGE, LAMBDA))
    d_res = torch.zeros(n_res * 16, dtype=torch.uint8, device="cuda")  # frame layout, the sub-tiles write disjoint slots
    work = []
    for c, e, ((a0, a1), (b0, b1)) in zip(ctxs, exts, subs):
        p = r.make_frame_me_params(W4K, H4K, RANGE, RANGE, LAMBDA, refine=1, use_satd=1, sb_rows=(a0, a1), sb_cols=(b0, b1))
        n = lib.r1c_frame_me_shard_count(c.h, C.byref(p))
        assert n == shard.shard_slots(W4K, H4K, r.ALL_SIZES_TO_64, a0, a1, b0, b1).size
        work.append((c, e, p, n, torch.zeros(max(n, 1) * 16, dtype=torch.uint8, device="cuda"),
                     torch.zeros(max(n, 1) * 16, dtype=torch.uint8).pin_memory()))
    n_mine = sum(w[3] for w in work)
    assert n_mine == shard.shard_slots(W4K, H4K, r.ALL_SIZES_TO_64, lo, hi, clo, chi).size
    up_done = torch.cuda.Event()

    def step_dev():
        for c, e, p, n, d_cmp, h_cmp in work:
            if n:
                c.check(lib.r1c_frame_me_dev(c.h, p_org.h, p_ref.h, C.byref(p), BD4K, d_res.data_ptr()))
                c.check(lib.r1c_frame_me_compact_dev(c.h, C.byref(p), d_res.data_ptr(), d_cmp.data_ptr()))
                with torch.cuda.stream(e):
                    h_cmp.copy_(d_cmp, non_blocking=True)

    # the rectangle of the planes this rank's tile can touch: one strided copy per plane
    up_rows = shard.tile_plane_rows(p_ref.cfg, lo, hi, RANGE)
    up_cols = shard.tile_plane_cols(p_ref.cfg, clo, chi, RANGE)

    def step_e2e():
        ctx.check(lib.r1c_plane_upload_rect_async(ctx.h, p_org.h, h_org.data_ptr(), up_rows[0], up_rows[1], up_cols[0], up_cols[1]))
        ctx.check(lib.r1c_plane_upload_rect_async(ctx.h, p_ref.h, h_ref.data_ptr(), up_rows[0], up_rows[1], up_cols[0], up_cols[1]))
        up_done.record(ext)
        for e in exts[1:]:
            e.wait_event(up_done)
        step_dev()

    def timed(fn):
        evs = [(torch.cuda.Event(enable_timing=True), [torch.cuda.Event(enable_timing=True) for _ in exts]) for _ in range(steps)]
        _barrier(torch, dist, world)
        for a, ends in evs:
            with torch.cuda.stream(ext):
                bench_kernels.flush_l2(torch, flush)
                a.record(ext)
            for e in exts[1:]:
                e.wait_event(a)
            fn()
            for e, b in zip(exts, ends):
                b.record(e)
            for b in ends[1:]:
                ext.wait_event(b)  # the next step's flush (and upload) starts when every stream is done
        _barrier(torch, dist, world)
        return _max_over_ranks(torch, dist, world, sum(max(a.elapsed_time(b) for b in ends) for a, ends in evs)) / steps

    for _ in range(warmup):
        step_dev()
    # the timed region is at least 2 s of device time on every N: a 3-step estimate (max over ranks) sets the count
    steps = 3
    est = timed(step_dev)
    steps = max(steps0, int(np.ceil(2300.0 / max(est, 1e-3))))
    ms = timed(step_dev)
    steps_dev = steps
    steps = 3
    est = timed(step_e2e)
    steps = max(steps0, int(np.ceil(2300.0 / max(est, 1e-3))))
    ms_e2e = timed(step_e2e)
    mpix = W4K * H4K / 1e6
    n_work = len(work)
    work.clear()  # the page-locked blocks were used on the extra streams: they go before those streams do
    torch.cuda.synchronize()
    for c in ctxs[1:]:
        c.close()
    return {
        "workload": "ONE 3840x2160 10-bit frame pair, full search +-64, 19 sizes, SATD sub-pel tail; the frame cut into N tiles "
                    "of whole superblocks (rav1e_b200.shard.split_tiles), one per rank, planes replicated; a rank runs its tile "
                    "as two sub-tiles on two contexts (streams), packs the (mv, cost) results of each and copies them to "
                    "page-locked host memory",
        "scaling": "strong", "n_gpus": world, "value": mpix / (ms / 1e3), "unit": "Mpix/s", "ms_per_frame": ms,
        "steps": steps_dev, "timed_region_s": steps_dev * ms / 1e3,
        "tiles_sb_rows_x_cols": ["%dx%d" % (b - a, d - c) for (a, b), (c, d) in tiles], "superblocks": [(H4K + 63) // 64, (W4K + 63) // 64],
        "d2h_bytes_this_rank": int(n_mine * 16), "d2h_copies_this_rank": n_work, "contexts_per_rank": n_work,
        "e2e": {"value": mpix / (ms_e2e / 1e3), "unit": "Mpix/s", "ms_per_frame": ms_e2e,
                "steps": steps, "timed_region_s": steps * ms_e2e / 1e3,
                "h2d_bytes_per_rank": int(2 * (up_rows[1] - up_rows[0]) * (up_cols[1] - up_cols[0]) * 2),
                "note": "every rank uploads the rectangle of both planes its tile can touch (the tile +- (range + 16) in both "
                        "directions, r1c_plane_upload_rect_async: one strided copy per plane)"},
        "note": "largest tile = %d of %d superblocks: the best possible speed-up is %.2f at N = %d"
                % (max((b - a) * (d - c) for (a, b), (c, d) in tiles), 34 * 60,
                   34 * 60 / max((b - a) * (d - c) for (a, b), (c, d) in tiles), world)}


def mc_tx_sharded(ctx, r, torch, dist, ext, flush, rank, world, reps=3):
    lib = ctx.lib
    rng = np.random.default_rng(3)
    src = r.Plane(ctx, W4K, H4K, 0, 0, 88, 88, 2)
    dst = r.Plane(ctx, W4K, H4K, 0, 0, 88, 88, 2)
    src.upload(rng.integers(0, 1024, size=src.shape, dtype=np.uint16))
    allj = np.concatenate(bench_kernels.mc_jobs(r, rng, lib, W4K, H4K, 1))
    tile_w = W4K // 8  # 8 tile columns (60 superblock columns -> 8, 8, ... : src/tiling/tiler.rs:87-109; 480 px here)
    col = np.minimum(allj["dst_x"] // tile_w, 7)
    mine = allj[(col * world) // 8 == rank]
    px_all = int((allj["w"].astype(np.int64) * allj["h"]).sum())
    px_mine = int((mine["w"].astype(np.int64) * mine["h"]).sum())
    d_jobs = torch.from_numpy(mine.view(np.uint8).copy()).cuda()

    def f_put():
        ctx.check(lib.r1c_put_8tap_batch_dev(ctx.h, dst.h, src.h, d_jobs.data_ptr(), mine.size, BD4K))

    def timed(fn):
        fn()
        _barrier(torch, dist, world)
        ms = bench_kernels._timed(torch, ext, flush, fn, reps)
        _barrier(torch, dist, world)
        return _max_over_ranks(torch, dist, world, ms)

    mc_ms = timed(f_put)
    n_samples = W4K * H4K * 3 // 2
    res = torch.from_numpy(rng.integers(-1023, 1024, size=n_samples, dtype=np.int16)).cuda()
    coef = torch.empty(n_samples, dtype=torch.int32, device="cuda")
    tx_ms, coeffs = 0.0, 0
    for ts in range(19):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        n = n_samples // (w * h)
        lo, hi = n * rank // world, n * (rank + 1) // world

        def f_tx():
            ctx.check(lib.r1c_fwd_txfm_batch_dev(ctx.h, ts, 0, BD4K, res.data_ptr() + lo * w * h * 2,
                                                 coef.data_ptr() + lo * w * h * 4, hi - lo))
        tx_ms += timed(f_tx)
        coeffs += n * w * h
    return {
        "mc_put_4k10_tile_columns": {
            "workload": "4K 10-bit luma, the 19 tilings as one mixed job list (%.0f Mpx), jobs split by tile column "
                        "(8 columns of 480 px), source plane on every rank; whole r1c_put_8tap_batch_dev call" % (px_all / 1e6),
            "scaling": "strong", "n_gpus": world, "value": px_all / 1e6 / (mc_ms / 1e3), "unit": "Mpix/s", "ms": mc_ms,
            "pixels_this_rank": px_mine},
        "fwd_tx_4k10_split": {
            "workload": "4K 10-bit 4:2:0 residual (12.4 M samples) tiled by each of the 19 TxSize, DCT_DCT, blocks split "
                        "evenly over the ranks; 19 r1c_fwd_txfm_batch_dev calls",
            "scaling": "strong", "n_gpus": world, "value": coeffs / 1e6 / (tx_ms / 1e3), "unit": "Mcoeff/s", "ms": tx_ms}}
