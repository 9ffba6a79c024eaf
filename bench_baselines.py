"""The measurement standard of the headline for the other kernel classes the metric names (8-tap MC, forward TX, SAD /
SATD): next to each `kernels.*` line of bench_kernels.py

  * `cpu_baseline`: the oracle (oracle/libr1o.so: C restatement of the reference's `_ref` functions, OpenMP over jobs, all
    host threads) on a bounded sample of the same workload -- kind "port"; the reference itself (Rust + NASM) cannot be
    built on this box.  This module and bench.py's CPU legs are the only bench code that touches oracle/;
  * `e2e`: the same work through the host-buffer C-ABI entry points (r1c_*_batch: host job / input / output arrays,
    copies inside the timed region), bytes declared;

and BASELINE.json configs[0]: the benches/dist.rs recipe (two 640x480 planes with padding 136, uniform random bytes, one
get_sad / get_satd call at (0, 0) per iteration, 22 block sizes x {8, 10}-bit) on the CPU oracle (one thread, ns per call)
and through r1c_sad / r1c_satd (per call) and r1c_sad_batch / r1c_satd_batch (per block of a 4096-job batch)."""
import ctypes as C
import os
import time

import numpy as np

import bench_kernels

W, H, BD = 3840, 2160, 10


def _cores(r1o):
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    r1o.lib().r1o_set_num_threads(avail)
    return r1o.lib().r1o_num_threads()


def _time_host(fn, reps=3):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e3  # ms


def config0(ctx, r, r1o):
    lib, L = ctx.lib, r1o.lib()
    rng = np.random.default_rng(0)
    out = {"recipe": "benches/dist.rs:70-137: Plane::new(640, 480, 0, 0, 136, 136) x 2, every pixel of the allocation a uniform "
                     "random byte (10-bit case: u16 storage, values 0..255), one call at region (0, 0) per iteration",
           "cpu": {"unit": "ns per call", "threads": 1, "kind": "port (oracle; SAD u8 = psadbw, the rest scalar C)"},
           "gpu_per_call": {"unit": "us per r1c_sad / r1c_satd call (one block per launch, result returned to the host)"},
           "gpu_batched": {"unit": "ns per block of one r1c_*_batch call with 4096 jobs (host arrays, copies inside)"}}
    for bd, pb in ((8, 1), (10, 2)):
        a = r1o.OPlane(640, 480, 0, 0, 136, 136, pb)
        b = r1o.OPlane(640, 480, 0, 0, 136, 136, pb)
        a.data[...] = rng.integers(0, 256, a.data.shape)
        b.data[...] = rng.integers(0, 256, b.data.shape)
        da = r.Plane(ctx, 640, 480, 0, 0, 136, 136, pb).upload(a.data)
        db = r.Plane(ctx, 640, 480, 0, 0, 136, 136, pb).upload(b.data)
        for satd, name in ((0, "get_sad"), (1, "get_satd")):
            cpu, per_call, batched = {}, {}, {}
            for bs in range(22):
                w, h = L.r1o_block_width(bs), L.r1o_block_height(bs)
                key = "%dx%d" % (w, h)
                iters = max(200, int(2e6 // (w * h * (8 if satd else 1))))
                cpu[key] = round(L.r1o_bench_dist(satd, a.origin_ptr(), a.cfg.stride, b.origin_ptr(), b.cfg.stride, w, h, pb, iters), 1)
                fn = r.get_satd if satd else r.get_sad
                want = (r1o.satd if satd else r1o.sad)(a, 0, 0, b, 0, 0, w, h)
                assert fn(ctx, da, (0, 0), db, (0, 0), bs, bd) == want, (name, key, bd)
                per_call[key] = round(_time_host(lambda: fn(ctx, da, (0, 0), db, (0, 0), bs, bd), 20) * 1e3, 1)
                jobs = np.zeros(4096, dtype=r.DIST_JOB)
                jobs["bsize"] = bs
                batched[key] = round(_time_host(lambda: r.sad_batch(ctx, da, db, jobs, bd, satd=bool(satd)), 3) * 1e6 / 4096, 1)
            out["cpu"]["%s_%dbit" % (name, bd)] = cpu
            out["gpu_per_call"]["%s_%dbit" % (name, bd)] = per_call
            out["gpu_batched"]["%s_%dbit" % (name, bd)] = batched
    out["note"] = ("a single block per call is the shape of the reference's per-call API (kept for a check_gpu mode); the GPU "
                   "path is the batched one: every per-call number is launch + copy latency, not kernel work")
    return out


def _sample(rng, n, k):
    return np.sort(rng.choice(n, size=min(n, k), replace=False))


def run(ctx, r, torch, ext, flush, kernels):
    """adds cpu_baseline / e2e to kernels['mc_4k10_*'], ['tx_4k10_fwd_dct_dct'], ['dist_4k10_*'] and the config0 line"""
    from oracle import r1o
    r1o.build()
    lib, L = ctx.lib, r1o.lib()
    cores = _cores(r1o)
    rng = np.random.default_rng(33)
    note = "oracle (C restatement of the reference's _ref functions), OpenMP over jobs"
    # ---------------------------------------------------------------------------------------------------- MC ----
    o_src = r1o.OPlane(W, H, 0, 0, 88, 88, 2)
    o_src.data[...] = rng.integers(0, 1024, o_src.data.shape)
    o_dst = r1o.OPlane(W, H, 0, 0, 88, 88, 2)
    src = r.Plane(ctx, W, H, 0, 0, 88, 88, 2).upload(o_src.data)
    dst = r.Plane(ctx, W, H, 0, 0, 88, 88, 2)
    allj = np.concatenate(bench_kernels.mc_jobs(r, rng, lib, W, H, 1))
    allj = torch.from_numpy(allj.view(np.uint8)).pin_memory().numpy().view(r.MC_JOB)  # page-locked host job list
    px_all = int((allj["w"].astype(np.int64) * allj["h"]).sum())
    sel = allj[_sample(rng, allj.size, 60000)]
    j10 = np.stack([sel[k].astype(np.int32) for k in ("dst_x", "dst_y", "src_x", "src_y", "w", "h", "col_frac", "row_frac",
                                                      "mode_x", "mode_y")], axis=1).copy()
    toff = np.concatenate([[0], np.cumsum(sel["w"].astype(np.int64) * sel["h"])[:-1]]).astype(np.int64)
    px_s = int((sel["w"].astype(np.int64) * sel["h"]).sum())
    tmp = np.zeros(px_s, dtype=np.int16)
    for kind, name in ((0, "put_8tap"), (1, "prep_8tap")):
        ms = _time_host(lambda: L.r1o_mc_batch(kind, o_dst.origin_ptr(), C.byref(o_dst.cfg), tmp.ctypes.data, toff.ctypes.data,
                                               o_src.origin_ptr(), C.byref(o_src.cfg), j10.ctypes.data, sel.size, BD, 2), 1)
        kernels["mc_4k10_" + name]["cpu_baseline"] = {
            "value": px_s / 1e6 / (ms / 1e3), "unit": "Mpix/s", "cores": cores, "kind": "port", "note": note,
            "sample": "%d jobs drawn from the 19 tilings (%.1f Mpx)" % (sel.size, px_s / 1e6)}
    # e2e put: source plane up, host job list in, predicted plane back
    h_src = torch.from_numpy(o_src.data).pin_memory()
    h_dst = torch.empty(dst.shape, dtype=torch.int16).pin_memory()

    def e2e_put():
        ctx.check(lib.r1c_plane_upload(ctx.h, src.h, h_src.data_ptr()))
        r.put_8tap_batch(ctx, dst, src, allj, BD)
        ctx.check(lib.r1c_plane_download(ctx.h, dst.h, h_dst.data_ptr()))
    ms = _time_host(e2e_put, 2)
    kernels["mc_4k10_put_8tap"]["e2e"] = {
        "value": px_all / 1e6 / (ms / 1e3), "unit": "Mpix/s", "ms": ms, "h2d_bytes": int(h_src.numel() * 2 + allj.nbytes),
        "d2h_bytes": int(h_dst.numel() * 2),
        "path": "r1c_plane_upload + r1c_put_8tap_batch (page-locked host job list: the 19 tilings, %.0f Mpx) + r1c_plane_download" % (px_all / 1e6)}
    t16 = [j for j in bench_kernels.mc_jobs(r, rng, lib, W, H, 1) if int(j["w"][0]) == 16 and int(j["h"][0]) == 16][0]
    t16["tmp_off"] = np.arange(t16.size, dtype=np.int64) * 256
    htmp = torch.zeros(t16.size * 256, dtype=torch.int16).pin_memory().numpy()  # page-locked: the DMA target itself
    ms = _time_host(lambda: r.prep_8tap_batch(ctx, htmp, src, t16, BD), 2)
    kernels["mc_4k10_prep_8tap"]["e2e"] = {
        "value": htmp.size / 1e6 / (ms / 1e3), "unit": "Mpix/s", "ms": ms, "h2d_bytes": int(t16.nbytes + htmp.nbytes),
        "d2h_bytes": int(htmp.nbytes), "path": "r1c_prep_8tap_batch (host job list: the 16x16 tiling, page-locked host i16 output; plane resident)"}
    # ---------------------------------------------------------------------------------------------------- TX ----
    n_samples = W * H * 3 // 2
    res = rng.integers(-1023, 1024, size=n_samples, dtype=np.int16)
    ms_cpu = ms_e2e = 0.0
    coeffs = cs = 0
    h_res = torch.from_numpy(res).pin_memory()  # page-locked host buffers: DMA source / target without a staging copy
    h_co = torch.empty(n_samples, dtype=torch.int32).pin_memory()
    for ts in range(19):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        n = n_samples // (w * h)
        ns = max(1, min(n, 400000 // (w * h)))
        blk = np.ascontiguousarray(res[:ns * w * h])
        co = np.empty(ns * w * h, dtype=np.int32)
        ms_cpu += _time_host(lambda: L.r1o_fht_batch(blk.ctypes.data, co.ctypes.data, ts, 0, BD, ns), 1)
        cs += ns * w * h
        ms_e2e += _time_host(lambda: ctx.check(lib.r1c_fwd_txfm_batch(ctx.h, ts, 0, BD, h_res.data_ptr(), h_co.data_ptr(), n)), 2)
        coeffs += n * w * h
    k = kernels["tx_4k10_fwd_dct_dct"]
    k["cpu_baseline"] = {"value": cs / 1e6 / (ms_cpu / 1e3), "unit": "Mcoeff/s", "cores": cores, "kind": "port", "note": note,
                         "sample": "0.4 M samples of each of the 19 TxSize (%.1f M coefficients)" % (cs / 1e6)}
    k["e2e"] = {"value": coeffs / 1e6 / (ms_e2e / 1e3), "unit": "Mcoeff/s", "ms": ms_e2e, "h2d_bytes": int(coeffs * 2),
                "d2h_bytes": int(coeffs * 4), "path": "r1c_fwd_txfm_batch per TxSize (page-locked host i16 residual in, page-locked host i32 coefficients out)"}
    # -------------------------------------------------------------------------------------------- SAD / SATD ----
    o_org = r1o.OPlane(W, H, 0, 0, 88, 88, 2)
    o_org.data[...] = rng.integers(0, 1024, o_org.data.shape)
    org = r.Plane(ctx, W, H, 0, 0, 88, 88, 2).upload(o_org.data)
    dj = np.zeros(allj.size, dtype=r.DIST_JOB)
    dj["org_x"], dj["org_y"], dj["ref_x"], dj["ref_y"] = allj["dst_x"], allj["dst_y"], allj["src_x"], allj["src_y"]
    bs_of = {(lib.r1c_block_width(b), lib.r1c_block_height(b)): b for b in bench_kernels.SIZES_TO_64}
    dj["bsize"] = [bs_of[(int(w), int(h))] for w, h in zip(allj["w"], allj["h"])]
    sj = dj[_sample(rng, dj.size, 200000)]
    j5 = np.stack([sj[k].astype(np.int32) for k in ("org_x", "org_y", "ref_x", "ref_y", "bsize")], axis=1).copy()
    px_s = int(sum(lib.r1c_block_width(int(b)) * lib.r1c_block_height(int(b)) for b in sj["bsize"]))
    outv = np.zeros(sj.size, dtype=np.uint32)
    h_org = torch.from_numpy(o_org.data).pin_memory()
    dj_pin = torch.from_numpy(dj.view(np.uint8)).pin_memory().numpy().view(r.DIST_JOB)  # page-locked job list and result array
    h_dout = torch.empty(dj.size, dtype=torch.int32).pin_memory()
    for satd, name in ((0, "sad"), (1, "satd")):
        ms = _time_host(lambda: L.r1o_dist_batch(satd, o_org.origin_ptr(), C.byref(o_org.cfg), o_src.origin_ptr(), C.byref(o_src.cfg),
                                                 j5.ctypes.data, sj.size, 2, outv.ctypes.data), 1)
        kernels["dist_4k10_" + name]["cpu_baseline"] = {
            "value": px_s / 1e6 / (ms / 1e3), "unit": "Mpix/s", "cores": cores, "kind": "port", "note": note,
            "sample": "%d jobs drawn from the 19 tilings (%.1f Mpx)" % (sj.size, px_s / 1e6)}

        def e2e_dist():
            ctx.check(lib.r1c_plane_upload(ctx.h, org.h, h_org.data_ptr()))
            ctx.check(lib.r1c_plane_upload(ctx.h, src.h, h_src.data_ptr()))
            fn = lib.r1c_satd_batch if satd else lib.r1c_sad_batch
            ctx.check(fn(ctx.h, org.h, src.h, dj_pin.ctypes.data, dj_pin.size, BD, h_dout.data_ptr()))
        ms = _time_host(e2e_dist, 2)
        kernels["dist_4k10_" + name]["e2e"] = {
            "value": px_all / 1e6 / (ms / 1e3), "unit": "Mpix/s", "ms": ms,
            "h2d_bytes": int(h_org.numel() * 2 + h_src.numel() * 2 + dj.nbytes), "d2h_bytes": int(dj.size * 4),
            "path": "r1c_plane_upload x2 + r1c_%s_batch (page-locked host job list: the 19 tilings as one list, page-locked host u32 results)" % name}
    kernels["config0_dist_bench"] = config0(ctx, r, r1o)
