"""bench_kernels.py -- the other kernel classes of BASELINE.json's metric, timed on
device-resident data and reported next to the headline ME line (bench.py "kernels"):

  mc_4k10  (configs[2]) 4K 10-bit luma: put_8tap of the frame tiled by every one of
           the 19 block sizes, random 1/16 fracs (all four (col, row) cases), the ten
           filter pairs of the asm set, integer offsets +-32; and the compound path
           prep_8tap x2 + mc_avg.  Mpix/s = output pixels / time; HBM-bound,
           4 B/px (put, prep) and 6 B/px (avg) algorithmic.
  tx_4k10  (configs[3]) 4K 10-bit 4:2:0 residual (12.4 M samples): forward_transform
           DCT_DCT of the frame tiled by every TxSize; 6 B/coeff algorithmic.

Inputs are larger than L2 in aggregate and L2 is flushed between timed launches (write 256 MiB, read 256 MiB).
"""
import ctypes as C
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
SIZES_TO_64 = [b for b in range(22) if b not in (13, 14, 15)]
FILTER_PAIRS = [(0, 0), (0, 1), (0, 2), (1, 0), (1, 1), (1, 2), (2, 0), (2, 1), (2, 2), (3, 3)]


def _hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def _kernel_ms(ctx, slot):
    """summed duration / launches of one kernel class since the last read (r1c_profile hooks)"""
    ms, n = C.c_double(), C.c_int64()
    ctx.check(ctx.lib.r1c_profile_read(ctx.h, slot, C.byref(ms), C.byref(n)))
    return ms.value, n.value


def flush_l2(torch, flush):
    """untimed: overwrite 256 MiB (> the 126 MB L2) and then read the other half of the buffer back, so the timed
    kernel starts with a cold L2 that holds no dirty lines of the flush itself (their write-back would otherwise
    be charged to the kernel as extra HBM traffic)"""
    half = flush.numel() // 2
    flush[:half].zero_()
    flush[half:].view(torch.int64).sum()  # 8-byte reads, no widening copy


def _timed(torch, ext, flush, fn, reps):
    """reps launches of fn, each event-bracketed on the library stream after an untimed L2 flush"""
    evs = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(ext):
            if not os.environ.get("R1C_BENCH_NOFLUSH"):  # probing aid only; the reported numbers are always flushed
                flush_l2(torch, flush)
            a.record(ext)
        fn()
        b.record(ext)
        evs.append((a, b))
    torch.cuda.synchronize()
    return sum(a.elapsed_time(b) for a, b in evs) / reps


def synth_frame_pair(seed, W, H, bit_depth=8):
    """SURVEY 8(d) cfg 1: cur = smooth random field; ref = cur moved by a random integer MV
    per 64x64 superblock (+-48 px) plus +-4 noise, so minima are unique and non-trivial.
    bit_depth > 8: the same field scaled to the deeper range (u16)."""
    rng = np.random.default_rng(seed)
    maxv = (1 << bit_depth) - 1
    sc = 1 << (bit_depth - 8)
    big = rng.integers(0, 256, size=(H + 256, W + 256)).astype(np.float32)
    for _ in range(2):  # separable box blur, radius 2
        big = (np.roll(big, 2, 0) + np.roll(big, 1, 0) + big + np.roll(big, -1, 0) + np.roll(big, -2, 0)) / 5
        big = (np.roll(big, 2, 1) + np.roll(big, 1, 1) + big + np.roll(big, -1, 1) + np.roll(big, -2, 1)) / 5
    big = ((big - big.mean()) * 3.0 + 128.0) * sc
    dt = np.uint8 if bit_depth == 8 else np.uint16
    cur = np.clip(np.rint(big[128:128 + H, 128:128 + W]), 0, maxv).astype(dt)
    ref = np.empty_like(cur)
    for sy in range(0, H, 64):
        for sx in range(0, W, 64):
            dy, dx = (int(v) for v in rng.integers(-48, 49, size=2))
            h, w = min(64, H - sy), min(64, W - sx)
            ref[sy:sy + h, sx:sx + w] = np.clip(np.rint(big[128 + sy + dy:128 + sy + dy + h,
                                                            128 + sx + dx:128 + sx + dx + w]), 0, maxv)
    noise = rng.integers(-4 * sc, 4 * sc + 1, size=ref.shape)
    ref = np.clip(ref.astype(np.int32) + noise, 0, maxv).astype(dt)
    return cur, ref


def padded(cfg, vis):
    """whole allocation with Plane::pad semantics (edge replication) -- numpy, input prep only"""
    a = np.empty((cfg.alloc_height, cfg.stride), dtype=vis.dtype)
    yo, xo, h, w = cfg.yorigin, cfg.xorigin, cfg.height, cfg.width
    a[yo:yo + h, xo:xo + w] = vis
    a[yo:yo + h, :xo] = vis[:, :1]
    a[yo:yo + h, xo + w:] = vis[:, -1:]
    a[:yo, :] = a[yo:yo + 1, :]
    a[yo + h:, :] = a[yo + h - 1:yo + h, :]
    return a


MC_CANDS = int(os.environ.get("R1C_BENCH_MC_CANDS", "4"))  # candidate MVs per block and launch


def mc_jobs(r, rng, lib, W, H, K=1):
    """per block size: the tiling of the frame by that size, K times over (K candidate MVs per block, as a motion
    search or a compound prediction hands them over in one launch); candidate c predicts into rows [c*H, (c+1)*H) of
    the destination"""
    jobs = []
    only = os.environ.get("R1C_BENCH_SIZES")  # profiling aid: restrict to some BlockSize values
    for b in ([int(v) for v in only.split(",")] if only else SIZES_TO_64):
        bw, bh = lib.r1c_block_width(b), lib.r1c_block_height(b)
        xs, ys = np.meshgrid(np.arange(0, W - bw + 1, bw), np.arange(0, H - bh + 1, bh))
        n = xs.size * K
        j = np.zeros(n, dtype=r.MC_JOB)
        j["dst_x"], j["dst_y"] = np.tile(xs.ravel(), K), np.tile(ys.ravel(), K)
        j["src_x"] = j["dst_x"] + rng.integers(-32, 33, size=n)
        j["src_y"] = j["dst_y"] + rng.integers(-32, 33, size=n)
        j["dst_y"] += np.repeat(np.arange(K), xs.size) * H
        j["w"], j["h"] = bw, bh
        j["col_frac"] = rng.integers(0, 16, size=n)
        j["row_frac"] = rng.integers(0, 16, size=n)
        fp = np.array(FILTER_PAIRS)[rng.integers(0, len(FILTER_PAIRS), size=n)]
        j["mode_x"], j["mode_y"] = fp[:, 0], fp[:, 1]
        jobs.append(j)
    return jobs


def run(ctx, r, torch, ext, flush, reps=3, only=None):
    lib = ctx.lib
    hbm, hbm_how = _hbm_peak()
    out = {}
    rng = np.random.default_rng(3)
    # ------------------------------------------------------------------ MC, 4K 10-bit luma
    W, H, BD = 3840, 2160, 10
    K = MC_CANDS
    src = r.Plane(ctx, W, H, 0, 0, 88, 88, 2)
    dst = r.Plane(ctx, W, K * H, 0, 0, 88, 88, 2)
    host = rng.integers(0, 1024, size=src.shape, dtype=np.uint16)
    src.upload(host)
    per_size = mc_jobs(r, rng, lib, W, H, K)
    put_ms = prep_ms = avg_ms = 0.0
    kput = kprep = kavg = 0.0
    px_total = 0
    by_size = {"put_8tap": {}, "prep_8tap": {}, "mc_avg": {}}
    ctx.check(lib.r1c_profile(ctx.h, 1))
    tmp1 = torch.empty(K * W * H, dtype=torch.int16, device="cuda")
    tmp2 = torch.empty(K * W * H, dtype=torch.int16, device="cuda")
    for j in per_size:
        n = j.size
        area = int(j["w"][0]) * int(j["h"][0])
        j["tmp_off"] = np.arange(n, dtype=np.int64) * area
        d_jobs = torch.from_numpy(j.view(np.uint8).copy()).cuda()
        aj = np.zeros(n, dtype=r.AVG_JOB)
        aj["dst_x"], aj["dst_y"], aj["w"], aj["h"], aj["tmp_off"] = j["dst_x"], j["dst_y"], j["w"], j["h"], j["tmp_off"]
        d_aj = torch.from_numpy(aj.view(np.uint8).copy()).cuda()
        px_total += n * area

        def f_put():
            ctx.check(lib.r1c_put_8tap_batch_dev(ctx.h, dst.h, src.h, d_jobs.data_ptr(), n, BD))

        def f_prep():
            ctx.check(lib.r1c_prep_8tap_batch_dev(ctx.h, tmp1.data_ptr(), src.h, d_jobs.data_ptr(), n, BD))

        def f_avg():
            ctx.check(lib.r1c_mc_avg_batch_dev(ctx.h, dst.h, tmp1.data_ptr(), tmp2.data_ptr(), d_aj.data_ptr(), n, BD))
        f_put()
        ctx.check(lib.r1c_prep_8tap_batch_dev(ctx.h, tmp2.data_ptr(), src.h, d_jobs.data_ptr(), n, BD))
        f_prep()
        f_avg()
        ctx.sync()
        _kernel_ms(ctx, 3)
        key = "%dx%d" % (int(j["w"][0]), int(j["h"][0]))
        put_ms += _timed(torch, ext, flush, f_put, reps)
        k = _kernel_ms(ctx, 3)[0] / reps
        kput += k
        by_size["put_8tap"][key] = round(k * 1e3, 1)
        prep_ms += _timed(torch, ext, flush, f_prep, reps)
        k = _kernel_ms(ctx, 3)[0] / reps
        kprep += k
        by_size["prep_8tap"][key] = round(k * 1e3, 1)
        avg_ms += _timed(torch, ext, flush, f_avg, reps)
        k = _kernel_ms(ctx, 3)[0] / reps
        kavg += k
        by_size["mc_avg"][key] = round(k * 1e3, 1)
    # the 19 tilings as ONE job list / one launch (157 Mpx, mixed block sizes: prefix-sum job lookup): the rate of
    # the kernel once the launch is long enough to hide its start-up chain (job -> rows) and its tail
    allj = np.concatenate(per_size)
    off = 0
    for j in per_size:
        a = int(j["w"][0]) * int(j["h"][0])
        allj["tmp_off"][off:off + j.size] = 0  # put only
        off += j.size
    d_all = torch.from_numpy(allj.view(np.uint8).copy()).cuda()

    def f_put_all():
        ctx.check(lib.r1c_put_8tap_batch_dev(ctx.h, dst.h, src.h, d_all.data_ptr(), allj.size, BD))
    f_put_all()
    ctx.sync()
    _kernel_ms(ctx, 3)
    one_ms = _timed(torch, ext, flush, f_put_all, reps)
    one_k = _kernel_ms(ctx, 3)[0] / reps
    try:
        ipeak = float(json.load(open(os.path.join(ROOT, "profiles", "int_peaks_r01.json")))["iadd"])
    except Exception:
        ipeak = 36000.0
    for name, call_ms, ms, bpp in (("put_8tap", put_ms, kput, 4), ("prep_8tap", prep_ms, kprep, 4),
                                   ("mc_avg", avg_ms, kavg, 6)):
        gbs = px_total * bpp / (ms / 1e3) / 1e9
        out["mc_4k10_" + name] = {
            "workload": "4K 10-bit luma tiled by each of the 19 block sizes, %d candidate MVs per block and launch (%.0f Mpx per launch), random fracs / filter pairs / +-32 px" % (K, K * W * H / 1e6),
            "value": px_total / 1e6 / (call_ms / 1e3), "unit": "Mpix/s", "ms": call_ms, "pixels": px_total,
            "kernel_ms": ms, "kernel_us_by_block_size": by_size[name], "note": ("value / ms = whole _dev call (one launch, no scratch); " if name == "mc_avg" else
                                      "value / ms = whole _dev call (unit count + scatter + kernel); ") +
                                     "roofline = the main kernel alone (19 launches, CUDA events inside the library)",
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                         "alg_bytes_per_px": bpp, "peak_source": hbm_how}}
        if name != "mc_avg":  # 16 MACs = 32 integer ops per output pixel (2-D case, halo not counted)
            out["mc_4k10_" + name]["roofline"]["int_frac"] = px_total * 32 / (ms / 1e3) / 1e9 / ipeak
    gbs = px_total * 4 / (one_k / 1e3) / 1e9
    out["mc_4k10_put_8tap_one_launch"] = {
        "workload": "the same 19 tilings as one mixed-size job list: one launch of %.0f Mpx" % (px_total / 1e6),
        "value": px_total / 1e6 / (one_ms / 1e3), "unit": "Mpix/s", "ms": one_ms, "pixels": px_total, "kernel_ms": one_k,
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm, "alg_bytes_per_px": 4,
                     "peak_source": hbm_how, "int_frac": px_total * 32 / (one_k / 1e3) / 1e9 / ipeak}}
    del tmp1, tmp2
    if only == "mc":
        ctx.check(lib.r1c_profile(ctx.h, 0))
        return out
    # ------------------------------------------ SAD / SATD, 4K 10-bit luma, one job per block
    org = r.Plane(ctx, W, H, 0, 0, 88, 88, 2)
    org.upload(rng.integers(0, 1024, size=src.shape, dtype=np.uint16))
    d_out = torch.empty(W * H // 16, dtype=torch.int32, device="cuda")
    sad_ms = satd_ms = 0.0
    dist_px = 0
    for j in per_size:
        n = j.size
        dj = np.zeros(n, dtype=r.DIST_JOB)
        dj["org_x"], dj["org_y"], dj["ref_x"], dj["ref_y"] = j["dst_x"], j["dst_y"], j["src_x"], j["src_y"]
        dj["bsize"] = [b for b in SIZES_TO_64 if lib.r1c_block_width(b) == int(j["w"][0])
                       and lib.r1c_block_height(b) == int(j["h"][0])][0]
        d_dj = torch.from_numpy(dj.view(np.uint8).copy()).cuda()
        dist_px += n * int(j["w"][0]) * int(j["h"][0])
        for satd in (0, 1):
            def f_dist():
                ctx.check(lib.r1c_dist_batch_dev(ctx.h, satd, org.h, src.h, d_dj.data_ptr(), n, BD, d_out.data_ptr()))
            f_dist()
            ctx.sync()
            _kernel_ms(ctx, 5)
            _timed(torch, ext, flush, f_dist, reps)
            ms = _kernel_ms(ctx, 5)[0] / reps
            if satd:
                satd_ms += ms
            else:
                sad_ms += ms
    for name, ms in (("sad", sad_ms), ("satd", satd_ms)):
        gbs = dist_px * 4 / (ms / 1e3) / 1e9
        out["dist_4k10_" + name] = {
            "workload": "4K 10-bit luma tiled by each of the 19 block sizes, %d candidate reference positions per block and launch (+-32 px; the org block is shared by its candidates, so DRAM reads less than the 4 B per pixel counted)" % MC_CANDS,
            "value": dist_px / 1e6 / (ms / 1e3), "unit": "Mpix/s", "ms": ms, "pixels": dist_px,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                         "alg_bytes_per_px": 4, "peak_source": hbm_how,
                         "note": "frac = algorithmic bytes (SURVEY 8d) over the copy peak, not DRAM utilisation: ncu "
                                 "(profiles/ncu_avg_dist_r02.md) counts 45 MB of DRAM reads per launch of 132 MB "
                                 "algorithmic at 8x8 (34 MB at 32x32); L2 serves the candidates' overlap"}}
    # ---------------------------------------------------- forward transforms, 4K 10-bit 4:2:0
    n_samples = 3840 * 2160 * 3 // 2
    res = torch.from_numpy(rng.integers(-1023, 1024, size=n_samples, dtype=np.int16)).cuda()
    coef = torch.empty(n_samples, dtype=torch.int32, device="cuda")
    tx_ms = 0.0
    coeffs = 0
    per = {}
    only = os.environ.get("R1C_BENCH_TXSIZES")
    for ts in ([int(v) for v in only.split(",")] if only else range(19)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        n = n_samples // (w * h)

        def f_tx():
            ctx.check(lib.r1c_fwd_txfm_batch_dev(ctx.h, ts, 0, BD, res.data_ptr(), coef.data_ptr(), n))
        f_tx()
        ctx.sync()
        _kernel_ms(ctx, 4)
        _timed(torch, ext, flush, f_tx, reps)
        ms = _kernel_ms(ctx, 4)[0] / reps
        tx_ms += ms
        coeffs += n * w * h
        per["%dx%d" % (w, h)] = round(n * w * h * 6 / (ms / 1e3) / 1e9 / hbm, 3)
    gbs = coeffs * 6 / (tx_ms / 1e3) / 1e9
    out["tx_4k10_fwd_dct_dct"] = {
        "workload": "4K 10-bit 4:2:0 residual (12.4 M samples) tiled by each of the 19 TxSize, DCT_DCT",
        "value": coeffs / 1e6 / (tx_ms / 1e3), "unit": "Mcoeff/s", "ms": tx_ms, "coeffs": coeffs,
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                     "alg_bytes_per_coeff": 6, "peak_source": hbm_how, "frac_by_tx_size": per}}
    # ------------------------------ quantize (+ reconstruction coefficients + tx-domain distortion), same coefficients
    qsum = {"q": 0.0, "fused": 0.0}
    qper = {"q": {}, "fused": {}}
    qcoeffs = 0
    qout = torch.empty(n_samples, dtype=torch.int32, device="cuda")
    rout = torch.empty(n_samples, dtype=torch.int32, device="cuda")
    dout = torch.empty(n_samples // 16, dtype=torch.int64, device="cuda")
    qin = torch.from_numpy(np.rint(rng.standard_normal(n_samples) * 400).astype(np.int32)).cuda()
    for ts in ([int(v) for v in only.split(",")] if only else range(19)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        n = n_samples // (w * h)
        qc = r.QuantizationContext().update(lib, ts, 0, 120, 147)
        scan = torch.from_numpy(np.concatenate([[0], 1 + rng.permutation(coded - 1)]).astype(np.int16)).cuda()
        for key, rp, dp in (("q", 0, 0), ("fused", rout.data_ptr(), dout.data_ptr())):
            def f_q():
                ctx.check(lib.r1c_quantize_batch_dev(ctx.h, C.byref(qc.c), ts, scan.data_ptr(), qin.data_ptr(),
                                                     qout.data_ptr(), rp, dp, n))
            f_q()
            ctx.sync()
            _kernel_ms(ctx, 6)
            _timed(torch, ext, flush, f_q, reps)
            k_ms = _kernel_ms(ctx, 6)[0] / reps
            qsum[key] += k_ms
            qper[key]["%dx%d" % (w, h)] = round(k_ms * 1e3, 1)
        qcoeffs += n * coded
    # the same pass on coefficients shaped like a real residual: energy decaying with frequency (sigma = 900 exp(-0.45 (r + c)))
    # under the zig-zag scan, so most of a large block lies past its end of block -- the chunks the quantizer skips
    ssum, sper = 0.0, {}
    for ts in ([int(v) for v in only.split(",")] if only else range(19)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        cw, ch = min(w, 32), min(h, 32)
        n = n_samples // (w * h)
        rr, cc = np.indices((h, w))
        sig = 900.0 * np.exp(-0.45 * (rr + cc))
        blk = np.rint(rng.standard_normal((min(n, 4096), h, w)) * sig).astype(np.int32)
        qs = torch.from_numpy(np.tile(blk.reshape(-1), -(-n // blk.shape[0]))[: n * w * h].copy()).cuda()
        order = sorted(range(cw * ch), key=lambda p_: (p_ // cw + p_ % cw, p_ // cw if (p_ // cw + p_ % cw) & 1 else p_ % cw))
        zz = torch.from_numpy(np.array(order, dtype=np.int16)).cuda()
        qc = r.QuantizationContext().update(lib, ts, 0, 120, 147)

        def f_qs():
            ctx.check(lib.r1c_quantize_batch_dev(ctx.h, C.byref(qc.c), ts, zz.data_ptr(), qs.data_ptr(), qout.data_ptr(),
                                                 rout.data_ptr(), dout.data_ptr(), n))
        f_qs()
        ctx.sync()
        _kernel_ms(ctx, 6)
        _timed(torch, ext, flush, f_qs, reps)
        k_ms = _kernel_ms(ctx, 6)[0] / reps
        ssum += k_ms
        sper["%dx%d" % (w, h)] = round(k_ms * 1e3, 1)
        del qs
    gbs = qcoeffs * 12 / (ssum / 1e3) / 1e9
    out["quant_4k10_fused_sparse"] = {
        "workload": "the fused pass on coefficients shaped like a real residual (sigma = 900 exp(-0.45 (row + col)), zig-zag scan, "
                    "qstep 147: most of a large block lies past its end of block)",
        "value": qcoeffs / 1e6 / (ssum / 1e3), "unit": "Mcoeff/s", "ms": ssum, "coeffs": qcoeffs, "kernel_us_by_tx_size": sper,
        "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                     "alg_bytes_per_coeff": 12, "peak_source": hbm_how}}
    for key, bpc, what in (("q", 8, "quantize"), ("fused", 12, "quantize + dequantize + tx-domain distortion in one pass")):
        gbs = qcoeffs * bpc / (qsum[key] / 1e3) / 1e9
        out["quant_4k10_" + key] = {
            "workload": "4K 4:2:0 coefficients (12.4 M, sigma 400, qstep 147) tiled by each of the 19 TxSize: " + what,
            "value": qcoeffs / 1e6 / (qsum[key] / 1e3), "unit": "Mcoeff/s", "ms": qsum[key], "coeffs": qcoeffs,
            "kernel_us_by_tx_size": qper[key],
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                         "alg_bytes_per_coeff": bpc, "peak_source": hbm_how}}
    # ------------------------------ the fused residual chain of encode_tx_block on a 4K 10-bit luma plane: source = `org`,
    # prediction = source + noise of +-24 held in `rec` (restored from page-locked memory before every launch, outside
    # the timed kernel), DCT_DCT, every TxSize; the same work through the four separate kernels
    # is timed next to it (diff + fht, quantize with tx_dist, dequantize + inverse add: three launches)
    rec = r.Plane(ctx, W, H, 0, 0, 88, 88, 2)
    org_host = org.download()
    h_pred = torch.from_numpy(np.clip(org_host.astype(np.int32) + rng.integers(-24, 25, size=org_host.shape), 0, 1023)
                              .astype(np.uint16)).pin_memory()
    del org_host
    # two lines: the 14 sizes up to 32x32 and the five with a 64-point side (enc_tx64.cuh)
    esum = {g: {"fused": 0.0, "chain": 0.0} for g in (0, 1)}
    eper = {0: {}, 1: {}}
    ebytes = {0: 0, 1: 0}
    epix = {0: 0, 1: 0}
    co_tmp = torch.empty(W * H, dtype=torch.int32, device="cuda")
    for ts in ([int(v) for v in only.split(",")] if only else range(19)):
        w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
        if not lib.r1c_encode_tx_block_valid(ts, 0):
            continue
        xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
        n = xs.size
        tj = np.zeros(n, dtype=r.TX_JOB)
        tj["in_x"] = tj["pred_x"] = xs.ravel()
        tj["in_y"] = tj["pred_y"] = ys.ravel()
        ij = np.zeros(n, dtype=r.ITX_JOB)
        ij["x"], ij["y"] = xs.ravel(), ys.ravel()
        d_tj = torch.from_numpy(tj.view(np.int32).reshape(-1)).cuda()
        d_ij = torch.from_numpy(ij.view(np.int32).reshape(-1)).cuda()
        qc = r.QuantizationContext().update(lib, ts, 0, 40, 52)
        coded = min(w, 32) * min(h, 32)
        g = int(max(w, h) == 64)
        scan = torch.from_numpy(np.concatenate([[0], 1 + rng.permutation(coded - 1)]).astype(np.int16)).cuda()

        def f_e():
            ctx.check(lib.r1c_plane_upload_async(ctx.h, rec.h, h_pred.data_ptr()))
            ctx.check(lib.r1c_encode_tx_block_batch_dev(ctx.h, org.h, rec.h, d_tj.data_ptr(), n, ts, 0, BD, C.byref(qc.c),
                                                        scan.data_ptr(), 1, qout.data_ptr(), dout.data_ptr()))

        def f_c():
            ctx.check(lib.r1c_plane_upload_async(ctx.h, rec.h, h_pred.data_ptr()))
            ctx.check(lib.r1c_diff_fwd_txfm_batch_dev(ctx.h, org.h, rec.h, d_tj.data_ptr(), n, ts, 0, BD, co_tmp.data_ptr()))
            ctx.check(lib.r1c_quantize_batch_dev(ctx.h, C.byref(qc.c), ts, scan.data_ptr(), co_tmp.data_ptr(), qout.data_ptr(),
                                                 0, dout.data_ptr(), n))
            ctx.check(lib.r1c_dequant_inv_txfm_add_batch_dev(ctx.h, rec.h, d_ij.data_ptr(), n, ts, 0, BD, qout.data_ptr(), 40, 52))
        k_ms = {}
        for key, fn in (("fused", f_e), ("chain", f_c)):
            fn()
            ctx.sync()
            for slot in (4, 6, 7):
                _kernel_ms(ctx, slot)
            _timed(torch, ext, flush, fn, reps)
            k_ms[key] = sum(_kernel_ms(ctx, slot)[0] for slot in (4, 6, 7)) / reps
            esum[g][key] += k_ms[key]
        b = n * (w * h * 6 + coded * 4 + 8)  # 2 + 2 B read and 2 B written per pixel, 4 B per coded qcoeff, 8 B tx_dist per block
        ebytes[g] += b
        epix[g] += n * w * h
        eper[g]["%dx%d" % (w, h)] = {"frac": round(b / (k_ms["fused"] / 1e3) / 1e9 / hbm, 3), "us": round(k_ms["fused"] * 1e3, 1),
                                  "separate_kernels_us": round(k_ms["chain"] * 1e3, 1)}
    for g, key, what, bytes_txt in ((0, "enc_tx_4k10_fused_dct_dct", "the 14 TxSize up to 32x32", "10 per pixel (2 + 2 read, 4 + 2 written) + 8 per block"),
                                    (1, "enc_tx_4k10_fused_dct_dct_64pt", "the 5 TxSize with a 64-point side (this fork's layouts: "
                                     "min(W,32) * min(H,32) coded coefficients per block)", "6 per pixel + 4 per coded coefficient + 8 per block")):
        if not epix[g]:
            continue
        gbs = ebytes[g] / (esum[g]["fused"] / 1e3) / 1e9
        out[key] = {
            "workload": "4K 10-bit luma tiled by each of " + what + ": diff + forward DCT_DCT + quantize (qstep 52) + "
                        "dequantize + inverse transform added back + tx-domain distortion as ONE kernel per size "
                        "(r1c_encode_tx_block_batch_dev); prediction = source +- 24",
            "value": epix[g] / 1e6 / (esum[g]["fused"] / 1e3), "unit": "Mpix/s", "ms": esum[g]["fused"], "pixels": epix[g],
            "separate_kernels_ms": esum[g]["chain"], "speedup_vs_separate_kernels": esum[g]["chain"] / esum[g]["fused"],
            "by_tx_size": eper[g],
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                         "alg_bytes": bytes_txt, "peak_source": hbm_how}}
    del rec, h_pred, co_tmp
    # ------------------------------ inverse transform + add onto a 4K luma plane (dequantized-like coefficients):
    # 10-bit (u16 plane `dst`) and 8-bit (u8 plane), and the 10-bit case with dequantize fused into the load
    rin = (qin // 147) * 147  # what dequantize leaves: multiples of the step, mostly zero
    qlv = qin // 147          # the quantized levels themselves, for the fused variant
    dst8 = r.Plane(ctx, W, H, 0, 0, 88, 88, 1)
    dst8.upload(rng.integers(0, 256, size=dst8.shape, dtype=np.uint8))
    for key, plane, bd_i, pxb, fused in (("itx_4k10_inv_dct_dct_add", dst, BD, 2, 0), ("itx_4k8_inv_dct_dct_add", dst8, 8, 1, 0),
                                         ("itx_4k10_dequant_inv_dct_dct_add", dst, BD, 2, 1)):
        isum = 0.0
        iper = {}
        ibytes = 0
        ipix = 0
        for ts in ([int(v) for v in only.split(",")] if only else range(19)):
            w, h = lib.r1c_tx_width(ts), lib.r1c_tx_height(ts)
            coded = min(w, 32) * min(h, 32)
            xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
            n = min(xs.size, n_samples // coded)
            ij = np.zeros(n, dtype=r.ITX_JOB)
            ij["x"], ij["y"] = xs.ravel()[:n], ys.ravel()[:n]
            d_ij = torch.from_numpy(ij.view(np.int32).reshape(-1)).cuda()

            def f_i():
                if fused:
                    ctx.check(lib.r1c_dequant_inv_txfm_add_batch_dev(ctx.h, plane.h, d_ij.data_ptr(), n, ts, 0, bd_i,
                                                                     qlv.data_ptr(), 120, 147))
                else:
                    ctx.check(lib.r1c_inv_txfm_add_batch_dev(ctx.h, plane.h, d_ij.data_ptr(), n, ts, 0, bd_i, rin.data_ptr()))
            f_i()
            ctx.sync()
            _kernel_ms(ctx, 7)
            _timed(torch, ext, flush, f_i, reps)
            k_ms = _kernel_ms(ctx, 7)[0] / reps
            isum += k_ms
            b = n * (coded * 4 + w * h * 2 * pxb)  # coded coefficients + destination read and written
            ibytes += b
            ipix += n * w * h
            iper["%dx%d" % (w, h)] = round(b / (k_ms / 1e3) / 1e9 / hbm, 3)
        gbs = ibytes / (isum / 1e3) / 1e9
        out[key] = {
            "workload": "4K %d-bit luma tiled by each of the 19 TxSize, DCT_DCT inverse transform added onto the plane%s"
                        % (bd_i, " (quantized levels in, dequantize fused)" if fused else ""),
            "value": ipix / 1e6 / (isum / 1e3), "unit": "Mpix/s", "ms": isum, "pixels": ipix,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s", "frac": gbs / hbm,
                         "alg_bytes": "4 per coded coefficient + 2 x %d per pixel" % pxb, "peak_source": hbm_how,
                         "frac_by_tx_size": iper}}
    del rin, qlv, dst8
    ctx.check(lib.r1c_profile(ctx.h, 0))
    del res, coef, org, src, dst, d_out, qout, rout, dout, qin
    # -------------------------- whole-frame motion estimation on a 4K 10-bit frame pair (u16 planes)
    if not os.environ.get("R1C_BENCH_SIZES") and not os.environ.get("R1C_BENCH_TXSIZES"):
        out["me_4k10_fullsearch_pm64"] = me_frame_line(ctx, r, torch, ext, flush, 3840, 2160, 10, 64)
        # 12-bit planes: the first-generation superblock search (me_sb.cu) and the generic sub-pel composition
        out["me_1080p12_fullsearch_pm64"] = me_frame_line(ctx, r, torch, ext, flush, 1920, 1080, 12, 64, reps=1)
    return out


def me_frame_line(ctx, r, torch, ext, flush, W, H, BD, RANGE, reps=2):
    """r1c_frame_me_dev on a synthetic frame pair, planes resident: the headline workload at another frame size /
    bit depth (same kernels; u16 planes take the packed-pair variants of me_sb2 / me_refine)"""
    lib = ctx.lib
    PB = 1 if BD == 8 else 2
    cur, ref = synth_frame_pair(11, W, H, BD)
    p_org = r.Plane(ctx, W, H, 0, 0, 88, 88, PB)
    p_ref = r.Plane(ctx, W, H, 0, 0, 88, 88, PB)
    p_org.upload(padded(p_org.cfg, cur))
    p_ref.upload(padded(p_ref.cfg, ref))
    params = r.make_frame_me_params(W, H, RANGE, RANGE, 160 << (BD - 8), refine=1, use_satd=1)
    n_res = r.frame_me_count(params)
    d_res = torch.empty(n_res * 16, dtype=torch.uint8, device="cuda")

    def step():
        ctx.check(lib.r1c_frame_me_dev(ctx.h, p_org.h, p_ref.h, C.byref(params), BD, d_res.data_ptr()))
    step()
    ctx.sync()
    ctx.check(lib.r1c_profile(ctx.h, 1))
    l0 = ctx.launches
    ms = _timed(torch, ext, flush, step, reps)
    launches = (ctx.launches - l0) // reps
    sb_ms = _kernel_ms(ctx, 0)[0] / reps
    mc_ms = _kernel_ms(ctx, 1)[0] / reps
    rf_ms = _kernel_ms(ctx, 2)[0] / reps
    ctx.check(lib.r1c_profile(ctx.h, 0))
    try:
        ipeak = float(json.load(open(os.path.join(ROOT, "profiles", "int_peaks_r01.json")))["iadd"])
        ihow = "measured (tools/int_peak.cu, IADD3 lane-ops/s)"
    except Exception:
        ipeak, ihow = 36000.0, "fallback"
    ops = 2.0 * W * H * (2 * RANGE + 1) ** 2
    return {"workload": "%dx%d %d-bit full-search ME +-%d, 19 sizes, SAD then SATD sub-pel (planes resident)" % (W, H, BD, RANGE),
            "value": W * H / 1e3 / ms, "unit": "Mpix/s", "ms": ms, "launches_per_frame": launches,
            "blocks_per_frame": int(n_res),
            "kernel_ms": {"me_sb2_kernel": sb_ms, "me_refine_kernel": rf_ms},
            "roofline": {"bound": "int32_alu", "kernel": "me_sb2_kernel (u16 variant)" if BD <= 10 else "me_sb2_kernel (u16, 12-bit variant: true SADs in the keys)", "achieved": ops / (sb_ms / 1e3) / 1e9 if sb_ms else None,
                         "peak": ipeak, "unit": "Gop/s", "frac": ops / (sb_ms / 1e3) / 1e9 / ipeak if sb_ms else None,
                         "alg_ops_per_launch": ops, "peak_source": ihow}}
