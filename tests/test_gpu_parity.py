"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU
oracle on the same seeded inputs.  Bar: bit-exact (integer / byte / index work).
Run on the B200 box: python -m pytest tests -m gpu"""
import ctypes as C

import numpy as np
import pytest

from gpu_util import (U64_MAX, full_pixel_window, oracle_full_search, oracle_subpel_cost,
                      oracle_telescopic, twin_plane)
from test_oracle_golden import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def r():
    import rav1e_b200 as mod
    return mod


@pytest.fixture(scope="module")
def ctx(r):
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    c = r.Context(0)
    yield c
    c.close()


BD = [(1, 8), (2, 10), (2, 12)]


# ------------------------------------------------------------------ planes -----
@pytest.mark.parametrize("pb,bd", BD)
def test_plane_roundtrip_pad_downsample(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(bd)
    o, d = twin_plane(r, oracle, ctx, 200, 120, pb, bd, rng, xpad=40, ypad=24, pad=False)
    assert np.array_equal(d.download(), o.data)
    o.pad()
    d.pad(200, 120)
    assert np.array_equal(d.download(), o.data)
    # chroma-style plane: pad(w, h) takes luma dimensions
    oc, dc = twin_plane(r, oracle, ctx, 100, 60, pb, bd, rng, xpad=20, ypad=12, xdec=1, ydec=1, pad=False)
    oc.pad(200, 120)
    dc.pad(200, 120)
    assert np.array_equal(dc.download(), oc.data)
    # downsample_from
    oh = oracle.OPlane(100, 60, 1, 1, 20, 12, pb)
    dh = r.Plane(ctx, 100, 60, 1, 1, 20, 12, pb)
    assert oracle.lib().r1o_plane_downsample(oh.data.ctypes.data, C.byref(oh.cfg), o.data.ctypes.data,
                                             C.byref(o.cfg), pb) == 0
    dh.downsample_from(d)
    assert np.array_equal(dh.download(), oh.data)
    with pytest.raises(r.Rav1eCudaError):
        dh.downsample_from(dc)  # not 2x


# ------------------------------------------------------------ SAD / SATD -------
@pytest.mark.parametrize("pb", [1, 2])
def test_sad_satd_reference_goldens(r, ctx, oracle, pb):
    """the 44 known answers of src/dist.rs:379-402,437-460 through the C ABI"""
    from test_oracle_golden import setup_planes
    inp, rec = setup_planes(oracle, pb)
    d_in = r.Plane(ctx, 640, 480, 0, 0, 136, 136, pb).upload(inp.data)
    d_rec = r.Plane(ctx, 640, 480, 0, 0, 264, 264, pb).upload(rec.data)
    for b, (name, w, h, sad, satd) in enumerate(GOLDEN):
        assert r.get_sad(ctx, d_in, (32, 40), d_rec, (32, 40), b, 8) == sad, name
        assert r.get_satd(ctx, d_in, (32, 40), d_rec, (32, 40), b, 8) == satd, name


@pytest.mark.parametrize("pb,bd", BD)
def test_sad_satd_batch_random(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(100 + bd)
    o_org, d_org = twin_plane(r, oracle, ctx, 320, 200, pb, bd, rng)
    o_ref, d_ref = twin_plane(r, oracle, ctx, 320, 200, pb, bd, rng, xpad=136, ypad=100)
    n = 3000
    jobs = np.zeros(n, dtype=r.DIST_JOB)
    L = oracle.lib()
    for i in range(n):
        b = int(rng.integers(0, 22))
        w, h = L.r1o_block_width(b), L.r1o_block_height(b)
        jobs[i] = (int(rng.integers(-88, 320 + 88 - w + 1)), int(rng.integers(-88, 200 + 88 - h + 1)),
                   int(rng.integers(-136, 320 + 136 - w + 1)), int(rng.integers(-100, 200 + 100 - h + 1)), b)
    jobs[0] = (-88, -88, -136, -100, 15)  # corner of the allocation, 128x128
    want_sad = np.empty(n, np.uint32)
    want_satd = np.empty(n, np.uint32)
    flat = np.ascontiguousarray(jobs.view(np.int32).reshape(n, 5))
    L.r1o_dist_batch(0, o_org.origin_ptr(), C.byref(o_org.cfg), o_ref.origin_ptr(), C.byref(o_ref.cfg),
                     flat.ctypes.data, n, pb, want_sad.ctypes.data)
    L.r1o_dist_batch(1, o_org.origin_ptr(), C.byref(o_org.cfg), o_ref.origin_ptr(), C.byref(o_ref.cfg),
                     flat.ctypes.data, n, pb, want_satd.ctypes.data)
    assert np.array_equal(r.sad_batch(ctx, d_org, d_ref, jobs, bd), want_sad)
    assert np.array_equal(r.satd_batch(ctx, d_org, d_ref, jobs, bd), want_satd)
    # the lanes-per-block choice follows the list length (32 / 4 / 1 lanes): same answers for any mix of sizes
    for reps in (3, 6, 12, 20, 30):  # 16, 8, 4, 2, 1 lanes per block on a 148-SM part
        big = np.tile(jobs, reps)
        assert np.array_equal(r.sad_batch(ctx, d_org, d_ref, big, bd), np.tile(want_sad, reps))
        assert np.array_equal(r.satd_batch(ctx, d_org, d_ref, big, bd), np.tile(want_satd, reps))
    # planes wrapped around memory that is only pixel-aligned take the word-load path (no 16-byte vectors)
    import torch
    raw_o = torch.zeros(o_org.data.nbytes + 64, dtype=torch.uint8, device="cuda")
    raw_r = torch.zeros(o_ref.data.nbytes + 64, dtype=torch.uint8, device="cuda")
    off = 4  # word aligned, not 16-byte aligned
    raw_o[off:off + o_org.data.nbytes] = torch.from_numpy(o_org.data.view(np.uint8).reshape(-1)).cuda()
    raw_r[off:off + o_ref.data.nbytes] = torch.from_numpy(o_ref.data.view(np.uint8).reshape(-1)).cuda()
    w_org = r.Plane(ctx, 0, 0, px_bytes=pb, cfg=d_org.cfg, device_ptr=raw_o.data_ptr() + off)
    w_ref = r.Plane(ctx, 0, 0, px_bytes=pb, cfg=d_ref.cfg, device_ptr=raw_r.data_ptr() + off)
    assert np.array_equal(r.sad_batch(ctx, w_org, w_ref, jobs, bd), want_sad)
    assert np.array_equal(r.satd_batch(ctx, w_org, w_ref, np.tile(jobs, 12), bd), np.tile(want_satd, 12))


def test_satd_extreme_amplitude(r, ctx, oracle):
    """12-bit checkerboard vs inverse: the i32 / u64 path of get_satd_ref, which the
    fork's generated i16 kernel would wrap"""
    yy, xx = np.indices((128, 128))
    a = (((xx + yy) & 1) * 4095).astype(np.uint16)
    o_a, d_a = twin_plane(r, oracle, ctx, 128, 128, 2, 12, data=a)
    o_b, d_b = twin_plane(r, oracle, ctx, 128, 128, 2, 12, data=4095 - a)
    for b in range(22):
        w, h = oracle.lib().r1o_block_width(b), oracle.lib().r1o_block_height(b)
        assert r.get_satd(ctx, d_a, (0, 0), d_b, (0, 0), b, 12) == oracle.satd(o_a, 0, 0, o_b, 0, 0, w, h)
        assert r.get_sad(ctx, d_a, (0, 0), d_b, (0, 0), b, 12) == w * h * 4095


def test_satd_compat_i16_wrap(r, ctx, oracle):
    """the `compat_i16_wrap` switch (the arithmetic of the fork's generated kernels, build/kernel/gen/dist.rs:171-350):
    equal to get_satd_ref for 8-bit planes and for 10- / 12-bit input of ordinary amplitude, different from it -- and
    equal to the oracle's restatement -- on high-amplitude 10- / 12-bit blocks"""
    rng = np.random.default_rng(16)
    L = oracle.lib()
    for pb, bd in BD:
        a = rng.integers(0, 1 << bd, size=(128, 128))
        cases = [("noise", a, np.clip(a + rng.integers(-60, 61, size=a.shape), 0, (1 << bd) - 1)),
                 ("random", a, rng.integers(0, 1 << bd, size=a.shape))]
        yy, xx = np.indices((128, 128))
        board = ((xx + yy) & 1) * ((1 << bd) - 1)
        cases.append(("checkerboard", board, (1 << bd) - 1 - board))
        for name, x, y in cases:
            o_a, d_a = twin_plane(r, oracle, ctx, 128, 128, pb, bd, data=x.astype(np.uint16 if pb == 2 else np.uint8))
            o_b, d_b = twin_plane(r, oracle, ctx, 128, 128, pb, bd, data=y.astype(np.uint16 if pb == 2 else np.uint8))
            jobs = np.zeros(22 * 3, dtype=r.DIST_JOB)
            for i in range(jobs.size):
                jobs[i] = (int(rng.integers(0, 64)), int(rng.integers(0, 64)), int(rng.integers(-8, 64)), int(rng.integers(-8, 64)), i % 22)
                if L.r1o_block_width(i % 22) > 64 or L.r1o_block_height(i % 22) > 64:
                    jobs[i] = (0, 0, 0, 0, i % 22)
            if name == "noise":  # co-located blocks: the difference is the +-60 noise, far from 2^15 after 64 terms
                jobs["ref_x"], jobs["ref_y"] = jobs["org_x"], jobs["org_y"]
            got = r.satd_batch(ctx, d_a, d_b, jobs, bd, compat_i16_wrap=True)
            ref = r.satd_batch(ctx, d_a, d_b, jobs, bd)
            want = np.array([oracle.satd_i16(o_a, int(j["org_x"]), int(j["org_y"]), o_b, int(j["ref_x"]), int(j["ref_y"]),
                                             L.r1o_block_width(int(j["bsize"])), L.r1o_block_height(int(j["bsize"]))) for j in jobs],
                            dtype=np.uint32)
            assert np.array_equal(got, want), (bd, name)
            if bd == 8 or name == "noise":
                assert np.array_equal(got, ref), (bd, name)
            elif name == "checkerboard":  # 8x8 tiles in phase: one term of 64 * max, its halves double to 2^16 - 64 and wrap
                assert (got != ref).any(), (bd, name)


def test_dist_errors(r, ctx, oracle):
    rng = np.random.default_rng(5)
    _, d = twin_plane(r, oracle, ctx, 64, 64, 1, 8, rng, xpad=8, ypad=8)
    _, d16 = twin_plane(r, oracle, ctx, 64, 64, 2, 10, rng, xpad=8, ypad=8)
    with pytest.raises(r.Rav1eCudaError) as e:
        r.get_sad(ctx, d, (60, 0), d, (0, 0), r.BLOCK["BLOCK_64X64"], 8)  # leaves the allocation
    assert e.value.code == r.capi.R1C_ERR_BOUNDS
    with pytest.raises(r.Rav1eCudaError) as e:
        r.get_sad(ctx, d, (0, 0), d16, (0, 0), 0, 8)  # mixed pixel types
    assert e.value.code == r.capi.R1C_ERR_ARG
    with pytest.raises(r.Rav1eCudaError):
        r.get_sad(ctx, d, (0, 0), d, (0, 0), 22, 8)  # BLOCK_INVALID
    assert r.sad_batch(ctx, d, d, np.zeros(0, dtype=r.DIST_JOB), 8).size == 0  # empty batch


# ----------------------------------------------------------- forward TX --------
def _residual(rng, n, h, w, bd, kind):
    amp = (1 << bd) - 1
    if kind == "random":
        return rng.integers(-amp, amp + 1, size=(n, h, w)).astype(np.int16)
    if kind == "max":
        return np.full((n, h, w), amp, np.int16)
    if kind == "min":
        return np.full((n, h, w), -amp, np.int16)
    yy, xx = np.indices((h, w))
    return np.broadcast_to((((xx + yy) & 1) * 2 - 1) * amp, (n, h, w)).astype(np.int16)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_fwd_txfm_all_pairs(r, ctx, oracle, bd):
    """every (TxSize, TxType) the reference can instantiate (193 pairs), random +
    saturating inputs, batch sizes that do not fill the last CTA"""
    rng = np.random.default_rng(bd)
    L = oracle.lib()
    pairs = 0
    for s in range(19):
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        for t in range(16):
            if not L.r1o_fht_valid(s, t):
                with pytest.raises(r.Rav1eCudaError) as e:
                    r.forward_transform(ctx, np.zeros((1, h, w), np.int16), s, t, bd)
                assert e.value.code == r.capi.R1C_ERR_UNIMPLEMENTED
                continue
            pairs += 1
            n = int(rng.integers(1, 70))
            res = np.concatenate([_residual(rng, n, h, w, bd, "random"), _residual(rng, 1, h, w, bd, "max"),
                                  _residual(rng, 1, h, w, bd, "min"), _residual(rng, 1, h, w, bd, "checker")])
            got = r.forward_transform(ctx, res, s, t, bd)
            want = oracle.fht_batch(res, s, t, bd).reshape(res.shape)
            assert np.array_equal(got, want), (r.capi.TX_SIZES[s], r.capi.TX_TYPES[t], bd)
    assert pairs == 193


def test_fwd_txfm_single_and_empty(r, ctx, oracle):
    rng = np.random.default_rng(3)
    res = rng.integers(-255, 256, size=(16, 16)).astype(np.int16)
    got = r.forward_transform(ctx, res, r.TX_SIZE["TX_16X16"], r.TX_TYPE["ADST_DCT"], 8)
    assert np.array_equal(got, oracle.fht(res, 2, 1, 8))
    assert r.forward_transform(ctx, np.zeros((0, 8, 8), np.int16), 1, 0, 8).size == 0
    with pytest.raises(r.Rav1eCudaError):
        r.forward_transform(ctx, res, r.TX_SIZE["TX_16X16"], 0, 9)  # bit depth


@pytest.mark.parametrize("pb,bd", BD)
def test_diff_fwd_txfm_fused(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(40 + bd)
    o_in, d_in = twin_plane(r, oracle, ctx, 256, 128, pb, bd, rng)
    o_pr, d_pr = twin_plane(r, oracle, ctx, 256, 128, pb, bd, rng, xpad=40, ypad=40)
    L = oracle.lib()
    for s, t in [(0, 0), (1, 3), (2, 0), (3, 0), (4, 0), (5, 1), (12, 0), (13, 9), (18, 0), (8, 11)]:
        w, h = L.r1o_tx_width(s), L.r1o_tx_height(s)
        n = 37
        jobs = np.zeros(n, dtype=r.TX_JOB)
        res = np.zeros((n, h, w), np.int16)
        for i in range(n):
            jobs[i] = (int(rng.integers(0, 256 - w + 1)), int(rng.integers(0, 128 - h + 1)),
                       int(rng.integers(-40, 256 + 40 - w + 1)), int(rng.integers(-40, 128 + 40 - h + 1)))
            L.r1o_diff(res[i].ctypes.data, o_in.ptr(int(jobs[i]["in_x"]), int(jobs[i]["in_y"])), o_in.cfg.stride,
                       o_pr.ptr(int(jobs[i]["pred_x"]), int(jobs[i]["pred_y"])), o_pr.cfg.stride, w, h, pb)
        got = r.diff_forward_transform(ctx, d_in, d_pr, jobs, s, t, bd)
        assert np.array_equal(got, oracle.fht_batch(res, s, t, bd).reshape(n, h, w)), (s, t)


# -------------------------------------------------------------------- MC -------
SIZES_MC = [(4, 4), (4, 8), (8, 4), (8, 8), (8, 16), (16, 8), (16, 16), (16, 32), (32, 16), (32, 32),
            (32, 64), (64, 32), (64, 64), (64, 128), (128, 64), (128, 128), (4, 16), (16, 4), (8, 32),
            (32, 8), (16, 64), (64, 16), (2, 2), (4, 2), (2, 4), (2, 8), (6, 10)]


def _mc_jobs(r, rng, n, W, H, pad, dst_wh=None):
    jobs = np.zeros(n, dtype=r.MC_JOB)
    off = 0
    for i in range(n):
        w, h = SIZES_MC[i % len(SIZES_MC)]
        case = (i // len(SIZES_MC)) % 4
        cf = 0 if case in (0, 1) else int(rng.integers(1, 16))
        rf = 0 if case in (0, 2) else int(rng.integers(1, 16))
        sx = int(rng.integers(-pad + 3, W + pad - w - 4))
        sy = int(rng.integers(-pad + 3, H + pad - h - 4))
        dx = int(rng.integers(0, (dst_wh or (W, H))[0] - w + 1))
        dy = int(rng.integers(0, (dst_wh or (W, H))[1] - h + 1))
        jobs[i] = (dx, dy, sx, sy, w, h, cf, rf, int(rng.integers(0, 4)), int(rng.integers(0, 4)), off)
        off += w * h
    return jobs, off


@pytest.mark.parametrize("pb,bd", BD)
def test_prep_8tap_batch(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(200 + bd)
    o_src, d_src = twin_plane(r, oracle, ctx, 384, 256, pb, bd, rng, xpad=24, ypad=24)
    jobs, total = _mc_jobs(r, rng, 4 * len(SIZES_MC) * 3, 384, 256, 24)
    got = np.full(total + 5, -12345, np.int16)
    r.prep_8tap_batch(ctx, got, d_src, jobs, bd)
    want = np.full(total + 5, -12345, np.int16)
    L = oracle.lib()
    for j in jobs:
        L.r1o_prep_8tap(want[int(j["tmp_off"]):].ctypes.data, o_src.ptr(int(j["src_x"]), int(j["src_y"])),
                        o_src.cfg.stride, int(j["w"]), int(j["h"]), int(j["col_frac"]), int(j["row_frac"]),
                        int(j["mode_x"]), int(j["mode_y"]), bd, pb)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("pb,bd", BD)
def test_put_8tap_and_avg(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(300 + bd)
    o_src, d_src = twin_plane(r, oracle, ctx, 384, 256, pb, bd, rng, xpad=24, ypad=24)
    L = oracle.lib()
    # put: jobs are applied one by one on a fresh dst each, so overlapping dst blocks of
    # a batch cannot hide an error
    jobs, _ = _mc_jobs(r, rng, 4 * len(SIZES_MC), 384, 256, 24)
    for j in jobs:
        o_dst, d_dst = twin_plane(r, oracle, ctx, 384, 256, pb, bd, pad=False, xpad=8, ypad=8)
        r.put_8tap_batch(ctx, d_dst, d_src, j.reshape(1), bd)
        L.r1o_put_8tap(o_dst.ptr(int(j["dst_x"]), int(j["dst_y"])), o_dst.cfg.stride,
                       o_src.ptr(int(j["src_x"]), int(j["src_y"])), o_src.cfg.stride, int(j["w"]), int(j["h"]),
                       int(j["col_frac"]), int(j["row_frac"]), int(j["mode_x"]), int(j["mode_y"]), bd, pb)
        assert np.array_equal(d_dst.download(), o_dst.data), tuple(j)
    # a non-overlapping tiling in ONE batch + compound average
    o_dst, d_dst = twin_plane(r, oracle, ctx, 384, 256, pb, bd, pad=False, xpad=8, ypad=8)
    tiles = [(x, y) for y in range(0, 256, 32) for x in range(0, 384, 32)]
    jb = np.zeros(len(tiles), dtype=r.MC_JOB)
    ab = np.zeros(len(tiles), dtype=r.AVG_JOB)
    for i, (x, y) in enumerate(tiles):
        jb[i] = (x, y, x + int(rng.integers(-8, 9)), y + int(rng.integers(-8, 9)), 32, 32,
                 int(rng.integers(0, 16)), int(rng.integers(0, 16)), i % 4, (i // 4) % 4, i * 1024)
        ab[i] = (x, y, 32, 32, i * 1024)
    r.put_8tap_batch(ctx, d_dst, d_src, jb, bd)
    for j in jb:
        L.r1o_put_8tap(o_dst.ptr(int(j["dst_x"]), int(j["dst_y"])), o_dst.cfg.stride,
                       o_src.ptr(int(j["src_x"]), int(j["src_y"])), o_src.cfg.stride, 32, 32,
                       int(j["col_frac"]), int(j["row_frac"]), int(j["mode_x"]), int(j["mode_y"]), bd, pb)
    assert np.array_equal(d_dst.download(), o_dst.data)
    t1 = np.zeros(len(tiles) * 1024, np.int16)
    t2 = np.zeros(len(tiles) * 1024, np.int16)
    r.prep_8tap_batch(ctx, t1, d_src, jb, bd)
    jb2 = jb.copy()
    jb2["src_x"] += 5
    jb2["row_frac"] = (jb2["row_frac"] + 7) % 16
    r.prep_8tap_batch(ctx, t2, d_src, jb2, bd)
    r.mc_avg_batch(ctx, d_dst, t1, t2, ab, bd)
    for a in ab:
        o = int(a["tmp_off"])
        L.r1o_mc_avg(o_dst.ptr(int(a["dst_x"]), int(a["dst_y"])), o_dst.cfg.stride, t1[o:].ctypes.data,
                     t2[o:].ctypes.data, 32, 32, bd, pb)
    assert np.array_equal(d_dst.download(), o_dst.data)
    # extreme intermediates: saturated tmp values must clamp, not wrap
    t1[:] = 32767
    t2[:] = 32767
    r.mc_avg_batch(ctx, d_dst, t1, t2, ab, bd)
    assert (d_dst.download()[8:8 + 256, o_dst.cfg.xorigin:o_dst.cfg.xorigin + 384] == (1 << bd) - 1).all()


@pytest.mark.parametrize("pb,bd", BD)
def test_mc_avg_mixed_and_long_lists(r, ctx, oracle, pb, bd):
    """mc_avg takes chunks of 1..32 consecutive jobs per warp (the chunk size follows the list length): a mixed
    tiling with every width class incl. ragged ones (chunks that must search), a list long enough for 32-job chunks
    of one size, and one whose last chunk is short"""
    rng = np.random.default_rng(2350 + bd)
    L = oracle.lib()
    W, H = 2048, 1280

    def run(ab, tlen):
        t1 = rng.integers(-4000, 20000, size=tlen).astype(np.int16)
        t2 = rng.integers(-4000, 20000, size=tlen).astype(np.int16)
        o_dst, d_dst = twin_plane(r, oracle, ctx, W, H, pb, bd, pad=False, xpad=8, ypad=8)
        r.mc_avg_batch(ctx, d_dst, t1, t2, ab, bd)
        got = d_dst.download()
        # the oracle one job at a time is slow for 10^5 jobs: mc_avg_ref is elementwise (src/mc.rs:389-408), so the
        # whole plane is checked against that formula in numpy and the formula against the oracle on 600 jobs
        ib = 2 if bd == 12 else 4
        x0, y0 = o_dst.cfg.xorigin, o_dst.cfg.yorigin
        want = o_dst.data.copy()
        for a in ab:
            x, y, w, h, o = (int(a[k]) for k in ("dst_x", "dst_y", "w", "h", "tmp_off"))
            s = t1[o:o + w * h].astype(np.int32) + t2[o:o + w * h]
            want[y0 + y:y0 + y + h, x0 + x:x0 + x + w] = np.clip((s + (1 << ib)) >> (ib + 1), 0, (1 << bd) - 1).reshape(h, w)
        for a in list(ab[:300]) + list(ab[rng.integers(0, len(ab), 300)]):
            x, y, w, h, o = (int(a[k]) for k in ("dst_x", "dst_y", "w", "h", "tmp_off"))
            L.r1o_mc_avg(o_dst.ptr(x, y), o_dst.cfg.stride, t1[o:].ctypes.data, t2[o:].ctypes.data, w, h, bd, pb)
            assert np.array_equal(o_dst.data[y0 + y:y0 + y + h, x0 + x:x0 + x + w], want[y0 + y:y0 + y + h, x0 + x:x0 + x + w])
        assert np.array_equal(got, want)

    # (a) bands of mixed heights, each filled with mixed widths: ~5 * 10^3 jobs, one per warp;
    # (a') the same list three times over (-> chunks of 2 jobs, rarely of one size)
    jobs, off, y = [], 0, 0
    heights = [4, 8, 4, 16, 8, 4, 12, 4, 32, 8, 4, 64, 4, 8, 20, 4, 128, 7]  # 7: 4-wide blocks of odd height leave the row-pair walker
    while True:
        h = heights[len(jobs) % len(heights)] if y else 4
        if y + h > H:
            break
        x = 0
        while x < W:
            w = int(rng.choice([4, 8, 16, 4, 8, 32, 12, 6, 5, 3, 64, 128]))
            if x + w > W:
                break
            jobs.append((x, y, w, h, off))
            off += w * h + int(rng.integers(0, 3))  # odd offsets: the 8-byte loads must fall back
            x += w
        y += h
    order = rng.permutation(len(jobs))
    ab = np.array([jobs[i] for i in order], dtype=r.AVG_JOB)
    run(ab, off + 8)
    ab = np.concatenate([ab, ab[::-1], ab[rng.permutation(len(ab))]])
    assert len(ab) > 148 * 32 * 2  # more than 32 warps per SM of 2-job chunks
    run(ab, off + 8)
    # (b) 4x4 blocks over the plane: 163840 jobs -> 32-job chunks of one size
    xs, ys = np.meshgrid(np.arange(0, W, 4), np.arange(0, H, 4))
    ab = np.zeros(xs.size, dtype=r.AVG_JOB)
    ab["dst_x"], ab["dst_y"], ab["w"], ab["h"] = xs.ravel(), ys.ravel(), 4, 4
    ab["tmp_off"] = rng.permutation(xs.size) * 16
    run(ab, xs.size * 16)
    # (c) the same list cut so that the last chunk is short, with a few blocks of another size mixed in
    ab = ab[:xs.size - 13].copy()
    big = (ab["dst_x"] % 64 == 0) & (ab["dst_y"] % 64 == 0) & (rng.random(ab.size) < 0.3)
    ab["w"][big] = 8  # covers its right-hand neighbour: both would write the plane, so compare only ...
    keep = np.ones(ab.size, bool)
    nb = {(int(a["dst_x"]) + 4, int(a["dst_y"])) for a in ab[big]}
    keep &= ~np.array([(int(a["dst_x"]), int(a["dst_y"])) in nb for a in ab])  # ... after dropping that neighbour
    ab = ab[keep]
    ab["tmp_off"] = np.arange(ab.size) * 32
    run(ab, ab.size * 32)


@pytest.mark.parametrize("pb,bd", BD)
def test_mc_one_block_size_lists(r, ctx, oracle, pb, bd):
    """job lists of ONE block size take the two-pass tile kernel (mc_tile_kernel): put into disjoint tiles, prep,
    all four frac cases, every block size incl. segments of 16 rows (h > 16) and short lists / ragged batches"""
    rng = np.random.default_rng(2300 + bd)
    W, H = 384, 256
    o_src, d_src = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, xpad=24, ypad=24)
    L = oracle.lib()
    for (w, h) in SIZES_MC[:22] + [(12, 12), (8, 24)]:  # (8, 24): 24 % 16 != 0 -> strip kernel, still one size
        xs, ys = np.meshgrid(np.arange(0, W - w + 1, w), np.arange(0, H - h + 1, h))
        keep = rng.permutation(xs.size)[: int(rng.integers(1, min(xs.size, 150) + 1))]
        n = keep.size
        jobs = np.zeros(n, dtype=r.MC_JOB)
        for i, k in enumerate(keep):
            case = i % 4
            cf = 0 if case in (0, 1) else int(rng.integers(1, 16))
            rf = 0 if case in (0, 2) else int(rng.integers(1, 16))
            jobs[i] = (int(xs.flat[k]), int(ys.flat[k]), int(rng.integers(-24 + 3, W + 24 - w - 4)),
                       int(rng.integers(-24 + 3, H + 24 - h - 4)), w, h, cf, rf, int(rng.integers(0, 4)), int(rng.integers(0, 4)),
                       i * w * h)
        o_dst, d_dst = twin_plane(r, oracle, ctx, W, H, pb, bd, pad=False, xpad=8, ypad=8)
        r.put_8tap_batch(ctx, d_dst, d_src, jobs, bd)
        got = np.full(n * w * h + 3, -777, np.int16)
        r.prep_8tap_batch(ctx, got, d_src, jobs, bd)
        want = np.full(n * w * h + 3, -777, np.int16)
        for j in jobs:
            a = (o_src.ptr(int(j["src_x"]), int(j["src_y"])), o_src.cfg.stride, w, h, int(j["col_frac"]), int(j["row_frac"]),
                 int(j["mode_x"]), int(j["mode_y"]), bd, pb)
            L.r1o_put_8tap(o_dst.ptr(int(j["dst_x"]), int(j["dst_y"])), o_dst.cfg.stride, *a)
            L.r1o_prep_8tap(want[int(j["tmp_off"]):].ctypes.data, *a)
        assert np.array_equal(d_dst.download(), o_dst.data), (w, h)
        assert np.array_equal(got, want), (w, h)


def test_mc_errors(r, ctx, oracle):
    rng = np.random.default_rng(9)
    _, d = twin_plane(r, oracle, ctx, 64, 64, 1, 8, rng, xpad=8, ypad=8)
    j = np.zeros(1, dtype=r.MC_JOB)
    xo = int(d.cfg.xorigin)  # Plane::new aligns the left padding up (8 -> 32 px)
    j[0] = (0, 0, -xo, 0, 8, 8, 3, 0, 0, 0, 0)  # horizontal filter needs 3 px left of the allocation
    with pytest.raises(r.Rav1eCudaError) as e:
        r.put_8tap_batch(ctx, d, d, j, 8)
    assert e.value.code == r.capi.R1C_ERR_BOUNDS
    j[0] = (0, 0, -xo, 0, 8, 8, 0, 0, 0, 0, 0)  # the copy case at the same origin is legal
    r.put_8tap_batch(ctx, d, d, j, 8)
    j[0] = (0, 0, 0, 0, 8, 8, 16, 0, 0, 0, 0)
    with pytest.raises(r.Rav1eCudaError):
        r.put_8tap_batch(ctx, d, d, j, 8)
    tmp = np.zeros(10, np.int16)
    j[0] = (0, 0, 0, 0, 8, 8, 0, 0, 0, 0, 0)
    with pytest.raises(r.Rav1eCudaError):  # tmp smaller than the compute block
        r.prep_8tap_batch(ctx, tmp, d, j, 8)


# ----------------------------------------------------------- full search -------
@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10)])
def test_full_search_batch(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(400 + bd)
    o_ref, d_ref = twin_plane(r, oracle, ctx, 256, 160, pb, bd, rng, smooth=True)
    shifted = np.roll(o_ref.visible, (3, -5), axis=(0, 1)).astype(np.int64) + rng.integers(-2, 3, size=(160, 256))
    o_org, d_org = twin_plane(r, oracle, ctx, 256, 160, pb, bd,
                              data=np.clip(shifted, 0, (1 << bd) - 1))
    L = oracle.lib()
    n = 120
    jobs = np.zeros(n, dtype=r.FS_JOB)
    init = np.zeros(n, dtype=r.ME_RESULT)
    init["cost"] = U64_MAX
    for i in range(n):
        b = int(rng.integers(0, 22))
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        po = (int(rng.integers(0, (256 - bw) // 4 + 1)) * 4, int(rng.integers(0, (160 - bh) // 4 + 1)) * 4)
        rx, ry = int(rng.integers(0, 20)), int(rng.integers(0, 12))
        x_lo, x_hi = max(po[0] - rx, -88), min(po[0] + rx, 256 + 88 - bw)
        y_lo, y_hi = max(po[1] - ry, -88), min(po[1] + ry, 160 + 88 - bh)
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize"] = po[0], po[1], b
        jobs[i]["x_lo"], jobs[i]["x_hi"], jobs[i]["y_lo"], jobs[i]["y_hi"] = x_lo, x_hi, y_lo, y_hi
        jobs[i]["step"] = int(rng.integers(1, 4))
        jobs[i]["lambda_"] = int(rng.choice([0, 1, 57, 4000, 1 << 20]))
        jobs[i]["pmv"][0] = (int(rng.integers(-64, 64)), int(rng.integers(-64, 64)))
        jobs[i]["pmv"][1] = (int(rng.integers(-64, 64)), int(rng.integers(-64, 64)))
        jobs[i]["allow_high_precision_mv"] = i & 1
        if i % 3 == 0:  # continue from an earlier result, like the loop over cmvs (src/me.rs:534-565)
            jobs[i]["has_init"] = 1
            init[i]["mv"] = (8, -8)
            init[i]["cost"] = [0, 5000, 10 ** 7, U64_MAX][int(rng.integers(0, 4))]
    got = r.full_search_batch(ctx, d_org, d_ref, jobs, bd, init.copy())
    for i in range(n):
        ini = ((int(init[i]["mv"]["row"]), int(init[i]["mv"]["col"])), int(init[i]["cost"])) if jobs[i]["has_init"] else None
        want = oracle_full_search(oracle, o_org, o_ref, jobs[i], ini)
        assert ((int(got[i]["mv"]["row"]), int(got[i]["mv"]["col"])), int(got[i]["cost"])) == want, i


def test_full_search_ties_and_errors(r, ctx, oracle):
    """flat planes: every candidate ties, the first raster position must win"""
    o, d = twin_plane(r, oracle, ctx, 128, 96, 1, 8)
    mv, cost = r.full_search(ctx, 30, 40, 28, 36, r.BLOCK["BLOCK_8X8"], d, d, (0, 0), U64_MAX, (32, 32), 2, 0,
                             [(0, 0), (0, 0)], False, 8)
    assert (mv, cost) == (((28 - 32) * 8, (30 - 32) * 8), 0)
    # with a rate term the zero MV wins instead
    mv, cost = r.full_search(ctx, 24, 40, 24, 40, r.BLOCK["BLOCK_8X8"], d, d, (0, 0), U64_MAX, (32, 32), 1, 10,
                             [(0, 0), (0, 0)], False, 8)
    assert (mv, cost) == ((0, 0), 0)
    with pytest.raises(r.Rav1eCudaError) as e:
        r.full_search(ctx, -200, 40, 28, 36, 3, d, d, (0, 0), U64_MAX, (32, 32), 1, 0, [(0, 0), (0, 0)], False, 8)
    assert e.value.code == r.capi.R1C_ERR_BOUNDS
    # quarter-resolution style job: 16x16 -> 4x4 blocks on a small plane (estimate_motion_ss4)
    rng = np.random.default_rng(8)
    o_q, d_q = twin_plane(r, oracle, ctx, 120, 68, 1, 8, rng, xpad=22, ypad=22, xdec=2, ydec=2)
    o_p, d_p = twin_plane(r, oracle, ctx, 120, 68, 1, 8, rng, xpad=22, ypad=22, xdec=2, ydec=2)
    j = np.zeros(1, dtype=r.FS_JOB)
    j[0]["po_x"], j[0]["po_y"], j[0]["bsize"] = 16, 16, 0
    j[0]["x_lo"], j[0]["x_hi"], j[0]["y_lo"], j[0]["y_hi"], j[0]["step"], j[0]["lambda_"] = -20, 64, 0, 32, 1, 3
    got = r.full_search_batch(ctx, d_q, d_p, j, 8)
    want = oracle_full_search(oracle, o_q, o_p, j[0])
    assert ((int(got[0]["mv"]["row"]), int(got[0]["mv"]["col"])), int(got[0]["cost"])) == want


# -------------------------------------------------- superblock full search -----
def _check_sb(r, oracle, o_org, o_ref, params, jobs, out, W4, H4):
    L = oracle.lib()
    checked = 0
    for ji, job in enumerate(jobs):
        for b in range(22):
            if r.sb_result_offset(b) < 0 or not (int(job["bsize_mask"]) >> b) & 1:
                continue
            bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
            k = 0
            for ry in range(0, 64, bh):
                for rx in range(0, 64, bw):
                    got = out[ji, r.sb_result_offset(b) + k]
                    k += 1
                    po = (int(job["po_x"]) + rx, int(job["po_y"]) + ry)
                    if po[0] + bw > W4 or po[1] + bh > H4:
                        assert int(got["cost"]) == 12345, "blocks outside the frame must stay untouched"
                        continue
                    cmv = (int(job["cmv"]["row"]), int(job["cmv"]["col"]))
                    x_lo, x_hi, y_lo, y_hi, _ = full_pixel_window(oracle, params.w_in_b, params.h_in_b, po, bw, bh,
                                                                  cmv, params.range_x, params.range_y)
                    fj = np.zeros(1, dtype=r.FS_JOB)[0]
                    fj["po_x"], fj["po_y"], fj["bsize"] = po[0], po[1], b
                    fj["x_lo"], fj["x_hi"], fj["y_lo"], fj["y_hi"] = x_lo, x_hi, y_lo, y_hi
                    fj["step"], fj["lambda_"] = params.step, params.lambda_
                    fj["pmv"] = job["pmv"]
                    fj["allow_high_precision_mv"] = params.allow_high_precision_mv
                    want = oracle_full_search(oracle, o_org, o_ref, fj)
                    assert want is not None
                    assert ((int(got["mv"]["row"]), int(got["mv"]["col"])), int(got["cost"])) == want, \
                        (ji, r.capi.BLOCK_SIZES[b], rx, ry)
                    checked += 1
    return checked


@pytest.mark.parametrize("pb,bd,step,lam", [(1, 8, 1, 0), (1, 8, 1, 700), (1, 8, 2, 33), (2, 10, 1, 90), (2, 12, 1, 1500),
                                            (2, 12, 1, 20000)])  # 20000: above the 12-bit limit of the fast search -> me_sb.cu
def test_me_sb_matches_per_block_full_search(r, ctx, oracle, pb, bd, step, lam):
    """every block of every size of every superblock (borders included) against the
    oracle's full_search run block by block with the reference's window rule"""
    rng = np.random.default_rng(500 + bd + step + lam)
    W, H = 192, 136  # 136 = 2 full superblock rows + a partial one
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    noisy = np.roll(o_ref.visible, (-2, 6), axis=(0, 1)).astype(np.int64) + rng.integers(-3, 4, size=(H, W))
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, (1 << bd) - 1))
    params = r.make_sb_params(W // 4, H // 4, 20, 12, step, lam)
    pos = [(x, y) for y in range(0, H, 64) for x in range(0, W, 64)] + [(64, H - 64)]  # + an adjust_bo style row
    jobs = np.zeros(len(pos), dtype=r.SB_JOB)
    for i, (x, y) in enumerate(pos):
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize_mask"] = x, y, 0x3FFFFF
        jobs[i]["pmv"][0] = (int(rng.integers(-40, 40)), int(rng.integers(-40, 40)))
        jobs[i]["pmv"][1] = (int(rng.integers(-40, 40)), int(rng.integers(-40, 40)))
    jobs[4]["cmv"] = (-16, 24)  # an off-centre search on an interior superblock
    res = np.zeros(len(pos) * r.SB_RESULTS, dtype=r.ME_RESULT)
    res["cost"] = 12345
    out = r.me_sb_batch(ctx, d_org, d_ref, params, jobs, bd, res)
    n = _check_sb(r, oracle, o_org, o_ref, params, jobs, out, W, H)
    assert n > 849 * 6


@pytest.mark.parametrize("pb,bd,kind", [(2, 10, "extremes"), (2, 10, "smooth"), (1, 8, "extremes"), (2, 12, "extremes"),
                                        (2, 12, "smooth")])
def test_me_sb_pm64_window(r, ctx, oracle, pb, bd, kind):
    """the +-64 px window of the headline workload (the specialised kernel instances), 8-, 10- and 12-bit; `extremes` =
    pixels at 0 / max only, so block SADs reach their largest values (key fields, 16-bit half accumulators)"""
    rng = np.random.default_rng(77 + bd)
    W, H = 256, 192
    maxv = (1 << bd) - 1
    if kind == "extremes":
        ref_px = rng.integers(0, 2, size=(H, W)) * maxv
        org_px = np.where(rng.random((H, W)) < 0.9, maxv - np.roll(ref_px, (3, -5), axis=(0, 1)), ref_px)
    else:
        ref_px = None
    if ref_px is None:
        o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
        noisy = np.roll(o_ref.visible, (-37, 51), axis=(0, 1)).astype(np.int64) + rng.integers(-3, 4, size=(H, W))
        o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, maxv))
    else:
        o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, data=ref_px)
        o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=org_px)
    params = r.make_sb_params(W // 4, H // 4, 64, 64, 1, 160)
    pos = [(64, 64), (192, 128)]  # interior, and the bottom-right corner (masked candidates)
    jobs = np.zeros(len(pos), dtype=r.SB_JOB)
    for i, (x, y) in enumerate(pos):
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize_mask"] = x, y, 0x3FFFFF
        jobs[i]["pmv"][0] = (int(rng.integers(-200, 200)), int(rng.integers(-200, 200)))
        jobs[i]["pmv"][1] = (int(rng.integers(-200, 200)), int(rng.integers(-200, 200)))
    res = np.zeros(len(pos) * r.SB_RESULTS, dtype=r.ME_RESULT)
    res["cost"] = 12345
    out = r.me_sb_batch(ctx, d_org, d_ref, params, jobs, bd, res)
    assert _check_sb(r, oracle, o_org, o_ref, params, jobs, out, W, H) == 849 * 2


def test_me_sb_flat_ties_and_mask(r, ctx, oracle):
    o, d = twin_plane(r, oracle, ctx, 128, 128, 1, 8)  # all 128: every candidate ties at SAD 0
    params = r.make_sb_params(32, 32, 9, 7, 1, 0)
    jobs = np.zeros(1, dtype=r.SB_JOB)
    jobs[0]["po_x"], jobs[0]["po_y"] = 64, 64
    jobs[0]["bsize_mask"] = (1 << r.BLOCK["BLOCK_8X8"]) | (1 << r.BLOCK["BLOCK_64X64"])
    res = np.zeros(r.SB_RESULTS, dtype=r.ME_RESULT)
    res["cost"] = 12345
    out = r.me_sb_batch(ctx, d, d, params, jobs, 8, res)[0]
    for b in range(22):
        off, cnt = r.sb_result_offset(b), r.sb_result_count(b)
        if off < 0:
            continue
        blk = out[off:off + cnt]
        if (int(jobs[0]["bsize_mask"]) >> b) & 1:  # first raster candidate: (-7, -9) px
            assert (blk["mv"]["row"] == -56).all() and (blk["mv"]["col"] == -72).all() and (blk["cost"] == 0).all()
        else:
            assert (blk["cost"] == 12345).all()
    with pytest.raises(r.Rav1eCudaError):
        jobs[0]["po_x"] = 62
        r.me_sb_batch(ctx, d, d, params, jobs, 8)


# ------------------------------------------------------------- sub-pel ---------
@pytest.mark.parametrize("pb,bd,use_satd,hp", [(1, 8, 1, 0), (1, 8, 0, 1), (2, 10, 1, 0), (2, 12, 1, 1)])
def test_subpel_cost_and_telescopic(r, ctx, oracle, pb, bd, use_satd, hp):
    rng = np.random.default_rng(600 + bd + use_satd)
    W, H = 192, 128
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    noisy = np.roll(o_ref.visible, (1, -2), axis=(0, 1)).astype(np.int64) + rng.integers(-2, 3, size=(H, W))
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, (1 << bd) - 1))
    L = oracle.lib()
    n = 60
    jobs = np.zeros(n, dtype=r.SUBPEL_JOB)
    start = np.zeros(n, dtype=r.ME_RESULT)
    for i in range(n):
        b = int(rng.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 16, 17, 18, 19, 20, 21]))
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        po = (int(rng.integers(0, (W - bw) // 4 + 1)) * 4, int(rng.integers(0, (H - bh) // 4 + 1)) * 4)
        rngv = (C.c_int64 * 4)()
        L.r1o_mv_range(W // 4, H // 4, po[0] // 4, po[1] // 4, bw, bh, C.byref(rngv))
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize"] = po[0], po[1], b
        jobs[i]["lambda_"] = int(rng.choice([0, 40, 900]))
        jobs[i]["pmv"][0] = (int(rng.integers(-30, 30)), int(rng.integers(-30, 30)))
        jobs[i]["mvx_min"], jobs[i]["mvx_max"], jobs[i]["mvy_min"], jobs[i]["mvy_max"] = tuple(rngv)
        jobs[i]["use_satd"], jobs[i]["allow_high_precision_mv"] = use_satd, hp
        start[i]["mv"] = (int(rng.integers(-3, 4)) * 8, int(rng.integers(-3, 4)) * 8)
        start[i]["cost"] = [0, 30000, U64_MAX][int(rng.integers(0, 3))]
    # a block in the corner whose candidates leave the legal MV range
    jobs[0]["po_x"], jobs[0]["po_y"], jobs[0]["bsize"] = 0, 0, 3
    rngv = (C.c_int64 * 4)()
    L.r1o_mv_range(W // 4, H // 4, 0, 0, 8, 8, C.byref(rngv))
    jobs[0]["mvx_min"], jobs[0]["mvx_max"], jobs[0]["mvy_min"], jobs[0]["mvy_max"] = tuple(rngv)
    start[0]["mv"] = (rngv[2], rngv[0])
    cand = np.zeros((n, 5), dtype=r.MV)
    for i in range(n):
        for k in range(5):
            cand[i, k] = (int(start[i]["mv"]["row"]) + int(rng.integers(-9, 10)),
                          int(start[i]["mv"]["col"]) + int(rng.integers(-9, 10)))
    got = r.subpel_cost_batch(ctx, d_org, d_ref, jobs, cand, bd)
    for i in range(n):
        for k in range(5):
            want = oracle_subpel_cost(oracle, o_org, o_ref, jobs[i], (int(cand[i, k]["row"]), int(cand[i, k]["col"])), bd)
            assert int(got[i, k]) == want, (i, k)
    ref = r.subpel_refine_batch(ctx, d_org, d_ref, jobs, start, bd)
    for i in range(n):
        want = oracle_telescopic(oracle, o_org, o_ref, jobs[i],
                                 (int(start[i]["mv"]["row"]), int(start[i]["mv"]["col"])), int(start[i]["cost"]), bd)
        assert ((int(ref[i]["mv"]["row"]), int(ref[i]["mv"]["col"])), int(ref[i]["cost"])) == want, i


# ------------------------------------------------- whole-frame motion search ----
def _frame_pair(r, oracle, ctx, W, H, pb, bd, seed):
    rng = np.random.default_rng(seed)
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    noisy = np.roll(o_ref.visible, (3, -5), axis=(0, 1)).astype(np.int64) + rng.integers(-3, 4, size=(H, W))
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, (1 << bd) - 1))
    return o_org, d_org, o_ref, d_ref


@pytest.mark.parametrize("pb,bd,refine,use_satd,hp", [(1, 8, 1, 1, 0), (1, 8, 0, 0, 0), (2, 10, 1, 1, 1), (1, 8, 1, 0, 0),
                                                      (2, 10, 1, 0, 0), (2, 12, 1, 1, 0), (2, 12, 1, 0, 1), (2, 12, 1, 1, 1)])
def test_frame_me_matches_per_block_motion_estimation(r, ctx, oracle, pb, bd, refine, use_satd, hp):
    """r1c_frame_me == motion_estimation (src/me.rs:193-271) block by block, every size,
    borders and the partial last superblock row included"""
    W, H = 200, 152
    o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, pb, bd, 900 + bd + refine + use_satd)
    params = r.make_frame_me_params(W, H, 14, 10, 77, refine=refine, use_satd=use_satd, allow_hp=hp)
    got = r.frame_me(ctx, d_org, d_ref, params, bd)
    mv, cost = oracle.frame_me(o_org, o_ref, 14, 10, 77, params.bsize_mask, refine, use_satd, hp, bd)
    assert got.size == cost.size == r.frame_me_count(params) > 5000
    assert np.array_equal(got["cost"], cost)
    assert np.array_equal(got["mv"]["row"], mv[:, 0]) and np.array_equal(got["mv"]["col"], mv[:, 1])


@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10), (2, 12)])
def test_frame_me_extreme_pixels(r, ctx, oracle, pb, bd):
    """pixels at 0 / max only: the largest differences the sub-pel filters, the packed pair Hadamard (4x4 tiles at 10
    bits, every tile at 8, none at 12) and the distortion accumulators can see; filter overshoot is clamped on both sides;
    12 bits also go through the fused sub-pel kernel, with intermediate_bits = 2"""
    rng = np.random.default_rng(4100 + bd)
    W, H = 136, 128
    maxv = (1 << bd) - 1
    blocks = rng.integers(0, 2, size=(H // 2 + 1, W // 2 + 1)) * maxv  # 2x2 checker noise: strong sub-pel response
    ref_px = np.kron(blocks, np.ones((2, 2), dtype=np.int64))[:H, :W]
    org_px = np.where(rng.random((H, W)) < 0.5, maxv - np.roll(ref_px, (1, -2), axis=(0, 1)), np.roll(ref_px, (1, -2), axis=(0, 1)))
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, data=ref_px)
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=org_px)
    params = r.make_frame_me_params(W, H, 6, 6, 20, refine=1, use_satd=1, allow_hp=1)
    got = r.frame_me(ctx, d_org, d_ref, params, bd)
    mv, cost = oracle.frame_me(o_org, o_ref, 6, 6, 20, params.bsize_mask, 1, 1, 1, bd)
    assert np.array_equal(got["cost"], cost)
    assert np.array_equal(got["mv"]["row"], mv[:, 0]) and np.array_equal(got["mv"]["col"], mv[:, 1])
    # the largest costs the search can see: every pixel differs by the full range, every candidate ties on the SAD
    # and the rate decides; lambda at the limit of the fast superblock search (60000, and 16000 at 12 bits, where
    # 256 * SAD of a 64x64 block + rate * lambda is just below 2^32), full-pel only so that the costs stay SAD-based
    lam = 16000 if bd == 12 else 60000
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.zeros((H, W), dtype=np.int64))
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.full((H, W), maxv, dtype=np.int64))
    params = r.make_frame_me_params(W, H, 6, 6, lam, refine=0, use_satd=0, allow_hp=0)
    got = r.frame_me(ctx, d_org, d_ref, params, bd)
    mv, cost = oracle.frame_me(o_org, o_ref, 6, 6, lam, params.bsize_mask, 0, 0, 0, bd)
    assert int(cost.max()) >= 256 * 64 * 64 * maxv
    assert np.array_equal(got["cost"], cost)
    assert np.array_equal(got["mv"]["row"], mv[:, 0]) and np.array_equal(got["mv"]["col"], mv[:, 1])


def test_frame_me_sharded_rows_equal_whole_frame(r, ctx, oracle):
    """multi-GPU sharding by superblock row: the shards of two 'ranks' fill one result
    array identical to the whole-frame call; untouched entries keep their values"""
    W, H = 192, 200
    o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, 1, 8, 77)
    mask = (1 << r.BLOCK["BLOCK_8X8"]) | (1 << r.BLOCK["BLOCK_16X16"]) | (1 << r.BLOCK["BLOCK_64X64"]) | 1
    whole = r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(W, H, 9, 9, 31, bsize_mask=mask), 8)
    part = np.zeros(whole.size, dtype=r.ME_RESULT)
    part["cost"] = 4242
    p0 = r.make_frame_me_params(W, H, 9, 9, 31, bsize_mask=mask, sb_rows=(0, 2))
    r.frame_me(ctx, d_org, d_ref, p0, 8, part)
    assert (part["cost"] == 4242).any() and not np.array_equal(part, whole)
    p1 = r.make_frame_me_params(W, H, 9, 9, 31, bsize_mask=mask, sb_rows=(2, 4))
    r.frame_me(ctx, d_org, d_ref, p1, 8, part)
    assert np.array_equal(part, whole)
    with pytest.raises(r.Rav1eCudaError):
        r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(W + 4, H, 9, 9, 31), 8)


def test_frame_me_tiles_equal_whole_frame_and_pack(r, ctx, oracle):
    """multi-GPU sharding by tile (superblock rows x columns, rav1e_b200.shard.split_tiles): the tiles of 2, 3, 4 and 6
    'ranks' partition the result array and fill it identically to the whole-frame call, for 8- and 10-bit planes with the
    sub-pel tail; r1c_frame_me_compact_dev packs a tile's results in shard_slots order"""
    import torch
    from rav1e_b200 import shard
    W, H = 256, 200
    for pb, bd in ((1, 8), (2, 10)):
        o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, pb, bd, 78)
        whole = r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(W, H, 9, 9, 31), bd)
        for world in (2, 3, 4, 6):
            tiles = shard.split_tiles(W, H, world)
            part = np.zeros(whole.size, dtype=r.ME_RESULT)
            part["cost"] = 4242
            seen = np.zeros(whole.size, dtype=np.int32)
            for (lo, hi), (clo, chi) in tiles:
                p = r.make_frame_me_params(W, H, 9, 9, 31, sb_rows=(lo, hi), sb_cols=(clo, chi))
                slots = shard.shard_slots(W, H, r.ALL_SIZES_TO_64, lo, hi, clo, chi)
                seen[slots] += 1
                before = part.copy()
                r.frame_me(ctx, d_org, d_ref, p, bd, part)
                untouched = np.ones(whole.size, dtype=bool)
                untouched[slots] = False
                assert np.array_equal(part[untouched], before[untouched]), (bd, world)
                assert ctx.lib.r1c_frame_me_shard_count(ctx.h, C.byref(p)) == slots.size
                # the packed form of this tile, from a device copy of the whole-frame result
                d_all = torch.from_numpy(whole.view(np.uint8).copy()).cuda()
                d_cmp = torch.zeros(slots.size * 16, dtype=torch.uint8, device="cuda")
                ctx.check(ctx.lib.r1c_frame_me_compact_dev(ctx.h, C.byref(p), d_all.data_ptr(), d_cmp.data_ptr()))
                ctx.sync()
                assert np.array_equal(d_cmp.cpu().numpy().view(r.ME_RESULT), whole[slots]), (bd, world)
            assert np.all(seen == 1), (bd, world)
            assert np.array_equal(part, whole), (bd, world)


def test_tile_search_needs_only_its_rows_of_the_planes(r, ctx, oracle):
    """r1c_plane_upload_rows_async + shard.tile_plane_rows: planes pre-filled with garbage, only the rows a tile can touch
    uploaded -- the tile's results equal the whole-frame search on the full planes (8- and 10-bit, sub-pel tail on)"""
    from rav1e_b200 import shard
    W, H, RNGE = 192, 448, 16
    for pb, bd in ((1, 8), (2, 10)):
        o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, pb, bd, 91)
        whole = r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(W, H, RNGE, RNGE, 31), bd)
        junk = np.full(o_org.data.shape, (1 << bd) - 1, dtype=o_org.data.dtype)
        junk[::2, 1::3] = 0
        for (lo, hi) in ((0, 2), (2, 5), (5, 7)):
            g_org = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(junk)
            g_ref = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(junk)
            r0, r1 = shard.tile_plane_rows(g_ref.cfg, lo, hi, RNGE)
            assert r1 - r0 < g_ref.cfg.alloc_height  # a real subset
            for g, o in ((g_org, o_org), (g_ref, o_ref)):
                host = np.ascontiguousarray(o.data)
                ctx.check(ctx.lib.r1c_plane_upload_rows_async(ctx.h, g.h, host.ctypes.data, r0, r1))
            ctx.sync()
            p = r.make_frame_me_params(W, H, RNGE, RNGE, 31, sb_rows=(lo, hi))
            slots = shard.shard_slots(W, H, r.ALL_SIZES_TO_64, lo, hi)
            got = r.frame_me(ctx, g_org, g_ref, p, bd)
            assert np.array_equal(got[slots], whole[slots]), (bd, lo, hi)
    with pytest.raises(r.Rav1eCudaError) as e:
        ctx.check(ctx.lib.r1c_plane_upload_rows_async(ctx.h, d_org.h, o_org.data.ctypes.data, 0, d_org.cfg.alloc_height + 1))
    assert e.value.code == r.capi.R1C_ERR_ARG


def test_tile_search_needs_only_its_rectangle_of_the_planes(r, ctx, oracle):
    """r1c_plane_upload_rect_async + shard.tile_plane_rows / tile_plane_cols: planes pre-filled with garbage, only the
    rectangle a tile of a 2 x 3 tile grid can touch uploaded -- the tile's results equal the whole-frame search on the
    full planes (8- and 10-bit, sub-pel tail on)"""
    from rav1e_b200 import shard
    W, H, RNGE = 704, 256, 16
    for pb, bd in ((1, 8), (2, 10)):
        o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, pb, bd, 93)
        whole = r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(W, H, RNGE, RNGE, 31), bd)
        junk = np.full(o_org.data.shape, (1 << bd) - 1, dtype=o_org.data.dtype)
        junk[::2, 1::3] = 0
        for (lo, hi) in ((0, 2), (2, 4)):
            for (clo, chi) in ((0, 4), (4, 8), (8, 11)):
                g_org = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(junk)
                g_ref = r.Plane(ctx, W, H, 0, 0, 88, 88, pb).upload(junk)
                r0, r1 = shard.tile_plane_rows(g_ref.cfg, lo, hi, RNGE)
                c0, c1 = shard.tile_plane_cols(g_ref.cfg, clo, chi, RNGE)
                assert c1 - c0 < g_ref.cfg.stride and c0 % 64 == 0  # a real subset, on 128-byte lines
                for g, o in ((g_org, o_org), (g_ref, o_ref)):
                    host = np.ascontiguousarray(o.data)
                    ctx.check(ctx.lib.r1c_plane_upload_rect_async(ctx.h, g.h, host.ctypes.data, r0, r1, c0, c1))
                ctx.sync()
                left = g_ref.download()
                assert np.array_equal(left[r0:r1, c0:c1], o_ref.data[r0:r1, c0:c1])
                keep = np.ones(left.shape, bool)
                keep[r0:r1, c0:c1] = False
                assert np.array_equal(left[keep], junk[keep])  # nothing else was written
                p = r.make_frame_me_params(W, H, RNGE, RNGE, 31, sb_rows=(lo, hi), sb_cols=(clo, chi))
                slots = shard.shard_slots(W, H, r.ALL_SIZES_TO_64, lo, hi, clo, chi)
                got = r.frame_me(ctx, g_org, g_ref, p, bd)
                assert np.array_equal(got[slots], whole[slots]), (bd, lo, hi, clo, chi)
    with pytest.raises(r.Rav1eCudaError) as e:
        ctx.check(ctx.lib.r1c_plane_upload_rect_async(ctx.h, d_org.h, o_org.data.ctypes.data, 0, 8, 0, d_org.cfg.stride + 1))
    assert e.value.code == r.capi.R1C_ERR_ARG


def test_frame_me_window_equals_frame_by_frame(r, ctx, oracle):
    """r1c_frame_me_window(_dev): a lookahead window of frame pairs over two streams == r1c_frame_me pair by pair; host
    form with the planes uploaded by the call (garbage in the planes before), device form with resident planes; odd and
    even window lengths, a window of one, an empty window; a bad pair is reported and nothing hangs"""
    import torch
    W, H = 192, 136
    for pb, bd in ((1, 8), (2, 10)):
        params = r.make_frame_me_params(W, H, 12, 12, 40)
        n_res = r.frame_me_count(params)
        pairs, want, hosts = [], [], []
        for k in range(5):
            o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, W, H, pb, bd, 300 + k)
            want.append(r.frame_me(ctx, d_org, d_ref, params, bd))
            pairs.append((d_org, d_ref))
            hosts.append((o_org.data.copy(), o_ref.data.copy()))
        assert not np.array_equal(want[0], want[1])
        for n in (5, 4, 1, 0):
            got = r.frame_me_window(ctx, pairs[:n], params, bd)
            assert len(got) == n and all(np.array_equal(g, w) for g, w in zip(got, want)), (bd, n)
        # host form with uploads: the planes hold garbage until the call uploads them
        junk = np.full(hosts[0][0].shape, (1 << bd) - 1, dtype=hosts[0][0].dtype)
        for d_org, d_ref in pairs:
            d_org.upload(junk)
            d_ref.upload(junk)
        got = r.frame_me_window(ctx, [(a, b, ho, hr) for (a, b), (ho, hr) in zip(pairs, hosts)], params, bd)
        assert all(np.array_equal(g, w) for g, w in zip(got, want)), bd
        # device form: results stay on the device; later work on the context stream sees both streams' results
        d_out = [torch.zeros(n_res * 16, dtype=torch.uint8, device="cuda") for _ in pairs]
        arr = (r.capi.FramePair * len(pairs))()
        for i, (a, b) in enumerate(pairs):
            arr[i].org, arr[i].ref, arr[i].results = a.h, b.h, d_out[i].data_ptr()
        ctx.check(ctx.lib.r1c_frame_me_window_dev(ctx.h, arr, len(pairs), C.byref(params), bd))
        ctx.sync()
        for i, w in enumerate(want):
            assert np.array_equal(d_out[i].cpu().numpy().view(r.ME_RESULT), w), (bd, i)
        # a pair whose plane does not match the frame size: error code, the call returns
        _, d_small = twin_plane(r, oracle, ctx, W - 64, H, pb, bd, np.random.default_rng(1))
        with pytest.raises(r.Rav1eCudaError) as e:
            r.frame_me_window(ctx, [pairs[0], (d_small, pairs[1][1]), pairs[2]], params, bd)
        assert e.value.code == r.capi.R1C_ERR_ARG
        assert np.array_equal(r.frame_me(ctx, pairs[3][0], pairs[3][1], params, bd), want[3])  # the context is still usable


def test_page_locked_host_buffers_are_dma_targets(r, ctx, oracle):
    """host-array entry points take page-locked buffers as the DMA source / target (no staging copy) and pageable ones
    through the pinned stage: same results either way (r1c_fwd_txfm_batch, r1c_prep_8tap_batch, r1c_frame_me)"""
    import torch
    rng = np.random.default_rng(8)
    res = rng.integers(-1023, 1024, size=(500, 16, 16)).astype(np.int16)
    want = r.forward_transform(ctx, res, 2, 0, 10)
    h_in = torch.from_numpy(res.copy()).pin_memory()
    h_out = torch.full((res.size,), -1, dtype=torch.int32).pin_memory()
    ctx.check(ctx.lib.r1c_fwd_txfm_batch(ctx.h, 2, 0, 10, h_in.data_ptr(), h_out.data_ptr(), 500))
    assert np.array_equal(h_out.numpy().reshape(want.shape), want)
    assert np.array_equal(want.reshape(-1), oracle.fht_batch(res, 2, 0, 10).reshape(-1))
    o_src, d_src = twin_plane(r, oracle, ctx, 256, 128, 2, 10, rng)
    jobs = np.zeros(40, dtype=r.MC_JOB)
    for i in range(40):
        jobs[i] = (0, 0, int(rng.integers(0, 240)), int(rng.integers(0, 112)), 16, 16, int(rng.integers(0, 16)), int(rng.integers(0, 16)),
                   0, 0, 300 * i)
    a = np.full(40 * 300, 77, dtype=np.int16)
    r.prep_8tap_batch(ctx, a, d_src, jobs, 10)
    b = torch.full((40 * 300,), 77, dtype=torch.int16).pin_memory().numpy()
    r.prep_8tap_batch(ctx, b, d_src, jobs, 10)
    assert np.array_equal(a, b) and (a.reshape(40, 300)[:, 256:] == 77).all() and (a.reshape(40, 300)[:, :256] != 77).any()


def test_profile_hooks_count_launches(r, ctx, oracle):
    o_org, d_org, o_ref, d_ref = _frame_pair(r, oracle, ctx, 128, 128, 1, 8, 5)
    ctx.check(ctx.lib.r1c_profile(ctx.h, 1))
    r.frame_me(ctx, d_org, d_ref, r.make_frame_me_params(128, 128, 8, 8, 10), 8)
    ms, n = C.c_double(), C.c_int64()
    ctx.check(ctx.lib.r1c_profile_read(ctx.h, 0, C.byref(ms), C.byref(n)))
    ctx.check(ctx.lib.r1c_profile(ctx.h, 0))
    assert n.value == 1 and ms.value > 0


# ------------------------------------------------------------ RDO distortion ----
def _rdo_jobs(r, rng, W, H, n, mult, max_log=5):
    """random regions on the MI grid of a W x H plane, sizes `mult` * 2^k up to 128"""
    jobs = np.zeros(n, dtype=r.RDO_JOB)
    for i in range(n):
        w = min(mult << int(rng.integers(0, max_log + 1)), 128, W)
        h = min(mult << int(rng.integers(0, max_log + 1)), 128, H)
        x = int(rng.integers(0, (W - w) // mult + 1)) * mult
        y = int(rng.integers(0, (H - h) // mult + 1)) * mult
        jobs[i] = (x, y, x, y, w, h)
    return jobs


@pytest.mark.parametrize("pb,bd", BD)
def test_sse_wxh_and_cdef_dist_match_oracle(r, ctx, oracle, pb, bd):
    rng = np.random.default_rng(1200 + bd)
    W, H = 256, 192
    o1, d1 = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, xpad=16, ypad=16)
    o2, d2 = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, xpad=48, ypad=32)  # different stride
    w_in_b, h_in_b = W // 4, H // 4
    w_imp = w_in_b // 2
    imp = (rng.random(w_imp * (h_in_b // 2)) * 5).astype(np.float32)
    for cdef in (False, True):
        jobs = _rdo_jobs(r, rng, W, H, 400, 8 if cdef else 4, 4 if cdef else 5)
        for use_imp in (False, True):
            kw = dict(block_importances=imp, w_in_imp_b=w_imp, w_in_b=w_in_b, h_in_b=h_in_b) if use_imp else {}
            got = r.sse_wxh(ctx, d1, d2, jobs, bd, cdef=cdef, **kw)
            for j, g in zip(jobs, got):
                a = (o1, int(j["src_x"]), int(j["src_y"]), o2, int(j["rec_x"]), int(j["rec_y"]), int(j["w"]), int(j["h"]))
                okw = {k: v for k, v in kw.items()}
                want = oracle.cdef_dist_wxh(*a, bd, **okw) if cdef else oracle.sse_wxh(*a, **okw)
                assert int(g) == want, (cdef, use_imp, j)
    # frame edge: the importance sum stops at w_in_b / h_in_b but divides by the full area
    jobs = np.array([(W - 8, H - 8, W - 8, H - 8, 8, 8)], dtype=r.RDO_JOB)
    got = r.sse_wxh(ctx, d1, d2, jobs, bd, imp, w_imp, w_in_b - 1, h_in_b - 1)
    assert int(got[0]) == oracle.sse_wxh(o1, W - 8, H - 8, o2, W - 8, H - 8, 8, 8, imp, w_imp, w_in_b - 1, h_in_b - 1)


def test_sse_wxh_chroma_plane_and_errors(r, ctx, oracle):
    rng = np.random.default_rng(31)
    o1, d1 = twin_plane(r, oracle, ctx, 128, 96, 2, 10, rng, xpad=8, ypad=8, xdec=1, ydec=1)
    o2, d2 = twin_plane(r, oracle, ctx, 128, 96, 2, 10, rng, xpad=8, ypad=8, xdec=1, ydec=1)
    w_in_b, h_in_b = 64, 48  # luma 256 x 192
    imp = (rng.random(32 * 24) * 3).astype(np.float32)
    jobs = _rdo_jobs(r, rng, 128, 96, 300, 4, 4)
    got = r.sse_wxh(ctx, d1, d2, jobs, 10, imp, 32, w_in_b, h_in_b)
    for j, g in zip(jobs, got):
        want = oracle.sse_wxh(o1, int(j["src_x"]), int(j["src_y"]), o2, int(j["rec_x"]), int(j["rec_y"]), int(j["w"]),
                              int(j["h"]), imp, 32, w_in_b, h_in_b)
        assert int(g) == want, j
    with pytest.raises(r.Rav1eCudaError):  # cdef_dist is luma-only
        r.cdef_dist_wxh(ctx, d1, d2, jobs[:1], 10)
    with pytest.raises(r.Rav1eCudaError):  # not a multiple of MI_SIZE
        r.sse_wxh(ctx, d1, d2, np.array([(0, 0, 0, 0, 6, 8)], dtype=r.RDO_JOB), 10)
    with pytest.raises(r.Rav1eCudaError):  # leaves the plane
        r.sse_wxh(ctx, d1, d2, np.array([(120, 90, 120, 90, 64, 64)], dtype=r.RDO_JOB), 10)
    assert r.sse_wxh(ctx, d1, d2, np.zeros(0, dtype=r.RDO_JOB), 10).size == 0


def test_tx_distortion_matches_oracle(r, ctx, oracle):
    rng = np.random.default_rng(8)
    L = oracle.lib()
    for ts in range(19):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        n = 37
        # large magnitudes: the i32 square wraps like the reference's release build
        c = rng.integers(-(1 << 17), 1 << 17, size=(n, w * h)).astype(np.int32)
        rc = rng.integers(-(1 << 17), 1 << 17, size=(n, coded)).astype(np.int32)
        got = r.tx_distortion(ctx, c, rc, ts)
        for i in range(n):
            assert int(got[i]) == oracle.tx_dist(c[i], rc[i], ts), (ts, i)


# ---------------------------------------------------------- diamond search -----
@pytest.mark.parametrize("pb,bd,subpixel,use_satd,hp", [(1, 8, 0, 0, 0), (1, 8, 1, 1, 0), (2, 10, 0, 1, 0), (2, 10, 1, 0, 1),
                                                         (2, 12, 1, 1, 1)])
def test_diamond_search_matches_oracle(r, ctx, oracle, pb, bd, subpixel, use_satd, hp):
    """diamond_me_search + get_best_predictor (src/me.rs:655-785): full-pel (radius 16 -> 8) and sub-pel
    (4 -> 2 / 1) variants, predictor lists with out-of-range entries and duplicates"""
    rng = np.random.default_rng(1500 + bd + 2 * subpixel + use_satd)
    W, H = 192, 128
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    noisy = np.roll(o_ref.visible, (5, -9), axis=(0, 1)).astype(np.int64) + rng.integers(-2, 3, size=(H, W))
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, (1 << bd) - 1))
    L = oracle.lib()
    n, n_pred = 80, 5
    jobs = np.zeros(n, dtype=r.SUBPEL_JOB)
    pred = np.zeros((n, n_pred), dtype=r.MV)
    for i in range(n):
        b = int(rng.choice([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 16, 17, 18, 19, 20, 21]))
        bw, bh = L.r1o_block_width(b), L.r1o_block_height(b)
        po = (int(rng.integers(0, (W - bw) // 4 + 1)) * 4, int(rng.integers(0, (H - bh) // 4 + 1)) * 4)
        rngv = (C.c_int64 * 4)()
        L.r1o_mv_range(W // 4, H // 4, po[0] // 4, po[1] // 4, bw, bh, C.byref(rngv))
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize"] = po[0], po[1], b
        jobs[i]["lambda_"] = int(rng.choice([0, 40, 900]))
        jobs[i]["pmv"][0] = (int(rng.integers(-30, 30)), int(rng.integers(-30, 30)))
        jobs[i]["pmv"][1] = (int(rng.integers(-30, 30)), int(rng.integers(-30, 30)))
        jobs[i]["mvx_min"], jobs[i]["mvx_max"], jobs[i]["mvy_min"], jobs[i]["mvy_max"] = tuple(rngv)
        jobs[i]["use_satd"], jobs[i]["allow_high_precision_mv"] = use_satd, hp
        pred[i, 0] = (0, 0)
        for k in range(1, n_pred - 1):  # sub-pel valued predictors also reach the full-pel path (mv / 8 truncates)
            pred[i, k] = (int(rng.integers(-80, 81)), int(rng.integers(-80, 81)))
        pred[i, n_pred - 1] = pred[i, 1]  # padding by repetition
        if i % 7 == 0:
            pred[i, 2] = (int(rngv[3]) + 8, 0)  # out of the legal range: u64::MAX, never chosen
    got = r.diamond_search_batch(ctx, d_org, d_ref, jobs, pred, bd, subpixel=bool(subpixel))
    moved = 0
    for i in range(n):
        j = jobs[i]
        bw, bh = L.r1o_block_width(int(j["bsize"])), L.r1o_block_height(int(j["bsize"]))
        mv, cost, iters = oracle.diamond_search(
            o_org, o_ref, (int(j["po_x"]), int(j["po_y"])), bw, bh, [(int(p["row"]), int(p["col"])) for p in pred[i]],
            [(int(j["pmv"][0]["row"]), int(j["pmv"][0]["col"])), (int(j["pmv"][1]["row"]), int(j["pmv"][1]["col"]))],
            int(j["lambda_"]), use_satd, hp, (int(j["mvx_min"]), int(j["mvx_max"]), int(j["mvy_min"]), int(j["mvy_max"])),
            subpixel, bd)
        assert ((int(got[i]["mv"]["row"]), int(got[i]["mv"]["col"])), int(got[i]["cost"])) == (mv, cost), i
        moved += iters > (3 if subpixel and hp else 2)
    assert moved > n // 4  # the search really walks (not only the shrink steps)


# ------------------------------------------------ pyramid stages of FullSearch ----
@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10)])
def test_estimate_motion_pyramid_matches_oracle(r, ctx, oracle, pb, bd):
    """estimate_motion_ss4 (quarter res), estimate_motion_ss2 (half res, <= 3 pmvs) and estimate_motion with
    FullSearch::full_pixel_me (full res, <= 7 cmvs, step 2) on a frame pyramid built by downsample_from"""
    rng = np.random.default_rng(1700 + bd)
    W, H = 256, 192
    planes = {}
    for name, shift in (("ref", (0, 0)), ("org", (6, -10))):
        o0 = oracle.OPlane(W, H, 0, 0, 88, 88, pb)
        if name == "ref":
            v = rng.integers(0, 1 << bd, size=(H, W)).astype(np.float64)
            for _ in range(4):
                v = (v + np.roll(v, 1, 0) + np.roll(v, 1, 1) + np.roll(v, -1, 0) + np.roll(v, -1, 1)) / 5
            base = np.clip(np.rint(v), 0, (1 << bd) - 1)
            o0.visible[...] = base.astype(o0.dtype)
        else:
            o0.visible[...] = np.clip(np.roll(base, shift, axis=(0, 1)) + rng.integers(-2, 3, size=(H, W)), 0,
                                      (1 << bd) - 1).astype(o0.dtype)
        o0.pad()
        levels = [o0]
        for lvl in (1, 2):  # hres / qres: Plane::new(w/2, h/2, 1, 1, pad/2, pad/2) ..., src/encoder.rs:361-376
            o = oracle.OPlane(W >> lvl, H >> lvl, lvl, lvl, 88 >> lvl, 88 >> lvl, pb)
            assert oracle.lib().r1o_plane_downsample(o.data.ctypes.data, C.byref(o.cfg), levels[-1].data.ctypes.data,
                                                     C.byref(levels[-1].cfg), pb) == 0
            o.pad(W, H)
            levels.append(o)
        planes[name] = [(o, r.Plane(ctx, int(o.cfg.width), int(o.cfg.height), int(o.cfg.xdec), int(o.cfg.ydec), int(o.cfg.xpad),
                                    int(o.cfg.ypad), pb).upload(o.data)) for o in levels]
    w_in_b, h_in_b = W // 4, H // 4
    L = oracle.lib()
    me_lambda = 1.37
    # ss4: 64x64 / 32x32 / 16x16 blocks (quarter-res 16x16 / 8x8 / 4x4), offsets incl. ones adjust_bo pulls back inside
    jobs = np.array([(bx, by, b) for b in (12, 9, 6) for by in range(0, h_in_b + 4, 8) for bx in range(0, w_in_b + 4, 8)],
                    dtype=r.BLOCK_JOB)
    got = r.estimate_motion_ss4_batch(ctx, planes["org"][2][1], planes["ref"][2][1], w_in_b, h_in_b, jobs, 1, me_lambda, bd)
    hits = 0
    for j, g in zip(jobs, got):
        bw, bh = L.r1o_block_width(int(j["bsize"])), L.r1o_block_height(int(j["bsize"]))
        want = oracle.estimate_motion_ss4(planes["org"][2][0], planes["ref"][2][0], w_in_b, h_in_b, (int(j["bo_x"]), int(j["bo_y"])),
                                          bw, bh, 1, me_lambda, False)
        assert (int(g["row"]), int(g["col"])) == want, j
        hits += abs(want[0] + 48) <= 32 and abs(want[1] - 80) <= 32  # org(y, x) = ref(y - 6, x + 10)
    assert hits > jobs.size // 2
    # ss2 and full res: predictors around the ss4 result, some slots empty, one block with none
    for ss2 in (True, False):
        slots = 3 if ss2 else 7
        jobs = np.array([(bx, by, b) for b in (12, 9, 6, 3) for by in range(0, h_in_b, 16) for bx in range(0, w_in_b + 4, 16)],
                        dtype=r.BLOCK_JOB)
        preds = np.zeros((jobs.size, slots), dtype=r.MV)
        valid = np.zeros((jobs.size, slots), dtype=np.uint8)
        for i in range(jobs.size):
            for s in range(slots):
                preds[i, s] = (-48 + int(rng.integers(-40, 41)), 80 + int(rng.integers(-40, 41)))
                valid[i, s] = i > 0 and rng.integers(0, 4) != 0
        lvl = 1 if ss2 else 0
        got = r.estimate_motion_batch(ctx, planes["org"][lvl][1], planes["ref"][lvl][1], w_in_b, h_in_b, jobs, preds, valid,
                                      me_lambda, bd, ss2=ss2)
        for i, (j, g) in enumerate(zip(jobs, got)):
            bw, bh = L.r1o_block_width(int(j["bsize"])), L.r1o_block_height(int(j["bsize"]))
            want = oracle.estimate_motion_pred(planes["org"][lvl][0], planes["ref"][lvl][0], w_in_b, h_in_b,
                                               (int(j["bo_x"]), int(j["bo_y"])), bw, bh,
                                               [(int(p["row"]), int(p["col"])) for p in preds[i]], valid[i], me_lambda, False, ss2)
            assert (int(g["row"]), int(g["col"])) == want, (ss2, j)
        assert (int(got[0]["row"]), int(got[0]["col"])) == (0, 0)  # no predictor: MotionVector::default()


# ------------------------------------------------------------ MV field ---------
def test_mv_field_save_and_subset_predictors(r, ctx, oracle):
    """save_block_motion (src/encoder.rs:1432-1444) and get_subset_predictors (src/me.rs:82-174) on a 48 x 32 MI field,
    then the predictor lists straight into the diamond search (n_pred = 17)"""
    rng = np.random.default_rng(1900)
    cols, rows = 48, 32
    L = oracle.lib()
    field = np.zeros((rows, cols), dtype=r.MV)
    want = np.zeros((rows, cols, 2), dtype=np.int16)
    # non-overlapping blocks: a 16x16-px grid with random block sizes that fit a cell, some hanging over the field edge
    jobs, mvs = [], []
    for by in range(0, rows, 4):
        for bx in range(0, cols, 4):
            b = int(rng.choice([0, 1, 2, 3, 4, 5, 6, 16, 17]))
            if bx == cols - 4 and by == rows - 4:
                b = 12  # 64x64 clipped by the field
            jobs.append((bx, by, b))
            mvs.append((int(rng.integers(-300, 301)), int(rng.integers(-300, 301))) if rng.integers(0, 5) else (0, 0))
    jobs = np.array(jobs, dtype=r.BLOCK_JOB)
    mvs = np.array(mvs, dtype=r.MV)
    r.save_block_motion(ctx, field, jobs, mvs)
    for j, m in zip(jobs, mvs):
        oracle.save_block_motion(want, (int(j["bo_x"]), int(j["bo_y"])), L.r1o_block_width(int(j["bsize"])),
                                 L.r1o_block_height(int(j["bsize"])), (int(m["row"]), int(m["col"])))
    assert np.array_equal(field["row"], want[..., 0]) and np.array_equal(field["col"], want[..., 1])
    prev = np.zeros((rows, cols), dtype=r.MV)
    prev["row"] = rng.integers(-200, 201, size=(rows, cols)) * (rng.integers(0, 3, size=(rows, cols)) > 0)
    prev["col"] = rng.integers(-200, 201, size=(rows, cols)) * (rng.integers(0, 3, size=(rows, cols)) > 0)
    prev3 = np.stack([prev["row"], prev["col"]], axis=-1)
    pj = np.array([(x, y, 3) for y in range(rows) for x in range(cols)], dtype=r.BLOCK_JOB)  # every cell incl. all borders
    cm = np.zeros((pj.size, 7), dtype=r.MV)
    cm["row"] = rng.integers(-500, 501, size=cm.shape)
    cm["col"] = rng.integers(-500, 501, size=cm.shape)
    ncm = rng.integers(0, 8, size=pj.size).astype(np.uint8)
    for use_prev in (True, False):
        got, cnt = r.get_subset_predictors(ctx, field, prev if use_prev else None, pj, cm, ncm)
        for i, j in enumerate(pj):
            w = oracle.get_subset_predictors((int(j["bo_x"]), int(j["bo_y"])), [(int(m["row"]), int(m["col"])) for m in cm[i, :ncm[i]]],
                                             want, prev3 if use_prev else None)
            assert cnt[i] == len(w) <= 17, i
            g = [(int(m["row"]), int(m["col"])) for m in got[i]]
            assert g[:len(w)] == w and all(m == w[-1] for m in g[len(w):]), i
    # chain: predictor lists -> diamond search
    o_ref, d_ref = twin_plane(r, oracle, ctx, cols * 4, rows * 4, 1, 8, rng, smooth=True)
    o_org, d_org = twin_plane(r, oracle, ctx, cols * 4, rows * 4, 1, 8, data=np.roll(o_ref.visible, (2, -3), axis=(0, 1)))
    sel = np.arange(0, pj.size, 37)
    sj = np.zeros(sel.size, dtype=r.SUBPEL_JOB)
    for k, i in enumerate(sel):
        rngv = (C.c_int64 * 4)()
        L.r1o_mv_range(cols, rows, int(pj[i]["bo_x"]), int(pj[i]["bo_y"]), 8, 8, C.byref(rngv))
        po = (min(int(pj[i]["bo_x"]) * 4, cols * 4 - 8), min(int(pj[i]["bo_y"]) * 4, rows * 4 - 8))
        sj[k]["po_x"], sj[k]["po_y"], sj[k]["bsize"], sj[k]["lambda_"] = po[0], po[1], 3, 50
        sj[k]["mvx_min"], sj[k]["mvx_max"], sj[k]["mvy_min"], sj[k]["mvy_max"] = tuple(rngv)
    got, cnt = r.get_subset_predictors(ctx, field, prev, pj[sel], cm[sel], ncm[sel])
    res = r.diamond_search_batch(ctx, d_org, d_ref, sj, got, 8, subpixel=False)
    for k in range(sel.size):
        j = sj[k]
        mv, cost, _ = oracle.diamond_search(o_org, o_ref, (int(j["po_x"]), int(j["po_y"])), 8, 8,
                                            [(int(m["row"]), int(m["col"])) for m in got[k, :cnt[k]]], [(0, 0), (0, 0)], 50, False, False,
                                            (int(j["mvx_min"]), int(j["mvx_max"]), int(j["mvy_min"]), int(j["mvy_max"])), False, 8)
        assert ((int(res[k]["mv"]["row"]), int(res[k]["mv"]["col"])), int(res[k]["cost"])) == (mv, cost), k


@pytest.mark.parametrize("pb,bd", BD)
def test_lookahead_intra_costs(r, ctx, oracle, pb, bd):
    """compute_lookahead_intra_costs (src/api/internal.rs:680-765): SATD of every 8x8 block vs pred_dc_128"""
    rng = np.random.default_rng(2100 + bd)
    W, H = 200, 120  # w_in_b = 50, h_in_b = 30 -> 25 x 15 importance blocks
    o, d = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    w_imp, h_imp = (W // 4) // 2, (H // 4) // 2
    got = r.lookahead_intra_costs(ctx, d, w_imp, h_imp, bd)
    assert np.array_equal(got, oracle.lookahead_intra_costs(o, w_imp, h_imp, bd))
    with pytest.raises(r.Rav1eCudaError):
        r.lookahead_intra_costs(ctx, d, 1000, 1000, bd)


# ------------------------------------------------ quantize / dequantize (8f.1) -----
@pytest.mark.parametrize("is_intra", [0, 1])
def test_quantize_matches_oracle(r, ctx, oracle, is_intra):
    from quant_util import scans_for, synth_coeffs
    L = oracle.lib()
    rng = np.random.default_rng(31 + is_intra)
    for ts in range(19):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        for (dcq, acq), kind, scan in zip(((8, 8), (37, 41), (256, 300), (1336, 1828)),
                                          ("dense", "decay", "sparse", "edge"), scans_for(rng, w, h)):
            n = 45 if w * h <= 256 else 11  # not a multiple of the blocks per warp / CTA
            co = synth_coeffs(rng, n, w * h, kind, acq)
            oq = oracle.quant_ctx(ts, is_intra, dcq, acq)
            want_q, want_r = oracle.quantize_batch(oq, ts, scan, co, want_rcoeffs=True)
            qc = r.QuantizationContext().update(ctx.lib, ts, is_intra, dcq, acq)
            assert bytes(qc.c) == bytes(oq), "QuantizationContext::update"
            got_q = qc.quantize(ctx, co, scan)
            assert got_q.shape == (n, coded)
            assert np.array_equal(got_q, want_q), (ts, kind)
            assert np.array_equal(r.dequantize(ctx, got_q, ts, dcq, acq), want_r), (ts, kind)


def test_quantize_fused_recon_and_tx_dist(r, ctx, oracle):
    """device entry with both optional outputs: rcoeffs = dequantize(quantize), tx_dist = src/encoder.rs:1210-1224"""
    import torch
    from quant_util import scans_for, synth_coeffs
    L = oracle.lib()
    rng = np.random.default_rng(41)
    for ts in (0, 1, 3, 4, 7, 12, 17):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        n, dcq, acq = 77, 120, 147
        scan = scans_for(rng, w, h)[0]
        co = synth_coeffs(rng, n, w * h, "decay", acq)
        co[0, :] = rng.integers(-(1 << 20), 1 << 20, size=w * h)  # squares that wrap i32
        oq = oracle.quant_ctx(ts, 0, dcq, acq)
        want_q, want_r = oracle.quantize_batch(oq, ts, scan, co, want_rcoeffs=True)
        qc = r.QuantizationContext().update(ctx.lib, ts, 0, dcq, acq)
        d_co, d_sc = torch.from_numpy(co).cuda(), torch.from_numpy(scan.view(np.int16)).cuda()
        d_q = torch.empty((n, coded), dtype=torch.int32, device="cuda")
        d_r = torch.empty((n, coded), dtype=torch.int32, device="cuda")
        d_d = torch.empty(n, dtype=torch.int64, device="cuda")
        torch.cuda.synchronize()
        ctx.check(ctx.lib.r1c_quantize_batch_dev(ctx.h, C.byref(qc.c), ts, d_sc.data_ptr(), d_co.data_ptr(), d_q.data_ptr(),
                                                 d_r.data_ptr(), d_d.data_ptr(), n))
        ctx.sync()
        assert np.array_equal(d_q.cpu().numpy(), want_q) and np.array_equal(d_r.cpu().numpy(), want_r), ts
        dist = d_d.cpu().numpy().view(np.uint64)
        for i in range(n):
            assert int(dist[i]) == oracle.tx_dist(co[i], want_r[i], ts), (ts, i)


def test_quantize_errors_and_empty(r, ctx, oracle):
    qc = r.QuantizationContext().update(ctx.lib, 1, 0, 100, 120)
    co = np.zeros((2, 64), dtype=np.int32)
    bad = np.arange(64, dtype=np.uint16)
    bad[5] = 6  # not a permutation
    with pytest.raises(r.Rav1eCudaError):
        qc.quantize(ctx, co, bad)
    bad = np.arange(64, dtype=np.uint16)[::-1].copy()  # scan[0] != 0
    with pytest.raises(r.Rav1eCudaError):
        qc.quantize(ctx, co, bad)
    with pytest.raises(r.Rav1eCudaError):  # context updated for another transform size class
        qc32 = r.QuantizationContext().update(ctx.lib, 3, 0, 100, 120)
        qc32.tx_size = 1
        qc32.quantize(ctx, co, np.arange(64, dtype=np.uint16))
    with pytest.raises(r.Rav1eCudaError):  # lossless / out-of-range quantizer
        r.QuantizationContext().update(ctx.lib, 1, 0, 0, 120)
    assert qc.quantize(ctx, np.zeros((0, 64), dtype=np.int32), np.arange(64, dtype=np.uint16)).shape == (0, 64)
    # all-zero blocks: eob = 1, everything zero
    assert not qc.quantize(ctx, co, np.arange(64, dtype=np.uint16)).any()


def test_quantize_mode_flipping_maps(r, ctx, oracle):
    """hand-made offsets (offset0 = +quant, offset1 = -quant) make coefficients whose map FLIPS the level mode
    (q0 > 1 and q1 == 0) -- never the case with update()'s offsets -- so the kernel's general composition of
    transition maps is exercised too"""
    from quant_util import scans_for, synth_coeffs
    L = oracle.lib()
    rng = np.random.default_rng(77)
    for ts in (0, 1, 2, 3, 9):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        acq = 64
        oq = oracle.quant_ctx(ts, 0, 50, acq)
        oq.ac_offset0, oq.ac_offset1 = acq, -acq
        qc = r.QuantizationContext().update(ctx.lib, ts, 0, 50, acq)
        qc.c.ac_offset0, qc.c.ac_offset1 = acq, -acq
        scan = scans_for(rng, w, h)[3]
        co = synth_coeffs(rng, 40, w * h, "edge", acq)
        want = oracle.quantize_batch(oq, ts, scan, co)
        assert np.array_equal(qc.quantize(ctx, co, scan), want), ts


# ------------------------------------------------ block importances (8f.4) -----
@pytest.mark.parametrize("pb,bd", [(1, 8), (2, 10)])
def test_block_importances_match_oracle(r, ctx, oracle, pb, bd):
    """compute_block_importances (src/api/internal.rs:823-1075): three chained propagations into one reference array
    (the f32 `+=` order matters), MVs of every phase incl. negative ones and blocks moved off the grid, then the
    final log2 pass"""
    rng = np.random.default_rng(900 + bd)
    W, H = 328, 200
    w, h = W // 8, H // 8
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    ref_imp_o = (rng.random((h, w)) * 50).astype(np.float32)
    ref_imp_g = ref_imp_o.copy()
    for it, n_refs in enumerate((1, 2, 3)):
        o_frm, d_frm = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
        mvs = rng.integers(-60 * 8, 60 * 8 + 1, size=(2 * h, 2 * w, 2)).astype(np.int16)
        mvs[::2, ::2][rng.random((h, w)) < 0.3] = 0
        intra = rng.integers(0, 4000, size=(h, w)).astype(np.uint32)  # zeros too: 0 / 0 and x / 0
        fut = (rng.random((h, w)) * (1000 if it else 0)).astype(np.float32)
        oracle.importance_propagate(o_frm, o_ref, mvs, intra, fut, n_refs, bd, ref_imp_o)
        mv_rec = np.zeros((2 * h, 2 * w), dtype=r.MV)
        mv_rec["row"], mv_rec["col"] = mvs[..., 0], mvs[..., 1]
        out = r.importance_propagate(ctx, d_frm, d_ref, mv_rec, intra, fut, n_refs, bd, ref_imp_g)
        assert out is ref_imp_g
        assert np.array_equal(ref_imp_g.view(np.uint32), ref_imp_o.view(np.uint32)), (it, "propagate")
    oracle.importance_finalize(intra, ref_imp_o)
    r.importance_finalize(ctx, intra, ref_imp_g)
    # log2 is the one transcendental of the path: the reference takes it from the libm it is linked with (not
    # correctly rounded, differs between platforms), the kernel rounds an f64 log2 once.  Tolerance: 1 ulp of f32.
    ulp = np.abs(ref_imp_g.view(np.int32).astype(np.int64) - ref_imp_o.view(np.int32).astype(np.int64))
    assert ulp.max() <= 1, int(ulp.max())
    assert np.array_equal(ref_imp_g == 0, ref_imp_o == 0)
    # an MV that leaves the allocated reference plane is refused like the reference's get_satd assert
    mv_rec[0, 0] = (-(88 + 1) * 8, 0)
    with pytest.raises(r.Rav1eCudaError):
        r.importance_propagate(ctx, d_frm, d_ref, mv_rec, intra, fut, 1, bd, ref_imp_g)
    with pytest.raises(r.Rav1eCudaError):
        r.importance_propagate(ctx, d_frm, d_ref, mv_rec[:3], intra, fut, 1, bd, ref_imp_g)
