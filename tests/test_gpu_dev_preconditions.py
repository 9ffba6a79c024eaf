"""The `_dev` entry points take job lists that live on the device, so the host cannot validate them the way the
host-array forms do.  What they do with jobs that break a precondition is part of the ABI (include/rav1e_cuda.h) and is
pinned here:
  * r1c_me_sb_batch(_dev): a search centre that is not a legal MV of the region -- host form R1C_ERR_ARG, device form
    (mv 0, cost UINT64_MAX) for the blocks whose whole window is illegal; r1c_full_search_batch with the reference's
    clamped window (src/me.rs:533-546) is the call that reproduces the reference there;
  * r1c_save_block_motion_batch_dev / r1c_subset_predictors_batch_dev: jobs outside the MV field are skipped on the
    device (nothing written / count 0), the host forms return R1C_ERR_ARG."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import full_pixel_window, oracle_full_search, twin_plane

pytestmark = pytest.mark.gpu
U64_MAX = np.iinfo(np.uint64).max


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


def test_sb_search_centre_outside_the_legal_range(r, ctx, oracle):
    import torch
    rng = np.random.default_rng(3)
    W = H = 128
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, 1, 8, rng, smooth=True)
    o_org, d_org = twin_plane(r, oracle, ctx, W, H, 1, 8, data=np.roll(o_ref.visible, (1, 2), axis=(0, 1)))
    params = r.make_sb_params(W // 4, H // 4, 8, 8, 1, 40)
    mask = (1 << r.BLOCK["BLOCK_8X8"]) | (1 << r.BLOCK["BLOCK_64X64"])
    jobs = np.zeros(2, dtype=r.SB_JOB)
    jobs[0]["po_x"], jobs[0]["po_y"], jobs[0]["bsize_mask"] = 0, 0, mask  # border superblock, centre 100 px to the left
    jobs[0]["cmv"]["row"], jobs[0]["cmv"]["col"] = 0, -100 * 8
    jobs[1]["po_x"], jobs[1]["po_y"], jobs[1]["bsize_mask"] = 64, 64, mask  # a legal neighbour in the same list
    with pytest.raises(r.Rav1eCudaError) as e:
        r.me_sb_batch(ctx, d_org, d_ref, params, jobs, 8)
    assert e.value.code == r.capi.R1C_ERR_ARG
    d_jobs = torch.from_numpy(jobs.view(np.uint8).copy()).cuda()
    res = np.zeros(2 * r.SB_RESULTS, dtype=r.ME_RESULT)
    res["cost"] = U64_MAX
    d_res = torch.from_numpy(res.view(np.uint8).copy()).cuda()
    ctx.check(ctx.lib.r1c_me_sb_batch_dev(ctx.h, d_org.h, d_ref.h, C.byref(params), d_jobs.data_ptr(), 2, 8, d_res.data_ptr()))
    ctx.sync()
    out = d_res.cpu().numpy().view(r.ME_RESULT).reshape(2, r.SB_RESULTS)
    for b in ("BLOCK_8X8", "BLOCK_64X64"):
        off, cnt = r.sb_result_offset(r.BLOCK[b]), ctx.lib.r1c_sb_result_count(r.BLOCK[b])
        bad = out[0, off:off + cnt]
        # -100 px is left of every block's range (>= -(16 + 64) - po_x), so is the whole window -108 .. -92
        assert np.all(bad["cost"] == U64_MAX) and np.all(bad["mv"]["row"] == 0) and np.all(bad["mv"]["col"] == 0), b
        good = out[1, off:off + cnt]
        assert np.all(good["cost"] != U64_MAX), b
    # the reference's answer for that centre: the clamped one-column window, through r1c_full_search_batch
    bw = 8
    win = full_pixel_window(oracle, W // 4, H // 4, (0, 0), bw, bw, (0, -100 * 8), 8, 8)[:4]
    fj = np.zeros(1, dtype=r.FS_JOB)
    fj[0]["po_x"], fj[0]["po_y"], fj[0]["bsize"] = 0, 0, r.BLOCK["BLOCK_8X8"]
    fj[0]["x_lo"], fj[0]["x_hi"], fj[0]["y_lo"], fj[0]["y_hi"] = win
    fj[0]["step"], fj[0]["lambda_"] = 1, 40
    assert win[0] == win[1] == -24
    got = r.full_search_batch(ctx, d_org, d_ref, fj, 8)
    want = oracle_full_search(oracle, o_org, o_ref, fj[0])
    assert want is not None and ((int(got[0]["mv"]["row"]), int(got[0]["mv"]["col"])), int(got[0]["cost"])) == want
    assert int(got[0]["mv"]["col"]) == -24 * 8


def test_mv_field_dev_forms_skip_jobs_outside_the_field(r, ctx):
    import torch
    cols, rows = 16, 12
    jobs = np.array([(2, 3, 3), (-1, 0, 3), (0, -2, 3), (cols, 0, 3), (0, rows, 3), (4, 4, 99), (5, 5, 0)], dtype=r.BLOCK_JOB)
    mvs = np.zeros(jobs.size, dtype=r.MV)
    mvs["row"], mvs["col"] = np.arange(1, jobs.size + 1) * 8, -np.arange(1, jobs.size + 1) * 8
    field = np.zeros((rows, cols), dtype=r.MV)
    with pytest.raises(r.Rav1eCudaError) as e:
        r.save_block_motion(ctx, field.copy(), jobs, mvs)
    assert e.value.code == r.capi.R1C_ERR_ARG
    d_field = torch.zeros(rows * cols * 4, dtype=torch.uint8, device="cuda")
    d_jobs = torch.from_numpy(jobs.view(np.uint8).copy()).cuda()
    d_mvs = torch.from_numpy(mvs.view(np.uint8).copy()).cuda()
    ctx.check(ctx.lib.r1c_save_block_motion_batch_dev(ctx.h, d_field.data_ptr(), cols, rows, d_jobs.data_ptr(), d_mvs.data_ptr(), jobs.size))
    ctx.sync()
    got = d_field.cpu().numpy().view(r.MV).reshape(rows, cols)
    want = np.zeros((rows, cols), dtype=r.MV)
    want[3:5, 2:4] = mvs[0]  # 8x8 = 2 x 2 cells
    want[5:6, 5:6] = mvs[6]  # 4x4
    assert np.array_equal(got, want)
    # predictors: the jobs outside the field get count 0 and zero predictors, the others their lists
    pj = jobs[[0, 1, 2, 3, 4, 6]]
    cm = np.zeros((pj.size, 7), dtype=r.MV)
    ncm = np.zeros(pj.size, dtype=np.uint8)
    with pytest.raises(r.Rav1eCudaError) as e:
        r.get_subset_predictors(ctx, want, None, pj, cm, ncm)
    assert e.value.code == r.capi.R1C_ERR_ARG
    d_pj = torch.from_numpy(pj.view(np.uint8).copy()).cuda()
    d_cm = torch.zeros(pj.size * 7 * 4, dtype=torch.uint8, device="cuda")
    d_ncm = torch.zeros(pj.size, dtype=torch.uint8, device="cuda")
    d_out = torch.full((pj.size * 17 * 4,), 0x55, dtype=torch.uint8, device="cuda")
    d_cnt = torch.full((pj.size,), -7, dtype=torch.int32, device="cuda")
    ctx.check(ctx.lib.r1c_subset_predictors_batch_dev(ctx.h, d_field.data_ptr(), None, cols, rows, d_pj.data_ptr(), pj.size,
                                                      d_cm.data_ptr(), d_ncm.data_ptr(), d_out.data_ptr(), d_cnt.data_ptr()))
    ctx.sync()
    cnt = d_cnt.cpu().numpy()
    out = d_out.cpu().numpy().view(r.MV).reshape(pj.size, 17)
    assert list(cnt[1:5]) == [0, 0, 0, 0] and np.all(out[1:5]["row"] == 0) and np.all(out[1:5]["col"] == 0)
    assert cnt[0] >= 1 and cnt[5] >= 1
    ok, okc = r.get_subset_predictors(ctx, want, None, pj[[0, 5]], cm[[0, 5]], ncm[[0, 5]])
    assert np.array_equal(out[[0, 5]], ok) and list(cnt[[0, 5]]) == list(okc)
