"""r1c_me_sb_batch_dev_noalloc (the allocation-free superblock search, scratch in a file-scope
`static __device__` array) against r1c_me_sb_batch_dev, through the C ABI, two calls in a row.
Bar: bit-exact, as for every other integer result of the library.  Collected last on purpose;
runs alone in a fresh process or after the other tests."""
import ctypes as C

import numpy as np
import pytest

from gpu_util import twin_plane

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


def _dev(arr):
    import torch
    return torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy()).cuda()


def _search(r, ctx, fn, d_org, d_ref, params, jobs, bd):
    """one call of `fn` (either entry point) with device job / result arrays -> host results"""
    import torch
    res = np.zeros(jobs.size * r.SB_RESULTS, dtype=r.ME_RESULT)
    res["cost"] = 12345  # entries the kernel does not emit keep the caller's values
    d_jobs, d_res = _dev(jobs), _dev(res)
    torch.cuda.synchronize()
    ctx.check(fn(ctx.h, d_org.h, d_ref.h, C.byref(params), d_jobs.data_ptr(), jobs.size, bd, d_res.data_ptr()))
    ctx.sync()
    return d_res.cpu().numpy().view(r.ME_RESULT).reshape(jobs.size, r.SB_RESULTS)


def _case(r, oracle, ctx, W, H, pb, bd, rng_x, rng_y, lam, seed):
    rng = np.random.default_rng(seed)
    o_ref, d_ref = twin_plane(r, oracle, ctx, W, H, pb, bd, rng, smooth=True)
    noisy = np.roll(o_ref.visible, (-5, 9), axis=(0, 1)).astype(np.int64) + rng.integers(-3, 4, size=(H, W))
    _, d_org = twin_plane(r, oracle, ctx, W, H, pb, bd, data=np.clip(noisy, 0, (1 << bd) - 1))
    params = r.make_sb_params(W // 4, H // 4, rng_x, rng_y, 1, lam)
    pos = [(x, y) for y in range(0, H, 64) for x in range(0, W, 64)]
    jobs = np.zeros(len(pos), dtype=r.SB_JOB)
    for i, (x, y) in enumerate(pos):
        jobs[i]["po_x"], jobs[i]["po_y"], jobs[i]["bsize_mask"] = x, y, 0x3FFFFF
        jobs[i]["pmv"][0] = (int(rng.integers(-40, 40)), int(rng.integers(-40, 40)))
        jobs[i]["pmv"][1] = (int(rng.integers(-40, 40)), int(rng.integers(-40, 40)))
    return d_org, d_ref, params, jobs


def test_noalloc_search_equals_allocating_entry_point_twice(r, ctx, oracle):
    lib = ctx.lib
    # call 1: 8-bit planes, +-64 window (2 chunks per superblock), 12 superblocks incl. partial ones
    a = _case(r, oracle, ctx, 256, 168, 1, 8, 64, 64, 160, 901)
    # call 2: 10-bit planes (the scratch then also carries the org sums behind the partial keys),
    # another window and job count, so that what call 1 left in the array lies under different slots
    b = _case(r, oracle, ctx, 192, 136, 2, 10, 20, 12, 90, 902)
    # call 3: 12-bit planes (true SADs in the keys: no org sums in the scratch, more chunks per superblock)
    c = _case(r, oracle, ctx, 192, 136, 2, 12, 24, 24, 2000, 903)
    for d_org, d_ref, params, jobs, bd in (a + (8,), b + (10,), c + (12,), a + (8,)):
        want = _search(r, ctx, lib.r1c_me_sb_batch_dev, d_org, d_ref, params, jobs, bd)
        got = _search(r, ctx, lib.r1c_me_sb_batch_dev_noalloc, d_org, d_ref, params, jobs, bd)
        assert (want["cost"] != 12345).any()
        assert np.array_equal(got["cost"], want["cost"])
        assert np.array_equal(got["mv"]["row"], want["mv"]["row"])
        assert np.array_equal(got["mv"]["col"], want["mv"]["col"])


def test_noalloc_search_rejects_what_it_cannot_hold(r, ctx, oracle):
    lib = ctx.lib
    d_org, d_ref, params, jobs = _case(r, oracle, ctx, 128, 128, 1, 8, 64, 64, 160, 903)
    d_jobs = _dev(jobs)
    # 2 chunks x 849 keys x 8 bytes per superblock: the 8 MiB array holds 617 superblocks at +-64
    per_sb = 2 * r.SB_RESULTS * 8
    too_many = r.capi.R1C_SB_NOALLOC_SCRATCH_BYTES // per_sb + 1
    rc = lib.r1c_me_sb_batch_dev_noalloc(ctx.h, d_org.h, d_ref.h, C.byref(params), d_jobs.data_ptr(), too_many, 8,
                                         d_jobs.data_ptr())
    assert rc == r.capi.R1C_ERR_NOMEM
    assert b"R1C_SB_NOALLOC_SCRATCH_BYTES" in lib.r1c_last_error(ctx.h)
    step2 = r.make_sb_params(32, 32, 16, 16, 2, 0)  # step-2 lattices belong to the first-generation kernel
    rc = lib.r1c_me_sb_batch_dev_noalloc(ctx.h, d_org.h, d_ref.h, C.byref(step2), d_jobs.data_ptr(), jobs.size, 8,
                                         d_jobs.data_ptr())
    assert rc == r.capi.R1C_ERR_UNIMPLEMENTED
