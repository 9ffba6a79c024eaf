"""one small put / prep batch through r1c_put_8tap_batch / r1c_prep_8tap_batch, checked against the oracle
(debugging aid for the MC kernels: python tools/mc_smoke.py [bit_depth])"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import rav1e_b200 as r  # noqa: E402
from oracle import r1o  # noqa: E402
from gpu_util import twin_plane  # noqa: E402

bd = int(sys.argv[1]) if len(sys.argv) > 1 else 10
pb = 1 if bd == 8 else 2
ctx = r.Context(0)
rng = np.random.default_rng(1)
W, H = 256, 192
o_src, d_src = twin_plane(r, r1o, ctx, W, H, pb, bd, rng, xpad=24, ypad=24)
sizes = [(4, 4), (8, 8), (16, 16), (32, 32), (64, 64), (4, 16), (16, 4), (8, 32), (32, 8), (128, 128), (2, 2), (6, 10), (64, 16), (16, 64)]
jobs = np.zeros(len(sizes) * 6, dtype=r.MC_JOB)
off = 0
for i in range(jobs.size):
    w, h = sizes[i % len(sizes)]
    jobs[i] = (0, 0, int(rng.integers(-8, W - w + 8)), int(rng.integers(-8, H - h + 8)), w, h, int(rng.integers(0, 16)) * (i % 3 != 0),
               int(rng.integers(0, 16)) * (i % 4 != 1), int(rng.integers(0, 4)), int(rng.integers(0, 4)), off)
    off += w * h
got = np.full(off + 3, -12345, np.int16)
r.prep_8tap_batch(ctx, got, d_src, jobs, bd)
want = np.full(off + 3, -12345, np.int16)
L = r1o.lib()
bad = 0
for j in jobs:
    L.r1o_prep_8tap(want[int(j["tmp_off"]):].ctypes.data, o_src.ptr(int(j["src_x"]), int(j["src_y"])), o_src.cfg.stride, int(j["w"]),
                    int(j["h"]), int(j["col_frac"]), int(j["row_frac"]), int(j["mode_x"]), int(j["mode_y"]), bd, pb)
    a, b = int(j["tmp_off"]), int(j["tmp_off"]) + int(j["w"]) * int(j["h"])
    if not np.array_equal(got[a:b], want[a:b]):
        bad += 1
        d = np.nonzero(got[a:b] != want[a:b])[0]
        print("MISMATCH", tuple(j), "first at", d[0], "of", b - a, "n", d.size, got[a:b][d[:4]], want[a:b][d[:4]])
print("prep:", jobs.size, "jobs,", bad, "bad; tail intact:", np.array_equal(got[off:], want[off:]))
ctx.sync()
