"""Times r1c_frame_me_dev on a synthetic frame pair of any size / bit depth (device-resident planes):
    python tools/run_me.py W H BD [range] [reps]
Prints one JSON line with ms per frame, Mpix/s and the per-kernel event times (r1c_profile)."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
import rav1e_b200 as r  # noqa: E402

W, H, BD = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
RANGE = int(sys.argv[4]) if len(sys.argv) > 4 else 64
REPS = int(sys.argv[5]) if len(sys.argv) > 5 else 3
PB = 1 if BD == 8 else 2

cur, ref = bench.bench_kernels.synth_frame_pair(1, W, H, BD)

ctx = r.Context(0)
p_org = r.Plane(ctx, W, H, 0, 0, 88, 88, PB)
p_ref = r.Plane(ctx, W, H, 0, 0, 88, 88, PB)
p_org.upload(bench.padded(p_org.cfg, cur))
p_ref.upload(bench.padded(p_ref.cfg, ref))
params = r.make_frame_me_params(W, H, RANGE, RANGE, 160, refine=1, use_satd=1)
n_res = r.frame_me_count(params)
d_res = torch.empty(n_res * 16, dtype=torch.uint8, device="cuda")
ext = torch.cuda.ExternalStream(ctx.stream, device=0)


def step():
    ctx.check(ctx.lib.r1c_frame_me_dev(ctx.h, p_org.h, p_ref.h, C.byref(params), BD, d_res.data_ptr()))


step()
ctx.sync()
ctx.check(ctx.lib.r1c_profile(ctx.h, 1))
l0 = ctx.launches
evs = []
for _ in range(REPS):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(ext)
    step()
    b.record(ext)
    evs.append((a, b))
torch.cuda.synchronize()
ms = sum(a.elapsed_time(b) for a, b in evs) / REPS
prof = {}
for name, slot in (("full_search_sb", 0), ("subpel_mc_generic", 1), ("me_refine", 2)):
    t, n = C.c_double(), C.c_int64()
    ctx.check(ctx.lib.r1c_profile_read(ctx.h, slot, C.byref(t), C.byref(n)))
    prof[name] = [round(t.value / REPS, 3), n.value // REPS]
print(json.dumps({"frame": "%dx%d" % (W, H), "bit_depth": BD, "range": RANGE, "ms": ms, "Mpix/s": W * H / 1e3 / ms,
                  "launches_per_frame": (ctx.launches - l0) // REPS, "kernels_ms_launches": prof}))
