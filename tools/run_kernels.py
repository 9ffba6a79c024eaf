"""Runs bench_kernels.run() on its own (MC + forward-TX kernel classes) and prints its JSON;
used for ncu captures of those kernels: python tools/run_kernels.py [reps] [mc]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench_kernels  # noqa: E402
import rav1e_b200 as r  # noqa: E402

ctx = r.Context(0)
ext = torch.cuda.ExternalStream(ctx.stream, device=0)
flush = torch.zeros(512 << 20, dtype=torch.uint8, device="cuda")
print(json.dumps(bench_kernels.run(ctx, r, torch, ext, flush, reps=int(sys.argv[1]) if len(sys.argv) > 1 else 3,
                                 only=sys.argv[2] if len(sys.argv) > 2 else None)))
