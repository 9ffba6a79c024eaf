#!/usr/bin/env python
"""bench.py -- throughput of the rav1e hot path on B200 (one JSON line).

Headline workload (BASELINE.json configs[1]): 1080p 8-bit full-search motion
estimation, +-64 px, every block of all 19 sizes 4x4..64x64, SAD search then the
SATD re-evaluation + telescopic sub-pel tail of MotionEstimation::motion_estimation
(src/me.rs:193-271).  A "step" is one lookahead window of 32 distinct frame pairs
(cur, ref) -> every block's (best_mv, lowest_cost), 3 kernel launches per pair.
Metric: Mpix/s = frame pixels / time.  A step is ONE call of the library's window
entry point (r1c_frame_me_window_dev; r1c_frame_me_window with host buffers for
`e2e`), which sends the pairs to two CUDA streams in turn: the tail of one
frame's kernels is filled by the next frame's, and on the host-buffer path the
copies of one frame overlap the kernels of another.  The same steps as a loop
of per-frame calls on ONE stream are timed first: that pass gives the per-kernel
durations of `roofline` and is reported as `one_stream`.

    python bench.py [--gpus N] [--steps K] [--warmup W]      # the CUDA path
    python bench.py --impl reference ...                     # CPU arm: the oracle
        (C restatement of the reference's scalar functions; the reference itself
        is Rust + NASM and cannot be built in this image) on all host threads.

N > 1 (torchrun): frames are sharded over ranks (lookahead motion search is
independent per frame) -- no collective on the data path; the time
is the max over ranks, value = N windows' pixels / time ("weak").  Every N also
reports the within-frame split (bench_scaling.py): one 4K 10-bit frame pair by
superblock rows ("strong_4k10_me"), MC by tile column and forward TX split evenly.

Only the `cpu_baseline` leg and `--impl reference` touch oracle/; the timed CUDA
path is librav1e_cuda.so through its C ABI (rav1e_b200.capi).
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

import numpy as np

import bench_kernels
import bench_baselines
import bench_scaling

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H, BD, PB = 1920, 1080, 8, 1
RANGE = 64
LAMBDA = 160  # (me_lambda * 256 * 0.5) as u32 at a mid-range quantizer
LUMA_PAD = 88  # src/frame/mod.rs:23,60
WORKLOAD = "me_1080p_8bit_fullsearch_pm64_19sizes_sad_then_satd_subpel"
# A step = the frame pairs of one lookahead window: 32 distinct 1080p pairs (170 MB of planes, more than the 126 MB L2),
# so that the driver's 20 steps time seconds, not milliseconds, of back-to-back work.
FRAMES_PER_STEP = 32
# the same dict in both arms (the driver compares them)
CONFIG = {"workload": WORKLOAD, "frame": "%dx%d" % (W, H), "bit_depth": BD, "range": RANGE, "lambda": LAMBDA,
          "frames_per_step": FRAMES_PER_STEP,
          "l2": "inputs larger than L2 (32 distinct frame pairs = 170 MB per step) and L2 flushed between steps",
          "sharding": "one lookahead window of frame pairs per GPU, no collective"}


def clipped_int_ops():
    """SURVEY 8(d) cfg 1 counts one abs-diff + one add per (pixel, candidate) of the 4x4 base layer; candidates that
    get_mv_range (src/me.rs:64-80) forbids for a 4x4 block at that position are not searched and not counted:
    dx in [max(-R, -(x + 20)), min(R, W - x - 4 + 20)], likewise dy."""
    xs, ys = np.arange(0, W, 4), np.arange(0, H, 4)
    nx = np.minimum(RANGE, W - xs - 4 + 20) - np.maximum(-RANGE, -(xs + 20)) + 1
    ny = np.minimum(RANGE, H - ys - 4 + 20) - np.maximum(-RANGE, -(ys + 20)) + 1
    return 2.0 * 16.0 * float(nx.sum()) * float(ny.sum())


ALG_INT_OPS = clipped_int_ops()


# ----------------------------------------------------------------- synthetic frames --
def synth_frame_pair(seed):
    return bench_kernels.synth_frame_pair(seed, W, H)


padded = bench_kernels.padded


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one me_sb2_kernel launch from the committed `ncu --set full`
    capture (profiles/ncu_me_sb2_r02.json), bytes per launch; None when the summary is missing"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "ncu_me_sb2_r02.json")))
        return int(d["dram__bytes_read.sum"]) + int(d["dram__bytes_write.sum"])
    except Exception:
        return None


# ----------------------------------------------------------------------- clocks --
class ClockSampler(threading.Thread):
    """samples SM clock and throttle reasons of one GPU (NVML) while the timed region runs"""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.sm, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {getattr(nv, n): lab for n, lab in (
            ("nvmlClocksThrottleReasonHwSlowdown", "hw_slowdown"),
            ("nvmlClocksThrottleReasonHwThermalSlowdown", "hw_thermal_slowdown"),
            ("nvmlClocksThrottleReasonSwThermalSlowdown", "sw_thermal_slowdown"),
            ("nvmlClocksThrottleReasonSwPowerCap", "sw_power_cap"),
            ("nvmlClocksThrottleReasonHwPowerBrakeSlowdown", "hw_power_brake"),
            ("nvmlClocksThrottleReasonApplicationsClocksSetting", "app_clocks")) if hasattr(nv, n)}
        while not self.stop_flag:
            try:
                self.sm.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, lab in names.items():
                    if r & bit:
                        self.reasons.add(lab)
            except Exception:
                pass
            time.sleep(0.02)

    def summary(self):
        self.stop_flag = True
        if self.is_alive():
            self.join()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


# ------------------------------------------------------------------ CPU arm (oracle) --
def cpu_sample_slots(rng, params_layout, frac):
    """a stratified sample of frame-layout slots: `frac` of the blocks of every size"""
    idx = []
    for off, cols, rows in params_layout:
        n = cols * rows
        k = max(1, int(round(n * frac)))
        idx.append(off + rng.choice(n, size=k, replace=False))
    return np.sort(np.concatenate(idx)).astype(np.int64)


def oracle_layout(r1o):
    L = r1o.lib()
    off, cols, rows = (C.c_int64 * 22)(), (C.c_int32 * 22)(), (C.c_int32 * 22)()
    mask = sum(1 << b for b in range(22) if b not in (13, 14, 15))
    L.r1o_frame_me_layout(W, H, mask, C.byref(off), C.byref(cols), C.byref(rows))
    return mask, [(off[b], cols[b], rows[b]) for b in range(22) if off[b] >= 0]


def run_cpu(seconds_target, steps=1, warmup=0):
    """times the oracle (all host threads) on a bounded sample of the workload ->
    (Mpix/s, cores, sample description, ms per step)"""
    from oracle import r1o
    r1o.build()
    cur, ref = synth_frame_pair(1)
    o_org = r1o.OPlane(W, H, 0, 0, LUMA_PAD, LUMA_PAD, PB)
    o_ref = r1o.OPlane(W, H, 0, 0, LUMA_PAD, LUMA_PAD, PB)
    o_org.visible[...] = cur
    o_ref.visible[...] = ref
    o_org.pad()
    o_ref.pad()
    mask, layout = oracle_layout(r1o)
    rng = np.random.default_rng(7)
    # all the host cores this process may use (torchrun exports OMP_NUM_THREADS=1: override it)
    avail = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    r1o.lib().r1o_set_num_threads(avail)
    cores = r1o.lib().r1o_num_threads()
    # calibrate on a tiny sample, then size the sample for the time target
    frac = 1.0 / 4000
    for _ in range(2):
        idx = cpu_sample_slots(rng, layout, frac)
        t0 = time.perf_counter()
        r1o.frame_me(o_org, o_ref, RANGE, RANGE, LAMBDA, mask, 1, 1, 0, BD, idx)
        dt = time.perf_counter() - t0
        frac = min(1.0, frac * max(1.0, min(40.0, seconds_target / max(dt, 1e-3))))
    idx = cpu_sample_slots(rng, layout, frac)
    for _ in range(warmup):
        r1o.frame_me(o_org, o_ref, RANGE, RANGE, LAMBDA, mask, 1, 1, 0, BD, idx[: max(19, idx.size // 8)])
    t0 = time.perf_counter()
    for _ in range(steps):
        r1o.frame_me(o_org, o_ref, RANGE, RANGE, LAMBDA, mask, 1, 1, 0, BD, idx)
    dt = (time.perf_counter() - t0) / steps
    # every size covers the frame once, so a fraction f of the blocks of every size is a
    # fraction f of the frame's work
    mpix = frac * W * H / 1e6
    sample = ("%.4f of the blocks of each of the 19 sizes (%d blocks) of one 1080p frame pair, "
              "full search +-64 + SATD/sub-pel tail, OpenMP over blocks" % (frac, idx.size))
    return mpix / dt, cores, sample, dt * 1e3


CPU_NOTE = ("C restatement of the reference's _ref functions, OpenMP over blocks; the u8 SAD inner loop is psadbw "
            "(_mm256_sad_epu8 / _mm_sad_epu8), the instruction the reference's AVX2 asm is built on, SATD / MC / search "
            "control are scalar; the reference's own Rust/NASM build is unavailable on this box (no rustc/cargo/nasm)")


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    v, cores, sample, ms = run_cpu(10.0, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": "Mpix/s", "value": v, "unit": "Mpix/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": dict(CONFIG),
        "cpu_baseline": {"value": v, "unit": "Mpix/s", "cores": cores, "kind": "port", "sample": sample, "note": CPU_NOTE},
        "e2e": {"value": v, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


# -------------------------------------------------------------------- CUDA arm --
def int_peak(ctx):
    """INT roofline denominators measured in this run on this device (r1c_int_peak: dependency-free streams of one
    instruction on every SM), G lane-ops/s"""
    names = (("iadd", 0), ("vabsdiff4_acc", 1), ("dp4a", 2), ("dp2a", 3), ("imad", 4), ("prmt", 5))
    out = {}
    for name, which in names:
        v = C.c_double()
        ctx.check(ctx.lib.r1c_int_peak(ctx.h, which, C.byref(v)))
        out[name] = v.value
    return out["iadd"], "measured in this run (r1c_int_peak: independent IADD3 streams on every SM, G lane-ops/s)", out


def hbm_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the MC / forward-TX kernel lines")
    ap.add_argument("--no-scaling", action="store_true", help="skip the within-frame sharding lines (4K ME by SB row, MC / TX by tile)")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_arm(args)
        return

    import torch
    import torch.distributed as dist

    import rav1e_b200 as r

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (rav1e_b200 has no CPU fallback)")
    torch.cuda.set_device(local)
    stdout_fd = None
    if world > 1:
        # stdout carries exactly one JSON line: while NCCL is in use, file descriptor 1 points at stderr (NCCL prints
        # its version banner to stdout on some boxes); it is put back just before rank 0 prints the line
        sys.stdout.flush()
        stdout_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W_ = max(args.warmup, 3)
    K = args.steps

    ctx = r.Context(local)
    lib = ctx.lib
    ext = torch.cuda.ExternalStream(ctx.stream, device=local)

    # this rank's lookahead window (frame-sharded lookahead: rank i takes window i): 32 distinct frame pairs derived
    # from one synthetic pair by moving both frames together (the block motion between cur and ref is kept)
    cur, ref = synth_frame_pair(1 + 2 * rank)
    F = FRAMES_PER_STEP
    p_org = [r.Plane(ctx, W, H, 0, 0, LUMA_PAD, LUMA_PAD, PB) for _ in range(F)]
    p_ref = [r.Plane(ctx, W, H, 0, 0, LUMA_PAD, LUMA_PAD, PB) for _ in range(F)]
    h_org, h_ref = [], []
    for i in range(F):
        sh = (37 * i % H, 91 * i % W)
        h_org.append(torch.from_numpy(padded(p_org[i].cfg, np.roll(cur, sh, axis=(0, 1)))).pin_memory())
        h_ref.append(torch.from_numpy(padded(p_ref[i].cfg, np.roll(ref, sh, axis=(0, 1)))).pin_memory())
        p_org[i].upload(h_org[i].numpy())
        p_ref[i].upload(h_ref[i].numpy())
    params = r.make_frame_me_params(W, H, RANGE, RANGE, LAMBDA, refine=1, use_satd=1)
    n_res = r.frame_me_count(params)
    d_res = torch.empty(n_res * 16, dtype=torch.uint8, device="cuda")
    h_res = torch.zeros(n_res * 16, dtype=torch.uint8).pin_memory().numpy().view(r.ME_RESULT)  # page-locked: DMA target
    flush = torch.zeros(512 << 20, dtype=torch.uint8, device="cuda")  # each half > 126 MB L2

    def step_dev():
        for i in range(F):
            ctx.check(lib.r1c_frame_me_dev(ctx.h, p_org[i].h, p_ref[i].h, C.byref(params), BD, d_res.data_ptr()))

    def step_e2e():
        for i in range(F):
            ctx.check(lib.r1c_plane_upload_async(ctx.h, p_org[i].h, h_org[i].data_ptr()))
            ctx.check(lib.r1c_plane_upload_async(ctx.h, p_ref[i].h, h_ref[i].data_ptr()))
            ctx.check(lib.r1c_frame_me(ctx.h, p_org[i].h, p_ref[i].h, C.byref(params), BD, h_res.ctypes.data))

    # the same step through the library's window entry points: ONE call per step hands the 32 pairs over, the library
    # sends them to two streams in turn (its own second context inside): the tail of one frame's kernels is filled by
    # the next frame's, and with host buffers the copies of one frame overlap the kernels of another
    d_ress = [d_res, torch.empty(n_res * 16, dtype=torch.uint8, device="cuda")]
    h_ress = [h_res, torch.zeros(n_res * 16, dtype=torch.uint8).pin_memory().numpy().view(r.ME_RESULT)]
    win_dev = (r.capi.FramePair * F)()
    win_e2e = (r.capi.FramePair * F)()
    for i in range(F):
        win_dev[i].org, win_dev[i].ref, win_dev[i].results = p_org[i].h, p_ref[i].h, d_ress[i % 2].data_ptr()
        win_e2e[i].org, win_e2e[i].ref, win_e2e[i].results = p_org[i].h, p_ref[i].h, h_ress[i % 2].ctypes.data
        win_e2e[i].org_host, win_e2e[i].ref_host = h_org[i].data_ptr(), h_ref[i].data_ptr()

    def step_dev_win():
        ctx.check(lib.r1c_frame_me_window_dev(ctx.h, win_dev, F, C.byref(params), BD))

    def step_e2e_win():
        ctx.check(lib.r1c_frame_me_window(ctx.h, win_e2e, F, C.byref(params), BD))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        """K steps, each bracketed by events on the library's stream, L2 flushed (untimed)
        between steps -> total ms (max over ranks)"""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(k)]
        barrier()
        for a, b in evs:
            with torch.cuda.stream(ext):
                bench_kernels.flush_l2(torch, flush)
                a.record(ext)
            fn()
            b.record(ext)
        barrier()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(W_):
        step_dev()
    ctx.sync()
    ipeak, ipeak_how, ipk = int_peak(ctx)  # a few ms of integer streams: also warms the clocks
    # ---- device-resident throughput + the dominant kernel's duration --------------
    ctx.check(lib.r1c_profile(ctx.h, 1))
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launches
    ms_serial = timed(step_dev, K)  # one stream, kernels one after the other: the per-kernel durations of the roofline
    launches_serial = ctx.launches - l0
    prof = {}
    for name, slot in (("me_sb2_kernel", 0), ("subpel_mc_generic", 1), ("me_refine_kernel", 2)):
        ms, n = C.c_double(), C.c_int64()
        ctx.check(lib.r1c_profile_read(ctx.h, slot, C.byref(ms), C.byref(n)))
        prof[name] = (ms.value, n.value)
    ctx.check(lib.r1c_profile(ctx.h, 0))
    # ---- the same K steps with alternate frames on two contexts (two streams): the tail of one frame's kernels is
    # filled by the other frame's -- the throughput a caller with a lookahead window gets, and the headline `value`
    step_dev_win()
    ctx.sync()
    l0 = ctx.launches  # r1c_launch_count: both streams of the context
    ms_total = timed(step_dev_win, K)
    launches = ctx.launches - l0
    # ---- end to end: host planes in, host results out, every step ------------------
    step_e2e()
    ms_e2e_serial = timed(step_e2e, K)
    step_e2e_win()
    l1 = ctx.launches
    ms_e2e = timed(step_e2e_win, K)
    launches_e2e = ctx.launches - l1
    clocks = sampler.summary()
    del d_ress[1:], h_ress[1:]

    mpix = F * W * H / 1e6
    value = world * mpix / (ms_total / K / 1e3)
    e2e = world * mpix / (ms_e2e / K / 1e3)
    h2d = int(sum(t.numel() for t in h_org) + sum(t.numel() for t in h_ref))
    d2h = int(F * n_res * 16)
    # ---- within-frame sharding (every N, N = 1 is the base point): 4K 10-bit ME by superblock rows, MC / TX by tile ---
    extra = {}
    if not args.no_scaling:
        p_org.clear()  # the planes of the headline workload are no longer needed
        p_ref.clear()
        try:
            extra["strong_4k10_me"] = bench_scaling.me_strong(ctx, r, torch, dist, ext, flush, rank, world, steps=max(5, min(K, 20)))
            extra.update(bench_scaling.mc_tx_sharded(ctx, r, torch, dist, ext, flush, rank, world))
        except Exception as e:  # the headline line must still print
            extra["scaling_error"] = repr(e)

    if rank == 0:
        hpeak, hpeak_how = hbm_peak()
        sb_ms = prof["me_sb2_kernel"][0] / max(1, prof["me_sb2_kernel"][1])
        n_sb = ((W + 63) // 64) * ((H + 63) // 64)
        win = (64 + 2 * RANGE) ** 2 + 64 * 64
        alg_bytes = n_sb * (win * PB + 849 * 16)  # window + org region per superblock, results out
        achieved = ALG_INT_OPS / (sb_ms / 1e3) / 1e9  # G int-ops/s
        roof = {"bound": "int32_alu", "kernel": "me_sb2_kernel<48> (superblock full search, every size from one 4x4 SAD layer)", "achieved": achieved, "peak": ipeak,
                "unit": "Gop/s", "frac": achieved / ipeak, "traffic": ncu_traffic(),
                "peak_source": ipeak_how,
                "alg_ops_per_launch": ALG_INT_OPS, "ms_per_launch": sb_ms,
                "share_of_step": prof["me_sb2_kernel"][0] / ms_serial,
                "hbm": {"achieved": alg_bytes / (sb_ms / 1e3) / 1e9, "peak": hpeak, "unit": "GB/s",
                        "frac": alg_bytes / (sb_ms / 1e3) / 1e9 / hpeak, "alg_bytes_per_launch": alg_bytes,
                        "peak_source": hpeak_how},
                "packed_simd": {"vabsdiff4_peak_Ginstr_s": ipk.get("vabsdiff4_acc"),
                                "note": "1 VABSDIFF4.ACC = 8 byte-ops; packed peak = 8 x this"},
                "kernels_ms_per_frame": {k: v[0] / (K * F) for k, v in prof.items()}}
        # second kernel of the step: fused sub-pel tail.  Integer work per pixel of every block: SATD re-evaluation
        # (1 diff + 6 Hadamard + 2 abs/add = 9) + the 8 full-pel neighbours of the first step (copies: 9 each) + 16
        # sub-pel candidates x (8 + 8 filter MACs + 9) = 481 lane-ops
        rf_ms = prof["me_refine_kernel"][0] / max(1, prof["me_refine_kernel"][1])
        rf_ops = 19 * W * H * (9 + 8 * 9 + 16 * 25)
        if rf_ms > 0:
            roof["me_refine_kernel"] = {"bound": "int32_alu", "achieved": rf_ops / (rf_ms / 1e3) / 1e9, "peak": ipeak,
                                        "unit": "Gop/s", "frac": rf_ops / (rf_ms / 1e3) / 1e9 / ipeak,
                                        "alg_ops_per_launch": rf_ops, "ms_per_launch": rf_ms,
                                        "share_of_step": prof["me_refine_kernel"][0] / ms_serial,
                                        "note": "MACs counted as one op each; the kernel issues them as DP4A (4 MACs)"}
        line = {
            "metric": "Mpix/s", "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": K, "warmup": W_,
            "ms_per_step": ms_total / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": dict(CONFIG),
            "one_stream": {"value": world * mpix / (ms_serial / K / 1e3), "ms_per_step": ms_serial / K, "gpu_launches": int(launches_serial),
                           "note": "the same K steps on ONE context / stream, kernels strictly one after the other: the pass the "
                                   "roofline's per-kernel durations and share_of_step are measured in; `value` is one "
                                   "r1c_frame_me_window_dev call per step (two streams inside the library), which fills the tail of "
                                   "each kernel"},
            "setup": {"blocks_per_frame": int(n_res), "frames_per_step": F, "launches_per_frame": 3, "streams": 2,
                      "flush": "256 MiB written, then 256 MiB read back, untimed, before every step"},
            "e2e": {"value": e2e, "unit": "Mpix/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / K, "gpu_launches": int(launches_e2e),
                    "path": "one r1c_frame_me_window call per step: per frame pair 2 plane uploads from page-locked host memory, "
                            "the three kernels, the results copied into a page-locked host array; the library sends the pairs "
                            "to two streams in turn, so the copies of one frame overlap the kernels of another",
                    "one_context": {"value": world * mpix / (ms_e2e_serial / K / 1e3), "ms_per_step": ms_e2e_serial / K,
                                    "note": "r1c_plane_upload_async x2 + r1c_frame_me per pair in a loop: copies and kernels serialised on one stream"}},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roof,
        }
        if world == 1 and not args.no_cpu_baseline:
            v, cores, sample, _ = run_cpu(12.0)
            line["cpu_baseline"] = {"value": v, "unit": "Mpix/s", "cores": cores, "kind": "port", "sample": sample, "note": CPU_NOTE}
        if world == 1 and not args.no_extras:
            try:
                line["kernels"] = bench_kernels.run(ctx, r, torch, ext, flush)
            except Exception as e:  # the headline line must still print
                line["kernels"] = {"error": repr(e)}
            if "error" not in line["kernels"] and not args.no_cpu_baseline:
                try:  # CPU baselines + host-buffer e2e of the MC / TX / SAD-SATD classes, configs[0]
                    bench_baselines.run(ctx, r, torch, ext, flush, line["kernels"])
                except Exception as e:
                    line["kernels"]["baselines_error"] = repr(e)
        line.update(extra)
        if stdout_fd is not None:
            sys.stdout.flush()
            os.dup2(stdout_fd, 1)
        print(json.dumps(line))
        sys.stdout.flush()
        if stdout_fd is not None:
            os.dup2(2, 1)  # NCCL's teardown messages, if any, go to stderr again
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
