"""Multi-GPU partitioning of the hot path: no collective on the data
path, only a gather of (mv, cost) results to the host / rank 0.

  * by frame  -- lookahead motion search is independent per frame
                 (src/api/internal.rs:557-573): frames round-robin over ranks;
  * by superblock row -- within one frame pair (r1c_frame_me sb_row_lo/hi), the
                 same split the reference makes by tile (src/encoder.rs:2772-2781).

Pure host logic (numpy + torch.distributed); works on gloo (CPU tests) and nccl.
"""
import numpy as np

from . import capi


def split_frames(n_frames, world):
    """-> list (per rank) of the frame indices it searches"""
    return [list(range(r, n_frames, world)) for r in range(world)]


def split_sb_rows(height, world):
    """-> list (per rank) of (sb_row_lo, sb_row_hi); contiguous, balanced, may be empty (lo == hi)"""
    rows = (height + 63) // 64
    base, extra = divmod(rows, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def shard_slots(width, height, bsize_mask, sb_row_lo, sb_row_hi):
    """frame_me result slots (include/rav1e_cuda.h layout) whose block origin lies in
    superblock rows [sb_row_lo, sb_row_hi)"""
    lib = capi.load()
    idx, off = [], 0
    for b in range(len(capi.BLOCK_SIZES)):
        bw, bh = lib.r1c_block_width(b), lib.r1c_block_height(b)
        if not (bsize_mask >> b) & 1 or bw > 64 or bh > 64:
            continue
        cols, rows = width // bw, height // bh
        by = np.arange(rows)
        sel = by[((by * bh) // 64 >= sb_row_lo) & ((by * bh) // 64 < sb_row_hi)]
        idx.append((off + sel[:, None] * cols + np.arange(cols)[None, :]).ravel())
        off += cols * rows
    return np.concatenate(idx).astype(np.int64) if idx else np.zeros(0, dtype=np.int64)


def gather_frame_results(local, slots, total, dist=None, dst=0):
    """Gather of the per-rank shards: every rank passes its result array (numpy
    structured, `total` entries, only `slots` valid) -> rank `dst` returns the merged
    array, the others None.  Without a process group: returns `local`."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    rank, world = dist.get_rank(), dist.get_world_size()
    payload = (np.ascontiguousarray(slots), np.ascontiguousarray(local[slots]))
    gathered = [None] * world if rank == dst else None
    dist.gather_object(payload, gathered, dst=dst)
    if rank != dst:
        return None
    merged = np.zeros(total, dtype=local.dtype)
    for s, vals in gathered:
        merged[s] = vals
    return merged
