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


def _split(n, parts):
    base, extra = divmod(n, parts)
    out, lo = [], 0
    for r in range(parts):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def split_tiles(width, height, world):
    """-> list (per rank) of ((sb_row_lo, sb_row_hi), (sb_col_lo, sb_col_hi)): the frame cut into a grid of `world`
    tiles of whole superblocks (AV1 tiles, src/tiling/tiler.rs:87-109), the grid shape that minimises the largest
    tile; rank = tile_row * tile_cols + tile_col.  3840x2160 over 8 ranks: 34 x 60 superblocks as 2 x 4 tiles of
    17 x 15 (superblock rows alone would give 5, 5, 4, 4, ... of 34: 6.8x at best)."""
    rows, cols = (height + 63) // 64, (width + 63) // 64
    best = None
    for tr in range(1, world + 1):
        if world % tr:
            continue
        tc = world // tr
        if tr > rows or tc > cols:
            continue
        area = -(-rows // tr) * -(-cols // tc)
        if best is None or area < best[0]:
            best = (area, tr, tc)
    if best is None:  # more ranks than superblocks along both axes: rows only, some ranks idle
        return [(rc, (0, cols)) for rc in split_sb_rows(height, world)]
    _, tr, tc = best
    return [(rr, cc) for rr in _split(rows, tr) for cc in _split(cols, tc)]


def tile_plane_rows(cfg, sb_row_lo, sb_row_hi, range_y):
    """rows [lo, hi) of a plane's ALLOCATION (r1c_plane_upload_rows_async) that the search of superblock rows
    [sb_row_lo, sb_row_hi) can touch: the tile's rows +- (range_y + 16): the full-pel window, the +-2 px the sub-pel
    steps add, and the 3 + 4 rows of the 8-tap filter"""
    lo = int(cfg.yorigin) + sb_row_lo * 64 - range_y - 16
    hi = int(cfg.yorigin) + min(sb_row_hi * 64, int(cfg.height)) + range_y + 16
    return max(lo, 0), min(hi, int(cfg.alloc_height))


def tile_plane_cols(cfg, sb_col_lo, sb_col_hi, range_x):
    """columns [lo, hi) of a plane's ALLOCATION (r1c_plane_upload_rect_async) that the search of superblock columns
    [sb_col_lo, sb_col_hi) can touch: the tile's columns +- (range_x + 16), like tile_plane_rows; widened to multiples
    of 64 pixels so that every row of the copy starts and ends on a 128-byte line"""
    lo = int(cfg.xorigin) + sb_col_lo * 64 - range_x - 16
    hi = int(cfg.xorigin) + min(sb_col_hi * 64, int(cfg.width)) + range_x + 16
    lo, hi = max(lo, 0) & ~63, min((hi + 63) & ~63, int(cfg.stride))
    return lo, hi


def shard_slots(width, height, bsize_mask, sb_row_lo, sb_row_hi, sb_col_lo=0, sb_col_hi=0):
    """frame_me result slots (include/rav1e_cuda.h layout) whose block origin lies in superblock rows
    [sb_row_lo, sb_row_hi) and columns [sb_col_lo, sb_col_hi) (0, 0 = all columns), in the order BlockSize, block row,
    block column -- the order r1c_frame_me_compact_dev packs them in"""
    lib = capi.load()
    idx, off = [], 0
    for b in range(len(capi.BLOCK_SIZES)):
        bw, bh = lib.r1c_block_width(b), lib.r1c_block_height(b)
        if not (bsize_mask >> b) & 1 or bw > 64 or bh > 64:
            continue
        cols, rows = width // bw, height // bh
        by = np.arange(rows)
        sel = by[((by * bh) // 64 >= sb_row_lo) & ((by * bh) // 64 < sb_row_hi)]
        bx = np.arange(cols)
        if sb_col_hi:
            bx = bx[((bx * bw) // 64 >= sb_col_lo) & ((bx * bw) // 64 < sb_col_hi)]
        idx.append((off + sel[:, None] * cols + bx[None, :]).ravel())
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
