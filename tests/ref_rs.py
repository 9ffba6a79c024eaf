"""Dev-container-only: the reference's OWN scalar functions, executed from their Rust text by tests/rustlite.py.

Loads the source files the hot path lives in (read from /root/reference at test time, nothing stored here):
src/util/mod.rs, src/frame/plane.rs, src/tiling/plane_region.rs (its `plane_region_common!` macro expanded
textually for PlaneRegion / PlaneRegionMut), src/partition.rs, src/mc.rs, src/dist.rs, src/me.rs, src/predict.rs,
src/quantize.rs, src/rdo.rs, src/transform/mod.rs, and offers thin Python entry points that build the argument
values (planes with the harness' pixels, MotionVector / PlaneOffset structs ...) and call the reference's functions
BY NAME.  What is executed is the reference's text: get_filter, put_8tap_ref, prep_8tap_ref, mc_avg_ref,
get_mv_range, get_mv_rate, full_search, get_mv_rd_cost / compute_mv_rd_cost, telescopic_subpel_search,
get_best_predictor, diamond_me_search, predict_inter's get_params, Plane::new / pad / downsample_from,
PlaneSlice / PlaneRegion indexing and windows, get_sad_ref / get_satd_ref, QuantizationContext::update / quantize,
dequantize, divu_gen / divu_pair, sse_wxh, cdef_dist_wxh(_8x8), compute_distortion_bias, RawDistortion * f64.
`Interp.calls` counts the calls per function, so a test can assert that the text really ran.

Harness-side stand-ins (the only restated pieces): PlaneData::new (a zeroed array), mem::size_of::<T>(),
FrameInvariants / TileStateMut reduced to the fields these functions read, `get_sad` / `get_satd` / `put_8tap`
dispatchers bound to their `_ref` bodies (the reference picks asm or `_ref` by CPU level; `_ref` is its ground truth).
"""
import os
import re

import numpy as np

import rustlite as R
from rustlite import Arr, Flt, Int, Opt, Struct, NONE

REF = "/root/reference/src"
FILES = ["context.rs", "encoder.rs", "util/mod.rs", "frame/plane.rs", "tiling/plane_region.rs", "tiling/tile.rs", "partition.rs", "transform/mod.rs", "mc.rs",
         "dist.rs", "me.rs", "predict.rs", "scan_order.rs", "quantize.rs", "rdo.rs"]


def available():
    return os.path.isdir(REF)


def _read(rel):
    return open(os.path.join(REF, rel)).read()


_CACHE = {}


def interp(pixel="u8"):
    """one interpreter per pixel type (the generic parameter T of Plane<T> etc.)"""
    if pixel in _CACHE:
        return _CACHE[pixel]
    I = R.Interp()
    I.tyenv["T"] = pixel
    I.tyenv["Self::Output"] = pixel
    for rel in FILES:
        text = _read(rel)
        if rel == "tiling/plane_region.rs":
            text += "\n" + R.expand_macro_rules(text, "plane_region_common", [
                {"name": "PlaneRegion", "as_ptr": "as_ptr"},
                {"name": "PlaneRegionMut", "as_ptr": "as_mut_ptr", "opt_mut": "mut"}])
        I.load(os.path.join(REF, rel), text)
    px_bytes = 1 if pixel == "u8" else 2

    # ---- stand-ins -------------------------------------------------------------------------------------
    I.methods.pop(("PlaneData", "new"), None)
    I.builtin_paths["PlaneData::new"] = lambda it, a: Arr([Int(0, pixel)] * int(R.deref(a[0])), pixel)
    I.builtin_paths["mem::size_of"] = lambda it, a: Int(px_bytes, "usize")
    I.builtin_fns["size_of"] = I.builtin_paths["mem::size_of"]
    # the dispatchers of src/dist.rs / src/mc.rs choose between asm and the `_ref` ground truth by CPU level:
    # bind them to `_ref` (check_asm asserts the asm against exactly these, src/asm/x86/dist.rs:116-117)
    for name, ref in (("get_sad", "get_sad_ref"), ("get_satd", "get_satd_ref"), ("put_8tap", "put_8tap_ref"),
                      ("prep_8tap", "prep_8tap_ref"), ("mc_avg", "mc_avg_ref")):
        I.fns[name] = I.fns[ref]
    _CACHE[pixel] = I
    return I


# ------------------------------------------------------------------------------------------ value helpers ----
def U(v):
    return Int(v, "usize")


def IS(v):
    return Int(v, "isize")


def I32(v):
    return Int(v, "i32")


def mv(row, col):
    return Struct("MotionVector", {"row": Int(row, "i16"), "col": Int(col, "i16")})


def plane_offset(x, y):
    return Struct("PlaneOffset", {"x": IS(x), "y": IS(y)})


def enum(I, name, idx_or_variant):
    for vn, kind, _f, disc in I.enums[name]:
        if vn == idx_or_variant or disc == idx_or_variant:
            return R.EnumVal(name, vn, disc)
    raise KeyError((name, idx_or_variant))


def block_size(I, bsize_index):
    return enum(I, "BlockSize", I.enums["BlockSize"][bsize_index][0])


class RPlane:
    """a reference Plane<T> made by the reference's own Plane::new, filled from / read back to numpy"""

    def __init__(self, I, width, height, xdec=0, ydec=0, xpad=0, ypad=0):
        self.I = I
        self.s = I.call("Plane::new", U(width), U(height), U(xdec), U(ydec), U(xpad), U(ypad))
        c = self.s.f["cfg"].f
        self.stride, self.alloc_height = int(c["stride"]), int(c["alloc_height"])
        self.xorigin, self.yorigin = int(c["xorigin"]), int(c["yorigin"])
        self.width, self.height = width, height

    def cfg(self, name):
        return int(self.s.f["cfg"].f[name])

    def set_all(self, arr2d):
        """whole allocation (alloc_height x stride)"""
        ty = self.I.tyenv["T"]
        flat = np.asarray(arr2d).reshape(-1)
        assert flat.size == self.stride * self.alloc_height
        self.s.f["data"].a[:] = [Int(int(v), ty) for v in flat]

    def set_visible(self, arr2d):
        ty = self.I.tyenv["T"]
        a = self.s.f["data"].a
        for y in range(self.height):
            base = (self.yorigin + y) * self.stride + self.xorigin
            a[base:base + self.width] = [Int(int(v), ty) for v in arr2d[y]]

    def all(self):
        return np.array([x.v for x in self.s.f["data"].a], dtype=np.int64).reshape(self.alloc_height, self.stride)

    def pad(self, w=None, h=None):
        self.I.call("Plane::pad", self.s, U(self.width if w is None else w), U(self.height if h is None else h))

    def slice(self, x, y):
        return self.I.call("Plane::slice", self.s, plane_offset(x, y))

    def region(self, x, y, w=None, h=None):
        I = self.I
        if w is None:
            area = Struct("Area::StartingAt", {"x": IS(x), "y": IS(y)})
            return I.call("Plane::region", self.s, area)
        area = Struct("Area::Rect", {"x": IS(x), "y": IS(y), "width": U(w), "height": U(h)})
        return I.call("Plane::region", self.s, area)

    def region_mut(self, x, y, w=None, h=None):
        I = self.I
        if w is None:
            area = Struct("Area::StartingAt", {"x": IS(x), "y": IS(y)})
        else:
            area = Struct("Area::Rect", {"x": IS(x), "y": IS(y), "width": U(w), "height": U(h)})
        return I.call("Plane::region_mut", self.s, area)


CPU = None  # CpuFeatureLevel argument of the `_ref` functions: unused by them


# ------------------------------------------------------------------------------------------------- MC --------
def put_8tap(I, dst, dst_xy, src, src_xy, w, h, col_frac, row_frac, mode_x, mode_y, bit_depth):
    """put_8tap_ref(dst region at dst_xy, src slice at src_xy, ..)   src/mc.rs:213-306"""
    I.call("put_8tap_ref", dst.region_mut(*dst_xy), src.slice(*src_xy), U(w), U(h), I32(col_frac), I32(row_frac),
           enum(I, "FilterMode", mode_x), enum(I, "FilterMode", mode_y), U(bit_depth), CPU)


def prep_8tap(I, src, src_xy, w, h, col_frac, row_frac, mode_x, mode_y, bit_depth):
    """prep_8tap_ref -> numpy int16 [h, w]   src/mc.rs:308-386"""
    tmp = Arr([Int(0, "i16")] * (w * h), "i16")
    I.call("prep_8tap_ref", tmp, src.slice(*src_xy), U(w), U(h), I32(col_frac), I32(row_frac),
           enum(I, "FilterMode", mode_x), enum(I, "FilterMode", mode_y), U(bit_depth), CPU)
    return np.array([x.v for x in tmp.a], dtype=np.int16).reshape(h, w)


def mc_avg(I, dst, dst_xy, tmp1, tmp2, w, h, bit_depth):
    """mc_avg_ref   src/mc.rs:388-408"""
    t1 = Arr([Int(int(v), "i16") for v in np.asarray(tmp1).reshape(-1)], "i16")
    t2 = Arr([Int(int(v), "i16") for v in np.asarray(tmp2).reshape(-1)], "i16")
    I.call("mc_avg_ref", dst.region_mut(*dst_xy), t1, t2, U(w), U(h), U(bit_depth), CPU)


# ------------------------------------------------------------------------------------------------- ME --------
class Cell:
    """&mut scalar owned by the harness"""

    def __init__(self, v):
        self.v = v
        self.ref = R.RefCell(lambda: self.v, self._set)

    def _set(self, v):
        self.v = v


def frame_invariants(I, bit_depth, allow_hp=False, rec_plane=None, w_in_b=0, h_in_b=0, default_filter=0,
                     block_importances=None, w_in_imp_b=0):
    """the fields of FrameInvariants the hot-path functions read"""
    f = {"sequence": Struct("Sequence", {"bit_depth": U(bit_depth)}), "cpu_feature_level": CPU,
         "allow_high_precision_mv": bool(allow_hp), "w_in_b": U(w_in_b), "h_in_b": U(h_in_b),
         "default_filter": enum(I, "FilterMode", default_filter), "w_in_imp_b": U(w_in_imp_b)}
    if rec_plane is not None:
        rf = Struct("ReferenceFrame", {"frame": Struct("Frame", {"planes": Arr([rec_plane.s])})})
        f["rec_buffer"] = Struct("ReferenceFramesSet", {"frames": Arr([Opt("Some", rf)] * 8)})
        f["ref_frames"] = Arr([Int(0, "u8")] * 7, "u8")
    if block_importances is not None:
        f["block_importances"] = Arr([Flt(float(v), "f32") for v in block_importances])
    return Struct("FrameInvariants", f)


def get_mv_range(I, w_in_b, h_in_b, bo_x, bo_y, blk_w, blk_h):
    """get_mv_range   src/me.rs:64-80"""
    bo = Struct("PlaneBlockOffset", {"0": Struct("BlockOffset", {"x": U(bo_x), "y": U(bo_y)})})
    r = I.call("get_mv_range", U(w_in_b), U(h_in_b), bo, U(blk_w), U(blk_h))
    return tuple(int(x) for x in r)


def get_mv_rate(I, a, b, allow_hp):
    """get_mv_rate   src/me.rs:1005-1021"""
    return int(I.call("get_mv_rate", mv(*a), mv(*b), bool(allow_hp)))


def full_search(I, org, ref, x_lo, x_hi, y_lo, y_hi, bsize, po, step, lambda_, pmv, allow_hp, bit_depth,
                best_mv=(0, 0), lowest_cost=(1 << 64) - 1):
    """full_search   src/me.rs:943-990 -> ((row, col), cost)"""
    fi = frame_invariants(I, bit_depth, allow_hp)
    best = mv(*best_mv)
    cost = Cell(Int(lowest_cost, "u64"))
    I.call("full_search", fi, IS(x_lo), IS(x_hi), IS(y_lo), IS(y_hi), block_size(I, bsize), org.s, ref.s, best, cost.ref,
           plane_offset(*po), U(step), Int(lambda_, "u32"), Arr([mv(*pmv[0]), mv(*pmv[1])]), bool(allow_hp))
    return (int(best.f["row"]), int(best.f["col"])), int(cost.v)


def get_mv_rd_cost(I, org, ref, po, bsize, cand, pmv, lambda_, use_satd, allow_hp, mv_range, bit_depth, subpel):
    """get_mv_rd_cost   src/me.rs:787-856 (sub-pel: through predict_inter + put_8tap_ref into a scratch plane)"""
    fi = frame_invariants(I, bit_depth, allow_hp, rec_plane=ref)
    bs = block_size(I, bsize)
    tmp = NONE
    if subpel:
        w, h = int(I.call("BlockSize::width", bs)), int(I.call("BlockSize::height", bs))
        tmp = Opt("Some", RPlane(I, w, h).s)
    tmp_cell = Cell(tmp)
    r = I.call("get_mv_rd_cost", fi, plane_offset(*po), org.s, ref.s, U(bit_depth), Arr([mv(*pmv[0]), mv(*pmv[1])]),
               Int(lambda_, "u32"), bool(use_satd), IS(mv_range[0]), IS(mv_range[1]), IS(mv_range[2]), IS(mv_range[3]),
               bs, mv(*cand), tmp_cell.ref, enum(I, "RefType", "LAST_FRAME"))
    return int(r)


def telescopic_subpel_search(I, org, ref, po, bsize, lambda_, pmv, mv_range, use_satd, allow_hp, bit_depth, best_mv,
                             lowest_cost):
    """telescopic_subpel_search   src/me.rs:858-941 -> ((row, col), cost)"""
    fi = frame_invariants(I, bit_depth, allow_hp, rec_plane=ref)
    ts = Struct("TileStateMut", {"input": Struct("Frame", {"planes": Arr([org.s])})})
    best = mv(*best_mv)
    cost = Cell(Int(lowest_cost, "u64"))
    I.call("telescopic_subpel_search", fi, ts, plane_offset(*po), Int(lambda_, "u32"), enum(I, "RefType", "LAST_FRAME"),
           Arr([mv(*pmv[0]), mv(*pmv[1])]), IS(mv_range[0]), IS(mv_range[1]), IS(mv_range[2]), IS(mv_range[3]),
           block_size(I, bsize), bool(use_satd), best, cost.ref)
    return (int(best.f["row"]), int(best.f["col"])), int(cost.v)


def diamond_me_search(I, org, ref, po, bsize, predictors, pmv, lambda_, mv_range, use_satd, allow_hp, bit_depth,
                      subpixel):
    """diamond_me_search (+ get_best_predictor)   src/me.rs:655-785 -> ((row, col), cost)"""
    fi = frame_invariants(I, bit_depth, allow_hp, rec_plane=ref)
    center = mv(0, 0)
    cost = Cell(Int((1 << 64) - 1, "u64"))
    I.call("diamond_me_search", fi, plane_offset(*po), org.s, ref.s, Arr([mv(*p) for p in predictors]), U(bit_depth),
           Arr([mv(*pmv[0]), mv(*pmv[1])]), Int(lambda_, "u32"), IS(mv_range[0]), IS(mv_range[1]), IS(mv_range[2]),
           IS(mv_range[3]), block_size(I, bsize), bool(use_satd), center, cost.ref, bool(subpixel),
           enum(I, "RefType", "LAST_FRAME"))
    return (int(center.f["row"]), int(center.f["col"])), int(cost.v)


def get_sad(I, org, oxy, ref, rxy, bsize, bit_depth):
    return int(I.call("get_sad_ref", org.region(*oxy), ref.region(*rxy), block_size(I, bsize), U(bit_depth), CPU))


def get_satd(I, org, oxy, ref, rxy, bsize, bit_depth):
    return int(I.call("get_satd_ref", org.region(*oxy), ref.region(*rxy), block_size(I, bsize), U(bit_depth), CPU))


# ----------------------------------------------------------------------------------------------- quantize ----
def tx_size_enum(I, tx_size):
    return enum(I, "TxSize", I.enums["TxSize"][tx_size][0])


def tx_type_enum(I, tx_type):
    return enum(I, "TxType", I.enums["TxType"][tx_type][0])


def scan_order(I, tx_size, tx_type):
    """av1_scan_orders[tx_size][tx_type].scan   src/scan_order.rs (the reference's own tables)"""
    so = I.const("av1_scan_orders")
    sc = R.seq_get(R.seq_get(so, tx_size), tx_type).f["scan"]
    return np.array([int(x) for x in R.seq_list(sc)], dtype=np.uint16)


def quant_update(I, qindex, tx_size, is_intra, bit_depth):
    """QuantizationContext::update   src/quantize.rs:206-255 -> the context struct (dc_quant / ac_quant inside)"""
    qc = I.default_of("QuantizationContext")
    I.call("QuantizationContext::update", qc, Int(qindex, "u8"), tx_size_enum(I, tx_size), bool(is_intra), U(bit_depth),
           Int(0, "i8"), Int(0, "i8"))
    return qc


def quantize(I, qc, coeffs, tx_size, tx_type):
    """QuantizationContext::quantize   src/quantize.rs:257-316 -> numpy i32 (same length as coeffs)"""
    c = Arr([Int(int(v), "i32") for v in np.asarray(coeffs).reshape(-1)], "i32")
    q = Arr([Int(0, "i32")] * len(c.a), "i32")
    I.call("QuantizationContext::quantize", qc, c, q, tx_size_enum(I, tx_size), tx_type_enum(I, tx_type))
    return np.array([x.v for x in q.a], dtype=np.int32)


def dequantize(I, qindex, qcoeffs, tx_size, bit_depth):
    """dequantize   src/quantize.rs:318-334"""
    q = Arr([Int(int(v), "i32") for v in np.asarray(qcoeffs).reshape(-1)], "i32")
    r = Arr([Int(0, "i32")] * len(q.a), "i32")
    I.call("dequantize", Int(qindex, "u8"), q, r, tx_size_enum(I, tx_size), U(bit_depth), Int(0, "i8"), Int(0, "i8"))
    return np.array([x.v for x in r.a], dtype=np.int32)


def divu_pair(I, x, d):
    return int(I.call("divu_pair", Int(x, "i32"), I.call("divu_gen", Int(d, "u32"))))


# ------------------------------------------------------------------------------------------ RDO distortion ----
def _bias_closure(I, fi, input_region):
    """the closure compute_distortion hands to sse_wxh / cdef_dist_wxh (src/rdo.rs:353-359): calls the
    reference's subregion / frame_block_offset / compute_distortion_bias"""
    def bias(it, args):
        area, bs = args
        sub = it.call("PlaneRegion::subregion", input_region, area)
        return it.call("compute_distortion_bias", fi, it.call("PlaneRegion::frame_block_offset", sub), bs)
    return bias


def _unit_bias(it, args):
    return Flt(1.0, "f64")


def sse_wxh(I, src1, xy1, src2, xy2, w, h, fi=None):
    """sse_wxh   src/rdo.rs:286-335 -> u64 (fi None: bias 1.0)"""
    r1, r2 = src1.region(*xy1), src2.region(*xy2)
    d = I.call("sse_wxh", r1, r2, U(w), U(h), _bias_closure(I, fi, r1) if fi is not None else _unit_bias)
    return int(d.f["0"])


def cdef_dist_wxh(I, src1, xy1, src2, xy2, w, h, bit_depth, fi=None):
    """cdef_dist_wxh (+ cdef_dist_wxh_8x8)   src/rdo.rs:219-283 -> u64"""
    r1, r2 = src1.region(*xy1), src2.region(*xy2)
    d = I.call("cdef_dist_wxh", r1, r2, U(w), U(h), U(bit_depth), _bias_closure(I, fi, r1) if fi is not None else _unit_bias)
    return int(d.f["0"])


# -------------------------------------------------------------------------------------- predict_inter ----
def mv_params(I, rec, po, mvec):
    """predict_inter's nested get_params (src/predict.rs:267-284), reached by running predict_inter itself on a
    1-pixel-high probe is not possible (it is a nested fn), so the harness runs predict_inter for a whole block
    instead: see predict_inter()."""
    raise NotImplementedError


def predict_inter(I, rec, po, w, h, mvec, bit_depth, default_filter=0):
    """PredictionMode::NEWMV.predict_inter, single reference (src/predict.rs:255-310): get_params + clamp + put_8tap
    into a fresh w x h plane -> numpy [h, w]"""
    fi = frame_invariants(I, bit_depth, False, rec_plane=rec, default_filter=default_filter)
    tmp = RPlane(I, w, h)
    tile_rect = Struct("TileRect", {"x": U(0), "y": U(0), "width": U(w), "height": U(h)})
    I.call("PredictionMode::predict_inter", enum(I, "PredictionMode", "NEWMV"), fi, tile_rect, U(0), plane_offset(*po),
           tmp.region_mut(0, 0), U(w), U(h),
           Arr([enum(I, "RefType", "LAST_FRAME"), enum(I, "RefType", "NONE_FRAME")]), Arr([mv(*mvec), mv(0, 0)]))
    return tmp.all()[:h, :w]
