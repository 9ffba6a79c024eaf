"""Pins the CPU oracle against the reference's own known-answer tests.

 * SAD  : 22 golden values, /root/reference src/dist.rs:379-402
 * SATD : 22 golden values, src/dist.rs:437-460
   inputs: setup_planes, src/dist.rs:342-375 (two planes 640x480 with pads
   136 / 264 -> different strides; patterns (j+i-xpad_off)&255 and
   (j-i-xpad_off)&255 over the WHOLE allocation; region at (32,40)); the
   reference asserts them for u8 and u16 storage (:426-434, :490-498).
 * Plane::pad golden, src/frame/plane.rs:709-752.
"""
import numpy as np
import pytest

# (BlockSize name, w, h, SAD golden, SATD golden) in the order of src/dist.rs; kept as data in tests/golden/
import json
import os

with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_dist_kat.json")) as _f:
    GOLDEN = [tuple(v) for v in json.load(_f)["values"]]
assert len(GOLDEN) == 22


def setup_planes(r1o, px_bytes):
    """src/dist.rs:342-375"""
    inp = r1o.OPlane(640, 480, 0, 0, 128 + 8, 128 + 8, px_bytes)
    rec = r1o.OPlane(640, 480, 0, 0, 2 * 128 + 8, 2 * 128 + 8, px_bytes)
    xpad_off = int(inp.cfg.xorigin - inp.cfg.xpad) - 8
    i, j = np.indices(inp.data.shape)
    inp.data[...] = ((j + i - xpad_off) & 255).astype(inp.dtype)
    i, j = np.indices(rec.data.shape)
    rec.data[...] = ((j - i - xpad_off) & 255).astype(rec.dtype)
    return inp, rec


@pytest.mark.parametrize("px_bytes", [1, 2])
def test_plane_geometry(oracle, px_bytes):
    inp, rec = setup_planes(oracle, px_bytes)
    # u8 xorigin 160/288, u16 144/272
    assert (inp.cfg.xorigin, rec.cfg.xorigin) == ((160, 288) if px_bytes == 1 else (144, 272))
    assert inp.cfg.stride != rec.cfg.stride


@pytest.mark.parametrize("px_bytes", [1, 2])
def test_sad_golden(oracle, px_bytes):
    inp, rec = setup_planes(oracle, px_bytes)
    for name, w, h, want, _ in GOLDEN:
        assert oracle.sad(inp, 32, 40, rec, 32, 40, w, h) == want, name


@pytest.mark.parametrize("px_bytes", [1, 2])
def test_satd_golden(oracle, px_bytes):
    inp, rec = setup_planes(oracle, px_bytes)
    for name, w, h, _, want in GOLDEN:
        assert oracle.satd(inp, 32, 40, rec, 32, 40, w, h) == want, name


def test_block_size_enum(oracle):
    L = oracle.lib()
    # enum order == GOLDEN order == src/partition.rs:116-140
    for idx, (_, w, h, _, _) in enumerate(GOLDEN):
        assert (L.r1o_block_width(idx), L.r1o_block_height(idx)) == (w, h)


def test_pad_golden(oracle):
    """src/frame/plane.rs:707-752 (test_plane_pad): 4x4 visible area inside an
    8x9 allocation with origin (2,3); pad(4,4) replicates the edges."""
    p = oracle.OPlane.__new__(oracle.OPlane)
    p.cfg = oracle.PlaneCfg(stride=8, alloc_height=9, width=4, height=4, xdec=0, ydec=0, xpad=0,
                            ypad=0, xorigin=2, yorigin=3)
    p.px_bytes, p.dtype = 1, np.uint8
    p.data = np.array([
        0, 0, 0, 0, 0, 0, 0, 0,
        0, 0, 0, 0, 0, 0, 0, 0,
        0, 0, 0, 0, 0, 0, 0, 0,
        0, 0, 1, 2, 3, 4, 0, 0,
        0, 0, 8, 7, 6, 5, 0, 0,
        0, 0, 9, 8, 7, 6, 0, 0,
        0, 0, 2, 3, 4, 5, 0, 0,
        0, 0, 0, 0, 0, 0, 0, 0,
        0, 0, 0, 0, 0, 0, 0, 0], dtype=np.uint8).reshape(9, 8)
    p.pad(4, 4)
    want = np.array([
        1, 1, 1, 2, 3, 4, 4, 4,
        1, 1, 1, 2, 3, 4, 4, 4,
        1, 1, 1, 2, 3, 4, 4, 4,
        1, 1, 1, 2, 3, 4, 4, 4,
        8, 8, 8, 7, 6, 5, 5, 5,
        9, 9, 9, 8, 7, 6, 6, 6,
        2, 2, 2, 3, 4, 5, 5, 5,
        2, 2, 2, 3, 4, 5, 5, 5,
        2, 2, 2, 3, 4, 5, 5, 5], dtype=np.uint8).reshape(9, 8)
    assert np.array_equal(p.data, want)
