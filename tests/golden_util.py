"""Loaders shared by the tests that consume tests/golden/*_r02.npz (reference-made known answers)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
G = {n: os.path.join(HERE, "golden", "%s_r02.npz" % n) for n in ("mc", "me", "quant", "rdo")}
_cache = {}


def load_golden(name):
    if name not in _cache:
        _cache[name] = dict(np.load(G[name]))
    return _cache[name]


def oplane(oracle, visible, px_bytes, xpad, ypad, xdec=0, ydec=0):
    """oracle plane with Plane::new(w, h, xdec, ydec, xpad, ypad) geometry, `visible` pixels, padded"""
    h, w = visible.shape
    o = oracle.OPlane(w, h, xdec, ydec, xpad, ypad, px_bytes)
    o.visible[...] = visible.astype(o.dtype)
    o.pad()
    return o
