"""Helpers of the quantize tests: scan orders and a second, pure-Python restatement of
QuantizationContext::quantize (src/quantize.rs:257-316) used to cross-check the C oracle."""
import numpy as np


def zigzag_scan(w, h):
    """anti-diagonal scan of a h x w block (the shape of AV1's default scan): position = r * w + c"""
    order = []
    for s in range(w + h - 1):
        diag = [(r, s - r) for r in range(h) if 0 <= s - r < w]
        if s & 1:
            diag.reverse()
        order += [r * w + c for r, c in diag]
    return np.array(order, dtype=np.uint16)


def row_scan(w, h):
    return np.arange(w * h, dtype=np.uint16)


def col_scan(w, h):
    return np.arange(w * h, dtype=np.uint16).reshape(h, w).T.reshape(-1).copy()


def random_scan(rng, n):
    s = np.concatenate([[0], 1 + rng.permutation(n - 1)]).astype(np.uint16)
    return s


def scans_for(rng, w, h):
    cw, ch = min(w, 32), min(h, 32)
    return [zigzag_scan(cw, ch), row_scan(cw, ch), col_scan(cw, ch), random_scan(rng, cw * ch)]


def divu_gen(d):
    """src/quantize.rs:123-138"""
    m = d.bit_length() - 1
    if d & (d - 1) == 0:
        return (0xFFFFFFFF, 0xFFFFFFFF, m)
    t = (1 << (m + 32)) // d
    r = (t * d + d) & 0xFFFFFFFF
    return (t + 1, 0, m) if r <= (1 << m) else (t, t, m)


def divu_pair(x, d):
    """src/quantize.rs:140-153"""
    a, b, s = d
    y = ((a * abs(x) + b) >> 32) >> s
    return -y if x < 0 else y


def sgn(v):
    return (v > 0) - (v < 0)


def quantize_py(coeffs, scan, lts, is_intra, dc_quant, ac_quant):
    """one block, Python integers (no wrap needed at test magnitudes)"""
    dc_off = dc_quant * (109 if is_intra else 108) // 256
    off0 = ac_quant * (98 if is_intra else 97) // 256
    off1 = ac_quant * (109 if is_intra else 108) // 256
    off_eob = ac_quant * (88 if is_intra else 44) // 256
    dcm, acm = divu_gen(dc_quant), divu_gen(ac_quant)
    deadzone = (ac_quant - off_eob + (1 << lts) - 1) >> lts
    n = len(scan)
    eob = 1
    for k in range(n - 2, -1, -1):
        if abs(int(coeffs[scan[1 + k]])) >= deadzone:
            eob = k + 1
            break
    q = [None] * n
    c0 = int(coeffs[0]) << lts
    q[0] = divu_pair(c0 + sgn(c0) * dc_off, dcm)
    mode = 1
    for pos in scan[1:][:eob]:
        c = int(coeffs[pos]) << lts
        level0 = divu_pair(c, acm)
        off = off1 if level0 > 1 - mode else off0
        qc = divu_pair(c + sgn(c) * off, acm)
        q[pos] = qc
        if mode != 0 and qc == 0:
            mode = 0
        elif qc > 1:
            mode = 1
    for pos in scan[eob + 1:]:
        q[pos] = 0
    return q


def synth_coeffs(rng, n, area, kind, ac_quant):
    """(n, area) int32 transform-like coefficients: 'dense' uniform, 'decay' energy falling with the index (long
    runs under the deadzone), 'sparse' mostly zero, 'edge' values around the rounding thresholds"""
    if kind == "dense":
        return rng.integers(-6 * ac_quant, 6 * ac_quant + 1, size=(n, area)).astype(np.int32)
    if kind == "decay":
        scale = 8.0 * ac_quant / (1.0 + np.arange(area) / 4.0)
        return np.rint(rng.standard_normal((n, area)) * scale).astype(np.int32)
    if kind == "sparse":
        c = rng.integers(-3 * ac_quant, 3 * ac_quant + 1, size=(n, area))
        return (c * (rng.random((n, area)) < 0.08)).astype(np.int32)
    base = rng.integers(-3, 4, size=(n, area)) * ac_quant  # 'edge'
    return (base + rng.integers(-ac_quant // 2 - 2, ac_quant // 2 + 3, size=(n, area))).astype(np.int32)
