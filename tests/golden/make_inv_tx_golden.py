"""Writes tests/golden/inv_tx_r01.npz: input / output vectors of the REFERENCE's inverse transforms,
made in the dev container by running the reference's own 1-D functions (their Rust text transliterated at
run time by tests/ref_inv_tx.py; nothing of that text is stored) under a restatement of its 2-D wrapper
(src/transform/inverse.rs:2061-2107).  Only the vectors are committed.

  python tests/golden/make_inv_tx_golden.py

One case per valid (tx_size, tx_type) pair, bit depth cycling 8 / 10 / 12:
  c_<size>_<type>_<bd>  coefficients, min(W,32)*min(H,32), row-major
  r_<size>_<type>_<bd>  residual, (H, W)
plus eight inv_txfm2d_add cases (`d_...` = the destination before, `o_...` after)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import ref_inv_tx as R  # noqa: E402

TXW = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TXH = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]


def valid(tx_size, tx_type):
    def ok(kind, n):
        return kind == "dct" or (kind == "identity" and n <= 32) or (kind in ("adst", "flipadst") and n <= 16)
    col, row = R.TX_TYPE_1D[tx_type]
    return ok(col, TXH[tx_size]) and ok(row, TXW[tx_size])


def coeffs_for(rng, tx_size, tx_type, bd):
    w, h = min(TXW[tx_size], 32), min(TXH[tx_size], 32)
    c = rng.integers(-(1 << (bd + 2)), 1 << (bd + 2), w * h)
    c = (c * rng.random(w * h) ** 3).astype(np.int64)  # mostly small, a few large
    eob = int(rng.integers(1, w * h + 1))
    if (tx_size + tx_type) % 4:
        c[eob:] = 0
    c[0] = int(rng.integers(-(1 << (bd + 4)), 1 << (bd + 4)))
    return c.astype(np.int32)


def main():
    rng = np.random.default_rng(20261105)
    out = {}
    k = 0
    for tx_size in range(19):
        for tx_type in range(16):
            if not valid(tx_size, tx_type):
                continue
            bd = (8, 10, 12)[k % 3]
            k += 1
            w, h = TXW[tx_size], TXH[tx_size]
            c = coeffs_for(rng, tx_size, tx_type, bd)
            r = np.array(R.inv_txfm2d([int(v) for v in c], w, h, tx_type, bd), dtype=np.int32).reshape(w, h).T
            out["c_%d_%d_%d" % (tx_size, tx_type, bd)] = c
            out["r_%d_%d_%d" % (tx_size, tx_type, bd)] = np.ascontiguousarray(r)
    for tx_size, tx_type, bd in [(0, 0, 8), (1, 3, 8), (2, 9, 10), (3, 0, 12), (4, 0, 8), (7, 1, 10), (16, 11, 8),
                                 (12, 0, 10)]:
        w, h = TXW[tx_size], TXH[tx_size]
        c = coeffs_for(rng, tx_size, tx_type, bd) * 4
        d = rng.integers(0, 1 << bd, (h, w + 5))
        d[: h // 2] = np.where(rng.random((h // 2, w + 5)) < 0.5, 0, (1 << bd) - 1)  # make both clamps bite
        flat = [int(v) for v in d.reshape(-1)]
        R.inv_txfm2d_add([int(v) for v in c], flat, w + 5, w, h, tx_type, bd)
        dt = np.uint8 if bd == 8 else np.uint16
        out["ac_%d_%d_%d" % (tx_size, tx_type, bd)] = c.astype(np.int32)
        out["d_%d_%d_%d" % (tx_size, tx_type, bd)] = d.astype(dt)
        out["o_%d_%d_%d" % (tx_size, tx_type, bd)] = np.array(flat).reshape(h, w + 5).astype(dt)
    np.savez_compressed(os.path.join(HERE, "inv_tx_r01.npz"), **out)
    print(len(out), "arrays")


if __name__ == "__main__":
    if not R.available():
        sys.exit("needs /root/reference (dev container only)")
    main()
