"""Writes tests/golden/fwd_tx_r01.npz: input / output vectors of the REFERENCE's forward transforms, made in the
dev container by running the reference's own 1-D functions (their Rust text transliterated at run time by
tests/ref_fwd_tx.py; nothing of that text is stored) under a restatement of its 2-D driver
(src/transform/forward.rs:1804-1899).  Only the vectors are committed.

  python tests/golden/make_fwd_tx_golden.py

One case per valid (tx_size, tx_type) pair, bit depth cycling 8 / 10 / 12:
  x_<size>_<type>_<bd>  residual, (H, W) int16         y_<size>_<type>_<bd>  coefficients, (H, W) int32"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
import ref_fwd_tx as F  # noqa: E402

TXW = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TXH = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]


def valid(tx_size, tx_type):
    def ok(kind, n):
        return kind == "dct" or (kind == "identity" and n <= 32) or (kind in ("adst", "flipadst") and n <= 16)
    col, row = F.TX_TYPE_1D[tx_type]
    return ok(col, TXH[tx_size]) and ok(row, TXW[tx_size])


def main():
    rng = np.random.default_rng(20261106)
    out = {}
    k = 0
    for tx_size in range(19):
        for tx_type in range(16):
            if not valid(tx_size, tx_type):
                continue
            bd = (8, 10, 12)[k % 3]
            w, h = TXW[tx_size], TXH[tx_size]
            lim = (1 << bd) - 1
            if k % 5 == 4:  # saturating checkerboard
                yy, xx = np.indices((h, w))
                x = ((((xx + yy) & 1) * 2 - 1) * lim).astype(np.int16)
            else:
                x = rng.integers(-lim, lim + 1, (h, w)).astype(np.int16)
            k += 1
            y = np.array(F.fht(x.reshape(-1), w, h, tx_type, bd), dtype=np.int32).reshape(h, w)
            out["x_%d_%d_%d" % (tx_size, tx_type, bd)] = x
            out["y_%d_%d_%d" % (tx_size, tx_type, bd)] = y
    np.savez_compressed(os.path.join(HERE, "fwd_tx_r01.npz"), **out)
    print(len(out), "arrays")


if __name__ == "__main__":
    if not F.available():
        sys.exit("needs /root/reference (dev container only)")
    main()
