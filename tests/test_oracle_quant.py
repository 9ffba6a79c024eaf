"""CPU tests of the quantize / dequantize oracle (oracle/r1o_quant.c): the reference's own unit tests
(src/quantize.rs:160-200) repeated against it, and a second pure-Python restatement of quantize."""
import ctypes as C

import numpy as np
import pytest

from quant_util import quantize_py, scans_for, synth_coeffs

# test_tx_log_scale, src/quantize.rs:176-200 (TxSize enum order)
LOG_TX_SCALE = [0, 0, 0, 1, 2, 0, 0, 0, 0, 1, 1, 2, 2, 0, 0, 0, 0, 1, 1]


def test_divu_pair_reference_unit_test(oracle):
    """test_divu_pair, src/quantize.rs:160-168: x / d == divu_pair(x, divu_gen(d)), d 1..1023, x -1000..999"""
    L = oracle.lib()
    xs = np.arange(-1000, 1000)
    for d in range(1, 1024):
        ab = oracle.divu_gen(d)
        got = np.array([L.r1o_divu_pair(int(x), C.byref(ab)) for x in xs[::3]])
        want = np.trunc(xs[::3] / d).astype(np.int64)  # Rust `/` truncates toward zero
        assert np.array_equal(got, want), d


def test_log_tx_scale_reference_unit_test(oracle):
    L = oracle.lib()
    assert [L.r1o_log_tx_scale(t) for t in range(19)] == LOG_TX_SCALE


@pytest.mark.parametrize("is_intra", [0, 1])
def test_quantize_oracle_matches_python_restatement(oracle, is_intra):
    L = oracle.lib()
    rng = np.random.default_rng(21 + is_intra)
    for ts in (0, 1, 2, 5, 9, 13, 3, 4):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        for dcq, acq in ((8, 8), (37, 41), (256, 300), (1336, 1828)):
            qc = oracle.quant_ctx(ts, is_intra, dcq, acq)
            for kind, scan in zip(("dense", "decay", "sparse", "edge"), scans_for(rng, w, h)):
                n = 3 if w * h > 256 else 6
                co = synth_coeffs(rng, n, w * h, kind, acq)
                q = oracle.quantize_batch(qc, ts, scan, co)
                for i in range(n):
                    want = quantize_py(co[i], [int(v) for v in scan], LOG_TX_SCALE[ts], is_intra, dcq, acq)
                    assert q[i].tolist() == want, (ts, dcq, acq, kind, i)


def test_quantize_dequantize_properties(oracle):
    """reconstruction error stays below one quantizer step; everything after the last coefficient that reaches the
    dead zone is zero; dequantize of zeros is zeros"""
    L = oracle.lib()
    rng = np.random.default_rng(5)
    for ts in range(19):
        w, h = L.r1o_tx_width(ts), L.r1o_tx_height(ts)
        coded = min(w, 32) * min(h, 32)
        lts = LOG_TX_SCALE[ts]
        dcq, acq = 52, 64
        qc = oracle.quant_ctx(ts, 0, dcq, acq)
        scan = scans_for(rng, w, h)[0]
        co = synth_coeffs(rng, 8, w * h, "decay", acq)
        q, r = oracle.quantize_batch(qc, ts, scan, co, want_rcoeffs=True)
        assert q.shape == (8, coded)
        err = np.abs(r.astype(np.int64) - co[:, :coded].astype(np.int64))
        assert err[:, 1:].max() <= (acq >> lts) + 1 and err[:, 0].max() <= (dcq >> lts) + 1
        assert np.array_equal(oracle.dequantize(np.zeros(coded, np.int32), dcq, acq, lts), np.zeros(coded, np.int32))
        neg = oracle.dequantize(np.array([-3, -3], np.int32), dcq, acq, lts)
        assert neg.tolist() == [(-3 * dcq + ((1 << lts) - 1)) >> lts, (-3 * acq + ((1 << lts) - 1)) >> lts]
