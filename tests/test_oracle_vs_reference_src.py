"""Dev-container-only cross-checks of the oracle's constant tables against the
reference SOURCE TEXT (skipped wherever /root/reference is absent, e.g. on the
GPU box).  They guard against transcription slips in the rotation constants
(src/transform/forward.rs) and the sub-pel filter bank (src/mc.rs:70-179)."""
import collections
import os
import re

import pytest

REF = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present")


def _strip_comments(text):
    return "\n".join(l.split("//")[0] for l in text.splitlines())


def test_rotation_constants_multiset():
    rust = _strip_comments(open(os.path.join(REF, "transform/forward.rs")).read())
    rust = rust[rust.index("fn daala_fdct_ii_2_asym"):rust.index("fn fidentity4")]
    # every (mult, shift) literal pair handed to tx_mul via a kernel call
    ref_pairs = collections.Counter(
        (int(a), int(b)) for a, b in re.findall(r"\((\d+),\s*(\d+)\)", rust))
    for i, j in enumerate([0, 8, 4, 12, 2, 10, 6, 14, 1, 9, 5, 13, 3, 11, 7, 15]):
        ref_pairs[(i, j)] -= 1  # the reorder_4(i, j) calls of daala_fdct64 are not constants
    ref_pairs = +ref_pairs
    c = open(os.path.join(HERE, "..", "oracle", "r1o_tx.c")).read()
    c = re.sub(r"/\*.*?\*/", "", c, flags=re.S)
    c = c[c.index("fdct_ii_2_asym"):c.index("1-D dispatch") if "1-D dispatch" in c else len(c)]
    mine = collections.Counter()
    for m in re.finditer(r"\b(rot_half|rot_neg|rot_pi4|rot|mul)\(([^;]*?)\);", c):
        nums = [int(x) for x in re.findall(r"(?<![\w\[])(\d+)(?![\w\]])", m.group(2))]
        if m.group(1) == "rot_neg":
            nums = nums[1:]  # leading avg flag
        if m.group(1) == "mul":
            nums = nums[-2:]
        assert len(nums) % 2 == 0, m.group(0)
        for k in range(0, len(nums), 2):
            mine[(nums[k], nums[k + 1])] += 1
    tab = re.search(r"ST0_32\[16\]\[6\] = \{(.*?)\};", c, flags=re.S).group(1)
    nums = [int(x) for x in re.findall(r"\d+", tab)]
    for k in range(0, len(nums), 2):
        mine[(nums[k], nums[k + 1])] += 1
    assert mine == ref_pairs, (mine - ref_pairs, ref_pairs - mine)


def test_subpel_filter_bank():
    rust = _strip_comments(open(os.path.join(REF, "mc.rs")).read())
    body = rust[rust.index("const SUBPEL_FILTERS"):rust.index("pub(crate) mod native")]
    body = body[body.index("= ["):]
    ref = [int(x) for x in re.findall(r"-?\d+", body)]
    assert len(ref) == 6 * 16 * 8
    c = open(os.path.join(HERE, "..", "oracle", "r1o_mc.c")).read()
    tab = re.search(r"FILT\[6\]\[16\]\[8\] = \{(.*?)\n\};", c, flags=re.S).group(1)
    tab = re.sub(r"/\*.*?\*/", "", tab, flags=re.S)
    mine = [int(x) for x in re.findall(r"-?\d+", tab)]
    assert mine == ref


# ---- inverse transforms: the reference's own 1-D functions (transliterated at test time) -------------
def _itx():
    import sys
    sys.path.insert(0, HERE)
    import ref_inv_tx
    sys.path.insert(0, os.path.join(HERE, ".."))
    from oracle import r1o
    return ref_inv_tx, r1o


@pytest.mark.parametrize("kind,t1d,sizes", [("dct", 0, (4, 8, 16, 32, 64)), ("adst", 1, (4, 8, 16)),
                                            ("flipadst", 2, (4, 8, 16)), ("identity", 3, (4, 8, 16, 32))])
def test_inverse_1d_matches_reference_text(kind, t1d, sizes):
    import numpy as np
    R, r1o = _itx()
    rng = np.random.default_rng(7 + t1d)
    for n in sizes:
        f = R.fn1d(kind, n)
        for trial in range(60):
            rbits = (16, 18, 20)[trial % 3]
            mag = (1 << (rbits - 1)) - 1 if trial % 5 == 0 else 1 << (4 + trial % 12)
            if (n == 4 and t1d in (1, 2)) or (t1d == 3 and n in (4, 16)):
                mag = min(mag, 1 << 16)  # plain i32 products, no clamps: beyond this they overflow
            x = rng.integers(-mag, mag + 1, n)
            if trial % 7 == 3:
                x[rng.integers(0, n):] = 0  # the usual shape: zeros after the eob
            ref = [0] * 64
            f([int(v) for v in x] + [0] * (64 - n), ref, rbits)
            got = r1o.inv_txfm1d(t1d, x, rbits)
            assert got.tolist() == ref[:n], (kind, n, trial)


def test_inverse_2d_matches_reference_text():
    import numpy as np
    R, r1o = _itx()
    rng = np.random.default_rng(11)
    for tx_size in range(19):
        w, h = r1o.lib().r1o_tx_width(tx_size), r1o.lib().r1o_tx_height(tx_size)
        for tx_type in range(16):
            if not r1o.inv_txfm_valid(tx_size, tx_type):
                continue
            if w * h > 1024 and tx_type not in (0, 9):
                continue  # keep the pure-Python side in seconds
            bd = (8, 10, 12)[(tx_size + tx_type) % 3]
            c = rng.integers(-(1 << (bd + 3)), 1 << (bd + 3), r1o.coded_area(tx_size))
            c[rng.integers(1, c.size + 1):] = 0
            ref = np.array(R.inv_txfm2d([int(v) for v in c], w, h, tx_type, bd)).reshape(w, h).T
            got = r1o.inv_txfm2d(c, tx_size, tx_type, bd)
            assert np.array_equal(got, ref), (tx_size, tx_type, bd)


# ---- forward transforms: the reference's own 1-D functions (transliterated at test time) -------------
def _ftx():
    import sys
    sys.path.insert(0, HERE)
    import ref_fwd_tx
    sys.path.insert(0, os.path.join(HERE, ".."))
    from oracle import r1o
    return ref_fwd_tx, r1o


@pytest.mark.parametrize("kind,t1d,sizes", [("dct", 0, (4, 8, 16, 32, 64)), ("adst", 1, (4, 8, 16)),
                                            ("identity", 2, (4, 8, 16, 32))])
def test_forward_1d_matches_reference_text(kind, t1d, sizes):
    """daala_fdct* / daala_fdst* / fidentity* (src/transform/forward.rs:338-1700) run from their own text, against
    the oracle: small, typical and 32-bit-wrapping magnitudes"""
    import numpy as np
    F, r1o = _ftx()
    rng = np.random.default_rng(17 + t1d)
    for n in sizes:
        f = F.fn1d(kind, n)
        for trial in range(40):
            mag = (1 << 30) if trial % 10 == 9 else 1 << (3 + trial % 16)
            x = rng.integers(-mag, mag + 1, n)
            if trial % 7 == 3:
                x[:] = mag if trial % 2 else -mag
            assert r1o.txfm1d(t1d, x).tolist() == f([int(v) for v in x]), (kind, n, trial)


def test_forward_2d_matches_reference_text():
    """every valid (TxSize, TxType) pair through the reference's 1-D text under a restatement of fht
    (forward.rs:1804-1899) with the shift tables parsed from the source (:22-40), against the oracle's fht"""
    import numpy as np
    F, r1o = _ftx()
    rng = np.random.default_rng(23)
    pairs = 0
    for tx_size in range(19):
        w, h = r1o.lib().r1o_tx_width(tx_size), r1o.lib().r1o_tx_height(tx_size)
        for tx_type in range(16):
            if not r1o.lib().r1o_fht_valid(tx_size, tx_type):
                continue
            pairs += 1
            if w * h > 1024 and tx_type != 0:
                continue  # keep the pure-Python side in seconds
            bd = (8, 10, 12)[(tx_size + tx_type) % 3]
            res = rng.integers(-(1 << bd) + 1, 1 << bd, (h, w)).astype(np.int16)
            ref = F.fht(res.reshape(-1), w, h, tx_type, bd)
            assert r1o.fht(res, tx_size, tx_type, bd).reshape(-1).tolist() == ref, (tx_size, tx_type, bd)
    assert pairs == 193


def test_inverse_clamp_shortcut_bound_holds_on_the_reference_text():
    """itx1d.cuh runs a 1-D inverse transform without its clamps when sum |input| + 128 <= range.hi.  The two facts that
    make the clamps no-ops then are measured here on the reference's OWN functions (transliterated from
    src/transform/inverse.rs by ref_inv_tx): (1) every value handed to clamp_value, and every output, is a linear
    combination of the inputs with weights of magnitude <= 1; (2) a rounding error injected at any half_btf reaches
    every later clamp site / output with gain <= 1, and there are at most 162 of them (drift <= 81 < 128)."""
    import ref_inv_tx as R
    ns = R._transliterate(open(R.REF).read())
    A = 1 << 20
    st = {"j": -1, "cnt": 0, "m": 0}
    orig = R.half_btf

    def hb(w0, in0, w1, in1, bit):
        v = orig(w0, in0, w1, in1, bit)
        if st["cnt"] == st["j"]:
            v += A
        st["cnt"] += 1
        return v

    def cv(v, bit):
        st["m"] = max(st["m"], abs(v))
        return v
    ns["half_btf"], ns["clamp_value"] = hb, cv
    names = sorted(k for k in ns if re.fullmatch(r"av1_i(dct|adst)\d+", k))
    assert len(names) == 8
    for name in names:
        n = int(re.search(r"(\d+)$", name).group(1))
        gain = 0.0
        for k in range(n):
            for sg in (1, -1):
                inp, out = [0] * n, [0] * n
                inp[k] = sg * A
                st.update(j=-1, cnt=0, m=0)
                ns[name](inp, out, 32)
                gain = max(gain, st["m"] / A, max(abs(o) for o in out) / A)
        assert gain <= 1.0, (name, gain)
        calls = st["cnt"]
        assert calls <= 162, (name, calls)
        prop = 0.0
        for j in range(calls):
            out = [0] * n
            st.update(j=j, cnt=0, m=0)
            ns[name]([0] * n, out, 32)
            prop = max(prop, st["m"] / A, max(abs(o) for o in out) / A)
        assert prop <= 1.0, (name, prop)
        assert 0.5 * calls * max(prop, 1.0) < 128
