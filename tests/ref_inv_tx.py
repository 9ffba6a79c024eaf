"""Dev-container-only: run the reference's OWN 1-D inverse transforms (src/transform/inverse.rs:35-1544)
by transliterating their Rust text to Python at test time.  Nothing of the reference is stored in this
repository: the text is read from /root/reference, rewritten by the handful of regular expressions below
(the functions are straight-line `let stgN = [..]` / `output[i] = ..` code) and exec'd.  Used by
tests/golden/make_inv_tx_golden.py (which commits only input / output VECTORS) and by
tests/test_oracle_vs_reference_src.py.  The 2-D wrapper below restates NativeInvTxfm2D::inv_txfm2d /
inv_txfm2d_add (inverse.rs:1954-2028, 2061-2107)."""
import os
import re

REF = "/root/reference/src/transform/inverse.rs"

SQRT2, INV_SQRT2, SQRT2_BITS = 5793, 2896, 12  # src/transform/mod.rs:52-54


def available():
    return os.path.isfile(REF)


def _i32(v):
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


def half_btf(w0, in0, w1, in1, bit):  # src/transform/mod.rs:478-488 (wrapping i32)
    r = _i32(_i32(w0 * in0) + _i32(w1 * in1))
    return r if bit == 0 else _i32(r + (1 << (bit - 1))) >> bit


def clamp_value(v, bit):  # src/transform/mod.rs:492-496
    return max(-(1 << (bit - 1)), min((1 << (bit - 1)) - 1, v))


def round_shift(v, bit):  # src/util/mod.rs:241-243
    return (v + ((1 << bit) >> 1)) >> bit


def _transliterate(text):
    ns = {"half_btf": half_btf, "clamp_value": clamp_value, "round_shift": round_shift, "SQRT2": SQRT2}
    for name in ("COSPI_INV", "SINPI_INV"):
        m = re.search(r"static %s: \[i32; \d+\] = \[(.*?)\];" % name, text, flags=re.S)
        ns[name] = [int(x) for x in re.findall(r"-?\d+", m.group(1))]
    ns["INV_COS_BIT"] = int(re.search(r"const INV_COS_BIT: usize = (\d+);", text).group(1))
    end = text.index("pub mod txfm_types")
    funcs = list(re.finditer(r"^(?:pub )?fn (av1_i\w+)\(input: &\[i32\], output: &mut \[i32\], _?range: usize\) \{$",
                             text[:end], flags=re.M))
    for k, m in enumerate(funcs):
        body = text[m.end():funcs[k + 1].start() if k + 1 < len(funcs) else end]
        body = body[:body.rindex("}")]
        body = re.sub(r"#\[inline\(always\)\]", "", body)
        ident = re.search(r"output\[\.\.(\d+)\]\s*\.iter_mut\(\)\s*\.zip\(input\[\.\.\d+\]\.iter\(\)\)\s*"
                          r"\.for_each\(\|\(outp, inp\)\| \*outp = (.*?)\);", body, flags=re.S)
        if ident:  # the four identity transforms: one closure over (outp, inp)
            n, expr = int(ident.group(1)), ident.group(2).replace("*inp", "x")
            src = "def %s(input, output, range):\n  for i in range_(%d):\n    x = input[i]\n    output[i] = %s\n" % (
                m.group(1), n, expr)
        else:
            lines = []
            for line in body.splitlines():
                line = line.split("//")[0].rstrip()
                if not line.strip() or "assert!" in line:
                    continue
                line = re.sub(r"^(\s*)let mut (\w+): \[i32; (\d+)\] = \[0; \d+\];", r"\1\2 = [0] * \3", line)
                line = re.sub(r"^(\s*)let (mut )?", r"\1", line)
                line = re.sub(r"&mut |&", "", line)
                line = re.sub(r"output\[\.\.(\d+)\]\.reverse\(\)", r"output[:\1] = output[:\1][::-1]", line)
                lines.append(line.rstrip(";"))
            # the bodies are indented by two spaces already; continuation lines sit inside brackets
            src = "def %s(input, output, range):\n%s\n" % (m.group(1), "\n".join(lines))
        ns["range_"] = range
        exec(src, ns)
    return ns


_NS = None


def fn1d(kind, n):
    """kind: 'dct' | 'adst' | 'flipadst' | 'identity' -> python callable (input, output, range)"""
    global _NS
    if _NS is None:
        _NS = _transliterate(open(REF).read())
    return _NS["av1_i%s%d" % (kind, n)]


# TxType -> (column 1-D type, row 1-D type), src/transform/mod.rs (tx_2d_types); order = TxType enum
TX_TYPE_1D = [("dct", "dct"), ("adst", "dct"), ("dct", "adst"), ("adst", "adst"),
              ("flipadst", "dct"), ("dct", "flipadst"), ("flipadst", "flipadst"), ("adst", "flipadst"),
              ("flipadst", "adst"), ("identity", "identity"), ("dct", "identity"), ("identity", "dct"),
              ("adst", "identity"), ("identity", "adst"), ("flipadst", "identity"), ("identity", "flipadst")]


def intermediate_shift(w, h):  # impl_inv_txs!, inverse.rs:1656-1665
    return {(4, 4): 0, (4, 8): 0, (8, 4): 0, (16, 16): 2, (16, 64): 2, (64, 16): 2, (64, 64): 2,
            (32, 32): 2, (8, 32): 2, (32, 8): 2}.get((w, h), 1)


def inv_txfm2d(coeffs, w, h, tx_type, bd):
    """inverse.rs:2061-2107; returns the W*H residual, COLUMN-major (t[c*H + r]) like the reference"""
    col_t, row_t = TX_TYPE_1D[tx_type]
    wi, hi = min(w, 32), min(h, 32)
    rect = abs(w.bit_length() - h.bit_length())
    buf = [[0] * w for _ in range(h)]
    rng = bd + 8
    row_fn, col_fn = fn1d(row_t, w), fn1d(col_t, h)
    for r in range(hi):
        tin = [0] * 64
        for c in range(wi):
            v = coeffs[r * wi + c]
            if rect == 1:
                v = round_shift(_i32(v * INV_SQRT2), SQRT2_BITS)
            tin[c] = clamp_value(v, rng)
        out = [0] * 64
        row_fn(tin, out, rng)
        buf[r] = out[:w]
    rng = max(bd + 6, 16)
    sh = intermediate_shift(w, h)
    res = [0] * (w * h)
    for c in range(w):
        tin = [0] * 64
        for r in range(h):
            tin[r] = clamp_value(round_shift(buf[r][c], sh), rng)
        out = [0] * 64
        col_fn(tin, out, rng)
        for r in range(h):
            res[c * h + r] = round_shift(out[r], 4)
    return res


def inv_txfm2d_add(coeffs, dst, stride, w, h, tx_type, bd):
    """inverse.rs:2017-2028: dst[r*stride + c] = clamp(dst + t, 0, 2^bd - 1), in place on a list"""
    t = inv_txfm2d(coeffs, w, h, tx_type, bd)
    for c in range(w):
        for r in range(h):
            dst[r * stride + c] = max(0, min((1 << bd) - 1, dst[r * stride + c] + t[c * h + r]))
