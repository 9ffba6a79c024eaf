"""Dev-container-only: run the reference's OWN 1-D forward transforms (src/transform/forward.rs:100-1700:
`TxOperations for i32`, the rotation traits, the butterflies and every `daala_f*` / `fidentity*` function) by
transliterating their Rust text to Python at test time.  Nothing of the reference is stored in this repository:
the text is read from /root/reference, rewritten by the small statement translator below (the functions are
straight-line `let` / tuple / method-call code over a generic T) and exec'd with T = a wrapping 32-bit integer.
The 2-D driver `fht` restates FwdTxfm2D::fht (forward.rs:1804-1899) the way the oracle does.
Used by tests/test_oracle_vs_reference_src.py and tests/golden/make_fwd_tx_golden.py."""
import os
import re

REF = "/root/reference/src/transform/forward.rs"


def available():
    return os.path.isfile(REF)


def _wrap(v):
    v &= 0xFFFFFFFF
    return v - (1 << 32) if v & 0x80000000 else v


class V(int):
    """i32 with wrapping arithmetic (a release build); the TxOperations methods are attached from the Rust text"""

    def __new__(cls, v=0):
        return int.__new__(cls, _wrap(int(v)))

    def __add__(self, o): return V(int(self) + int(o))
    __radd__ = __add__
    def __sub__(self, o): return V(int(self) - int(o))
    def __rsub__(self, o): return V(int(o) - int(self))
    def __mul__(self, o): return V(int(self) * int(o))
    __rmul__ = __mul__
    def __rshift__(self, o): return V(int(self) >> int(o))
    def __lshift__(self, o): return V(int(self) << int(o))
    def __neg__(self): return V(-int(self))
    def copy_fn(self): return self  # trait default, forward.rs:109-111
    def into(self): return self


class View:
    """&mut out[a..b]"""

    def __init__(self, base, a, b):
        self.base, self.a, self.b = base, a, b

    def __getitem__(self, i): return self.base[self.a + i]
    def __setitem__(self, i, v): self.base[self.a + i] = v
    def __len__(self): return self.b - self.a

    def reverse(self):
        vals = [self.base[self.a + i] for i in range(self.b - self.a)][::-1]
        for i, v in enumerate(vals):
            self.base[self.a + i] = v


OPS = {"Add::add": lambda a, b: a + b, "Sub::sub": lambda a, b: a - b, "T::sub": lambda a, b: a - b,
       "T::add_avg": lambda a, b: a.add_avg(b), "T::sub_avg": lambda a, b: a.sub_avg(b),
       "T::into": lambda a: a, "T::copy_fn": lambda a: a.copy_fn(), "T::rshift1": lambda a: a.rshift1()}


def _strip(text):
    return "\n".join(l.split("//")[0].rstrip() for l in text.splitlines())


def _match_brace(text, i):
    """index just past the brace block opening at text[i] == '{'"""
    d = 0
    for k in range(i, len(text)):
        if text[k] == "{":
            d += 1
        elif text[k] == "}":
            d -= 1
            if d == 0:
                return k + 1
    raise ValueError("unbalanced")


def _expr(e):
    e = re.sub(r"::<[^>]*>", "", e)
    e = re.sub(r"if (.+?) \{ (.+?) \} else \{ (.+?) \}", r"(\2 if \1 else \3)", e)
    e = re.sub(r"&mut (\w+)\[(\d+)\.\.(\d+)\]", r"View(\1, \2, \3)", e)
    e = re.sub(r"(\w+)\[(\d+)\.\.(\d+)\]\.reverse\(\)", r"View(\1, \2, \3).reverse()", e)
    e = re.sub(r"&mut |&", "", e)
    e = re.sub(r"\b([a-z_]\w*)\.(\d)\b", r"\1[\2]", e)  # tuple fields
    e = re.sub(r"Self::", "Self.", e)
    e = re.sub(r"\b([A-Z]\w*)::(\w+)\(", r"\1.\2(", e)    # RotateAdd::kernel( -> RotateAdd.kernel(
    return e


def _statements(body):
    """split a brace-free-at-depth-0 body into statements at ';' (depth 0 of (), [], {})"""
    out, cur, d = [], "", 0
    for ch in body:
        if ch in "([{":
            d += 1
        elif ch in ")]}":
            d -= 1
        if ch == ";" and d == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
            if ch == "}" and d == 0 and cur.strip().startswith("{"):  # a bare block: flatten it
                inner, rest = _statements(cur.strip()[1:-1])
                out.extend(inner + ([rest] if rest else []))
                cur = ""
    tail = cur.strip()
    return [s for s in out if s], tail


def _body(body, indent="  "):
    stmts, tail = _statements(" ".join(body.split()))
    lines = []
    for s in stmts:
        if s.startswith(("assert!", "debug_assert!")):
            continue
        m = re.match(r"let (?:mut )?(\w+) = \|([^|]*)\| \{(.*)\}$", s)
        if m:  # closures of daala_fdct64
            args = ", ".join(a.split(":")[0].strip() for a in m.group(2).split(","))
            lines.append("%sdef %s(%s):" % (indent, m.group(1), args))
            lines.extend(_body(m.group(3), indent + "  "))
            continue
        m = re.match(r"let mut (\w+): \[\(T, T\); (\d+)\] = .*$", s)
        if m:
            lines.append("%s%s = [(V(0), V(0))] * %s" % (indent, m.group(1), m.group(2)))
            continue
        m = re.match(r"let mut (\w+): \[T; (\d+)\] = .*$", s)
        if m:
            lines.append("%s%s = [V(0)] * %s" % (indent, m.group(1), m.group(2)))
            continue
        m = re.match(r"store_coeffs!\(\s*(\w+), (.*?),?\s*\)$", s)
        if m:
            for i, x in enumerate(a.strip() for a in m.group(2).split(",")):
                lines.append("%s%s[%d] = %s" % (indent, m.group(1), i, x))
            continue
        m = re.match(r"output\[\.\.(\d+)\]\.copy_from_slice\(&input\[\.\.\d+\]\)$", s)
        if m:
            lines.append("%sfor i_ in range(%s): output[i_] = input[i_]" % (indent, m.group(1)))
            continue
        s = re.sub(r"^let (mut )?", "", s)
        lines.append(indent + _expr(s))
    if tail:
        lines.append(indent + "return " + _expr(tail))
    return lines


def _params(sig):
    names = []
    d = 0
    cur = ""
    for ch in sig + ",":
        if ch in "([<":
            d += 1
        elif ch in ")]>":
            d -= 1
        if ch == "," and d == 0:
            if cur.strip():
                names.append(cur.strip().split(":")[0].strip())
            cur = ""
        else:
            cur += ch
    return [n for n in names if n]


def _functions(text, ns, indent="", first="", decorator=""):
    src = []
    for m in re.finditer(r"\bfn (\w+)(?:<[^>]*>)?\(", text):
        close = text.index(")", m.end())
        # parameter lists nest parentheses ((T, T)): find the matching one
        d, k = 1, m.end()
        while d:
            d += {"(": 1, ")": -1}.get(text[k], 0)
            k += 1
        close = k - 1
        brace = text.index("{", close)
        if ";" in text[close:brace]:
            continue  # a declaration without body
        end = _match_brace(text, brace)
        params = [p for p in _params(text[m.end():close]) if p != "self"]
        has_self = "self" in text[m.end():close].split(",")[0]
        head = ([first] if first else []) + (["self"] if has_self else []) + params
        if decorator:
            src.append(indent + decorator)
        src.append("%sdef %s(%s):" % (indent, m.group(1), ", ".join(head)))
        src.extend(_body(text[brace + 1:end - 1], indent + "  ") or [indent + "  pass"])
    return src


def _transliterate(text):
    text = _strip(text)
    text = text[text.index("impl TxOperations for i32"):text.index("pub mod txfm_types")]
    ns = {"V": V, "View": View}
    # TxOperations for i32 -> methods of V
    a = text.index("{", text.index("impl TxOperations for i32"))
    b = _match_brace(text, a)
    exec("\n".join(_functions(text[a + 1:b - 1], ns)), ns)
    for name in ("tx_mul", "add_avg", "sub_avg", "rshift1"):
        f = ns[name]
        setattr(V, name, (lambda g: lambda self, *args: V(g(self, *args)))(f))
    # drop the SIMD macro
    a = text.index("macro_rules! impl_tx_ops_simd")
    text = text[:a] + text[text.index("impl_tx_ops_simd!(", a):]
    # traits with default methods -> classes with classmethods; impls -> subclasses with the consts
    for m in list(re.finditer(r"\btrait (\w+)<T: TxOperations> \{", text)):
        end = _match_brace(text, m.end() - 1)
        src = ["class %s:" % m.group(1)] + _functions(text[m.end():end - 1], ns, "  ", "Self", "@classmethod")
        exec("\n".join(src), ns)
    for m in re.finditer(r"impl<T: TxOperations> (\w+)<T> for (\w+) \{(.*?)\}", text, flags=re.S):
        cls = type(m.group(2), (ns[m.group(1)],), {})
        for c in re.finditer(r"const (\w+): [^=]*= ([\w:]+);", m.group(3)):
            setattr(cls, c.group(1), staticmethod(OPS[c.group(2)]))
        ns[m.group(2)] = cls
    # free functions (butterflies, daala_f*, fidentity*)
    a = text.index("fn butterfly_add")
    free = re.sub(r"macro_rules! store_coeffs \{.*?\n\}\n", "", text[a:], flags=re.S)
    exec("\n".join(_functions(free, ns)), ns)
    return ns


_NS = None
_NAMES = {("dct", 4): "daala_fdct4", ("dct", 8): "daala_fdct8", ("dct", 16): "daala_fdct16", ("dct", 32): "daala_fdct32",
          ("dct", 64): "daala_fdct64", ("adst", 4): "daala_fdst_vii_4", ("adst", 8): "daala_fdst8",
          ("adst", 16): "daala_fdst16", ("identity", 4): "fidentity4", ("identity", 8): "fidentity8",
          ("identity", 16): "fidentity16", ("identity", 32): "fidentity32"}  # forward.rs txfm_types, :1702-1769


def fn1d(kind, n):
    """-> callable(list of n ints) -> list of n ints, running the reference's function for (kind, n)"""
    global _NS
    if _NS is None:
        _NS = _transliterate(open(REF).read())
    f = _NS[_NAMES[(kind, n)]]

    def run(vec):
        out = [V(0)] * n
        f([V(v) for v in vec], out)
        return [int(v) for v in out]
    return run


def shift_table(w, h):
    """FWD_SHIFT_<W>X<H>, forward.rs:22-40, parsed from the text"""
    m = re.search(r"const FWD_SHIFT_%dX%d: TxfmShifts = (\[\[.*?\]\]);" % (w, h), open(REF).read())
    return eval(m.group(1))


# TxType -> (column 1-D type, row 1-D type), src/transform/mod.rs:196-213; order = TxType enum
TX_TYPE_1D = [("dct", "dct"), ("adst", "dct"), ("dct", "adst"), ("adst", "adst"),
              ("flipadst", "dct"), ("dct", "flipadst"), ("flipadst", "flipadst"), ("adst", "flipadst"),
              ("flipadst", "adst"), ("identity", "identity"), ("dct", "identity"), ("identity", "dct"),
              ("adst", "identity"), ("identity", "adst"), ("flipadst", "identity"), ("identity", "flipadst")]


def _shift(v, entry):  # round_shift_array(arr, n, -entry), src/transform/mod.rs:499-521
    if entry > 0:
        return _wrap(v << entry)
    if entry < 0:
        return (v + ((1 << -entry) >> 1)) >> -entry
    return v


def fht(residual, w, h, tx_type, bd):
    """FwdTxfm2D::fht, forward.rs:1804-1899, one column / one row at a time (the lane width changes no value);
    residual: h*w ints row-major -> h*w coefficients row-major.  Row-flipped types write column c to w-1-c
    (the intended semantic; the fork's SIMD branch :1858-1862 is unreachable from the encoder)."""
    col_t, row_t = TX_TYPE_1D[tx_type]
    cf, rf = col_t == "flipadst", row_t == "flipadst"
    col_fn = fn1d("adst" if cf else col_t, h)
    row_fn = fn1d("adst" if rf else row_t, w)
    sh = shift_table(w, h)[(bd - 8) // 2]
    buf = [0] * (w * h)
    for c in range(w):
        a = [_shift(int(residual[((h - 1 - r) if cf else r) * w + c]), sh[0]) for r in range(h)]
        b = col_fn(a)
        cc = (w - 1 - c) if rf else c
        for r in range(h):
            buf[r * w + cc] = _shift(b[r], sh[1])
    out = [0] * (w * h)
    for r in range(h):
        b = row_fn(buf[r * w:(r + 1) * w])
        for c in range(w):
            out[r * w + c] = _shift(b[c], sh[2])
    return out
