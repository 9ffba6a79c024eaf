"""rustlite -- a small interpreter for the subset of Rust that the reference's scalar ground-truth functions are
written in (`let` / `for` / `match` / closures / iterator chains / structs, impls and operator traits / raw
pointer arithmetic of the plane-region views), with typed wrapping integers (release-build semantics) and IEEE
f32 / f64.

Test infrastructure, dev container only: it READS the reference's source text from /root/reference at test time,
parses the functions it is asked for and executes them -- nothing of the reference is stored in this repository.
tests/ref_rs.py drives it for src/mc.rs, src/me.rs, src/quantize.rs, src/rdo.rs, src/predict.rs, src/dist.rs and
the plane / plane-region views those functions index through; tests/test_oracle_vs_reference_src.py compares the
oracle (oracle/*.c) with the results, tests/golden/make_ref_golden.py freezes vectors for the GPU box.

Deliberate simplifications (none changes a result of the functions above):
  * lifetimes, generics bounds, visibility, attributes and `where` clauses are parsed and ignored; a generic type
    parameter (T) is bound by the harness through `Interp.tyenv` (e.g. T = u8 / u16);
  * integer literals without a suffix take the type of the other operand (or stay untyped until they meet one);
  * borrows: `&mut place` is a reference cell, `&expr` is the value itself; Copy structs are copied on binding;
  * iterator adaptors are evaluated eagerly where the source is a slice or a range.
"""
import math
import re
import struct as _struct

# ------------------------------------------------------------------------------------------------ lexer ----
_TOKEN_RE = re.compile(r"""
    (?P<ws>\s+|//[^\n]*|/\*.*?\*/)
  | (?P<float>\d[\d_]*\.\d[\d_]*(?:[eE][+-]?\d+)?(?:_?f32|_?f64)?|\d[\d_]*(?:_f32|_f64|f32|f64)|\d[\d_]*\.(?![\.\w]))
  | (?P<int>0x[0-9a-fA-F_]+(?:[iu](?:8|16|32|64|128|size))?|0b[01_]+(?:[iu](?:8|16|32|64|128|size))?|\d[\d_]*(?:[iu](?:8|16|32|64|128|size))?)
  | (?P<life>'[A-Za-z_]\w*(?!'))
  | (?P<char>'(?:\\.|[^\\'])')
  | (?P<str>"(?:\\.|[^\\"])*")
  | (?P<id>[A-Za-z_]\w*)
  | (?P<op><<=|>>=|\.\.=|\.\.\.|::|->|=>|==|!=|<=|>=|&&|\|\||\+=|-=|\*=|/=|%=|\^=|&=|\|=|<<|>>|\.\.|[-+*/%^!&|=<>.,;:()\[\]{}#?@$~])
""", re.X | re.S)


class Tok:
    __slots__ = ("k", "v", "pos")

    def __init__(self, k, v, pos):
        self.k, self.v, self.pos = k, v, pos

    def __repr__(self):
        return "%s:%s" % (self.k, self.v)


def lex(text):
    out, i, n = [], 0, len(text)
    while i < n:
        m = _TOKEN_RE.match(text, i)
        if not m:
            raise SyntaxError("cannot lex at %d: %r" % (i, text[i:i + 30]))
        i = m.end()
        k = m.lastgroup
        if k == "ws":
            continue
        out.append(Tok(k, m.group(k), m.start()))
    out.append(Tok("eof", "", n))
    return out


# ------------------------------------------------------------------------------------------------ values ----
INT_TYPES = {"i8": (8, True), "i16": (16, True), "i32": (32, True), "i64": (64, True), "i128": (128, True),
             "isize": (64, True), "u8": (8, False), "u16": (16, False), "u32": (32, False), "u64": (64, False),
             "u128": (128, False), "usize": (64, False)}
FLOAT_TYPES = ("f32", "f64")
OVERFLOWS = []  # (type, op) of every integer operation that wrapped (a debug build would panic there)


def _wrap(v, ty):
    bits, signed = INT_TYPES[ty]
    m = (1 << bits) - 1
    w = v & m
    if signed and w >> (bits - 1):
        w -= 1 << bits
    return w


class Int:
    __slots__ = ("v", "ty")

    def __init__(self, v, ty=None):
        self.v = int(v) if ty is None else _wrap(int(v), ty)
        self.ty = ty

    def __repr__(self):
        return "%d%s" % (self.v, self.ty or "")

    def __index__(self):
        return self.v

    def __int__(self):
        return self.v

    def __eq__(self, o):
        return isinstance(o, Int) and self.v == o.v

    def __hash__(self):
        return hash(self.v)


class Flt:
    __slots__ = ("v", "ty")

    def __init__(self, v, ty=None):
        v = float(v)
        if ty == "f32":
            v = _f32(v)
        self.v, self.ty = v, ty

    def __repr__(self):
        return "%r%s" % (self.v, self.ty or "")

    def __float__(self):
        return self.v

    def __eq__(self, o):
        return isinstance(o, Flt) and self.v == o.v

    def __hash__(self):
        return hash(self.v)


def _f32(v):
    if math.isnan(v) or math.isinf(v):
        return v
    try:
        return _struct.unpack("f", _struct.pack("f", v))[0]
    except OverflowError:
        return math.copysign(math.inf, v)


class Struct:
    __slots__ = ("name", "f")

    def __init__(self, name, fields):
        self.name, self.f = name, fields

    def __repr__(self):
        return "%s%r" % (self.name, self.f)

    def __eq__(self, o):
        return isinstance(o, Struct) and self.name == o.name and self.f == o.f

    def __hash__(self):
        return id(self)


class EnumVal:
    __slots__ = ("enum", "variant", "disc")

    def __init__(self, enum, variant, disc):
        self.enum, self.variant, self.disc = enum, variant, disc

    def __repr__(self):
        return "%s::%s" % (self.enum, self.variant)

    def __eq__(self, o):
        return isinstance(o, EnumVal) and self.enum == o.enum and self.variant == o.variant

    def __hash__(self):
        return hash((self.enum, self.variant))


class Opt:
    """Some(v) / Ok(v) / Err(v); None is the singleton NONE"""
    __slots__ = ("tag", "v")

    def __init__(self, tag, v=None):
        self.tag, self.v = tag, v

    def __repr__(self):
        return "%s(%r)" % (self.tag, self.v) if self.tag != "None" else "None"

    def __eq__(self, o):
        return isinstance(o, Opt) and self.tag == o.tag and self.v == o.v

    def __hash__(self):
        return hash(self.tag)


NONE = Opt("None")
UNIT = ()


class Arr:
    """Vec / array / boxed slice: elements are interpreter values; `ety` = declared element type or None"""
    __slots__ = ("a", "ety")

    def __init__(self, a, ety=None):
        self.a, self.ety = a, ety

    def __len__(self):
        return len(self.a)

    def __repr__(self):
        return "Arr(len %d)" % len(self.a)

    def __eq__(self, o):
        return isinstance(o, (Arr, View)) and seq_list(self) == seq_list(o)

    def __hash__(self):
        return id(self)


class View:
    """&[T] / &mut [T]: a window of an Arr"""
    __slots__ = ("base", "lo", "hi")

    def __init__(self, base, lo, hi):
        if isinstance(base, View):
            lo, hi, base = base.lo + lo, base.lo + hi, base.base
        if not (0 <= lo <= hi <= len(base.a)):
            raise RustPanic("slice [%d..%d] out of range for length %d" % (lo, hi, len(base.a)))
        self.base, self.lo, self.hi = base, lo, hi

    def __len__(self):
        return self.hi - self.lo

    def __repr__(self):
        return "View(%d..%d)" % (self.lo, self.hi)

    def __eq__(self, o):
        return isinstance(o, (Arr, View)) and seq_list(self) == seq_list(o)

    def __hash__(self):
        return id(self)


def seq_list(s):
    return s.a if isinstance(s, Arr) else s.base.a[s.lo:s.hi]


def seq_get(s, i):
    if isinstance(s, Arr):
        if not 0 <= i < len(s.a):
            raise RustPanic("index %d out of range for length %d" % (i, len(s.a)))
        return s.a[i]
    if not 0 <= i < s.hi - s.lo:
        raise RustPanic("index %d out of range for slice of length %d" % (i, s.hi - s.lo))
    return s.base.a[s.lo + i]


def seq_set(s, i, v):
    if isinstance(s, Arr):
        if not 0 <= i < len(s.a):
            raise RustPanic("index %d out of range for length %d" % (i, len(s.a)))
        if s.ety and isinstance(v, Int) and v.ty is None:
            v = Int(v.v, s.ety)
        s.a[i] = v
    else:
        if not 0 <= i < s.hi - s.lo:
            raise RustPanic("index %d out of range for slice of length %d" % (i, s.hi - s.lo))
        if s.base.ety and isinstance(v, Int) and v.ty is None:
            v = Int(v.v, s.base.ety)
        s.base.a[s.lo + i] = v


class Ptr:
    """*const T / *mut T into an Arr"""
    __slots__ = ("base", "off")

    def __init__(self, base, off):
        if isinstance(base, View):
            base, off = base.base, base.lo + off
        self.base, self.off = base, off

    def __repr__(self):
        return "Ptr(+%d)" % self.off


class Range:
    __slots__ = ("lo", "hi", "incl")

    def __init__(self, lo, hi, incl=False):
        self.lo, self.hi, self.incl = lo, hi, incl

    def bounds(self, n=None):
        lo = 0 if self.lo is None else int(self.lo)
        hi = n if self.hi is None else int(self.hi) + (1 if self.incl else 0)
        return lo, hi


class RefCell:
    """&mut place"""
    __slots__ = ("getf", "setf")

    def __init__(self, getf, setf):
        self.getf, self.setf = getf, setf

    def get(self):
        return self.getf()

    def set(self, v):
        self.setf(v)


class Closure:
    __slots__ = ("params", "body", "env", "interp")

    def __init__(self, params, body, env, interp):
        self.params, self.body, self.env, self.interp = params, body, env, interp


class Iter:
    """eager iterator over a Python list (slices, ranges and everything adapted from them)"""
    __slots__ = ("items", "i")

    def __init__(self, items):
        self.items, self.i = items, 0


class RustPanic(Exception):
    pass


class _Break(Exception):
    def __init__(self, v=UNIT):
        self.v = v


class _Continue(Exception):
    pass


class _Return(Exception):
    def __init__(self, v):
        self.v = v


def deref(v):
    while isinstance(v, RefCell):
        v = v.get()
    return v


def copyval(v):
    """Copy semantics for plain-data structs and tuples (arrays and pointers stay shared)"""
    if isinstance(v, Struct):
        for x in v.f.values():
            if isinstance(x, (Arr, View, RefCell)):
                return v  # owns or borrows storage: a move, never a copy
        return Struct(v.name, {k: copyval(x) for k, x in v.f.items()})
    if isinstance(v, tuple) and v:
        return tuple(copyval(x) for x in v)
    return v


# ------------------------------------------------------------------------------------------------ parser ----
# AST nodes are tuples: (kind, ...).
BINPREC = {"||": 2, "&&": 3, "==": 4, "!=": 4, "<": 4, ">": 4, "<=": 4, ">=": 4, "|": 5, "^": 6, "&": 7,
           "<<": 8, ">>": 8, "+": 9, "-": 9, "*": 10, "/": 10, "%": 10}
ASSIGN_OPS = ("=", "+=", "-=", "*=", "/=", "%=", "^=", "&=", "|=", "<<=", ">>=")


class Parser:
    def __init__(self, toks, i=0):
        self.t, self.i = toks, i

    # -- token helpers
    def peek(self, k=0):
        return self.t[self.i + k]

    def at(self, v, k=0):
        t = self.t[self.i + k]
        return t.v == v and t.k in ("op", "id")

    def eat(self, v):
        if self.at(v):
            self.i += 1
            return True
        return False

    def expect(self, v):
        if not self.at(v):
            raise SyntaxError("expected %r, found %r near token %d (%s)" % (
                v, self.peek().v, self.i, " ".join(t.v for t in self.t[max(0, self.i - 8):self.i + 4])))
        self.i += 1

    def ident(self):
        t = self.peek()
        if t.k != "id":
            raise SyntaxError("expected identifier, found %r (%s)" % (
                t.v, " ".join(x.v for x in self.t[max(0, self.i - 8):self.i + 4])))
        self.i += 1
        return t.v

    def skip_attrs(self):
        while self.at("#"):
            self.i += 1
            self.eat("!")
            self.skip_balanced("[", "]")

    def skip_balanced(self, o, c):
        self.expect(o)
        d = 1
        while d:
            t = self.peek()
            if t.k == "eof":
                raise SyntaxError("unbalanced %s" % o)
            if t.k == "op":
                if t.v == o:
                    d += 1
                elif t.v == c:
                    d -= 1
            self.i += 1

    def skip_generics(self):
        """at '<': skip to the matching '>' ('>>' closes two, '->' and '>=' are single tokens)"""
        if self.at("<<"):
            self.i += 1
            d = 2
        else:
            self.expect("<")
            d = 1
        while d:
            t = self.peek()
            if t.k == "eof":
                raise SyntaxError("unbalanced <")
            if t.k == "op":
                if t.v == "<":
                    d += 1
                elif t.v == ">":
                    d -= 1
                elif t.v == ">>":
                    d -= 2
                elif t.v == "<<":
                    d += 2
                elif t.v in ("(", "["):
                    self.skip_balanced(t.v, ")" if t.v == "(" else "]")
                    continue
            self.i += 1

    def generic_args(self):
        """::<A, B> -> list of type strings"""
        a = self.i
        self.skip_generics()
        inner = self.t[a + 1:self.i - 1]
        out, cur, d = [], [], 0
        for t in inner:
            if t.v in ("<", "(", "["):
                d += 1
            elif t.v in (">", ")", "]"):
                d -= 1
            if t.v == "," and d == 0:
                out.append(" ".join(cur))
                cur = []
            else:
                cur.append(t.v)
        if cur:
            out.append(" ".join(cur))
        return out

    # -- types: returned as a list of token strings
    def type_(self):
        a = self.i
        self._type()
        return [t.v for t in self.t[a:self.i]]

    def _type(self):
        while True:
            if self.at("&") or self.at("&&"):
                self.i += 1
                if self.peek().k == "life":
                    self.i += 1
                self.eat("mut")
                continue
            if self.at("*") and (self.at("const", 1) or self.at("mut", 1)):
                self.i += 2
                continue
            break
        if self.at("["):
            self.skip_balanced("[", "]")
            return
        if self.at("("):
            self.skip_balanced("(", ")")
            return
        bounds = False
        if self.at("impl") or self.at("dyn"):
            self.i += 1
            bounds = True
        if self.at("<"):  # <T as Trait>::X
            self.skip_generics()
            while self.eat("::"):
                self.ident()
            return
        self.eat("::")
        while True:
            self.ident()
            if self.at("<") or self.at("<<"):
                self.skip_generics()
            if self.at("(") and self.t[self.i - 1].v in ("Fn", "FnMut", "FnOnce", "fn"):
                self.skip_balanced("(", ")")
                if self.eat("->"):
                    self._type()
            if self.at("::") and not self.at("<", 1):
                self.i += 1
                continue
            if self.at("::") and self.at("<", 1):
                self.i += 1
                self.skip_generics()
                if self.eat("::"):
                    continue
            break
        while bounds and self.at("+"):  # bounds in `impl A + B`
            self.i += 1
            if self.peek().k == "life":
                self.i += 1
            else:
                self._type()

    # -- patterns
    def pattern(self):
        p = self._pattern1()
        if self.at("|") and not self.at("||"):
            alts = [p]
            while self.eat("|"):
                alts.append(self._pattern1())
            return ("p_or", alts)
        return p

    def _pattern1(self):
        t = self.peek()
        if self.eat("&"):
            self.eat("mut")
            return ("p_ref", self._pattern1())
        if self.eat("&&"):
            return ("p_ref", ("p_ref", self._pattern1()))
        if self.eat("("):
            ps = []
            while not self.at(")"):
                ps.append(self.pattern())
                if not self.eat(","):
                    break
            self.expect(")")
            return ("p_tuple", ps)
        if self.eat("["):
            ps = []
            while not self.at("]"):
                ps.append(self.pattern())
                if not self.eat(","):
                    break
            self.expect("]")
            return ("p_slice", ps)
        if t.k in ("int", "float") or (self.at("-") and self.peek(1).k in ("int", "float")):
            lo = self._lit_pattern()
            if self.at("..=") or self.at("..."):
                self.i += 1
                hi = self._lit_pattern()
                return ("p_range", lo, hi)
            return ("p_lit", lo)
        if t.k == "id" and t.v == "_":
            self.i += 1
            return ("p_wild",)
        if self.at("ref") or self.at("mut"):
            self.eat("ref")
            self.eat("mut")
            return ("p_bind", self.ident())
        if t.k == "id" and t.v in ("true", "false"):
            self.i += 1
            return ("p_lit", t.v == "true")
        # path / binding
        path = [self.ident()]
        while self.at("::"):
            self.i += 1
            if self.at("<"):
                self.skip_generics()
                continue
            path.append(self.ident())
        if self.eat("("):
            ps = []
            while not self.at(")"):
                ps.append(self.pattern())
                if not self.eat(","):
                    break
            self.expect(")")
            return ("p_tstruct", path, ps)
        if self.at("{"):
            self.i += 1
            fields = []
            while not self.at("}"):
                if self.eat(".."):
                    break
                self.eat("ref")
                self.eat("mut")
                name = self.ident()
                if self.eat(":"):
                    fields.append((name, self.pattern()))
                else:
                    fields.append((name, ("p_bind", name)))
                if not self.eat(","):
                    break
            self.expect("}")
            return ("p_struct", path, fields)
        if len(path) == 1:
            if self.eat("@"):
                return ("p_at", path[0], self._pattern1())
            return ("p_ident", path[0])
        return ("p_path", path)

    def _lit_pattern(self):
        neg = self.eat("-")
        t = self.peek()
        self.i += 1
        v = _literal(t)
        if neg:
            v = Int(-v.v, v.ty) if isinstance(v, Int) else Flt(-v.v, v.ty)
        return v

    # -- blocks and statements
    def block(self):
        self.expect("{")
        stmts, tail = [], None
        while not self.at("}"):
            self.skip_attrs()
            if self.at("}"):
                break
            if self.eat(";"):
                continue
            if self.at("let"):
                self.i += 1
                pat = self.pattern()
                ty = None
                if self.eat(":"):
                    ty = self.type_()
                init = None
                if self.eat("="):
                    init = self.expr()
                self.expect(";")
                stmts.append(("let", pat, ty, init))
                continue
            if self._at_item():
                stmts.append(("item", self.item()))
                continue
            if (self.at("{") or self.at("if") or self.at("match") or self.at("for") or self.at("while") or
                    self.at("loop") or (self.at("unsafe") and self.at("{", 1))):
                e = self._primary(False)  # a block-like expression statement ends at its closing brace
                if self.at(".") or self.at("?"):
                    raise SyntaxError("postfix after a block-like statement is not supported")
            else:
                e = self.expr(stmt=True)
            if self.eat(";"):
                stmts.append(("expr", e))
            elif self.at("}"):
                tail = e
            elif e[0] in ("if", "iflet", "match", "for", "while", "whilelet", "loop", "block", "unsafe"):
                stmts.append(("expr", e))
            else:
                raise SyntaxError("expected ; or } after expression, found %r (%s)" % (
                    self.peek().v, " ".join(t.v for t in self.t[max(0, self.i - 10):self.i + 3])))
        self.expect("}")
        return ("block", stmts, tail)

    def _at_item(self):
        k = 0
        while self.at("pub", k) or self.at("unsafe", k) and self.at("fn", k + 1) or self.at("const", k) and self.at("fn", k + 1):
            if self.at("pub", k) and self.at("(", k + 1):
                k += 4
            else:
                k += 1
        return (self.at("fn", k) or self.at("struct", k) or self.at("enum", k) or self.at("impl", k) or
                self.at("use", k) or self.at("static", k) or (self.at("const", k) and self.peek(k + 1).k == "id" and self.peek(k + 2).v == ":"))

    # -- items (function bodies are kept as token positions and parsed on first call)
    def item(self):
        self.skip_attrs()
        if self.eat("pub"):
            if self.at("("):
                self.skip_balanced("(", ")")
        while self.at("unsafe") or (self.at("const") and self.at("fn", 1)) or self.at("extern"):
            if self.at("extern"):
                self.i += 1
                if self.peek().k == "str":
                    self.i += 1
            else:
                self.i += 1
        if self.at("fn"):
            return self.fn_item()
        if self.at("const") or self.at("static"):
            self.i += 1
            self.eat("mut")
            name = self.ident()
            self.expect(":")
            ty = self.type_()
            self.expect("=")
            a = self.i
            self._skip_to_semicolon()
            return ("const", name, ty, a)
        if self.at("use"):
            self._skip_to_semicolon()
            return ("skip",)
        if self.at("struct"):
            self.i += 1
            name = self.ident()
            if self.at("<"):
                self.skip_generics()
            fields = []
            if self.at("("):
                self.i += 1
                k = 0
                while not self.at(")"):
                    self.skip_attrs()
                    self.eat("pub")
                    if self.at("("):
                        self.skip_balanced("(", ")")
                    fields.append((str(k), self.type_()))
                    k += 1
                    if not self.eat(","):
                        break
                self.expect(")")
                if self.at("where"):
                    while not self.at(";"):
                        self.i += 1
                self.expect(";")
                return ("struct", name, fields, True)
            if self.at("where"):
                while not self.at("{"):
                    self.i += 1
            if self.eat(";"):
                return ("struct", name, [], False)
            self.expect("{")
            while not self.at("}"):
                self.skip_attrs()
                if self.eat("pub"):
                    if self.at("("):
                        self.skip_balanced("(", ")")
                fname = self.ident()
                self.expect(":")
                fields.append((fname, self.type_()))
                if not self.eat(","):
                    break
            self.expect("}")
            return ("struct", name, fields, False)
        if self.at("enum"):
            self.i += 1
            name = self.ident()
            if self.at("<"):
                self.skip_generics()
            self.expect("{")
            variants, disc = [], 0
            while not self.at("}"):
                self.skip_attrs()
                vname = self.ident()
                kind, fields = "unit", []
                if self.at("{"):
                    kind = "struct"
                    self.i += 1
                    while not self.at("}"):
                        fn_ = self.ident()
                        self.expect(":")
                        fields.append((fn_, self.type_()))
                        if not self.eat(","):
                            break
                    self.expect("}")
                elif self.at("("):
                    kind = "tuple"
                    self.i += 1
                    k = 0
                    while not self.at(")"):
                        fields.append((str(k), self.type_()))
                        k += 1
                        if not self.eat(","):
                            break
                    self.expect(")")
                if self.eat("="):
                    neg = self.eat("-")
                    t = self.peek()
                    self.i += 1
                    disc = -_literal(t).v if neg else _literal(t).v
                variants.append((vname, kind, fields, disc))
                disc += 1
                if not self.eat(","):
                    break
            self.expect("}")
            return ("enum", name, variants)
        raise SyntaxError("unsupported item at %r (%s)" % (self.peek().v, " ".join(t.v for t in self.t[max(0, self.i - 12):self.i + 6])))

    def _skip_to_semicolon(self):
        d = 0
        while True:
            t = self.peek()
            if t.k == "eof":
                raise SyntaxError("no ; found")
            if t.k == "op":
                if t.v in "([{":
                    d += 1
                elif t.v in ")]}":
                    d -= 1
                elif t.v == ";" and d == 0:
                    self.i += 1
                    return
            self.i += 1

    def fn_item(self):
        self.expect("fn")
        name = self.ident()
        if self.at("<"):
            self.skip_generics()
        self.expect("(")
        params = []
        while not self.at(")"):
            self.skip_attrs()
            # self forms
            if self.at("self") or (self.at("&") and (self.at("self", 1) or self.at("mut", 1) and self.at("self", 2) or
                                                     self.peek(1).k == "life")) or (self.at("mut") and self.at("self", 1)):
                kind = "ref" if self.at("&") else "val"
                while not self.at("self"):
                    self.i += 1
                self.i += 1
                if self.eat(":"):
                    self.type_()
                params.append((("p_bind", "self"), [kind]))
            else:
                pat = self.pattern()
                self.expect(":")
                params.append((pat, self.type_()))
            if not self.eat(","):
                break
        self.expect(")")
        ret = None
        if self.eat("->"):
            ret = self.type_()
        if self.at("where"):
            while not (self.at("{") or self.at(";")):
                if self.at("<"):
                    self.skip_generics()
                else:
                    self.i += 1
        if self.eat(";"):
            return ("fn", name, params, ret, None)
        body_at = self.i
        self.skip_balanced("{", "}")
        return ("fn", name, params, ret, body_at)

    # -- expressions
    def expr(self, stmt=False, nostruct=False):
        return self._assign(nostruct)

    def _assign(self, ns):
        lhs = self._range(ns)
        t = self.peek()
        if t.k == "op" and t.v in ASSIGN_OPS:
            self.i += 1
            rhs = self._assign(ns)
            return ("assign", t.v, lhs, rhs)
        return lhs

    def _range(self, ns):
        if self.at("..") or self.at("..="):
            incl = self.peek().v == "..="
            self.i += 1
            hi = None if self._range_end(ns) else self._binary(2, ns)
            return ("range", None, hi, incl)
        lo = self._binary(2, ns)
        if self.at("..") or self.at("..="):
            incl = self.peek().v == "..="
            self.i += 1
            hi = None if self._range_end(ns) else self._binary(2, ns)
            return ("range", lo, hi, incl)
        return lo

    def _range_end(self, ns):
        t = self.peek()
        return t.k == "eof" or (t.k == "op" and t.v in (")", "]", "}", ",", ";", "=>")) or (ns and t.v == "{")

    def _binary(self, minprec, ns):
        lhs = self._cast(ns)
        while True:
            t = self.peek()
            if t.k != "op" or t.v not in BINPREC or BINPREC[t.v] < minprec:
                return lhs
            prec = BINPREC[t.v]
            self.i += 1
            rhs = self._binary(prec + 1, ns)
            lhs = ("bin", t.v, lhs, rhs)

    def _cast(self, ns):
        e = self._unary(ns)
        while self.at("as"):
            self.i += 1
            e = ("cast", e, self.type_())
        return e

    def _unary(self, ns):
        t = self.peek()
        if t.k == "op":
            if t.v == "-":
                self.i += 1
                return ("neg", self._unary(ns))
            if t.v == "!":
                self.i += 1
                return ("not", self._unary(ns))
            if t.v == "*":
                self.i += 1
                return ("deref", self._unary(ns))
            if t.v in ("&", "&&"):
                self.i += 1
                mut = self.eat("mut")
                inner = self._unary(ns)
                if t.v == "&&":
                    return ("addr", False, ("addr", mut, inner))
                return ("addr", mut, inner)
        return self._postfix(ns)

    def _postfix(self, ns):
        e = self._primary(ns)
        while True:
            if self.at("?"):
                self.i += 1
                continue
            if self.at("."):
                nt = self.peek(1)
                if nt.k == "int":  # tuple field
                    self.i += 2
                    e = ("field", e, nt.v)
                    continue
                if nt.k == "float":  # a.0.1 lexes as float "0.1"
                    self.i += 2
                    for part in nt.v.split("."):
                        e = ("field", e, part)
                    continue
                if nt.k == "id":
                    self.i += 2
                    name = nt.v
                    targs = None
                    if self.at("::") and self.at("<", 1):
                        self.i += 1
                        targs = self.generic_args()
                    if self.at("("):
                        args = self._args()
                        e = ("mcall", e, name, args, targs)
                    else:
                        e = ("field", e, name)
                    continue
                raise SyntaxError("bad token after . : %r" % nt.v)
            if self.at("("):
                e = ("call", e, self._args())
                continue
            if self.at("["):
                self.i += 1
                idx = self.expr()
                self.expect("]")
                e = ("index", e, idx)
                continue
            return e

    def _args(self):
        self.expect("(")
        args = []
        while not self.at(")"):
            args.append(self.expr())
            if not self.eat(","):
                break
        self.expect(")")
        return args

    def _primary(self, ns):
        t = self.peek()
        if t.k in ("int", "float"):
            self.i += 1
            return ("lit", _literal(t))
        if t.k == "str":
            self.i += 1
            return ("lit", t.v)
        if t.k == "char":
            self.i += 1
            return ("lit", t.v)
        if t.k == "life":  # loop label
            self.i += 1
            self.expect(":")
            return self._primary(ns)
        if t.k == "op":
            if t.v == "(":
                self.i += 1
                if self.eat(")"):
                    return ("tuple", [])
                first = self.expr()
                if self.eat(")"):
                    return ("paren", first)
                items = [first]
                while self.eat(","):
                    if self.at(")"):
                        break
                    items.append(self.expr())
                self.expect(")")
                return ("tuple", items)
            if t.v == "[":
                self.i += 1
                if self.eat("]"):
                    return ("array", [])
                first = self.expr()
                if self.eat(";"):
                    n = self.expr()
                    self.expect("]")
                    return ("repeat", first, n)
                items = [first]
                while self.eat(","):
                    if self.at("]"):
                        break
                    items.append(self.expr())
                self.expect("]")
                return ("array", items)
            if t.v == "{":
                return self.block()
            if t.v in ("|", "||"):
                return self._closure()
            if t.v == "<":  # <T as Trait>::name
                self.skip_generics()
                path = ["<qualified>"]
                while self.eat("::"):
                    path.append(self.ident())
                return ("path", path, None)
        if t.k != "id":
            raise SyntaxError("unexpected token %r (%s)" % (t.v, " ".join(x.v for x in self.t[max(0, self.i - 10):self.i + 4])))
        v = t.v
        if v in ("true", "false"):
            self.i += 1
            return ("lit", v == "true")
        if v == "unsafe":
            self.i += 1
            return self.block()
        if v == "move" and (self.at("|", 1) or self.at("||", 1)):
            self.i += 1
            return self._closure()
        if v == "if":
            return self._if()
        if v == "match":
            self.i += 1
            scrut = self.expr(nostruct=True)
            self.expect("{")
            arms = []
            while not self.at("}"):
                self.skip_attrs()
                self.eat("|")
                pat = self.pattern()
                guard = None
                if self.eat("if"):
                    guard = self.expr()
                self.expect("=>")
                body = self.block() if self.at("{") else self.expr()
                arms.append((pat, guard, body))
                if not self.eat(","):
                    if not self.at("}") and body[0] not in ("block", "if", "match", "iflet"):
                        raise SyntaxError("expected , in match")
            self.expect("}")
            return ("match", scrut, arms)
        if v == "loop":
            self.i += 1
            return ("loop", self.block())
        if v == "while":
            self.i += 1
            if self.eat("let"):
                pat = self.pattern()
                self.expect("=")
                e = self.expr(nostruct=True)
                return ("whilelet", pat, e, self.block())
            c = self.expr(nostruct=True)
            return ("while", c, self.block())
        if v == "for":
            self.i += 1
            pat = self.pattern()
            self.expect("in")
            it = self.expr(nostruct=True)
            return ("for", pat, it, self.block())
        if v == "return":
            self.i += 1
            if self._range_end(False):
                return ("return", None)
            return ("return", self.expr())
        if v == "break":
            self.i += 1
            if self.peek().k == "life":
                self.i += 1
            if self._range_end(False):
                return ("break", None)
            return ("break", self.expr())
        if v == "continue":
            self.i += 1
            if self.peek().k == "life":
                self.i += 1
            return ("continue",)
        # path, macro, struct literal
        path, targs = [self.ident()], None
        if self.at("!") and not self.at("=", 1) and self.peek(1).v in ("(", "[", "{"):
            self.i += 1
            return self._macro(path[0])
        while self.at("::"):
            self.i += 1
            if self.at("<"):
                targs = self.generic_args()
                continue
            path.append(self.ident())
        if self.at("{") and not ns and self._looks_like_struct_lit(path):
            self.i += 1
            fields, base = [], None
            while not self.at("}"):
                if self.eat(".."):
                    base = self.expr()
                    break
                nt = self.peek()
                self.i += 1
                name = nt.v
                if self.eat(":"):
                    fields.append((name, self.expr()))
                else:
                    fields.append((name, ("path", [name], None)))
                if not self.eat(","):
                    break
            self.expect("}")
            return ("structlit", path, fields, base)
        return ("path", path, targs)

    def _looks_like_struct_lit(self, path):
        if not path[-1][0].isupper():
            return False
        a, b = self.peek(1), self.peek(2)
        if a.v == "}":
            return True
        if a.v == "..":
            return True
        return a.k in ("id", "int") and b.v in (":", ",", "}") and not (b.v == ":" and self.peek(3).v == ":")

    def _closure(self):
        params = []
        if self.eat("||"):
            pass
        else:
            self.expect("|")
            while not self.at("|"):
                pat = self._pattern1()
                if self.eat(":"):
                    self.type_()
                params.append(pat)
                if not self.eat(","):
                    break
            self.expect("|")
        if self.eat("->"):
            self.type_()
            body = self.block()
        else:
            body = self.expr()
        return ("closure", params, body)

    def _if(self):
        self.expect("if")
        if self.eat("let"):
            pat = self.pattern()
            self.expect("=")
            e = self.expr(nostruct=True)
            then = self.block()
            els = None
            if self.eat("else"):
                els = self._if() if self.at("if") else self.block()
            return ("iflet", pat, e, then, els)
        c = self.expr(nostruct=True)
        then = self.block()
        els = None
        if self.eat("else"):
            els = self._if() if self.at("if") else self.block()
        return ("if", c, then, els)

    def _macro(self, name):
        o = self.peek().v
        c = {"(": ")", "[": "]", "{": "}"}[o]
        self.i += 1
        args = []
        rep = None
        while not self.at(c):
            args.append(self.expr())
            if self.eat(";"):
                rep = self.expr()
                break
            if not self.eat(","):
                break
        self.expect(c)
        return ("macro", name, args, rep)


def _literal(t):
    s = t.v.replace("_", "")
    if t.k == "float":
        ty = None
        for f in FLOAT_TYPES:
            if s.endswith(f):
                ty, s = f, s[:-3]
        return Flt(float(s), ty)
    ty = None
    m = re.search(r"([iu](?:8|16|32|64|128|size))$", s)
    if m and not s.startswith("0x"):
        ty, s = m.group(1), s[:m.start()]
    elif m and s.startswith("0x") and re.search(r"[iu](8|16|32|64|128|size)$", s) and not re.fullmatch(r"0x[0-9a-fA-F]+", s):
        ty, s = m.group(1), s[:m.start()]
    return Int(int(s, 0), ty)


# ------------------------------------------------------------------------------------------- interpreter ----
class FnDef:
    __slots__ = ("name", "params", "ret", "toks", "body_at", "body", "impl_ty", "src", "scope")

    def __init__(self, name, params, ret, toks, body_at, impl_ty, src, scope=None):
        self.name, self.params, self.ret, self.toks, self.body_at = name, params, ret, toks, body_at
        self.body, self.impl_ty, self.src, self.scope = None, impl_ty, src, scope


class Env:
    __slots__ = ("vars", "parent")

    def __init__(self, parent=None):
        self.vars, self.parent = {}, parent

    def lookup(self, name):
        e = self
        while e is not None:
            if name in e.vars:
                return e
            e = e.parent
        return None


class Interp:
    def __init__(self):
        self.fns = {}        # free functions by name
        self.methods = {}    # (type name, method) -> FnDef
        self.consts = {}     # name -> ("lazy", ty, toks, at) | ("val", v)
        self.assoc_consts = {}
        self.enums = {}      # enum -> [(variant, kind, fields, disc)]
        self.variants = {}   # bare variant name -> EnumVal (C-like enums, for `use Enum::*`)
        self.structs = {}    # struct -> ([(field, type tokens)], is_tuple)
        self.tyenv = {}      # generic parameter / alias -> concrete type name
        self.builtin_fns = {}    # name -> python callable(interp, args)
        self.builtin_methods = {}  # (kind, method) -> python callable(interp, recv, args, targs)
        self.builtin_paths = {}  # "A::B" -> value or callable
        self.calls = {}      # name -> call count (what was actually executed from the reference's text)
        _install_builtins(self)

    # ---- loading source text
    def load(self, path, text=None):
        if text is None:
            text = open(path).read()
        self.load_tokens(lex(text), path)

    def load_tokens(self, toks, src):
        p = Parser(toks)
        self._items(p, src, until_eof=True)

    def _items(self, p, src, until_eof=False, impl_ty=None):
        while True:
            p.skip_attrs()
            if p.peek().k == "eof":
                return
            if p.at("}") and not until_eof:
                return
            # things that are skipped wholesale
            k = 0
            if p.at("pub"):
                k = 1
                if p.at("(", 1):
                    k = 4
            if p.at("mod", k):
                p.i += k + 1
                p.ident()
                if p.eat(";"):
                    continue
                p.expect("{")
                self._items(p, src)
                p.expect("}")
                continue
            if p.at("trait", k) or (p.at("unsafe", k) and p.at("trait", k + 1)):
                while not p.at("{"):
                    if p.at("<"):
                        p.skip_generics()
                    else:
                        p.i += 1
                p.skip_balanced("{", "}")
                continue
            if p.at("type", k):
                p._skip_to_semicolon()
                continue
            if p.at("macro_rules", k):
                p.i += k + 2
                p.ident()
                p.skip_balanced(p.peek().v, {"{": "}", "(": ")", "[": "]"}[p.peek().v])
                p.eat(";")
                continue
            if p.at("extern", k) and p.at("crate", k + 1):
                p._skip_to_semicolon()
                continue
            if p.at("extern", k) and (p.at("{", k + 1) or p.peek(k + 1).k == "str" and p.at("{", k + 2)):
                while not p.at("{"):
                    p.i += 1
                p.skip_balanced("{", "}")
                continue
            if p.at("impl", k) or (p.at("unsafe", k) and p.at("impl", k + 1)):
                while not p.at("impl"):
                    p.i += 1
                p.i += 1
                if p.at("<"):
                    p.skip_generics()
                p.eat("!")
                t1 = p.type_()
                ty = t1
                if p.eat("for"):
                    ty = p.type_()
                if p.at("where"):
                    while not p.at("{"):
                        if p.at("<"):
                            p.skip_generics()
                        else:
                            p.i += 1
                tname = _type_head(ty)
                p.expect("{")
                while not p.at("}"):
                    p.skip_attrs()
                    if p.at("}"):
                        break
                    if p.at("type"):
                        p._skip_to_semicolon()
                        continue
                    kk = 0
                    if p.at("pub"):
                        kk = 4 if p.at("(", 1) else 1
                    if p.at("const", kk) and not p.at("fn", kk + 1):
                        it = p.item()
                        self.assoc_consts[(tname, it[1])] = ("lazy", it[2], p.t, it[3])
                        continue
                    it = p.item()
                    if it[0] == "fn":
                        self.methods[(tname, it[1])] = FnDef(it[1], it[2], it[3], p.t, it[4], tname, src)
                p.expect("}")
                continue
            # an ident followed by ! at item level: a macro invocation (e.g. plane_region_common!(..);)
            if p.peek().k == "id" and p.at("::", 1) and p.peek(2).k == "id" and p.at("!", 3):
                p.i += 2
            if p.peek().k == "id" and p.at("!", 1):
                p.i += 2
                p.skip_balanced(p.peek().v, {"{": "}", "(": ")", "[": "]"}[p.peek().v])
                p.eat(";")
                continue
            it = p.item()
            self._register(it, p.t, src)

    def _register(self, it, toks, src):
        if it[0] == "fn":
            self.fns[it[1]] = FnDef(it[1], it[2], it[3], toks, it[4], None, src)
        elif it[0] == "const":
            self.consts[it[1]] = ("lazy", it[2], toks, it[3])
        elif it[0] == "struct":
            self.structs[it[1]] = (it[2], it[3])
        elif it[0] == "enum":
            self.enums[it[1]] = it[2]
            for vname, kind, _f, disc in it[2]:
                if kind == "unit":
                    self.variants.setdefault(vname, EnumVal(it[1], vname, disc))

    # ---- calling
    def call(self, name, *args):
        """call a free function (or 'Type::method') of the loaded text with interpreter values"""
        if "::" in name:
            ty, m = name.split("::")
            return self._invoke(self.methods[(ty, m)], list(args))
        return self._invoke(self.fns[name], list(args))

    def _invoke(self, fd, args, closure_env=None):
        if fd.body is None:
            if fd.body_at is None:
                raise RustPanic("function %s has no body" % fd.name)
            fd.body = Parser(fd.toks, fd.body_at).block()
        self.calls[fd.name] = self.calls.get(fd.name, 0) + 1
        env = Env(fd.scope)  # nested fns see the sibling items of their block (never its locals, in valid Rust)
        env.vars["%impl"] = fd.impl_ty
        env.vars["%toks"] = fd.toks
        if len(args) != len(fd.params):
            raise RustPanic("%s: %d arguments for %d parameters" % (fd.name, len(args), len(fd.params)))
        for (pat, ty), a in zip(fd.params, args):
            if pat == ("p_bind", "self"):
                env.vars["self"] = deref(a) if not isinstance(deref(a), (Int, Flt)) or ty == ["val"] else a
                if ty == ["val"]:
                    env.vars["self"] = copyval(deref(a))
                continue
            if ty and ty[0] not in ("&", "&&"):
                a = copyval(deref(a)) if not (ty[0] == "*") else a
            a = self.coerce(a, ty)
            self._bind(pat, a, env)
        try:
            v = self.eval(fd.body, env)
        except _Return as r:
            v = r.v
        if fd.ret:
            v = self.coerce(v, fd.ret)
        return v

    def call_value(self, f, args):
        f = deref(f)
        if isinstance(f, Closure):
            env = Env(f.env)
            if len(args) != len(f.params):
                raise RustPanic("closure arity")
            for pat, a in zip(f.params, args):
                self._bind(pat, a, env)
            try:
                return self.eval(f.body, env)
            except _Return as r:
                return r.v
        if isinstance(f, FnDef):
            return self._invoke(f, args)
        if callable(f):
            return f(self, args)
        raise RustPanic("not callable: %r" % (f,))

    # ---- types
    def resolve_ty(self, name):
        seen = 0
        while name in self.tyenv and seen < 8:
            name = self.tyenv[name]
            seen += 1
        return name

    def coerce(self, v, ty):
        """give untyped literals the declared type; arrays of a declared element type remember it"""
        if not ty:
            return v
        ty = [t for t in ty if t not in ("&", "&&", "mut") and not t.startswith("'")]
        if not ty:
            return v
        head = self.resolve_ty(ty[0])
        if isinstance(v, Int):
            if head in INT_TYPES and len(ty) == 1:
                if v.ty is None:
                    return Int(v.v, head)
                return v
            if head in FLOAT_TYPES and v.ty is None:
                return Flt(v.v, head)
        elif isinstance(v, Flt):
            if head in FLOAT_TYPES and v.ty is None and len(ty) == 1:
                return Flt(v.v, head)
        elif isinstance(v, Arr) and ty[0] == "[":
            inner = ty[1:]
            # [T; N] or [T]
            d, cut = 0, len(inner) - 1
            for k, t in enumerate(inner):
                if t in ("[", "(", "<"):
                    d += 1
                elif t in ("]", ")", ">"):
                    if d == 0:
                        cut = k
                        break
                    d -= 1
                elif t == ";" and d == 0:
                    cut = k
                    break
            ety = inner[:cut]
            eh = self.resolve_ty(ety[0]) if ety else None
            if len(ety) == 1 and eh in INT_TYPES:
                v.ety = eh
                v.a = [Int(x.v, eh) if isinstance(x, Int) and x.ty is None else x for x in v.a]
            elif len(ety) == 1 and eh in FLOAT_TYPES:
                v.a = [Flt(x.v, eh) if isinstance(x, (Flt, Int)) and x.ty is None else x for x in v.a]
            else:
                v.a = [self.coerce(x, ety) for x in v.a]
        elif isinstance(v, tuple) and ty[0] == "(":
            parts, cur, d = [], [], 0
            for t in ty[1:-1]:
                if t in ("[", "(", "<"):
                    d += 1
                elif t in ("]", ")", ">"):
                    d -= 1
                if t == "," and d == 0:
                    parts.append(cur)
                    cur = []
                else:
                    cur.append(t)
            if cur:
                parts.append(cur)
            if len(parts) == len(v):
                return tuple(self.coerce(x, p) for x, p in zip(v, parts))
        elif isinstance(v, Opt) and ty[0] == "Option" and v.tag == "Some" and len(ty) > 3:
            return Opt("Some", self.coerce(v.v, ty[2:-1]))
        return v

    def cast(self, v, ty):
        v = deref(v)
        ty = [t for t in ty if t not in ("&", "&&", "mut")]
        if ty[0] == "*":  # pointer casts keep the pointer
            return v
        head = self.resolve_ty(ty[0])
        if head == "_":
            return v
        if head in INT_TYPES:
            if isinstance(v, Int):
                return Int(v.v, head)
            if isinstance(v, bool):
                return Int(int(v), head)
            if isinstance(v, EnumVal):
                return Int(v.disc, head)
            if isinstance(v, Flt):
                bits, signed = INT_TYPES[head]
                lo, hi = (-(1 << (bits - 1)), (1 << (bits - 1)) - 1) if signed else (0, (1 << bits) - 1)
                if math.isnan(v.v):
                    return Int(0, head)
                if math.isinf(v.v):
                    return Int(hi if v.v > 0 else lo, head)
                return Int(min(max(int(v.v), lo), hi), head)  # `as` saturates since Rust 1.45
            if isinstance(v, str) and v.startswith("'"):
                return Int(ord(v[1:-1].encode().decode("unicode_escape")), head)
        if head in FLOAT_TYPES:
            if isinstance(v, Int):
                return Flt(float(v.v), head)  # int -> float rounds to nearest even, as Python does
            if isinstance(v, Flt):
                return Flt(v.v, head)
        raise RustPanic("unsupported cast of %r to %s" % (v, " ".join(ty)))

    def type_name(self, v):
        if isinstance(v, Struct):
            return v.name.split("::")[0]
        if isinstance(v, EnumVal):
            return v.enum
        if isinstance(v, Int):
            return v.ty or "i32"
        if isinstance(v, Flt):
            return v.ty or "f64"
        if isinstance(v, bool):
            return "bool"
        return None

    def default_of(self, tyname):
        tyname = self.resolve_ty(tyname)
        if tyname in INT_TYPES:
            return Int(0, tyname)
        if tyname in FLOAT_TYPES:
            return Flt(0.0, tyname)
        if tyname == "bool":
            return False
        if tyname in self.structs:
            fields, _tup = self.structs[tyname]
            return Struct(tyname, {n: self.default_of(_type_head(t)) for n, t in fields})
        if (tyname, "default") in self.methods:
            return self._invoke(self.methods[(tyname, "default")], [])
        raise RustPanic("no default for %s" % tyname)

    # ---- constants
    def const(self, name):
        c = self.consts[name]
        if c[0] == "lazy":
            _k, ty, toks, at = c
            e = Parser(toks, at).expr()
            v = self.coerce(self.eval(e, Env(None)), ty)
            self.consts[name] = ("val", v)
            return v
        return c[1]

    # ---- pattern binding
    def _bind(self, pat, v, env):
        if not self._match(pat, v, env):
            raise RustPanic("irrefutable pattern did not match: %r vs %r" % (pat, v))

    def _match(self, pat, v, env):
        k = pat[0]
        if k == "p_wild":
            return True
        if k == "p_bind":
            env.vars[pat[1]] = v
            return True
        if k == "p_ident":
            name = pat[1]
            if name == "None":
                return deref(v) is NONE
            if name in self.variants and isinstance(deref(v), EnumVal):
                return deref(v) == self.variants[name]
            if name in self.consts and name.isupper():
                return _values_equal(deref(v), self.const(name))
            env.vars[name] = v
            return True
        if k == "p_at":
            if self._match(pat[2], v, env):
                env.vars[pat[1]] = v
                return True
            return False
        if k == "p_ref":
            return self._match(pat[1], deref(v) if isinstance(v, RefCell) else v, env)
        v = deref(v)
        if k == "p_lit":
            return _values_equal(v, pat[1])
        if k == "p_range":
            return pat[1].v <= v.v <= pat[2].v
        if k == "p_or":
            return any(self._match(p, v, env) for p in pat[1])
        if k == "p_tuple":
            if not isinstance(v, tuple) or len(v) != len(pat[1]):
                return False
            return all(self._match(p, x, env) for p, x in zip(pat[1], v))
        if k == "p_slice":
            xs = seq_list(v)
            return len(xs) == len(pat[1]) and all(self._match(p, x, env) for p, x in zip(pat[1], xs))
        if k == "p_path":
            want = self.eval(("path", pat[1], None), env)
            return _values_equal(v, want)
        if k == "p_tstruct":
            name = pat[1][-1]
            if name in ("Some", "Ok", "Err"):
                if not isinstance(v, Opt) or v.tag != name:
                    return False
                return self._match(pat[2][0], v.v, env)
            if not isinstance(v, Struct):
                return False
            full = "::".join(pat[1][-2:]) if len(pat[1]) > 1 else name
            if v.name != full and v.name != name and v.name.split("::")[-1] != name:
                return False
            return all(self._match(p, v.f[str(i)], env) for i, p in enumerate(pat[2]))
        if k == "p_struct":
            if not isinstance(v, Struct):
                return False
            name = pat[1][-1]
            if name != "Self" and v.name.split("::")[-1] != name:
                return False
            return all(self._match(p, v.f[n], env) for n, p in pat[2])
        raise RustPanic("pattern %r" % (pat,))

    # ---- evaluation
    def eval(self, e, env):
        return getattr(self, "_e_" + e[0])(e, env)

    def _e_lit(self, e, env):
        return e[1]

    def _e_paren(self, e, env):
        return self.eval(e[1], env)

    def _e_tuple(self, e, env):
        return tuple(copyval(deref_if_val(self.eval(x, env))) for x in e[1])

    def _e_array(self, e, env):
        return Arr([copyval(self.eval(x, env)) for x in e[1]])

    def _e_repeat(self, e, env):
        v = self.eval(e[1], env)
        n = int(deref(self.eval(e[2], env)))
        if isinstance(v, (Arr, Struct)):
            return Arr([_deepcopy(v) for _ in range(n)])
        ety = v.ty if isinstance(v, Int) else None
        return Arr([v] * n, ety)

    def _e_block(self, e, env, new_scope=True):
        env2 = Env(env) if new_scope else env
        for s in e[1]:  # items are visible in the whole block
            if s[0] == "item" and s[1][0] == "fn":
                it = s[1]
                env2.vars[it[1]] = FnDef(it[1], it[2], it[3], self._cur_toks(env2), it[4], None, "nested", env2)
        for s in e[1]:
            if s[0] == "let":
                v = UNIT
                if s[3] is not None:
                    v = self.eval(s[3], env2)
                    if _is_place(s[3]):
                        v = copyval(v)
                    if s[2]:
                        v = self.coerce(v, s[2])
                self._bind(s[1], v, env2)
            elif s[0] == "expr":
                self.eval(s[1], env2)
            elif s[0] == "item":
                it = s[1]
                if it[0] == "const":
                    ex = Parser(self._cur_toks(env2), it[3]).expr()
                    env2.vars[it[1]] = self.coerce(self.eval(ex, env2), it[2])
        if e[2] is not None:
            return self.eval(e[2], env2)
        return UNIT

    def _cur_toks(self, env):
        e = env.lookup("%toks")
        return e.vars["%toks"] if e else None

    def _e_if(self, e, env):
        c = deref(self.eval(e[1], env))
        if not isinstance(c, bool):
            raise RustPanic("if on non-bool %r" % (c,))
        if c:
            return self.eval(e[2], env)
        if e[3] is not None:
            return self.eval(e[3], env)
        return UNIT

    def _e_iflet(self, e, env):
        v = self.eval(e[2], env)
        env2 = Env(env)
        if self._match(e[1], v, env2):
            return self._e_block(e[3], env2, new_scope=False)
        if e[4] is not None:
            return self.eval(e[4], env)
        return UNIT

    def _e_match(self, e, env):
        v = self.eval(e[1], env)
        for pat, guard, body in e[2]:
            env2 = Env(env)
            if self._match(pat, v, env2):
                if guard is not None and not deref(self.eval(guard, env2)):
                    continue
                return self.eval(body, env2)
        raise RustPanic("no match arm for %r" % (v,))

    def _e_loop(self, e, env):
        while True:
            try:
                self.eval(e[1], env)
            except _Break as b:
                return b.v
            except _Continue:
                pass

    def _e_while(self, e, env):
        while deref(self.eval(e[1], env)):
            try:
                self.eval(e[2], env)
            except _Break:
                break
            except _Continue:
                pass
        return UNIT

    def _e_whilelet(self, e, env):
        while True:
            env2 = Env(env)
            if not self._match(e[1], self.eval(e[2], env), env2):
                break
            try:
                self._e_block(e[3], env2, new_scope=False)
            except _Break:
                break
            except _Continue:
                pass
        return UNIT

    def _e_for(self, e, env):
        if e[2][0] == "addr" and e[2][1]:  # for x in &mut seq: the elements are places
            seq = deref(self.eval(e[2][2], env))
            it = self.builtin_methods[("seq", "iter_mut")](self, None, seq, [], None)
        else:
            it = self.eval(e[2], env)
        for x in self.iterate(it):
            env2 = Env(env)
            self._bind(e[1], x, env2)
            try:
                self._e_block(e[3], env2, new_scope=False)
            except _Break:
                break
            except _Continue:
                pass
        return UNIT

    def iterate(self, it):
        """python iterator over anything a `for` loop / adaptor accepts"""
        it = deref(it)
        if isinstance(it, Iter):
            while it.i < len(it.items):
                x = it.items[it.i]
                it.i += 1
                yield x
            return
        if isinstance(it, Range):
            lo, hi = it.bounds()
            ty = (it.lo.ty if it.lo is not None else None) or (it.hi.ty if it.hi is not None else None)
            for k in range(lo, hi):
                yield Int(k, ty)
            return
        if isinstance(it, (Arr, View)):
            for x in list(seq_list(it)):
                yield x
            return
        if isinstance(it, Struct) and (self.type_name(it), "next") in self.methods:
            nxt = self.methods[(self.type_name(it), "next")]
            while True:
                r = self._invoke(nxt, [it])
                if r is NONE:
                    return
                yield r.v
        if hasattr(it, "__next__"):
            for x in it:
                yield x
            return
        raise RustPanic("cannot iterate %r" % (it,))

    def _e_return(self, e, env):
        raise _Return(UNIT if e[1] is None else copyval(self.eval(e[1], env)))

    def _e_break(self, e, env):
        raise _Break(UNIT if e[1] is None else self.eval(e[1], env))

    def _e_continue(self, e, env):
        raise _Continue()

    def _e_closure(self, e, env):
        return Closure(e[1], e[2], env, self)

    def _e_range(self, e, env):
        lo = None if e[1] is None else deref(self.eval(e[1], env))
        hi = None if e[2] is None else deref(self.eval(e[2], env))
        if lo is not None and hi is not None:
            if lo.ty is None and hi.ty is not None:
                lo = Int(lo.v, hi.ty)
            elif hi.ty is None and lo.ty is not None:
                hi = Int(hi.v, lo.ty)
        return Range(lo, hi, e[3])

    def _e_cast(self, e, env):
        return self.cast(self.eval(e[1], env), e[2])

    def _e_neg(self, e, env):
        v = deref(self.eval(e[1], env))
        if isinstance(v, Int):
            r = Int(-v.v, v.ty)
            if v.ty and r.v != -v.v:
                OVERFLOWS.append((v.ty, "neg"))
            return r
        if isinstance(v, Flt):
            return Flt(-v.v, v.ty)
        if isinstance(v, Struct) and (self.type_name(v), "neg") in self.methods:
            return self._invoke(self.methods[(self.type_name(v), "neg")], [v])
        raise RustPanic("neg of %r" % (v,))

    def _e_not(self, e, env):
        v = deref(self.eval(e[1], env))
        if isinstance(v, bool):
            return not v
        if isinstance(v, Int):
            if v.ty is None:
                return Int(~v.v)
            return Int(~v.v, v.ty)
        raise RustPanic("not of %r" % (v,))

    def _e_deref(self, e, env):
        v = self.eval(e[1], env)
        if isinstance(v, RefCell):
            return copyval(v.get())
        if isinstance(v, Ptr):
            return seq_get(v.base, v.off)
        return v  # `*x` on something bound by value (borrows are values here)

    def _e_addr(self, e, env):
        mut, inner = e[1], e[2]
        if mut:
            return self.place_ref(inner, env)
        return self.eval(inner, env)

    def place_ref(self, e, env):
        """&mut place -> RefCell (or the shared object itself for aggregates)"""
        k = e[0]
        if k == "paren":
            return self.place_ref(e[1], env)
        if k == "path" and len(e[1]) == 1:
            name = e[1][0]
            scope = env.lookup(name)
            if scope is None:
                raise RustPanic("&mut of unknown variable %s" % name)
            cur = scope.vars[name]
            if isinstance(cur, (RefCell, Arr, View)) or (isinstance(cur, Struct)):
                return cur  # already a reference / a shared mutable object
            return RefCell(lambda: scope.vars[name], lambda v: scope.vars.__setitem__(name, v))
        if k == "field":
            obj = deref(self.eval(e[1], env))
            cur = obj.f[e[2]] if isinstance(obj, Struct) else obj[int(e[2])]
            if isinstance(cur, (RefCell, Arr, View, Struct)):
                return cur
            if isinstance(obj, Struct):
                fname = e[2]
                return RefCell(lambda: obj.f[fname], lambda v: obj.f.__setitem__(fname, v))
        if k == "index":
            r = self.eval(e, env)  # a View for range indices, else an element
            if isinstance(r, (View, Arr, Struct, RefCell)):
                return r
            cont = deref(self.eval(e[1], env))
            idx = int(deref(self.eval(e[2], env)))
            return RefCell(lambda: seq_get(cont, idx), lambda v: seq_set(cont, idx, v))
        if k == "deref":
            return self.eval(e[1], env)
        return self.eval(e, env)

    def _e_assign(self, e, env):
        op, lhs, rhs = e[1], e[2], e[3]
        v = self.eval(rhs, env)
        if _is_place(rhs):
            v = copyval(v)
        if op != "=":
            cur = self.eval(lhs, env)
            curd = deref(cur)
            if isinstance(curd, Struct):
                m = {"+=": "add_assign", "-=": "sub_assign", "*=": "mul_assign", "/=": "div_assign"}[op]
                if (self.type_name(curd), m) in self.methods:
                    self._invoke(self.methods[(self.type_name(curd), m)], [curd, v])
                    return UNIT
            v = self.binop(op[:-1], curd, deref(v))
        self.assign(lhs, v, env)
        return UNIT

    def assign(self, lhs, v, env):
        k = lhs[0]
        if k == "paren":
            return self.assign(lhs[1], v, env)
        if k == "path" and len(lhs[1]) == 1:
            name = lhs[1][0]
            scope = env.lookup(name)
            if scope is None:
                raise RustPanic("assignment to unknown variable %s" % name)
            cur = scope.vars[name]
            if isinstance(cur, Int) and isinstance(v, Int) and v.ty is None and cur.ty:
                v = Int(v.v, cur.ty)
            if isinstance(cur, Flt) and isinstance(v, (Flt, Int)) and v.ty is None and cur.ty:
                v = Flt(v.v, cur.ty)
            scope.vars[name] = v
            return
        if k == "field":
            obj = deref(self.eval(lhs[1], env))
            if not isinstance(obj, Struct):
                raise RustPanic("field assignment on %r" % (obj,))
            cur = obj.f.get(lhs[2])
            if isinstance(cur, Int) and isinstance(v, Int) and v.ty is None and cur.ty:
                v = Int(v.v, cur.ty)
            obj.f[lhs[2]] = v
            return
        if k == "index":
            cont = deref(self.eval(lhs[1], env))
            idx = deref(self.eval(lhs[2], env))
            if isinstance(cont, Struct):
                tn = self.type_name(cont)
                m = self.methods.get((tn, "index_mut")) or self.methods.get((tn, "index"))
                cont = deref(self._invoke(m, [cont, idx]))
                raise RustPanic("assignment through user Index yields a place, not supported here")
            if isinstance(idx, Range):
                raise RustPanic("slice assignment")
            seq_set(cont, int(idx), v)
            return
        if k == "deref":
            r = self.eval(lhs[1], env)
            if isinstance(r, RefCell):
                cur = r.get()
                if isinstance(cur, Int) and isinstance(v, Int) and v.ty is None and cur.ty:
                    v = Int(v.v, cur.ty)
                r.set(v)
                return
            if isinstance(r, Ptr):
                seq_set(r.base, r.off, v)
                return
            if isinstance(r, Struct) and isinstance(v, Struct):
                r.f = dict(v.f)  # *self = other
                r.name = v.name
                return
            # `*x = v` where x was bound by value from an iterator of element refs is not supported
            raise RustPanic("deref assignment to %r" % (r,))
        raise RustPanic("cannot assign to %r" % (lhs[0],))

    def _e_bin(self, e, env):
        op = e[1]
        if op == "&&":
            a = deref(self.eval(e[2], env))
            return a and deref(self.eval(e[3], env))
        if op == "||":
            a = deref(self.eval(e[2], env))
            return a or deref(self.eval(e[3], env))
        a = deref(self.eval(e[2], env))
        b = deref(self.eval(e[3], env))
        return self.binop(op, a, b)

    def binop(self, op, a, b):
        if isinstance(a, Int) and isinstance(b, Int):
            if op in ("<<", ">>"):
                ty = a.ty
                sh = b.v
                if op == "<<":
                    r = a.v << sh
                else:
                    r = a.v >> sh
                if ty is None:
                    return Int(r)
                out = Int(r, ty)
                return out
            ty = a.ty or b.ty
            if a.ty and b.ty and a.ty != b.ty:
                ra, rb = self.resolve_ty(a.ty), self.resolve_ty(b.ty)
                if INT_TYPES[ra] != INT_TYPES[rb]:
                    raise RustPanic("mismatched integer types %s %s %s" % (a.ty, op, b.ty))
            x, y = a.v, b.v
            if op == "+":
                r = x + y
            elif op == "-":
                r = x - y
            elif op == "*":
                r = x * y
            elif op == "/":
                if y == 0:
                    raise RustPanic("division by zero")
                r = abs(x) // abs(y)
                if (x < 0) != (y < 0):
                    r = -r
            elif op == "%":
                if y == 0:
                    raise RustPanic("remainder by zero")
                r = abs(x) % abs(y)
                if x < 0:
                    r = -r
            elif op == "&":
                r = x & y
            elif op == "|":
                r = x | y
            elif op == "^":
                r = x ^ y
            elif op == "==":
                return x == y
            elif op == "!=":
                return x != y
            elif op == "<":
                return x < y
            elif op == ">":
                return x > y
            elif op == "<=":
                return x <= y
            elif op == ">=":
                return x >= y
            else:
                raise RustPanic("int op %s" % op)
            if ty is None:
                return Int(r)
            out = Int(r, ty)
            if out.v != r and op in ("+", "-", "*"):
                OVERFLOWS.append((ty, op))
            return out
        if isinstance(a, (Flt, Int)) and isinstance(b, (Flt, Int)) and (isinstance(a, Flt) or isinstance(b, Flt)):
            if isinstance(a, Int) and a.ty is not None or isinstance(b, Int) and b.ty is not None:
                raise RustPanic("mixed int / float arithmetic %r %s %r" % (a, op, b))
            ty = (a.ty if isinstance(a, Flt) else None) or (b.ty if isinstance(b, Flt) else None)
            x, y = float(a.v), float(b.v)
            if op == "+":
                r = x + y
            elif op == "-":
                r = x - y
            elif op == "*":
                r = x * y
            elif op == "/":
                r = x / y if y != 0 else (math.nan if x == 0 or math.isnan(x) else math.copysign(math.inf, x) * math.copysign(1, y))
            elif op == "==":
                return x == y
            elif op == "!=":
                return x != y
            elif op == "<":
                return x < y
            elif op == ">":
                return x > y
            elif op == "<=":
                return x <= y
            elif op == ">=":
                return x >= y
            else:
                raise RustPanic("float op %s" % op)
            return Flt(r, ty)
        if isinstance(a, bool) and isinstance(b, bool):
            return {"==": a == b, "!=": a != b, "&": a and b, "|": a or b, "^": a != b}[op]
        if isinstance(a, EnumVal) and isinstance(b, EnumVal):
            if op == "==":
                return a == b
            if op == "!=":
                return a != b
            return {"<": a.disc < b.disc, ">": a.disc > b.disc, "<=": a.disc <= b.disc, ">=": a.disc >= b.disc}[op]
        if isinstance(a, Struct):
            m = {"+": "add", "-": "sub", "*": "mul", "/": "div", "<<": "shl", ">>": "shr"}.get(op)
            if m and (self.type_name(a), m) in self.methods:
                return self._invoke(self.methods[(self.type_name(a), m)], [a, b])
        if op == "==":
            return _values_equal(a, b)
        if op == "!=":
            return not _values_equal(a, b)
        if isinstance(a, tuple) and isinstance(b, tuple) and op in ("<", ">", "<=", ">="):
            ka, kb = _order_key(a), _order_key(b)
            return {"<": ka < kb, ">": ka > kb, "<=": ka <= kb, ">=": ka >= kb}[op]
        raise RustPanic("unsupported operands %r %s %r" % (a, op, b))

    def _e_field(self, e, env):
        obj = deref(self.eval(e[1], env))
        name = e[2]
        if isinstance(obj, tuple):
            return obj[int(name)]
        if isinstance(obj, Struct):
            if name not in obj.f:
                raise RustPanic("no field %s on %s" % (name, obj.name))
            return obj.f[name]
        raise RustPanic("field %s of %r" % (name, obj))

    def _e_index(self, e, env):
        cont = deref(self.eval(e[1], env))
        idx = deref(self.eval(e[2], env))
        if isinstance(cont, Struct):
            tn = self.type_name(cont)
            m = self.methods.get((tn, "index")) or self.methods.get((tn, "index_mut"))
            if m is None:
                raise RustPanic("%s is not indexable" % tn)
            return self._invoke(m, [cont, idx])
        if isinstance(cont, (Arr, View)):
            if isinstance(idx, Range):
                lo, hi = idx.bounds(len(cont))
                return View(cont, lo, hi)
            return seq_get(cont, int(idx))
        if isinstance(cont, tuple):
            return cont[int(idx)]
        raise RustPanic("index into %r" % (cont,))

    def _e_structlit(self, e, env):
        path = e[1]
        name = path[-1]
        if name == "Self":
            name = env.lookup("%impl").vars["%impl"] if env.lookup("%impl") else None
            name = self._impl_of(env)
        full = name
        if len(path) > 1 and path[-2] in self.enums:
            full = path[-2] + "::" + name
            decl = next((f for vn, _k, f, _d in self.enums[path[-2]] if vn == name), [])
        else:
            decl = self.structs.get(name, ([], False))[0]
        decl = dict(decl)
        fields = {}
        for fname, fe in e[2]:
            v = copyval(self.eval(fe, env))
            if fname in decl:
                v = self.coerce(v, decl[fname])
            fields[fname] = v
        if e[3] is not None:
            base = deref(self.eval(e[3], env))
            for k2, v2 in base.f.items():
                fields.setdefault(k2, copyval(v2))
        return Struct(full, fields)

    def _impl_of(self, env):
        s = env.lookup("%impl")
        return s.vars["%impl"] if s else None

    def _e_macro(self, e, env):
        name, args, rep = e[1], e[2], e[3]
        if name in ("assert", "debug_assert"):
            if not deref(self.eval(args[0], env)):
                raise RustPanic("assertion failed in reference text")
            return UNIT
        if name in ("assert_eq", "debug_assert_eq"):
            a, b = deref(self.eval(args[0], env)), deref(self.eval(args[1], env))
            if not _values_equal(a, b):
                raise RustPanic("assert_eq failed: %r vs %r" % (a, b))
            return UNIT
        if name == "vec":
            if rep is not None:
                v = self.eval(args[0], env)
                n = int(deref(self.eval(rep, env)))
                return Arr([_deepcopy(v) if isinstance(v, (Arr, Struct)) else v for _ in range(n)],
                           v.ty if isinstance(v, Int) else None)
            return Arr([copyval(self.eval(a, env)) for a in args])
        if name in ("unimplemented", "unreachable", "panic"):
            raise RustPanic("%s!() reached in reference text" % name)
        if name in ("println", "eprintln", "print", "debug", "trace"):
            return UNIT
        raise RustPanic("macro %s! is not supported" % name)

    def _e_path(self, e, env):
        path, targs = e[1], e[2]
        if len(path) == 1:
            name = path[0]
            scope = env.lookup(name)
            if scope is not None:
                return scope.vars[name]
            if name in self.consts:
                return self.const(name)
            if name == "None":
                return NONE
            if name in self.variants:
                return self.variants[name]
            if name in self.fns:
                return self.fns[name]
            if name in self.builtin_fns:
                return self.builtin_fns[name]
            if name in self.builtin_paths:
                return self.builtin_paths[name]
            if name in self.structs and not self.structs[name][0]:
                return Struct(name, {})  # unit struct
            raise RustPanic("unknown name %s" % name)
        # strip leading crate / module qualifiers down to the last two segments
        a, b = path[-2], path[-1]
        if a == "Self":
            a = self._impl_of(env)
        a = self.resolve_ty(a)
        if a in self.enums:
            for vn, kind, fields, disc in self.enums[a]:
                if vn == b:
                    if kind == "unit":
                        return EnumVal(a, vn, disc)
                    en = a
                    return lambda it, args, _en=en, _vn=vn: Struct(_en + "::" + _vn, {str(i): x for i, x in enumerate(args)})
        if (a, b) in self.methods:
            return self.methods[(a, b)]
        if (a, b) in self.assoc_consts:
            c = self.assoc_consts[(a, b)]
            if c[0] == "lazy":
                ex = Parser(c[2], c[3]).expr()
                v = self.coerce(self.eval(ex, Env(None)), c[1])
                self.assoc_consts[(a, b)] = ("val", v)
                return v
            return c[1]
        key = a + "::" + b
        if key in self.builtin_paths:
            return self.builtin_paths[key]
        if a in INT_TYPES or a in FLOAT_TYPES:
            f = _prim_static(self, a, b)
            if f is not None:
                return f
        if b in self.consts:
            return self.const(b)
        if b in self.fns:
            return self.fns[b]
        if b in self.builtin_fns:
            return self.builtin_fns[b]
        if b == "default" and a in self.structs:
            return lambda it, args, _a=a: it.default_of(_a)
        if b in ("new", "with_capacity") and a in ("Vec", "ArrayVec"):
            return lambda it, args: Arr([])
        if b == "into" and a == "Into":
            t = self.resolve_ty(targs[0]) if targs else None
            return lambda it, args, _t=t: it.cast(args[0], [_t])
        if b == "from" and (a in INT_TYPES or a in FLOAT_TYPES):
            return lambda it, args, _a=a: it.cast(args[0], [_a])
        raise RustPanic("unknown path %s" % "::".join(path))

    def _e_call(self, e, env):
        fexpr = e[1]
        # tuple-struct constructor / Some / Ok / Err
        if fexpr[0] == "path":
            p = fexpr[1]
            last = p[-1]
            if last in ("Some", "Ok", "Err") and len(p) == 1:
                return Opt(last, copyval(self.eval(e[2][0], env)))
            if len(p) == 1 and env.lookup(last) is None and last in self.structs and self.structs[last][1]:
                decl = self.structs[last][0]
                vals = [copyval(self.eval(a, env)) for a in e[2]]
                return Struct(last, {str(i): self.coerce(v, decl[i][1]) for i, v in enumerate(vals)})
            if len(p) == 1 and last == "Self" and self._impl_of(env) in self.structs:
                nm = self._impl_of(env)
                decl = self.structs[nm][0]
                vals = [copyval(self.eval(a, env)) for a in e[2]]
                return Struct(nm, {str(i): self.coerce(v, decl[i][1]) for i, v in enumerate(vals)})
        f = self.eval(fexpr, env)
        typed = isinstance(deref(f), FnDef)  # _invoke copies by-value parameters itself, by declared type
        args = []
        for a in e[2]:
            v = self.eval(a, env)
            if not typed and _is_place(a):
                v = copyval(v)
            args.append(v)
        if isinstance(f, FnDef) and f.impl_ty is not None and f.params and f.params[0][0] == ("p_bind", "self"):
            return self._invoke(f, args)  # Type::method(self_value, ..)
        return self.call_value(f, args)

    def _e_mcall(self, e, env):
        recv_e, name, arg_es, targs = e[1], e[2], e[3], e[4]
        recv = self.eval(recv_e, env)
        r = deref(recv)
        tn = self.type_name(r)
        m = None
        if tn is not None:
            m = self.methods.get((tn, name))
            if m is None and isinstance(r, (Int, Flt)):
                m = self.methods.get((self.resolve_ty(tn), name))
        args = []
        for a in arg_es:
            v = self.eval(a, env)
            if m is None and _is_place(a):
                v = copyval(v)
            args.append(v)
        if tn is not None:
            if m is not None:
                # &mut self on a scalar held in a variable: hand the place over
                if isinstance(r, (Int, Flt)) and m.params and m.params[0][1] == ["ref"] and not isinstance(recv, RefCell):
                    recv = r
                return self._invoke(m, [recv if isinstance(recv, RefCell) and isinstance(r, (Int, Flt)) else r] + args)
            if isinstance(r, Struct) and name in r.f and isinstance(deref(r.f[name]), (Closure, FnDef)):
                return self.call_value(r.f[name], args)
        return self.builtin_method(recv, r, name, args, targs, env)

    def builtin_method(self, recv, r, name, args, targs, env):
        kind = _kind(r)
        f = self.builtin_methods.get((kind, name)) or self.builtin_methods.get(("any", name))
        if f is None:
            raise RustPanic("no method %s on %s (%r)" % (name, kind, r))
        return f(self, recv, r, args, targs)


def deref_if_val(v):
    return v


def _is_place(e):
    k = e[0]
    if k == "path":
        return len(e[1]) == 1
    if k in ("field", "index", "deref"):
        return True
    if k == "paren":
        return _is_place(e[1])
    return False


def _deepcopy(v):
    if isinstance(v, Arr):
        return Arr([_deepcopy(x) for x in v.a], v.ety)
    if isinstance(v, Struct):
        return Struct(v.name, {k: _deepcopy(x) for k, x in v.f.items()})
    if isinstance(v, tuple):
        return tuple(_deepcopy(x) for x in v)
    return v


def _type_head(ty):
    """first path identifier of a type token list (skipping references / pointers / generics)"""
    names = [t for t in ty if re.match(r"[A-Za-z_]\w*$", t) and t not in ("mut", "const", "dyn", "impl", "crate", "super", "self", "std", "ops", "core")]
    # path a::b::C<..> -> C ; stop at the first generic bracket
    out = None
    for t in ty:
        if t == "<" and out is not None:
            break
        if t in names:
            out = t
    return out


def _values_equal(a, b):
    a, b = deref(a), deref(b)
    if isinstance(a, Int) and isinstance(b, Int):
        return a.v == b.v
    if isinstance(a, Flt) and isinstance(b, (Flt, Int)) or isinstance(b, Flt) and isinstance(a, (Flt, Int)):
        return float(a.v) == float(b.v)
    if isinstance(a, tuple) and isinstance(b, tuple):
        return len(a) == len(b) and all(_values_equal(x, y) for x, y in zip(a, b))
    if isinstance(a, (Arr, View)) and isinstance(b, (Arr, View)):
        la, lb = seq_list(a), seq_list(b)
        return len(la) == len(lb) and all(_values_equal(x, y) for x, y in zip(la, lb))
    if isinstance(a, Struct) and isinstance(b, Struct):
        return a.name == b.name and a.f.keys() == b.f.keys() and all(_values_equal(a.f[k], b.f[k]) for k in a.f)
    if isinstance(a, Opt) and isinstance(b, Opt):
        return a.tag == b.tag and (a.tag == "None" or _values_equal(a.v, b.v))
    return a == b


def _order_key(v):
    v = deref(v)
    if isinstance(v, (Int, Flt)):
        return v.v
    if isinstance(v, EnumVal):
        return v.disc
    if isinstance(v, tuple):
        return tuple(_order_key(x) for x in v)
    if isinstance(v, bool):
        return int(v)
    raise RustPanic("no ordering for %r" % (v,))


def _kind(r):
    if isinstance(r, Int):
        return "int"
    if isinstance(r, Flt):
        return "float"
    if isinstance(r, bool):
        return "bool"
    if isinstance(r, (Arr, View)):
        return "seq"
    if isinstance(r, Iter):
        return "iter"
    if isinstance(r, Range):
        return "range"
    if isinstance(r, Opt):
        return "opt"
    if isinstance(r, Ptr):
        return "ptr"
    if isinstance(r, tuple):
        return "tuple"
    if isinstance(r, Struct):
        return "struct"
    if isinstance(r, EnumVal):
        return "enum"
    if isinstance(r, Closure):
        return "closure"
    return "other"


# ---------------------------------------------------------------------------------------------- builtins ----
def _prim_static(it, ty, name):
    if ty in INT_TYPES:
        bits, signed = INT_TYPES[ty]
        if name in ("MAX", "max_value"):
            v = Int((1 << (bits - 1)) - 1 if signed else (1 << bits) - 1, ty)
            return v if name == "MAX" else (lambda _it, a: v)
        if name in ("MIN", "min_value"):
            v = Int(-(1 << (bits - 1)) if signed else 0, ty)
            return v if name == "MIN" else (lambda _it, a: v)
        if name in ("cast_from", "from"):
            return lambda _it, a: _it.cast(a[0], [ty])
    else:
        if name == "sqrt":
            return lambda _it, a: Flt(math.sqrt(deref(a[0]).v), ty)
        if name in ("cast_from", "from"):
            return lambda _it, a: _it.cast(a[0], [ty])
        if name == "MAX":
            return Flt(3.4028234663852886e38 if ty == "f32" else 1.7976931348623157e308, ty)
    return None


def _mk_iter(it, r):
    """iterator value over a sequence / range / user iterator -> Iter (eager)"""
    if isinstance(r, Iter):
        return r
    return Iter(list(it.iterate(r)))


def _install_builtins(I):
    B = I.builtin_methods

    def reg(kind, names, f):
        for n in names.split():
            B[(kind, n)] = f

    # ---- integers
    def int_minmax(which):
        def f(it, recv, r, a, t):
            o = deref(a[0])
            ty = r.ty or o.ty
            v = min(r.v, o.v) if which == "min" else max(r.v, o.v)
            return Int(v, ty)
        return f
    reg("int", "min", int_minmax("min"))
    reg("int", "max", int_minmax("max"))

    def int_abs(it, recv, r, a, t):
        out = Int(abs(r.v), r.ty)
        return out
    reg("int", "abs", int_abs)
    reg("int", "wrapping_abs", int_abs)

    def need_ty(r):
        if r.ty is None:
            return "i32"
        return I.resolve_ty(r.ty)

    def int_lz(it, recv, r, a, t):
        bits, _s = INT_TYPES[need_ty(r)]
        u = r.v & ((1 << bits) - 1)
        return Int(bits - u.bit_length(), "u32")
    reg("int", "leading_zeros", int_lz)

    def int_tz(it, recv, r, a, t):
        bits, _s = INT_TYPES[need_ty(r)]
        u = r.v & ((1 << bits) - 1)
        return Int(bits if u == 0 else (u & -u).bit_length() - 1, "u32")
    reg("int", "trailing_zeros", int_tz)
    reg("int", "count_ones", lambda it, recv, r, a, t: Int(bin(r.v & ((1 << INT_TYPES[need_ty(r)][0]) - 1)).count("1"), "u32"))
    reg("int", "signum", lambda it, recv, r, a, t: Int((r.v > 0) - (r.v < 0), r.ty))
    reg("int", "pow", lambda it, recv, r, a, t: Int(r.v ** deref(a[0]).v, r.ty))
    reg("int", "is_power_of_two", lambda it, recv, r, a, t: r.v > 0 and r.v & (r.v - 1) == 0)
    reg("int", "as_", lambda it, recv, r, a, t: Int(r.v, "i32"))  # AsPrimitive<i32>: the only instantiation on these paths
    reg("int", "into clone to_owned", lambda it, recv, r, a, t: r)
    reg("int", "wrapping_add", lambda it, recv, r, a, t: Int(r.v + deref(a[0]).v, r.ty or deref(a[0]).ty))
    reg("int", "wrapping_sub", lambda it, recv, r, a, t: Int(r.v - deref(a[0]).v, r.ty or deref(a[0]).ty))
    reg("int", "wrapping_mul", lambda it, recv, r, a, t: Int(r.v * deref(a[0]).v, r.ty or deref(a[0]).ty))
    reg("int", "saturating_sub", lambda it, recv, r, a, t: Int(max(r.v - deref(a[0]).v, -(1 << 63) if INT_TYPES[need_ty(r)][1] else 0), r.ty))
    reg("int", "to_i32", lambda it, recv, r, a, t: Opt("Some", Int(r.v, "i32")))
    reg("int", "is_zero", lambda it, recv, r, a, t: r.v == 0)

    def int_clamp(it, recv, r, a, t):
        lo, hi = deref(a[0]), deref(a[1])
        return Int(min(max(r.v, lo.v), hi.v), r.ty or lo.ty or hi.ty)
    reg("int", "clamp", int_clamp)

    # Fixed for usize (src/util/mod.rs:89-106) is interpreted from the text when loaded; no builtin here.

    # ---- floats
    reg("float", "sqrt", lambda it, recv, r, a, t: Flt(math.sqrt(r.v), r.ty))
    reg("float", "abs", lambda it, recv, r, a, t: Flt(abs(r.v), r.ty))
    reg("float", "is_finite", lambda it, recv, r, a, t: math.isfinite(r.v))
    reg("float", "max", lambda it, recv, r, a, t: Flt(max(r.v, deref(a[0]).v), r.ty or deref(a[0]).ty))
    reg("float", "min", lambda it, recv, r, a, t: Flt(min(r.v, deref(a[0]).v), r.ty or deref(a[0]).ty))
    reg("float", "into clone", lambda it, recv, r, a, t: r)
    reg("float", "round", lambda it, recv, r, a, t: Flt(math.floor(abs(r.v) + 0.5) * (1 if r.v >= 0 else -1), r.ty))
    reg("float", "floor", lambda it, recv, r, a, t: Flt(math.floor(r.v), r.ty))
    reg("float", "log2", lambda it, recv, r, a, t: Flt(math.log2(r.v), r.ty))

    # ---- sequences
    reg("seq", "len", lambda it, recv, r, a, t: Int(len(r), "usize"))
    reg("seq", "is_empty", lambda it, recv, r, a, t: len(r) == 0)
    reg("seq", "iter into_iter", lambda it, recv, r, a, t: Iter(list(seq_list(r))))

    def seq_iter_mut(it, recv, r, a, t):
        out = []
        for k in range(len(r)):
            x = seq_get(r, k)
            if isinstance(x, (Struct, Arr)):
                out.append(x)
            else:
                out.append(RefCell((lambda _k: lambda: seq_get(r, _k))(k), (lambda _k: lambda v: seq_set(r, _k, v))(k)))
        return Iter(out)
    reg("seq", "iter_mut", seq_iter_mut)
    reg("seq", "as_ptr as_mut_ptr", lambda it, recv, r, a, t: Ptr(r, 0))

    def seq_copy_from(it, recv, r, a, t):
        src = seq_list(deref(a[0]))
        if len(src) != len(r):
            raise RustPanic("copy_from_slice length mismatch")
        for k, x in enumerate(src):
            seq_set(r, k, x)
        return UNIT
    reg("seq", "copy_from_slice clone_from_slice", seq_copy_from)

    def seq_push(it, recv, r, a, t):
        v = copyval(deref(a[0])) if not isinstance(a[0], RefCell) else a[0]
        r.a.append(v)
        return UNIT
    reg("seq", "push", seq_push)

    def seq_extend(it, recv, r, a, t):
        for x in it.iterate(a[0]):
            r.a.append(copyval(x))
        return UNIT
    reg("seq", "extend", seq_extend)
    reg("seq", "clone to_vec to_owned", lambda it, recv, r, a, t: Arr([copyval(x) for x in seq_list(r)], r.ety if isinstance(r, Arr) else r.base.ety))
    reg("seq", "as_slice as_mut_slice as_ref as_mut deref", lambda it, recv, r, a, t: r)

    def seq_bsearch(it, recv, r, a, t):
        key = deref(a[0]).v
        xs = [x.v for x in seq_list(r)]
        lo, hi = 0, len(xs)
        # same contract as slice::binary_search: Ok(any matching index) / Err(insertion point)
        import bisect
        k = bisect.bisect_left(xs, key)
        if k < len(xs) and xs[k] == key:
            return Opt("Ok", Int(k, "usize"))
        return Opt("Err", Int(k, "usize"))
    reg("seq", "binary_search", seq_bsearch)

    def seq_chunks(it, recv, r, a, t):
        n = int(deref(a[0]))
        return Iter([View(r, k, min(k + n, len(r))) for k in range(0, len(r), n)])
    reg("seq", "chunks chunks_mut", seq_chunks)
    reg("seq", "first", lambda it, recv, r, a, t: Opt("Some", seq_get(r, 0)) if len(r) else NONE)
    reg("seq", "last", lambda it, recv, r, a, t: Opt("Some", seq_get(r, len(r) - 1)) if len(r) else NONE)
    reg("seq", "fill", lambda it, recv, r, a, t: [seq_set(r, k, deref(a[0])) for k in range(len(r))] and UNIT)
    reg("seq", "contains", lambda it, recv, r, a, t: any(_values_equal(x, a[0]) for x in seq_list(r)))
    reg("seq", "swap", lambda it, recv, r, a, t: _seq_swap(r, int(deref(a[0])), int(deref(a[1]))))
    reg("seq", "split_at split_at_mut", lambda it, recv, r, a, t: (View(r, 0, int(deref(a[0]))), View(r, int(deref(a[0])), len(r))))

    def _seq_swap(r, i, j):
        x, y = seq_get(r, i), seq_get(r, j)
        seq_set(r, i, y)
        seq_set(r, j, x)
        return UNIT

    # ---- ranges
    reg("range", "step_by rev map enumerate zip take skip filter sum collect into_iter count min max fold all any position for_each len",
        lambda it, recv, r, a, t: None)  # placeholder, replaced below by delegation

    def delegate(name):
        def f(it, recv, r, a, t):
            return B[("iter", name)](it, None, _mk_iter(it, r), a, t)
        return f
    for n in ("step_by rev map enumerate zip take skip filter sum collect count min max fold all any position rposition "
              "for_each last chain cloned copied flat_map filter_map take_while skip_while next nth peekable min_by_key max_by_key").split():
        B[("range", n)] = delegate(n)
        B[("seq", n)] = B.get(("seq", n)) or delegate(n)
    reg("range", "into_iter", lambda it, recv, r, a, t: _mk_iter(it, r))
    reg("range", "len", lambda it, recv, r, a, t: Int(max(0, r.bounds()[1] - r.bounds()[0]), "usize"))
    reg("range", "contains", lambda it, recv, r, a, t: r.bounds()[0] <= deref(a[0]).v < r.bounds()[1])

    # ---- iterators (eager lists)
    def rest(r):
        return r.items[r.i:]
    reg("iter", "iter into_iter cloned copied by_ref peekable fuse", lambda it, recv, r, a, t: Iter([copyval(deref(x)) if False else x for x in rest(r)]))
    reg("iter", "cloned copied", lambda it, recv, r, a, t: Iter([copyval(deref(x)) for x in rest(r)]))
    reg("iter", "map", lambda it, recv, r, a, t: Iter([it.call_value(a[0], [x]) for x in rest(r)]))
    reg("iter", "for_each", lambda it, recv, r, a, t: [it.call_value(a[0], [x]) for x in rest(r)] and UNIT)
    reg("iter", "enumerate", lambda it, recv, r, a, t: Iter([(Int(k, "usize"), x) for k, x in enumerate(rest(r))]))
    def iter_zip(it, recv, r, a, t):
        out, other = [], it.iterate(a[0])  # the second iterator is advanced only as far as the first reaches
        for x in rest(r):
            try:
                y = next(other)
            except StopIteration:
                break
            out.append((x, y))
        return Iter(out)
    reg("iter", "zip", iter_zip)
    reg("iter", "rev", lambda it, recv, r, a, t: Iter(rest(r)[::-1]))
    reg("iter", "skip", lambda it, recv, r, a, t: Iter(rest(r)[int(deref(a[0])):]))
    reg("iter", "take", lambda it, recv, r, a, t: Iter(rest(r)[:int(deref(a[0]))]))
    reg("iter", "step_by", lambda it, recv, r, a, t: Iter(rest(r)[::int(deref(a[0]))]))
    reg("iter", "chain", lambda it, recv, r, a, t: Iter(rest(r) + list(it.iterate(a[0]))))
    reg("iter", "filter", lambda it, recv, r, a, t: Iter([x for x in rest(r) if deref(it.call_value(a[0], [x]))]))
    reg("iter", "take_while", lambda it, recv, r, a, t: Iter(_take_while(it, rest(r), a[0])))
    reg("iter", "count len", lambda it, recv, r, a, t: Int(len(rest(r)), "usize"))
    reg("iter", "collect", lambda it, recv, r, a, t: Arr([copyval(deref(x)) if not isinstance(deref(x), (Struct, Arr)) else x for x in rest(r)]))
    reg("iter", "last", lambda it, recv, r, a, t: Opt("Some", rest(r)[-1]) if rest(r) else NONE)

    def _take_while(it, xs, f):
        out = []
        for x in xs:
            if not deref(it.call_value(f, [x])):
                break
            out.append(x)
        return out

    def iter_next(it, recv, r, a, t):
        if r.i < len(r.items):
            r.i += 1
            return Opt("Some", r.items[r.i - 1])
        return NONE
    reg("iter", "next", iter_next)

    def iter_nth(it, recv, r, a, t):
        r.i += int(deref(a[0]))
        return iter_next(it, recv, r, a, t)
    reg("iter", "nth", iter_nth)

    def iter_sum(it, recv, r, a, t):
        ty = it.resolve_ty(t[0]) if t else None
        xs = [deref(x) for x in rest(r)]
        if xs and isinstance(xs[0], Flt) or ty in FLOAT_TYPES:
            acc = Flt(0.0, ty or (xs[0].ty if xs else None))
            for x in xs:
                acc = it.binop("+", acc, x)
            return acc
        acc = Int(0, ty or (xs[0].ty if xs else None))
        for x in xs:
            acc = it.binop("+", acc, x)
        return acc
    reg("iter", "sum", iter_sum)

    def iter_fold(it, recv, r, a, t):
        acc = a[0]
        for x in rest(r):
            acc = it.call_value(a[1], [acc, x])
        return acc
    reg("iter", "fold", iter_fold)

    def iter_position(it, recv, r, a, t):
        for k, x in enumerate(rest(r)):
            if deref(it.call_value(a[0], [x])):
                return Opt("Some", Int(k, "usize"))
        return NONE
    reg("iter", "position", iter_position)

    def iter_rposition(it, recv, r, a, t):
        xs = rest(r)
        for k in range(len(xs) - 1, -1, -1):
            if deref(it.call_value(a[0], [xs[k]])):
                return Opt("Some", Int(k, "usize"))
        return NONE
    reg("iter", "rposition", iter_rposition)
    reg("iter", "all", lambda it, recv, r, a, t: all(deref(it.call_value(a[0], [x])) for x in rest(r)))
    reg("iter", "any", lambda it, recv, r, a, t: any(deref(it.call_value(a[0], [x])) for x in rest(r)))

    def iter_minmax(which, bykey):
        def f(it, recv, r, a, t):
            xs = rest(r)
            if not xs:
                return NONE
            key = (lambda x: _order_key(it.call_value(a[0], [x]))) if bykey else _order_key
            best = xs[0]
            bk = key(best)
            for x in xs[1:]:
                k = key(x)
                # Iterator::min returns the FIRST minimum, Iterator::max the LAST maximum
                if (which == "min" and k < bk) or (which == "max" and k >= bk):
                    best, bk = x, k
            return Opt("Some", best)
        return f
    reg("iter", "min", iter_minmax("min", False))
    reg("iter", "max", iter_minmax("max", False))
    reg("iter", "min_by_key", iter_minmax("min", True))
    reg("iter", "max_by_key", iter_minmax("max", True))

    # ---- user iterators (structs with a `next` method): adaptors go through an eager list, except step_by,
    # which std implements as next() then nth(step - 1) -- reproduced so that a type's own `nth` is exercised
    def struct_step_by(it, recv, r, a, t):
        step = int(deref(a[0]))
        tn = it.type_name(r)
        nxt, nth = it.methods.get((tn, "next")), it.methods.get((tn, "nth"))
        if nxt is None:
            raise RustPanic("step_by on %s" % tn)
        out = []
        x = it._invoke(nxt, [r])
        while x is not NONE:
            out.append(x.v)
            if nth is not None:
                x = it._invoke(nth, [r, Int(step - 1, "usize")])
            else:
                for _ in range(step - 1):
                    x = it._invoke(nxt, [r])
                    if x is NONE:
                        break
                if x is not NONE:
                    x = it._invoke(nxt, [r])
        return Iter(out)
    reg("struct", "step_by", struct_step_by)

    def struct_take(it, recv, r, a, t):
        n, out = int(deref(a[0])), []
        for x in it.iterate(r):
            if len(out) >= n:
                break
            out.append(x)
            if len(out) >= n:
                break
        return Iter(out)
    B[("struct", "take")] = struct_take
    for n in "map enumerate zip skip filter sum collect count fold all any position for_each last rev".split():
        B[("struct", n)] = delegate(n)

    # ---- Option / Result
    def opt_map(it, recv, r, a, t):
        if r.tag in ("Some", "Ok"):
            return Opt(r.tag, it.call_value(a[0], [r.v]))
        return r
    reg("opt", "map", opt_map)
    reg("opt", "unwrap expect", lambda it, recv, r, a, t: r.v if r.tag in ("Some", "Ok") else _panic("unwrap on %r" % (r,)))
    reg("opt", "unwrap_or", lambda it, recv, r, a, t: r.v if r.tag in ("Some", "Ok") else a[0])
    reg("opt", "unwrap_or_else", lambda it, recv, r, a, t: r.v if r.tag in ("Some", "Ok") else it.call_value(a[0], []))
    reg("opt", "unwrap_or_default", lambda it, recv, r, a, t: r.v if r.tag in ("Some", "Ok") else Int(0))
    reg("opt", "is_some is_ok", lambda it, recv, r, a, t: r.tag in ("Some", "Ok"))
    reg("opt", "is_none is_err", lambda it, recv, r, a, t: r.tag in ("None", "Err"))
    reg("opt", "as_ref as_mut clone cloned copied", lambda it, recv, r, a, t: r)
    reg("opt", "ok", lambda it, recv, r, a, t: Opt("Some", r.v) if r.tag == "Ok" else NONE)
    reg("opt", "and_then", lambda it, recv, r, a, t: it.call_value(a[0], [r.v]) if r.tag in ("Some", "Ok") else r)
    reg("opt", "iter into_iter", lambda it, recv, r, a, t: Iter([r.v] if r.tag in ("Some", "Ok") else []))

    def _panic(msg):
        raise RustPanic(msg)

    # ---- raw pointers
    reg("ptr", "add", lambda it, recv, r, a, t: Ptr(r.base, r.off + int(deref(a[0]))))
    reg("ptr", "offset", lambda it, recv, r, a, t: Ptr(r.base, r.off + int(deref(a[0]))))
    reg("ptr", "sub", lambda it, recv, r, a, t: Ptr(r.base, r.off - int(deref(a[0]))))
    reg("ptr", "read", lambda it, recv, r, a, t: seq_get(r.base, r.off))

    # ---- anything
    reg("any", "clone into to_owned borrow borrow_mut as_ref as_mut", lambda it, recv, r, a, t: copyval(r))
    reg("tuple", "clone into", lambda it, recv, r, a, t: copyval(r))
    reg("enum", "clone into", lambda it, recv, r, a, t: r)
    reg("bool", "clone into", lambda it, recv, r, a, t: r)

    # ---- free functions / paths
    F = I.builtin_fns

    def from_raw_parts(it, a):
        p, n = deref(a[0]), int(deref(a[1]))
        return View(p.base, p.off, p.off + n)
    F["from_raw_parts"] = from_raw_parts
    F["from_raw_parts_mut"] = from_raw_parts
    I.builtin_paths["slice::from_raw_parts"] = from_raw_parts
    I.builtin_paths["slice::from_raw_parts_mut"] = from_raw_parts
    F["size_of_val"] = lambda it, a: Int(INT_TYPES[it.resolve_ty(deref(a[0]).ty or "i32")][0] // 8, "usize")
    I.builtin_paths["mem::size_of_val"] = F["size_of_val"]
    I.builtin_paths["PhantomData"] = Struct("PhantomData", {})
    F["drop"] = lambda it, a: UNIT
    I.builtin_paths["cmp::min"] = lambda it, a: B[("int", "min")](it, None, deref(a[0]), [a[1]], None)
    I.builtin_paths["cmp::max"] = lambda it, a: B[("int", "max")](it, None, deref(a[0]), [a[1]], None)
    F["min"] = I.builtin_paths["cmp::min"]
    F["max"] = I.builtin_paths["cmp::max"]


# ------------------------------------------------------------------------------------------ conveniences ----
def expand_macro_rules(text, macro_name, invocations):
    """macro_rules! NAME { (pattern) => { body } } with one arm: returns the body text once per invocation with
    the `$x` metavariables substituted (`$($opt)?` groups take the argument or nothing).  `invocations` is a list
    of dicts metavariable -> replacement text."""
    m = re.search(r"macro_rules!\s*" + macro_name + r"\s*\{", text)
    if not m:
        raise ValueError("macro %s not found" % macro_name)
    i = text.index("=>", m.end())
    j = text.index("{", i)
    d, k = 0, j
    while True:
        if text[k] == "{":
            d += 1
        elif text[k] == "}":
            d -= 1
            if d == 0:
                break
        k += 1
    body = text[j + 1:k]
    out = []
    for inv in invocations:
        b = body
        b = re.sub(r"\$\(\s*\$(\w+)\s*\)\?", lambda mm: inv.get(mm.group(1), ""), b)
        b = re.sub(r"\$(\w+)", lambda mm: inv.get(mm.group(1), ""), b)
        out.append(b)
    return "\n".join(out)
