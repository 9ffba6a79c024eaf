"""Generates the Rust side of the drop-in boundary from include/rav1e_cuda.h:

    python tools/gen_rust_ffi.py            -> writes bindings/rust/rav1e_cuda_sys.rs
    python tools/gen_rust_ffi.py --check    -> exit 1 if the committed file is not what the header gives

The output is the `rav1e-cuda-sys` crate's lib.rs a maintainer of the reference would add (INTEGRATION.md section 1
shows how the encoder calls it): every `#define`d constant, enum, `#[repr(C)]` struct and `extern "C"` function of the
header, nothing else.  There is no Rust toolchain in the build image, so the file is generated mechanically from the C
declarations (pycparser) instead of by bindgen, and tests/test_boundary_cpu.py keeps it in step with the header and with
the symbols the shared library exports."""
import os
import re
import sys

from pycparser import c_ast, c_parser

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rav1e_cuda.h")
OUT = os.path.join(ROOT, "bindings", "rust", "rav1e_cuda_sys.rs")

PRELUDE = """
typedef signed char int8_t; typedef unsigned char uint8_t; typedef short int16_t; typedef unsigned short uint16_t;
typedef int int32_t; typedef unsigned int uint32_t; typedef long long int64_t; typedef unsigned long long uint64_t;
typedef unsigned long size_t;
"""
SCALARS = {
    "int8_t": "i8", "uint8_t": "u8", "int16_t": "i16", "uint16_t": "u16", "int32_t": "i32", "uint32_t": "u32",
    "int64_t": "i64", "uint64_t": "u64", "size_t": "usize", "int": "c_int", "unsigned int": "c_uint", "unsigned": "c_uint",
    "char": "c_char", "float": "f32", "double": "f64", "void": "c_void", "long long": "c_longlong",
    "unsigned long long": "c_ulonglong", "short": "c_short", "unsigned char": "c_uchar", "signed char": "c_schar",
}
RUST_KEYWORDS = {"ref", "type", "in", "box", "move", "fn", "mod", "use", "loop", "match", "self", "impl", "where", "as"}


def strip(text):
    """the header without comments, includes and preprocessor conditionals (the extern "C" guard goes with them)"""
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    keep = []
    for line in text.splitlines():
        s = line.strip()
        if s.startswith("#") or s == 'extern "C" {' or s == "}":
            continue
        keep.append(line)
    return "\n".join(keep)


def defines(text):
    out = []
    for m in re.finditer(r"^#define\s+(R1C_[A-Z0-9_]+)\s+(.+?)\s*(?:/\*.*)?$", text, flags=re.M):
        name, val = m.group(1), m.group(2).strip()
        unsigned = bool(re.search(r"\d[uU]\b", val))
        expr = re.sub(r"(\d)[uUlL]+\b", r"\1", val)
        if not re.fullmatch(r"[0-9xXa-fA-F\s()<>+\-*|&]+", expr):
            continue
        v = eval(expr, {"__builtins__": {}})  # integer literals and shifts only (checked above)
        out.append((name, "u32" if unsigned else "c_int", v))
    return out


def name_of(n):
    return "r#" + n if n in RUST_KEYWORDS else n


class Gen:
    def __init__(self):
        self.opaque, self.structs, self.enums, self.funcs = [], [], [], []
        self.known = set()

    def rtype(self, t, const=False):
        """Rust spelling of a C declarator type"""
        if isinstance(t, c_ast.PtrDecl):
            inner = t.type
            is_const = "const" in getattr(inner, "quals", []) or (isinstance(inner, c_ast.TypeDecl) and "const" in inner.quals)
            if isinstance(inner, c_ast.FuncDecl):
                raise NotImplementedError("function pointers are not part of this header")
            return ("*const " if is_const else "*mut ") + self.rtype(inner)
        if isinstance(t, c_ast.ArrayDecl):
            return "[%s; %s]" % (self.rtype(t.type), self.const_expr(t.dim))
        if isinstance(t, c_ast.TypeDecl):
            tt = t.type
            if isinstance(tt, c_ast.IdentifierType):
                nm = " ".join(tt.names)
                if nm in SCALARS:
                    return SCALARS[nm]
                if nm in self.known:
                    return nm
                raise KeyError("unknown type " + nm)
            if isinstance(tt, c_ast.Struct):
                return tt.name
            if isinstance(tt, c_ast.Enum):
                return "c_int"
        raise NotImplementedError(type(t))

    def const_expr(self, e):
        if isinstance(e, c_ast.Constant):
            return re.sub(r"[uUlL]+$", "", e.value)
        if isinstance(e, c_ast.ID):
            return e.name
        if isinstance(e, c_ast.BinaryOp):
            return "(%s %s %s)" % (self.const_expr(e.left), e.op, self.const_expr(e.right))
        raise NotImplementedError(type(e))

    def visit(self, ast):
        for ext in ast.ext:
            if isinstance(ext, c_ast.Typedef):
                t = ext.type.type if isinstance(ext.type, c_ast.TypeDecl) else None
                if isinstance(t, c_ast.Struct):
                    if t.decls is None:
                        self.opaque.append(ext.name)
                    else:
                        fields = []
                        for d in t.decls:
                            fields.append((name_of(d.name), self.rtype(d.type)))
                        self.structs.append((ext.name, fields))
                    self.known.add(ext.name)
                elif isinstance(t, c_ast.Enum):
                    vals, nxt = [], 0
                    for en in t.values.enumerators:
                        if en.value is not None:
                            nxt = int(self.const_expr(en.value), 0)
                        vals.append((en.name, nxt))
                        nxt += 1
                    self.enums.append((ext.name, vals))
                    self.known.add(ext.name)
                elif ext.name not in ("int8_t", "uint8_t", "int16_t", "uint16_t", "int32_t", "uint32_t", "int64_t", "uint64_t", "size_t"):
                    raise NotImplementedError("typedef " + ext.name)
            elif isinstance(ext, c_ast.Decl) and isinstance(ext.type, c_ast.FuncDecl):
                fd = ext.type
                args = []
                if fd.args:
                    for i, p in enumerate(fd.args.params):
                        if isinstance(p, c_ast.Typename) or p.name is None:
                            if isinstance(p.type, c_ast.TypeDecl) and isinstance(p.type.type, c_ast.IdentifierType) and p.type.type.names == ["void"]:
                                continue
                            args.append(("arg%d" % i, self.rtype(p.type)))
                        elif isinstance(p.type, c_ast.ArrayDecl):  # `T out[4]` is a pointer in a parameter list
                            q = p.type.type
                            const = isinstance(q, c_ast.TypeDecl) and "const" in q.quals
                            args.append((name_of(p.name), ("*const " if const else "*mut ") + self.rtype(q)))
                        else:
                            args.append((name_of(p.name), self.rtype(p.type)))
                ret = self.rtype(fd.type)
                self.funcs.append((ext.name, args, None if ret == "c_void" else ret))

    def enum_type(self, name):
        return name


def generate():
    raw = open(HEADER).read()
    g = Gen()
    g.known.update(["r1c_ctx", "r1c_plane"])
    ast = c_parser.CParser().parse(PRELUDE + strip(raw), filename="rav1e_cuda.h")
    g.visit(ast)
    enum_names = {n for n, _ in g.enums}
    o = []
    o.append("// rav1e_cuda_sys.rs -- GENERATED from include/rav1e_cuda.h by tools/gen_rust_ffi.py; do not edit.")
    o.append("// The `extern \"C\"` surface of librav1e_cuda.so as the `rav1e-cuda-sys` crate of INTEGRATION.md section 1.")
    o.append("#![allow(non_camel_case_types, non_upper_case_globals, dead_code)]")
    o.append("use std::os::raw::{c_char, c_int, c_longlong, c_uint, c_void};")
    o.append("")
    for name, ty, v in defines(raw):
        o.append("pub const %s: %s = %d;" % (name, ty, v))
    o.append("")
    for name, vals in g.enums:
        o.append("pub type %s = c_int;" % name)
        for en, v in vals:
            o.append("pub const %s: %s = %d;" % (en, name, v))
        o.append("")
    for name in g.opaque:
        o.append("#[repr(C)] pub struct %s { _private: [u8; 0] }" % name)
    o.append("")
    for name, fields in g.structs:
        o.append("#[repr(C)] #[derive(Clone, Copy)]")
        o.append("pub struct %s {" % name)
        for fn, ft in fields:
            o.append("  pub %s: %s," % (fn, ft))
        o.append("}")
        o.append("")
    o.append('#[link(name = "rav1e_cuda")]')
    o.append('extern "C" {')
    for name, args, ret in g.funcs:
        sig = "  pub fn %s(%s)" % (name, ", ".join("%s: %s" % a for a in args))
        if ret:
            sig += " -> " + ret
        o.append(sig + ";")
    o.append("}")
    text = "\n".join(o) + "\n"
    unused = [t for t in ("c_char", "c_longlong", "c_uint") if not re.search(r"\b%s\b" % t, text.split("use std::os::raw", 1)[1].split("\n", 1)[1])]
    if unused:
        imports = [t for t in ("c_char", "c_int", "c_longlong", "c_uint", "c_void") if t not in unused]
        text = text.replace("use std::os::raw::{c_char, c_int, c_longlong, c_uint, c_void};", "use std::os::raw::{%s};" % ", ".join(imports))
    return text, g


def main():
    text, g = generate()
    if "--check" in sys.argv:
        have = open(OUT).read() if os.path.exists(OUT) else ""
        if have != text:
            print("bindings/rust/rav1e_cuda_sys.rs is out of date: run python tools/gen_rust_ffi.py")
            return 1
        return 0
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    open(OUT, "w").write(text)
    print("%s: %d functions, %d structs, %d enums" % (os.path.relpath(OUT, ROOT), len(g.funcs), len(g.structs), len(g.enums)))
    return 0


if __name__ == "__main__":
    sys.exit(main())
