"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports
every symbol include/rav1e_cuda.h declares, the Python struct mirrors match the
C layout, and the pure-host helpers agree with the oracle.  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    from rav1e_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        g.build()
    return capi.load()


def test_every_header_symbol_is_exported(lib):
    from rav1e_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "rav1e_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(r1c_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations found"
    assert declared == set(capi.SIGNATURES), declared ^ set(capi.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name


def test_header_is_plain_c99(lib, tmp_path):
    """the boundary is a C header: it compiles as strict C99 (and as C++11) with every warning an error, and a C
    program linked against the library can call it"""
    import subprocess
    from rav1e_b200 import capi
    src = tmp_path / "c99.c"
    src.write_text('#include "rav1e_cuda.h"\n'
                   'int main(void) { r1c_frame_me_params p; r1c_quant_ctx q; (void)p; (void)q;\n'
                   '  return r1c_block_width(R1C_BLOCK_8X8) == 8 && r1c_tx_height(R1C_TX_16X64) == 64 ? 0 : 1; }\n')
    libdir = os.path.dirname(capi.LIB_PATH)
    exe = str(tmp_path / "c99")
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                           str(src), "-L", libdir, "-lrav1e_cuda", "-Wl,-rpath," + libdir, "-o", exe])
    assert subprocess.call([exe]) == 0
    subprocess.check_call(["g++", "-std=c++11", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c++",
                           "-I", os.path.join(ROOT, "include"), str(src)])


def test_rust_binding_is_in_step_with_the_header(lib):
    """bindings/rust/rav1e_cuda_sys.rs (the crate INTEGRATION.md section 1 describes; no Rust toolchain here, so it is
    generated from the header by tools/gen_rust_ffi.py): regenerating gives the committed text, it binds exactly the
    functions the header declares and the library exports, and its #[repr(C)] structs have the fields, in order, of
    the ctypes mirrors the GPU tests call through (whose layout test_struct_layouts_match pins)."""
    import subprocess
    import sys
    from rav1e_b200 import capi
    assert subprocess.call([sys.executable, os.path.join(ROOT, "tools", "gen_rust_ffi.py"), "--check"]) == 0
    rs = open(os.path.join(ROOT, "bindings", "rust", "rav1e_cuda_sys.rs")).read()
    fns = set(re.findall(r"pub fn (r1c_[a-z0-9_]+)\(", rs))
    assert fns == set(capi.SIGNATURES), fns ^ set(capi.SIGNATURES)
    for name in fns:
        assert hasattr(lib, name), name
    structs = {m.group(1): re.findall(r"pub (?:r#)?([a-z_0-9]+):", m.group(2))
               for m in re.finditer(r"pub struct (r1c_[a-z0-9_]+) \{\n(.*?)\n\}", rs, flags=re.S)}
    checked = 0
    for cname in dir(capi):
        ct = getattr(capi, cname)
        if isinstance(ct, type) and issubclass(ct, C.Structure) and ct is not C.Structure and hasattr(ct, "_fields_"):
            c_name = "r1c_" + re.sub(r"(?<!^)(?=[A-Z])", "_", cname).lower()  # FrameMeParams -> r1c_frame_me_params
            assert structs[c_name] == [f[0].rstrip("_") for f in ct._fields_], c_name
            checked += 1
    assert checked >= 15 and checked == len(structs)


def test_struct_layouts_match():
    from rav1e_b200 import api
    for npdt, ct in api.STRUCT_PAIRS:
        assert npdt.itemsize == C.sizeof(ct), (npdt, ct)
        for (fname, _), in zip(ct._fields_):
            pass
        for fname, _ in ct._fields_:
            assert npdt.fields[fname][1] == getattr(ct, fname).offset, (ct, fname)


def test_enums_and_tables(lib, oracle):
    L = oracle.lib()
    for b in range(22):
        assert lib.r1c_block_width(b) == L.r1o_block_width(b)
        assert lib.r1c_block_height(b) == L.r1o_block_height(b)
    assert lib.r1c_block_width(22) == -1
    for s in range(19):
        assert lib.r1c_tx_width(s) == L.r1o_tx_width(s)
        assert lib.r1c_tx_height(s) == L.r1o_tx_height(s)
        for t in range(16):
            assert bool(lib.r1c_fwd_txfm_valid(s, t)) == bool(L.r1o_fht_valid(s, t))
    assert sum(lib.r1c_sb_result_count(b) for b in range(22)) == 849
    assert lib.r1c_sb_result_offset(13) == -1  # BLOCK_64X128


def test_plane_cfg_matches_oracle(lib, oracle):
    from rav1e_b200 import capi
    for (w, h, xd, yd, xp, yp, pb) in [(640, 480, 0, 0, 136, 136, 1), (640, 480, 0, 0, 264, 264, 2),
                                        (1920, 1080, 0, 0, 88, 88, 1), (3840, 2160, 0, 0, 88, 88, 2),
                                        (1920, 1080, 1, 1, 44, 44, 2), (480, 270, 2, 2, 22, 22, 1)]:
        a = capi.PlaneCfg()
        assert lib.r1c_plane_cfg_new(C.byref(a), w, h, xd, yd, xp, yp, pb) == 0
        b = oracle.OPlane(w, h, xd, yd, xp, yp, pb).cfg
        for f, _ in capi.PlaneCfg._fields_:
            assert getattr(a, f) == getattr(b, f), f
    # 1080p 8-bit luma stride 2112, xorigin 96; 4K 10-bit stride 4032
    a = capi.PlaneCfg()
    lib.r1c_plane_cfg_new(C.byref(a), 1920, 1080, 0, 0, 88, 88, 1)
    assert (a.stride, a.xorigin, a.alloc_height) == (2112, 96, 1256)
    lib.r1c_plane_cfg_new(C.byref(a), 3840, 2160, 0, 0, 88, 88, 2)
    assert (a.stride, a.xorigin, a.alloc_height) == (4032, 96, 2336)


def test_host_helpers_match_oracle(lib, oracle):
    from rav1e_b200 import api, capi
    rng = np.random.default_rng(1)
    L = oracle.lib()
    for _ in range(500):
        a = (int(rng.integers(-2000, 2000)), int(rng.integers(-2000, 2000)))
        b = (int(rng.integers(-2000, 2000)), int(rng.integers(-2000, 2000)))
        for hp in (0, 1):
            assert api.get_mv_rate(a, b, hp) == L.r1o_mv_rate(oracle.Mv(*a), oracle.Mv(*b), hp)
    for _ in range(200):
        w_in_b, h_in_b = int(rng.integers(16, 1000)), int(rng.integers(16, 600))
        bo = (int(rng.integers(0, w_in_b)), int(rng.integers(0, h_in_b)))
        bs = int(rng.integers(0, 22))
        bw, bh = lib.r1c_block_width(bs), lib.r1c_block_height(bs)
        out = (C.c_int64 * 4)()
        L.r1o_mv_range(w_in_b, h_in_b, bo[0], bo[1], bw, bh, C.byref(out))
        assert api.get_mv_range(w_in_b, h_in_b, bo, bw, bh) == tuple(out)
    cfg = oracle.OPlane(64, 48, 0, 0, 16, 16, 1).cfg
    ccfg = capi.PlaneCfg()
    C.memmove(C.byref(ccfg), C.byref(cfg), C.sizeof(ccfg))
    for _ in range(500):
        po = (int(rng.integers(0, 64)), int(rng.integers(0, 48)))
        mv = (int(rng.integers(-900, 900)), int(rng.integers(-900, 900)))
        rf, cf, sx, sy = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
        L.r1o_mv_params(C.byref(cfg), po[0], po[1], mv[0], mv[1], C.byref(rf), C.byref(cf),
                        C.byref(sx), C.byref(sy))
        assert api.mv_params(ccfg, po, mv) == (rf.value, cf.value, sx.value, sy.value)


def test_no_context_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from rav1e_b200 import api
    with pytest.raises(api.Rav1eCudaError):
        api.Context(0)  # no CPU fallback: must fail loudly


def test_frame_me_layout_matches_oracle(lib):
    """pure host arithmetic: result layout of r1c_frame_me vs the oracle's"""
    from oracle import r1o
    from rav1e_b200 import api
    L = r1o.lib()
    for (w, h, mask) in [(1920, 1080, api.ALL_SIZES_TO_64), (200, 152, api.ALL_SIZES_TO_64), (64, 64, 0b1001)]:
        p = api.make_frame_me_params(w, h, 8, 8, 1, bsize_mask=mask)
        off, cols, rows = (C.c_int64 * 22)(), (C.c_int32 * 22)(), (C.c_int32 * 22)()
        total = L.r1o_frame_me_layout(w, h, mask, C.byref(off), C.byref(cols), C.byref(rows))
        assert api.frame_me_count(p) == total
        for b in range(22):
            o, c, r_ = api.frame_me_layout(p, b)
            assert (o, c, r_) == (off[b], cols[b], rows[b]), b
    assert api.frame_me_count(api.make_frame_me_params(1922, 1080, 8, 8, 1)) == -1
    assert api.frame_me_count(api.make_frame_me_params(64, 64, 8, 8, 1, bsize_mask=0)) == -1
