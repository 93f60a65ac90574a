"""CPU: the Rust shim shipped under integration/ (quantization/b200_w4a16.rs + build.rs) cannot drift
from the C header.  There is no Rust toolchain in this image, so the files are not compiled; what is checked:
  * integration/b200_w4a16_ffi.rs is exactly what tools/gen_rust_ffi.py generates from include/w4a16_b200.h
    (every prototype, every parameter type);
  * the generated block declares every symbol the header declares and the library exports, once;
  * every `ffi::w4a16_*` call in integration/b200_w4a16.rs names a declared function and passes as many arguments as
    the declaration has parameters."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ffi_block_is_generated_from_the_header():
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "gen_rust_ffi.py"), "--check"])


def _decls():
    src = open(os.path.join(ROOT, "integration", "b200_w4a16_ffi.rs")).read()
    return {m.group(1): [a for a in m.group(2).split(",") if a.strip()]
            for m in re.finditer(r"pub fn (w4a16_\w+)\((.*?)\)", src)}


def test_ffi_block_covers_every_header_symbol():
    from test_abi import header_symbols
    decls = _decls()
    assert sorted(decls) == header_symbols(os.path.join(ROOT, "include", "w4a16_b200.h"))


def _split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur)
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def test_shim_calls_match_the_declarations():
    decls = _decls()
    src = open(os.path.join(ROOT, "integration", "b200_w4a16.rs")).read()
    calls = 0
    for m in re.finditer(r"ffi::(w4a16_\w+)\(", src):
        name = m.group(1)
        assert name in decls, f"{name} is not declared in b200_w4a16_ffi.rs"
        i, depth = m.end(), 1
        while depth:
            depth += {"(": 1, ")": -1}.get(src[i], 0)
            i += 1
        args = _split_args(src[m.end():i - 1])
        assert len(args) == len(decls[name]), f"{name}: {len(args)} arguments passed, {len(decls[name])} declared"
        calls += 1
    assert calls >= 20
    assert "impl QuantizedLinear for B200W4A16Linear" in src
    build = open(os.path.join(ROOT, "integration", "build.rs")).read()
    assert "rustc-link-lib=dylib=w4a16_b200" in build
