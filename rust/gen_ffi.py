#!/usr/bin/env python
"""Generates rust/brush-b200-sys/src/ffi.rs -- the `extern "C"` block, `#[repr(C)]` structs and constants of
include/brush_b200.h -- so the Rust binding cannot drift from the header (tests/test_abi_host.py re-runs this and
diffs the result against the committed file, and checks every prototype of the header against it).

    python rust/gen_ffi.py            # rewrite ffi.rs
    python rust/gen_ffi.py --check    # exit 1 if ffi.rs is stale

No Rust toolchain exists in the build image, so nothing here is compiled; the generator is a plain C-declaration
parser for the subset of C the header uses (typedef structs, one anonymous enum, #define constants, prototypes)."""
from __future__ import annotations

import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(HERE, "..", "include", "brush_b200.h")
OUT = os.path.join(HERE, "brush-b200-sys", "src", "ffi.rs")

SCALARS = {"float": "f32", "double": "f64", "int": "c_int", "uint32_t": "u32", "int32_t": "i32", "uint64_t": "u64",
           "int64_t": "i64", "uint8_t": "u8", "char": "c_char", "void": "c_void", "size_t": "usize"}


def strip_comments(src: str) -> str:
    return re.sub(r"/\*.*?\*/", " ", src, flags=re.S)


def rust_type(base: str, const: bool, stars: int, structs) -> str:
    base = base.strip()
    t = SCALARS.get(base) or (base if base in structs or base == "brush_ctx" else None)
    if t is None:
        raise ValueError(f"unknown C type {base!r}")
    for level in range(stars):
        # `const T*` makes the innermost pointer const; outer levels are mutable
        t = ("*const " if (const and level == 0) else "*mut ") + t
    return t


def parse_decl_list(decl: str, consts, structs):
    """`const float* a`, `float *a, *b`, `uint32_t a, b[3]`, `float* rows[N]` -> [(name, rust type)]"""
    decl = decl.strip()
    const = bool(re.match(r"^const\b", decl))
    decl = re.sub(r"^const\s+", "", decl)
    m = re.match(r"^(struct\s+)?([A-Za-z_][A-Za-z_0-9]*)\s*(.*)$", decl, flags=re.S)
    base, rest = m.group(2), m.group(3)
    out = []
    for part in rest.split(","):
        part = part.strip()
        pm = re.match(r"^(\**)\s*([A-Za-z_][A-Za-z_0-9]*)?\s*(\[\s*([A-Za-z_0-9]+)\s*\])?$", part)
        if not pm:
            raise ValueError(f"cannot parse declarator {part!r} in {decl!r}")
        stars, name, _, dim = pm.group(1), pm.group(2), pm.group(3), pm.group(4)
        t = rust_type(base, const, len(stars), structs)
        if dim is not None:
            n = consts.get(dim, dim)
            t = f"[{t}; {n}]"
        out.append((name, t))
    return out


def parse(src: str):
    src = strip_comments(src)
    consts = {}
    for m in re.finditer(r"^#define\s+(BRUSH_[A-Z_0-9]+)\s+(.+)$", src, flags=re.M):
        name, val = m.group(1), m.group(2).strip()
        if name == "BRUSH_B200_H":
            continue
        consts[name] = val
    enum_vals = []
    em = re.search(r"enum\s*\{(.*?)\}\s*;", src, flags=re.S)
    if em:
        nxt = 0
        for item in em.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            if "=" in item:
                nm, v = [x.strip() for x in item.split("=")]
                nxt = int(v, 0)
            else:
                nm = item
            enum_vals.append((nm, nxt))
            nxt += 1
    struct_names = set(re.findall(r"typedef\s+struct\s+([A-Za-z_0-9]+)\s*\{", src))
    structs = []
    clean_consts = {}
    for k, v in consts.items():
        vv = v.strip("()").rstrip("u")
        clean_consts[k] = vv
    for m in re.finditer(r"typedef\s+struct\s+([A-Za-z_0-9]+)\s*\{(.*?)\}\s*\1\s*;", src, flags=re.S):
        fields = []
        for stmt in m.group(2).split(";"):
            if stmt.strip():
                fields += parse_decl_list(stmt, clean_consts, struct_names)
        structs.append((m.group(1), fields))
    body = src.replace("typedef struct brush_ctx brush_ctx;", " ")
    body = re.sub(r"typedef\s+struct\s+[A-Za-z_0-9]+\s*\{.*?\}\s*[A-Za-z_0-9]+\s*;", " ", body, flags=re.S)
    body = re.sub(r"enum\s*\{.*?\}\s*;", " ", body, flags=re.S)
    body = re.sub(r"^#.*$", " ", body, flags=re.M)
    body = body.replace('extern "C" {', " ")
    funcs = []
    for m in re.finditer(r"([A-Za-z_][A-Za-z_0-9\s\*]*?)\b(brush_[a-z_0-9]+)\s*\((.*?)\)\s*;", body, flags=re.S):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        rconst = ret.startswith("const ")
        rbase = re.sub(r"^const\s+", "", ret)
        stars = rbase.count("*")
        rbase = rbase.replace("*", "").strip()
        rtype = None if (rbase == "void" and stars == 0) else rust_type(rbase, rconst, stars, struct_names)
        params = []
        if args and args != "void":
            for a in re.split(r",(?![^\[]*\])", args):
                (nm, t), = parse_decl_list(" ".join(a.split()), clean_consts, struct_names)
                params.append((nm, t))
        funcs.append((name, params, rtype))
    return consts, enum_vals, structs, funcs


def rust_ident(name: str) -> str:
    return {"in": "input", "type": "ty", "ref": "reference"}.get(name, name)


def generate() -> str:
    with open(HEADER) as f:
        consts, enum_vals, structs, funcs = parse(f.read())
    o = ["// @generated by rust/gen_ffi.py from include/brush_b200.h -- DO NOT EDIT (tests/test_abi_host.py diffs it).",
         "// Raw FFI surface of libbrush_b200.so: every pointer is a DEVICE pointer unless its name ends in `_host`;",
         "// `stream` is a cudaStream_t passed as *mut c_void.",
         "#![allow(non_camel_case_types, non_snake_case, dead_code)]",
         "use std::os::raw::{c_char, c_int, c_void};",
         ""]
    for name, val in consts.items():
        v = val.strip()
        m = re.match(r"^\(?(-?\d+)\)?u?$", v)
        if not m:
            continue
        iv = int(m.group(1))
        ty = "c_int" if name.startswith("BRUSH_E_") or name == "BRUSH_OK" else ("usize" if name == "BRUSH_DP_MAX_WORLD" else "u32")
        o.append(f"pub const {name}: {ty} = {iv};")
    o.append("")
    for nm, v in enum_vals:
        o.append(f"pub const {nm}: usize = {v};")
    o += ["", "#[repr(C)]", "pub struct brush_ctx {", "    _opaque: [u8; 0],", "}", ""]
    for name, fields in structs:
        o += ["#[repr(C)]", "#[derive(Clone, Copy)]", f"pub struct {name} {{"]
        for fn_, ft in fields:
            o.append(f"    pub {rust_ident(fn_)}: {ft},")
        o += ["}", ""]
    o += ['#[link(name = "brush_b200")]', 'extern "C" {']
    for name, params, rtype in funcs:
        ps = ", ".join(f"{rust_ident(n)}: {t}" for n, t in params)
        line = f"    pub fn {name}({ps})" + (f" -> {rtype}" if rtype else "") + ";"
        if len(line) > 118:
            line = f"    pub fn {name}(\n" + "".join(f"        {rust_ident(n)}: {t},\n" for n, t in params) + "    )" + \
                   (f" -> {rtype}" if rtype else "") + ";"
        o.append(line)
    o += ["}", ""]
    return "\n".join(o)


def main():
    text = generate()
    if "--check" in sys.argv:
        cur = open(OUT).read() if os.path.exists(OUT) else ""
        if cur != text:
            print("rust/brush-b200-sys/src/ffi.rs is stale: run `python rust/gen_ffi.py`", file=sys.stderr)
            sys.exit(1)
        return
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    with open(OUT, "w") as f:
        f.write(text)
    print(OUT)


if __name__ == "__main__":
    main()
