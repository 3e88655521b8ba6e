#!/usr/bin/env python
"""gen_rust_sys.py — generate rust/rtx-cuda-sys/src/lib.rs from include/rtx_cuda.h, 1:1.

The reference is a Rust crate (Cargo.toml:1-13, src/lib.rs:1-5); its toolchain is absent from this image, so the
FFI crate ships as source.  To keep that source from drifting behind the C header (round 1: 37 of 56 functions),
it is GENERATED: every function, struct, enum constant and numeric #define of the header becomes its Rust
declaration, with compile-time size assertions computed here from the C layout rules.

  python tools/gen_rust_sys.py            writes lib.rs
  python tools/gen_rust_sys.py --check    exits 1 if lib.rs is not what the header generates (tests/test_rust_crate.py)

parse_header() is also what the CPU tests use to compare the header with the shared library's exports.
"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "rtx_cuda.h")
LIB_RS = os.path.join(ROOT, "rust-raytracing_b200", "rust", "rtx-cuda-sys", "src", "lib.rs")

SCALARS = {"double": ("f64", 8), "float": ("f32", 4), "uint64_t": ("u64", 8), "int64_t": ("i64", 8),
           "uint32_t": ("u32", 4), "int32_t": ("i32", 4), "uint8_t": ("u8", 1), "size_t": ("usize", 8),
           "char": ("c_char", 1), "void": ("c_void", 0)}


def strip_comments(text):
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    return re.sub(r"//[^\n]*", " ", text)


def parse_header(path=HEADER):
    """-> dict(defines=[(name, value)], enums=[(name, value)], structs=[(name, [(ctype, field, array_len)])],
    opaque=[name], aliases=[(name, ctype)], functions=[(ret_ctype, name, [(ctype, argname)])])"""
    text = strip_comments(open(path).read())
    out = dict(defines=[], enums=[], structs=[], opaque=[], aliases=[], functions=[])
    for m in re.finditer(r"^[ \t]*#define[ \t]+(RTX_\w+)[ \t]+(\d+)[uU]?[ \t]*$", text, flags=re.M):
        out["defines"].append((m.group(1), int(m.group(2))))
    body = re.sub(r"^[ \t]*#.*$", " ", text, flags=re.M)  # preprocessor lines are done
    body = body.replace('extern "C" {', " ")
    for m in re.finditer(r"\benum\s*\{(.*?)\}\s*;", body, flags=re.S):
        for item in m.group(1).split(","):
            item = item.strip()
            if not item:
                continue
            name, value = (s.strip() for s in item.split("="))
            out["enums"].append((name, int(value.rstrip("uU"), 0)))
    body = re.sub(r"\benum\s*\{.*?\}\s*;", " ", body, flags=re.S)
    for m in re.finditer(r"typedef\s+struct\s*\{(.*?)\}\s*(\w+)\s*;", body, flags=re.S):
        fields = []
        for decl in m.group(1).split(";"):
            decl = " ".join(decl.split())
            if not decl:
                continue
            mm = re.match(r"((?:const\s+)?\w+\s*\**)\s*(.*)$", decl)
            ctype, names = mm.group(1).replace(" *", "*").strip(), mm.group(2)
            for nm in names.split(","):
                nm = nm.strip()
                stars = len(nm) - len(nm.lstrip("*"))
                nm = nm.lstrip("* ")
                arr = re.match(r"(\w+)\[(\d+)\]$", nm)
                fields.append((ctype + "*" * stars, arr.group(1) if arr else nm, int(arr.group(2)) if arr else 0))
        out["structs"].append((m.group(2), fields))
    body = re.sub(r"typedef\s+struct\s*\{.*?\}\s*\w+\s*;", " ", body, flags=re.S)
    for m in re.finditer(r"typedef\s+struct\s+(\w+)\s+(\w+)\s*;", body):
        out["opaque"].append(m.group(2))
    body = re.sub(r"typedef\s+struct\s+\w+\s+\w+\s*;", " ", body)
    for m in re.finditer(r"typedef\s+(\w+)\s+(\w+)\s*;", body):
        out["aliases"].append((m.group(2), m.group(1)))
    body = re.sub(r"typedef\s+\w+\s+\w+\s*;", " ", body)
    for m in re.finditer(r"([\w\s\*]+?)\b(rtx_\w+)\s*\(([^()]*)\)\s*;", body):
        ret = " ".join(m.group(1).split()).replace(" *", "*")
        args = []
        a = " ".join(m.group(3).split())
        if a and a != "void":
            for arg in a.split(","):
                arg = arg.strip()
                mm = re.match(r"(.*?)(\w+)$", arg)
                args.append((mm.group(1).strip().replace(" *", "*").replace("* ", "*"), mm.group(2)))
        out["functions"].append((ret, m.group(2), args))
    return out


def rust_type(ctype):
    c = ctype.strip()
    const = c.startswith("const ")
    if const:
        c = c[len("const "):].strip()
    stars = len(c) - len(c.rstrip("*"))
    base = c.rstrip("*").strip()
    t = SCALARS[base][0] if base in SCALARS else base
    for i in range(stars):
        t = ("*const " if (const and i == 0) else "*mut ") + t
    return t


def layout(h):
    """sizes/alignments of the header's structs by the C rules (LP64)."""
    sz = {k: (v[1], v[1] or 1) for k, v in SCALARS.items()}
    for name, ctype in h["aliases"]:
        sz[name] = sz[ctype]
    for name, fields in h["structs"]:
        off, al = 0, 1
        for ctype, _, arr in fields:
            s, a = (8, 8) if "*" in ctype else sz[ctype.replace("const ", "").strip()]
            off = (off + a - 1) // a * a + s * (arr or 1)
            al = max(al, a)
        sz[name] = ((off + al - 1) // al * al, al)
    return sz


def generate(h):
    sz = layout(h)
    L = []
    w = L.append
    w("//! rtx-cuda-sys — `extern \"C\"` declarations for include/rtx_cuda.h with `#[repr(C)]` mirrors of the `rtx`")
    w("//! crate's types (math/vector.rs:12-20, math/mat.rs:10-18, raytracing/camera.rs:7-15, scene.rs:16-28,")
    w("//! object.rs:79-89, object/{sphere,plane,triangle}.rs:9-12).")
    w("//! GENERATED by tools/gen_rust_sys.py from the header — do not edit; tests/test_rust_crate.py fails when the two")
    w("//! drift apart.  UNBUILT IN THIS ENVIRONMENT (no cargo/rustc); layouts are asserted at compile time.")
    w("#![allow(non_camel_case_types)]")
    w("use std::os::raw::{c_char, c_void};")
    w("")
    for name, ctype in h["aliases"]:
        w(f"pub type {name} = {rust_type(ctype)};")
    w("")
    for name, fields in h["structs"]:
        ptr = any("*" in f[0] for f in fields)
        w(f"#[repr(C)] #[derive(Clone, Copy, Debug{'' if ptr else ', PartialEq'})]")
        w(f"pub struct {name} {{")
        for ctype, fname, arr in fields:
            t = rust_type(ctype)
            w(f"    pub {fname}: {f'[{t}; {arr}]' if arr else t},")
        w("}")
        w(f"const _: () = assert!(std::mem::size_of::<{name}>() == {sz[name][0]});")
    for name in h["opaque"]:
        w(f"#[repr(C)] pub struct {name} {{ _private: [u8; 0] }}")
    w("")
    for name, value in h["defines"]:
        w(f"pub const {name}: {'usize' if name.endswith('_BYTES') else 'u32'} = {value};")
    for name, value in h["enums"]:
        t = "rtx_status" if (name == "RTX_OK" or name.startswith("RTX_ERR_")) else "u32"
        w(f"pub const {name}: {t} = {value};")
    w("")
    w('extern "C" {')
    for ret, name, args in h["functions"]:
        a = ", ".join(f"{an}: {rust_type(ct)}" for ct, an in args)
        r = "" if ret == "void" else f" -> {rust_type(ret)}"
        w(f"    pub fn {name}({a}){r};")
    w("}")
    return "\n".join(L) + "\n"


def main():
    text = generate(parse_header())
    if "--check" in sys.argv:
        cur = open(LIB_RS).read() if os.path.exists(LIB_RS) else ""
        if cur != text:
            print("lib.rs is out of date with include/rtx_cuda.h: run python tools/gen_rust_sys.py", file=sys.stderr)
            sys.exit(1)
        return
    open(LIB_RS, "w").write(text)
    print(f"wrote {LIB_RS}")


if __name__ == "__main__":
    main()
