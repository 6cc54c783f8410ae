#!/usr/bin/env python3
"""Rust `extern "C"` declarations for every function include/bn254_b200.h declares (the block INTEGRATION.md section 3 shows).
Generated so that it cannot drift from the header; tests/test_abi_exports.py checks INTEGRATION.md against the header.
usage: gen_rust_extern.py            print the block
       gen_rust_extern.py --update   rewrite the block between the BEGIN/END markers of INTEGRATION.md"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BEGIN, END = "// BEGIN GENERATED extern block (tools/gen_rust_extern.py)", "// END GENERATED extern block"

TYPES = {
    "int": "i32", "size_t": "usize", "uint64_t": "u64", "uint32_t": "u32", "int32_t": "i32",
    "void*": "*mut c_void", "const void*": "*const c_void", "uint8_t*": "*mut u8", "const uint8_t*": "*const u8",
    "float*": "*mut f32", "double*": "*mut f64", "const char*": "*const c_char",
    "bn254_ctx*": "*mut BnCtx", "const bn254_ctx*": "*const BnCtx", "bn254_ctx**": "*mut *mut BnCtx",
    "bn254_g2precomp*": "*mut BnG2Precomp", "const bn254_g2precomp*": "*const BnG2Precomp", "bn254_g2precomp**": "*mut *mut BnG2Precomp",
    "bn254_ctx_options*": "*mut BnCtxOptions", "const bn254_ctx_options*": "*const BnCtxOptions",
}


def prototypes():
    src = open(os.path.join(ROOT, "include", "bn254_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"^\s*#.*$", "", src, flags=re.M)
    src = re.sub(r"typedef struct bn254_ctx_options \{.*?\} bn254_ctx_options;", "", src, flags=re.S)
    out = []
    for m in re.finditer(r"([A-Za-z_][A-Za-z0-9_ ]*?[\s\*]+)(bn254_[a-z0-9_]+)\s*\(([^;{}]*?)\)\s*;", src, flags=re.S):
        ret, name, args = " ".join(m.group(1).split()), m.group(2), " ".join(m.group(3).split())
        out.append((ret.replace(" *", "*"), name, args))
    return out


def rust_type(t):
    t = " ".join(t.replace("*", "* ").split()).replace(" *", "*").replace("* ", "*").strip()
    t = re.sub(r"\s+\*", "*", t)
    if t not in TYPES:
        raise SystemExit("no Rust mapping for C type %r" % t)
    return TYPES[t]


def block():
    lines = ["extern \"C\" {"]
    for ret, name, args in prototypes():
        params = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                m = re.match(r"(.*?)([A-Za-z_][A-Za-z0-9_]*)$", a)
                ctype, pname = m.group(1).strip(), m.group(2)
                if pname in ("fn", "in", "type", "ref", "move"):
                    pname += "_"
                params.append("%s: %s" % (pname, rust_type(ctype)))
        r = "" if ret == "void" else " -> %s" % rust_type(ret)
        lines.append("    pub fn %s(%s)%s;" % (name, ", ".join(params), r))
    lines.append("}")
    return "\n".join(lines)


if __name__ == "__main__":
    b = block()
    if "--update" in sys.argv:
        p = os.path.join(ROOT, "INTEGRATION.md")
        s = open(p).read()
        i, j = s.index(BEGIN), s.index(END)
        s = s[:i] + BEGIN + "\n" + b + "\n" + s[j:]
        open(p, "w").write(s)
        print("INTEGRATION.md updated: %d functions" % (len(b.splitlines()) - 2))
    else:
        print(b)
