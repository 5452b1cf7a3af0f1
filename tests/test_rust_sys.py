"""rust/minipath-b200-sys (the FFI crate the north_star asks for; no cargo in this image, so it cannot be compiled here) is kept
in step with include/minipath_b200.h mechanically: same function set and argument counts, same constants, and every
#[repr(C)] struct has the field names, order and C-layout size of the ctypes mirror that the GPU tests exercise."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_RS = os.path.join(ROOT, "rust", "minipath-b200-sys", "src", "lib.rs")
HEADER = os.path.join(ROOT, "include", "minipath_b200.h")


def _strip_c(text):
    return re.sub(r"/\*.*?\*/", "", text, flags=re.S)


def _strip_rs(text):
    return re.sub(r"//[^\n]*", "", text)


def _split_args(s):
    s = s.strip()
    if s in ("", "void"):
        return []
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "(<[":
            depth += 1
        elif ch in ")>]":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def header_functions():
    text = _strip_c(open(HEADER).read())
    return {m.group(1): len(_split_args(m.group(2))) for m in re.finditer(r"\b(mp_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S)}


def rust_functions():
    text = _strip_rs(open(LIB_RS).read())
    block = re.search(r'extern "C" \{(.*?)\n\}', text, flags=re.S).group(1)
    return {m.group(1): len(_split_args(m.group(2))) for m in re.finditer(r"pub fn (mp_[a-z0-9_]+)\s*\(([^;]*?)\)\s*(?:->[^;]*)?;", block, flags=re.S)}


def test_every_header_function_is_bound_with_the_same_arity():
    h, r = header_functions(), rust_functions()
    assert len(h) > 60
    assert sorted(h) == sorted(r)
    assert {k: v for k, v in h.items() if r[k] != v} == {}


def test_constants_match_the_header():
    htext = _strip_c(open(HEADER).read())
    hdr = {m.group(1): int(m.group(2).rstrip("u"), 0) for m in re.finditer(r"#define (MP_[A-Z0-9_]+) \(?(-?(?:0x[0-9A-Fa-f]+|\d+)u?)\)?(?![.\w])", htext)}
    rs = {m.group(1): int(m.group(2).replace("_", ""), 0) for m in re.finditer(r"pub const (MP_[A-Z0-9_]+): [a-z0-9_]+ = (-?(?:0x[0-9A-Fa-f_]+|\d+));", open(LIB_RS).read())}
    assert len(hdr) >= 30
    assert {k: (v, rs.get(k)) for k, v in hdr.items() if rs.get(k) != v} == {}


_PRIM = {"u8": (1, 1), "u16": (2, 2), "u32": (4, 4), "i32": (4, 4), "f32": (4, 4), "u64": (8, 8), "f64": (8, 8), "usize": (8, 8)}


def _size_align(ty):
    ty = ty.strip()
    if ty.startswith("*"):
        return 8, 8
    m = re.fullmatch(r"\[(.*);\s*(\d+)\]", ty)
    if m:
        s, a = _size_align(m.group(1))
        return s * int(m.group(2)), a
    return _PRIM[ty]


def rust_structs():
    text = _strip_rs(open(LIB_RS).read())
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^\]]*\)\]\s*)?pub struct (Mp[A-Za-z0-9]+) \{(.*?)\n\}", text, flags=re.S):
        fields = [(f.group(1), f.group(2)) for f in re.finditer(r"pub ([a-z_0-9]+):\s*([^\n]+?),\s*(?=\n|$)", m.group(2) + "\n")]
        out[m.group(1)] = fields
    return out


def test_repr_c_structs_match_the_ctypes_mirror():
    from minipath_b200 import _ffi

    structs = rust_structs()
    checked = 0
    for name, fields in structs.items():
        if not fields:  # opaque handles
            continue
        mirror = getattr(_ffi, name)
        assert [f for f, _ in fields] == [f[0] for f in mirror._fields_], name
        off, align = 0, 1
        for fname, ty in fields:
            s, a = _size_align(ty)
            off = (off + a - 1) // a * a
            assert off == getattr(mirror, fname).offset, (name, fname)
            off += s
            align = max(align, a)
        assert (off + align - 1) // align * align == C.sizeof(mirror), name
        checked += 1
    assert checked >= 16


def test_build_rs_uses_the_makefile_flags():
    b = open(os.path.join(ROOT, "rust", "minipath-b200-sys", "build.rs")).read()
    mk = open(os.path.join(ROOT, "Makefile")).read()
    for flag in ["arch=compute_100a,code=sm_100a", "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ffp-contract=off", "-lineinfo"]:
        assert flag in b and flag in mk, flag
    for src in re.findall(r"\$\(CS\)/(host_[a-z_]+)\.o:", mk):
        assert f'"{src}"' in b, src


def test_reference_patch_only_uses_symbols_the_sys_crate_defines():
    """rust/reference-patch/*.rs (flatten(), render() over mp_render) name nothing in `sys::` that lib.rs does not declare,
    and every struct literal they build names exactly the struct's fields."""
    lib = _strip_rs(open(LIB_RS).read())
    defined = set(re.findall(r"pub (?:fn|struct|const|type) ([A-Za-z_0-9]+)", lib))
    structs = rust_structs()
    patch_dir = os.path.join(ROOT, "rust", "reference-patch")
    used = set()
    for name in sorted(os.listdir(patch_dir)):
        text = _strip_rs(open(os.path.join(patch_dir, name)).read())
        used |= set(re.findall(r"\bsys::([A-Za-z_0-9]+)", text))
        for m in re.finditer(r"sys::(Mp[A-Za-z0-9]+) \{(.*?)\n\s*\}", text, flags=re.S):
            body = m.group(2)
            if ".." in body:  # ..Default::default()
                continue
            depth, top = 0, ""
            for ch in body:  # keep only the literal's own nesting level
                if ch in "([{":
                    depth += 1
                elif ch in ")]}":
                    depth -= 1
                elif depth == 0:
                    top += ch
            fields = re.findall(r"(?:^|,)\s*([a-z_0-9]+)\s*(?::|,|$)", top, flags=re.M)
            want = [f for f, _ in structs[m.group(1)]]
            assert sorted(set(fields)) == sorted(want), (name, m.group(1), fields, want)
    assert used and used <= defined, sorted(used - defined)
