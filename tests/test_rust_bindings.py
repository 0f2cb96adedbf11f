"""bindings/rust: no Rust toolchain exists in this image, so the crate cannot be compiled — but it can be CHECKED.
This test parses every `#[repr(C)]` struct, every `extern "C"` function and every constant out of
`bindings/rust/rtgpu-sys/src/lib.rs` and compares them with `include/rtgpu.h`: struct names, field order, field
types (incl. array extents and pointer constness), function names, argument order and types, return types, and
the values of the enums / #defines.  The flattener (`bindings/rust/reference_patch/flatten.rs`) is checked against the
binding (every rtgpu_sys item it uses exists) and against the reference's types (one impl per Hitable / Material /
Texture implementor the surveyed snapshot has), and the hook-in script's line numbers against that snapshot when it
is present.  Status of the crate: UNCOMPILED, not untested."""
import os
import re
import shutil
import subprocess

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(REPO, "include", "rtgpu.h")
LIB_RS = os.path.join(REPO, "bindings", "rust", "rtgpu-sys", "src", "lib.rs")
FLATTEN_RS = os.path.join(REPO, "bindings", "rust", "reference_patch", "flatten.rs")

C2RUST = {"uint32_t": "u32", "int32_t": "i32", "uint64_t": "u64", "uint8_t": "u8", "float": "f32", "double": "f64",
          "int": "c_int", "size_t": "usize", "char": "c_char", "void": "c_void"}


def _strip_c_comments(src):
    return re.sub(r"/\*.*?\*/", " ", src, flags=re.S)


def _c_type_to_rust(ctype, arrays=()):
    """'const float*' -> '*const f32', 'RtScene* const*' -> '*const *mut RtScene', arrays=(256, 4) -> [[f32; 4]; 256]"""
    t = ctype.strip()
    # split into base + pointer levels with their constness
    toks = re.findall(r"const|\*|[A-Za-z_][A-Za-z_0-9]*", t)
    base, base_const, levels = None, False, []   # levels[k] = constness of the k-th pointer itself (T* const)
    for tk in toks:
        if tk == "const":
            if levels:
                levels[-1] = True
            else:
                base_const = True                  # `const T` or `T const`
        elif tk == "*":
            levels.append(False)
        else:
            base = tk
    rust = C2RUST.get(base, base)
    # innermost pointee constness is base_const; pointer k's pointee constness is levels[k-1]
    pointee_const = base_const
    for lvl_const in levels:
        rust = ("*const " if pointee_const else "*mut ") + rust
        pointee_const = lvl_const
    for n in reversed(arrays):
        rust = f"[{rust}; {n}]"
    return rust


def parse_header():
    src = _strip_c_comments(open(HEADER).read())
    structs, funcs, consts = {}, {}, {}
    for m in re.finditer(r"typedef struct (\w+) \{(.*?)\} \1;", src, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            fm = re.match(r"(.+?)\s*(\w+)((?:\[\d+\])*)$", decl)
            ctype, name, arr = fm.group(1), fm.group(2), [int(x) for x in re.findall(r"\[(\d+)\]", fm.group(3))]
            fields.append((name, _c_type_to_rust(ctype, arr)))
        structs[m.group(1)] = fields
    for m in re.finditer(r"typedef enum \w+ \{(.*?)\} \w+;", src, flags=re.S):
        for item in m.group(1).split(","):
            im = re.match(r"\s*(\w+)\s*=\s*(-?\d+)", item)
            if im:
                consts[im.group(1)] = int(im.group(2))
    for m in re.finditer(r"#define\s+(RT_\w+)\s+(\d+)u?\b", src):
        consts[m.group(1)] = int(m.group(2))
    for m in re.finditer(r"^\s*((?:const\s+)?\w+\s*\**)\s*(rt_\w+)\s*\((.*?)\)\s*;", src, flags=re.S | re.M):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        alist = []
        if args and args != "void":
            for a in re.split(r",(?![^\[]*\])", args):
                a = a.strip()
                am = re.match(r"(.+?)\s*(\w+)((?:\[\w+\])*)$", a)
                ctype, aname, arr = am.group(1), am.group(2), am.group(3)
                if arr:  # an array parameter is a pointer
                    ctype = ctype + "*"
                alist.append((aname, _c_type_to_rust(ctype)))
        funcs[name] = (alist, None if ret == "void" else _c_type_to_rust(ret))
    return structs, funcs, consts


def parse_lib_rs():
    src = re.sub(r"//.*", "", open(LIB_RS).read())
    structs, funcs, consts = {}, {}, {}
    for m in re.finditer(r"#\[repr\(C\)\]\s*(?:#\[derive\([^)]*\)\]\s*)?pub struct (\w+) \{(.*?)\n\}", src, flags=re.S):
        fields = [(f.group(1), re.sub(r"\s+", " ", f.group(2).strip())) for f in re.finditer(r"pub (\w+):\s*([^,\n]+),", m.group(2))]
        structs[m.group(1)] = fields
    ext = re.search(r'extern "C" \{(.*?)\n\}', src, flags=re.S).group(1)
    for m in re.finditer(r"pub fn (\w+)\((.*?)\)\s*(?:->\s*([^;]+))?;", ext, flags=re.S):
        args = []
        if m.group(2).strip():
            for a in m.group(2).split(","):
                an, at = a.split(":", 1)
                args.append((an.strip(), re.sub(r"\s+", " ", at.strip())))
        funcs[m.group(1)] = (args, m.group(3).strip() if m.group(3) else None)
    for m in re.finditer(r"pub const (\w+):\s*\w+\s*=\s*(-?\d+);", src):
        consts[m.group(1)] = int(m.group(2))
    opaque = set(re.findall(r"pub enum (\w+) \{\}", src))
    return structs, funcs, consts, opaque


def test_type_mapping_helper():
    assert _c_type_to_rust("const float*") == "*const f32"
    assert _c_type_to_rust("float*") == "*mut f32"
    assert _c_type_to_rust("RtScene* const*") == "*const *mut RtScene"
    assert _c_type_to_rust("void* const*") == "*const *mut c_void"
    assert _c_type_to_rust("float", [256, 4]) == "[[f32; 4]; 256]"
    assert _c_type_to_rust("const RtFlatScene*") == "*const RtFlatScene"
    assert _c_type_to_rust("RtScene**") == "*mut *mut RtScene"


def test_rust_structs_mirror_the_header_field_for_field():
    h_structs, _, _ = parse_header()
    r_structs, _, _, opaque = parse_lib_rs()
    assert set(h_structs) == set(r_structs), set(h_structs) ^ set(r_structs)
    for name, fields in h_structs.items():
        assert r_structs[name] == fields, (name, r_structs[name], fields)
    assert opaque == {"RtScene", "RtComm", "RtHostScene"}


def test_rust_functions_mirror_the_header():
    _, h_funcs, _ = parse_header()
    _, r_funcs, _, _ = parse_lib_rs()
    assert set(h_funcs) == set(r_funcs), set(h_funcs) ^ set(r_funcs)
    for name, (args, ret) in h_funcs.items():
        r_args, r_ret = r_funcs[name]
        assert [a[1] for a in r_args] == [a[1] for a in args], (name, r_args, args)
        assert [a[0] for a in r_args] == [a[0] for a in args], (name, r_args, args)
        assert r_ret == ret, (name, r_ret, ret)


def test_rust_constants_match_the_header():
    _, _, h_consts = parse_header()
    _, _, r_consts, _ = parse_lib_rs()
    for name, val in h_consts.items():
        assert r_consts.get(name) == val, (name, r_consts.get(name), val)
    assert r_consts["RT_ABI_VERSION"] == 2


def test_rust_layout_equals_ctypes_layout(rtlib):
    """Sizes implied by the Rust field lists (repr(C) rules) equal the ctypes mirror's, which the GPU tests exercise."""
    import ctypes as C
    r_structs, _, _, _ = parse_lib_rs()
    prim = {"u32": (4, 4), "i32": (4, 4), "f32": (4, 4), "u64": (8, 8), "f64": (8, 8), "u8": (1, 1)}

    def size_align(t):
        t = t.strip()
        if t.startswith("*"):
            return 8, 8
        m = re.match(r"\[(.+); (\d+)\]$", t)
        if m:
            s, a = size_align(m.group(1))
            return s * int(m.group(2)), a
        if t in prim:
            return prim[t]
        return struct_size(t)

    def struct_size(name):
        off, align = 0, 1
        for _, t in r_structs[name]:
            s, a = size_align(t)
            off = (off + a - 1) // a * a + s
            align = max(align, a)
        return (off + align - 1) // align * align, align

    api = rtlib.api
    for name in r_structs:
        assert struct_size(name)[0] == C.sizeof(getattr(api, name)), name


def test_flattener_uses_only_what_the_binding_declares():
    src = re.sub(r"//.*", "", open(FLATTEN_RS).read())
    r_structs, r_funcs, r_consts, opaque = parse_lib_rs()
    known = set(r_structs) | set(r_funcs) | set(r_consts) | opaque
    used = set(re.findall(r"\b(RT_[A-Z0-9_]+|Rt[A-Z]\w+|rt_[a-z0-9_]+)\b", src))
    assert used <= known, used - known
    # struct literals name existing fields only
    for m in re.finditer(r"\b(Rt\w+) \{([^{}]*)\}", src):
        if m.group(1) not in r_structs:
            continue
        names = {f[0] for f in r_structs[m.group(1)]}
        for f in re.findall(r"(\w+)\s*:(?!:)", m.group(2)):
            assert f in names, (m.group(1), f)
    # one FlattenHitable / FlattenMaterial / FlattenTexture impl per implementor of the reference's traits
    impls = {t: set(re.findall(rf"impl {t} for (\w+)", src)) for t in ("FlattenHitable", "FlattenMaterial", "FlattenTexture")}
    assert impls["FlattenHitable"] == {"Sphere", "XYRect", "XZRect", "YZRect", "HitableList", "BvhNode", "GBox", "Translate",
                                       "RotateY", "ConstantMedium"}
    assert impls["FlattenMaterial"] == {"Emission", "Diffuse", "Lambert", "Metal", "Dielectric", "Isotropic", "OrenNayar",
                                        "BurleyDiffuse", "RoughPlastic", "DisneyDiffuse", "DisneyMetal", "DisneySheen",
                                        "DisneyClearcoat"}
    assert impls["FlattenTexture"] == {"ConstantTex", "CheckerTex", "PerlinTex", "ImageTex"}


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="the reference snapshot is not present on this machine")
def test_hook_in_script_applies_to_the_surveyed_snapshot(tmp_path):
    """apply.sh (sed edits guarded by `expect` checks) runs clean against the reference snapshot, and every type the
    flattener implements its traits for is an implementor of the reference's Hitable / Material / Texture."""
    dst = tmp_path / "ref"
    shutil.copytree("/root/reference", dst)
    r = subprocess.run(["sh", os.path.join(REPO, "bindings", "rust", "reference_patch", "apply.sh")], cwd=dst, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    hit = open(dst / "src" / "hitable.rs").read()
    assert "pub trait Hitable: Send + Sync + crate::flatten::FlattenHitable" in hit
    assert "pub(crate) left: Arc<dyn Hitable>" in hit and "pub(crate) neg_inv_density: f32" in hit
    assert "pub(crate) struct Perlin" in open(dst / "src" / "texture.rs").read()
    assert "mod flatten;" in open(dst / "src" / "main.rs").read() and (dst / "src" / "flatten.rs").exists()
    ref_src = "".join(open(dst / "src" / f).read() for f in ("hitable.rs", "material.rs", "pbr.rs", "texture.rs"))
    flat = open(FLATTEN_RS).read()
    for trait, ftrait in (("Hitable", "FlattenHitable"), ("Material", "FlattenMaterial"), ("Texture", "FlattenTexture")):
        ref_impls = set(re.findall(rf"impl {trait} for (\w+)", ref_src))
        assert ref_impls == set(re.findall(rf"impl {ftrait} for (\w+)", flat)), (trait, ref_impls)
    # every field the flattener reads exists in the reference struct of that name
    for ty, fields in {"Sphere": ["c", "r", "mat"], "BvhNode": ["left", "right"], "RotateY": ["ptr", "sin_theta", "cos_theta", "angle"],
                       "ConstantMedium": ["boundary", "phase_fn", "neg_inv_density"], "Translate": ["offset", "ptr"], "GBox": ["sides"],
                       "PerlinTex": ["perlin", "scale"], "CheckerTex": ["odd", "even"], "ImageTex": ["img"]}.items():
        body = re.search(rf"pub struct {ty} \{{(.*?)\n\}}", ref_src, flags=re.S).group(1)
        for f in fields:
            assert re.search(rf"\b{f}:", body), (ty, f)
