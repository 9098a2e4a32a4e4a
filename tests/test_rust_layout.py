"""Layout proof for the Rust binding (bindings/rust/trt-b200-sys/src/lib.rs) without a Rust toolchain.

Every `#[repr(C)]` struct of the crate is parsed, laid out by the C ABI rules repr(C) promises (fields in order, each
aligned to its own alignment, size rounded up to the largest alignment) and compared FIELD BY FIELD -- name, offset,
size -- and in total size with what gcc reports (offsetof / sizeof) for the struct of the same name in
include/trt_b200.h.  A field added, reordered, renamed or retyped on one side only fails here."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HDR = os.path.join(ROOT, "include", "trt_b200.h")
RS = os.path.join(ROOT, "bindings", "rust", "trt-b200-sys", "src", "lib.rs")

PRIM = {"u8": (1, 1), "i8": (1, 1), "u16": (2, 2), "i16": (2, 2), "u32": (4, 4), "i32": (4, 4), "f32": (4, 4), "u64": (8, 8),
        "i64": (8, 8), "f64": (8, 8), "usize": (8, 8), "isize": (8, 8), "c_int": (4, 4), "c_char": (1, 1)}


def rust_structs():
    src = re.sub(r"//.*", "", open(RS).read())
    out = {}
    for m in re.finditer(r"#\[repr\(C\)\][^{;]*?pub struct (\w+)\s*\{(.*?)\}", src, flags=re.S):
        name, body = m.group(1), m.group(2)
        fields = []
        for f in re.finditer(r"(?:pub\s+)?(\w+)\s*:\s*([^,]+?)\s*(?:,|$)", body.strip(), flags=re.S):
            fields.append((f.group(1), f.group(2).strip()))
        out[name] = fields
    return out


def rust_layout(structs, name, memo):
    """-> (size, align, [(field, offset, size)])"""
    if name in memo:
        return memo[name]

    def type_layout(t):
        t = t.strip()
        if t.startswith("*"):
            return 8, 8
        a = re.match(r"\[\s*(.+?)\s*;\s*(\d+)\s*\]$", t)
        if a:
            s, al = type_layout(a.group(1))
            return s * int(a.group(2)), al
        if t in PRIM:
            return PRIM[t]
        s, al, _ = rust_layout(structs, t, memo)
        return s, al

    off, align, rows = 0, 1, []
    for fname, ftype in structs[name]:
        s, al = type_layout(ftype)
        off = (off + al - 1) // al * al
        rows.append((fname, off, s))
        off += s
        align = max(align, al)
    size = (off + align - 1) // align * align
    memo[name] = (size, align, rows)
    return memo[name]


def c_structs():
    """struct name -> field names, parsed from the header (declarators only; types are gcc's business)"""
    src = re.sub(r"/\*.*?\*/", "", open(HDR).read(), flags=re.S)
    out = {}
    for m in re.finditer(r"typedef struct (\w+)\s*\{(.*?)\}\s*(\w+)\s*;", src, flags=re.S):
        fields = []
        for decl in m.group(2).split(";"):
            decl = decl.strip()
            if not decl:
                continue
            decl = re.sub(r"^(const\s+)?(unsigned\s+)?\w+\s+", "", decl, count=1)      # drop the type
            for d in decl.split(","):
                fields.append(re.sub(r"[\*\s]|\[.*?\]", "", d))
        out[m.group(3)] = fields
    return out


def c_layout(structs, tmp_path):
    lines = ['#include <stddef.h>', '#include <stdio.h>', f'#include "{HDR}"', "int main(void) {"]
    for s, fields in structs.items():
        lines.append(f'  printf("S {s} %zu\\n", sizeof({s}));')
        for f in fields:
            lines.append(f'  printf("F {s} {f} %zu %zu\\n", offsetof({s}, {f}), sizeof((({s} *)0)->{f}));')
    lines += ["  return 0;", "}"]
    c = tmp_path / "layout.c"
    c.write_text("\n".join(lines))
    exe = tmp_path / "layout"
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(c)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout
    sizes, fields = {}, {}
    for ln in out.splitlines():
        p = ln.split()
        if p[0] == "S":
            sizes[p[1]] = int(p[2])
        else:
            fields.setdefault(p[1], []).append((p[2], int(p[3]), int(p[4])))
    return sizes, fields


def test_rust_structs_match_the_header_field_by_field(tmp_path):
    rs = rust_structs()
    cs = c_structs()
    opaque = {n for n, f in rs.items() if [x[0] for x in f] == ["_private"]}
    assert {"trt_ctx", "trt_scene", "trt_multi"} <= opaque
    shared = [n for n in cs if n in rs and n not in opaque]
    # every struct the header defines by value is restated in the crate
    assert sorted(shared) == sorted(cs), sorted(set(cs) - set(shared))
    assert {"trt_vec3", "trt_primitive", "trt_uniforms", "trt_variance", "trt_config", "trt_counters", "trt_timings",
            "trt_frame_stats", "trt_camera", "trt_scene_info", "trt_run_options"} <= set(shared)
    sizes, fields = c_layout({n: cs[n] for n in shared}, tmp_path)
    memo = {}
    for n in shared:
        size, _, rows = rust_layout(rs, n, memo)
        assert rows == fields[n], f"{n}: Rust {rows} != C {fields[n]}"
        assert size == sizes[n], f"{n}: Rust size {size} != C sizeof {sizes[n]}"
    assert sizes["trt_vec3"] == 16 and sizes["trt_primitive"] == 180 and sizes["trt_uniforms"] == 160 and sizes["trt_variance"] == 8


def test_python_ctypes_structs_match_the_header(pkg, tmp_path):
    """the ctypes face (api.py) against the same gcc dump"""
    import ctypes as C
    cs = c_structs()
    py = {"trt_vec3": pkg.api.Vec3, "trt_primitive": pkg.api.Primitive, "trt_uniforms": pkg.api.Uniforms, "trt_config": pkg.api.Config,
          "trt_counters": pkg.api.Counters, "trt_timings": pkg.api.Timings, "trt_frame_stats": pkg.api.FrameStats,
          "trt_camera": pkg.api.CameraStruct, "trt_scene_info": pkg.api.SceneInfo, "trt_run_options": pkg.api.RunOptions}
    sizes, fields = c_layout({n: cs[n] for n in py}, tmp_path)
    for n, t in py.items():
        assert C.sizeof(t) == sizes[n], n
        got = [(f[0], getattr(t, f[0]).offset, getattr(t, f[0]).size) for f in t._fields_]
        assert got == fields[n], f"{n}: ctypes {got} != C {fields[n]}"
