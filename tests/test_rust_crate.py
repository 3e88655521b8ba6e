"""The Rust side of the boundary (rust/rtx-cuda-sys) cannot be compiled here — the reference's toolchain (cargo/rustc,
Cargo.toml:1-13) is absent — so what CAN be checked without it is checked: the crate's `extern "C"` block is exactly what
include/rtx_cuda.h generates (every function with its argument count, every flag/status constant, every struct with its
size assertion), the header's functions are exactly the shared library's exports, and build.rs compiles every csrc/*.cu
with the flags __graft_entry__.build() uses."""
import importlib.util
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CRATE = os.path.join(ROOT, "rust-raytracing_b200", "rust", "rtx-cuda-sys")


def _gen():
    spec = importlib.util.spec_from_file_location("gen_rust_sys", os.path.join(ROOT, "tools", "gen_rust_sys.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_lib_rs_is_what_the_header_generates():
    g = _gen()
    h = g.parse_header()
    assert open(g.LIB_RS).read() == g.generate(h), "run python tools/gen_rust_sys.py"
    # and, independently of the generator's own output format: names and argument counts
    rs = open(g.LIB_RS).read()
    block = rs[rs.index('extern "C" {'):]
    rust_fns = {m.group(1): (m.group(2).count(":") if m.group(2).strip() else 0)
                for m in re.finditer(r"pub fn (rtx_\w+)\((.*?)\)", block)}
    c_fns = {name: len(args) for _, name, args in h["functions"]}
    assert rust_fns == c_fns
    assert len(c_fns) >= 56
    for name, value in h["enums"] + h["defines"]:
        assert re.search(rf"pub const {name}: \w+ = {value};", rs), name
    for must in ("RTX_ACCEL_BRUTE", "RTX_ACCEL_LBVH", "RTX_ERR_CUDA", "RTX_ERR_UNSUPPORTED_SHAPE", "RTX_TILE",
                 "rtx_host_alloc", "rtx_host_free", "rtx_shard_tile_count", "rtx_deinterleave_tiles",
                 "rtx_quantize_rgb8_device", "rtx_camera_to_cam_space", "rtx_scene_get_camera", "rtx_abi_version",
                 "rtx_stream_state"):
        assert must in rs, must
    assert "size_of::<rtx_stats>() == 72" in rs and "size_of::<rtx_camera>() == 200" in rs


def test_header_functions_are_the_library_exports(rtx):
    g = _gen()
    declared = {name for _, name, _ in g.parse_header()["functions"]}
    lib_path = os.path.join(ROOT, "rust-raytracing_b200", "lib", "librtx_cuda.so")
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path], capture_output=True, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if re.search(r" T rtx_\w+$", ln)}
    assert declared == exported, declared ^ exported


def test_struct_sizes_match_ctypes(rtx):
    import ctypes as C
    g = _gen()
    sz = g.layout(g.parse_header())
    a = rtx._abi
    pairs = {"rtx_vec3": a.Vec3, "rtx_mat3": a.Mat3, "rtx_camera": a.CameraC, "rtx_config": a.ConfigC,
             "rtx_material": a.MaterialC, "rtx_sphere": a.SphereC, "rtx_plane": a.PlaneC, "rtx_triangle": a.TriangleC,
             "rtx_stats": a.Stats, "rtx_shard": a.Shard}
    for name, t in pairs.items():
        assert sz[name][0] == C.sizeof(t), name


def test_build_rs_compiles_every_translation_unit():
    import __graft_entry__ as ge
    text = open(os.path.join(CRATE, "build.rs")).read()
    # no hand-kept source list: the crate globs csrc/*.cu like __graft_entry__.build()
    assert "read_dir(&csrc)" in text and 'x == "cu"' in text
    assert not re.search(r'"rtx_\w+\.cu"', text), "build.rs names individual .cu files again"
    for flag in ge.NVCC_FLAGS:
        for part in flag.split(","):
            assert part in text, f"nvcc flag {part} of __graft_entry__.NVCC_FLAGS is missing from build.rs"
    cu = [f for f in os.listdir(ge.CSRC) if f.endswith(".cu")]
    assert len(cu) >= 6
