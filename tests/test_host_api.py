"""Host side of the drop-in boundary, no GPU needed: the C-ABI library loads and exports every symbol the header
declares, the scene model / scene-text reader (serialize::deserialize, src/serialize.rs) / BMP sink (src/bmp.rs) behave
like the reference, and every compute entry point fails loudly (RT_ERR_CUDA) instead of falling back to a CPU path.
"""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from test_gpu_parity import REF_TEXT_SCENE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "raytrace_b200.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"RT_API\s+[^;(]*?\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(rt):
    syms = header_symbols()
    assert len(syms) >= 45, syms
    L = C.CDLL(rt.LIB_PATH)
    for s in syms:
        assert hasattr(L, s), "libraytrace_b200.so does not export %s" % s
    B = __import__("importlib").import_module("rust-raytrace_b200._binding")
    assert sorted(B.API) == syms, "binding table and header disagree: %s" % (set(B.API) ^ set(syms))
    # nothing but the C ABI is visible (-fvisibility=hidden): no C++ or oracle symbols leak
    out = subprocess.run(["nm", "-D", "--defined-only", rt.LIB_PATH], capture_output=True, text=True).stdout
    exported = [ln.split()[-1] for ln in out.splitlines() if " T " in ln]
    assert sorted(exported) == syms, set(exported) ^ set(syms)
    assert not any("orc_" in ln or "emu_" in ln for ln in out.splitlines())


def test_product_does_not_reference_the_oracle():
    pkg = os.path.join(ROOT, "rust-raytrace_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.path.basename(dirpath) in ("build", "lib", "__pycache__"):
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                if f.endswith(".py"):
                    bad = re.search(r"^\s*(import|from)\s+(oracle|tests|emu)\b|liboracle|libemu", src, flags=re.M)
                else:  # comments may mention the test harness; includes and loaded libraries may not
                    bad = re.search(r"#\s*include[^\n]*(oracle|emu)|liboracle|libemu", src)
                    # the one library the product loads at run time is NCCL (context.cu, RT_GATHER_NCCL): every shared-object
                    # name that appears in a file calling dlopen must be libnccl
                    if "dlopen" in src:
                        assert f == "context.cu"
                        names = re.findall(r'"([^"\n]*\.so[^"\n]*)"', src)
                        assert names and all(re.fullmatch(r"libnccl\.so(\.2)?", n) or "libnccl.so.2" in n for n in names), names
                assert not bad, "%s refers to test infrastructure" % f


def test_version_and_defaults(rt):
    assert rt.version().startswith("raytrace_b200") and "sm_100a" in rt.version()
    B = __import__("importlib").import_module("rust-raytrace_b200._binding")
    o = B.make_opts(rt.lib())
    assert o.max_depth == 4 and o.min_significance == 1.0 / 512.0  # src/raytrace.rs:17-18
    assert o.antialias == 0xFFFFFFFF and o.width == 0 and o.height == 0
    assert o.world == 1 and o.rank == 0 and o.accel == rt.ACCEL_AUTO and o.fp_mode == rt.FP_STRICT


def test_no_cpu_fallback(rt, scenes):
    if rt.device_count() > 0:
        pytest.skip("a GPU is visible: the failure path cannot be observed")
    sc = rt.Scene(scenes.materials_showcase(16, 12, 1))
    for call in (lambda: rt.raytrace.render(sc), lambda: rt.Context(sc), lambda: rt.raytrace.intersect(sc, np.zeros((1, 6))),
                 lambda: rt.raytrace.project(sc, np.zeros((1, 2))), lambda: rt.color.write_bgr(np.zeros((1, 3)))):
        with pytest.raises(rt.RtError) as e:
            call()
        assert e.value.code == -4 and "no CPU fallback" in e.value.message
    for call in (lambda: rt.raytrace.render_multi(sc, 2), lambda: rt.MultiContext(sc, 1)):
        with pytest.raises(rt.RtError) as e:
            call()
        assert e.value.code == -4 and "no CPU fallback" in e.value.message


def test_new_entry_points_check_their_arguments_before_they_touch_a_device(rt):
    """rt_context_render_begin / _end and rt_intersect_spheres_noalloc: bad arguments are RT_ERR_INVALID with a message (whether or
    not a GPU is visible); without a GPU a well-formed no-alloc call is RT_ERR_CUDA -- no CPU path."""
    import ctypes as C
    L = rt.lib()
    assert L.rt_context_render_begin(None) == -1 and b"null context" in L.rt_last_error()
    assert L.rt_context_render_end(None, None) == -1
    fake = C.c_void_p(4096)   # never dereferenced: every check below fails before a launch
    assert L.rt_intersect_spheres_noalloc(0, None, None, 5, fake, 10, fake, fake) == -1 and b"null argument" in L.rt_last_error()
    assert L.rt_intersect_spheres_noalloc(0, None, fake, 5, None, 10, fake, fake) == -1
    assert L.rt_intersect_spheres_noalloc(0, None, fake, 5, fake, 10, None, fake) == -1
    assert L.rt_intersect_spheres_noalloc(0, None, fake, rt.NOALLOC_MAX_SPHERES + 1, fake, 10, fake, fake) == -1
    msg = L.rt_last_error().decode()
    assert "RT_NOALLOC_MAX_SPHERES" in msg and "65536" in msg and "rt_trace_rays" in msg
    assert rt.NOALLOC_SCRATCH_BYTES == 65536 * 32 + 256
    assert L.rt_intersect_spheres_noalloc(0, None, C.c_void_p(4096 + 8), 5, fake, 10, fake, fake) == -1 and b"misaligned" in L.rt_last_error()
    assert L.rt_intersect_spheres_noalloc(0, None, fake, 5, fake, 1 << 31, fake, fake) == -1 and b"too many rays" in L.rt_last_error()
    if rt.device_count() == 0:
        assert L.rt_intersect_spheres_noalloc(0, None, fake, 5, fake, 10, fake, fake) == -4 and b"no CPU fallback" in L.rt_last_error()
    else:
        assert L.rt_intersect_spheres_noalloc(rt.device_count(), None, fake, 5, fake, 10, fake, fake) == -1 and b"no such device" in L.rt_last_error()


# ---- serialize::deserialize -------------------------------------------------------------------------------
def test_parse_reference_scene(rt):
    sc = rt.serialize.deserialize(REF_TEXT_SCENE)
    assert sc.options() == (800, 800, 1024)
    info = sc.info()
    assert info["n_objects"] == 7 and info["n_planes"] == 5 and info["n_spheres"] == 2 and info["n_lights"] == 0
    assert info["camera_kind"] == 0 and info["background_kind"] == 0
    pos, m = sc.camera()
    assert pos.tolist() == [0, 3, 17] and np.array_equal(m, [[1, 0, 0], [0, 1, 0], [0, 0, -3.6]])


def test_text_and_builder_scenes_are_identical(rt, scenes):
    for desc in (scenes.materials_showcase(64, 48, 2), scenes.random_spheres(200, 3, 9.0, 32, 32)):
        a = rt.Scene(desc).dump()
        b = rt.serialize.deserialize(rt.scene.to_text(desc)).dump()
        assert a == b
        assert len(a) > 500


def test_parser_agrees_with_python_restatement(rt, scenes, orc):
    """oracle/scene_text.py restates src/serialize.rs independently; both readers must build the same scene."""
    from oracle import scene_text
    text = rt.scene.to_text(scenes.materials_showcase(40, 30, 3))
    text = "# leading comment\n" + text.replace("objects:", "/* block\n comment */ objects:").replace("lights:", "// line comment\n lights:")
    desc = scene_text.deserialize(text)
    assert rt.Scene(desc).dump() == rt.serialize.deserialize(text).dump()
    # camera matrix computed by the host library == the oracle's (src/camera.rs:51-63)
    pos, m = rt.serialize.deserialize(text).camera()
    opos, om = orc.OracleScene(desc).camera()
    assert np.array_equal(pos, opos) and np.array_equal(m, om)


BAD_SCENES = [
    "", "{", "}", "{ objects: [ } ", "{ objects: [] lights: [] }", "{ objects: [], }", "[", "{ objects: 5 }",
    "{ objects: [ { bounds: Cube { } } ] }",
    "{ objects: [ { bounds: Sphere { center: (0, 0) radius: 1 } } ] }",
    "{ objects: [ { bounds: Sphere { center: (0, 0, 0) radius: abc } material: PhongMaterial {} } ] }",
    "{ objects: [ { bounds: Sphere { radius: 1 center: (0, 0, 0) } } ] }",
    '{ objects: [] lights: [] camera: SimplePerspectiveCamera new((0,0,0),(0,0,-1),(0,1,0)) }',
    '{ objects: [] lights: [] camera: SimplePerspectiveCamera look_at((0,0,0),(0,0,-1),(0,1,0), 45 furlongs, 1) }',
    '{ objects: [] lights: [] camera: Pinhole new() }',
    '{ objects: [] lights: [ { model: SpotLight { } color: rgb(1,1,1) } ] }',
    '{ objects: [] lights: [] camera: SimplePerspectiveCamera new((0,0,0),(0,0,-1),(0,1,0),1) background: SolidColorBackground { color: rgb(1,1) } }',
    '{ objects: [] lights: [] camera: SimplePerspectiveCamera new((0,0,0),(0,0,-1),(0,1,0),1) background: SolidColorBackground { color: rgb(1,1,1) } options: { width: -3 height: 2 antialias: 1 } }',
    '{ objects: [] lights: [] camera: SimplePerspectiveCamera new((0,0,0),(0,0,-1),(0,1,0),1) background: SolidColorBackground { color: rgb(1,1,1) } options: { width: 3 height: 2 antialias: 1 } } trailing',
    '{ objects: [ "unterminated ] }', "{ objects: [ /* unterminated ] }", '{ objects: [ "bad \\q escape" ] }', "{ objects: @ }", "{ objects: [ 1e400 ] }",
]


@pytest.mark.parametrize("i", range(len(BAD_SCENES)))
def test_syntax_errors_match_reference_format(rt, i):
    """error text is upstream's `row:col: message` (src/serialize.rs:244-272); compared with the Python restatement"""
    from oracle import scene_text
    text = BAD_SCENES[i]
    want = None
    try:
        scene_text.deserialize(text)
    except scene_text.SyntaxError_ as e:
        want = str(e)
    if want is None:
        rt.serialize.deserialize(text)  # both accept
        return
    with pytest.raises(rt.serialize.SyntaxError) as e:
        rt.serialize.deserialize(text)
    assert e.value.code == -2
    assert e.value.message == want
    assert re.match(r"^\d+:\d+: ", e.value.message)


def test_parse_warnings_and_units(rt):
    from oracle import scene_text
    base = ('{ objects: [] lights: [] camera: SimplePerspectiveCamera look_at((0,1,0),(0,0,-1),(0,1,0), %s, 2.0) '
            'background: SolidColorBackground { color: rgb(0,0,0) } options: { width: 4.0 height: 3.7 antialias: 1 } }')
    a = rt.serialize.deserialize(base % "90 deg")
    b = rt.serialize.deserialize(base % ("%r rad" % (np.pi / 2)))
    assert np.allclose(a.camera()[1], b.camera()[1], rtol=0, atol=1e-15)
    d, warns = scene_text.deserialize(base % "90 deg", return_warnings=True)
    assert a.options() == (d["options"]["width"], d["options"]["height"], d["options"]["antialias"])
    assert a.warnings().strip().splitlines() == [w.strip() for w in warns]


def test_parse_file_and_missing_file(rt, tmp_path):
    p = tmp_path / "scene.txt"
    p.write_text(REF_TEXT_SCENE)
    L = rt.lib()
    out = C.c_void_p()
    err = C.create_string_buffer(256)
    assert L.rt_scene_parse_file(os.fsencode(str(p)), C.byref(out), err, 256) == 0
    L.rt_scene_free(out)
    assert L.rt_scene_parse_file(os.fsencode(str(tmp_path / "nope.txt")), C.byref(out), err, 256) == -3


def test_builder_validation(rt):
    L = rt.lib()
    B = __import__("importlib").import_module("rust-raytrace_b200._binding")
    h = C.c_void_p(L.rt_scene_new())
    one = np.array([[0.0, 0.0, 0.0, 1.0]])
    mat = np.array([3], dtype=np.uint32)
    assert L.rt_scene_add_spheres(h, 1, B.ptr(one, C.c_double), B.ptr(mat, C.c_uint32)) == -1  # material index out of range
    assert b"material" in L.rt_last_error()
    assert L.rt_scene_add_plane(h, B.d3((0, 0, 0)), B.d3((0, 1, 0)), 0) == -1
    assert L.rt_scene_set_camera_dof(h, 1.0, 0.1, 2) == -1                                          # no perspective camera yet
    assert L.rt_scene_set_skybox_face(h, 6, 2, 2, None) == -1
    assert L.rt_scene_add_material_phong(h, B.d3((1, 1, 1)), B.d3((0, 0, 0)), 1.0, B.d3((0, 0, 0))) == 0
    mat[0] = 0
    assert L.rt_scene_add_spheres(h, 1, B.ptr(one, C.c_double), B.ptr(mat, C.c_uint32)) == 0
    L.rt_scene_free(h)


# ---- skybox textures (src/texture.rs:34-37 for the formats decodable without the `image` crate) --------------
def test_skybox_face_loaders(rt, scenes, tmp_path):
    faces = scenes.skybox_faces(16, 4)
    S = rt.scene
    desc = scenes.random_spheres(5, 1, 3.0, 8, 8, background=S.SkyboxBackground(*faces))
    paths = []
    for i, f in enumerate(faces):
        if i % 2 == 0:  # binary PPM
            p = tmp_path / ("f%d.ppm" % i)
            p.write_bytes(b"P6\n# c\n16 16\n255\n" + f.tobytes())
        else:           # 24-bit BMP, bottom-up BGR rows
            p = tmp_path / ("f%d.bmp" % i)
            rows = np.zeros((16, rt.bmp.bytewidth(16)), dtype=np.uint8)
            rows[:, :48] = f[::-1, :, ::-1].reshape(16, 48)
            rt.bmp.write(str(p), 16, 16, rows)
        paths.append(str(p))
    from_text = rt.serialize.deserialize(S.to_text(desc, texture_paths=paths))
    assert from_text.dump() == rt.Scene(desc).dump()
    assert from_text.info()["background_kind"] == 1
    with pytest.raises(rt.RtError):
        rt.serialize.deserialize(S.to_text(desc, texture_paths=[str(tmp_path / "missing.ppm")] * 6))


# ---- bmp (src/bmp.rs:10-61) -------------------------------------------------------------------------------
def test_bmp_header_and_file(rt, orc, tmp_path):
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "outbmp_block_means.npz"))
    assert rt.bmp.header(800, 800) == g["header"].tobytes()
    for w, h in [(1, 1), (2, 3), (3, 2), (5, 5), (800, 800), (3840, 2160), (7680, 4320), (1023, 7)]:
        assert rt.bmp.header(w, h) == orc.bmp_header(w, h)
        assert rt.bmp.bytewidth(w) == orc.bytewidth(w) == (3 * w + 3) // 4 * 4
    rows = np.random.default_rng(0).integers(0, 256, size=(7, rt.bmp.bytewidth(5)), dtype=np.uint8)
    p = tmp_path / "o.bmp"
    rt.bmp.write(str(p), 5, 7, rows)
    raw = p.read_bytes()
    assert raw[:122] == orc.bmp_header(5, 7) and raw[122:] == rows.tobytes() and len(raw) == 122 + 16 * 7
    with pytest.raises(ValueError):
        rt.bmp.write(str(p), 5, 7, rows[:, :15])
    with pytest.raises(rt.RtError):
        rt.bmp.write(str(tmp_path / "no_such_dir" / "o.bmp"), 5, 7, rows)


CLI = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rust-raytrace_b200", "lib", "raytrace_b200")


def test_cli_reports_like_the_reference_binary(rt, tmp_path):
    """csrc/cli.cpp = src/main.rs over the C ABI: default file names; open and syntax errors are ONE stdout line
    `error: <Display of the error>` with exit status 0, exactly `println!("error: {}", e); return` (src/main.rs:18,23,28);
    without a GPU the render step fails loudly (no CPU path)"""
    assert os.path.exists(CLI), "rust-raytrace_b200/build.py builds the CLI next to the library"
    r = subprocess.run([CLI, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "test_scene.txt -> out.bmp" in r.stderr
    bad = tmp_path / "bad.txt"
    bad.write_text("camera { bogus }")
    r = subprocess.run([CLI, str(bad)], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0
    assert len(r.stdout.splitlines()) == 1 and re.match(r"^error: 1:\d+: ", r.stdout)
    from oracle import scene_text
    with pytest.raises(scene_text.SyntaxError_) as ei:   # the independent Python restatement of the grammar agrees on the text
        scene_text.deserialize("camera { bogus }")
    assert r.stdout == "error: %s\n" % ei.value
    r = subprocess.run([CLI], capture_output=True, text=True, cwd=tmp_path)   # no test_scene.txt in the working directory
    assert r.returncode == 0 and r.stdout == "error: No such file or directory (os error 2)\n"   # io::Error's Display
    notutf8 = tmp_path / "latin1.txt"
    notutf8.write_bytes(b"# caf\xe9\n")
    r = subprocess.run([CLI, str(notutf8)], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0 and r.stdout == "error: stream did not contain valid UTF-8\n"    # read_to_string, src/main.rs:22-24
    assert not (tmp_path / "out.bmp").exists()
    if rt.device_count() == 0:
        good = tmp_path / "test_scene.txt"
        good.write_text(REF_TEXT_SCENE)
        r = subprocess.run([CLI, "--width", "16", "--height", "16", "--antialias", "1"], capture_output=True, text=True, cwd=tmp_path)
        assert r.returncode == 1 and "render failed" in r.stderr
        assert not (tmp_path / "out.bmp").exists()


# ---- the Rust shim crate (shim/): source only, but its FFI block cannot drift from the header ---------------------------
def _rust_layout(fields):
    """size / offsets of a #[repr(C)] struct of primitives and fixed arrays, by the C rules Rust's repr(C) follows"""
    prim = {"u8": 1, "i8": 1, "u16": 2, "i16": 2, "u32": 4, "i32": 4, "f32": 4, "u64": 8, "i64": 8, "f64": 8, "usize": 8, "c_int": 4}
    off, align, out = 0, 1, []
    for name, ty in fields:
        m = re.fullmatch(r"\[(\w+); (\d+)\]", ty)
        elem, n = (m.group(1), int(m.group(2))) if m else (ty, 1)
        sz = prim[elem]
        off = (off + sz - 1) // sz * sz
        out.append((name, off, sz * n))
        off += sz * n
        align = max(align, sz)
    return out, (off + align - 1) // align * align


def test_rust_shim_ffi_is_generated_from_the_header_and_matches_the_c_layout(rt):
    import ctypes as C
    import importlib.util
    spec = importlib.util.spec_from_file_location("gen_rust_ffi", os.path.join(ROOT, "tools", "gen_rust_ffi.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    text, structs = gen.generate()
    committed = open(os.path.join(ROOT, "shim", "src", "ffi.rs")).read()
    assert committed == text, "shim/src/ffi.rs is stale: run python tools/gen_rust_ffi.py"
    B = __import__("importlib").import_module("rust-raytrace_b200._binding")
    mirror = {"RtRenderOpts": B.RenderOpts, "RtStats": B.Stats, "RtSceneInfo": B.SceneInfo, "RtPeaks": B.Peaks}
    found = 0
    for m in re.finditer(r"#\[repr\(C\)\]\s*#\[derive\([^)]*\)\]\s*pub struct (\w+) \{(.*?)\n\}", committed, flags=re.S):
        name = m.group(1)
        fields = re.findall(r"pub (\w+): ([^,\n]+),", m.group(2))
        assert name in mirror, "no ctypes mirror for %s" % name
        ct = mirror[name]
        layout, size = _rust_layout(fields)
        assert [f for f, _, _ in layout] == [f for f, _ in ct._fields_], name
        for f, off, sz in layout:
            d = getattr(ct, f)
            assert (d.offset, d.size) == (off, sz), "%s.%s: Rust (%d, %d) vs C (%d, %d)" % (name, f, off, sz, d.offset, d.size)
        assert size == C.sizeof(ct), "%s: %d bytes in Rust, %d in C" % (name, size, C.sizeof(ct))
        found += 1
    assert found == len(mirror)
    # every exported function is declared, with as many parameters as the ctypes table has
    decls = dict(re.findall(r"pub fn (rt_\w+)\((.*?)\)", committed))
    assert sorted(decls) == sorted(B.API)
    for fn, args in decls.items():
        n = 0 if not args.strip() else args.count(":")
        assert n == len(B.API[fn][1]), fn
    assert "pub bvh_node_bytes: u64" in committed and "pub scan_sphere_tests: [u64; 2]" in committed   # r01's drift


def test_rust_shim_exports_the_crates_modules():
    """shim/src/lib.rs keeps the nine module names of upstream's lib.rs (src/lib.rs:4-13), texture behind the same feature"""
    lib = open(os.path.join(ROOT, "shim", "src", "lib.rs")).read()
    mods = re.findall(r"^pub mod (\w+);", lib, flags=re.M)
    assert mods == ["ffi", "shapes", "types", "color", "scene", "camera", "raytrace", "bmp", "serialize", "texture"]
    assert re.search(r'#\[cfg\(feature = "skybox"\)\]\s*pub mod texture;', lib)
    for m in mods:
        assert os.path.exists(os.path.join(ROOT, "shim", "src", m + ".rs")), m
    cargo = open(os.path.join(ROOT, "shim", "Cargo.toml")).read()
    assert 'name = "libraytrace"' in cargo and 'name = "raytrace"' in cargo      # upstream's lib and bin names (Cargo.toml:14-21)
    main = open(os.path.join(ROOT, "shim", "src", "bin", "raytrace.rs")).read()
    assert main.count('println!("error: {}", e); return') == 3                     # src/main.rs:18,23,28
    assert '"test_scene.txt"' in main and '"out.bmp"' in main
    # every ffi function the hand-written modules call exists in the generated block
    ffi = open(os.path.join(ROOT, "shim", "src", "ffi.rs")).read()
    declared = set(re.findall(r"pub fn (rt_\w+)\(", ffi))
    consts = set(re.findall(r"pub const (RT_\w+):", ffi))
    for m in mods[1:]:
        src = open(os.path.join(ROOT, "shim", "src", m + ".rs")).read()
        for call in re.findall(r"ffi::(rt_\w+)", src):
            assert call in declared, "%s.rs calls undeclared %s" % (m, call)
        for c in re.findall(r"ffi::(RT_\w+)", src):
            if c.endswith("_"):      # `ffi::RT_ACCEL_*` in a doc comment
                continue
            assert c in consts, "%s.rs uses undeclared %s" % (m, c)
