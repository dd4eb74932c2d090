"""CPU: the C-ABI library loads without a GPU and exports exactly what include/rptx.h declares.
No compute call is made here (there is no device in this container)."""
import ctypes as C
import os
import re

import pytest

from rpt_b200 import _abi as A

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "rptx.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rptx_[a-z0-9_]+)\s*\(", text)))


def test_header_and_prototypes_agree():
    syms = _header_symbols()
    assert len(syms) >= 25
    assert sorted(A.PROTOTYPES) == syms


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(A.LIB_PATH)
    for name in _header_symbols():
        assert hasattr(lib, name), f"librptx.so does not export {name}"
    assert A.load().rptx_abi_version() == 3


def test_struct_layouts_match_the_header():
    # sizes follow from the C declarations in include/rptx.h (natural alignment, LP64)
    assert C.sizeof(A.ShapeDesc) == 4 + 4 + 128 + 4 + 4 + 24 + 8 + 16 + 8
    assert C.sizeof(A.MaterialDesc) == 8 + 24 + 24
    assert C.sizeof(A.LightDesc) == 8 + 48 + C.sizeof(A.ShapeDesc) + C.sizeof(A.MaterialDesc)
    assert C.sizeof(A.CameraDesc) == 96
    assert C.sizeof(A.Params) == 8 + 8 + 16 + 16 + 8 + 8 + 8 + 8


def test_no_device_is_an_error_not_a_fallback():
    """Without a CUDA device the product must fail loudly (no CPU path exists)."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a device is present")
    from rpt_b200 import Context, RptxError

    with pytest.raises(RptxError, match="no CUDA device"):
        Context(0)


def test_product_does_not_import_the_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg = os.path.join(ROOT, "rpt_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f)).read()
                for needle in ("liborc", "import orc", "oracle/", "orc_"):
                    assert needle not in text, f"rpt_b200/{f} references the oracle ({needle})"


def test_rust_sys_crate_declares_the_same_abi():
    """rust/rpt-cuda-sys/src/lib.rs (the binding a Rust `rpt` mirror links; not compiled here: no rustc) names exactly
    the header's entry points, with the header's argument counts, and its #[repr(C)] records have the header's fields."""
    rs = open(os.path.join(ROOT, "rust", "rpt-cuda-sys", "src", "lib.rs")).read()
    rs_fns = {m.group(1): m.group(2) for m in re.finditer(r"pub fn (rptx_[a-z0-9_]+)\(([^;]*?)\)\s*(?:->[^;]+)?;", rs, flags=re.S)}
    assert sorted(rs_fns) == _header_symbols()
    htext = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "rptx.h")).read(), flags=re.S)
    for name, args in rs_fns.items():
        m = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", htext, flags=re.S)
        assert m, name
        hargs = m.group(1).strip()
        n_h = 0 if hargs in ("", "void") else hargs.count(",") + 1
        n_r = 0 if not args.strip() else args.count(":")
        assert n_h == n_r, (name, n_h, n_r)
    assert f"RPTX_ABI_VERSION: c_int = {A.load().rptx_abi_version()}" in rs
    for struct, ctype in (("rptx_shape_desc", A.ShapeDesc), ("rptx_material", A.MaterialDesc), ("rptx_light", A.LightDesc),
                          ("rptx_camera", A.CameraDesc), ("rptx_params", A.Params)):
        body = re.search(r"pub struct " + struct + r" \{(.*?)\n\}", rs, flags=re.S).group(1)
        fields = re.findall(r"pub (\w+):", body)
        assert fields == [f[0] for f in ctype._fields_], struct


def _split_top_level(args):
    """Number of top-level comma-separated arguments of a call's argument text."""
    depth, n, seen = 0, 0, False
    for ch in args:
        if ch in "([{<":
            depth += 1
        elif ch in ")]}>":
            depth -= 1
        elif ch == "," and depth == 0:
            n += 1
            continue
        if not ch.isspace():
            seen = True
    return n + 1 if seen else 0


def test_rust_mirror_crate_calls_the_declared_abi():
    """rust/rpt (the `rpt` API mirror; source only, no rustc here): every `sys::rptx_*` call names an entry point of
    include/rptx.h with the header's argument count, the hot-path entry points are all reached, and the reference's public
    builder / render methods exist by name."""
    src_dir = os.path.join(ROOT, "rust", "rpt", "src")
    text = "".join(open(os.path.join(src_dir, f)).read() for f in sorted(os.listdir(src_dir)) if f.endswith(".rs"))
    htext = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "rptx.h")).read(), flags=re.S)
    called = set()
    for m in re.finditer(r"sys::(rptx_[a-z0-9_]+)\s*\(", text):
        name, i, depth = m.group(1), m.end(), 1
        j = i
        while depth:
            depth += {"(": 1, ")": -1}.get(text[j], 0)
            j += 1
        n_call = _split_top_level(text[i:j - 1])
        h = re.search(r"\b" + name + r"\s*\(([^;]*?)\)\s*;", htext, flags=re.S)
        assert h, f"rust/rpt calls {name}, which include/rptx.h does not declare"
        hargs = h.group(1).strip()
        n_h = 0 if hargs in ("", "void") else hargs.count(",") + 1
        assert n_call == n_h, (name, n_call, n_h)
        called.add(name)
    for must in ("rptx_ctx_create", "rptx_scene_create", "rptx_scene_add_mesh", "rptx_scene_add_object", "rptx_scene_add_light", "rptx_scene_set_medium",
                 "rptx_scene_set_environment_color", "rptx_scene_set_environment_hdri", "rptx_scene_commit", "rptx_render_begin", "rptx_render_sample",
                 "rptx_film_image_rgb8", "rptx_film_variance", "rptx_photon_render", "rptx_intersect_batch", "rptx_comm_init", "rptx_last_error"):
        assert must in called, must
    # the reference's API surface (src/renderer.rs:60-156, src/photon.rs:642-655, src/scene.rs:34-81, src/shape.rs:180-314)
    for name in ("pub fn new(scene: &'a Scene, camera: Camera)", "pub fn width(", "pub fn height(", "pub fn exposure_value(", "pub fn stepsize(", "pub fn filter(",
                 "pub fn max_bounces(", "pub fn num_samples(", "pub fn gather_size(", "pub fn gather_size_volume(", "pub fn watts(", "pub fn render(&self) -> RgbImage",
                 "pub fn iterative_render<F>(&self, callback_interval: u32, mut callback: F)", "pub fn get_closest_hit(", "pub fn photon_map_render(",
                 "pub fn photon_point_query_beam_render(", "pub fn photon_beam_query_beam_render(", "pub fn photon_render(", "impl SceneAdd<LightAndMeshObject> for Scene",
                 "impl SceneAdd<(Transformed<Cube>, Material)> for Scene", "pub trait Transformable<T>", "pub fn polygon(", "pub fn homogeneous_isotropic(",
                 "pub fn colored_glowing_fog(", "pub fn look_at(", "pub fn focus(", "pub fn load_obj(", "pub fn load_stl(", "pub fn image(&self) -> RgbImage",
                 "pub fn variance(&self) -> f64"):
        assert name in text, name
