"""CPU-side checks of the product's host code: the C-ABI library loads and exports every symbol
include/rayca_cuda.h declares, POD struct sizes agree with the ctypes mirror, and the binned BLAS
builder reproduces the reference builder (as restated verbatim by the oracle) node for node."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, golden_scenes, load_desc
from rayca_b200 import model as M
from rayca_b200 import renderer as R


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "rayca_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rayca_cuda_[a-z_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(cuda_lib):
    names = declared_symbols()
    assert len(names) >= 14
    for n in names:
        assert hasattr(cuda_lib, n), f"{n} is declared in include/rayca_cuda.h but not exported"
    assert cuda_lib.rayca_cuda_abi_version() == 1
    assert cuda_lib.rayca_cuda_strerror(-2) == b"unsupported configuration"


def test_pod_struct_sizes():
    assert C.sizeof(M.rayca_trs) == 40
    assert C.sizeof(M.rayca_node) == 64
    assert C.sizeof(M.rayca_geometry) == 36
    assert C.sizeof(M.rayca_config) == 40
    assert C.sizeof(M.rayca_options) == 32
    assert C.sizeof(M.rayca_light) == 64
    assert C.sizeof(M.rayca_model) == 40 + 8 + 12 * 8
    assert M.VERTEX_FLOATS * 4 == 72


def test_invalid_arguments_return_codes(cuda_lib):
    h = C.c_void_p()
    assert cuda_lib.rayca_cuda_scene_create_host(None, 1, C.byref(h), None) == -1
    # a scene with no geometry: Tlas::new would assert (tlas.rs:75) -> INVALID_ARG, never a crash
    s = M.Scene()
    s.push_model(M.create_default_model())
    with pytest.raises(R.RaycaError) as e:
        R.DeviceScene(None, s.to_desc())
    assert e.value.code == -1


@pytest.mark.parametrize("scene", golden_scenes())
@pytest.mark.parametrize("bvh", [True, False])
def test_builder_matches_reference_builder(oracle, cuda_lib, scene, bvh):
    if scene == "scene7" and not bvh:
        pytest.skip("covered by the bvh=True case")
    desc = load_desc(scene)
    osc = oracle.OracleScene(desc, bvh=bvh)
    psc = R.DeviceScene(None, desc, bvh=bvh)
    oi, on, ono, opo, osrc = osc.bvh()
    pi, pn, pno, ppo, psrc = psc.bvh()
    assert (oi.blas_count, oi.node_count, oi.triangle_count) == (pi.blas_count, pi.node_count, pi.triangle_count)
    assert np.array_equal(ono, pno) and np.array_equal(opo, ppo)
    assert np.array_equal(on.view(np.uint32), pn.view(np.uint32)), "node boxes / links differ"
    assert np.array_equal(osrc, psrc), "primitive order differs"
    if not bvh:
        assert oi.node_count == 2 * oi.blas_count  # max_depth(0): every BLAS root stays a leaf


def test_builder_on_random_soup(oracle, cuda_lib):
    rng = np.random.default_rng(7)
    n = 3000
    c = rng.uniform(-5, 5, (n, 1, 3))
    tris = (c + rng.normal(0, 0.3, (n, 3, 3))).astype(np.float32)
    tris[:50] = tris[0]  # duplicates: identical centroids must not break the partition
    s = M.Scene()
    s.push_model(M.triangle_soup_model(tris))
    s.push_model(M.create_default_model())
    d = s.to_desc()
    oi, on, _, _, osrc = oracle.OracleScene(d).bvh()
    pi, pn, _, _, psrc = R.DeviceScene(None, d).bvh()
    assert oi.node_count == pi.node_count
    assert np.array_equal(on.view(np.uint32), pn.view(np.uint32)) and np.array_equal(osrc, psrc)


def test_product_does_not_import_the_oracle():
    """The oracle is test infrastructure: nothing under rayca_b200/ may import, link or execute it."""
    for root, _, files in os.walk(os.path.join(ROOT, "rayca_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(root, f)).read()
                for needle in ("import oracle", "from oracle", "librayca_oracle", "rayca_oracle_", "oracle_math"):
                    assert needle not in text, f"{f} references the oracle ({needle})"


def test_header_compiles_as_c_and_layout_is_pinned(tmp_path):
    """include/rayca_cuda.h compiles as plain C11 and every struct has the layout tests/abi_layout.c freezes with
    _Static_asserts; the offsets / sizes the C compiler reports equal the ctypes mirror field by field (what the
    #[repr(C)] structs of rayca-cuda/src/ffi.rs must reproduce too)."""
    import subprocess

    exe = tmp_path / "abi_layout"
    cc = "/usr/bin/gcc" if os.path.exists("/usr/bin/gcc") else "gcc"
    subprocess.check_call([cc, "-std=c11", "-Wall", "-Werror", "-pedantic", "-o", str(exe), os.path.join(ROOT, "tests", "abi_layout.c")])
    lines = subprocess.check_output([str(exe)], text=True).split("\n")
    seen = set()
    for ln in lines:
        if not ln:
            continue
        name, field, off, size = ln.split()
        seen.add(name)
        if name == "rayca_vertex":
            assert int(size) == M.VERTEX_FLOATS * 4
            continue
        t = getattr(M, name)
        if field == ".":
            assert C.sizeof(t) == int(size), name
        else:
            d = getattr(t, field)
            assert (d.offset, d.size) == (int(off), int(size)), f"{name}.{field}"
    # every struct the header typedefs is covered (opaque handles excepted)
    text = open(os.path.join(ROOT, "include", "rayca_cuda.h")).read()
    declared = set(re.findall(r"typedef struct (rayca_[a-z_]+) \{", text))
    assert declared == seen, declared ^ seen


def test_glb_container_and_embedded_buffers(tmp_path):
    """the harness loader reads binary glTF (.glb), buffers without a uri (the BIN chunk,
    `Source::Bin` — `unimplemented!()` in rayca-model/src/loader/gltf.rs:239-241) and images stored in a bufferView;
    the model equals the one loaded from the equivalent .gltf with base64 data URIs."""
    import base64
    import io
    import json
    import struct

    from PIL import Image as PILImage

    pos = np.array([[-1, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32)
    uv = np.array([[0, 0], [1, 0], [0.5, 1]], dtype=np.float32)
    idx = np.array([0, 1, 2], dtype=np.uint16)
    png = io.BytesIO()
    PILImage.fromarray((np.arange(4 * 4 * 3) % 251).astype(np.uint8).reshape(4, 4, 3), "RGB").save(png, format="PNG")
    png = png.getvalue()
    parts = [pos.tobytes(), uv.tobytes(), idx.tobytes() + b"\0\0", png + b"\0" * ((4 - len(png) % 4) % 4)]
    offs = np.cumsum([0] + [len(p) for p in parts])
    blob = b"".join(parts)
    doc = {"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"mesh": 0, "translation": [0, 0.5, 0]}],
           "meshes": [{"primitives": [{"attributes": {"POSITION": 0, "TEXCOORD_0": 1}, "indices": 2, "material": 0}]}],
           "materials": [{"pbrMetallicRoughness": {"baseColorTexture": {"index": 0}, "metallicFactor": 0.0}}],
           "textures": [{"source": 0}], "images": [{"bufferView": 3, "mimeType": "image/png"}],
           "buffers": [{"byteLength": len(blob)}],
           "bufferViews": [{"buffer": 0, "byteOffset": int(offs[0]), "byteLength": 36}, {"buffer": 0, "byteOffset": int(offs[1]), "byteLength": 24},
                           {"buffer": 0, "byteOffset": int(offs[2]), "byteLength": 6}, {"buffer": 0, "byteOffset": int(offs[3]), "byteLength": len(png)}],
           "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"},
                         {"bufferView": 1, "componentType": 5126, "count": 3, "type": "VEC2"},
                         {"bufferView": 2, "componentType": 5123, "count": 3, "type": "SCALAR"}]}
    js = json.dumps(doc).encode()
    js += b" " * ((4 - len(js) % 4) % 4)
    glb = b"glTF" + struct.pack("<II", 2, 12 + 8 + len(js) + 8 + len(blob)) + struct.pack("<II", len(js), 0x4E4F534A) + js + \
        struct.pack("<II", len(blob), 0x004E4942) + blob
    (tmp_path / "tri.glb").write_bytes(glb)
    doc2 = json.loads(json.dumps(doc))
    doc2["buffers"][0]["uri"] = "data:application/octet-stream;base64," + base64.b64encode(blob).decode()
    doc2["images"][0] = {"uri": "data:image/png;base64," + base64.b64encode(png).decode()}
    (tmp_path / "tri.gltf").write_text(json.dumps(doc2))
    a, b = M.load_gltf_path(tmp_path / "tri.glb"), M.load_gltf_path(tmp_path / "tri.gltf")
    for m in (a, b):
        assert len(m.geometries) == 1 and len(m.images) == 1 and m.images[0].width == 4 and m.pbr_materials[0].albedo == 0
    assert np.array_equal(a.geometries[0].vertices, b.geometries[0].vertices)
    assert np.array_equal(a.geometries[0].indices, b.geometries[0].indices)
    assert np.array_equal(a.images[0].data, b.images[0].data)
    assert np.array_equal(a.geometries[0].vertices[:, :3], pos)
    (tmp_path / "bad.glb").write_bytes(b"glTF" + struct.pack("<II", 1, 12) + b"\0" * 8)
    with pytest.raises(ValueError):
        M.load_gltf_path(tmp_path / "bad.glb")


def test_rust_ffi_mirror_lists_the_same_fields():
    """rayca-cuda/src/ffi.rs (generated by tools/gen_ffi_rs.py; uncompiled here — no Rust toolchain) declares a
    #[repr(C)] struct for every struct of the header with the same fields in the same order, and an extern "C"
    declaration for every exported entry point."""
    ffi = open(os.path.join(ROOT, "rayca-cuda", "src", "ffi.rs")).read()
    text = open(os.path.join(ROOT, "include", "rayca_cuda.h")).read()
    declared = set(re.findall(r"typedef struct (rayca_[a-z_]+) \{", text))
    for name in sorted(declared):
        rust = "".join(p.capitalize() for p in name.split("_"))  # rayca_pbr_material -> RaycaPbrMaterial
        m = re.search(r"#\[repr\(C\)\]\s*#\[derive\([^)]*\)\]\s*pub struct " + rust + r" \{(.*?)\n\}", ffi, flags=re.S)
        assert m, f"{rust} is missing from rayca-cuda/src/ffi.rs"
        rust_fields = re.findall(r"pub ([a-z0-9_]+):", m.group(1))
        if name == "rayca_vertex":
            assert rust_fields == ["pos", "color", "normal", "tangent", "bitangent", "uv"]
        else:
            t = getattr(M, name)
            assert rust_fields == [f for f, _ in t._fields_], name
            assert f"size_of::<{rust}>() == {C.sizeof(t)}" in ffi, f"{rust}: size assertion"
    rust_fns = set(re.findall(r"pub fn (rayca_cuda_[a-z_]+)\(", ffi))
    assert rust_fns == set(declared_symbols()), rust_fns ^ set(declared_symbols())


def test_png_encode_round_trips(cuda_lib, tmp_path):
    """`Image::dump_png` (rayca-model/src/image.rs:160-169) inside the library.  The decoded file
    holds the frame's bytes exactly (smooth gradients, noise and flat areas exercise all five scanline filters);
    sizes that are not multiples of anything; the two-call size protocol of rayca_cuda_encode_png."""
    import io

    from PIL import Image as PILImage

    rng = np.random.default_rng(3)
    for w, h in ((1, 1), (7, 3), (129, 65), (640, 360)):
        img = R.Image(w, h)
        yy, xx = np.mgrid[0:h, 0:w]
        img.buffer[..., 0] = (xx * 255 // max(w - 1, 1)).astype(np.uint8)
        img.buffer[..., 1] = ((xx + yy) % 256).astype(np.uint8)
        img.buffer[..., 2] = rng.integers(0, 256, (h, w), dtype=np.uint8)
        img.buffer[..., 3] = 255
        img.buffer[h // 2:, : w // 2] = (10, 20, 30, 128)
        png = img.encode_png()
        assert png[:8] == b"\x89PNG\r\n\x1a\n"
        back = PILImage.open(io.BytesIO(png))
        assert back.mode == "RGBA" and back.size == (w, h)
        assert np.array_equal(np.asarray(back), img.buffer)
        path = tmp_path / f"f{w}x{h}.png"
        img.dump_png(path)
        assert np.array_equal(np.asarray(PILImage.open(path)), img.buffer)
    assert len(png) < 640 * 360 * 4  # compressed
    n = C.c_size_t()
    buf = np.zeros((2, 2, 4), dtype=np.uint8)
    assert cuda_lib.rayca_cuda_encode_png(buf.ctypes.data, 2, 2, None, 0, C.byref(n)) == -1 and n.value > 50
    assert cuda_lib.rayca_cuda_encode_png(None, 2, 2, None, 0, C.byref(n)) == -1
    assert cuda_lib.rayca_cuda_write_png(str(tmp_path / "no" / "such" / "dir.png").encode(), buf.ctypes.data, 2, 2) == -1
