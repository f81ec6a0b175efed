#!/usr/bin/env python
"""Writes rayca-cuda/src/ffi.rs: `#[repr(C)]` mirrors of every struct of include/rayca_cuda.h (generated from the ctypes
mirror in rayca_b200/model.py, which tests/abi_layout.c pins against the C compiler) plus the extern "C" block.

    python tools/gen_ffi_rs.py
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rayca_b200 import model as M  # noqa: E402

STRUCTS = ["rayca_trs", "rayca_node", "rayca_geometry", "rayca_primitive", "rayca_mesh", "rayca_material",
           "rayca_pbr_material", "rayca_phong_material", "rayca_ggx_material", "rayca_texture", "rayca_image",
           "rayca_light", "rayca_camera", "rayca_model", "rayca_scene_desc", "rayca_config", "rayca_options",
           "rayca_aux_out", "rayca_stats", "rayca_bvh_info"]
SCALAR = {C.c_float: "f32", C.c_double: "f64", C.c_uint32: "u32", C.c_int32: "i32", C.c_uint64: "u64", C.c_void_p: "*const c_void"}
# pointer fields that ctypes declares as c_void_p: their Rust pointee types
VOID_PTR = {("rayca_scene_desc", "vertices"): "*const RaycaVertex", ("rayca_scene_desc", "indices"): "*const u32",
            ("rayca_image", "data"): "*const c_void",
            ("rayca_aux_out", "hit_blas"): "*mut u32", ("rayca_aux_out", "hit_prim"): "*mut u32", ("rayca_aux_out", "hit_src"): "*mut u32",
            ("rayca_aux_out", "depth"): "*mut f32", ("rayca_aux_out", "uv"): "*mut f32", ("rayca_aux_out", "rgba_f32"): "*mut f32"}


def rust_name(c_name):
    return "".join(p.capitalize() for p in c_name.split("_"))


def rust_type(struct, field, t):
    if (struct, field) in VOID_PTR:
        return VOID_PTR[(struct, field)]
    if t in SCALAR:
        return SCALAR[t]
    if issubclass(t, C.Array):
        return f"[{rust_type(struct, field, t._type_)}; {t._length_}]"
    if issubclass(t, C.Structure):
        return rust_name(t.__name__)
    if issubclass(t, C._Pointer):
        inner = t._type_
        return "*const " + (SCALAR[inner] if inner in SCALAR else rust_name(inner.__name__))
    raise TypeError((struct, field, t))


def main():
    out = [open(os.path.join(ROOT, "tools", "ffi_rs_head.txt")).read()]
    for name in STRUCTS:
        t = getattr(M, name)
        out.append(f"/// `{name}` (include/rayca_cuda.h), {C.sizeof(t)} bytes\n#[repr(C)]\n#[derive(Clone, Copy)]\npub struct {rust_name(name)} {{")
        for f, ft in t._fields_:
            out.append(f"    pub {f}: {rust_type(name, f, ft)},")
        out.append("}\n")
        out.append(f"const _: () = assert!(core::mem::size_of::<{rust_name(name)}>() == {C.sizeof(t)});\n")
    out.append(open(os.path.join(ROOT, "tools", "ffi_rs_tail.txt")).read())
    with open(os.path.join(ROOT, "rayca-cuda", "src", "ffi.rs"), "w") as f:
        f.write("\n".join(out))


if __name__ == "__main__":
    main()
