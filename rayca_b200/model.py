"""Host-side mirror of the `rayca-model` data types the renderer consumes.

These classes are the Python stand-in for what a Rust `rayca-cuda` crate would receive as
`&rayca_model::Scene` (rayca-model/src/scene.rs:47-53): scene nodes, models, model nodes,
meshes, primitives, geometries, materials, textures, images, lights, cameras.  They carry
no rendering logic: `Scene.to_desc()` serialises the graph into the POD
`rayca_scene_desc` of include/rayca_cuda.h, and everything from scene flattening onwards
happens inside librayca_cuda.

The two loaders (glTF, SDTF) restate only as much of rayca-model/src/loader/{gltf,sdtf}.rs
as the repository's fixtures need; asset loading is outside the hot path (SURVEY §2 rows
15-16) and exists here so the test-suite can read the same files the reference tests read.
All arithmetic the loaders perform (bitangents, SDTF face normals, look-at camera,
transform stack) is done in float32 in the reference's operation order.
"""
from __future__ import annotations

import base64
import ctypes as C
import json
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

f32 = np.float32
NONE = -1

# ----------------------------------------------------------------------------- ctypes POD
# field-for-field copies of include/rayca_cuda.h


class rayca_trs(C.Structure):
    _fields_ = [("translation", C.c_float * 3), ("rotation", C.c_float * 4), ("scale", C.c_float * 3)]


class rayca_node(C.Structure):
    _fields_ = [("trs", rayca_trs), ("mesh", C.c_int32), ("camera", C.c_int32), ("light", C.c_int32),
                ("model", C.c_int32), ("first_child", C.c_uint32), ("child_count", C.c_uint32)]


class rayca_geometry(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("first_vertex", C.c_uint32), ("vertex_count", C.c_uint32),
                ("first_index", C.c_uint32), ("index_count", C.c_uint32), ("center", C.c_float * 3),
                ("radius", C.c_float)]


class rayca_primitive(C.Structure):
    _fields_ = [("geometry", C.c_uint32), ("material", C.c_int32)]


class rayca_mesh(C.Structure):
    _fields_ = [("first_primitive", C.c_uint32), ("primitive_count", C.c_uint32)]


class rayca_material(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("index", C.c_int32)]


class rayca_pbr_material(C.Structure):
    _fields_ = [("color", C.c_float * 4), ("albedo", C.c_int32), ("normal", C.c_int32),
                ("metallic_factor", C.c_float), ("roughness_factor", C.c_float), ("metallic_roughness", C.c_int32)]


class rayca_phong_material(C.Structure):
    _fields_ = [("ambient", C.c_float * 4), ("emission", C.c_float * 4), ("diffuse", C.c_float * 4),
                ("specular", C.c_float * 4), ("shininess", C.c_float)]


class rayca_ggx_material(C.Structure):
    _fields_ = [("diffuse", C.c_float * 4), ("specular", C.c_float * 4), ("roughness", C.c_float)]


class rayca_texture(C.Structure):
    _fields_ = [("image", C.c_int32)]


class rayca_image(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("color_type", C.c_uint32),
                ("reserved", C.c_uint32), ("data", C.c_void_p)]


class rayca_light(C.Structure):
    _fields_ = [("kind", C.c_uint32), ("color", C.c_float * 4), ("intensity", C.c_float),
                ("attenuation", C.c_float * 3), ("ab", C.c_float * 3), ("ac", C.c_float * 3),
                ("material", C.c_int32)]


class rayca_camera(C.Structure):
    _fields_ = [("yfov_radians", C.c_float)]


class rayca_model(C.Structure):
    _fields_ = [("root_trs", rayca_trs), ("root_first_child", C.c_uint32), ("root_child_count", C.c_uint32)] + [
        (f"{n}_{k}", C.c_uint32)
        for n in ("node", "mesh", "primitive", "geometry", "material", "pbr", "phong", "ggx", "texture", "image",
                  "light", "camera")
        for k in ("offset", "count")
    ]


class rayca_scene_desc(C.Structure):
    _fields_ = [
        ("root_trs", rayca_trs), ("root_first_child", C.c_uint32), ("root_child_count", C.c_uint32),
        ("scene_nodes", C.POINTER(rayca_node)), ("scene_node_count", C.c_uint32),
        ("models", C.POINTER(rayca_model)), ("model_count", C.c_uint32),
        ("nodes", C.POINTER(rayca_node)), ("node_count", C.c_uint32),
        ("children", C.POINTER(C.c_uint32)), ("children_count", C.c_uint32),
        ("meshes", C.POINTER(rayca_mesh)), ("mesh_count", C.c_uint32),
        ("mesh_primitives", C.POINTER(C.c_uint32)), ("mesh_primitive_count", C.c_uint32),
        ("primitives", C.POINTER(rayca_primitive)), ("primitive_count", C.c_uint32),
        ("geometries", C.POINTER(rayca_geometry)), ("geometry_count", C.c_uint32),
        ("vertices", C.c_void_p), ("vertex_count", C.c_uint64),
        ("indices", C.c_void_p), ("index_count", C.c_uint64),
        ("materials", C.POINTER(rayca_material)), ("material_count", C.c_uint32),
        ("pbr_materials", C.POINTER(rayca_pbr_material)), ("pbr_count", C.c_uint32),
        ("phong_materials", C.POINTER(rayca_phong_material)), ("phong_count", C.c_uint32),
        ("ggx_materials", C.POINTER(rayca_ggx_material)), ("ggx_count", C.c_uint32),
        ("textures", C.POINTER(rayca_texture)), ("texture_count", C.c_uint32),
        ("images", C.POINTER(rayca_image)), ("image_count", C.c_uint32),
        ("lights", C.POINTER(rayca_light)), ("light_count", C.c_uint32),
        ("cameras", C.POINTER(rayca_camera)), ("camera_count", C.c_uint32),
    ]


class rayca_config(C.Structure):
    _fields_ = [("bvh", C.c_uint32), ("light_samples", C.c_uint32), ("light_stratify", C.c_uint32),
                ("samples_per_pixel", C.c_uint32), ("russian_roulette", C.c_uint32), ("direct_sampler", C.c_uint32),
                ("indirect_sampler", C.c_uint32), ("integrator", C.c_uint32), ("max_depth", C.c_uint32),
                ("gamma", C.c_float)]


class rayca_options(C.Structure):
    _fields_ = [("rng_seed", C.c_uint64), ("traversal", C.c_uint32), ("shard_index", C.c_uint32),
                ("shard_count", C.c_uint32), ("tile_rows", C.c_uint32), ("flags", C.c_uint32),
                ("reserved", C.c_uint32)]


class rayca_aux_out(C.Structure):
    _fields_ = [("hit_blas", C.c_void_p), ("hit_prim", C.c_void_p), ("hit_src", C.c_void_p),
                ("depth", C.c_void_p), ("uv", C.c_void_p), ("rgba_f32", C.c_void_p)]


class rayca_stats(C.Structure):
    _fields_ = [("build_ms", C.c_double), ("upload_ms", C.c_double), ("trace_ms", C.c_double),
                ("readback_ms", C.c_double), ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64),
                ("rays_secondary", C.c_uint64), ("node_bytes", C.c_uint64), ("triangle_bytes", C.c_uint64),
                ("kernel_launches", C.c_uint32), ("reserved", C.c_uint32)]


class rayca_bvh_info(C.Structure):
    _fields_ = [("blas_count", C.c_uint32), ("node_count", C.c_uint32), ("triangle_count", C.c_uint32),
                ("tlas_node_count", C.c_uint32)]


# enums (include/rayca_cuda.h)
GEOMETRY_TRIANGLE_MESH, GEOMETRY_SPHERE = 0, 1
MATERIAL_PBR, MATERIAL_PHONG, MATERIAL_GGX = 0, 1, 2
COLOR_RGB8, COLOR_RGBA8, COLOR_RGBA32F = 0, 1, 2
LIGHT_DIRECTIONAL, LIGHT_POINT, LIGHT_QUAD = 0, 1, 2
VERTEX_FLOATS = 18  # pos3 color4 normal3 tangent3 bitangent3 uv2

# ----------------------------------------------------------------------- float32 helpers
# rayca-math restated for the loaders only (look-at camera, SDTF transform stack).


def _v(x, y, z):
    return np.array([x, y, z], dtype=f32)


def _dot(a, b):
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2])) + f32(0.0)


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], dtype=f32)


EPS = f32(np.finfo(np.float32).eps) * f32(8192.0)


def _normalized(a):
    l = np.sqrt(_dot(a, a))
    return (a / l).astype(f32) if l > EPS else a


def quat_rotate(q, v):
    """rayca-math/src/vec3.rs:148-159"""
    u = np.array(q[:3], dtype=f32)
    s = f32(q[3])
    a = (f32(2.0) * _dot(u, v)) * u
    b = (s * s - _dot(u, u)) * v
    c = (f32(2.0) * s) * _cross(u, v)
    return ((a + b) + c).astype(f32)


def quat_mul(a, b):
    """rayca-math/src/quat.rs:236-258"""
    ax, ay, az, aw = [f32(t) for t in a]
    bx, by, bz, bw = [f32(t) for t in b]
    return np.array([
        ax * bw + ay * bz - az * by + aw * bx,
        -ax * bz + ay * bw + az * bx + aw * by,
        ax * by - ay * bx + az * bw + aw * bz,
        -ax * bx - ay * by - az * bz + aw * bw,
    ], dtype=f32)


def quat_conjugate(q):
    return np.array([-q[0], -q[1], -q[2], q[3]], dtype=f32)


def quat_from_rows(m):
    """rayca-math/src/quat.rs:186-228 (m[row][col])."""
    m = np.asarray(m, dtype=f32)
    t = m[0][0] + m[1][1] + m[2][2]
    if t > 0:
        s = f32(0.5) / np.sqrt(t + f32(1.0))
        q = [(m[2][1] - m[1][2]) * s, (m[0][2] - m[2][0]) * s, (m[1][0] - m[0][1]) * s, f32(0.25) / s]
    elif m[0][0] > m[1][1] and m[0][0] > m[2][2]:
        s = f32(2.0) * np.sqrt(f32(1.0) + m[0][0] - m[1][1] - m[2][2])
        q = [f32(0.25) * s, (m[0][1] + m[1][0]) / s, (m[0][2] + m[2][0]) / s, (m[2][1] - m[1][2]) / s]
    elif m[1][1] > m[2][2]:
        s = f32(2.0) * np.sqrt(f32(1.0) + m[1][1] - m[0][0] - m[2][2])
        q = [(m[0][1] + m[1][0]) / s, f32(0.25) * s, (m[1][2] + m[2][1]) / s, (m[0][2] - m[2][0]) / s]
    else:
        s = f32(2.0) * np.sqrt(f32(1.0) + m[2][2] - m[0][0] - m[1][1])
        q = [(m[0][2] + m[2][0]) / s, (m[1][2] + m[2][1]) / s, f32(0.25) * s, (m[1][0] - m[0][1]) / s]
    q = np.array(q, dtype=f32)
    l = np.sqrt(f32(f32(f32(q[0] * q[0]) + f32(q[1] * q[1])) + f32(q[2] * q[2])) + f32(q[3] * q[3]))
    return (q / l).astype(f32)


def quat_axis_angle(axis, angle_radians):
    """rayca-math/src/quat.rs:66-75"""
    half = f32(angle_radians) / f32(2.0)
    factor = f32(np.sin(half))
    q = np.array([axis[0] * factor, axis[1] * factor, axis[2] * factor, f32(np.cos(half))], dtype=f32)
    l = np.sqrt(f32(f32(f32(q[0] * q[0]) + f32(q[1] * q[1])) + f32(q[2] * q[2])) + f32(q[3] * q[3]))
    return (q / l).astype(f32)


@dataclass
class Trs:
    """rayca-math/src/trs.rs:77-86"""
    translation: np.ndarray = field(default_factory=lambda: _v(0, 0, 0))
    rotation: np.ndarray = field(default_factory=lambda: np.array([0, 0, 0, 1], dtype=f32))
    scale: np.ndarray = field(default_factory=lambda: _v(1, 1, 1))

    def copy(self):
        return Trs(self.translation.copy(), self.rotation.copy(), self.scale.copy())

    def left_mul(self, rhs: "Trs"):
        """trs.rs:111-118"""
        translation = self.translation + quat_rotate(self.rotation, self.scale * rhs.translation)
        rotation = quat_mul(self.rotation, rhs.rotation)
        scale = quat_rotate(quat_conjugate(rhs.rotation), self.scale * quat_rotate(rhs.rotation, rhs.scale))
        self.translation, self.rotation, self.scale = translation.astype(f32), rotation, scale.astype(f32)

    def to_c(self) -> rayca_trs:
        t = rayca_trs()
        t.translation[:] = [float(x) for x in self.translation]
        t.rotation[:] = [float(x) for x in self.rotation]
        t.scale[:] = [float(x) for x in self.scale]
        return t


# ------------------------------------------------------------------------- scene classes


@dataclass
class Node:
    """rayca-model/src/node.rs:10-31"""
    trs: Trs = field(default_factory=Trs)
    children: List[int] = field(default_factory=list)
    mesh: int = NONE
    camera: int = NONE
    light: int = NONE
    model: int = NONE
    name: str = "Unknown"


@dataclass
class TriangleMesh:
    vertices: np.ndarray  # (n, 18) float32
    indices: np.ndarray   # (m,) uint32


@dataclass
class Sphere:
    center: np.ndarray
    radius: float


@dataclass
class Primitive:
    geometry: int
    material: int = NONE


@dataclass
class Mesh:
    primitives: List[int] = field(default_factory=list)


@dataclass
class Material:
    kind: int = MATERIAL_PBR
    index: int = NONE


@dataclass
class PbrMaterial:
    """rayca-model/src/material/pbr.rs:63-72; `Default` derives give colour WHITE and factors 0."""
    color: tuple = (1.0, 1.0, 1.0, 1.0)
    albedo: int = NONE
    normal: int = NONE
    metallic_factor: float = 0.0
    roughness_factor: float = 0.0
    metallic_roughness: int = NONE


@dataclass
class PhongMaterial:
    ambient: tuple = (0.0, 0.0, 0.0, 1.0)
    emission: tuple = (0.0, 0.0, 0.0, 1.0)
    diffuse: tuple = (0.0, 0.0, 0.0, 1.0)
    specular: tuple = (0.0, 0.0, 0.0, 1.0)
    shininess: float = 0.0


@dataclass
class GgxMaterial:
    diffuse: tuple = (0.0, 0.0, 0.0, 1.0)
    specular: tuple = (0.0, 0.0, 0.0, 1.0)
    roughness: float = 0.0


@dataclass
class Texture:
    image: int = NONE


@dataclass
class Image:
    """rayca-model/src/image.rs:26-37"""
    width: int
    height: int
    color_type: int
    data: np.ndarray  # contiguous uint8 (or float32 for RGBA32F)


@dataclass
class Light:
    kind: int = LIGHT_POINT
    color: tuple = (1.0, 1.0, 1.0, 1.0)
    intensity: float = 1.0
    attenuation: tuple = (0.0, 0.0, 1.0)  # PointLight::new (light/point.rs:16-22)
    ab: tuple = (0.0, 0.0, 0.0)
    ac: tuple = (0.0, 0.0, 0.0)
    material: int = NONE


@dataclass
class Camera:
    yfov_radians: float = float(f32(np.pi / 4))  # Camera::default (camera.rs:79-83)


@dataclass
class Model:
    """rayca-model/src/model.rs:31-51"""
    root: Node = field(default_factory=Node)
    nodes: List[Node] = field(default_factory=list)
    meshes: List[Mesh] = field(default_factory=list)
    primitives: List[Primitive] = field(default_factory=list)
    geometries: list = field(default_factory=list)
    materials: List[Material] = field(default_factory=list)
    pbr_materials: List[PbrMaterial] = field(default_factory=list)
    phong_materials: List[PhongMaterial] = field(default_factory=list)
    ggx_materials: List[GgxMaterial] = field(default_factory=list)
    textures: List[Texture] = field(default_factory=list)
    images: List[Image] = field(default_factory=list)
    lights: List[Light] = field(default_factory=list)
    cameras: List[Camera] = field(default_factory=list)

    def get_mut_last_camera(self) -> Optional[Node]:
        """rayca-model/src/loader/gltf.rs:580-586"""
        for n in reversed(self.nodes):
            if n.camera != NONE:
                return n
        return None


def create_default_model() -> Model:
    """SoftRenderer::create_default_model (rayca-soft/src/scene.rs:18-55): one camera at z=4 and
    two point lights of intensity 1024 at (-1,4,3) and (1,4,3) sharing one Light."""
    m = Model()
    m.cameras.append(Camera())
    m.nodes.append(Node(camera=0, trs=Trs(translation=_v(0.0, 0.0, 4.0))))
    m.root.children.append(0)
    m.lights.append(Light(kind=LIGHT_POINT, intensity=1024.0))
    m.nodes.append(Node(light=0, trs=Trs(translation=_v(-1.0, 4.0, 3.0))))
    m.root.children.append(1)
    m.nodes.append(Node(light=0, trs=Trs(translation=_v(1.0, 4.0, 3.0))))
    m.root.children.append(2)
    return m


_DESC_ARRAYS = [  # (pointer field, count field, element ctype or None for raw numpy)
    ("scene_nodes", "scene_node_count", rayca_node), ("models", "model_count", rayca_model),
    ("nodes", "node_count", rayca_node), ("children", "children_count", C.c_uint32),
    ("meshes", "mesh_count", rayca_mesh), ("mesh_primitives", "mesh_primitive_count", C.c_uint32),
    ("primitives", "primitive_count", rayca_primitive), ("geometries", "geometry_count", rayca_geometry),
    ("materials", "material_count", rayca_material), ("pbr_materials", "pbr_count", rayca_pbr_material),
    ("phong_materials", "phong_count", rayca_phong_material), ("ggx_materials", "ggx_count", rayca_ggx_material),
    ("textures", "texture_count", rayca_texture), ("lights", "light_count", rayca_light),
    ("cameras", "camera_count", rayca_camera),
]
_DESC_FORMAT = 1


class SceneDesc:
    """Owns the ctypes arrays a `rayca_scene_desc` points into."""

    def __init__(self):
        self.desc = rayca_scene_desc()
        self._keep = []
        self.triangle_count = 0

    def ptr(self):
        return C.byref(self.desc)

    # ---- compact on-disk form used for the committed test fixtures (tests/golden/*.npz)
    def save(self, path) -> None:
        d = self.desc
        out = {"format": np.array([_DESC_FORMAT], dtype=np.uint32),
               "root": np.frombuffer(bytes(d.root_trs), dtype=np.uint8),
               "root_children": np.array([d.root_first_child, d.root_child_count], dtype=np.uint32)}
        for ptr, cnt, ctype in _DESC_ARRAYS:
            n = getattr(d, cnt)
            raw = C.string_at(getattr(d, ptr), n * C.sizeof(ctype)) if n else b""
            out[ptr] = np.frombuffer(raw, dtype=np.uint8)
        out["vertices"] = np.frombuffer(C.string_at(d.vertices, d.vertex_count * VERTEX_FLOATS * 4), dtype=f32) \
            if d.vertex_count else np.zeros(0, dtype=f32)
        out["indices"] = np.frombuffer(C.string_at(d.indices, d.index_count * 4), dtype=np.uint32) \
            if d.index_count else np.zeros(0, dtype=np.uint32)
        meta = []
        for i in range(d.image_count):
            im = d.images[i]
            bpp = {COLOR_RGB8: 3, COLOR_RGBA8: 4, COLOR_RGBA32F: 16}[im.color_type]
            out[f"image_{i}"] = np.frombuffer(C.string_at(im.data, im.width * im.height * bpp), dtype=np.uint8)
            meta.append((im.width, im.height, im.color_type))
        out["image_meta"] = np.array(meta, dtype=np.uint32).reshape(-1, 3)
        np.savez_compressed(path, **out)

    @staticmethod
    def load(path) -> "SceneDesc":
        z = np.load(path)
        assert int(z["format"][0]) == _DESC_FORMAT, "fixture written by another descriptor layout"
        sd = SceneDesc()
        d = sd.desc
        C.memmove(C.byref(d.root_trs), z["root"].tobytes(), C.sizeof(rayca_trs))
        d.root_first_child, d.root_child_count = [int(x) for x in z["root_children"]]
        for ptr, cnt, ctype in _DESC_ARRAYS:
            raw = z[ptr].tobytes()
            n = len(raw) // C.sizeof(ctype)
            a = (ctype * max(n, 1))()
            if n:
                C.memmove(a, raw, len(raw))
            sd._keep.append(a)
            setattr(d, ptr, a)
            setattr(d, cnt, n)
        verts = np.ascontiguousarray(z["vertices"], dtype=f32)
        idx = np.ascontiguousarray(z["indices"], dtype=np.uint32)
        sd._keep += [verts, idx]
        d.vertices, d.vertex_count = verts.ctypes.data, len(verts) // VERTEX_FLOATS
        d.indices, d.index_count = idx.ctypes.data, len(idx)
        meta = z["image_meta"]
        imgs = (rayca_image * max(len(meta), 1))()
        for i, (w, h, ct) in enumerate(meta):
            data = np.ascontiguousarray(z[f"image_{i}"])
            sd._keep.append(data)
            imgs[i] = rayca_image(int(w), int(h), int(ct), 0, data.ctypes.data)
        sd._keep.append(imgs)
        d.images, d.image_count = imgs, len(meta)
        sd.triangle_count = len(idx) // 3
        return sd


@dataclass
class Scene:
    """rayca-model/src/scene.rs:47-53"""
    root: Node = field(default_factory=Node)
    nodes: List[Node] = field(default_factory=list)
    models: List[Model] = field(default_factory=list)

    def push_model(self, model: Model) -> int:
        """scene.rs:109-115"""
        self.models.append(model)
        self.nodes.append(Node(model=len(self.models) - 1))
        self.root.children.append(len(self.nodes) - 1)
        return len(self.nodes) - 1

    def push_gltf_from_path(self, path, image_hook=None) -> int:
        return self.push_model(load_gltf_path(path, image_hook=image_hook))

    def push_sdtf_from_path(self, path):
        model, config = load_sdtf_path(path)
        return self.push_model(model), config

    # ---- serialisation to the POD descriptor
    def to_desc(self) -> SceneDesc:
        sd = SceneDesc()
        d = sd.desc
        children: List[int] = []

        def conv_node(n: Node) -> rayca_node:
            r = rayca_node()
            r.trs = n.trs.to_c()
            r.mesh, r.camera, r.light, r.model = n.mesh, n.camera, n.light, n.model
            r.first_child = len(children)
            r.child_count = len(n.children)
            children.extend(n.children)
            return r

        d.root_trs = self.root.trs.to_c()
        d.root_first_child = len(children)
        d.root_child_count = len(self.root.children)
        children.extend(self.root.children)
        scene_nodes = [conv_node(n) for n in self.nodes]

        models, nodes, meshes, mesh_prims, prims, geoms = [], [], [], [], [], []
        mats, pbrs, phongs, ggxs, texs, imgs, lights, cams = [], [], [], [], [], [], [], []
        vert_chunks, index_chunks = [], []
        nvert = nindex = 0
        for m in self.models:
            cm = rayca_model()
            cm.root_trs = m.root.trs.to_c()
            cm.root_first_child = len(children)
            cm.root_child_count = len(m.root.children)
            children.extend(m.root.children)
            cm.node_offset, cm.node_count = len(nodes), len(m.nodes)
            nodes.extend(conv_node(n) for n in m.nodes)
            cm.mesh_offset, cm.mesh_count = len(meshes), len(m.meshes)
            for mesh in m.meshes:
                meshes.append(rayca_mesh(len(mesh_prims), len(mesh.primitives)))
                mesh_prims.extend(mesh.primitives)
            cm.primitive_offset, cm.primitive_count = len(prims), len(m.primitives)
            prims.extend(rayca_primitive(p.geometry, p.material) for p in m.primitives)
            cm.geometry_offset, cm.geometry_count = len(geoms), len(m.geometries)
            for g in m.geometries:
                cg = rayca_geometry()
                if isinstance(g, TriangleMesh):
                    v = np.ascontiguousarray(g.vertices, dtype=f32).reshape(-1, VERTEX_FLOATS)
                    i = np.ascontiguousarray(g.indices, dtype=np.uint32).reshape(-1)
                    cg.kind = GEOMETRY_TRIANGLE_MESH
                    cg.first_vertex, cg.vertex_count = nvert, len(v)
                    cg.first_index, cg.index_count = nindex, len(i)
                    vert_chunks.append(v)
                    index_chunks.append(i)
                    nvert += len(v)
                    nindex += len(i)
                else:
                    cg.kind = GEOMETRY_SPHERE
                    cg.center[:] = [float(x) for x in g.center]
                    cg.radius = float(g.radius)
                geoms.append(cg)
            cm.material_offset, cm.material_count = len(mats), len(m.materials)
            mats.extend(rayca_material(x.kind, x.index) for x in m.materials)
            cm.pbr_offset, cm.pbr_count = len(pbrs), len(m.pbr_materials)
            for p in m.pbr_materials:
                pbrs.append(rayca_pbr_material((C.c_float * 4)(*p.color), p.albedo, p.normal, p.metallic_factor,
                                               p.roughness_factor, p.metallic_roughness))
            cm.phong_offset, cm.phong_count = len(phongs), len(m.phong_materials)
            for p in m.phong_materials:
                phongs.append(rayca_phong_material((C.c_float * 4)(*p.ambient), (C.c_float * 4)(*p.emission),
                                                   (C.c_float * 4)(*p.diffuse), (C.c_float * 4)(*p.specular),
                                                   p.shininess))
            cm.ggx_offset, cm.ggx_count = len(ggxs), len(m.ggx_materials)
            for p in m.ggx_materials:
                ggxs.append(rayca_ggx_material((C.c_float * 4)(*p.diffuse), (C.c_float * 4)(*p.specular), p.roughness))
            cm.texture_offset, cm.texture_count = len(texs), len(m.textures)
            texs.extend(rayca_texture(t.image) for t in m.textures)
            cm.image_offset, cm.image_count = len(imgs), len(m.images)
            for im in m.images:
                data = np.ascontiguousarray(im.data)
                sd._keep.append(data)
                imgs.append(rayca_image(im.width, im.height, im.color_type, 0, data.ctypes.data))
            cm.light_offset, cm.light_count = len(lights), len(m.lights)
            for l in m.lights:
                lights.append(rayca_light(l.kind, (C.c_float * 4)(*l.color), l.intensity,
                                          (C.c_float * 3)(*l.attenuation), (C.c_float * 3)(*l.ab),
                                          (C.c_float * 3)(*l.ac), l.material))
            cm.camera_offset, cm.camera_count = len(cams), len(m.cameras)
            cams.extend(rayca_camera(c.yfov_radians) for c in m.cameras)
            models.append(cm)

        def arr(ctype, items):
            a = (ctype * max(len(items), 1))(*items)
            sd._keep.append(a)
            return a

        d.scene_nodes, d.scene_node_count = arr(rayca_node, scene_nodes), len(scene_nodes)
        d.models, d.model_count = arr(rayca_model, models), len(models)
        d.nodes, d.node_count = arr(rayca_node, nodes), len(nodes)
        d.children, d.children_count = arr(C.c_uint32, children), len(children)
        d.meshes, d.mesh_count = arr(rayca_mesh, meshes), len(meshes)
        d.mesh_primitives, d.mesh_primitive_count = arr(C.c_uint32, mesh_prims), len(mesh_prims)
        d.primitives, d.primitive_count = arr(rayca_primitive, prims), len(prims)
        d.geometries, d.geometry_count = arr(rayca_geometry, geoms), len(geoms)
        verts = np.concatenate(vert_chunks) if vert_chunks else np.zeros((1, VERTEX_FLOATS), dtype=f32)
        idx = np.concatenate(index_chunks) if index_chunks else np.zeros((1,), dtype=np.uint32)
        verts = np.ascontiguousarray(verts, dtype=f32)
        idx = np.ascontiguousarray(idx, dtype=np.uint32)
        sd._keep += [verts, idx]
        d.vertices, d.vertex_count = verts.ctypes.data, nvert
        d.indices, d.index_count = idx.ctypes.data, nindex
        d.materials, d.material_count = arr(rayca_material, mats), len(mats)
        d.pbr_materials, d.pbr_count = arr(rayca_pbr_material, pbrs), len(pbrs)
        d.phong_materials, d.phong_count = arr(rayca_phong_material, phongs), len(phongs)
        d.ggx_materials, d.ggx_count = arr(rayca_ggx_material, ggxs), len(ggxs)
        d.textures, d.texture_count = arr(rayca_texture, texs), len(texs)
        d.images, d.image_count = arr(rayca_image, imgs), len(imgs)
        d.lights, d.light_count = arr(rayca_light, lights), len(lights)
        d.cameras, d.camera_count = arr(rayca_camera, cams), len(cams)
        sd.triangle_count = nindex // 3
        return sd


def make_vertices(pos, normal=None, uv=None, color=None, tangent=None, bitangent=None) -> np.ndarray:
    """Vertex::default (rayca-geometry/src/vertex.rs:165-176): colour white, normal +Z, rest zero."""
    pos = np.asarray(pos, dtype=f32).reshape(-1, 3)
    n = len(pos)
    v = np.zeros((n, VERTEX_FLOATS), dtype=f32)
    v[:, 0:3] = pos
    v[:, 3:7] = 1.0 if color is None else np.asarray(color, dtype=f32).reshape(n, 4)
    v[:, 7:10] = (0.0, 0.0, 1.0) if normal is None else np.asarray(normal, dtype=f32).reshape(n, 3)
    if tangent is not None:
        v[:, 10:13] = np.asarray(tangent, dtype=f32).reshape(n, 3)
    if bitangent is not None:
        v[:, 13:16] = np.asarray(bitangent, dtype=f32).reshape(n, 3)
    if uv is not None:
        v[:, 16:18] = np.asarray(uv, dtype=f32).reshape(n, 2)
    return v


def triangle_soup_model(positions, material: Optional[PbrMaterial] = None, normals=None) -> Model:
    """A model with one mesh node holding de-indexed triangles; positions (n,3,3)."""
    positions = np.asarray(positions, dtype=f32).reshape(-1, 3)
    if normals is None:
        p = positions.reshape(-1, 3, 3)
        ab = p[:, 1] - p[:, 0]
        ac = p[:, 2] - p[:, 0]
        n = _face_normals(ab, ac)
        normals = np.repeat(n, 3, axis=0)
    m = Model()
    m.geometries.append(TriangleMesh(make_vertices(positions, normal=normals),
                                     np.arange(len(positions), dtype=np.uint32)))
    mat = NONE
    if material is not None:
        m.pbr_materials.append(material)
        m.materials.append(Material(MATERIAL_PBR, 0))
        mat = 0
    m.primitives.append(Primitive(0, mat))
    m.meshes.append(Mesh([0]))
    m.nodes.append(Node(mesh=0))
    m.root.children.append(0)
    return m


def _face_normals(ab, ac):
    """n = ab.cross(ac).get_normalized() in float32, vectorised (sdtf.rs:289-291)."""
    ab = ab.astype(f32)
    ac = ac.astype(f32)
    n = np.stack([
        ab[:, 1] * ac[:, 2] - ab[:, 2] * ac[:, 1],
        ab[:, 2] * ac[:, 0] - ab[:, 0] * ac[:, 2],
        ab[:, 0] * ac[:, 1] - ab[:, 1] * ac[:, 0],
    ], axis=1).astype(f32)
    l2 = ((n[:, 0] * n[:, 0] + n[:, 1] * n[:, 1]) + n[:, 2] * n[:, 2]) + f32(0.0)
    l = np.sqrt(l2).astype(f32)
    ok = l > EPS
    out = n.copy()
    out[ok] = (n[ok] / l[ok, None]).astype(f32)
    return out


# ------------------------------------------------------------------------------ glTF

_COMP = {5120: np.int8, 5121: np.uint8, 5122: np.int16, 5123: np.uint16, 5125: np.uint32, 5126: np.float32}
_NCOMP = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4, "MAT4": 16}


def _decode_image(data: bytes) -> Image:
    """Image::load_data (image.rs:143-158).  Decoding is done by PIL here (the reference uses the
    `image` crate); oracle and CUDA path both receive exactly these bytes."""
    import io

    from PIL import Image as PILImage

    im = PILImage.open(io.BytesIO(data))
    if im.mode == "P":
        im = im.convert("RGBA" if "transparency" in im.info else "RGB")
    if im.mode == "RGBA":
        a = np.asarray(im, dtype=np.uint8)
        return Image(im.width, im.height, COLOR_RGBA8, np.ascontiguousarray(a))
    if im.mode != "RGB":
        raise ValueError(f"Unsupported image color type: {im.mode}")  # from_image_color_type panics
    a = np.asarray(im, dtype=np.uint8)
    return Image(im.width, im.height, COLOR_RGB8, np.ascontiguousarray(a))


def _decompose(mat16):
    """`gltf` 1.4.1 Transform::decomposed for matrix nodes — a third-party dependency that is not
    vendored in the reference checkout; the published algorithm (column scale, rotation matrix ->
    quaternion) is restated.  PARITY UNPINNED: affects box, duck, orientation only."""
    m = np.array(mat16, dtype=f32).reshape(4, 4)  # m[col][row]
    translation = m[3][:3].copy()
    cols = [m[0][:3].copy(), m[1][:3].copy(), m[2][:3].copy()]

    def mag(c):
        return np.sqrt(f32(f32(c[0] * c[0] + c[1] * c[1]) + c[2] * c[2]))

    det = (cols[0][0] * (cols[1][1] * cols[2][2] - cols[2][1] * cols[1][2])
           - cols[1][0] * (cols[0][1] * cols[2][2] - cols[2][1] * cols[0][2])
           + cols[2][0] * (cols[0][1] * cols[1][2] - cols[1][1] * cols[0][2]))
    sx, sy = mag(cols[0]), mag(cols[1])
    sz = f32(np.sign(det) if det != 0 else 1.0) * mag(cols[2])
    cx, cy, cz = cols[0] * (f32(1.0) / sx), cols[1] * (f32(1.0) / sy), cols[2] * (f32(1.0) / sz)
    trace = cx[0] + cy[1] + cz[2]
    if trace >= 0:
        s = np.sqrt(f32(1.0) + trace)
        w = f32(0.5) * s
        s = f32(0.5) / s
        q = [(cy[2] - cz[1]) * s, (cz[0] - cx[2]) * s, (cx[1] - cy[0]) * s, w]
    elif cx[0] > cy[1] and cx[0] > cz[2]:
        s = np.sqrt((cx[0] - cy[1] - cz[2]) + f32(1.0))
        x = f32(0.5) * s
        s = f32(0.5) / s
        q = [x, (cy[0] + cx[1]) * s, (cx[2] + cz[0]) * s, (cy[2] - cz[1]) * s]
    elif cy[1] > cz[2]:
        s = np.sqrt((cy[1] - cx[0] - cz[2]) + f32(1.0))
        y = f32(0.5) * s
        s = f32(0.5) / s
        q = [(cy[0] + cx[1]) * s, y, (cz[1] + cy[2]) * s, (cz[0] - cx[2]) * s]
    else:
        s = np.sqrt((cz[2] - cx[0] - cy[1]) + f32(1.0))
        z = f32(0.5) * s
        s = f32(0.5) / s
        q = [(cx[2] + cz[0]) * s, (cz[1] + cy[2]) * s, z, (cx[1] - cy[0]) * s]
    return translation.astype(f32), np.array(q, dtype=f32), np.array([sx, sy, sz], dtype=f32)


def _read_glb(data: bytes):
    """Binary glTF container (glTF 2.0 spec 4.4): 12-byte header `glTF`, version 2, total length; then chunks
    {u32 length, u32 type, payload} — the first is JSON (0x4E4F534A), the optional second BIN (0x004E4942).
    Returns (json dict, bin chunk or None).  The reference leaves this case `unimplemented!()`
    (rayca-model/src/loader/gltf.rs:239-241, buffer `Source::Bin`); SURVEY 8(f) rank 4 asks for it."""
    import struct

    if len(data) < 20 or data[:4] != b"glTF":
        raise ValueError("not a GLB container")
    version, total = struct.unpack_from("<II", data, 4)
    if version != 2 or total > len(data):
        raise ValueError(f"unsupported GLB (version {version}, length {total} of {len(data)})")
    off, doc, blob = 12, None, None
    while off + 8 <= total:
        n, kind = struct.unpack_from("<II", data, off)
        payload = data[off + 8: off + 8 + n]
        if len(payload) != n:
            raise ValueError("truncated GLB chunk")
        if kind == 0x4E4F534A and doc is None:
            doc = json.loads(payload.decode("utf-8"))
        elif kind == 0x004E4942 and blob is None:
            blob = payload
        off += 8 + n  # chunk lengths are multiples of 4 (spec 4.4.3)
    if doc is None:
        raise ValueError("GLB without a JSON chunk")
    return doc, blob


def load_gltf_path(path, image_hook=None) -> Model:
    """Model::load_gltf_path (rayca-model/src/loader/gltf.rs:256-299).  Also reads `.glb` containers, buffers without
    a `uri` (the GLB BIN chunk) and images stored in a bufferView — the `Source::Bin` / `Source::View` cases the
    reference leaves unimplemented (gltf.rs:239-241,318-324).
    `image_hook(uri_or_None, load)` (fixtures only) may replace an image: `load()` decodes the original."""
    path = os.fspath(path)
    parent = os.path.dirname(path)
    with open(path, "rb") as f:
        raw = f.read()
    glb_bin = None
    if raw[:4] == b"glTF":
        g, glb_bin = _read_glb(raw)
    else:
        g = json.loads(raw.decode("utf-8"))

    def load_uri(uri, prefix):
        if uri.startswith("data:"):
            head, _, payload = uri.partition(",")
            if not head.endswith(";base64"):
                raise ValueError("only base64 data URIs are supported")
            return base64.b64decode(payload)
        with open(os.path.join(parent, uri), "rb") as fh:
            return fh.read()

    def load_buffer(b):
        if "uri" in b:
            return load_uri(b["uri"], "data:application/octet-stream;base64,")
        if glb_bin is None:
            raise ValueError("buffer without a uri outside a GLB container")
        return glb_bin  # Source::Bin

    buffers = [load_buffer(b) for b in g.get("buffers", [])]

    def view_bytes(index):
        view = g["bufferViews"][index]
        o = view.get("byteOffset", 0)
        return bytes(buffers[view["buffer"]][o:o + view["byteLength"]])

    m = Model()
    # load_images (gltf.rs:306-350)
    for gi in g.get("images", []):
        def load(gi=gi):
            data = load_uri(gi["uri"], "data:image/png;base64,") if "uri" in gi else view_bytes(gi["bufferView"])
            return _decode_image(data)
        m.images.append(image_hook(gi.get("uri"), load) if image_hook else load())
    # load_textures (gltf.rs:352-362)
    for gt in g.get("textures", []):
        m.textures.append(Texture(gt["source"]))
    # load_materials (gltf.rs:364-406): always PBR
    for gm in g.get("materials", []):
        pbr = gm.get("pbrMetallicRoughness", {})
        p = PbrMaterial(color=tuple(float(x) for x in pbr.get("baseColorFactor", [1.0, 1.0, 1.0, 1.0])),
                        metallic_factor=float(pbr.get("metallicFactor", 1.0)),
                        roughness_factor=float(pbr.get("roughnessFactor", 1.0)))
        if "baseColorTexture" in pbr:
            p.albedo = pbr["baseColorTexture"]["index"]
        if "normalTexture" in gm:
            p.normal = gm["normalTexture"]["index"]
        if "metallicRoughnessTexture" in pbr:
            p.metallic_roughness = pbr["metallicRoughnessTexture"]["index"]
        m.pbr_materials.append(p)
        m.materials.append(Material(MATERIAL_PBR, len(m.pbr_materials) - 1))

    def read_accessor(index):
        acc = g["accessors"][index]
        view = g["bufferViews"][acc["bufferView"]]
        comp = np.dtype(_COMP[acc["componentType"]])
        ncomp = _NCOMP[acc["type"]]
        offset = view.get("byteOffset", 0) + acc.get("byteOffset", 0)
        stride = view.get("byteStride", 0) or comp.itemsize * ncomp
        count = acc["count"]
        buf = buffers[view["buffer"]]
        out = np.ndarray((count, ncomp), dtype=comp, buffer=buf, offset=offset, strides=(stride, comp.itemsize))
        return np.array(out)

    # load_meshes / load_primitive / load_vertices (gltf.rs:408-502)
    for gmesh in g.get("meshes", []):
        handles = []
        for gp in gmesh["primitives"]:
            assert gp.get("mode", 4) == 4, "only triangle lists"
            attrs = gp["attributes"]
            n = g["accessors"][attrs["POSITION"]]["count"]
            normal = read_accessor(attrs["NORMAL"]).astype(f32) if "NORMAL" in attrs else None
            pos = read_accessor(attrs["POSITION"]).astype(f32)
            uv = color = tangent = bitangent = None
            # BTreeMap order of gltf::Semantic: later TEXCOORD_n / COLOR_n sets overwrite earlier ones
            for key in sorted(k for k in attrs if k.startswith("TEXCOORD_")):
                uv = read_accessor(attrs[key]).astype(f32)
            for key in sorted(k for k in attrs if k.startswith("COLOR_")):
                c = read_accessor(attrs[key]).astype(f32)
                color = np.concatenate([c, np.ones((len(c), 1), dtype=f32)], axis=1) if c.shape[1] == 3 else c
            if "TANGENT" in attrs:
                t4 = read_accessor(attrs["TANGENT"]).astype(f32)
                tangent = t4[:, :3]
                nn = normal if normal is not None else np.tile(np.array([0, 0, 1], dtype=f32), (n, 1))
                # bitangent = normal.cross(tangent) * w (gltf.rs:230), Vec3::cross op order
                cr = np.stack([
                    nn[:, 1] * tangent[:, 2] - nn[:, 2] * tangent[:, 1],
                    nn[:, 2] * tangent[:, 0] - nn[:, 0] * tangent[:, 2],
                    nn[:, 0] * tangent[:, 1] - nn[:, 1] * tangent[:, 0],
                ], axis=1).astype(f32)
                bitangent = (cr * t4[:, 3:4]).astype(f32)
            verts = make_vertices(pos, normal=normal, uv=uv, color=color, tangent=tangent, bitangent=bitangent)
            if "indices" in gp:
                idx = read_accessor(gp["indices"]).reshape(-1).astype(np.uint32)
            else:
                idx = np.zeros((0,), dtype=np.uint32)  # load_indices leaves an empty index list
            m.geometries.append(TriangleMesh(verts, idx))
            m.primitives.append(Primitive(len(m.geometries) - 1, gp.get("material", NONE)))
            handles.append(len(m.primitives) - 1)
        m.meshes.append(Mesh(handles))

    # load_cameras (gltf.rs:504-527)
    for gc in g.get("cameras", []):
        if gc["type"] == "perspective":
            m.cameras.append(Camera(float(f32(gc["perspective"]["yfov"]))))
        else:
            m.cameras.append(Camera(1.0))  # Camera::orthographic sets yfov_radians = 1.0

    # load_nodes (gltf.rs:529-578)
    scene = g["scenes"][0]
    m.root.children = list(scene.get("nodes", []))
    for gn in g.get("nodes", []):
        if "matrix" in gn:
            t, r, s = _decompose(gn["matrix"])
        else:
            t = np.array(gn.get("translation", [0, 0, 0]), dtype=f32)
            r = np.array(gn.get("rotation", [0, 0, 0, 1]), dtype=f32)
            s = np.array(gn.get("scale", [1, 1, 1]), dtype=f32)
        m.nodes.append(Node(trs=Trs(t, r, s), children=list(gn.get("children", [])), mesh=gn.get("mesh", NONE),
                            camera=gn.get("camera", NONE), name=gn.get("name", "Unknown")))
    return m


# ------------------------------------------------------------------------------ SDTF


@dataclass
class SdtfConfig:
    """rayca-model/src/loader/sdtf.rs:87-120"""
    width: int = 0
    height: int = 0
    max_depth: int = 5
    light_samples: int = 1
    light_stratify: bool = False
    samples_per_pixel: int = 1
    direct_sampler: str = "none"
    russian_roulette: bool = False
    indirect_sampler: str = "hemisphere"
    integrator: str = "raytracer"
    brdf: str = "phong"
    gamma: float = 1.0


def load_sdtf_path(path):
    """Model::load_sdtf_path (sdtf.rs:903-907) for the command subset the repository's fixtures use:
    size, camera, maxverts, vertex, tri, ambient/emission/diffuse/specular/shininess/roughness/brdf,
    push/popTransform, translate/rotate/scale, point, directional, attenuation, maxdepth, integrator,
    lightsamples, lightstratify, spp, nexteventestimation, russianroulette, importancesampling, gamma.
    `sphere` and `quadLight` are parsed into the model."""
    cfg = SdtfConfig()
    model = Model()
    vertices: List[np.ndarray] = []
    transform: List[Trs] = []
    phong = PhongMaterial()
    ggx = GgxMaterial()
    attenuation = (1.0, 0.0, 0.0)
    pending = {"tris": None, "sphere": None}

    def process_primitive():
        """sdtf.rs:833-857"""
        if pending["tris"] is None and pending["sphere"] is None:
            return
        if pending["tris"] is not None:
            tri_idx = np.array(pending["tris"], dtype=np.int64).reshape(-1, 3)
            vp = np.array(vertices, dtype=f32).reshape(-1, 3)
            p = vp[tri_idx]  # (n,3,3)
            ab = p[:, 1] - p[:, 0]
            ac = p[:, 2] - p[:, 0]
            n = _face_normals(ab, ac)
            verts = make_vertices(p.reshape(-1, 3), normal=np.repeat(n, 3, axis=0))
            geom = TriangleMesh(verts, np.arange(len(verts), dtype=np.uint32))
        else:
            geom = pending["sphere"]
        pending["tris"] = pending["sphere"] = None
        model.geometries.append(geom)
        # process_material (sdtf.rs:816-830)
        if cfg.brdf == "phong":
            model.phong_materials.append(PhongMaterial(phong.ambient, phong.emission, phong.diffuse, phong.specular,
                                                       phong.shininess))
            model.materials.append(Material(MATERIAL_PHONG, len(model.phong_materials) - 1))
        else:
            model.ggx_materials.append(GgxMaterial(ggx.diffuse, ggx.specular, ggx.roughness))
            model.materials.append(Material(MATERIAL_GGX, len(model.ggx_materials) - 1))
        model.primitives.append(Primitive(len(model.geometries) - 1, len(model.materials) - 1))
        model.meshes.append(Mesh([len(model.primitives) - 1]))
        trs = Trs()
        for t in transform:
            trs.left_mul(t)
        model.nodes.append(Node(trs=trs, mesh=len(model.meshes) - 1))
        model.root.children.append(len(model.nodes) - 1)

    def color3(w):
        return (float(f32(w[0])), float(f32(w[1])), float(f32(w[2])), 1.0)

    with open(path, "r") as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line.startswith("#") or not any(ch.isalnum() for ch in line):
                continue
            words = [w for w in line.split(" ") if w]
            cmd, w = words[0], words[1:]
            if cmd == "size":
                cfg.width, cfg.height = int(w[0]), int(w[1])
            elif cmd == "camera":
                # parse_camera (sdtf.rs:172-234) with Mat4::look_at (mat4.rs:81-95)
                eye, target, up = _v(*map(f32, w[0:3])), _v(*map(f32, w[3:6])), _v(*map(f32, w[6:9]))
                yfov = f32(w[9]) * f32(np.pi) / f32(180.0)
                z_axis = _normalized(eye - target)
                x_axis = _normalized(_cross(up, z_axis))
                y_axis = _cross(z_axis, x_axis)
                rot = quat_conjugate(quat_from_rows([x_axis, y_axis, z_axis]))
                model.cameras.append(Camera(float(yfov)))
                model.nodes.append(Node(camera=len(model.cameras) - 1, trs=Trs(translation=eye, rotation=rot)))
                model.root.children.append(len(model.nodes) - 1)
            elif cmd == "maxverts":
                pass
            elif cmd == "vertex":
                vertices.append(np.array([f32(w[0]), f32(w[1]), f32(w[2])], dtype=f32))
            elif cmd == "tri":
                if pending["tris"] is None and pending["sphere"] is None:
                    pending["tris"] = []
                if pending["tris"] is not None:
                    pending["tris"].append((int(w[0]), int(w[1]), int(w[2])))
            elif cmd == "sphere":
                process_primitive()
                pending["sphere"] = Sphere(_v(f32(w[0]), f32(w[1]), f32(w[2])), float(f32(w[3])))
            elif cmd == "ambient":
                process_primitive()
                phong.ambient = color3(w)
            elif cmd == "emission":
                process_primitive()
                phong.emission = color3(w)
            elif cmd == "diffuse":
                process_primitive()
                phong.diffuse = ggx.diffuse = color3(w)
            elif cmd == "specular":
                process_primitive()
                phong.specular = ggx.specular = color3(w)
            elif cmd == "shininess":
                process_primitive()
                phong.shininess = float(f32(w[0]))
            elif cmd == "roughness":
                process_primitive()
                ggx.roughness = float(f32(w[0]))
            elif cmd == "brdf":
                process_primitive()
                cfg.brdf = w[0]
            elif cmd == "translate":
                transform[-1].left_mul(Trs(translation=_v(f32(w[0]), f32(w[1]), f32(w[2]))))
            elif cmd == "rotate":
                ang = f32(np.deg2rad(f32(w[3])))
                transform[-1].left_mul(Trs(rotation=quat_axis_angle(_v(f32(w[0]), f32(w[1]), f32(w[2])), ang)))
            elif cmd == "scale":
                transform[-1].left_mul(Trs(scale=_v(f32(w[0]), f32(w[1]), f32(w[2]))))
            elif cmd == "pushTransform":
                transform.append(Trs())
            elif cmd == "popTransform":
                process_primitive()
                if transform:
                    transform.pop()
            elif cmd == "point":
                model.lights.append(Light(kind=LIGHT_POINT, color=color3(w[3:6]), intensity=1.0,
                                          attenuation=attenuation))
                model.nodes.append(Node(light=len(model.lights) - 1,
                                        trs=Trs(translation=_v(f32(w[0]), f32(w[1]), f32(w[2])))))
                model.root.children.append(len(model.nodes) - 1)
            elif cmd == "directional":
                # Quat::angle_between(X_AXIS, -dir) (quat.rs:118-126)
                a = _v(1, 0, 0)
                b = -_v(f32(w[0]), f32(w[1]), f32(w[2]))
                c = _cross(a, b)
                ww = np.sqrt(_dot(a, a) * _dot(b, b)) + _dot(a, b)
                q = np.array([c[0], c[1], c[2], ww], dtype=f32)
                q = (q / np.sqrt(f32(f32(f32(q[0] * q[0]) + f32(q[1] * q[1])) + f32(q[2] * q[2])) + f32(q[3] * q[3]))).astype(f32)
                model.lights.append(Light(kind=LIGHT_DIRECTIONAL, color=color3(w[3:6]), intensity=1.0))
                model.nodes.append(Node(light=len(model.lights) - 1, trs=Trs(rotation=q)))
                model.root.children.append(len(model.nodes) - 1)
            elif cmd == "attenuation":
                attenuation = (float(f32(w[0])), float(f32(w[1])), float(f32(w[2])))
            elif cmd == "maxdepth":
                cfg.max_depth = int(w[0])
            elif cmd == "integrator":
                cfg.integrator = w[0]
            elif cmd == "quadLight":
                model.phong_materials.append(PhongMaterial(emission=color3(w[9:12])))
                model.materials.append(Material(MATERIAL_PHONG, len(model.phong_materials) - 1))
                model.lights.append(Light(kind=LIGHT_QUAD, color=color3(w[9:12]), intensity=1.0,
                                          ab=tuple(float(f32(x)) for x in w[3:6]),
                                          ac=tuple(float(f32(x)) for x in w[6:9]), material=len(model.materials) - 1))
                model.nodes.append(Node(light=len(model.lights) - 1,
                                        trs=Trs(translation=_v(f32(w[0]), f32(w[1]), f32(w[2])))))
                model.root.children.append(len(model.nodes) - 1)
            elif cmd == "lightsamples":
                cfg.light_samples = int(w[0])
            elif cmd == "lightstratify":
                cfg.light_stratify = w[0] == "on"
            elif cmd == "spp":
                cfg.samples_per_pixel = int(w[0])
            elif cmd == "nexteventestimation":
                cfg.direct_sampler = {"on": "nee", "mis": "mis"}[w[0]]
            elif cmd == "russianroulette":
                cfg.russian_roulette = w[0] == "on"
            elif cmd == "importancesampling":
                cfg.indirect_sampler = w[0]
            elif cmd == "gamma":
                cfg.gamma = float(f32(w[0]))
    process_primitive()
    return model, cfg
