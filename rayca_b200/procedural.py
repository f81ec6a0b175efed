"""Deterministic procedural scenes for bench.py and the full-size property tests.

The reference's headline assets are not available to this repository: `sponza.bin` is missing from
the reference checkout (`.MISSING_LARGE_BLOBS`) and DamagedHelmet is not in it at all, and nothing
under /root/reference exists on the GPU box.  These generators build stand-ins of the same class:

  atrium(...)  — an architectural interior of ~262k triangles in one mesh node / one BLAS (Sponza
                 ships 262,267 triangles in one mesh): floor, walls, vaulted ceiling, two
                 storeys of colonnades with arches, hanging drapes; textured PBR materials with
                 albedo / normal / metallic-roughness maps; the reference test's root rotation and
                 camera placement (rayca-soft/tests/gltf.rs:319-341) and the default model's lights.
  soup(n, ...) — BASELINE.md workload M4: n random triangles, numpy.random.default_rng(12345).

Everything is float32 and seeded, so every run (and every rank) builds bit-identical input.
"""
from __future__ import annotations

import numpy as np

from . import model as M

f32 = np.float32


def _patch(fn, nu, nv, flip=False):
    """Tessellates a parametric surface fn(u, v) -> (x, y, z) over [0,1]^2 into 2*nu*nv triangles.
    Returns (positions (n,3), normals, uvs, tangents, bitangents, indices)."""
    u = np.linspace(0.0, 1.0, nu + 1, dtype=np.float64)
    v = np.linspace(0.0, 1.0, nv + 1, dtype=np.float64)
    uu, vv = np.meshgrid(u, v, indexing="ij")
    p = np.stack(fn(uu, vv), axis=-1)
    e = 1e-4
    du = (np.stack(fn(uu + e, vv), axis=-1) - np.stack(fn(uu - e, vv), axis=-1)) / (2 * e)
    dv = (np.stack(fn(uu, vv + e), axis=-1) - np.stack(fn(uu, vv - e), axis=-1)) / (2 * e)
    n = np.cross(du, dv)
    if flip:
        n = -n
    n /= np.maximum(np.linalg.norm(n, axis=-1, keepdims=True), 1e-12)
    t = du / np.maximum(np.linalg.norm(du, axis=-1, keepdims=True), 1e-12)
    b = np.cross(n, t)
    idx = np.arange((nu + 1) * (nv + 1)).reshape(nu + 1, nv + 1)
    a, bq, c, d = idx[:-1, :-1], idx[1:, :-1], idx[1:, 1:], idx[:-1, 1:]
    if flip:
        tri = np.stack([a, c, bq, a, d, c], axis=-1)
    else:
        tri = np.stack([a, bq, c, a, c, d], axis=-1)
    uv = np.stack([uu, vv], axis=-1)
    r = lambda x, k: x.reshape(-1, k).astype(f32)  # noqa: E731
    return r(p, 3), r(n, 3), r(uv, 2), r(t, 3), r(b, 3), tri.reshape(-1).astype(np.uint32)


class _MeshBuilder:
    def __init__(self):
        self.parts = {}  # material -> list of (verts(n,18), idx)

    def add(self, material, patch, uv_scale=(1.0, 1.0)):
        p, n, uv, t, b, idx = patch
        verts = M.make_vertices(p, normal=n, uv=uv * np.array(uv_scale, dtype=f32), tangent=t, bitangent=b)
        self.parts.setdefault(material, []).append((verts, idx))

    def triangles(self):
        return sum(len(i) // 3 for parts in self.parts.values() for _, i in parts)

    def geometries(self):
        out = []
        for mat, parts in sorted(self.parts.items()):
            off, vs, is_ = 0, [], []
            for v, i in parts:
                vs.append(v)
                is_.append(i + off)
                off += len(v)
            out.append((mat, M.TriangleMesh(np.concatenate(vs), np.concatenate(is_).astype(np.uint32))))
        return out


def _textures(rng, size=512):
    """A small set of procedural RGBA8 / RGB8 maps: albedo (stone, brick, fabric), normal, metal-rough."""
    y, x = np.mgrid[0:size, 0:size].astype(np.float64) / size
    noise = rng.random((size, size))
    smooth = (noise + np.roll(noise, 1, 0) + np.roll(noise, 1, 1) + np.roll(noise, (1, 1), (0, 1))) / 4

    def rgba(r, g, b):
        a = np.stack([r, g, b, np.ones_like(r)], axis=-1)
        return np.ascontiguousarray((np.clip(a, 0, 1) * 255).astype(np.uint8))

    brick = ((np.floor(y * 16) % 2) * 0.5 + x * 8) % 1.0
    mortar = ((brick < 0.06) | ((y * 16) % 1.0 < 0.1)).astype(np.float64)
    stone = 0.55 + 0.25 * smooth
    imgs = [
        rgba(stone, stone * 0.95, stone * 0.85),                                          # 0 stone albedo
        rgba(0.55 - 0.3 * mortar + 0.1 * smooth, 0.25 + 0.2 * mortar, 0.2 + 0.2 * mortar),  # 1 brick albedo
        rgba(0.6 + 0.3 * np.sin(x * 80) ** 2, 0.1 + 0.1 * smooth, 0.12 + 0.1 * smooth),    # 2 red fabric
        rgba(0.1 + 0.1 * smooth, 0.25 + 0.1 * smooth, 0.5 + 0.3 * np.sin(y * 60) ** 2),    # 3 blue fabric
    ]
    gx = np.roll(smooth, -1, 1) - np.roll(smooth, 1, 1)
    gy = np.roll(smooth, -1, 0) - np.roll(smooth, 1, 0)
    nz = np.ones_like(gx)
    nn = np.stack([-gx * 4, -gy * 4, nz], axis=-1)
    nn /= np.linalg.norm(nn, axis=-1, keepdims=True)
    normal = np.ascontiguousarray(((nn * 0.5 + 0.5) * 255).astype(np.uint8))               # 4 normal map (RGB8)
    mr = rgba(0.55 + 0.4 * smooth, np.zeros_like(smooth), 0.05 + 0.1 * mortar)             # 5 R=roughness, B=metallic
    out = [M.Image(size, size, M.COLOR_RGBA8, i) for i in imgs]
    out.append(M.Image(size, size, M.COLOR_RGB8, normal))
    out.append(M.Image(size, size, M.COLOR_RGBA8, mr))
    return out


def atrium(target_triangles: int = 262_267, seed: int = 2025, textured: bool = True) -> M.Scene:
    """Sponza-class stand-in.  One model, one mesh node, one BLAS; root rotation and camera as in the
    reference's `sponza` test (rayca-soft/tests/gltf.rs:319-341)."""
    rng = np.random.default_rng(seed)
    L, W, H = 28.0, 18.0, 12.0  # long axis X before the root rotation, like Sponza
    detail = 1.0
    mb = None
    for _ in range(12):  # scale the tessellation until the triangle count is just under the target
        mb = _MeshBuilder()
        k = lambda n: max(2, int(round(n * detail)))  # noqa: E731
        hx, hz = L / 2, W / 2
        mb.add(0, _patch(lambda u, v: ((u - 0.5) * L, 0 * u, (0.5 - v) * W), k(96), k(64)), (14, 9))       # floor
        mb.add(1, _patch(lambda u, v: ((u - 0.5) * L, v * H, 0 * u - hz), k(96), k(48)), (10, 4))         # wall -z
        mb.add(1, _patch(lambda u, v: ((0.5 - u) * L, v * H, 0 * u + hz), k(96), k(48)), (10, 4))         # wall +z
        mb.add(1, _patch(lambda u, v: (0 * u - hx, v * H, (0.5 - u) * W), k(64), k(48)), (6, 4))          # end wall
        mb.add(1, _patch(lambda u, v: (0 * u + hx, v * H, (u - 0.5) * W), k(64), k(48)), (6, 4))          # end wall
        # barrel vault ceiling
        mb.add(0, _patch(lambda u, v: ((u - 0.5) * L, H + 2.0 * np.sin(np.pi * v), (v - 0.5) * W), k(96), k(64), flip=False), (14, 9))
        # two storeys of colonnades along both sides of the nave, with arches between the columns
        ncol = 9
        for side in (-1.0, 1.0):
            zc = side * 3.2
            for storey, (y0, hcol, rad) in enumerate(((0.0, 4.2, 0.32), (5.0, 3.4, 0.26))):
                for i in range(ncol):
                    xc = -hx + 2.0 + i * (L - 4.0) / (ncol - 1)
                    mb.add(0, _patch(lambda u, v, xc=xc, zc=zc, y0=y0, hcol=hcol, rad=rad:
                                     (xc + rad * (1 + 0.08 * np.cos(16 * 2 * np.pi * u)) * np.cos(2 * np.pi * u),
                                      y0 + v * hcol,
                                      zc - rad * (1 + 0.08 * np.cos(16 * 2 * np.pi * u)) * np.sin(2 * np.pi * u)),
                                     k(40), k(24)), (2, 3))
                    # capital (a squashed torus-like ring)
                    mb.add(0, _patch(lambda u, v, xc=xc, zc=zc, y0=y0, hcol=hcol, rad=rad:
                                     (xc + (rad * 1.5 + 0.12 * np.cos(2 * np.pi * v)) * np.cos(2 * np.pi * u),
                                      y0 + hcol + 0.12 * np.sin(2 * np.pi * v),
                                      zc - (rad * 1.5 + 0.12 * np.cos(2 * np.pi * v)) * np.sin(2 * np.pi * u)),
                                     k(32), k(10)))
                    if i + 1 < ncol:  # arch to the next column: a half-tube
                        xn = -hx + 2.0 + (i + 1) * (L - 4.0) / (ncol - 1)
                        mid, span = (xc + xn) / 2, (xn - xc) / 2 - rad
                        mb.add(1, _patch(lambda u, v, mid=mid, span=span, zc=zc, y0=y0, hcol=hcol:
                                         (mid - span * np.cos(np.pi * u), y0 + hcol + 0.2 + span * 0.8 * np.sin(np.pi * u),
                                          zc + (v - 0.5) * 0.7), k(28), k(6)), (3, 1))
                # gallery floor slab between the storeys
                if storey == 0:
                    mb.add(0, _patch(lambda u, v, side=side: ((u - 0.5) * L, 0 * u + 4.9,
                                                              side * (3.2 + v * (hz - 3.2))), k(96), k(24),
                                     flip=(side < 0)), (14, 3))
        # drapes: wavy sheets hanging across the nave
        for j, xd in enumerate(np.linspace(-hx + 4.0, hx - 4.0, 6)):
            mat = 2 + (j % 2)
            mb.add(mat, _patch(lambda u, v, xd=xd, j=j: (xd + 0.25 * np.sin(9 * np.pi * u + j) * (1 - v),
                                                         9.5 - 3.2 * v - 0.4 * np.sin(np.pi * u),
                                                         (u - 0.5) * 5.6), k(72), k(40)), (2, 2))
        n = mb.triangles()
        if abs(n - target_triangles) / target_triangles < 0.004 or detail <= 0:
            break
        detail *= (target_triangles / n) ** 0.5
    model = M.Model()
    if textured:
        model.images = _textures(rng)
        model.textures = [M.Texture(i) for i in range(len(model.images))]
        mats = [M.PbrMaterial((1, 1, 1, 1), 0, 4, 0.0, 0.8, 5), M.PbrMaterial((1, 1, 1, 1), 1, 4, 0.0, 0.9, 5),
                M.PbrMaterial((1, 1, 1, 1), 2, M.NONE, 0.0, 1.0, M.NONE),
                M.PbrMaterial((1, 1, 1, 1), 3, M.NONE, 0.0, 1.0, M.NONE)]
    else:
        mats = [M.PbrMaterial((0.8, 0.8, 0.75, 1), M.NONE, M.NONE, 0.0, 1.0, M.NONE) for _ in range(4)]
    model.pbr_materials = mats
    model.materials = [M.Material(M.MATERIAL_PBR, i) for i in range(len(mats))]
    handles = []
    for mat, geom in mb.geometries():
        model.geometries.append(geom)
        model.primitives.append(M.Primitive(len(model.geometries) - 1, mat))
        handles.append(len(model.primitives) - 1)
    model.meshes.append(M.Mesh(handles))
    model.nodes.append(M.Node(mesh=0))
    model.root.children.append(0)
    scene = M.Scene()
    h = scene.push_model(model)
    scene.nodes[h].trs.rotation = np.array([0.0, -0.707, 0.0, 0.707], dtype=f32)  # gltf.rs:326-327
    dm = M.create_default_model()
    dm.get_mut_last_camera().trs.translation = np.array([0.2, 1.0, 0.0], dtype=f32)  # gltf.rs:330-331
    scene.push_model(dm)
    return scene


def soup(n: int = 10_000_000, seed: int = 12345) -> M.Scene:
    """centroids U[-50,50]^3, edge scale logU[0.05,0.5], random orientation of a unit
    equilateral triangle, one white PBR material (metallic 0, roughness 1), camera (0,0,120) looking -z
    with yfov 45 degrees, two point lights of intensity 1e6 at (-/+30, 80, 90)."""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-50.0, 50.0, (n, 3))
    s = np.exp(rng.uniform(np.log(0.05), np.log(0.5), (n, 1)))
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    base = np.array([[1.0, 0.0, 0.0], [-0.5, np.sqrt(3) / 2, 0.0], [-0.5, -np.sqrt(3) / 2, 0.0]])
    x, y, z, w = q[:, 0:1], q[:, 1:2], q[:, 2:3], q[:, 3:4]
    r0 = np.concatenate([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)], axis=1)
    r1 = np.concatenate([2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)], axis=1)
    r2 = np.concatenate([2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1)
    rot = np.stack([r0, r1, r2], axis=1)  # (n,3,3)
    verts = np.einsum("nij,kj->nki", rot, base) * s[:, None, :] + c[:, None, :]
    model = M.triangle_soup_model(verts.astype(f32), M.PbrMaterial((1, 1, 1, 1), M.NONE, M.NONE, 0.0, 1.0, M.NONE))
    scene = M.Scene()
    scene.push_model(model)
    dm = M.Model()
    dm.cameras.append(M.Camera(float(f32(np.pi / 4))))
    dm.nodes.append(M.Node(camera=0, trs=M.Trs(translation=np.array([0.0, 0.0, 120.0], dtype=f32))))
    dm.root.children.append(0)
    dm.lights.append(M.Light(kind=M.LIGHT_POINT, intensity=1.0e6))
    for k, px in enumerate((-30.0, 30.0)):
        dm.nodes.append(M.Node(light=0, trs=M.Trs(translation=np.array([px, 80.0, 90.0], dtype=f32))))
        dm.root.children.append(1 + k)
    scene.push_model(dm)
    return scene
