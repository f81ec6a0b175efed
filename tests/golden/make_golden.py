"""Regenerates tests/golden/*.npz.  Run in the authoring container only:

    python tests/golden/make_golden.py [name ...]

It reads the reference's own test assets from /root/reference/tests/model (glTF / SDTF files of the
`rayca-models` submodule), builds the same scenes the reference's integration tests build
(rayca-soft/tests/gltf.rs, rayca-soft/tests/sdtf.rs), serialises each scene as the POD descriptor of
include/rayca_cuda.h (`<scene>.scene.npz`) and stores what the CPU oracle renders for each test case
(`<scene>.<case>.npz`: rgba8, rgba_f32, hit_src, hit_prim, hit_blas, depth, uv, ray counts).

/root/reference does not exist on the GPU box; the tests only ever read the committed .npz files.
The reference itself cannot be executed here (Rust nightly, no toolchain), so these vectors pin the
ORACLE's behaviour over time — they are not outputs of the reference (image-level parity unpinned).
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

from oracle import binding as O  # noqa: E402
from rayca_b200 import model as M  # noqa: E402
from rayca_b200.renderer import Config, IntegratorStrategy, SamplerStrategy  # noqa: E402

ASSETS = "/root/reference/tests/model"
f32 = np.float32


def gltf_scene(rel, tweak=None, camera=None):
    s = M.Scene()
    h = s.push_gltf_from_path(os.path.join(ASSETS, rel))
    if tweak:
        tweak(s, h)
    dm = M.create_default_model()
    if camera is not None:
        dm.get_mut_last_camera().trs.translation = np.array(camera, dtype=f32)
    s.push_model(dm)
    return s


def unit_triangle_scene():
    """rayca-soft/tests/gltf.rs:47-83 `triangle`: TriangleMesh::unit with RGB vertex colours under a
    node translated (0,-1,0) and scaled (1,2,1)."""
    m = M.Model()
    col = np.array([[1, 0, 0, 1], [0, 1, 0, 1], [0, 0, 1, 1]], dtype=f32)
    verts = M.make_vertices([[-1, 0, 0], [1, 0, 0], [0, 1, 0]], color=col)
    m.geometries.append(M.TriangleMesh(verts, np.array([0, 1, 2], dtype=np.uint32)))
    m.primitives.append(M.Primitive(0))
    m.meshes.append(M.Mesh([0]))
    m.nodes.append(M.Node(mesh=0, trs=M.Trs(translation=np.array([0, -1, 0], dtype=f32),
                                            scale=np.array([1, 2, 1], dtype=f32))))
    m.root.children.append(0)
    s = M.Scene()
    s.push_model(m)
    s.push_model(M.create_default_model())
    return s


def cubes_over_plane_scene():
    """rayca-soft/tests/gltf.rs:85-183 `cube_over_plane` without its sphere (spheres are a later row):
    four box.gltf instances => four BLAS under one TLAS."""
    s = M.Scene()
    box = os.path.join(ASSETS, "box/box.gltf")
    h = s.push_gltf_from_path(box)  # blue floor
    s.nodes[h].trs.scale = np.array([16.0, 0.125, 16.0], dtype=f32)
    s.nodes[h].trs.translation = s.nodes[h].trs.translation + np.array([0, -1, 0], dtype=f32)
    floor = s.models[s.nodes[h].model]
    floor.nodes[1].trs.rotation = np.array([0, 0, 0, 1], dtype=f32)
    floor.pbr_materials[0].color = (0.1, 0.2, 0.7, 1.0)
    floor.pbr_materials[0].metallic_factor = 1.0
    for shift in ([1.0, 1.0, -2.0], [0.0, 0.0, -1.0], [-1.5, 0.0, -4.0]):
        h = s.push_gltf_from_path(box)
        s.nodes[h].trs.translation = s.nodes[h].trs.translation + np.array(shift, dtype=f32)
    s.push_model(M.create_default_model())
    return s


def scratcher_sphere_scene():
    """rayca-soft/tests/gltf.rs:9-45 `sphere`: Sphere::unit under a node translated (0,0,-1) and scaled (1,2,1),
    drawn by the Scratcher integrator with Config::bvh(false)."""
    m = M.Model()
    m.geometries.append(M.Sphere(np.array([0, 0, 0], dtype=f32), 1.0))
    m.primitives.append(M.Primitive(0))
    m.meshes.append(M.Mesh([0]))
    m.nodes.append(M.Node(mesh=0, trs=M.Trs(translation=np.array([0, 0, -1], dtype=f32),
                                            scale=np.array([1, 2, 1], dtype=f32))))
    m.root.children.append(0)
    s = M.Scene()
    s.push_model(m)
    s.push_model(M.create_default_model())
    return s


def glass_scene():
    """Transmission branch of the Scratcher (scratcher.rs:34-43, Color::over): two half-transparent triangles
    (vertex alpha 0.35 / 0.6) one behind the other in front of an opaque checker of four triangles."""
    m = M.Model()
    back = M.make_vertices([[-3, -3, 0], [3, -3, 0], [3, 3, 0], [-3, 3, 0], [0, 0, 0]],
                           color=[[0.9, 0.2, 0.2, 1], [0.2, 0.9, 0.2, 1], [0.2, 0.2, 0.9, 1], [0.9, 0.9, 0.2, 1], [1, 1, 1, 1]])
    m.geometries.append(M.TriangleMesh(back, np.array([0, 1, 4, 1, 2, 4, 2, 3, 4, 3, 0, 4], dtype=np.uint32)))
    near = M.make_vertices([[-1, -0.8, 0], [1, -0.8, 0], [0, 1, 0]], color=[[0.1, 0.6, 0.9, 0.35]] * 3)
    m.geometries.append(M.TriangleMesh(near, np.array([0, 1, 2], dtype=np.uint32)))
    mid = M.make_vertices([[-1.2, 0.9, 0], [0.2, -1.1, 0], [1.4, 0.7, 0]], color=[[0.9, 0.5, 0.1, 0.6]] * 3)
    m.geometries.append(M.TriangleMesh(mid, np.array([0, 1, 2], dtype=np.uint32)))
    for g in range(3):
        m.primitives.append(M.Primitive(g))
        m.meshes.append(M.Mesh([g]))
    for mesh, z in ((0, -3.0), (1, -1.0), (2, -2.0)):
        m.nodes.append(M.Node(mesh=mesh, trs=M.Trs(translation=np.array([0, 0, z], dtype=f32))))
        m.root.children.append(len(m.nodes) - 1)
    s = M.Scene()
    s.push_model(m)
    s.push_model(M.create_default_model())
    return s


def normalquad_scene():
    """Row a12 (normal mapping): a 2 x 2 quad with tangents, bitangents and three textures — albedo, metallic /
    roughness, and a procedural NORMAL map of overlapping ripples — under a rotated, non-uniformly scaled node, so
    that PbrMaterial::get_normal (rayca-model/src/material/pbr.rs:94-123), Mat3::tbn (rayca-math/src/mat3.rs:57-65)
    and the tangent / bitangent transforms of BvhPrimitive (rayca-soft/src/bvh/primitive.rs:209-226) all take part.
    The uv range (0..1.5) wraps the sampler."""
    n = 64
    yy, xx = np.mgrid[0:n, 0:n].astype(np.float64)
    hx = 0.45 * np.cos(2 * np.pi * xx / 16.0) + 0.25 * np.sin(2 * np.pi * (xx + 2 * yy) / 21.0)
    hy = 0.45 * np.sin(2 * np.pi * yy / 12.0) + 0.25 * np.cos(2 * np.pi * (xx - yy) / 17.0)
    nz = np.sqrt(np.maximum(1.0 - np.minimum(hx * hx + hy * hy, 0.95), 0.05))
    nrm = np.stack([hx, hy, nz], axis=-1)
    normal_tex = np.concatenate([np.clip((nrm * 0.5 + 0.5) * 255.0 + 0.5, 0, 255), np.full((n, n, 1), 255.0)], axis=-1).astype(np.uint8)
    albedo = np.zeros((32, 32, 3), dtype=np.uint8)
    cy, cx = np.mgrid[0:32, 0:32]
    albedo[..., 0] = 120 + 100 * (((cx // 4) + (cy // 4)) % 2)
    albedo[..., 1] = 90 + 5 * cx
    albedo[..., 2] = 200 - 4 * cy
    mr = np.zeros((16, 16, 4), dtype=np.uint8)
    my, mx = np.mgrid[0:16, 0:16]
    mr[..., 0] = 60 + 12 * mx      # roughness = R
    mr[..., 2] = 16 * ((my // 4) % 2) * 8  # metallic = B
    mr[..., 3] = 255
    m = M.Model()
    m.images = [M.Image(32, 32, M.COLOR_RGB8, np.ascontiguousarray(albedo)), M.Image(n, n, M.COLOR_RGBA8, np.ascontiguousarray(normal_tex)),
                M.Image(16, 16, M.COLOR_RGBA8, np.ascontiguousarray(mr))]
    m.textures = [M.Texture(0), M.Texture(1), M.Texture(2)]
    m.pbr_materials.append(M.PbrMaterial(color=(0.9, 0.85, 0.8, 1.0), albedo=0, normal=1, metallic_factor=1.0, roughness_factor=1.0,
                                         metallic_roughness=2))
    m.materials.append(M.Material(M.MATERIAL_PBR, 0))
    pos = [[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]]
    uv = [[0, 1.5], [1.5, 1.5], [1.5, 0], [0, 0]]
    # a slightly tilted tangent frame per vertex (not orthonormal on purpose: the reference does not re-orthogonalise)
    nor = [[0, 0, 1], [0.1, 0, 0.995], [0, 0.1, 0.995], [-0.1, -0.1, 0.99]]
    tan = [[1, 0, 0], [0.995, 0, -0.1], [1, 0.05, 0], [0.99, 0, 0.1]]
    bit = [[0, 1, 0], [0, 1, 0.02], [0, 0.995, -0.1], [0.03, 0.99, 0.1]]
    verts = M.make_vertices(pos, normal=nor, uv=uv, tangent=tan, bitangent=bit)
    m.geometries.append(M.TriangleMesh(verts, np.array([0, 1, 2, 0, 2, 3], dtype=np.uint32)))
    m.primitives.append(M.Primitive(0, 0))
    m.meshes.append(M.Mesh([0]))
    q = np.array([0.18, -0.25, 0.07, 0.0], dtype=f32)
    q[3] = np.sqrt(f32(1.0) - (q[:3] * q[:3]).sum())
    m.nodes.append(M.Node(mesh=0, trs=M.Trs(translation=np.array([0.1, -0.2, 0.5], dtype=f32), rotation=q,
                                            scale=np.array([1.4, 0.9, 1.0], dtype=f32))))
    m.root.children.append(0)
    s = M.Scene()
    s.push_model(m)
    s.push_model(M.create_default_model())
    return s


def flighthelmet_scene():
    """rayca-soft/tests/gltf.rs:296-317 `flight`: FlightHelmet.gltf (94,722 triangles with the asset's own tangents, six
    PBR materials with albedo + NORMAL + metallic/roughness maps) and the default model with its camera moved to
    (0, 0.32, 1).  Two reductions keep the fixture committable: every texture is point-subsampled to 256 x 256
    (stride 8 / 4, no filtering), and the three leather textures missing from the reference checkout
    (.MISSING_LARGE_BLOBS) are replaced by flat 4 x 4 textures (albedo grey-brown, normal (0,0,1),
    occlusion/rough/metal (255, 160, 0))."""
    def hook(uri, load):
        try:
            im = load()
        except FileNotFoundError:
            flat = {"BaseColor": (150, 110, 80, 255), "Normal": (128, 128, 255, 255)}
            c = next((v for k, v in flat.items() if k in (uri or "")), (255, 160, 0, 255))
            return M.Image(4, 4, M.COLOR_RGBA8, np.ascontiguousarray(np.tile(np.array(c, dtype=np.uint8), (4, 4, 1))))
        st = max(1, im.width // 256)
        d = np.ascontiguousarray(im.data[::st, ::st])
        return M.Image(d.shape[1], d.shape[0], im.color_type, d)

    s = M.Scene()
    s.push_gltf_from_path(os.path.join(ASSETS, "flight-helmet/FlightHelmet.gltf"), image_hook=hook)
    dm = M.create_default_model()
    dm.get_mut_last_camera().trs.translation = np.array([0.0, 0.32, 1.0], dtype=f32)
    s.push_model(dm)
    return s


def scale_root(f):
    def tweak(s, h):
        s.nodes[h].trs.scale = (s.nodes[h].trs.scale * np.array(f, dtype=f32)).astype(f32)
    return tweak


def sdtf_scene(rel):
    s = M.Scene()
    _, cfg = s.push_sdtf_from_path(os.path.join(ASSETS, rel))
    return s, cfg


PT1 = dict(integrator=IntegratorStrategy.Pathtracer, max_depth=1)          # deterministic NEE slice
PT2 = dict(integrator=IntegratorStrategy.Pathtracer, max_depth=2)          # + one cosine bounce (keyed RNG)
RT0 = dict(integrator=IntegratorStrategy.Raytracer, max_depth=0)
RT3 = dict(integrator=IntegratorStrategy.Raytracer, max_depth=3)
FLAT = dict(integrator=IntegratorStrategy.Flat)
PTC = dict(integrator=IntegratorStrategy.Pathtracer, indirect_sampler=SamplerStrategy.Hemisphere, max_depth=5, samples_per_pixel=4,
           light_samples=1, light_stratify=False)
SC3 = dict(integrator=IntegratorStrategy.Scratcher, max_depth=3)

# scene name -> (builder, {case name -> (config kwargs, width, height)})
SCENES = {
    "triangle": (lambda: gltf_scene("triangle/triangle.gltf"),
                 {"pt1": (PT1, 128, 128), "rt0": (RT0, 128, 128), "flat": (FLAT, 64, 64),
                  "pt1_spp4": (dict(PT1, samples_per_pixel=4), 64, 64)}),
    "unit_triangle": (unit_triangle_scene, {"pt1": (PT1, 128, 128), "pt5": (dict(integrator=IntegratorStrategy.Pathtracer), 64, 64)}),
    "box": (lambda: gltf_scene("box/box.gltf"),
            {"pt1": (PT1, 128, 128), "rt0": (RT0, 128, 128), "rt3": (RT3, 96, 96), "pt2": (PT2, 96, 96),
             "pt1_nobvh": (dict(PT1, bvh=False), 64, 64), "pt1_spp2": (dict(PT1, samples_per_pixel=2), 64, 64),
             "pt1_gamma": (dict(PT1, gamma=2.2), 64, 64), "wide": (PT1, 160, 48)}),
    "orientation": (lambda: gltf_scene("orientation/OrientationTest.gltf", scale_root([0.2, 0.2, 0.2])),
                    {"pt1": (PT1, 128, 128), "rt0": (RT0, 96, 96)}),
    "cubes_over_plane": (cubes_over_plane_scene, {"pt1": (PT1, 128, 128), "rt3": (RT3, 96, 96), "sc3": (SC3, 96, 96), "pt3": (dict(integrator=IntegratorStrategy.Pathtracer, max_depth=3), 64, 64),
                                                  # branching bounces under point-light NEE (PBR materials, cosine sampler)
                                                  "pt3_ls2": (dict(integrator=IntegratorStrategy.Pathtracer, max_depth=3, light_samples=2), 64, 64)}),
    # normal-mapped shading (row a12 had no oracle evidence): a synthetic quad with a tilted tangent frame
    # under a rotated / scaled node, and the reference's own `flight` test scene
    "normalquad": (normalquad_scene, {"pt1": (PT1, 128, 128), "rt0": (RT0, 96, 96), "pt2": (PT2, 64, 64),
                                      "pt1_spp4": (dict(PT1, samples_per_pixel=4), 64, 64)}),
    "flighthelmet": (flighthelmet_scene, {"pt1": (PT1, 128, 128), "rt2": (dict(integrator=IntegratorStrategy.Raytracer, max_depth=2), 64, 64)}),
    "suzanne": (lambda: gltf_scene("suzanne/suzanne.gltf"), {"pt1": (PT1, 128, 128), "rt0": (RT0, 96, 96), "pt2": (PT2, 64, 64)}),
    # Scratcher (integrator/scratcher.rs): the reference's own fixture, the transmission branch, reflections
    "scratcher_sphere": (scratcher_sphere_scene, {"nobvh": (dict(integrator=IntegratorStrategy.Scratcher, bvh=False), 128, 128),
                                                  "sc3": (SC3, 64, 64)}),
    "glass": (glass_scene, {"sc3": (SC3, 128, 128), "sc0": (dict(integrator=IntegratorStrategy.Scratcher, max_depth=0), 64, 64),
                            "sc5_spp4": (dict(integrator=IntegratorStrategy.Scratcher, max_depth=5, samples_per_pixel=4), 48, 48)}),
    "scene7": ("sdtf/1/scene7.sdtf", {"sdtf": (None, 80, 60)}),
    # spheres (BvhGeometry::Sphere) under the Raytracer with the files' own cameras, lights and Phong materials
    # (rayca-soft/tests/sdtf.rs scene2..scene6); `nobvh` = the Config::bvh(false) those tests use
    "dice": ("sdtf/1/scene2.sdtf", {"sdtf": (None, 160, 120), "nobvh": (dict(bvh=False), 96, 72)}),
    "scene4": ("sdtf/1/scene4-specular.sdtf", {"sdtf": (None, 160, 120)}),
    "scene5": ("sdtf/1/scene5.sdtf", {"sdtf": (None, 160, 120)}),
    "scene6": ("sdtf/1/scene6.sdtf", {"sdtf": (None, 160, 120), "spp4": (dict(samples_per_pixel=4), 64, 48)}),
    # quad lights (two emissive triangles each) under AnalyticDirect / Direct (rayca-soft/tests/sdtf.rs, sdtf/2)
    "analytic": ("sdtf/2/analytic.sdtf", {"sdtf": (None, 160, 120)}),
    "direct9": ("sdtf/2/direct9.sdtf", {"sdtf": (None, 120, 90)}),
    "direct3x3": ("sdtf/2/direct3x3.sdtf", {"sdtf": (None, 120, 90)}),
    "cornell": ("sdtf/2/cornell.sdtf", {"sdtf": (None, 120, 120), "analytic": (dict(integrator=IntegratorStrategy.AnalyticDirect), 96, 96),
                                        # the path-tracing settings of sdtf/3 and sdtf/4 on the same box (cornellSimple, cornellNEE,
                                        # cornellRR, cornellCosine) at 4 spp: hemisphere / cosine bounces, quad-light NEE, roulette
                                        "pt_simple": (dict(PTC, direct_sampler=SamplerStrategy.NONE), 48, 48),
                                        "pt_nee": (dict(PTC, direct_sampler=SamplerStrategy.Nee), 48, 48),
                                        "pt_rr": (dict(PTC, direct_sampler=SamplerStrategy.Nee, russian_roulette=True, max_depth=16), 48, 48),
                                        "pt_brdf_rr": (dict(PTC, direct_sampler=SamplerStrategy.Nee, indirect_sampler=SamplerStrategy.Brdf,
                                                            russian_roulette=True, max_depth=16), 40, 40),
                                        "pt_mis_rr": (dict(PTC, direct_sampler=SamplerStrategy.Mis, indirect_sampler=SamplerStrategy.Brdf,
                                                           russian_roulette=True, max_depth=16), 40, 40),
                                        "pt_cosine_rr": (dict(PTC, direct_sampler=SamplerStrategy.Nee, indirect_sampler=SamplerStrategy.Cosine,
                                                              russian_roulette=True, max_depth=16), 40, 40),
                                        # light_samples > 1 WITH bounces: every hit spawns light_samples indirect rays
                                        # (pathtracer.rs:23-66), a tree of paths per pixel sample
                                        "pt_branch_nee": (dict(PTC, samples_per_pixel=2, direct_sampler=SamplerStrategy.Nee,
                                                               indirect_sampler=SamplerStrategy.Cosine, light_samples=3, max_depth=3), 40, 40),
                                        "pt_branch_simple": (dict(PTC, samples_per_pixel=2, direct_sampler=SamplerStrategy.NONE,
                                                                  light_samples=2, max_depth=3), 40, 40)}),
    # sdtf/4/ggx.sdtf: spheres with Phong and GGX materials of growing roughness under the Brdf sampler (its own settings,
    # 2 spp instead of 64; russian roulette as the file asks, which Config::apply drops: both variants)
    "ggx": ("sdtf/4/ggx.sdtf", {"brdf": (dict(samples_per_pixel=2), 160, 35),
                                "brdf_rr": (dict(samples_per_pixel=2, russian_roulette=True), 160, 35)}),
    # sdtf/4/mis.sdtf: glossy plates under four quad lights, multiple importance sampling, its own settings at 2 spp
    "mis": ("sdtf/4/mis.sdtf", {"sdtf": (dict(samples_per_pixel=2), 96, 64)}),
    "quadsphere": ("sdtf/2/sphere.sdtf", {"sdtf": (None, 120, 90)}),
    "dragon": ("sdtf/2/dragon.sdtf", {"ls4": (dict(light_samples=4), 48, 36)}),  # 100,002 triangles + 2 quad lights
}


def render_case(osc, cfg: Config, w, h, seed=0x5EED):
    opts = M.rayca_options(seed, 0, 0, 0, 0, 0, 0)
    rgba8, aux, st = osc.render(cfg.to_c(), w, h, options=opts, flags=O.COUNT)
    return dict(rgba8=rgba8, rgba_f32=aux.rgba_f32.reshape(h, w, 4), hit_src=aux.hit_src.reshape(h, w),
                hit_prim=aux.hit_prim.reshape(h, w), hit_blas=aux.hit_blas.reshape(h, w),
                depth=aux.depth.reshape(h, w), uv=aux.uv.reshape(h, w, 2),
                rays=np.array([st.rays_primary, st.rays_shadow, st.rays_secondary], dtype=np.uint64),
                walk=np.array([st.ord_nodes, st.ord_tris, st.exh_nodes, st.exh_tris], dtype=np.uint64),
                config=np.array([int(cfg.bvh), cfg.light_samples, int(cfg.light_stratify), cfg.samples_per_pixel,
                                 int(cfg.russian_roulette), int(cfg.direct_sampler), int(cfg.indirect_sampler),
                                 int(cfg.integrator), cfg.max_depth], dtype=np.uint32),
                gamma=np.array([cfg.gamma], dtype=f32), seed=np.array([seed], dtype=np.uint64))


def main(names):
    """`names`: scene names (all their cases) and / or `scene.case` entries (only those cases are rewritten)."""
    only = {n for n in names if "." in n}
    names = [n.split(".")[0] for n in names]
    for name, (builder, cases) in SCENES.items():
        if names and name not in names:
            continue
        if only and not any(o.startswith(name + ".") for o in only) :
            pass
        elif only:
            cases = {c: v for c, v in cases.items() if f"{name}.{c}" in only}
        t0 = time.time()
        sdtf_cfg = None
        if isinstance(builder, str):
            scene, sdtf_cfg = sdtf_scene(builder)
        else:
            scene = builder()
        desc = scene.to_desc()
        desc.save(os.path.join(HERE, f"{name}.scene.npz"))
        by_bvh = {}
        for case, (kw, w, h) in cases.items():
            cfg = Config()
            if sdtf_cfg is not None:
                cfg.apply(sdtf_cfg)
            for k, v in (kw or {}).items():
                setattr(cfg, k, v)
            if cfg.bvh not in by_bvh:
                by_bvh[cfg.bvh] = O.OracleScene(desc, bvh=cfg.bvh)
            out = render_case(by_bvh[cfg.bvh], cfg, w, h)
            np.savez_compressed(os.path.join(HERE, f"{name}.{case}.npz"), **out)
            print(f"{name}.{case}: {w}x{h} hits={(out['hit_src'] != 0xFFFFFFFF).sum()} rays={out['rays'].tolist()} "
                  f"mean rgb={out['rgba_f32'][..., :3].mean():.5f}")
        print(f"{name}: {desc.triangle_count} triangles, {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main(sys.argv[1:])
