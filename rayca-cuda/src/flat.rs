// flat.rs — `FlatScene::from(&Scene)`: walks a `rayca_model::Scene` (rayca-model/src/scene.rs:47-53: `Pack<Node>`,
// `Pack<Model>`, root) into the plain arrays `rayca_scene_desc` points at (include/rayca_cuda.h).  Nothing is computed
// here: flattening the graph, emitting primitives and building the BVH all happen inside the library
// (rayca_cuda_scene_create replaces rayca-soft/src/scene.rs:90-99).  UNCOMPILED in the authoring container.
//
// Handles: a `Handle<T>` indexes `Pack::indices`, which points into the contiguous `Pack::vec`
// (rayca-util/src/pack.rs:144-205).  The C arrays are emitted in `vec` order and every handle is translated through
// `get_indices()`, so packs with removed elements flatten correctly.
use rayca_model::*;

use crate::ffi::*;

/// Owns every array a `RaycaSceneDesc` borrows; `desc()` is valid for as long as `self` (and the `Scene` the image
/// pointers came from) lives.
#[derive(Default)]
pub struct FlatScene {
    pub root_trs: Option<RaycaTrs>,
    pub root_children: (u32, u32),
    pub scene_nodes: Vec<RaycaNode>,
    pub models: Vec<RaycaModel>,
    pub nodes: Vec<RaycaNode>,
    pub children: Vec<u32>,
    pub meshes: Vec<RaycaMesh>,
    pub mesh_primitives: Vec<u32>,
    pub primitives: Vec<RaycaPrimitive>,
    pub geometries: Vec<RaycaGeometry>,
    pub vertices: Vec<RaycaVertex>,
    pub indices: Vec<u32>,
    pub materials: Vec<RaycaMaterial>,
    pub pbr_materials: Vec<RaycaPbrMaterial>,
    pub phong_materials: Vec<RaycaPhongMaterial>,
    pub ggx_materials: Vec<RaycaGgxMaterial>,
    pub textures: Vec<RaycaTexture>,
    pub images: Vec<RaycaImage>,
    pub lights: Vec<RaycaLight>,
    pub cameras: Vec<RaycaCamera>,
}

/// position of `handle`'s element in `pack.as_slice()`, or RAYCA_NONE
fn slot<T>(pack: &Pack<T>, handle: Handle<T>) -> i32 {
    if !handle.is_valid() {
        return RAYCA_NONE;
    }
    match pack.get_indices().get(handle.as_index()) {
        Some(i) => *i as i32,
        None => RAYCA_NONE,
    }
}

fn opt_slot<T>(pack: &Pack<T>, handle: Option<Handle<T>>) -> i32 {
    handle.map_or(RAYCA_NONE, |h| slot(pack, h))
}

fn trs(t: &Trs) -> RaycaTrs {
    RaycaTrs { translation: t.translation.to_array(), rotation: t.rotation.to_array(), scale: t.scale.to_array() }
}

fn color(c: &Color) -> [f32; 4] {
    [c.r, c.g, c.b, c.a]
}

fn point(p: &Point3) -> [f32; 3] {
    [p.get_x(), p.get_y(), p.get_z()]
}

impl FlatScene {
    /// appends `node`'s child handles (translated to `vec` positions of `pack`) and returns (first_child, child_count)
    fn push_children(&mut self, pack: &Pack<Node>, node: &Node) -> (u32, u32) {
        let first = self.children.len() as u32;
        for c in &node.children {
            let s = slot(pack, *c);
            assert!(s >= 0, "dangling child handle");
            self.children.push(s as u32);
        }
        (first, node.children.len() as u32)
    }

    /// rayca-model/src/node.rs:10-31.  `graph` is the pack the node's children live in; `model`, when given, resolves
    /// the node's mesh / camera / light handles (scene-level nodes only carry `model` and children).
    fn node(&mut self, graph: &Pack<Node>, node: &Node, model: Option<&Model>, scene_models: Option<&Pack<Model>>) -> RaycaNode {
        let (first_child, child_count) = self.push_children(graph, node);
        RaycaNode {
            trs: trs(&node.trs),
            mesh: model.map_or(RAYCA_NONE, |m| opt_slot(&m.meshes, node.mesh)),
            camera: model.map_or(RAYCA_NONE, |m| opt_slot(&m.cameras, node.camera)),
            light: model.map_or(RAYCA_NONE, |m| opt_slot(&m.lights, node.light)),
            model: scene_models.map_or(RAYCA_NONE, |p| opt_slot(p, node.model)),
            first_child,
            child_count,
        }
    }
}

impl FlatScene {
    /// rayca-geometry/src/geometry.rs: TriangleMesh (vertices + u8 / u16 / u32 indices, widened here) or Sphere
    fn geometry(&mut self, g: &Geometry) -> RaycaGeometry {
        match g {
            Geometry::TriangleMesh(mesh) => {
                let first_vertex = self.vertices.len() as u32;
                for v in &mesh.vertices {
                    // Vertex { pos: Point3, ext: VertexExt { color, normal, tangent, bitangent, uv } } (vertex.rs:64-72,139-143)
                    self.vertices.push(RaycaVertex {
                        pos: point(&v.pos),
                        color: color(&v.ext.color),
                        normal: v.ext.normal.to_array(),
                        tangent: v.ext.tangent.to_array(),
                        bitangent: v.ext.bitangent.to_array(),
                        uv: [v.ext.uv.x, v.ext.uv.y],
                    });
                }
                let first_index = self.indices.len() as u32;
                // TriangleIndices keeps raw bytes + a ComponentType (triangle.rs:215-306); the same three cases
                // BvhPrimitive::from_triangle_mesh dispatches on (bvh/primitive.rs:239-268)
                match mesh.indices.index_type {
                    ComponentType::U8 => self.indices.extend(mesh.indices.get_indices::<u8>().iter().map(|i| *i as u32)),
                    ComponentType::U16 => self.indices.extend(mesh.indices.get_indices::<u16>().iter().map(|i| *i as u32)),
                    ComponentType::U32 => self.indices.extend_from_slice(mesh.indices.get_indices::<u32>()),
                    _ => panic!("Unsupported index type"),
                }
                RaycaGeometry {
                    kind: RAYCA_GEOMETRY_TRIANGLE_MESH,
                    first_vertex,
                    vertex_count: mesh.vertices.len() as u32,
                    first_index,
                    index_count: self.indices.len() as u32 - first_index,
                    center: [0.0; 3],
                    radius: 0.0,
                }
            }
            Geometry::Sphere(s) => RaycaGeometry {
                kind: RAYCA_GEOMETRY_SPHERE,
                first_vertex: 0,
                vertex_count: 0,
                first_index: 0,
                index_count: 0,
                center: point(&s.get_model_center()),
                radius: s.get_model_radius(),
            },
        }
    }

    /// rayca-model/src/light/{mod,point,directional,quad}.rs.  DirectionalLight keeps colour and intensity private:
    /// `get_intensity()` = intensity * color is passed as the colour with intensity 1 (x * 1.0 is exact).
    fn light(model: &Model, l: &Light) -> RaycaLight {
        let mut r = RaycaLight { kind: 0, color: [1.0; 4], intensity: 1.0, attenuation: [0.0; 3], ab: [0.0; 3], ac: [0.0; 3], material: RAYCA_NONE };
        match l {
            Light::Directional(d) => {
                r.kind = RAYCA_LIGHT_DIRECTIONAL;
                r.color = color(&d.get_intensity());
            }
            Light::Point(p) => {
                r.kind = RAYCA_LIGHT_POINT;
                r.color = color(&p.color);
                r.intensity = p.intensity;
                r.attenuation = p.attenuation.to_array();
            }
            Light::Quad(q) => {
                r.kind = RAYCA_LIGHT_QUAD;
                r.color = color(&q.color);
                r.intensity = q.intensity;
                r.ab = q.ab.to_array();
                r.ac = q.ac.to_array();
                r.material = slot(&model.materials, q.material);
            }
        }
        r
    }

    /// rayca-model/src/model.rs:31-51: one model's packs appended to the global arrays; offsets recorded in RaycaModel
    fn model(&mut self, m: &Model) -> RaycaModel {
        let (root_first_child, root_child_count) = self.push_children(&m.nodes, &m.root);
        let mut r: RaycaModel = unsafe { core::mem::zeroed() };
        r.root_trs = trs(&m.root.trs);
        r.root_first_child = root_first_child;
        r.root_child_count = root_child_count;

        r.node_offset = self.nodes.len() as u32;
        r.node_count = m.nodes.len() as u32;
        for n in m.nodes.as_slice() {
            let cn = self.node(&m.nodes, n, Some(m), None);
            self.nodes.push(cn);
        }
        r.mesh_offset = self.meshes.len() as u32;
        r.mesh_count = m.meshes.len() as u32;
        for mesh in m.meshes.as_slice() {
            let first_primitive = self.mesh_primitives.len() as u32;
            for p in &mesh.primitives {
                self.mesh_primitives.push(slot(&m.primitives, *p) as u32);
            }
            self.meshes.push(RaycaMesh { first_primitive, primitive_count: mesh.primitives.len() as u32 });
        }
        r.primitive_offset = self.primitives.len() as u32;
        r.primitive_count = m.primitives.len() as u32;
        for p in m.primitives.as_slice() {
            self.primitives.push(RaycaPrimitive { geometry: slot(&m.geometries, p.geometry) as u32, material: slot(&m.materials, p.material) });
        }
        r.geometry_offset = self.geometries.len() as u32;
        r.geometry_count = m.geometries.len() as u32;
        for g in m.geometries.as_slice() {
            let cg = self.geometry(g);
            self.geometries.push(cg);
        }
        self.model_shading(m, &mut r);
        r
    }

    /// materials, textures, images, lights and cameras of one model
    fn model_shading(&mut self, m: &Model, r: &mut RaycaModel) {
        r.material_offset = self.materials.len() as u32;
        r.material_count = m.materials.len() as u32;
        for mat in m.materials.as_slice() {
            // rayca-model/src/material/mod.rs:15-20
            self.materials.push(match mat {
                Material::Pbr(h) => RaycaMaterial { kind: RAYCA_MATERIAL_PBR, index: slot(&m.pbr_materials, *h) },
                Material::Phong(h) => RaycaMaterial { kind: RAYCA_MATERIAL_PHONG, index: slot(&m.phong_materials, *h) },
                Material::Ggx(h) => RaycaMaterial { kind: RAYCA_MATERIAL_GGX, index: slot(&m.ggx_materials, *h) },
            });
        }
        r.pbr_offset = self.pbr_materials.len() as u32;
        r.pbr_count = m.pbr_materials.len() as u32;
        for p in m.pbr_materials.as_slice() {
            self.pbr_materials.push(RaycaPbrMaterial {
                color: color(&p.color),
                albedo: slot(&m.textures, p.albedo),
                normal: slot(&m.textures, p.normal),
                metallic_factor: p.metallic_factor,
                roughness_factor: p.roughness_factor,
                metallic_roughness: slot(&m.textures, p.metallic_roughness),
            });
        }
        r.phong_offset = self.phong_materials.len() as u32;
        r.phong_count = m.phong_materials.len() as u32;
        for p in m.phong_materials.as_slice() {
            self.phong_materials.push(RaycaPhongMaterial {
                ambient: color(&p.ambient),
                emission: color(&p.emission),
                diffuse: color(&p.diffuse),
                specular: color(&p.specular),
                shininess: p.shininess,
            });
        }
        r.ggx_offset = self.ggx_materials.len() as u32;
        r.ggx_count = m.ggx_materials.len() as u32;
        for g in m.ggx_materials.as_slice() {
            self.ggx_materials.push(RaycaGgxMaterial { diffuse: color(&g.diffuse), specular: color(&g.specular), roughness: g.roughness });
        }
        r.texture_offset = self.textures.len() as u32;
        r.texture_count = m.textures.len() as u32;
        for t in m.textures.as_slice() {
            self.textures.push(RaycaTexture { image: slot(&m.images, t.image) });
        }
        r.image_offset = self.images.len() as u32;
        r.image_count = m.images.len() as u32;
        for im in m.images.as_slice() {
            // rayca-model/src/image.rs:26-37: decoded, row-major, top-left origin; the bytes are borrowed from the Scene
            self.images.push(RaycaImage {
                width: im.width(),
                height: im.height(),
                color_type: match im.color_type {
                    ColorType::RGB8 => RAYCA_COLOR_RGB8,
                    ColorType::RGBA8 => RAYCA_COLOR_RGBA8,
                    ColorType::RGBA32F => RAYCA_COLOR_RGBA32F,
                },
                reserved: 0,
                data: im.bytes().as_ptr() as *const core::ffi::c_void,
            });
        }
        r.light_offset = self.lights.len() as u32;
        r.light_count = m.lights.len() as u32;
        for l in m.lights.as_slice() {
            self.lights.push(Self::light(m, l));
        }
        r.camera_offset = self.cameras.len() as u32;
        r.camera_count = m.cameras.len() as u32;
        for c in m.cameras.as_slice() {
            self.cameras.push(RaycaCamera { yfov_radians: c.yfov_radians });
        }
    }
}

impl From<&Scene> for FlatScene {
    /// rayca-model/src/scene.rs:47-53: scene-level nodes reference models; each model brings its own packs
    fn from(scene: &Scene) -> Self {
        let mut f = FlatScene::default();
        f.root_trs = Some(trs(&scene.root.trs));
        f.root_children = f.push_children(&scene.nodes, &scene.root);
        for n in scene.nodes.as_slice() {
            let cn = f.node(&scene.nodes, n, None, Some(&scene.models));
            f.scene_nodes.push(cn);
        }
        for m in scene.models.as_slice() {
            let cm = f.model(m);
            f.models.push(cm);
        }
        f
    }
}

impl FlatScene {
    /// The descriptor the C ABI takes; borrows `self` (and, through the image pointers, the `Scene`).
    pub fn desc(&self) -> RaycaSceneDesc {
        fn ptr<T>(v: &Vec<T>) -> *const T {
            if v.is_empty() { core::ptr::NonNull::dangling().as_ptr() } else { v.as_ptr() }
        }
        RaycaSceneDesc {
            root_trs: self.root_trs.expect("FlatScene::from"),
            root_first_child: self.root_children.0,
            root_child_count: self.root_children.1,
            scene_nodes: ptr(&self.scene_nodes),
            scene_node_count: self.scene_nodes.len() as u32,
            models: ptr(&self.models),
            model_count: self.models.len() as u32,
            nodes: ptr(&self.nodes),
            node_count: self.nodes.len() as u32,
            children: ptr(&self.children),
            children_count: self.children.len() as u32,
            meshes: ptr(&self.meshes),
            mesh_count: self.meshes.len() as u32,
            mesh_primitives: ptr(&self.mesh_primitives),
            mesh_primitive_count: self.mesh_primitives.len() as u32,
            primitives: ptr(&self.primitives),
            primitive_count: self.primitives.len() as u32,
            geometries: ptr(&self.geometries),
            geometry_count: self.geometries.len() as u32,
            vertices: ptr(&self.vertices),
            vertex_count: self.vertices.len() as u64,
            indices: ptr(&self.indices),
            index_count: self.indices.len() as u64,
            materials: ptr(&self.materials),
            material_count: self.materials.len() as u32,
            pbr_materials: ptr(&self.pbr_materials),
            pbr_count: self.pbr_materials.len() as u32,
            phong_materials: ptr(&self.phong_materials),
            phong_count: self.phong_materials.len() as u32,
            ggx_materials: ptr(&self.ggx_materials),
            ggx_count: self.ggx_materials.len() as u32,
            textures: ptr(&self.textures),
            texture_count: self.textures.len() as u32,
            images: ptr(&self.images),
            image_count: self.images.len() as u32,
            lights: ptr(&self.lights),
            light_count: self.lights.len() as u32,
            cameras: ptr(&self.cameras),
            camera_count: self.cameras.len() as u32,
        }
    }
}
