//! Lowering: the closed-world walk over the reference's open-world scene (`Box<dyn Shape>`, `Transformed<T>`,
//! `KdTree<Box<dyn Bounded>>`, `Arc<Mesh>`) into the POD records of include/rptx.h.  This is the one piece of new
//! host logic of the mirror; rpt_b200/world.py:lower_scene is the same function in the Python host.
use std::collections::HashMap;
use std::ptr;
use std::sync::Arc;

use rpt_cuda_sys as sys;

use crate::environment::Environment;
use crate::ffi::{check, Ctx};
use crate::kdtree::{Bounded, KdTree};
use crate::light::Light;
use crate::scene::Scene;
use crate::shape::{Cube, Mesh, MonomialSurface, Plane, Shape, Sphere, Transformed, Triangle};

/// One lowered shape: a leaf record, or a group of leaf records (its children own their storage here until the
/// call that passes them to the library returns).
pub struct ShapeDesc {
    pub(crate) desc: sys::rptx_shape_desc,
    pub(crate) children: Vec<sys::rptx_shape_desc>,
}

const IDENTITY: [f64; 16] = [1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 0.0, 1.0];

impl ShapeDesc {
    fn leaf(kind: i32) -> Self {
        let desc = sys::rptx_shape_desc {
            kind,
            has_transform: 0,
            transform: IDENTITY,
            mesh_id: 0,
            n_children: 0,
            plane_normal: [0.0; 3],
            plane_value: 0.0,
            mono_height: 0.0,
            mono_exp: 0.0,
            children: ptr::null(),
        };
        Self { desc, children: Vec::new() }
    }

    /// Transformed<T>::new(shape, m) over an already lowered shape: compose on the left, keep one level
    /// (src/shape.rs:102-127, 238-289)
    fn pre_multiply(mut self, m: &glm::DMat4) -> Option<Self> {
        if self.desc.kind == sys::RPTX_SHAPE_GROUP {
            return None; // the reference never transforms a KdTree of boxed shapes; the library rejects it too
        }
        let cur = glm::DMat4::from_row_slice(&self.desc.transform);
        let out = if self.desc.has_transform != 0 { m * cur } else { *m };
        for r in 0..4 {
            for c in 0..4 {
                self.desc.transform[4 * r + c] = out[(r, c)];
            }
        }
        self.desc.has_transform = 1;
        Some(self)
    }
}

/// State of one scene lowering: the library scene being filled and the meshes already uploaded.
pub struct Lowering {
    pub(crate) scene: *mut sys::rptx_scene,
    pub(crate) ctx: *mut sys::rptx_ctx,
    meshes: HashMap<*const Vec<Triangle>, u32>,
}

impl Lowering {
    fn mesh_id(&mut self, mesh: &Mesh) -> u32 {
        let key = Arc::as_ptr(&mesh.triangles);
        if let Some(&id) = self.meshes.get(&key) {
            return id;
        }
        let mut v = Vec::with_capacity(9 * mesh.triangles.len());
        let mut n = Vec::with_capacity(9 * mesh.triangles.len());
        for t in mesh.triangles.iter() {
            for p in &[t.v1, t.v2, t.v3] {
                v.extend_from_slice(&[p.x, p.y, p.z]);
            }
            for p in &[t.n1, t.n2, t.n3] {
                n.extend_from_slice(&[p.x, p.y, p.z]);
            }
        }
        let mut id = 0u32;
        check(self.ctx, unsafe { sys::rptx_scene_add_mesh(self.scene, v.as_ptr(), n.as_ptr(), mesh.triangles.len(), &mut id) });
        self.meshes.insert(key, id);
        id
    }
}

impl Shape for Sphere {
    fn lower(&self, _: &mut Lowering) -> Option<ShapeDesc> {
        Some(ShapeDesc::leaf(sys::RPTX_SHAPE_SPHERE))
    }
}
impl Shape for Cube {
    fn lower(&self, _: &mut Lowering) -> Option<ShapeDesc> {
        Some(ShapeDesc::leaf(sys::RPTX_SHAPE_CUBE))
    }
}
impl Shape for Plane {
    fn lower(&self, _: &mut Lowering) -> Option<ShapeDesc> {
        let mut d = ShapeDesc::leaf(sys::RPTX_SHAPE_PLANE);
        d.desc.plane_normal = [self.normal.x, self.normal.y, self.normal.z];
        d.desc.plane_value = self.value;
        Some(d)
    }
}
impl Shape for MonomialSurface {
    fn lower(&self, _: &mut Lowering) -> Option<ShapeDesc> {
        let mut d = ShapeDesc::leaf(sys::RPTX_SHAPE_MONOMIAL);
        d.desc.mono_height = self.height;
        d.desc.mono_exp = self.exp;
        Some(d)
    }
}
impl Shape for Triangle {
    /// a lone triangle is a one-triangle mesh
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        Mesh::new(vec![*self]).lower(cx)
    }
}
impl Shape for Mesh {
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        let mut d = ShapeDesc::leaf(sys::RPTX_SHAPE_MESH);
        d.desc.mesh_id = cx.mesh_id(self);
        Some(d)
    }
}
impl<T: Shape> Shape for Transformed<T> {
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        self.shape.lower(cx)?.pre_multiply(&self.transform)
    }
}
impl<T: Bounded> Shape for KdTree<T> {
    /// KdTree<Box<dyn Bounded>> of transformed instances (examples/fractal_teapots.rs:56): one GROUP object whose
    /// children are bounded leaf shapes.  A KdTree<Triangle> built by hand lowers the same way as a Mesh would,
    /// through its triangles' one-triangle meshes.
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        let mut children = Vec::with_capacity(self.objects.len());
        for o in &self.objects {
            let c = o.lower(cx)?;
            if c.desc.kind == sys::RPTX_SHAPE_GROUP || c.desc.kind == sys::RPTX_SHAPE_PLANE {
                return None; // nested groups / unbounded children: not expressible (the reference's Plane is not Bounded either)
            }
            children.push(c.desc);
        }
        let mut d = ShapeDesc::leaf(sys::RPTX_SHAPE_GROUP);
        d.desc.n_children = children.len() as u32;
        d.children = children;
        Some(d)
    }
}

/// Lower and commit a whole scene (objects in order, lights in order, media[0], environment): `rptx_scene_*`.
pub(crate) fn lower_scene(ctx: &Ctx, scene: &Scene) -> *mut sys::rptx_scene {
    let mut handle: *mut sys::rptx_scene = ptr::null_mut();
    check(ctx.raw, unsafe { sys::rptx_scene_create(ctx.raw, &mut handle) });
    let mut cx = Lowering { scene: handle, ctx: ctx.raw, meshes: HashMap::new() };
    for (i, object) in scene.objects.iter().enumerate() {
        let mut d = object.shape.lower(&mut cx).unwrap_or_else(|| panic!("object {}: shape cannot be lowered to the CUDA core", i));
        d.desc.children = if d.children.is_empty() { ptr::null() } else { d.children.as_ptr() };
        let m = object.material.lower();
        check(ctx.raw, unsafe { sys::rptx_scene_add_object(handle, &d.desc, &m, ptr::null_mut()) });
    }
    for (i, light) in scene.lights.iter().enumerate() {
        let mut l = sys::rptx_light {
            kind: 0,
            _pad: 0,
            color: [0.0; 3],
            vec: [0.0; 3],
            shape: ShapeDesc::leaf(sys::RPTX_SHAPE_SPHERE).desc,
            material: crate::material::Material::default().lower(),
        };
        match light {
            Light::Point(c, p) => {
                l.kind = sys::RPTX_LIGHT_POINT;
                l.color = [c.x, c.y, c.z];
                l.vec = [p.x, p.y, p.z];
            }
            Light::Ambient(c) => {
                l.kind = sys::RPTX_LIGHT_AMBIENT;
                l.color = [c.x, c.y, c.z];
            }
            Light::Directional(c, d) => {
                l.kind = sys::RPTX_LIGHT_DIRECTIONAL;
                l.color = [c.x, c.y, c.z];
                l.vec = [d.x, d.y, d.z];
            }
            Light::Object(object) => {
                let d = object.shape.lower(&mut cx).unwrap_or_else(|| panic!("light {}: shape cannot be lowered to the CUDA core", i));
                assert!(d.children.is_empty(), "light {}: a group cannot be an object light", i);
                l.kind = sys::RPTX_LIGHT_OBJECT;
                l.shape = d.desc;
                l.material = object.material.lower();
            }
        }
        check(ctx.raw, unsafe { sys::rptx_scene_add_light(handle, &l) });
    }
    if let Some(m) = scene.media.first() {
        // only media[0] is ever used by the reference (src/renderer.rs:190)
        check(ctx.raw, unsafe { sys::rptx_scene_set_medium(handle, m.lower_kind(), m.absorption, m.scattering) });
    }
    match &scene.environment {
        Environment::Color(c) => {
            let rgb = [c.x, c.y, c.z];
            check(ctx.raw, unsafe { sys::rptx_scene_set_environment_color(handle, rgb.as_ptr()) });
        }
        Environment::Hdri(h) => {
            check(ctx.raw, unsafe { sys::rptx_scene_set_environment_hdri(handle, h.width, h.height, h.pixels.as_ptr()) });
        }
    }
    check(ctx.raw, unsafe { sys::rptx_scene_commit(handle) });
    handle
}
