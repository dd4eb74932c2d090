//! src/scene.rs:12-81 of the reference, unchanged: containers plus the SceneAdd impls, including the one that adds
//! a (Mesh, Material) as an object AND a Light::Object copy.
use crate::environment::Environment;
use crate::light::Light;
use crate::medium::Medium;
use crate::object::Object;
use crate::{Cube, Material, Mesh, Transformed};

/// Type for adding light and object at the same time
pub type LightAndMeshObject = (Mesh, Material);

/// Object representing a scene that can be rendered
#[derive(Default)]
pub struct Scene {
    pub objects: Vec<Object>,
    pub lights: Vec<Light>,
    pub media: Vec<Medium>,
    pub environment: Environment,
}

impl Scene {
    /// Construct a new, empty scene
    pub fn new() -> Self {
        Default::default()
    }
}

/// Trait that allows adding an object or light to a scene
pub trait SceneAdd<T> {
    fn add(&mut self, node: T);
}

impl SceneAdd<Object> for Scene {
    fn add(&mut self, object: Object) {
        self.objects.push(object);
    }
}

impl SceneAdd<Light> for Scene {
    fn add(&mut self, light: Light) {
        self.lights.push(light);
    }
}

impl SceneAdd<LightAndMeshObject> for Scene {
    fn add(&mut self, m: LightAndMeshObject) {
        let (mesh, material) = m;
        let obj = Object::new(mesh.clone()).material(material);
        let light = Light::Object(Object::new(mesh).material(material));
        self.add(obj);
        self.add(light);
    }
}

impl SceneAdd<(Transformed<Cube>, Material)> for Scene {
    fn add(&mut self, m: (Transformed<Cube>, Material)) {
        let (cube, material) = m;
        let obj = Object::new(cube).material(material);
        let light = Light::Object(Object::new(cube).material(material));
        self.add(obj);
        self.add(light);
    }
}

impl SceneAdd<Medium> for Scene {
    fn add(&mut self, medium: Medium) {
        self.media.push(medium);
    }
}
