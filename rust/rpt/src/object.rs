//! src/object.rs:10-31 of the reference.
use crate::material::Material;
use crate::shape::Shape;

/// An object rendered in a scene
pub struct Object {
    /// Basic geometry of the object
    pub shape: Box<dyn Shape>,
    /// Material of the object
    pub material: Material,
}

impl Object {
    /// Create a new object from a shape, with default material
    pub fn new<T: Shape + 'static>(shape: T) -> Self {
        Self { shape: Box::new(shape), material: Material::default() }
    }

    /// Set the material of the object (builder pattern)
    pub fn material(mut self, material: Material) -> Self {
        self.material = material;
        self
    }
}
