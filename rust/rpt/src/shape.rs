//! src/shape.rs + src/shape/*.rs of the reference: the shape types and the Transformable chain.
//!
//! The reference's `trait Shape { intersect, sample }` is the code the CUDA kernels replace (`closest_hit`,
//! `sample_light_shape`).  What a shape has to do here is describe itself to the device: `Shape::lower`
//! (lower.rs).  The set is closed — Sphere, Plane, Cube, Triangle/Mesh, MonomialSurface, Transformed<T>,
//! KdTree<T>, and Box/Arc of those — exactly the shapes the reference ships.
use std::sync::Arc;

use crate::kdtree::{Bounded, BoundingBox};
use crate::lower::{Lowering, ShapeDesc};

/// Represents a physical shape (src/shape.rs:19-26)
pub trait Shape: Send + Sync {
    /// Describe the shape to the CUDA core.  `None`: the shape is outside the closed set the device knows
    /// (a user-defined `Shape`); `Renderer::render` then panics with a message naming the object.
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc>;
}

impl<T: Shape + ?Sized> Shape for Box<T> {
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        self.as_ref().lower(cx)
    }
}

impl<T: Shape + ?Sized> Shape for Arc<T> {
    fn lower(&self, cx: &mut Lowering) -> Option<ShapeDesc> {
        self.as_ref().lower(cx)
    }
}

/// An infinite ray in one direction (src/shape.rs:50-73)
#[derive(Copy, Clone)]
pub struct Ray {
    pub origin: glm::DVec3,
    pub dir: glm::DVec3,
}

impl Ray {
    pub fn at(&self, time: f64) -> glm::DVec3 {
        self.origin + time * self.dir
    }
}

/// Record of when a hit occurs, and the corresponding normal (src/shape.rs:76-99)
pub struct HitRecord {
    pub time: f64,
    pub normal: glm::DVec3,
}

impl Default for HitRecord {
    fn default() -> Self {
        Self { time: f64::INFINITY, normal: glm::vec3(0.0, 0.0, 0.0) }
    }
}

impl HitRecord {
    pub fn new() -> Self {
        Default::default()
    }
}

/// A unit sphere centered at the origin (src/shape/sphere.rs:10)
#[derive(Copy, Clone)]
pub struct Sphere;

/// A plane represented by the linear equation x . normal = value (src/shape/plane.rs:7-13)
#[derive(Copy, Clone)]
pub struct Plane {
    pub normal: glm::DVec3,
    pub value: f64,
}

/// A unit cube centered at the origin (src/shape/cube.rs:9)
#[derive(Copy, Clone)]
pub struct Cube;

/// y = height * (x^2 + z^2)^(exp / 2) over the unit disc (src/shape/monomial_surface.rs:8-19)
#[derive(Copy, Clone)]
pub struct MonomialSurface {
    pub height: f64,
    pub exp: f64,
}

/// A triangle with three vertices and three normals (src/shape/mesh.rs:9-23)
#[derive(Copy, Clone)]
pub struct Triangle {
    pub v1: glm::DVec3,
    pub v2: glm::DVec3,
    pub v3: glm::DVec3,
    pub n1: glm::DVec3,
    pub n2: glm::DVec3,
    pub n3: glm::DVec3,
}

impl Triangle {
    /// Construct a triangle from three vertices, inferring the normals (src/shape/mesh.rs:27-38)
    pub fn from_vertices(v1: glm::DVec3, v2: glm::DVec3, v3: glm::DVec3) -> Self {
        let n = (v2 - v1).cross(&(v3 - v1)).normalize();
        Self { v1, v2, v3, n1: n, n2: n, n3: n }
    }
}

/// A triangle mesh (the reference's `Mesh = KdTree<Triangle>`, src/shape/mesh.rs:103).  The kd-tree is gone: the
/// device builds its own BVH over the triangles at `rptx_scene_commit`.  Clones share the triangle storage, and a
/// shared mesh is uploaded once (`Arc` identity is the cache key, as `Arc<Mesh>` instancing in
/// examples/fractal_teapots.rs:17,49 expects).
#[derive(Clone)]
pub struct Mesh {
    pub(crate) triangles: Arc<Vec<Triangle>>,
}

impl Mesh {
    pub fn new(triangles: Vec<Triangle>) -> Self {
        Self { triangles: Arc::new(triangles) }
    }
    pub fn triangles(&self) -> &[Triangle] {
        &self.triangles
    }
}

impl Bounded for Sphere {
    fn bounding_box(&self) -> BoundingBox {
        BoundingBox { p_min: glm::vec3(-1.0, -1.0, -1.0), p_max: glm::vec3(1.0, 1.0, 1.0) }
    }
}
impl Bounded for Cube {
    fn bounding_box(&self) -> BoundingBox {
        BoundingBox { p_min: glm::vec3(-0.5, -0.5, -0.5), p_max: glm::vec3(0.5, 0.5, 0.5) }
    }
}
impl Bounded for MonomialSurface {
    fn bounding_box(&self) -> BoundingBox {
        BoundingBox { p_min: glm::vec3(-1.0, 0.0, -1.0), p_max: glm::vec3(1.0, self.height, 1.0) }
    }
}
impl Bounded for Triangle {
    fn bounding_box(&self) -> BoundingBox {
        BoundingBox { p_min: glm::min2(&glm::min2(&self.v1, &self.v2), &self.v3), p_max: glm::max2(&glm::max2(&self.v1, &self.v2), &self.v3) }
    }
}
impl Bounded for Mesh {
    fn bounding_box(&self) -> BoundingBox {
        self.triangles.iter().map(Triangle::bounding_box).fold(BoundingBox::default(), |a, b| a.merge(&b))
    }
}

/// A shape that has been composed with a transformation (src/shape.rs:102-127).  Only the forward matrix is
/// kept: the inverse, the normal matrix and the area scale are derived by the library at commit.
#[derive(Copy, Clone)]
pub struct Transformed<T> {
    pub(crate) shape: T,
    pub(crate) transform: glm::DMat4,
}

impl<T> Transformed<T> {
    fn new(shape: T, transform: glm::DMat4) -> Self {
        Self { shape, transform }
    }
}

impl<T: Bounded> Bounded for Transformed<T> {
    /// src/shape.rs:154-173: box of the eight transformed corners
    fn bounding_box(&self) -> BoundingBox {
        let BoundingBox { p_min, p_max } = self.shape.bounding_box();
        let mut out = BoundingBox::default();
        for &x in &[p_min.x, p_max.x] {
            for &y in &[p_min.y, p_max.y] {
                for &z in &[p_min.z, p_max.z] {
                    let p = (self.transform * glm::vec4(x, y, z, 1.0)).xyz();
                    out = out.merge(&BoundingBox { p_min: p, p_max: p });
                }
            }
        }
        out
    }
}

/// An object that can be transformed (src/shape.rs:180-236)
pub trait Transformable<T> {
    fn translate(self, v: &glm::DVec3) -> Transformed<T>;
    fn scale(self, v: &glm::DVec3) -> Transformed<T>;
    fn rotate(self, angle: f64, axis: &glm::DVec3) -> Transformed<T>;
    fn rotate_x(self, angle: f64) -> Transformed<T>;
    fn rotate_y(self, angle: f64) -> Transformed<T>;
    fn rotate_z(self, angle: f64) -> Transformed<T>;
    fn transform(self, transform: glm::DMat4) -> Transformed<T>;
}

impl<T: Shape> Transformable<T> for T {
    fn translate(self, v: &glm::DVec3) -> Transformed<T> {
        Transformed::new(self, glm::translate(&glm::identity(), v))
    }
    fn scale(self, v: &glm::DVec3) -> Transformed<T> {
        Transformed::new(self, glm::scale(&glm::identity(), v))
    }
    fn rotate(self, angle: f64, axis: &glm::DVec3) -> Transformed<T> {
        Transformed::new(self, glm::rotate(&glm::identity(), angle, axis))
    }
    fn rotate_x(self, angle: f64) -> Transformed<T> {
        Transformed::new(self, glm::rotate_x(&glm::identity(), angle))
    }
    fn rotate_y(self, angle: f64) -> Transformed<T> {
        Transformed::new(self, glm::rotate_y(&glm::identity(), angle))
    }
    fn rotate_z(self, angle: f64) -> Transformed<T> {
        Transformed::new(self, glm::rotate_z(&glm::identity(), angle))
    }
    fn transform(self, transform: glm::DMat4) -> Transformed<T> {
        Transformed::new(self, transform)
    }
}

/// Chaining transforms keeps ONE matrix instead of nesting Transformed<Transformed<..>> (src/shape.rs:238-289)
impl<T: Shape> Transformed<T> {
    pub fn translate(self, v: &glm::DVec3) -> Transformed<T> {
        Self::new(self.shape, glm::translate(&glm::identity(), v) * self.transform)
    }
    pub fn scale(self, v: &glm::DVec3) -> Transformed<T> {
        Self::new(self.shape, glm::scale(&glm::identity(), v) * self.transform)
    }
    pub fn rotate(self, angle: f64, axis: &glm::DVec3) -> Transformed<T> {
        Self::new(self.shape, glm::rotate(&glm::identity(), angle, axis) * self.transform)
    }
    pub fn rotate_x(self, angle: f64) -> Transformed<T> {
        Self::new(self.shape, glm::rotate_x(&glm::identity(), angle) * self.transform)
    }
    pub fn rotate_y(self, angle: f64) -> Transformed<T> {
        Self::new(self.shape, glm::rotate_y(&glm::identity(), angle) * self.transform)
    }
    pub fn rotate_z(self, angle: f64) -> Transformed<T> {
        Self::new(self.shape, glm::rotate_z(&glm::identity(), angle) * self.transform)
    }
    pub fn transform(self, transform: glm::DMat4) -> Transformed<T> {
        Self::new(self.shape, transform * self.transform)
    }
}

/// Helper function to construct a sphere
pub fn sphere() -> Sphere {
    Sphere
}
/// Helper function to construct a glass-like monomial surface
pub fn monomial_surface(height: f64, exp: f64) -> MonomialSurface {
    MonomialSurface { height, exp }
}
/// Helper function to construct a plane
pub fn plane(normal: glm::DVec3, value: f64) -> Plane {
    Plane { normal, value }
}
/// Helper function to construct a cube
pub fn cube() -> Cube {
    Cube
}
/// Helper function to construct a simple polygon made from triangles (src/shape.rs:304-311)
pub fn polygon(verts: &[glm::DVec3]) -> Mesh {
    let mut tris = Vec::new();
    for i in 1..(verts.len() - 1) {
        tris.push(Triangle::from_vertices(verts[0], verts[i], verts[i + 1]));
    }
    Mesh::new(tris)
}
