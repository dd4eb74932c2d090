//! src/kdtree.rs:12-122 of the reference: Bounded, BoundingBox and the KdTree container.  The reference's
//! median-split build and recursive traversal (`construct`, `intersect_subtree`) are replaced by the device BVH;
//! a `KdTree<T>` here only groups bounded shapes so that they lower to ONE object (RPTX_SHAPE_GROUP), which keeps
//! the reference's object indices, per-object materials and tie-breaking order.
use std::sync::Arc;

use crate::shape::Shape;

/// A geometric shape with a bounding box
pub trait Bounded: Shape {
    fn bounding_box(&self) -> BoundingBox;
}

impl<T: Bounded + ?Sized> Bounded for Box<T> {
    fn bounding_box(&self) -> BoundingBox {
        self.as_ref().bounding_box()
    }
}

impl<T: Bounded + ?Sized> Bounded for Arc<T> {
    fn bounding_box(&self) -> BoundingBox {
        self.as_ref().bounding_box()
    }
}

/// A 3D axis-aligned bounding box (src/kdtree.rs:31-101)
#[derive(Copy, Clone, Debug)]
pub struct BoundingBox {
    pub p_min: glm::DVec3,
    pub p_max: glm::DVec3,
}

impl Default for BoundingBox {
    fn default() -> Self {
        Self {
            p_min: glm::vec3(f64::INFINITY, f64::INFINITY, f64::INFINITY),
            p_max: glm::vec3(f64::NEG_INFINITY, f64::NEG_INFINITY, f64::NEG_INFINITY),
        }
    }
}

impl BoundingBox {
    pub fn merge(&self, other: &BoundingBox) -> BoundingBox {
        BoundingBox { p_min: glm::min2(&self.p_min, &other.p_min), p_max: glm::max2(&self.p_max, &other.p_max) }
    }
}

/// A collection of bounded shapes behind one object (src/kdtree.rs:103-122)
#[derive(Clone)]
pub struct KdTree<T> {
    pub(crate) objects: Vec<T>,
    bounds: BoundingBox,
}

impl<T: Bounded> KdTree<T> {
    /// Construct a new kd-tree from a collection of objects
    pub fn new(objects: Vec<T>) -> Self {
        let bounds = objects.iter().map(T::bounding_box).fold(BoundingBox::default(), |b1, b2| b1.merge(&b2));
        Self { objects, bounds }
    }
}

impl<T: Bounded> Bounded for KdTree<T> {
    fn bounding_box(&self) -> BoundingBox {
        self.bounds
    }
}
