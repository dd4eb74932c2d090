//! src/light.rs:7-19 of the reference (`illuminate` runs on the device).
use crate::color::Color;
use crate::object::Object;

/// Type representing various forms of lighting
pub enum Light {
    /// Point light represented as (color, location)
    Point(Color, glm::DVec3),
    /// Ambient light represented as (color)
    Ambient(Color),
    /// Directional light represented as (color, direction)
    Directional(Color, glm::DVec3),
    /// Light from an invisible, emissive object
    Object(Object),
}
