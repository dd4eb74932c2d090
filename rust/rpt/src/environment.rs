//! src/environment.rs:5-77 of the reference (lookup runs on the device: `env_color`).
use image::{ImageBuffer, Rgb};

use crate::color::Color;

/// A high-dynamic-range equirectangular image for lighting 3D scenes
pub struct Hdri {
    pub(crate) width: u32,
    pub(crate) height: u32,
    /// [height][width][3] linear f64, the layout rptx_scene_set_environment_hdri takes
    pub(crate) pixels: Vec<f64>,
}

impl Hdri {
    /// Construct a new HDRI from an image buffer (src/environment.rs:13-22)
    pub fn new(buf: &ImageBuffer<Rgb<f32>, Vec<f32>>) -> Self {
        let (width, height) = buf.dimensions();
        let pixels = buf.pixels().flat_map(|p| vec![p[0] as f64, p[1] as f64, p[2] as f64]).collect();
        Self { width, height, pixels }
    }
}

/// An environment map for lighting 3D scenes
pub enum Environment {
    /// Solid-color environment lighting
    Color(Color),
    /// High-dynamic-range image environment lighting
    Hdri(Hdri),
}

impl Default for Environment {
    fn default() -> Self {
        Environment::Color(glm::vec3(0.0, 0.0, 0.0))
    }
}
