//! src/color.rs of the reference, unchanged in meaning (host-side helpers; the device encodes images itself).
pub type Color = glm::DVec3;

const SRGB_GAMMA: f64 = 2.2;

/// Construct a linear color from a gamma-encoded hex code (src/color.rs:10-15)
pub fn hex_color(x: u32) -> Color {
    let r = ((x >> 16) & 0xff) as f64 / 255.0;
    let g = ((x >> 8) & 0xff) as f64 / 255.0;
    let b = (x & 0xff) as f64 / 255.0;
    glm::vec3(r.powf(SRGB_GAMMA), g.powf(SRGB_GAMMA), b.powf(SRGB_GAMMA))
}

/// Gamma-encode a linear color into bytes (src/color.rs:18-24)
pub fn color_bytes(color: &Color) -> [u8; 3] {
    let c = color.map(|x| (x.max(0.0).min(1.0).powf(1.0 / SRGB_GAMMA) * 255.0) as u8);
    [c.x, c.y, c.z]
}
