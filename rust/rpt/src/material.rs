//! src/material.rs:8-98 of the reference: the enum and its constructors.  sample_f / bsdf run on the device
//! (`mat_sample_f`, `mat_bsdf`).
use rpt_cuda_sys as sys;

use crate::color::Color;

#[derive(Copy, Clone)]
pub enum Material {
    Lambertian { albedo: Color, emittance: f64 },
    Phong { albedo: Color, shininess: f64, emittance: f64 },
    Mirror,
    Transmissive { albedo: Color, ior: f64 },
}

impl Default for Material {
    fn default() -> Self {
        Self::Lambertian { albedo: glm::vec3(0.5, 0.5, 0.5), emittance: 0. }
    }
}

impl Material {
    pub fn diffuse(color: Color) -> Material {
        Self::Lambertian { albedo: color, emittance: 0. }
    }
    /// `roughness` is stored as the Phong exponent, as in the reference (src/material.rs:44-50)
    pub fn specular(color: Color, roughness: f64) -> Material {
        Self::Phong { albedo: color, shininess: roughness, emittance: 0. }
    }
    pub fn mirror() -> Material {
        Self::Mirror
    }
    pub fn transmissive(ior: f64) -> Material {
        Self::Transmissive { ior, albedo: Color::new(0., 0., 0.) }
    }
    pub fn clear(index: f64, _roughness: f64) -> Self {
        Self::Transmissive { albedo: glm::vec3(0., 0., 0.), ior: index }
    }
    pub fn transparent(color: Color, index: f64, _roughness: f64) -> Self {
        Self::Transmissive { albedo: color, ior: index }
    }
    pub fn metallic(color: Color, roughness: f64) -> Self {
        Self::Phong { albedo: color, shininess: roughness, emittance: 0. }
    }
    pub fn light(color: Color, emittance: f64) -> Self {
        Self::Lambertian { albedo: color, emittance }
    }

    pub(crate) fn lower(&self) -> sys::rptx_material {
        let mut m = sys::rptx_material { kind: 0, _pad: 0, albedo: [0.0; 3], emittance: 0.0, shininess: 0.0, ior: 1.0 };
        match *self {
            Material::Lambertian { albedo, emittance } => {
                m.kind = sys::RPTX_MAT_LAMBERTIAN;
                m.albedo = [albedo.x, albedo.y, albedo.z];
                m.emittance = emittance;
            }
            Material::Phong { albedo, shininess, emittance } => {
                m.kind = sys::RPTX_MAT_PHONG;
                m.albedo = [albedo.x, albedo.y, albedo.z];
                m.shininess = shininess;
                m.emittance = emittance;
            }
            Material::Mirror => m.kind = sys::RPTX_MAT_MIRROR,
            Material::Transmissive { albedo, ior } => {
                m.kind = sys::RPTX_MAT_TRANSMISSIVE;
                m.albedo = [albedo.x, albedo.y, albedo.z];
                m.ior = ior;
            }
        }
        m
    }
}
