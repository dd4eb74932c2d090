//! src/medium.rs:80-121 of the reference.  The reference's Medium is six boxed closures behind private fields and
//! exactly two public constructors; the closed set {homogeneous_isotropic, colored_glowing_fog} is what the device
//! implements (`MediumDev`), so the mirror stores the constructor that was used.
use rpt_cuda_sys as sys;

#[derive(Copy, Clone, Debug)]
pub(crate) enum MediumKind {
    HomogeneousIsotropic,
    ColoredGlowingFog,
}

/// Represents volumetric media
#[derive(Copy, Clone, Debug)]
pub struct Medium {
    pub(crate) kind: MediumKind,
    pub(crate) absorption: f64,
    pub(crate) scattering: f64,
}

impl Medium {
    /// Create a default homogeneous isotropic medium, i.e. uniform fog (src/medium.rs:80-96)
    pub fn homogeneous_isotropic(absorption: f64, scattering: f64) -> Medium {
        Medium { kind: MediumKind::HomogeneousIsotropic, absorption, scattering }
    }

    /// Create a colored, emissive homogeneous isotropic medium (src/medium.rs:99-121)
    pub fn colored_glowing_fog(absorption: f64, scattering: f64) -> Medium {
        Medium { kind: MediumKind::ColoredGlowingFog, absorption, scattering }
    }

    pub(crate) fn lower_kind(&self) -> i32 {
        match self.kind {
            MediumKind::HomogeneousIsotropic => sys::RPTX_MEDIUM_HOMOGENEOUS_ISOTROPIC,
            MediumKind::ColoredGlowingFog => sys::RPTX_MEDIUM_COLORED_GLOWING_FOG,
        }
    }
}
