//! src/camera.rs:9-63 of the reference (data + constructors; `cast_ray` runs on the device: `primary_ray`).
use rpt_cuda_sys as sys;

/// A simple thin-lens perspective camera
#[derive(Copy, Clone, Debug)]
pub struct Camera {
    pub eye: glm::DVec3,
    pub direction: glm::DVec3,
    /// Direction of "up" for screen, must be orthogonal to `direction`
    pub up: glm::DVec3,
    /// Field of view in the longer direction as an angle in radians, in (0, pi)
    pub fov: f64,
    pub aperture: f64,
    pub focal_distance: f64,
}

impl Default for Camera {
    fn default() -> Self {
        Self {
            eye: glm::vec3(0.0, 0.0, 10.0),
            direction: glm::vec3(0.0, 0.0, -1.0),
            up: glm::vec3(0.0, 1.0, 0.0),
            fov: std::f64::consts::FRAC_PI_6,
            aperture: 0.0,
            focal_distance: 0.0,
        }
    }
}

impl Camera {
    /// Perspective camera looking at a point, with a given field of view (src/camera.rs:43-54)
    pub fn look_at(eye: glm::DVec3, center: glm::DVec3, up: glm::DVec3, fov: f64) -> Self {
        let direction = (center - eye).normalize();
        let up = (up - up.dot(&direction) * direction).normalize();
        Self { eye, direction, up, fov, aperture: 0.0, focal_distance: 0.0 }
    }

    /// Focus the camera on a position, with simulated depth-of-field (src/camera.rs:57-61)
    pub fn focus(mut self, focal_point: glm::DVec3, aperture: f64) -> Self {
        self.focal_distance = (focal_point - self.eye).dot(&self.direction);
        self.aperture = aperture;
        self
    }

    pub(crate) fn lower(&self) -> sys::rptx_camera {
        sys::rptx_camera {
            eye: [self.eye.x, self.eye.y, self.eye.z],
            direction: [self.direction.x, self.direction.y, self.direction.z],
            up: [self.up.x, self.up.y, self.up.z],
            fov: self.fov,
            aperture: self.aperture,
            focal_distance: self.focal_distance,
        }
    }
}
