//! src/renderer.rs:23-156,416-425 and src/photon.rs:629-655 of the reference: the Renderer builder and its render
//! entry points.  Bodies: lower the scene, call the CUDA core.
use std::ptr;

use image::{ImageBuffer, RgbImage};
use rpt_cuda_sys as sys;

use crate::buffer::{Buffer, Filter};
use crate::camera::Camera;
use crate::ffi::{check, comm_info, context, Ctx};
use crate::lower::lower_scene;
use crate::object::Object;
use crate::scene::Scene;
use crate::shape::{HitRecord, Ray};

/// src/photon.rs:629-637
#[derive(Copy, Clone, Debug, PartialEq)]
pub enum PhotonRenderKind {
    PhotonMap,
    PhotonPointBeam,
    PhotonBeamBeam,
}

/// Builder object for rendering a scene
pub struct Renderer<'a> {
    pub scene: &'a Scene,
    pub camera: Camera,
    pub width: u32,
    pub height: u32,
    pub exposure_value: f64,
    pub filter: Filter,
    /// stored, never read — as in the reference (src/renderer.rs:42-43)
    pub stepsize: f64,
    pub max_bounces: u32,
    pub num_samples: u32,
    pub gather_size: usize,
    pub gather_size_volume: usize,
    pub watts: f64,
    /// addition: the reference seeds from OS entropy; runs here are reproducible
    pub seed: u64,
}

/// The lowered scene of one render call; destroyed with it.
struct DeviceScene {
    raw: *mut sys::rptx_scene,
}

impl Drop for DeviceScene {
    fn drop(&mut self) {
        unsafe { sys::rptx_scene_destroy(self.raw) }
    }
}

impl<'a> Renderer<'a> {
    /// Construct a new renderer for a scene (defaults of src/renderer.rs:60-76)
    pub fn new(scene: &'a Scene, camera: Camera) -> Self {
        Self {
            scene,
            camera,
            width: 800,
            height: 600,
            exposure_value: 0.0,
            filter: Filter::default(),
            stepsize: 0.0,
            max_bounces: 0,
            num_samples: 1,
            gather_size: 50,
            gather_size_volume: 50,
            watts: 100.,
            seed: 0,
        }
    }

    pub fn width(mut self, width: u32) -> Self {
        self.width = width;
        self
    }
    pub fn height(mut self, height: u32) -> Self {
        self.height = height;
        self
    }
    pub fn exposure_value(mut self, exposure_value: f64) -> Self {
        self.exposure_value = exposure_value;
        self
    }
    pub fn stepsize(mut self, stepsize: f64) -> Self {
        self.stepsize = stepsize;
        self
    }
    pub fn filter(mut self, filter: Filter) -> Self {
        self.filter = filter;
        self
    }
    pub fn max_bounces(mut self, max_bounces: u32) -> Self {
        self.max_bounces = max_bounces;
        self
    }
    pub fn num_samples(mut self, num_samples: u32) -> Self {
        self.num_samples = num_samples;
        self
    }
    pub fn gather_size(mut self, gather_size: usize) -> Self {
        self.gather_size = gather_size;
        self
    }
    pub fn gather_size_volume(mut self, gather_size_volume: usize) -> Self {
        self.gather_size_volume = gather_size_volume;
        self
    }
    pub fn watts(mut self, watts: f64) -> Self {
        self.watts = watts;
        self
    }
    pub fn seed(mut self, seed: u64) -> Self {
        self.seed = seed;
        self
    }

    fn params(&self, ctx: &Ctx) -> sys::rptx_params {
        let Filter::Box(radius) = self.filter;
        let (world, rank) = comm_info(ctx); // with a communicator every rank traces its sample shard
        sys::rptx_params {
            width: self.width,
            height: self.height,
            exposure_value: self.exposure_value,
            filter_radius: radius,
            max_bounces: self.max_bounces,
            num_samples: self.num_samples,
            _pad: 0,
            gather_size: self.gather_size as u64,
            gather_size_volume: self.gather_size_volume as u64,
            watts: self.watts,
            stepsize: self.stepsize,
            seed: self.seed,
            shard_count: world as u32,
            shard_index: rank as u32,
        }
    }

    fn begin(&self, ctx: &Ctx, scene: &DeviceScene) -> Buffer {
        let (cam, params) = (self.camera.lower(), self.params(ctx));
        let mut film: *mut sys::rptx_film = ptr::null_mut();
        check(ctx.raw, unsafe { sys::rptx_render_begin(ctx.raw, scene.raw, &cam, &params, &mut film) });
        Buffer { width: self.width, height: self.height, film, ctx: ctx.raw }
    }

    /// Render the scene by path tracing (src/renderer.rs:137-141)
    pub fn render(&self) -> RgbImage {
        let guard = context();
        let ctx = guard.as_ref().unwrap();
        let scene = DeviceScene { raw: lower_scene(ctx, self.scene) };
        let buffer = self.begin(ctx, &scene);
        self.sample(ctx, self.num_samples, &buffer);
        buffer.image()
    }

    /// Render the scene iteratively, calling a callback after every k samples (src/renderer.rs:144-156)
    pub fn iterative_render<F>(&self, callback_interval: u32, mut callback: F)
    where
        F: FnMut(u32, &Buffer),
    {
        let guard = context();
        let ctx = guard.as_ref().unwrap();
        let scene = DeviceScene { raw: lower_scene(ctx, self.scene) };
        let buffer = self.begin(ctx, &scene);
        let mut iteration = 0;
        while iteration < self.num_samples {
            let steps = std::cmp::min(self.num_samples - iteration, callback_interval);
            self.sample(ctx, steps, &buffer);
            iteration += steps;
            callback(iteration, &buffer);
        }
    }

    /// Renderer::sample (src/renderer.rs:158-171): one batch mean per pixel pushed into the buffer.  On a context
    /// with a communicator the library traces this rank's shard and all-reduces the batch sums before the commit.
    fn sample(&self, ctx: &Ctx, iterations: u32, buffer: &Buffer) {
        check(ctx.raw, unsafe { sys::rptx_render_sample(buffer.film, iterations) });
    }

    /// Loop through all objects in the scene to find the closest hit (src/renderer.rs:416-425)
    pub fn get_closest_hit(&self, ray: Ray) -> Option<(HitRecord, &'a Object)> {
        let guard = context();
        let ctx = guard.as_ref().unwrap();
        let scene = DeviceScene { raw: lower_scene(ctx, self.scene) };
        let (o, d) = ([ray.origin.x, ray.origin.y, ray.origin.z], [ray.dir.x, ray.dir.y, ray.dir.z]);
        let (mut t, mut object, mut normal) = (0.0f64, 0u32, [0.0f64; 3]);
        check(ctx.raw, unsafe {
            sys::rptx_intersect_batch(scene.raw, o.as_ptr(), d.as_ptr(), 1, 1e-12, &mut t, &mut object, ptr::null_mut(), ptr::null_mut(), normal.as_mut_ptr())
        });
        if object == u32::MAX {
            return None;
        }
        Some((HitRecord { time: t, normal: glm::vec3(normal[0], normal[1], normal[2]) }, &self.scene.objects[object as usize]))
    }

    /// src/photon.rs:642-653
    pub fn photon_point_query_beam_render(&self, photon_count: usize) -> RgbImage {
        self.photon_render(photon_count, PhotonRenderKind::PhotonPointBeam)
    }
    pub fn photon_beam_query_beam_render(&self, photon_count: usize) -> RgbImage {
        self.photon_render(photon_count, PhotonRenderKind::PhotonBeamBeam)
    }
    pub fn photon_map_render(&self, photon_count: usize) -> RgbImage {
        self.photon_render(photon_count, PhotonRenderKind::PhotonMap)
    }

    /// photon_render (src/photon.rs:655-720): emit -> merge -> build -> gather, `num_samples` camera samples per pixel.
    /// Panics "Only found non-object lights while photon mapping" like the reference (:798).
    pub fn photon_render(&self, photon_count: usize, kind: PhotonRenderKind) -> RgbImage {
        let guard = context();
        let ctx = guard.as_ref().unwrap();
        let scene = DeviceScene { raw: lower_scene(ctx, self.scene) };
        let (cam, params) = (self.camera.lower(), self.params(ctx));
        let kind = match kind {
            PhotonRenderKind::PhotonMap => sys::RPTX_PHOTON_MAP,
            PhotonRenderKind::PhotonPointBeam => sys::RPTX_PHOTON_POINT_BEAM,
            PhotonRenderKind::PhotonBeamBeam => sys::RPTX_PHOTON_BEAM_BEAM,
        };
        let mut buf = vec![0u8; (self.width * self.height * 3) as usize];
        check(ctx.raw, unsafe {
            sys::rptx_photon_render(ctx.raw, scene.raw, &cam, &params, photon_count as u64, kind, buf.as_mut_ptr(), ptr::null_mut())
        });
        ImageBuffer::from_raw(self.width, self.height, buf).expect("Image buffer has incorrect size")
    }
}
