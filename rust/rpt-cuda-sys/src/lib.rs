//! `extern "C"` declarations of `include/rptx.h` (ABI version 3), one to one.
//!
//! The reference (neevparikh/rpt) has no FFI; its boundary is the Rust API of `src/renderer.rs` /
//! `src/photon.rs`.  A drop-in `rpt` mirror keeps those types and lowers them into the POD records
//! below.  All reals are f64 as in the reference (`src/color.rs:2`).
#![allow(non_camel_case_types)]

use std::os::raw::{c_char, c_int, c_void};

pub const RPTX_ABI_VERSION: c_int = 3;

// shapes: src/shape.rs:19-26 (trait Shape), :102-109 (Transformed<T>)
pub const RPTX_SHAPE_SPHERE: i32 = 0;
pub const RPTX_SHAPE_PLANE: i32 = 1;
pub const RPTX_SHAPE_CUBE: i32 = 2;
pub const RPTX_SHAPE_MESH: i32 = 3;
pub const RPTX_SHAPE_GROUP: i32 = 4;
pub const RPTX_SHAPE_MONOMIAL: i32 = 5;
// materials: src/material.rs:8-23
pub const RPTX_MAT_LAMBERTIAN: i32 = 0;
pub const RPTX_MAT_PHONG: i32 = 1;
pub const RPTX_MAT_MIRROR: i32 = 2;
pub const RPTX_MAT_TRANSMISSIVE: i32 = 3;
// lights: src/light.rs:7-19
pub const RPTX_LIGHT_POINT: i32 = 0;
pub const RPTX_LIGHT_AMBIENT: i32 = 1;
pub const RPTX_LIGHT_DIRECTIONAL: i32 = 2;
pub const RPTX_LIGHT_OBJECT: i32 = 3;
// media: src/medium.rs:80-121
pub const RPTX_MEDIUM_NONE: i32 = 0;
pub const RPTX_MEDIUM_HOMOGENEOUS_ISOTROPIC: i32 = 1;
pub const RPTX_MEDIUM_COLORED_GLOWING_FOG: i32 = 2;
// PhotonRenderKind: src/photon.rs:629-637
pub const RPTX_PHOTON_MAP: i32 = 0;
pub const RPTX_PHOTON_POINT_BEAM: i32 = 1;
pub const RPTX_PHOTON_BEAM_BEAM: i32 = 2;
/// One photon record = position, direction, power, starting_position (f64 x 3 each) + u64 key.
pub const RPTX_PHOTON_ROW_BYTES: usize = 104;
/// Size of the communicator id rank 0 hands to the other ranks (an ncclUniqueId).
pub const RPTX_INTERSECT_STATIC_MAX_RAYS: usize = 65536;
pub const RPTX_COMM_ID_BYTES: usize = 128;

#[repr(C)] pub struct rptx_ctx { _p: [u8; 0] }
#[repr(C)] pub struct rptx_scene { _p: [u8; 0] }
#[repr(C)] pub struct rptx_film { _p: [u8; 0] }
#[repr(C)] pub struct rptx_photons { _p: [u8; 0] }
#[repr(C)] pub struct rptx_photon_map { _p: [u8; 0] }

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rptx_shape_desc {
    pub kind: i32,
    pub has_transform: i32,
    pub transform: [f64; 16], // row-major 4x4
    pub mesh_id: u32,
    pub n_children: u32,
    pub plane_normal: [f64; 3],
    pub plane_value: f64,
    pub mono_height: f64,
    pub mono_exp: f64,
    pub children: *const rptx_shape_desc,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rptx_material {
    pub kind: i32,
    pub _pad: i32,
    pub albedo: [f64; 3],
    pub emittance: f64,
    pub shininess: f64,
    pub ior: f64,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rptx_light {
    pub kind: i32,
    pub _pad: i32,
    pub color: [f64; 3],
    pub vec: [f64; 3],
    pub shape: rptx_shape_desc,
    pub material: rptx_material,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rptx_camera {
    pub eye: [f64; 3],
    pub direction: [f64; 3],
    pub up: [f64; 3],
    pub fov: f64,
    pub aperture: f64,
    pub focal_distance: f64,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct rptx_params {
    pub width: u32,
    pub height: u32,
    pub exposure_value: f64,
    pub filter_radius: u32,
    pub max_bounces: u32,
    pub num_samples: u32,
    pub _pad: u32,
    pub gather_size: u64,
    pub gather_size_volume: u64,
    pub watts: f64,
    pub stepsize: f64,
    pub seed: u64,
    pub shard_count: u32,
    pub shard_index: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct rptx_stats {
    pub path_samples: u64,
    pub extend_rays: u64,
    pub shadow_rays: u64,
    pub shadow_traced: u64,
    pub shadow_occlusion: u64,
    pub waves: u64,
    pub kernel_launches: u64,
    pub device_ms: f64,
    pub trace_ms: f64,
    pub extend_ms: f64,
    pub shade_ms: f64,
    pub connect_ms: f64,
    pub shadow_ms: f64,
    pub photons_emitted: u64,
    pub photons_surface: u64,
    pub photons_volume: u64,
    pub gather_queries: u64,
    pub gather_visibility: u64,
    pub gather_candidates: u64,
    pub path_mode: u64,
}

extern "C" {
    pub fn rptx_abi_version() -> c_int;

    // context = one CUDA device + its stream (one process per GPU)
    pub fn rptx_ctx_create(device_id: c_int, out: *mut *mut rptx_ctx) -> c_int;
    pub fn rptx_ctx_destroy(ctx: *mut rptx_ctx);
    pub fn rptx_last_error(ctx: *const rptx_ctx) -> *const c_char;
    pub fn rptx_ctx_trim(ctx: *mut rptx_ctx) -> c_int;

    // multi-GPU: one context per rank joined into an NCCL communicator; the exchanges replace `add_samples`
    // (src/renderer.rs:169-170) and PhotonList::merge (src/photon.rs:153-160,675)
    pub fn rptx_comm_unique_id(id_out: *mut c_void) -> c_int;
    pub fn rptx_comm_init(ctx: *mut rptx_ctx, id: *const c_void, world: c_int, rank: c_int) -> c_int;
    pub fn rptx_comm_info(ctx: *const rptx_ctx, world: *mut c_int, rank: *mut c_int) -> c_int;
    pub fn rptx_comm_destroy(ctx: *mut rptx_ctx);

    // Scene, src/scene.rs:12-81
    pub fn rptx_scene_create(ctx: *mut rptx_ctx, out: *mut *mut rptx_scene) -> c_int;
    pub fn rptx_scene_destroy(scene: *mut rptx_scene);
    pub fn rptx_scene_add_mesh(scene: *mut rptx_scene, vertices: *const f64, normals: *const f64, ntri: usize, mesh_id: *mut u32) -> c_int;
    pub fn rptx_scene_add_object(scene: *mut rptx_scene, shape: *const rptx_shape_desc, material: *const rptx_material, object_id: *mut u32) -> c_int;
    pub fn rptx_scene_add_light(scene: *mut rptx_scene, light: *const rptx_light) -> c_int;
    pub fn rptx_scene_set_medium(scene: *mut rptx_scene, kind: i32, absorption: f64, scattering: f64) -> c_int;
    pub fn rptx_scene_set_environment_color(scene: *mut rptx_scene, rgb: *const f64) -> c_int;
    pub fn rptx_scene_set_environment_hdri(scene: *mut rptx_scene, width: u32, height: u32, pixels: *const f64) -> c_int;
    pub fn rptx_scene_commit(scene: *mut rptx_scene) -> c_int;

    // Renderer::render / iterative_render / sample, Buffer (src/renderer.rs:137-171, src/buffer.rs)
    pub fn rptx_render_begin(ctx: *mut rptx_ctx, scene: *mut rptx_scene, camera: *const rptx_camera, params: *const rptx_params, out: *mut *mut rptx_film) -> c_int;
    pub fn rptx_film_destroy(film: *mut rptx_film);
    pub fn rptx_render_sample(film: *mut rptx_film, iterations: u32) -> c_int;
    pub fn rptx_render_sample_trace(film: *mut rptx_film, iterations: u32) -> c_int;
    pub fn rptx_render_sample_commit(film: *mut rptx_film, iterations: u32) -> c_int;
    pub fn rptx_render_sample_allreduce(film: *mut rptx_film) -> c_int;
    pub fn rptx_film_batch_sum_device_ptr(film: *mut rptx_film) -> *mut c_void;
    pub fn rptx_film_image_rgb8(film: *mut rptx_film, out_rgb: *mut u8) -> c_int;
    pub fn rptx_film_variance(film: *mut rptx_film, out: *mut f64) -> c_int;
    pub fn rptx_film_mean(film: *mut rptx_film, out: *mut f64) -> c_int;

    // photon integrators, src/photon.rs:655-720 pass by pass
    pub fn rptx_photons_emit(ctx: *mut rptx_ctx, scene: *mut rptx_scene, params: *const rptx_params, photon_count_total: u64, first: u64, count: u64, kind: i32, out: *mut *mut rptx_photons) -> c_int;
    pub fn rptx_photons_create(ctx: *mut rptx_ctx, out: *mut *mut rptx_photons) -> c_int;
    pub fn rptx_photons_destroy(photons: *mut rptx_photons);
    pub fn rptx_photons_append_host(photons: *mut rptx_photons, which: i32, rows: *const c_void, n: usize) -> c_int;
    pub fn rptx_photons_append_device(photons: *mut rptx_photons, which: i32, d_rows: *const c_void, n: usize) -> c_int;
    pub fn rptx_photons_allgather(local: *mut rptx_photons, out: *mut *mut rptx_photons) -> c_int;
    pub fn rptx_photons_count(photons: *const rptx_photons, counts: *mut u64) -> c_int;
    pub fn rptx_photons_device_ptr(photons: *mut rptx_photons, which: i32) -> *mut c_void;
    pub fn rptx_photons_download(photons: *mut rptx_photons, which: i32, rows: *mut c_void) -> c_int;
    pub fn rptx_photons_stats(photons: *const rptx_photons, emitted: *mut u64, rays: *mut u64, emit_ms: *mut f64) -> c_int;
    pub fn rptx_photon_map_build(ctx: *mut rptx_ctx, scene: *mut rptx_scene, photons: *mut rptx_photons, kind: i32, out: *mut *mut rptx_photon_map) -> c_int;
    pub fn rptx_photon_map_build_sharded(ctx: *mut rptx_ctx, scene: *mut rptx_scene, local_photons: *mut rptx_photons, kind: i32, out: *mut *mut rptx_photon_map) -> c_int;
    pub fn rptx_photon_map_destroy(map: *mut rptx_photon_map);
    pub fn rptx_photon_map_stats(map: *const rptx_photon_map, counts: *mut u64, build_ms: *mut f64) -> c_int;
    pub fn rptx_photon_gather_begin(ctx: *mut rptx_ctx, scene: *mut rptx_scene, camera: *const rptx_camera, params: *const rptx_params, out: *mut *mut rptx_film) -> c_int;
    pub fn rptx_photon_gather_trace(film: *mut rptx_film, map: *mut rptx_photon_map, iterations: u32) -> c_int;
    pub fn rptx_photon_render(ctx: *mut rptx_ctx, scene: *mut rptx_scene, camera: *const rptx_camera, params: *const rptx_params, photon_count: u64, kind: i32, out_rgb8: *mut u8, out_mean: *mut f64) -> c_int;

    // parity hooks
    pub fn rptx_photon_map_knn(map: *mut rptx_photon_map, which: i32, q: *const f64, n: usize, k: u32, keys: *mut u64, d2: *mut f64, counts: *mut u32) -> c_int;
    pub fn rptx_photon_map_spheres(map: *mut rptx_photon_map, center: *mut f64, radius: *mut f64) -> c_int;
    pub fn rptx_intersect_batch(scene: *mut rptx_scene, o: *const f64, d: *const f64, n: usize, t_min: f64, t: *mut f64, object_id: *mut u32, prim_id: *mut u32, inst_id: *mut u32, normal: *mut f64) -> c_int;
    pub fn rptx_intersect_batch_device(scene: *mut rptx_scene, d_o: *const c_void, d_d: *const c_void, n: usize, t_min: f64, d_t: *mut c_void, d_object_id: *mut c_void, d_prim_id: *mut c_void, d_inst_id: *mut c_void, d_normal: *mut c_void) -> c_int;
    pub fn rptx_stats_get(film: *const rptx_film, out: *mut rptx_stats) -> c_int;
    /// allocation-free rptx_intersect_batch (static device scratch, n <= RPTX_INTERSECT_STATIC_MAX_RAYS); no counterpart in rpt
    pub fn rptx_intersect_batch_static(scene: *mut rptx_scene, o: *const f64, d: *const f64, n: usize, t_min: f64, t: *mut f64, object_id: *mut u32, prim_id: *mut u32, inst_id: *mut u32, normal: *mut f64, n_hits: *mut u64) -> c_int;
}
