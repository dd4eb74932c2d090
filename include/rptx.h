/* rptx.h — C ABI of the B200 light-transport core for neevparikh/rpt.
 *
 * The reference has no FFI: its boundary is the Rust library API
 * (Renderer builder src/renderer.rs:23-134, render()/iterative_render() :137-156,
 * photon_render src/photon.rs:642-655, get_closest_hit src/renderer.rs:416-425,
 * Scene/SceneAdd src/scene.rs:12-81, Object src/object.rs:10-31, Light src/light.rs:7-19,
 * Material src/material.rs:8-98, Medium src/medium.rs:80-121, Camera src/camera.rs:9-27).
 * A Rust `rpt` mirror lowers those types into the POD records below and calls
 * these entry points (binding shown in INTEGRATION.md).  Everything is plain
 * pointers and sizes; all reals are f64 as in the reference (src/color.rs:2).
 *
 * Every function returns 0 on success, non-zero on error; the message is
 * available from rptx_last_error().  Nothing aborts across the boundary
 * (the reference panics: src/buffer.rs:26,33,55,89, src/photon.rs:798).
 *
 * Threading: one host thread per context at a time; one context per device
 * (one process per GPU).  Contexts are independent.
 */
#ifndef RPTX_H
#define RPTX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RPTX_ABI_VERSION 3

/* ---- shapes: src/shape.rs:19-26 (trait Shape), :102-109 (Transformed<T>) ---- */
enum {
    RPTX_SHAPE_SPHERE = 0, /* unit sphere, src/shape/sphere.rs:10 */
    RPTX_SHAPE_PLANE  = 1, /* x.normal = value, src/shape/plane.rs:7-13 */
    RPTX_SHAPE_CUBE   = 2, /* unit cube [-.5,.5]^3, src/shape/cube.rs:9 */
    RPTX_SHAPE_MESH   = 3, /* Mesh = KdTree<Triangle>, src/shape/mesh.rs:103 */
    RPTX_SHAPE_GROUP  = 4, /* KdTree<Box<dyn Bounded>> of transformed children, src/kdtree.rs:103 */
    RPTX_SHAPE_MONOMIAL = 5 /* y = height * (x^2 + z^2)^2, x^2 + z^2 <= 1, src/shape/monomial_surface.rs:8-19 */
};

typedef struct rptx_shape_desc {
    int32_t  kind;
    int32_t  has_transform;     /* 1: wrapped in Transformed<T> with `transform` */
    double   transform[16];     /* row-major 4x4, M[r][c] = transform[4*r+c] */
    uint32_t mesh_id;           /* MESH: id returned by rptx_scene_add_mesh */
    uint32_t n_children;        /* GROUP */
    double   plane_normal[3];   /* PLANE */
    double   plane_value;       /* PLANE */
    double   mono_height;       /* MONOMIAL: MonomialSurface::height */
    double   mono_exp;          /* MONOMIAL: MonomialSurface::exp (must be 4: the reference's intersect hard-codes it) */
    const struct rptx_shape_desc* children; /* GROUP: n_children leaf shapes (SPHERE/CUBE/MESH/MONOMIAL) */
} rptx_shape_desc;

/* ---- materials: src/material.rs:8-23 ---- */
enum {
    RPTX_MAT_LAMBERTIAN   = 0,
    RPTX_MAT_PHONG        = 1,
    RPTX_MAT_MIRROR       = 2,
    RPTX_MAT_TRANSMISSIVE = 3
};

typedef struct rptx_material {
    int32_t kind;
    int32_t _pad;
    double  albedo[3];
    double  emittance;
    double  shininess;  /* Phong exponent (the reference stores "roughness" here, src/material.rs:44-50) */
    double  ior;
} rptx_material;

/* ---- lights: src/light.rs:7-19 ---- */
enum {
    RPTX_LIGHT_POINT       = 0, /* color, vec = location */
    RPTX_LIGHT_AMBIENT     = 1, /* color */
    RPTX_LIGHT_DIRECTIONAL = 2, /* color, vec = direction */
    RPTX_LIGHT_OBJECT      = 3  /* shape + material (emissive object, invisible unless also added as object) */
};

typedef struct rptx_light {
    int32_t         kind;
    int32_t         _pad;
    double          color[3];
    double          vec[3];
    rptx_shape_desc shape;
    rptx_material   material;
} rptx_light;

/* ---- media: src/medium.rs:80-121 (the only two constructors) ---- */
enum {
    RPTX_MEDIUM_NONE                  = 0,
    RPTX_MEDIUM_HOMOGENEOUS_ISOTROPIC = 1,
    RPTX_MEDIUM_COLORED_GLOWING_FOG   = 2
};

/* ---- camera: src/camera.rs:9-27 ---- */
typedef struct rptx_camera {
    double eye[3];
    double direction[3];
    double up[3];
    double fov;
    double aperture;
    double focal_distance;
} rptx_camera;

/* ---- Renderer builder fields: src/renderer.rs:23-56, defaults :60-76 ---- */
typedef struct rptx_params {
    uint32_t width, height;
    double   exposure_value;
    uint32_t filter_radius;       /* Filter::Box(r), src/buffer.rs:98-108 */
    uint32_t max_bounces;
    uint32_t num_samples;
    uint32_t _pad;
    uint64_t gather_size;         /* k of the surface estimate; limits of this library: 1..1024 (the per-sample heap is a */
    uint64_t gather_size_volume;  /* compile-time local array: 32 / 128 / 1024 variants); above 1024 the gather returns an error */
    double   watts;
    double   stepsize;            /* stored, never read (src/renderer.rs:42-43) */
    uint64_t seed;                /* the reference seeds from OS entropy; here runs are reproducible */
    /* Multi-GPU sample sharding: this rank renders sample indices
     * {s : s % shard_count == shard_index} of every batch.  (1,0) = all. */
    uint32_t shard_count, shard_index;
} rptx_params;

enum { /* PhotonRenderKind, src/photon.rs:629-637 */
    RPTX_PHOTON_MAP        = 0,
    RPTX_PHOTON_POINT_BEAM = 1,
    RPTX_PHOTON_BEAM_BEAM  = 2
};

typedef struct rptx_ctx   rptx_ctx;
typedef struct rptx_scene rptx_scene;
typedef struct rptx_film  rptx_film;

typedef struct rptx_stats {
    uint64_t path_samples;     /* camera paths started */
    uint64_t extend_rays;      /* closest-hit queries for path segments */
    uint64_t shadow_rays;      /* NEE visibility queries (one get_closest_hit each in the reference) */
    uint64_t shadow_traced;    /* of which probed on the device (non-zero contribution) */
    uint64_t shadow_occlusion; /* of which also needed the occlusion-segment search */
    uint64_t waves;            /* wavefront iterations */
    uint64_t kernel_launches;  /* kernels of this library launched */
    double   device_ms;        /* CUDA-event time of the last render_sample call */
    double   trace_ms;         /* of which: extend + shadow kernels (all ray queries) */
    double   extend_ms, shade_ms, connect_ms, shadow_ms; /* per-kernel CUDA-event sums of the last call */
    uint64_t photons_emitted, photons_surface, photons_volume;
    uint64_t gather_queries;   /* kNN queries of the photon camera pass */
    uint64_t gather_visibility; /* visibility rays of surface estimates */
    uint64_t gather_candidates; /* sphere/beam candidates returned by the BVH */
    uint64_t path_mode;        /* 0 wavefront kernels, 1 fused per-thread kernel, 2 photon gather film */
} rptx_stats;

int         rptx_abi_version(void);

/* context = one CUDA device + its streams */
int         rptx_ctx_create(int device_id, rptx_ctx** out);
void        rptx_ctx_destroy(rptx_ctx* ctx);
const char* rptx_last_error(const rptx_ctx* ctx);
/* Release the device memory the context keeps cached between calls (stream-ordered pool blocks, the spare film slab).  No
 * counterpart in the reference (its allocator is the process heap); for hosts that switch workloads. */
int         rptx_ctx_trim(rptx_ctx* ctx);

/* ---- multi-GPU: one process (or host thread) per GPU, one context each ------------------------------
 * The reference's two exchange points are `add_samples` after the rows' collect() (src/renderer.rs:169-170)
 * and PhotonList::merge (src/photon.rs:153-160,675).  rptx_comm_init joins this rank's context into an NCCL
 * communicator (libnccl.so.2 is loaded on first use; NVLink / NVSwitch inside a box).  On a context with a
 * communicator
 *   - rptx_render_sample on a film whose (shard_count, shard_index) equal (world, rank) traces this rank's
 *     sample shard, all-reduces the f64 batch sums and commits the same batch mean on every rank;
 *   - rptx_photon_render emits this rank's photon-index range, all-gathers the lists, builds the map on every
 *     rank, gathers this rank's sample shard of the camera pass and all-reduces it.
 * Rank 0 makes the id with rptx_comm_unique_id and hands its 128 bytes to the other ranks by whatever channel
 * the host has (the Python host uses torch.distributed's store); every rank then calls rptx_comm_init. */
#define RPTX_COMM_ID_BYTES 128
int  rptx_comm_unique_id(void* id_out /* RPTX_COMM_ID_BYTES */);
int  rptx_comm_init(rptx_ctx* ctx, const void* id, int world, int rank);
int  rptx_comm_info(const rptx_ctx* ctx, int* world, int* rank);   /* (1, 0) without a communicator */
void rptx_comm_destroy(rptx_ctx* ctx);                             /* before rptx_ctx_destroy */

/* Scene (src/scene.rs:12-24).  Caller-owned arrays are copied. */
int  rptx_scene_create(rptx_ctx* ctx, rptx_scene** out);
void rptx_scene_destroy(rptx_scene* scene);
/* triangles as [ntri][3 vertices][xyz]; normals same layout or NULL (=> face normals,
 * Triangle::from_vertices src/shape/mesh.rs:27-38) */
int  rptx_scene_add_mesh(rptx_scene* scene, const double* vertices, const double* normals,
                         size_t ntri, uint32_t* mesh_id);
/* SceneAdd<Object> src/scene.rs:40-44 */
int  rptx_scene_add_object(rptx_scene* scene, const rptx_shape_desc* shape,
                           const rptx_material* material, uint32_t* object_id);
/* SceneAdd<Light> src/scene.rs:46-50 */
int  rptx_scene_add_light(rptx_scene* scene, const rptx_light* light);
/* SceneAdd<Medium> src/scene.rs:77-81; only media[0] is ever used (src/renderer.rs:190) */
int  rptx_scene_set_medium(rptx_scene* scene, int32_t kind, double absorption, double scattering);
/* Environment::Color src/environment.rs:56-68 */
int  rptx_scene_set_environment_color(rptx_scene* scene, const double rgb[3]);
/* Environment::Hdri src/environment.rs:5-52; pixels as [height][width][3] f64 */
int  rptx_scene_set_environment_hdri(rptx_scene* scene, uint32_t width, uint32_t height,
                                     const double* pixels);
/* flatten, upload, build the BVHs on the device */
int  rptx_scene_commit(rptx_scene* scene);

/* Renderer::new(...).width(..)... ; Buffer::new (src/renderer.rs:138,148) */
int  rptx_render_begin(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera,
                       const rptx_params* params, rptx_film** out);
void rptx_film_destroy(rptx_film* film);
/* Renderer::sample(iterations, &mut buffer) src/renderer.rs:158-171: one batch = one
 * mean colour per pixel pushed into the Buffer. */
int  rptx_render_sample(rptx_film* film, uint32_t iterations);
/* Split form used for multi-GPU: trace this rank's shard of the batch into the batch-sum
 * buffer (device, f64 [h][w][3]); the caller may all-reduce that buffer (NCCL) between the
 * two calls; commit divides by `iterations`, applies 2^EV and pushes the batch mean. */
int  rptx_render_sample_trace(rptx_film* film, uint32_t iterations);
int  rptx_render_sample_commit(rptx_film* film, uint32_t iterations);
/* the exchange between the two, inside the library: NCCL all-reduce (sum) of the batch-sum film over the
 * context's communicator, enqueued on the library's stream (no host synchronisation) */
int  rptx_render_sample_allreduce(rptx_film* film);
void* rptx_film_batch_sum_device_ptr(rptx_film* film);  /* device pointer, 3*w*h doubles */
/* Buffer::image() src/buffer.rs:43-56 (box filter, clamp, gamma 2.2, truncating u8) */
int  rptx_film_image_rgb8(rptx_film* film, uint8_t* out_rgb /* [h][w][3] */);
/* Buffer::variance() src/buffer.rs:59-73 */
int  rptx_film_variance(rptx_film* film, double* out);
/* mean over batches of the batch means, unfiltered, linear; [h][w][3] f64 */
int  rptx_film_mean(rptx_film* film, double* out);

/* ---- photon integrators, src/photon.rs -------------------------------------------------
 * photon_render (:655-720) is three passes with a barrier between them; they are exposed
 * one by one so that a multi-GPU host can shard pass 1 by photon index, all-gather the
 * records (NCCL) and shard pass 3 by sample, and so that each pass can be checked alone. */
typedef struct rptx_photons    rptx_photons;     /* PhotonList (:141-171) on the device */
typedef struct rptx_photon_map rptx_photon_map;  /* PhotonMap (:185-189) on the device */

/* One photon record = 13 x 8 bytes: position[3], direction[3] (= wo at the deposit),
 * power[3], starting_position[3] as f64 (Photon, :25-34) + a u64 key
 * (photon index << 16 | bounce depth) that makes list order canonical. */
#define RPTX_PHOTON_ROW_BYTES 104

/* PASS 1 (:663-675): shoot photons [first, first + count) of a run of photon_count_total
 * (power = watts / photon_count_total each, :671).  kind matters for the beam-beam
 * thinning of volume photons (:773-793).  Error if the scene has no Light::Object (:798). */
int  rptx_photons_emit(rptx_ctx* ctx, rptx_scene* scene, const rptx_params* params,
                       uint64_t photon_count_total, uint64_t first, uint64_t count,
                       int32_t kind, rptx_photons** out);
int  rptx_photons_create(rptx_ctx* ctx, rptx_photons** out);            /* empty list */
void rptx_photons_destroy(rptx_photons* photons);
/* PhotonList::merge (:153-160): append rows (which: 0 surface, 1 volume) from host or device memory */
int  rptx_photons_append_host(rptx_photons* photons, int32_t which, const void* rows, size_t n);
int  rptx_photons_append_device(rptx_photons* photons, int32_t which, const void* d_rows, size_t n);
/* PhotonList::merge across the communicator's ranks: *out = the concatenation, in rank order, of every rank's
 * `local` list (ragged; rows go straight to their offset, no padding).  `local` stays valid. */
int  rptx_photons_allgather(rptx_photons* local, rptx_photons** out);
int  rptx_photons_count(const rptx_photons* photons, uint64_t counts[2]);
void* rptx_photons_device_ptr(rptx_photons* photons, int32_t which);      /* rows on the device */
int  rptx_photons_download(rptx_photons* photons, int32_t which, void* rows /* [n][13 x 8 B] */);
int  rptx_photons_stats(const rptx_photons* photons, uint64_t* emitted, uint64_t* rays, double* emit_ms);

/* PASS 2 (:693-702): PhotonMap::new_point_map_for_point_estimate / _for_beam_estimate /
 * new_beam_map_for_beam_estimate (:193-305).  The list may be destroyed afterwards. */
int  rptx_photon_map_build(rptx_ctx* ctx, rptx_scene* scene, rptx_photons* photons, int32_t kind,
                           rptx_photon_map** out);
/* The same map from lists that are spread over the ranks of the context's communicator (rptx_comm_init):
 * `local_photons` is THIS rank's part (e.g. its rptx_photons_emit range).  The ranks split the build of the
 * two kd-trees (:193-199) by Morton range - global splitters, one personalised row exchange, each rank sorts
 * and lays out 1/N of the photons, the finished segments are all-gathered - and the 10-NN radius pass
 * (:212-236) by photon index.  Every rank returns the complete map, identical to rptx_photon_map_build over
 * the merged lists (rptx_photons_allgather).  Collective: every rank of the communicator must call it.
 * Without a communicator it is rptx_photon_map_build. */
int  rptx_photon_map_build_sharded(rptx_ctx* ctx, rptx_scene* scene, rptx_photons* local_photons,
                                   int32_t kind, rptx_photon_map** out);
void rptx_photon_map_destroy(rptx_photon_map* map);
/* counts: surface photons, volume photons in the point map, spheres/beams in the BVH */
int  rptx_photon_map_stats(const rptx_photon_map* map, uint64_t counts[3], double* build_ms);

/* PASS 3 (:706-716): camera pass.  gather_begin makes a film without a path pool;
 * gather_trace evaluates this rank's shard of `iterations` samples per pixel
 * (get_color_with_photon_map, :950-985) into the film's batch-sum buffer; finish the batch
 * with rptx_render_sample_commit (after an all-reduce of the batch sums when sharded). */
int  rptx_photon_gather_begin(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera,
                              const rptx_params* params, rptx_film** out);
int  rptx_photon_gather_trace(rptx_film* film, rptx_photon_map* map, uint32_t iterations);

/* photon_render (:655-720): the three passes on one device, num_samples per pixel. */
int  rptx_photon_render(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera,
                        const rptx_params* params, uint64_t photon_count, int32_t kind,
                        uint8_t* out_rgb8 /* [h][w][3] or NULL */,
                        double* out_mean /* [h][w][3] or NULL */);

/* Parity hook: KdTree::nearests (kd-tree 0.4.1; call sites :329-332, :401-404, :216-225) over
 * host queries q[n][3]: keys[n][k] of the k nearest photons in ascending distance (~0 where
 * the map holds fewer), d2[n][k] squared distances (+inf), counts[n]. */
int  rptx_photon_map_knn(rptx_photon_map* map, int32_t which, const double* q, size_t n, uint32_t k,
                         uint64_t* keys, double* d2, uint32_t* counts);
/* Parity hook: PhotonSphere centres and radii (:212-236) of a point-beam map, BVH leaf order. */
int  rptx_photon_map_spheres(rptx_photon_map* map, double* center /* [n][3] */, double* radius /* [n] */);

/* Parity hooks.
 * get_closest_hit (src/renderer.rs:416-425) over a batch of rays given as host arrays
 * o[n][3], d[n][3].  Outputs (host): t[n] (+inf on miss), object_id[n] (UINT32_MAX on miss),
 * prim_id[n] (triangle index within its mesh, 0 for analytic shapes), inst_id[n] (child index
 * within a GROUP, else 0), normal[n][3].  Any output pointer may be NULL. */
int  rptx_intersect_batch(rptx_scene* scene, const double* o, const double* d, size_t n,
                          double t_min, double* t, uint32_t* object_id, uint32_t* prim_id,
                          uint32_t* inst_id, double* normal);
/* Same query with all arrays resident on the device (used by bench.py for the kernel-only
 * roofline number). */
int  rptx_intersect_batch_device(rptx_scene* scene, const void* d_o, const void* d_d, size_t n,
                                 double t_min, void* d_t, void* d_object_id, void* d_prim_id,
                                 void* d_inst_id, void* d_normal);
int  rptx_stats_get(const rptx_film* film, rptx_stats* out);

/* Allocation-free variant of rptx_intersect_batch.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as
 * rptx_intersect_batch (host arrays; any output pointer may be NULL), plus n_hits (may be NULL):
 * the number of rays that hit something.  The device staging that rptx_intersect_batch
 * cudaMalloc's per call lives in ONE fixed `static __device__` array of the library (about
 * 5.8 MiB, held by every GPU that loads librptx.so until the process exits), so
 *   - n > RPTX_INTERSECT_STATIC_MAX_RAYS is rejected (returns 1, message in rptx_last_error);
 *   - the call performs no cudaMalloc / cudaFree / cudaMallocAsync;
 *   - all calls on one GPU share the array: calls that overlap on the GPU (two contexts or two
 *     host threads on the same device) race; calls through one context are serialised. */
#define RPTX_INTERSECT_STATIC_MAX_RAYS 65536
int  rptx_intersect_batch_static(rptx_scene* scene, const double* o, const double* d, size_t n,
                                 double t_min, double* t, uint32_t* object_id, uint32_t* prim_id,
                                 uint32_t* inst_id, double* normal, uint64_t* n_hits);

#ifdef __cplusplus
}
#endif
#endif /* RPTX_H */
