// Device-side records and f64 vector helpers of the B200 light-transport core.
// Compiled with -fmad=false: every f64 expression below is evaluated unfused, in the
// operation order of the reference (nalgebra: dot = x*x + y*y + z*z left to right,
// normalize = v / norm), so primitive tests reproduce rpt's f64 results bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#define RPTX_HD __host__ __device__ __forceinline__
// -DRPTX_CHECKED: device-side assertions on every shared-memory list, queue and traversal-stack write of the kernels (the pool
// refuses compute-sanitizer; `make -C rpt_b200/csrc checked` builds librptx_checked.so and tools/gpu.sh checked runs the parity
// suite on it).  A failed assertion traps the kernel and the next CUDA call of the host reports it.
#ifdef RPTX_CHECKED
#include <cassert>
#define RPTX_ASSERT(c) assert(c)
#else
#define RPTX_ASSERT(c) ((void)0)
#endif
#define RPTX_D __device__ __forceinline__
// big bodies that kernels call from several sites: one copy each keeps the kernels inside the
// instruction cache (the ray kernels were 100-400 KB of SASS when everything was inlined)
#define RPTX_DN static __device__ __noinline__

namespace rptx {

constexpr double kPi = 3.14159265358979323846264338327950288;
constexpr double kEpsilon = 1e-12;       // src/renderer.rs:17
constexpr double kFireflyClamp = 100.0;  // src/renderer.rs:18
constexpr double kRrMedium = 0.8;        // src/renderer.rs:193
constexpr double kBackgroundDist = 400.0;  // src/renderer.rs:199
constexpr uint32_t kNoHit = 0xffffffffu;

struct d3 {
    double x, y, z;
};
// f64 division and square root expand inline to ~24 / ~20 instructions each; a kernel holding the whole path (persist.cuh)
// has ~160 + ~45 such sites, a third of its code bytes, and is bound by instruction fetch.  With RPTX_NOINLINE_MATH the
// translation unit shares ONE copy of each (same IEEE results: the bodies are the same operations).
#if defined(__CUDA_ARCH__) && defined(RPTX_NOINLINE_MATH)
static __device__ __noinline__ double rptx_nm_div(double a, double b) { return a / b; }
static __device__ __noinline__ double rptx_nm_sqrt(double a) { return sqrt(a); }
#define DDIV(a, b) rptx_nm_div((a), (b))
#define DSQRT(a) rptx_nm_sqrt((a))
#else
#define DDIV(a, b) ((a) / (b))
#define DSQRT(a) sqrt((a))
#endif

RPTX_HD d3 mk3(double x, double y, double z) { d3 r; r.x = x; r.y = y; r.z = z; return r; }
RPTX_HD d3 operator+(d3 a, d3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RPTX_HD d3 operator-(d3 a, d3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RPTX_HD d3 operator-(d3 a) { return mk3(-a.x, -a.y, -a.z); }
RPTX_HD d3 operator*(d3 a, double s) { return mk3(a.x * s, a.y * s, a.z * s); }
RPTX_HD d3 operator*(double s, d3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
// a / s, componentwise, IEEE results.  On the device a ZERO numerator is kept away from the divider: nvcc's
// inlined f64 division sends every quotient of zero magnitude to its ~90-instruction slow path, and
// axis-aligned normals (two zero components) made that 23 % of k_shade's instructions on the Cornell box.
// 0 / s for finite non-zero s is a zero whose sign is sign(a) ^ sign(s), i.e. a * s.
RPTX_HD double div_component(double a, double s, bool s_regular) {
    const bool z = s_regular && a == 0.0;
    double num = z ? 1.0 : a;
#ifdef __CUDA_ARCH__
    asm("" : "+d"(num));  // opaque: otherwise nvcc divides a / s anyway (slow path included) and selects afterwards
#endif
    const double q = DDIV(num, s);
    return z ? a * s : q;
}
// scalar form, for quotients whose numerator is often exactly zero (a ray leaving a wall at coordinate 0
// has dot(pn, v1 - o) == 0 against that wall's own triangles)
RPTX_HD double fdiv(double a, double b) {
#ifdef __CUDA_ARCH__
    return div_component(a, b, fabs(b) > 0.0 && fabs(b) < CUDART_INF);
#else
    return a / b;
#endif
}
RPTX_HD d3 operator/(d3 a, double s) {
#ifdef __CUDA_ARCH__
    const bool reg = fabs(s) > 0.0 && fabs(s) < CUDART_INF;
    return mk3(div_component(a.x, s, reg), div_component(a.y, s, reg), div_component(a.z, s, reg));
#else
    return mk3(a.x / s, a.y / s, a.z / s);
#endif
}
RPTX_HD double dot(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
RPTX_HD d3 cross(d3 a, d3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
RPTX_HD d3 cmul(d3 a, d3 b) { return mk3(a.x * b.x, a.y * b.y, a.z * b.z); }
RPTX_HD double length(d3 a) { return DSQRT(dot(a, a)); }
RPTX_HD d3 normalize(d3 a) { return a / length(a); }
RPTX_HD d3 reflect_vec(d3 i, d3 n) { return i - n * (dot(n, i) * 2.0); }  // glm::reflect_vec
RPTX_HD double clampd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }
RPTX_HD double comp(d3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// 3x4 affine rows (the reference multiplies the full 4x4 by (v,w); with an affine last row the
// xyz result is the same sums in the same order)
struct Affine {
    double m[12];  // row-major rows 0..2
};
RPTX_HD d3 xform_point(const Affine& a, d3 v) {
    return mk3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z + a.m[3] * 1.0,
               a.m[4] * v.x + a.m[5] * v.y + a.m[6] * v.z + a.m[7] * 1.0,
               a.m[8] * v.x + a.m[9] * v.y + a.m[10] * v.z + a.m[11] * 1.0);
}
RPTX_HD d3 xform_dir(const Affine& a, d3 v) {
    return mk3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z + a.m[3] * 0.0,
               a.m[4] * v.x + a.m[5] * v.y + a.m[6] * v.z + a.m[7] * 0.0,
               a.m[8] * v.x + a.m[9] * v.y + a.m[10] * v.z + a.m[11] * 0.0);
}
struct Mat3 {
    double m[9];
};
RPTX_HD d3 mul3(const Mat3& a, d3 v) {
    return mk3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[3] * v.x + a.m[4] * v.y + a.m[5] * v.z,
               a.m[6] * v.x + a.m[7] * v.y + a.m[8] * v.z);
}

// ---- geometry records (HBM layout; all 16-byte aligned for vector loads) ----

// BVH2 node, 64 B = 4 x float4: both children's f32 boxes (already padded outward) + refs.
// ref >= 0: inner node index; ref < 0: leaf, ~ref = (first << 3) | (count - 1).
struct __align__(16) BvhNode {
    float lmin[3]; int left;
    float lmax[3]; int pad0;
    float rmin[3]; int right;
    float rmax[3]; int pad1;
};
constexpr int kLeafMax = 4;
RPTX_HD int make_leaf_ref(uint32_t first, uint32_t count) { return ~(int)((first << 3) | (count - 1)); }

// Prepared triangle, 128 B: everything Triangle::intersect (src/shape/mesh.rs:50-83) derives
// from the vertices, precomputed once with the same f64 operations.
struct __align__(16) TriRec {
    double v1[3], d0[3], d1[3], pn[3];
    double d00, d01, d11, denom;
};
struct __align__(16) TriNormals {  // fetched only for the winning hit
    double n1[3], n2[3], n3[3];
};
// f32 copy for the conservative pre-test in front of the exact f64 test: 3 x float4
struct __align__(16) TriF {
    float v0[3], pad0;
    float e1[3], pad1;
    float e2[3], pad2;
};
struct RawTri {  // upload format
    double v[9], n[9];
};

enum InstKind : int { kSphere = 0, kPlane = 1, kCube = 2, kMesh = 3, kMonomial = 4 };

struct __align__(16) Instance {
    Affine inv;        // inverse transform (world -> local), src/shape.rs:130
    Mat3 nrm;          // M^-T, src/shape.rs:133
    double lmin[3], lmax[3];  // local f64 bounds of the mesh (padded) for the entry clip
    double bmin[3], bmax[3];  // exact local bounds (KdTree::bounds), see mesh_bounds_nan_reject
    double plane_n[3], plane_v;  // PLANE; MONOMIAL: plane_v = height, plane_n[0] = exp
    int kind;
    int has_transform;
    uint32_t object_id, inst_id;
    int root_ref;        // MESH: BLAS root ref
    uint32_t node_base;  // MESH: first node of the BLAS in the node pool
    uint32_t tri_base;   // MESH: first triangle of the BLAS in the triangle pool
    uint32_t ntri;
    float lpad;          // MESH: absolute pad baked into the BLAS boxes
    int pad[3];
};

struct Material {
    int kind;
    int pad;
    double albedo[3];
    double emittance, shininess, ior;
};

struct Light {
    int kind;
    int shape_kind;      // OBJECT lights: kSphere/kCube/kMesh/kMonomial (vec[0] = height, vec[1] = exp)
    int has_transform;
    uint32_t tri_base, ntri;  // OBJECT mesh lights: ORIGINAL-order raw triangles (light pool)
    int pad;
    double color[3], vec[3];
    Affine fwd, inv;     // src/shape.rs:140-151
    Mat3 lin, nrm;
    double scale;
    Material material;
};

struct Camera {
    d3 eye, direction, up;
    double fov, aperture, focal_distance;
    double inv_tan_half_fov;  // 1 / tan(fov / 2) (src/camera.rs:66), evaluated once on the host
};

constexpr uint32_t kSmallSceneMax = 16;

// What phase 1 of closest_hit_small (trace.cuh) reads, as part of the kernel parameters: every lane of a warp walks
// these tables in step, so they are warp-uniform loads from the constant bank instead of dependent L1 hits
// (31.8 cycles each at 16-32 warps per SM was what bound the small-scene kernels).  5.4 KB.
struct SmallTables {
    float4 boxes[2 * kSmallSceneMax];  // padded f32 world box of each bounded instance
    float4 trif[3 * 64];               // TriF records (v0, e1, e2) in slot order
    Affine inv[kSmallSceneMax];        // world -> local of each bounded instance
    int kind[kSmallSceneMax];
    int root_ref[kSmallSceneMax];
    int has_transform[kSmallSceneMax];
};

struct SceneDev {
    const BvhNode* nodes;       // node pool: TLAS first, then BLASes
    const Instance* instances;  // sorted in TLAS leaf order, then unbounded ones
    const TriRec* tris;         // per mesh, BLAS leaf order
    const TriF* trif;           // same order, f32 pre-test records
    const TriNormals* tri_normals;
    const uint32_t* tri_orig;   // original triangle index within its mesh
    const Material* materials;  // per object
    const Light* lights;
    const RawTri* light_tris;
    const uint32_t* cube_instances;  // indices of the kCube instances (see closest_hit: rays lying in a face plane)
    uint32_t n_cube_instances;
    int tlas_root_ref;
    uint32_t n_bounded, n_unbounded, n_lights, n_shadow_lights;
    int small_scene;             // <= kSmallSceneMax instances, <= 64 triangles, every mesh a single leaf: closest_hit_small applies
    const uint8_t* small_slots;  // [<= 64] triangle slot -> instance << 3 | triangle within its mesh leaf
    const float4* small_boxes;   // [n_bounded][2] padded f32 box of each bounded instance
    uint32_t small_kind_mask[3]; // instances of kSphere / kPlane / kCube, bit = instance index
    uint32_t small_unbounded_mask;
    SmallTables small;           // valid when small_scene
    uint32_t shadow_probe_mask;  // bit k: k-th shadow-casting light is probed at its sample first (trace.cuh nee_visible)
    double wmin[3], wmax[3];    // padded world bounds of the bounded instances
    int medium_kind;
    double sigma_a, sigma_s;
    int env_is_hdri;
    double env_color[3];
    uint32_t env_w, env_h;
    const double* env_pixels;
};

}  // namespace rptx
