// C ABI of the B200 light-transport core (include/rptx.h): scene lowering + upload, GPU BVH
// builds, the wavefront render loop and the film.  Host logic only; the kernels live in
// path_kernels.cuh / trace.cuh / bvh_build.cu.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cub/cub.cuh>
#include <cstring>
#include <memory>
#include <string>
#include <map>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "../../include/rptx.h"
#include "bvh_build.cuh"
#include "bvh_sah.hpp"
#include "host_types.hpp"
#include "path_kernels.cuh"
#include "persist.cuh"

using namespace rptx;

namespace {

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

using rptx_host::ctx_fail;

// ---- small host linear algebra (same closed forms nalgebra uses) ----
struct M4 { double m[4][4]; };
struct M3 { double m[3][3]; };

M4 m4_from(const double* t) {
    M4 r;
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) r.m[i][j] = t[4 * i + j];
    return r;
}
M4 m4_inverse(const M4& a) {  // cofactor expansion on the column-major array (glm::inverse)
    double m[16], inv[16];
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) m[c * 4 + r] = a.m[r][c];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    const double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    M4 out;
    if (det == 0.0) { std::memset(&out, 0, sizeof(out)); return out; }
    const double inv_det = 1.0 / det;
    for (int c = 0; c < 4; c++) for (int r = 0; r < 4; r++) out.m[r][c] = inv[c * 4 + r] * inv_det;
    return out;
}
M3 m3_of(const M4& a) {
    M3 r;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.m[i][j] = a.m[i][j];
    return r;
}
void m3_minors(const M3& a, double* p, double* q, double* r, double* det) {
    *p = a.m[1][1] * a.m[2][2] - a.m[2][1] * a.m[1][2];
    *q = a.m[1][0] * a.m[2][2] - a.m[2][0] * a.m[1][2];
    *r = a.m[1][0] * a.m[2][1] - a.m[2][0] * a.m[1][1];
    *det = a.m[0][0] * *p - a.m[0][1] * *q + a.m[0][2] * *r;
}
double m3_det(const M3& a) {
    double p, q, r, d;
    m3_minors(a, &p, &q, &r, &d);
    return d;
}
M3 m3_inverse_transpose(const M3& a) {
    double p, q, r, det;
    m3_minors(a, &p, &q, &r, &det);
    const double m11 = a.m[0][0], m12 = a.m[0][1], m13 = a.m[0][2], m21 = a.m[1][0], m22 = a.m[1][1], m23 = a.m[1][2],
                 m31 = a.m[2][0], m32 = a.m[2][1], m33 = a.m[2][2];
    M3 inv;
    inv.m[0][0] = p / det;
    inv.m[0][1] = (m13 * m32 - m33 * m12) / det;
    inv.m[0][2] = (m12 * m23 - m22 * m13) / det;
    inv.m[1][0] = -q / det;
    inv.m[1][1] = (m11 * m33 - m31 * m13) / det;
    inv.m[1][2] = (m13 * m21 - m23 * m11) / det;
    inv.m[2][0] = r / det;
    inv.m[2][1] = (m12 * m31 - m32 * m11) / det;
    inv.m[2][2] = (m11 * m22 - m21 * m12) / det;
    M3 t;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) t.m[i][j] = inv.m[j][i];
    return t;
}
Affine affine_of(const M4& a) {
    Affine r;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 4; j++) r.m[4 * i + j] = a.m[i][j];
    return r;
}
Mat3 mat3_of(const M3& a) {
    Mat3 r;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) r.m[3 * i + j] = a.m[i][j];
    return r;
}
Affine affine_identity() {
    Affine r;
    std::memset(&r, 0, sizeof(r));
    r.m[0] = r.m[5] = r.m[10] = 1.0;
    return r;
}
Mat3 mat3_identity() {
    Mat3 r;
    std::memset(&r, 0, sizeof(r));
    r.m[0] = r.m[4] = r.m[8] = 1.0;
    return r;
}
float f32_down(double x) {
    float f = (float)x;
    if ((double)f > x) f = std::nextafterf(f, -INFINITY);
    return f;
}
float f32_upf(double x) {
    float f = (float)x;
    if ((double)f < x) f = std::nextafterf(f, INFINITY);
    return f;
}

// Host->device copy ordered on the context's (non-blocking) stream.  A plain cudaMemcpy from
// pageable memory may return before the DMA lands and does not order against that stream.
}  // namespace (reopened below): the block cache has external linkage (host_types.hpp)

namespace {
struct BlockCache {
    std::mutex mu;
    std::vector<size_t> sizes;                                        // sticky sizes, ascending
    std::unordered_map<const void*, std::pair<size_t, cudaStream_t>> live;  // blocks handed out
    std::map<cudaStream_t, std::multimap<size_t, void*>> idle;        // cached whole blocks per stream
};
BlockCache& block_cache() { static BlockCache c; return c; }
constexpr size_t kBigBlock = 32ull << 20;
}  // namespace

cudaError_t rptx_alloc_async(void** p, size_t bytes, cudaStream_t st) {
    if (bytes < kBigBlock) return cudaMallocAsync(p, bytes ? bytes : 16, st);
    BlockCache& c = block_cache();
    std::lock_guard<std::mutex> lock(c.mu);
    size_t want = 0;
    for (size_t r : c.sizes)
        if (r >= bytes) { if (r <= bytes + bytes / 10 * 3) want = r; break; }
    if (!want) {
        const size_t cls = 256ull << 20;
        want = (bytes + bytes / 10 + cls - 1) / cls * cls;
        c.sizes.insert(std::upper_bound(c.sizes.begin(), c.sizes.end(), want), want);
    }
    auto& idle = c.idle[st];
    auto it = idle.lower_bound(want);
    if (it != idle.end() && it->first <= want + want / 10 * 3) {
        *p = it->second;
        c.live[*p] = {it->first, st};
        idle.erase(it);
        return cudaSuccess;
    }
    cudaError_t e = cudaMalloc(p, want);
    if (e != cudaSuccess) {  // out of memory: give the idle blocks of this stream back and try once more
        cudaGetLastError();
        cudaStreamSynchronize(st);
        for (auto& kv : idle) cudaFree(kv.second);
        idle.clear();
        e = cudaMalloc(p, want);
    }
    if (e == cudaSuccess) c.live[*p] = {want, st};
    return e;
}

void rptx_free_async(const void* p, cudaStream_t st) {
    if (!p) return;
    BlockCache& c = block_cache();
    {
        std::lock_guard<std::mutex> lock(c.mu);
        auto it = c.live.find(p);
        if (it != c.live.end()) {
            c.idle[it->second.second].emplace(it->second.first, const_cast<void*>(p));
            c.live.erase(it);
            return;
        }
    }
    cudaFreeAsync(const_cast<void*>(p), st);
}

void rptx_block_cache_release(cudaStream_t st) {
    BlockCache& c = block_cache();
    std::lock_guard<std::mutex> lock(c.mu);
    auto it = c.idle.find(st);
    if (it == c.idle.end()) return;
    for (auto& kv : it->second) cudaFree(kv.second);
    c.idle.erase(it);
}

namespace {

cudaError_t copy_h2d(void* dst, const void* src, size_t bytes, cudaStream_t st) {
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    return e;
}

template <class T>
cudaError_t upload(T** dst, const std::vector<T>& src, cudaStream_t st) {
    *dst = nullptr;
    cudaError_t e = rptx_alloc_async((void**)dst, std::max<size_t>(src.size(), 1) * sizeof(T), st);
    if (e != cudaSuccess) return e;
    if (!src.empty()) e = copy_h2d(*dst, src.data(), src.size() * sizeof(T), st);
    return e;
}

int grid_for(size_t n, int block) { return (int)std::max<size_t>(1, (n + block - 1) / block); }

void free_scene_device(rptx_scene* s) {
    rptx_free_async((void*)s->d_nodes, s->ctx->stream); rptx_free_async((void*)s->d_instances, s->ctx->stream); rptx_free_async((void*)s->d_tris, s->ctx->stream);
    rptx_free_async((void*)s->d_tri_normals, s->ctx->stream); rptx_free_async((void*)s->d_trif, s->ctx->stream); s->d_trif = nullptr; rptx_free_async((void*)s->d_tri_orig, s->ctx->stream); rptx_free_async((void*)s->d_materials, s->ctx->stream);
    rptx_free_async((void*)s->d_lights, s->ctx->stream); rptx_free_async((void*)s->d_raw, s->ctx->stream); rptx_free_async((void*)s->d_env, s->ctx->stream); rptx_free_async((void*)s->d_cube_idx, s->ctx->stream); s->d_cube_idx = nullptr;
    rptx_free_async(s->d_small_slots, s->ctx->stream); s->d_small_slots = nullptr; rptx_free_async(s->d_small_boxes, s->ctx->stream); s->d_small_boxes = nullptr;
    s->d_nodes = nullptr; s->d_instances = nullptr; s->d_tris = nullptr; s->d_tri_normals = nullptr;
    s->d_tri_orig = nullptr; s->d_materials = nullptr; s->d_lights = nullptr; s->d_raw = nullptr; s->d_env = nullptr;
    s->committed = false;
}

Material to_material(const rptx_material& m) {
    Material r;
    std::memset(&r, 0, sizeof(r));
    r.kind = m.kind;
    for (int k = 0; k < 3; k++) r.albedo[k] = m.albedo[k];
    r.emittance = m.emittance; r.shininess = m.shininess; r.ior = m.ior;
    return r;
}

// local bounds of a leaf shape (Bounded::bounding_box of the reference)
bool local_bounds(const rptx_scene* s, const rptx_shape_desc& d, double mn[3], double mx[3]) {
    if (d.kind == RPTX_SHAPE_SPHERE) { for (int k = 0; k < 3; k++) { mn[k] = -1; mx[k] = 1; } return true; }
    if (d.kind == RPTX_SHAPE_CUBE) { for (int k = 0; k < 3; k++) { mn[k] = -0.5; mx[k] = 0.5; } return true; }
    if (d.kind == RPTX_SHAPE_MONOMIAL) {  // src/shape/monomial_surface.rs:175-181
        mn[0] = -1; mn[1] = 0; mn[2] = -1; mx[0] = 1; mx[1] = d.mono_height; mx[2] = 1;
        return true;
    }
    if (d.kind == RPTX_SHAPE_MESH) {
        const HostMesh& m = s->meshes[d.mesh_id];
        for (int k = 0; k < 3; k++) { mn[k] = m.bmin[k]; mx[k] = m.bmax[k]; }
        return true;
    }
    return false;
}

// lower one leaf shape (SPHERE/PLANE/CUBE/MESH) into an instance record + world box
int lower_leaf(rptx_scene* s, const rptx_shape_desc& d, uint32_t object_id, uint32_t inst_id, Instance* out, double wmn[3],
               double wmx[3], bool* bounded) {
    rptx_ctx* ctx = s->ctx;
    Instance in;
    std::memset(&in, 0, sizeof(in));
    in.object_id = object_id;
    in.inst_id = inst_id;
    in.has_transform = d.has_transform ? 1 : 0;
    in.inv = affine_identity();
    in.nrm = mat3_identity();
    M4 fwd = m4_from(d.transform);
    if (d.has_transform) {
        in.inv = affine_of(m4_inverse(fwd));
        in.nrm = mat3_of(m3_inverse_transpose(m3_of(fwd)));
    }
    switch (d.kind) {
        case RPTX_SHAPE_SPHERE: in.kind = kSphere; break;
        case RPTX_SHAPE_CUBE: in.kind = kCube; break;
        case RPTX_SHAPE_MONOMIAL:
            if (d.mono_exp != 4.0) return ctx_fail(ctx, "MonomialSurface: exp must be 4 (src/shape/monomial_surface.rs:17)");
            if (!(d.mono_height > 0.0)) return ctx_fail(ctx, "MonomialSurface: height must be positive");
            in.kind = kMonomial;
            in.plane_v = d.mono_height; in.plane_n[0] = d.mono_exp;
            break;
        case RPTX_SHAPE_PLANE:
            in.kind = kPlane;
            for (int k = 0; k < 3; k++) in.plane_n[k] = d.plane_normal[k];
            in.plane_v = d.plane_value;
            break;
        case RPTX_SHAPE_MESH: {
            if (d.mesh_id >= s->meshes.size()) return ctx_fail(ctx, "shape refers to an unknown mesh id");
            const HostMesh& m = s->meshes[d.mesh_id];
            if (m.tris.empty()) return ctx_fail(ctx, "empty mesh cannot be instanced");
            in.kind = kMesh;
            in.root_ref = m.root_ref; in.node_base = m.node_base; in.tri_base = m.tri_base; in.ntri = (uint32_t)m.tris.size();
            in.lpad = m.pad;
            for (int k = 0; k < 3; k++) { in.lmin[k] = m.bmin[k] - (double)m.pad; in.lmax[k] = m.bmax[k] + (double)m.pad; in.bmin[k] = m.bmin[k]; in.bmax[k] = m.bmax[k]; }
            break;
        }
        default: return ctx_fail(ctx, "unsupported leaf shape kind");
    }
    double mn[3], mx[3];
    *bounded = local_bounds(s, d, mn, mx);
    if (*bounded) {
        for (int k = 0; k < 3; k++) { wmn[k] = INFINITY; wmx[k] = -INFINITY; }
        for (int c = 0; c < 8; c++) {
            const double p[3] = {c & 4 ? mx[0] : mn[0], c & 2 ? mx[1] : mn[1], c & 1 ? mx[2] : mn[2]};
            double w[3] = {p[0], p[1], p[2]};
            if (d.has_transform)
                for (int r = 0; r < 3; r++) w[r] = fwd.m[r][0] * p[0] + fwd.m[r][1] * p[1] + fwd.m[r][2] * p[2] + fwd.m[r][3];
            for (int k = 0; k < 3; k++) { wmn[k] = std::fmin(wmn[k], w[k]); wmx[k] = std::fmax(wmx[k], w[k]); }
        }
    }
    *out = in;
    return 0;
}

int lower_light(rptx_scene* s, const rptx_light& l, Light* out) {
    rptx_ctx* ctx = s->ctx;
    Light r;
    std::memset(&r, 0, sizeof(r));
    r.kind = l.kind;
    for (int k = 0; k < 3; k++) { r.color[k] = l.color[k]; r.vec[k] = l.vec[k]; }
    r.fwd = affine_identity(); r.inv = affine_identity(); r.lin = mat3_identity(); r.nrm = mat3_identity(); r.scale = 1.0;
    if (l.kind == RPTX_LIGHT_OBJECT) {
        const rptx_shape_desc& d = l.shape;
        r.material = to_material(l.material);
        r.has_transform = d.has_transform ? 1 : 0;
        if (d.has_transform) {
            const M4 fwd = m4_from(d.transform);
            r.fwd = affine_of(fwd);
            r.inv = affine_of(m4_inverse(fwd));
            const M3 lin = m3_of(fwd);
            r.lin = mat3_of(lin);
            r.scale = m3_det(lin);
            r.nrm = mat3_of(m3_inverse_transpose(lin));
        }
        if (d.kind == RPTX_SHAPE_MESH) {
            if (d.mesh_id >= s->meshes.size()) return ctx_fail(ctx, "light refers to an unknown mesh id");
            r.shape_kind = kMesh;
            r.tri_base = s->meshes[d.mesh_id].raw_base;
            r.ntri = (uint32_t)s->meshes[d.mesh_id].tris.size();
        } else if (d.kind == RPTX_SHAPE_CUBE) r.shape_kind = kCube;
        else if (d.kind == RPTX_SHAPE_SPHERE) r.shape_kind = kSphere;
        else if (d.kind == RPTX_SHAPE_MONOMIAL) { r.shape_kind = kMonomial; r.vec[0] = d.mono_height; r.vec[1] = d.mono_exp; }
        else return ctx_fail(ctx, "object lights must be mesh, cube, sphere or monomial surface (Plane::sample is unimplemented, src/shape/plane.rs:34)");
    }
    *out = r;
    return 0;
}

}  // namespace

namespace rptx_host {

int ctx_fail(rptx_ctx* ctx, const std::string& msg) {
    if (ctx) ctx->err = msg;
    return 1;
}

MediumDev medium_dev(const rptx_scene* s) {
    MediumDev md;
    std::memset(&md, 0, sizeof(md));
    md.kind = s->medium_kind;
    md.sigma_a = s->sigma_a; md.sigma_s = s->sigma_s;
    auto hex = [](uint32_t x) {
        return mk3(std::pow((double)((x >> 16) & 0xff) / 255.0, 2.2), std::pow((double)((x >> 8) & 0xff) / 255.0, 2.2),
                   std::pow((double)(x & 0xff) / 255.0, 2.2));
    };
    if (s->medium_kind == RPTX_MEDIUM_COLORED_GLOWING_FOG) {  // src/medium.rs:99-121
        md.color_lo = hex(0x0000FF); md.color_hi = hex(0xFF0000);
        md.emission = 10.0;
        md.phase = 1.0 / 4.0 * kPi;
    } else {  // src/medium.rs:80-96
        md.color_lo = hex(0xD2B48C); md.color_hi = md.color_lo;
        md.emission = 0.0;
        md.phase = 1.0 / (4.0 * kPi);
    }
    return md;
}

SceneDev scene_dev(const rptx_scene* s) {
    SceneDev sc;
    std::memset(&sc, 0, sizeof(sc));
    sc.nodes = s->d_nodes; sc.instances = s->d_instances; sc.tris = s->d_tris; sc.trif = s->d_trif; sc.tri_normals = s->d_tri_normals;
    sc.tri_orig = s->d_tri_orig; sc.materials = s->d_materials; sc.lights = s->d_lights; sc.light_tris = s->d_raw;
    sc.tlas_root_ref = s->tlas_root_ref;
    sc.cube_instances = s->d_cube_idx; sc.n_cube_instances = s->n_cubes;
    sc.n_bounded = s->n_bounded; sc.n_unbounded = s->n_unbounded;
    sc.n_lights = (uint32_t)s->lights.size(); sc.n_shadow_lights = s->n_shadow_lights;
    sc.shadow_probe_mask = s->shadow_probe_mask;
    sc.small_scene = s->small_scene && !std::getenv("RPTX_NO_SMALL_SCENE");
    sc.small_slots = s->d_small_slots; sc.small_boxes = s->d_small_boxes;
    for (int k = 0; k < 3; k++) sc.small_kind_mask[k] = s->small_kind_mask[k];
    sc.small_unbounded_mask = s->small_unbounded_mask;
    sc.small = s->small_tables;
    for (int k = 0; k < 3; k++) { sc.wmin[k] = s->wmin[k]; sc.wmax[k] = s->wmax[k]; sc.env_color[k] = s->env_color[k]; }
    sc.medium_kind = s->medium_kind; sc.sigma_a = s->sigma_a; sc.sigma_s = s->sigma_s;
    sc.env_is_hdri = s->env_is_hdri ? 1 : 0; sc.env_w = s->env_w; sc.env_h = s->env_h; sc.env_pixels = s->d_env;
    return sc;
}


}  // namespace rptx_host

using rptx_host::medium_dev;
using rptx_host::scene_dev;

namespace rptx_host {
cudaError_t launch_path_persist(bool medium, bool probe, bool small, int blocks, const SceneDev& sc, const MediumDev& md, const PersistArgs& a,
                                unsigned long long* stats, cudaStream_t st);
}

extern "C" {

int rptx_abi_version(void) { return RPTX_ABI_VERSION; }

int rptx_ctx_create(int device_id, rptx_ctx** out) {
    if (!out) return 1;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    rptx_ctx* ctx = new rptx_ctx();
    *out = ctx;
    if (e != cudaSuccess || n == 0) return ctx_fail(ctx, std::string("no CUDA device: ") + cudaGetErrorString(e));
    if (device_id < 0 || device_id >= n) return ctx_fail(ctx, "device id out of range");
    ctx->device = device_id;
    CK(cudaSetDevice(device_id));
    CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device_id));
    ctx->sm_count = prop.multiProcessorCount;
    CK(cudaMallocHost((void**)&ctx->h_pinned, 64));
    {  // scene tables, batch means and every allocation under 32 MB come from the stream-ordered pool (larger ones from the block
       // cache above: rptx_alloc_async).  The pool never trims: with a finite release threshold (16 GB in round 1) it unmapped and
       // re-mapped the excess at every synchronisation (beam-beam step 3.9 -> 6.2 s once another workload had run in the process)
        cudaMemPool_t pool;
        CK(cudaDeviceGetDefaultMemPool(&pool, device_id));
        uint64_t keep = UINT64_MAX;
        CK(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    return 0;
}

void rptx_ctx_destroy(rptx_ctx* ctx) {
    if (!ctx) return;
    if (ctx->comm) rptx_comm_destroy(ctx);
    if (ctx->spare_arena) cudaFree(ctx->spare_arena);
    if (ctx->stream) { cudaStreamSynchronize(ctx->stream); rptx_block_cache_release(ctx->stream); cudaStreamDestroy(ctx->stream); }
    if (ctx->h_pinned) cudaFreeHost(ctx->h_pinned);
    delete ctx;
}

const char* rptx_last_error(const rptx_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

// Hands the memory this context keeps cached (the stream-ordered pool never trims on its own, and a destroyed film leaves its
// slab for the next one) back to the device.  For hosts that switch between workloads with different footprints: a 1e8-photon
// render caches ~45 GB, and the next workload's first allocations would otherwise have to squeeze them out one by one.
int rptx_ctx_trim(rptx_ctx* ctx) {
    if (!ctx) return 1;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->spare_arena) { cudaFree(ctx->spare_arena); ctx->spare_arena = nullptr; ctx->spare_arena_bytes = 0; }
    rptx_block_cache_release(ctx->stream);
    cudaMemPool_t pool;
    CK(cudaDeviceGetDefaultMemPool(&pool, ctx->device));
    CK(cudaMemPoolTrimTo(pool, 0));
    return 0;
}

int rptx_scene_create(rptx_ctx* ctx, rptx_scene** out) {
    if (!ctx || !out) return 1;
    rptx_scene* s = new rptx_scene();
    s->ctx = ctx;
    *out = s;
    return 0;
}

void rptx_scene_destroy(rptx_scene* s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    free_scene_device(s);
    delete s;
}

int rptx_scene_add_mesh(rptx_scene* s, const double* v, const double* n, size_t ntri, uint32_t* mesh_id) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    if (!v || !mesh_id) return ctx_fail(ctx, "rptx_scene_add_mesh: null argument");
    if (ntri >= (1u << 28)) return ctx_fail(ctx, "mesh too large (>= 2^28 triangles)");
    HostMesh m;
    m.has_normals = n != nullptr;
    m.tris.resize(ntri);
    for (int k = 0; k < 3; k++) { m.bmin[k] = INFINITY; m.bmax[k] = -INFINITY; }
    for (size_t i = 0; i < ntri; i++) {
        std::memcpy(m.tris[i].v, v + 9 * i, 9 * sizeof(double));
        if (n) std::memcpy(m.tris[i].n, n + 9 * i, 9 * sizeof(double));
        else std::memset(m.tris[i].n, 0, 9 * sizeof(double));
        for (int c = 0; c < 3; c++)
            for (int k = 0; k < 3; k++) {
                const double x = v[9 * i + 3 * c + k];
                if (!(x == x) || std::isinf(x)) return ctx_fail(ctx, "mesh vertex is not finite");
                m.bmin[k] = std::fmin(m.bmin[k], x); m.bmax[k] = std::fmax(m.bmax[k], x);
            }
    }
    s->meshes.push_back(std::move(m));
    *mesh_id = (uint32_t)s->meshes.size() - 1;
    s->committed = false;
    return 0;
}

int rptx_scene_add_object(rptx_scene* s, const rptx_shape_desc* shape, const rptx_material* material, uint32_t* object_id) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    if (!shape || !material) return ctx_fail(ctx, "rptx_scene_add_object: null argument");
    HostObject o;
    o.desc = *shape;
    o.material = *material;
    if (shape->kind == RPTX_SHAPE_GROUP) {
        if (shape->has_transform) return ctx_fail(ctx, "transformed groups are not supported");
        if (shape->n_children == 0 || !shape->children) return ctx_fail(ctx, "empty group");
        o.children.assign(shape->children, shape->children + shape->n_children);
        for (const auto& c : o.children)
            if (c.kind != RPTX_SHAPE_SPHERE && c.kind != RPTX_SHAPE_CUBE && c.kind != RPTX_SHAPE_MESH && c.kind != RPTX_SHAPE_MONOMIAL)
                return ctx_fail(ctx, "group children must be bounded leaf shapes (sphere, cube, mesh, monomial surface)");
        o.desc.children = nullptr;
    } else if (shape->kind < RPTX_SHAPE_SPHERE || shape->kind > RPTX_SHAPE_MONOMIAL) {
        return ctx_fail(ctx, "unknown shape kind");
    }
    if (shape->kind == RPTX_SHAPE_MESH && shape->mesh_id >= s->meshes.size()) return ctx_fail(ctx, "unknown mesh id");
    s->objects.push_back(std::move(o));
    if (object_id) *object_id = (uint32_t)s->objects.size() - 1;
    s->committed = false;
    return 0;
}

int rptx_scene_add_light(rptx_scene* s, const rptx_light* light) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    if (!light) return ctx_fail(ctx, "rptx_scene_add_light: null argument");
    if (light->kind < RPTX_LIGHT_POINT || light->kind > RPTX_LIGHT_OBJECT) return ctx_fail(ctx, "unknown light kind");
    s->lights.push_back(*light);
    s->committed = false;
    return 0;
}

int rptx_scene_set_medium(rptx_scene* s, int32_t kind, double absorption, double scattering) {
    if (!s) return 1;
    if (kind < RPTX_MEDIUM_NONE || kind > RPTX_MEDIUM_COLORED_GLOWING_FOG) return ctx_fail(s->ctx, "unknown medium kind");
    s->medium_kind = kind; s->sigma_a = absorption; s->sigma_s = scattering;
    return 0;
}

int rptx_scene_set_environment_color(rptx_scene* s, const double rgb[3]) {
    if (!s || !rgb) return 1;
    s->env_is_hdri = false;
    for (int k = 0; k < 3; k++) s->env_color[k] = rgb[k];
    return 0;
}

int rptx_scene_set_environment_hdri(rptx_scene* s, uint32_t w, uint32_t h, const double* px) {
    if (!s) return 1;
    if (!px || w == 0 || h == 0) return ctx_fail(s->ctx, "bad HDRI");
    s->env_is_hdri = true; s->env_w = w; s->env_h = h;
    s->env_pixels.assign(px, px + (size_t)w * h * 3);
    s->committed = false;
    return 0;
}

int rptx_scene_commit(rptx_scene* s) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    CK(cudaSetDevice(ctx->device));
    free_scene_device(s);
    cudaStream_t st = ctx->stream;

    // 1. meshes: upload raw triangles, build one BLAS per mesh (GPU: PLOC, or the LBVH)
    const bool lbvh_only = std::getenv("RPTX_LBVH_ONLY") != nullptr;  // developer knob: the GPU LBVH for every tree
    // object trees are built on the GPU with the surface-area criterion (build_ploc; an LBVH if PLOC declines).
    // RPTX_BVH_BUILDER=host: binned-SAH trees from the host for up to 65 K items (bvh_sah.hpp) - 2-4 % faster to traverse on
    // fractal_teapots (extend 33.4 vs 34.9 ms per 64 spp at max_bounces 0, 276 vs 284 at 5), kept for comparison
    bool gpu_sah = true;
    if (const char* e = std::getenv("RPTX_BVH_BUILDER")) gpu_sah = std::string(e) != "host";
    if (lbvh_only) gpu_sah = false;
    constexpr int kPlocMaxDepth = 56;  // two levels + sentinel must fit kStackSize = 128 pending entries
    std::vector<RawTri> raw_all;
    uint32_t tri_total = 0, node_total = 0;
    for (auto& m : s->meshes) {
        m.raw_base = (uint32_t)raw_all.size();
        m.tri_base = tri_total;
        raw_all.insert(raw_all.end(), m.tris.begin(), m.tris.end());
        tri_total += (uint32_t)m.tris.size();
    }
    // count bounded instances for the TLAS node range
    uint32_t n_leaf_shapes = 0;
    for (const auto& o : s->objects) n_leaf_shapes += o.desc.kind == RPTX_SHAPE_GROUP ? (uint32_t)o.children.size() : 1u;
    node_total = std::max<uint32_t>(n_leaf_shapes, 1);
    for (auto& m : s->meshes) { m.node_base = node_total; node_total += std::max<uint32_t>((uint32_t)m.tris.size(), 1); }

    RawTri* d_raw = nullptr;
    CK(upload(&d_raw, raw_all, st));
    s->d_raw = d_raw;
    BvhNode* d_nodes = nullptr;
    CK(rptx_alloc_async((void**)&d_nodes, sizeof(BvhNode) * node_total, st));
    CK(cudaMemsetAsync(d_nodes, 0, sizeof(BvhNode) * node_total, st));
    s->d_nodes = d_nodes;
    TriRec* d_tris = nullptr; TriNormals* d_tn = nullptr; uint32_t* d_orig = nullptr; TriF* d_trif = nullptr;
    CK(rptx_alloc_async((void**)&d_trif, sizeof(TriF) * std::max<uint32_t>(tri_total, 1), st));
    s->d_trif = d_trif;
    CK(rptx_alloc_async((void**)&d_tris, sizeof(TriRec) * std::max<uint32_t>(tri_total, 1), st));
    CK(rptx_alloc_async((void**)&d_tn, sizeof(TriNormals) * std::max<uint32_t>(tri_total, 1), st));
    CK(rptx_alloc_async((void**)&d_orig, sizeof(uint32_t) * std::max<uint32_t>(tri_total, 1), st));
    s->d_tris = d_tris; s->d_tri_normals = d_tn; s->d_tri_orig = d_orig;

    for (auto& m : s->meshes) {
        const uint32_t n = (uint32_t)m.tris.size();
        if (n == 0) continue;
        if (!m.has_normals) {
            k_face_normals<<<grid_for(n, 128), 128, 0, st>>>(d_raw + m.raw_base, n);
            ctx->launches++;
        }
        PrimBox* d_boxes = nullptr; uint32_t* d_perm = nullptr;
        double e_abs = 0.0;
        for (int k = 0; k < 3; k++) e_abs = std::fmax(e_abs, std::fmax(std::fabs(m.bmin[k]), std::fabs(m.bmax[k])));
        m.pad = f32_upf(std::fmax(e_abs, 1e-30) * (1.0 / 2097152.0));  // E * 2^-21
        if (n <= rptx_host::kSahMaxPrims && !lbvh_only && !gpu_sah) {
            // small mesh: binned SAH on the host (bvh_sah.hpp), same node format, same padding
            std::vector<PrimBox> hb(n);
            for (uint32_t i = 0; i < n; i++)
                for (int k = 0; k < 3; k++) {
                    const double* v = m.tris[i].v;
                    hb[i].mn[k] = f32_down(std::fmin(std::fmin(v[k], v[3 + k]), v[6 + k]));
                    hb[i].mx[k] = f32_upf(std::fmax(std::fmax(v[k], v[3 + k]), v[6 + k]));
                }
            std::vector<BvhNode> hn;
            std::vector<uint32_t> hp;
            int leaf_max = kLeafMax;
            if (const char* e = std::getenv("RPTX_SAH_LEAF")) leaf_max = std::min(8, std::max(1, std::atoi(e)));  // developer knob
            rptx_host::build_sah_host(hb, m.pad, leaf_max, hn, hp, &m.root_ref);
            if (!hn.empty()) CK(copy_h2d(d_nodes + m.node_base, hn.data(), sizeof(BvhNode) * hn.size(), st));
            CK(upload(&d_perm, hp, st));
            CK(cudaStreamSynchronize(st));  // hn / hp go out of scope
        } else {
            CK(rptx_alloc_async((void**)&d_boxes, sizeof(PrimBox) * n, st));
            CK(rptx_alloc_async((void**)&d_perm, sizeof(uint32_t) * n, st));
            k_tri_boxes<<<grid_for(n, 128), 128, 0, st>>>(d_raw + m.raw_base, n, d_boxes);
            ctx->launches++;
            int rc = (int)cudaErrorNotSupported;
            if (gpu_sah) rc = build_ploc(d_boxes, n, m.pad, kLeafMax, m.bmin, m.bmax, d_nodes + m.node_base, d_perm, &m.root_ref, kPlocMaxDepth, st, &ctx->launches);
            if (rc == (int)cudaErrorNotSupported) rc = build_lbvh(d_boxes, n, m.pad, kLeafMax, m.bmin, m.bmax, d_nodes + m.node_base, d_perm, &m.root_ref, st, &ctx->launches);
            if (rc != 0) { rptx_free_async(d_boxes, st); rptx_free_async(d_perm, st); return ctx_fail(ctx, std::string("BLAS build: ") + cudaGetErrorString((cudaError_t)rc)); }
        }
        k_tri_prepare<<<grid_for(n, 128), 128, 0, st>>>(d_raw + m.raw_base, d_perm, n, d_tris + m.tri_base, d_trif + m.tri_base, d_tn + m.tri_base,
                                                         d_orig + m.tri_base);
        ctx->launches++;
        CK(cudaStreamSynchronize(st));
        rptx_free_async(d_boxes, st); rptx_free_async(d_perm, st);
    }
    // 2. instances
    std::vector<Instance> bounded, unbounded;
    std::vector<PrimBox> boxes;
    double wmn[3] = {INFINITY, INFINITY, INFINITY}, wmx[3] = {-INFINITY, -INFINITY, -INFINITY};
    std::vector<std::array<double, 6>> wboxes;
    for (uint32_t oi = 0; oi < s->objects.size(); oi++) {
        const HostObject& o = s->objects[oi];
        const uint32_t nc = o.desc.kind == RPTX_SHAPE_GROUP ? (uint32_t)o.children.size() : 1u;
        for (uint32_t c = 0; c < nc; c++) {
            const rptx_shape_desc& d = o.desc.kind == RPTX_SHAPE_GROUP ? o.children[c] : o.desc;
            Instance in;
            double bmn[3], bmx[3];
            bool is_bounded = false;
            if (lower_leaf(s, d, oi, o.desc.kind == RPTX_SHAPE_GROUP ? c : 0u, &in, bmn, bmx, &is_bounded)) return 1;
            if (is_bounded) {
                bounded.push_back(in);
                wboxes.push_back({bmn[0], bmn[1], bmn[2], bmx[0], bmx[1], bmx[2]});
                for (int k = 0; k < 3; k++) { wmn[k] = std::fmin(wmn[k], bmn[k]); wmx[k] = std::fmax(wmx[k], bmx[k]); }
            } else {
                unbounded.push_back(in);
            }
        }
    }
    s->n_bounded = (uint32_t)bounded.size();
    s->n_unbounded = (uint32_t)unbounded.size();
    std::vector<Instance> ordered;
    float wpad_used = 0.0f;
    std::vector<std::array<double, 6>> wboxes_ordered;  // world bounds of the bounded instances, TLAS order
    s->tlas_root_ref = 0;
    if (!bounded.empty()) {
        double e_abs = 0.0;
        for (int k = 0; k < 3; k++) e_abs = std::fmax(e_abs, std::fmax(std::fabs(wmn[k]), std::fabs(wmx[k])));
        const float wpad = f32_upf(std::fmax(e_abs, 1e-30) * (1.0 / 2097152.0));
        wpad_used = wpad;
        for (const auto& b : wboxes) {
            PrimBox pb;
            for (int k = 0; k < 3; k++) { pb.mn[k] = f32_down(b[k]); pb.mx[k] = f32_upf(b[3 + k]); }
            boxes.push_back(pb);
        }
        std::vector<uint32_t> perm(boxes.size());
        if (boxes.size() <= rptx_host::kSahMaxPrims && !lbvh_only && !gpu_sah) {
            std::vector<BvhNode> hn;
            rptx_host::build_sah_host(boxes, wpad, 1, hn, perm, &s->tlas_root_ref);
            if (!hn.empty()) CK(copy_h2d(d_nodes, hn.data(), sizeof(BvhNode) * hn.size(), st));
            CK(cudaStreamSynchronize(st));
        } else {
            PrimBox* d_boxes = nullptr; uint32_t* d_perm = nullptr;
            CK(upload(&d_boxes, boxes, st));
            CK(rptx_alloc_async((void**)&d_perm, sizeof(uint32_t) * boxes.size(), st));
            int rc = (int)cudaErrorNotSupported;
            if (gpu_sah) rc = build_ploc(d_boxes, (uint32_t)boxes.size(), wpad, 1, wmn, wmx, d_nodes, d_perm, &s->tlas_root_ref, kPlocMaxDepth, st, &ctx->launches);
            if (rc == (int)cudaErrorNotSupported) rc = build_lbvh(d_boxes, (uint32_t)boxes.size(), wpad, 1, wmn, wmx, d_nodes, d_perm, &s->tlas_root_ref, st, &ctx->launches);
            if (rc != 0) { rptx_free_async(d_boxes, st); rptx_free_async(d_perm, st); return ctx_fail(ctx, std::string("TLAS build: ") + cudaGetErrorString((cudaError_t)rc)); }
            CK(cudaMemcpyAsync(perm.data(), d_perm, sizeof(uint32_t) * perm.size(), cudaMemcpyDeviceToHost, st));
            CK(cudaStreamSynchronize(st));
            rptx_free_async(d_boxes, st); rptx_free_async(d_perm, st);
        }
        for (uint32_t k : perm) { ordered.push_back(bounded[k]); wboxes_ordered.push_back(wboxes[k]); }
        for (int k = 0; k < 3; k++) { s->wmin[k] = wmn[k] - 2.0 * (double)wpad; s->wmax[k] = wmx[k] + 2.0 * (double)wpad; }
    }
    ordered.insert(ordered.end(), unbounded.begin(), unbounded.end());
    Instance* d_inst = nullptr;
    CK(upload(&d_inst, ordered, st));
    s->d_instances = d_inst;
    std::vector<uint32_t> cubes;
    for (uint32_t i = 0; i < s->n_bounded; i++)
        if (ordered[i].kind == kCube) cubes.push_back(i);
    // small scenes: candidate tables of closest_hit_small (trace.cuh)
    {
        s->small_scene = ordered.size() <= kSmallSceneMax;
        std::vector<uint8_t> slots;
        std::vector<float4> sboxes;
        for (int k = 0; k < 3; k++) s->small_kind_mask[k] = 0;
        s->small_unbounded_mask = 0;
        const float spad = wpad_used;
        for (uint32_t i = 0; i < ordered.size() && s->small_scene; i++) {
            const Instance& in = ordered[i];
            if (in.kind == kMesh) {
                if (in.root_ref >= 0) { s->small_scene = 0; break; }
                const uint32_t count = ((uint32_t)(~in.root_ref) & 7u) + 1u;
                for (uint32_t j = 0; j < count; j++) slots.push_back((uint8_t)(i << 3 | j));
            } else if (in.kind == kMonomial) {
                s->small_scene = 0;  // the root finder is too long for the in-step candidate walk
                break;
            } else {
                s->small_kind_mask[in.kind] |= 1u << i;
                if (i >= s->n_bounded) s->small_unbounded_mask |= 1u << i;
            }
            if (i < s->n_bounded) {
                const auto& wb = wboxes_ordered[i];
                sboxes.push_back(make_float4(f32_down(wb[0]) - spad, f32_down(wb[1]) - spad, f32_down(wb[2]) - spad, 0.0f));
                sboxes.push_back(make_float4(f32_upf(wb[3]) + spad, f32_upf(wb[4]) + spad, f32_upf(wb[5]) + spad, 0.0f));
            }
        }
        if (slots.size() > 64) s->small_scene = 0;
        if (s->small_scene) {
            uint8_t* d_slots = nullptr; float4* d_sboxes = nullptr;
            CK(upload(&d_slots, slots, st));
            CK(upload(&d_sboxes, sboxes, st));
            s->d_small_slots = d_slots; s->d_small_boxes = d_sboxes;
            // kernel-parameter copy of the tables phase 1 of closest_hit_small walks (SmallTables, common.cuh)
            SmallTables& t = s->small_tables;
            std::memset(&t, 0, sizeof(t));
            uint32_t slot = 0;
            for (uint32_t i = 0; i < s->n_bounded; i++) {
                const Instance& in = ordered[i];
                t.kind[i] = in.kind; t.root_ref[i] = in.root_ref; t.has_transform[i] = in.has_transform; t.inv[i] = in.inv;
                t.boxes[2 * i] = sboxes[2 * i]; t.boxes[2 * i + 1] = sboxes[2 * i + 1];
                if (in.kind == kMesh) {
                    const uint32_t code = (uint32_t)(~in.root_ref);
                    const uint32_t first = code >> 3, count = (code & 7u) + 1u;
                    CK(cudaMemcpyAsync(&t.trif[3 * slot], d_trif + in.tri_base + first, sizeof(TriF) * count, cudaMemcpyDeviceToHost, st));
                    slot += count;
                }
            }
            CK(cudaStreamSynchronize(st));
        }
    }
    uint32_t* d_cubes = nullptr;
    CK(upload(&d_cubes, cubes, st));
    s->d_cube_idx = d_cubes;
    s->n_cubes = (uint32_t)cubes.size();

    // 3. materials, lights, environment
    std::vector<Material> mats;
    for (const auto& o : s->objects) mats.push_back(to_material(o.material));
    Material* d_mats = nullptr;
    CK(upload(&d_mats, mats, st));
    s->d_materials = d_mats;
    std::vector<Light> lights;
    s->n_shadow_lights = 0;
    s->shadow_probe_mask = 0;
    for (const auto& l : s->lights) {
        Light r;
        if (lower_light(s, l, &r)) return 1;
        if (l.kind == RPTX_LIGHT_POINT || l.kind == RPTX_LIGHT_OBJECT) {
            if (l.kind == RPTX_LIGHT_POINT && s->n_shadow_lights < 32) s->shadow_probe_mask |= 1u << s->n_shadow_lights;
            s->n_shadow_lights++;
        }
        lights.push_back(r);
    }
    Light* d_lights = nullptr;
    CK(upload(&d_lights, lights, st));
    s->d_lights = d_lights;
    if (s->env_is_hdri) {
        double* d_env = nullptr;
        CK(upload(&d_env, s->env_pixels, st));
        s->d_env = d_env;
    }
    CK(cudaStreamSynchronize(st));
    s->committed = true;
    return 0;
}

// ---------------------------------------------------------------- film / render loop

// the film of a photon camera pass or of the fused path kernel needs the lane sums and the Buffer, not the path pool
// Per-slot path state is carved from one slab (ctx->spare_arena when a previous film left one that is large
// enough).  256-byte alignment per array.
extern "C++" {
namespace {
struct Carver {
    char* base;
    size_t off = 0;
    template <class T>
    void take(T** out, size_t n) {
        off = (off + 255) & ~(size_t)255;
        if (base) *out = reinterpret_cast<T*>(base + off);
        off += n * sizeof(T);
    }
};
}  // namespace
}  // extern "C++"

static void film_carve(rptx_film* f, Carver& c) {
    const bool gather_only = f->mode != kFilmWavefront;
    PathPool& p = f->pool;
    const size_t ns = p.n_slots;
    std::vector<double**> dptrs = {&p.ox, &p.oy, &p.oz, &p.dx, &p.dy, &p.dz, &p.ht, &p.hnx, &p.hny, &p.hnz, &p.ax, &p.ay, &p.az};
    if (gather_only) dptrs = {&p.ax, &p.ay, &p.az};
    for (double** d : dptrs) c.take(d, ns);
    if (!gather_only) {
        c.take(&p.sjob, ns * std::max<uint32_t>(p.n_shadow, 1) * 8);
        if (p.record_mode) {  // surface mode: the vertex lives in its 64-byte level record
            c.take(&p.stack, ns * p.rec_stride * std::max<uint32_t>(p.stack_depth, 1));
        } else {              // medium mode: forward accumulation
            c.take(&p.mvtx, ns * 8);
            c.take(&p.mpath, ns * 8);
        }
        c.take(&p.hobj, ns); c.take(&p.rng, ns); c.take(&p.sample_j, ns); c.take(&p.depth, ns); c.take(&p.flags, ns);
        c.take(&f->q.cur, ns); c.take(&f->q.next, ns); c.take(&f->q.keep, ns);
        c.take(&f->d_select_tmp, f->select_tmp_bytes);
    }
    if (f->mode == kFilmPersist && p.record_mode)
        c.take(&f->d_pstack, (size_t)f->persist_blocks * kPersistBlock * 8 * std::max<uint32_t>(p.stack_depth, 1));
    c.take(&f->q.counters, 8);
    c.take(&f->q.stats, 8);
    const size_t n3 = (size_t)p.n_pixels * 3;
    c.take(&f->d_batch_sum, n3);
    c.take(&f->d_film_sum, n3);
    c.take(&f->d_scratch, n3);  // read-back staging (rgb8 image, mean): no cudaMalloc / cudaFree on the read-back calls
    // room for the first batch means (Buffer keeps every batch for variance()); later ones come from the stream-ordered pool
    f->n_mean_slots = std::min<uint32_t>(std::max<uint32_t>(f->params.num_samples, 1), 16);
    c.take(&f->d_mean_slots, n3 * f->n_mean_slots);
}

static int film_alloc(rptx_film* f) {
    rptx_ctx* ctx = f->ctx;
    if (f->mode == kFilmWavefront) {
        uint32_t* u = nullptr; uint8_t* b = nullptr;
        CK(cub::DeviceSelect::Flagged(nullptr, f->select_tmp_bytes, u, b, u, u, (int)f->pool.n_slots, ctx->stream));
    }
    Carver size_pass{nullptr};
    film_carve(f, size_pass);
    const size_t need = size_pass.off + 256;
    if (ctx->spare_arena && ctx->spare_arena_bytes >= need) {
        f->arena = ctx->spare_arena; f->arena_bytes = ctx->spare_arena_bytes;
        ctx->spare_arena = nullptr; ctx->spare_arena_bytes = 0;
    } else {
        if (std::getenv("RPTX_TRACE_ALLOC")) fprintf(stderr, "[rptx] film slab: need %zu B, spare %zu B -> cudaMalloc\n", need, ctx->spare_arena_bytes);
        if (ctx->spare_arena) { cudaFree(ctx->spare_arena); ctx->spare_arena = nullptr; ctx->spare_arena_bytes = 0; }
        CK(cudaMalloc(&f->arena, need));
        f->arena_bytes = need;
    }
    Carver carve{static_cast<char*>(f->arena)};
    film_carve(f, carve);
    const size_t n3 = (size_t)f->pool.n_pixels * 3;
    CK(cudaMemsetAsync(f->q.stats, 0, sizeof(unsigned long long) * 8, ctx->stream));
    CK(cudaMemsetAsync(f->d_film_sum, 0, sizeof(double) * n3, ctx->stream));
    CK(cudaMemsetAsync(f->d_batch_sum, 0, sizeof(double) * n3, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaEventCreate(&f->ev0)); CK(cudaEventCreate(&f->ev1));
    for (int r = 0; r < 4; r++) {
        for (int k = 0; k < 5; k++) CK(cudaEventCreate(&f->evw[r][k]));
        CK(cudaEventCreate(&f->ev_done[r]));
    }
    f->no_one_vertex = std::getenv("RPTX_NO_ONE_VERTEX") != nullptr;  // developer knob, read once per film
    if (const char* e = std::getenv("RPTX_SHADE_TILE_ROUNDS")) f->shade_rounds = (uint32_t)std::min(std::max(std::atoi(e), 0), kShadeRoundsMax);  // experiments: fixes the k_shade tile
    return 0;
}

// Path-tracing films pick their kernel set here: the fused per-thread kernel when the whole
// accelerator is a handful of nodes (L1-resident), the wavefront otherwise.  RPTX_PATH_MODE =
// wavefront | fused overrides (benchmarks compare both).
static int choose_path_mode(const rptx_scene* s, const rptx_params* p) {
    const bool fits = p->max_bounces + 1 <= (uint32_t)kFusedMaxDepth || s->medium_kind != RPTX_MEDIUM_NONE;
    const char* env = std::getenv("RPTX_PATH_MODE");
    if (env && std::strcmp(env, "wavefront") == 0) return kFilmWavefront;
    if (env && std::strcmp(env, "fused") == 0 && fits) return kFilmFused;
    if (env && std::strcmp(env, "persist") == 0) return kFilmPersist;
    // measured (profiles/r1_v2_summary.md): the fused kernel wins only for one-vertex paths on small
    // scenes (teapot, max_bounces 0: +33 %); with bounces the wavefront's compaction wins 3-7x
    size_t tris = 0;
    for (const auto& m : s->meshes) tris += m.tris.size();
    const bool one_vertex = s->medium_kind == RPTX_MEDIUM_NONE && p->max_bounces == 0;
    return (one_vertex && tris <= 4096 && s->n_bounded + s->n_unbounded <= 64) ? kFilmFused : kFilmWavefront;
}

static int film_begin(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera, const rptx_params* params, bool gather_only,
                      rptx_film** out) {
    if (!ctx || !out) return 1;
    *out = nullptr;
    if (!scene || !camera || !params) return ctx_fail(ctx, "render begin: null argument");
    if (scene->ctx != ctx) return ctx_fail(ctx, "scene belongs to another context");
    if (!scene->committed) return ctx_fail(ctx, "scene not committed");
    if (params->width == 0 || params->height == 0) return ctx_fail(ctx, "empty image");
    if (params->shard_count == 0 || params->shard_index >= params->shard_count) return ctx_fail(ctx, "bad shard");
    if (params->max_bounces > kMaxBounces) return ctx_fail(ctx, "max_bounces above 4096 is not supported");
    CK(cudaSetDevice(ctx->device));
    rptx_film* f = new rptx_film();
    f->ctx = ctx; f->scene = scene; f->params = *params;
    f->mode = gather_only ? kFilmGather : choose_path_mode(scene, params);
    f->cam.eye = mk3(camera->eye[0], camera->eye[1], camera->eye[2]);
    f->cam.direction = mk3(camera->direction[0], camera->direction[1], camera->direction[2]);
    f->cam.up = mk3(camera->up[0], camera->up[1], camera->up[2]);
    f->cam.fov = camera->fov; f->cam.aperture = camera->aperture; f->cam.focal_distance = camera->focal_distance;
    f->cam.inv_tan_half_fov = 1.0 / std::tan(camera->fov / 2.0);
    PathPool& p = f->pool;
    std::memset(&p, 0, sizeof(p));
    const uint64_t n_pixels = (uint64_t)params->width * params->height;
    if (n_pixels >= (1ull << 31)) { delete f; return ctx_fail(ctx, "image too large"); }
    p.n_pixels = (uint32_t)n_pixels;
    p.width = params->width;
    p.tiled = (f->mode == kFilmWavefront || f->mode == kFilmPersist) && params->width % 8 == 0 && params->height % 4 == 0 && !std::getenv("RPTX_LINEAR_PIXELS");
    p.n_shadow = scene->n_shadow_lights;
    p.stack_depth = scene->medium_kind == RPTX_MEDIUM_NONE ? params->max_bounces + 1 : 1;
    p.record_mode = scene->medium_kind == RPTX_MEDIUM_NONE;
    p.rec_stride = p.stack_depth == 1 ? 4 : 8;
    // Paths in flight.  Wavefront: as many as a quarter of the free HBM holds, up to 32 M (~0.7 KB per slot on
    // the Cornell box).  The pool size is not about occupancy (2 M paths already fill the GPU) but about
    // coherence: a slot that regenerates mid-batch puts depth-0 and depth-5 vertices into the same warps, so
    // the fewer regenerations, the closer every wave is to one bounce depth of neighbouring pixels.  Measured
    // (profiles/r1_v2_summary.md): 2 M -> 32 M slots is +40 % on cornell, +26 % on lampshade, +9 % on fractal.
    // Fused / photon gather: ~0.5 M threads, each walking its samples.
    uint64_t target = 1ull << 19;
    if (f->mode == kFilmWavefront) {
        const uint64_t slot_bytes = 137 + 56ull * std::max<uint32_t>(p.n_shadow, 1) + (p.record_mode ? 64ull * std::max<uint32_t>(p.stack_depth, 1) : 96ull);
        size_t free_b = 0, total_b = 0;
        CK(cudaMemGetInfo(&free_b, &total_b));
        target = 1ull << 25;
        // deep clamp stacks (64 B x (max_bounces + 1) per slot) may take the pool below 2^21 slots, down to one lane
        while (target > n_pixels && target * slot_bytes > free_b / 4) target >>= 1;
    }
    if (f->mode == kFilmPersist) {
        // slots (pixel, sample lane) are claimed by the resident threads of ONE launch; ~128 per thread keep the tail
        // (threads that run out of slots while others still walk theirs) under a percent of the batch
        const bool small = scene->small_scene && !std::getenv("RPTX_NO_SMALL_SCENE");
        f->persist_blocks = rptx_host::persist_grid(ctx->sm_count, scene->medium_kind != RPTX_MEDIUM_NONE, scene->shadow_probe_mask != 0, small);
        if (const char* e = std::getenv("RPTX_PERSIST_BLOCKS")) f->persist_blocks = std::max(1, std::atoi(e));  // developer knob
        target = 128ull * (uint64_t)f->persist_blocks * kPersistBlock;
    }
    if (const char* e = std::getenv("RPTX_POOL_LOG2")) {  // developer knob: paths in flight
        const int v = std::atoi(e);
        if (v >= 10 && v <= 26) target = 1ull << v;
    }
    uint64_t lanes = std::max<uint64_t>(1, target / n_pixels);
    if (f->mode == kFilmPersist) lanes = std::max<uint64_t>(1, (target + n_pixels - 1) / n_pixels);
    lanes = std::min<uint64_t>(lanes, 4096);
    lanes = std::min<uint64_t>(lanes, std::max<uint32_t>(params->num_samples, 1));
    p.n_lanes = (uint32_t)lanes;
    p.n_slots = (uint32_t)(n_pixels * lanes);
    if (film_alloc(f)) { rptx_film_destroy(f); return 1; }
    *out = f;
    return 0;
}

int rptx_render_begin(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera, const rptx_params* params, rptx_film** out) {
    return film_begin(ctx, scene, camera, params, false, out);
}

int rptx_photon_gather_begin(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera, const rptx_params* params, rptx_film** out) {
    return film_begin(ctx, scene, camera, params, true, out);
}

void rptx_film_destroy(rptx_film* f) {
    if (!f) return;
    cudaSetDevice(f->ctx->device);
    for (void* p : f->owned) cudaFree(p);
    for (double* b : f->pooled_means) cudaFree(b);
    if (f->arena) {  // the slab goes back to the context for the next film (one spare at most, the larger one)
        cudaStreamSynchronize(f->ctx->stream);
        rptx_ctx* c = f->ctx;
        if (f->arena_bytes > c->spare_arena_bytes) {
            if (c->spare_arena) cudaFree(c->spare_arena);
            c->spare_arena = f->arena; c->spare_arena_bytes = f->arena_bytes;
        } else {
            if (std::getenv("RPTX_TRACE_ALLOC")) fprintf(stderr, "[rptx] film slab: %zu B freed (spare holds %zu B)\n", f->arena_bytes, c->spare_arena_bytes);
            cudaFree(f->arena);
        }
    }
    if (f->d_batch_ptrs) cudaFree(f->d_batch_ptrs);
    if (f->ev0) cudaEventDestroy(f->ev0);
    if (f->ev1) cudaEventDestroy(f->ev1);
    for (int r = 0; r < 4; r++) {
        for (int k = 0; k < 5; k++) if (f->evw[r][k]) cudaEventDestroy(f->evw[r][k]);
        if (f->ev_done[r]) cudaEventDestroy(f->ev_done[r]);
    }
    delete f;
}

}  // extern "C"

// The wave loop.  The host never waits for the wave it just launched: the live count of wave w is copied to pinned memory
// behind it and read kWaveLag waves later, so the launches of waves w+1 .. w+kWaveLag are already queued when wave w ends
// (round 1 synchronised the stream once per wave: a launch-latency bubble per wave, 97 waves per lampshade batch).  Counts
// never grow (a finished path's slot either regenerates in place or leaves the queue), so a count that is kWaveLag waves old
// is an upper bound for the grid of the next wave; every kernel bounds itself by the device-side count, and k_connect /
// k_shade_final clear the keep flags between the true count and the bound for the in-order compaction.  The loop ends when
// a lagged count reads zero, i.e. after at most kWaveLag empty waves.
constexpr int kWaveLag = 2;
constexpr int kWaveRing = 4;

template <bool MEDIUM>
static int run_waves(rptx_film* f, const BatchParams& bp) {
    rptx_ctx* ctx = f->ctx;
    cudaStream_t st = ctx->stream;
    const SceneDev sc = scene_dev(f->scene);
    const MediumDev md = medium_dev(f->scene);
    PathPool& p = f->pool;
    Queues q = f->q;
    CK(cudaMemsetAsync(q.counters, 0, sizeof(uint32_t) * 8, st));
    k_generate<MEDIUM><<<grid_for(p.n_slots, kBlock), kBlock, 0, st>>>(p, q, bp, f->cam);
    ctx->launches++;
    const int persistent_blocks = ctx->sm_count * 8;
    const bool small = sc.small_scene != 0;
    const bool one_vertex = !MEDIUM && bp.max_bounces == 0 && !f->no_one_vertex;  // k_shade_final
    uint32_t* h_count = (uint32_t*)ctx->h_pinned;  // ring of kWaveRing counts
    uint32_t bound = std::min<uint64_t>(p.n_slots, (uint64_t)p.n_pixels * std::min(p.n_lanes, bp.n_local));
    auto collect = [&](int slot) -> uint32_t {  // wait for the wave in ring slot `slot`, add its kernel times, return its live count
        cudaEventSynchronize(f->ev_done[slot]);
        cudaEvent_t* e = f->evw[slot];
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, e[0], e[1]) == cudaSuccess) f->stats.extend_ms += ms;
        if (cudaEventElapsedTime(&ms, e[1], e[2]) == cudaSuccess) f->stats.shade_ms += ms;
        if (cudaEventElapsedTime(&ms, e[2], e[3]) == cudaSuccess) f->stats.shadow_ms += ms;
        if (cudaEventElapsedTime(&ms, e[3], e[4]) == cudaSuccess) f->stats.connect_ms += ms;
        return h_count[slot];
    };
    const uint32_t shade_rounds_env = f->shade_rounds;
    int launched = 0, collected = 0;
    while (bound > 0) {
        const int slot = launched % kWaveRing;
        cudaEvent_t* e = f->evw[slot];
        const int blocks = (int)std::min<uint64_t>(persistent_blocks, grid_for(bound, kBlock));
        cudaEventRecord(e[0], st);
        if (small) k_extend<true><<<blocks, kBlock, 0, st>>>(sc, p, q);
        else k_extend<false><<<blocks, kBlock, 0, st>>>(sc, p, q);
        cudaEventRecord(e[1], st);
        if (one_vertex) {
            // the vertex ends the path: shade + NEE visibility + film add + regeneration in one kernel
            const int g = (int)grid_for(bound, kBlock);
            if (small) {
                if (sc.shadow_probe_mask) k_shade_final<true, true><<<g, kBlock, 0, st>>>(sc, p, q, bp, f->cam);
                else k_shade_final<false, true><<<g, kBlock, 0, st>>>(sc, p, q, bp, f->cam);
            } else {
                if (sc.shadow_probe_mask) k_shade_final<true, false><<<g, kBlock, 0, st>>>(sc, p, q, bp, f->cam);
                else k_shade_final<false, false><<<g, kBlock, 0, st>>>(sc, p, q, bp, f->cam);
            }
            cudaEventRecord(e[2], st);
            cudaEventRecord(e[3], st);
            ctx->launches += 2;
        } else {
            // class-sorted tile of `rounds` x kShadeBlock entries per block; a wave too small for ~4 rounds of resident blocks
            // at that tile size keeps shorter tiles
            // (measured, 1xB200: lampshade k_shade 368 / 284 / 259 / 235 ms per 256 spp at 1 / 2 / 4 / 8 rounds, 24.7 live lanes
            // at 8; without a medium nearly every Cornell vertex is a surface hit and longer tiles only cost parallelism:
            // 2.8 -> 3.1 ms, so that mode keeps one round)
            const uint32_t rounds = shade_rounds_env ? shade_rounds_env
                                    : !MEDIUM        ? 1u
                                                     : (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>(kShadeRoundsMax, bound / ((uint64_t)ctx->sm_count * RPTX_OCC_SHADE * 4 * kShadeBlock)));
            k_shade<MEDIUM><<<grid_for(bound, kShadeBlock * rounds), kShadeBlock, 0, st>>>(sc, p, q, bp, md, rounds);
            cudaEventRecord(e[2], st);
            if (p.n_shadow) {
                if (small) {
                    if (sc.shadow_probe_mask) k_shadow_warp<true, true><<<blocks, kBlock, 0, st>>>(sc, p, q);
                    else k_shadow_warp<false, true><<<blocks, kBlock, 0, st>>>(sc, p, q);
                } else {
                    if (sc.shadow_probe_mask) k_shadow_block<true, false><<<blocks, kBlock, 0, st>>>(sc, p, q);
                    else k_shadow_block<false, false><<<blocks, kBlock, 0, st>>>(sc, p, q);
                }
            }
            cudaEventRecord(e[3], st);
            k_connect<MEDIUM><<<grid_for(bound, kBlock), kBlock, 0, st>>>(p, q, bp, f->cam);
            ctx->launches += p.n_shadow ? 4 : 3;
        }
        {  // next = kept entries of cur, in order; the count lands in counters[1]
            size_t bytes = f->select_tmp_bytes;
            CK(cub::DeviceSelect::Flagged(f->d_select_tmp, bytes, q.cur, q.keep, q.next, q.counters + 1, (int)bound, st));
            ctx->launches += 2;
        }
        cudaEventRecord(e[4], st);
        k_wave_end<<<1, 1, 0, st>>>(q);
        ctx->launches++;
        f->stats.waves++;
        std::swap(q.cur, q.next);
        CK(cudaMemcpyAsync(h_count + slot, q.counters, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        CK(cudaEventRecord(f->ev_done[slot], st));
        launched++;
        if (launched - collected > kWaveLag - 1) bound = std::min(bound, collect(collected++ % kWaveRing));
    }
    while (collected < launched) collect(collected++ % kWaveRing);
    CK(cudaGetLastError());
    return 0;
}

extern "C" {

static void film_collect_timing(rptx_film* f);

// `sync`: the public split form hands the batch sums to the caller's own exchange, so it returns with the stream idle;
// rptx_render_sample keeps everything stream-ordered (trace -> NCCL all-reduce -> commit) and synchronises once, in commit
static int sample_trace_impl(rptx_film* f, uint32_t iterations, bool sync) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (iterations == 0) return ctx_fail(ctx, "iterations must be > 0");
    if (f->mode == kFilmGather) return ctx_fail(ctx, "this film belongs to a photon gather pass (rptx_photon_gather_begin)");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    BatchParams bp;
    bp.width = f->params.width; bp.height = f->params.height; bp.max_bounces = f->params.max_bounces;
    bp.seed = f->params.seed; bp.batch_base = f->next_sample; bp.iterations = iterations;
    bp.shard_count = f->params.shard_count; bp.shard_index = f->params.shard_index;
    bp.n_local = iterations > bp.shard_index ? (iterations - bp.shard_index + bp.shard_count - 1) / bp.shard_count : 0;
    CK(cudaEventRecord(f->ev0, st));
    f->stats.extend_ms = f->stats.shade_ms = f->stats.connect_ms = f->stats.shadow_ms = 0.0;
    PathPool& p = f->pool;
    const uint32_t full_lanes = p.n_lanes;
    p.n_lanes = std::max<uint32_t>(1, std::min(full_lanes, bp.n_local));
    int rc = 0;
    if (f->mode == kFilmPersist) {
        const SceneDev sc = scene_dev(f->scene);
        const MediumDev md = medium_dev(f->scene);
        PersistArgs pa;
        std::memset(&pa, 0, sizeof(pa));
        pa.cam = f->cam; pa.bp = bp; pa.n_pixels = p.n_pixels; pa.n_lanes = p.n_lanes; pa.n_slots = p.n_pixels * p.n_lanes;
        pa.width = p.width; pa.tiled = p.tiled; pa.stack_depth = p.stack_depth;
        pa.ax = p.ax; pa.ay = p.ay; pa.az = p.az; pa.stack = f->d_pstack; pa.next_slot = f->q.counters;
        if (bp.n_local == 0) {  // this rank holds no sample of the batch
            for (double* a : {p.ax, p.ay, p.az}) CK(cudaMemsetAsync(a, 0, sizeof(double) * pa.n_slots, st));
        } else {
            CK(cudaMemsetAsync(f->q.counters, 0, sizeof(uint32_t) * 8, st));
            cudaError_t e = rptx_host::launch_path_persist(f->scene->medium_kind != RPTX_MEDIUM_NONE, sc.shadow_probe_mask != 0, sc.small_scene != 0,
                                                           f->persist_blocks, sc, md, pa, f->q.stats, st);
            ctx->launches++;
            f->stats.waves++;
            if (e != cudaSuccess) rc = ctx_fail(ctx, std::string("k_path_persist: ") + cudaGetErrorString(e));
        }
    } else if (f->mode == kFilmFused) {
        FusedArgs fa;
        fa.cam = f->cam; fa.bp = bp; fa.n_pixels = p.n_pixels; fa.n_lanes = p.n_lanes;
        fa.ax = p.ax; fa.ay = p.ay; fa.az = p.az;
        const SceneDev sc = scene_dev(f->scene);
        const MediumDev md = medium_dev(f->scene);
        const size_t threads = (size_t)p.n_pixels * p.n_lanes;
        if (f->scene->medium_kind == RPTX_MEDIUM_NONE) k_path_fused<false><<<grid_for(threads, kBlock), kBlock, 0, st>>>(sc, md, fa, f->q.stats);
        else k_path_fused<true><<<grid_for(threads, kBlock), kBlock, 0, st>>>(sc, md, fa, f->q.stats);
        ctx->launches++;
        f->stats.waves++;
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) rc = ctx_fail(ctx, std::string("k_path_fused: ") + cudaGetErrorString(e));
    } else {
        rc = f->scene->medium_kind == RPTX_MEDIUM_NONE ? run_waves<false>(f, bp) : run_waves<true>(f, bp);
    }
    if (rc == 0) {
        k_batch_reduce<<<grid_for(p.n_pixels, 128), 128, 0, st>>>(p, f->d_batch_sum);
        ctx->launches++;
    }
    p.n_lanes = full_lanes;
    if (rc) return rc;
    CK(cudaEventRecord(f->ev1, st));
    f->timing_pending = true;  // read at the next synchronisation point (commit / stats): no host sync between trace and exchange
    if (sync) { CK(cudaStreamSynchronize(st)); film_collect_timing(f); }
    return 0;
}

int rptx_render_sample_trace(rptx_film* f, uint32_t iterations) { return sample_trace_impl(f, iterations, true); }

// device time of the last traced batch, once its events have completed
static void film_collect_timing(rptx_film* f) {
    if (!f->timing_pending) return;
    f->timing_pending = false;
    float ms = 0.f;
    if (cudaEventSynchronize(f->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, f->ev0, f->ev1) == cudaSuccess) f->stats.device_ms = ms;
    f->stats.trace_ms = f->stats.extend_ms + f->stats.shadow_ms;
}

int rptx_render_sample_commit(rptx_film* f, uint32_t iterations) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (iterations == 0) return ctx_fail(ctx, "iterations must be > 0");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t n3 = (size_t)f->pool.n_pixels * 3;
    double* mean = nullptr;
    if (f->batches.size() < f->n_mean_slots) {
        mean = f->d_mean_slots + f->batches.size() * n3;
    } else {
        CK(cudaMallocAsync((void**)&mean, sizeof(double) * n3, st));  // stream-ordered pool: no device-wide lock on the hot path
        f->pooled_means.push_back(mean);
    }
    f->batches.push_back(mean);
    const double exposure = std::pow(2.0, f->params.exposure_value);
    k_batch_commit<<<grid_for(n3, 256), 256, 0, st>>>(f->d_batch_sum, n3, (double)iterations, exposure, mean, f->d_film_sum);
    ctx->launches++;
    CK(cudaStreamSynchronize(st));
    film_collect_timing(f);
    f->next_sample += iterations;
    return 0;
}

int rptx_render_sample(rptx_film* f, uint32_t iterations) {
    // a sharded film holds 1 / shard_count of the batch: committing it without the exchange would dim the image
    if (f && f->params.shard_count > 1 && !f->ctx->comm)
        return ctx_fail(f->ctx, "sharded films must use rptx_render_sample_trace + all-reduce + rptx_render_sample_commit, or a context with rptx_comm_init");
    if (int rc = sample_trace_impl(f, iterations, false)) return rc;
    if (f->params.shard_count > 1)
        if (int rc = rptx_render_sample_allreduce(f)) return rc;
    return rptx_render_sample_commit(f, iterations);
}

void* rptx_film_batch_sum_device_ptr(rptx_film* f) { return f ? (void*)f->d_batch_sum : nullptr; }

int rptx_film_image_rgb8(rptx_film* f, uint8_t* out) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (!out) return ctx_fail(ctx, "null output");
    if (f->batches.empty()) return ctx_fail(ctx, "Pixel found with no samples");  // src/buffer.rs:89
    CK(cudaSetDevice(ctx->device));
    const uint32_t w = f->params.width, h = f->params.height;
    uint8_t* d_out = reinterpret_cast<uint8_t*>(f->d_scratch);
    dim3 blk(16, 16), grd((w + 15) / 16, (h + 15) / 16);
    k_image<<<grd, blk, 0, ctx->stream>>>(f->d_film_sum, w, h, f->params.filter_radius, (uint32_t)f->batches.size(), d_out);
    ctx->launches++;
    cudaError_t e = cudaMemcpyAsync(out, d_out, (size_t)w * h * 3, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    
    if (e != cudaSuccess) return ctx_fail(ctx, cudaGetErrorString(e));
    return 0;
}

int rptx_film_variance(rptx_film* f, double* out) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (!out) return ctx_fail(ctx, "null output");
    CK(cudaSetDevice(ctx->device));
    const uint32_t nb = (uint32_t)f->batches.size();
    if (nb == 0) { *out = NAN; return 0; }
    if (f->d_batch_ptrs) { rptx_free_async(f->d_batch_ptrs, ctx->stream); f->d_batch_ptrs = nullptr; }
    CK(rptx_alloc_async((void**)&f->d_batch_ptrs, sizeof(double*) * nb, ctx->stream));
    CK(copy_h2d(f->d_batch_ptrs, f->batches.data(), sizeof(double*) * nb, ctx->stream));
    double* d_acc = nullptr;
    CK(rptx_alloc_async((void**)&d_acc, sizeof(double), ctx->stream));
    CK(cudaMemsetAsync(d_acc, 0, sizeof(double), ctx->stream));
    k_variance<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(f->d_batch_ptrs, nb, f->d_film_sum, f->pool.n_pixels, d_acc);
    ctx->launches++;
    double acc = 0.0;
    cudaError_t e = cudaMemcpyAsync(&acc, d_acc, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    rptx_free_async(d_acc, ctx->stream);
    if (e != cudaSuccess) return ctx_fail(ctx, cudaGetErrorString(e));
    *out = acc / (double)f->pool.n_pixels;
    return 0;
}

int rptx_film_mean(rptx_film* f, double* out) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (!out) return ctx_fail(ctx, "null output");
    if (f->batches.empty()) return ctx_fail(ctx, "no batches rendered");
    CK(cudaSetDevice(ctx->device));
    const size_t n3 = (size_t)f->pool.n_pixels * 3;
    double* d_out = f->d_scratch;
    k_film_mean<<<grid_for(n3, 256), 256, 0, ctx->stream>>>(f->d_film_sum, n3, (double)f->batches.size(), d_out);
    ctx->launches++;
    cudaError_t e = cudaMemcpyAsync(out, d_out, sizeof(double) * n3, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    
    if (e != cudaSuccess) return ctx_fail(ctx, cudaGetErrorString(e));
    return 0;
}

int rptx_stats_get(const rptx_film* f, rptx_stats* out) {
    if (!f || !out) return 1;
    cudaSetDevice(f->ctx->device);
    unsigned long long st[8] = {0};
    cudaStreamSynchronize(f->ctx->stream);
    film_collect_timing(const_cast<rptx_film*>(f));
    cudaMemcpy(st, f->q.stats, sizeof(st), cudaMemcpyDeviceToHost);
    *out = f->stats;
    out->path_samples = st[0]; out->extend_rays = st[1]; out->shadow_rays = st[2];
    out->shadow_traced = st[3]; out->shadow_occlusion = st[4];
    out->gather_queries = st[5]; out->gather_visibility = st[6]; out->gather_candidates = st[7];
    out->path_mode = (uint64_t)f->mode;
    out->kernel_launches = f->ctx->launches;
    return 0;
}

// ---------------------------------------------------------------- parity hooks

int rptx_intersect_batch_device(rptx_scene* s, const void* d_o, const void* d_d, size_t n, double t_min, void* d_t,
                                void* d_object_id, void* d_prim_id, void* d_inst_id, void* d_normal) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    if (!s->committed) return ctx_fail(ctx, "scene not committed");
    if (n == 0) return 0;
    if (!d_o || !d_d) return ctx_fail(ctx, "null ray arrays");
    CK(cudaSetDevice(ctx->device));
    const SceneDev sc = scene_dev(s);
    k_intersect_batch<<<grid_for(n, kBlock), kBlock, 0, ctx->stream>>>(sc, (const double*)d_o, (const double*)d_d, n, t_min, (double*)d_t,
                                                                        (uint32_t*)d_object_id, (uint32_t*)d_prim_id,
                                                                        (uint32_t*)d_inst_id, (double*)d_normal);
    ctx->launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rptx_intersect_batch(rptx_scene* s, const double* o, const double* d, size_t n, double t_min, double* t, uint32_t* object_id,
                         uint32_t* prim_id, uint32_t* inst_id, double* normal) {
    if (!s) return 1;
    rptx_ctx* ctx = s->ctx;
    if (!s->committed) return ctx_fail(ctx, "scene not committed");
    if (n == 0) return 0;
    if (!o || !d) return ctx_fail(ctx, "null ray arrays");
    CK(cudaSetDevice(ctx->device));
    double *d_o = nullptr, *d_d = nullptr, *d_t = nullptr, *d_n = nullptr;
    uint32_t *d_obj = nullptr, *d_prim = nullptr, *d_inst = nullptr;
    int rc = 0;
    auto cleanup = [&] { cudaFree(d_o); cudaFree(d_d); cudaFree(d_t); cudaFree(d_n); cudaFree(d_obj); cudaFree(d_prim); cudaFree(d_inst); };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while (0)
    CKC(cudaMalloc((void**)&d_o, sizeof(double) * 3 * n));
    CKC(cudaMalloc((void**)&d_d, sizeof(double) * 3 * n));
    CKC(cudaMalloc((void**)&d_t, sizeof(double) * n));
    CKC(cudaMalloc((void**)&d_n, sizeof(double) * 3 * n));
    CKC(cudaMalloc((void**)&d_obj, sizeof(uint32_t) * n));
    CKC(cudaMalloc((void**)&d_prim, sizeof(uint32_t) * n));
    CKC(cudaMalloc((void**)&d_inst, sizeof(uint32_t) * n));
    CKC(copy_h2d(d_o, o, sizeof(double) * 3 * n, ctx->stream));
    CKC(copy_h2d(d_d, d, sizeof(double) * 3 * n, ctx->stream));
    rc = rptx_intersect_batch_device(s, d_o, d_d, n, t_min, d_t, d_obj, d_prim, d_inst, d_n);
    if (rc == 0) {
        if (t) CKC(cudaMemcpy(t, d_t, sizeof(double) * n, cudaMemcpyDeviceToHost));
        if (object_id) CKC(cudaMemcpy(object_id, d_obj, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
        if (prim_id) CKC(cudaMemcpy(prim_id, d_prim, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
        if (inst_id) CKC(cudaMemcpy(inst_id, d_inst, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost));
        if (normal) CKC(cudaMemcpy(normal, d_n, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost));
    }
#undef CKC
    cleanup();
    return rc;
}

}  // extern "C"
