// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// C entry points over the CPU restatement, consumed through ctypes by tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs ONLY.
// They take the same POD records as include/rptx.h so that one scene description
// drives both the CUDA library and this checker.
#include <chrono>
#include <cstring>
#include <string>

#include "../include/rptx.h"
#include "builtin_scenes.hpp"
#include "photon.hpp"
#include "render.hpp"

using namespace orc;

namespace {

struct OScene {
    Scene scene;
    std::vector<std::shared_ptr<const Shape>> meshes;
    std::vector<bool> object_is_group;
    std::string err;
};

struct OFilm {
    OScene* scene;
    Renderer r;
    Buffer buf;
    uint64_t next_sample = 0;
    Counters ctr;
    double last_seconds = 0;
    OFilm(OScene* s, uint32_t w, uint32_t h, uint32_t fr) : scene(s), buf(w, h, fr) {}
};

Mat4 to_mat4(const double* t) {
    Mat4 m;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) m.m[r][c] = t[4 * r + c];
    return m;
}

Material to_material(const rptx_material* m) {
    Material out;
    out.kind = (Material::Kind)m->kind;
    out.albedo = Vec3(m->albedo[0], m->albedo[1], m->albedo[2]);
    out.emittance_ = m->emittance;
    out.shininess = m->shininess;
    out.ior = m->ior;
    return out;
}

std::shared_ptr<const Shape> build_shape(OScene* s, const rptx_shape_desc* d, bool* is_group) {
    std::shared_ptr<const Shape> base;
    if (is_group) *is_group = false;
    switch (d->kind) {
        case RPTX_SHAPE_SPHERE: base = std::make_shared<Sphere>(); break;
        case RPTX_SHAPE_CUBE: base = std::make_shared<Cube>(); break;
        case RPTX_SHAPE_PLANE:
            base = std::make_shared<Plane>(Vec3(d->plane_normal[0], d->plane_normal[1], d->plane_normal[2]), d->plane_value);
            break;
        case RPTX_SHAPE_MONOMIAL: base = std::make_shared<MonomialSurface>(d->mono_height, d->mono_exp); break;
        case RPTX_SHAPE_MESH:
            if (d->mesh_id >= s->meshes.size()) { s->err = "bad mesh id"; return nullptr; }
            base = s->meshes[d->mesh_id];
            break;
        case RPTX_SHAPE_GROUP: {
            std::vector<std::shared_ptr<const Shape>> kids;
            for (uint32_t i = 0; i < d->n_children; i++) {
                if (d->children[i].kind == RPTX_SHAPE_GROUP || d->children[i].kind == RPTX_SHAPE_PLANE) {
                    s->err = "group children must be bounded leaf shapes";
                    return nullptr;
                }
                auto k = build_shape(s, &d->children[i], nullptr);
                if (!k) return nullptr;
                kids.push_back(k);
            }
            base = std::make_shared<KdTree>(std::move(kids), true);
            if (is_group) *is_group = true;
            break;
        }
        default: s->err = "bad shape kind"; return nullptr;
    }
    if (d->has_transform) base = std::make_shared<Transformed>(base, to_mat4(d->transform));
    return base;
}

}  // namespace

extern "C" {

void* orc_scene_create() { return new OScene(); }
void orc_scene_destroy(void* s) { delete (OScene*)s; }
const char* orc_last_error(void* s) { return ((OScene*)s)->err.c_str(); }

// A BASELINE scene built natively from the example sources' literals (builtin_scenes.hpp), not from lowered records.
// name: cornell | teapot | fractal_teapots | lampshade_{pathtrace,photonphoton,beamphoton,beambeam}; the teapot mesh (the
// reference's examples/teapot.obj as triangles) is input data for the two teapot scenes.
void* orc_scene_builtin(const char* name, const double* teapot_v, const double* teapot_n, size_t ntri) {
    OScene* s = new OScene();
    const std::string n = name ? name : "";
    const char* variants[4] = {"lampshade_pathtrace", "lampshade_photonphoton", "lampshade_beamphoton", "lampshade_beambeam"};
    bool ok = true;
    if (n == "cornell") {
        builtin::cornell(s->scene);
    } else if (n == "teapot" || n == "fractal_teapots") {
        if (!teapot_v || !teapot_n || !ntri) { delete s; return nullptr; }
        auto mesh = builtin::mesh_from(teapot_v, teapot_n, ntri);
        if (n == "teapot") builtin::teapot(s->scene, mesh);
        else builtin::fractal_teapots(s->scene, mesh, &s->object_is_group);
    } else {
        ok = false;
        for (int v = 0; v < 4; v++)
            if (n == variants[v]) { builtin::lampshade(s->scene, v); ok = true; }
    }
    if (!ok) { delete s; return nullptr; }
    s->object_is_group.resize(s->scene.objects.size(), false);
    return s;
}

int orc_scene_add_mesh(void* sp, const double* v, const double* n, size_t ntri, uint32_t* mesh_id) {
    OScene* s = (OScene*)sp;
    std::vector<std::shared_ptr<const Shape>> tris;
    tris.reserve(ntri);
    for (size_t i = 0; i < ntri; i++) {
        const double* p = v + 9 * i;
        auto t = std::make_shared<Triangle>();
        if (n) {
            const double* q = n + 9 * i;
            t->v1 = Vec3(p[0], p[1], p[2]); t->v2 = Vec3(p[3], p[4], p[5]); t->v3 = Vec3(p[6], p[7], p[8]);
            t->n1 = Vec3(q[0], q[1], q[2]); t->n2 = Vec3(q[3], q[4], q[5]); t->n3 = Vec3(q[6], q[7], q[8]);
        } else {
            *t = Triangle::from_vertices(Vec3(p[0], p[1], p[2]), Vec3(p[3], p[4], p[5]), Vec3(p[6], p[7], p[8]));
        }
        tris.push_back(t);
    }
    s->meshes.push_back(std::make_shared<KdTree>(std::move(tris), false));
    *mesh_id = (uint32_t)s->meshes.size() - 1;
    return 0;
}

int orc_scene_add_object(void* sp, const rptx_shape_desc* d, const rptx_material* m, uint32_t* id) {
    OScene* s = (OScene*)sp;
    bool grp = false;
    auto sh = build_shape(s, d, &grp);
    if (!sh) return 1;
    s->scene.objects.push_back({sh, to_material(m)});
    s->object_is_group.push_back(grp);
    if (id) *id = (uint32_t)s->scene.objects.size() - 1;
    return 0;
}

int orc_scene_add_light(void* sp, const rptx_light* l) {
    OScene* s = (OScene*)sp;
    Light out;
    out.kind = (Light::Kind)l->kind;
    out.color = Vec3(l->color[0], l->color[1], l->color[2]);
    out.vec = Vec3(l->vec[0], l->vec[1], l->vec[2]);
    if (l->kind == RPTX_LIGHT_OBJECT) {
        auto sh = build_shape(s, &l->shape, nullptr);
        if (!sh) return 1;
        out.object = {sh, to_material(&l->material)};
    }
    s->scene.lights.push_back(out);
    return 0;
}

int orc_scene_set_medium(void* sp, int32_t kind, double a, double sc) {
    OScene* s = (OScene*)sp;
    s->scene.media.clear();
    if (kind == RPTX_MEDIUM_NONE) return 0;
    Medium m;
    m.kind = (Medium::Kind)kind; m.absorption_ = a; m.scattering_ = sc;
    s->scene.media.push_back(m);
    return 0;
}

int orc_scene_set_environment_color(void* sp, const double rgb[3]) {
    OScene* s = (OScene*)sp;
    s->scene.environment.is_hdri = false;
    s->scene.environment.color = Vec3(rgb[0], rgb[1], rgb[2]);
    return 0;
}

int orc_scene_set_environment_hdri(void* sp, uint32_t w, uint32_t h, const double* px) {
    OScene* s = (OScene*)sp;
    auto& e = s->scene.environment;
    e.is_hdri = true; e.width = w; e.height = h;
    e.buf.resize((size_t)w * h);
    for (size_t i = 0; i < e.buf.size(); i++) e.buf[i] = Vec3(px[3 * i], px[3 * i + 1], px[3 * i + 2]);
    return 0;
}

// kd-tree statistics of a mesh (nodes, leaves, leaf items) — informational
int orc_mesh_kd_stats(void* sp, uint32_t mesh_id, uint64_t out[3]) {
    OScene* s = (OScene*)sp;
    if (mesh_id >= s->meshes.size()) return 1;
    auto* k = (const KdTree*)s->meshes[mesh_id].get();
    out[0] = k->n_nodes; out[1] = k->n_leaves; out[2] = k->n_leaf_items;
    return 0;
}

static void fill_renderer(Renderer& r, OScene* s, const rptx_camera* c, const rptx_params* p, int threads) {
    r.scene = &s->scene;
    r.camera.eye = Vec3(c->eye[0], c->eye[1], c->eye[2]);
    r.camera.direction = Vec3(c->direction[0], c->direction[1], c->direction[2]);
    r.camera.up = Vec3(c->up[0], c->up[1], c->up[2]);
    r.camera.fov = c->fov; r.camera.aperture = c->aperture; r.camera.focal_distance = c->focal_distance;
    r.width = p->width; r.height = p->height; r.exposure_value = p->exposure_value;
    r.filter_radius = p->filter_radius; r.max_bounces = p->max_bounces; r.num_samples = p->num_samples;
    r.gather_size = p->gather_size; r.gather_size_volume = p->gather_size_volume; r.watts = p->watts;
    r.seed = p->seed; r.threads = threads;
}

void* orc_render_begin(void* sp, const rptx_camera* c, const rptx_params* p, int threads) {
    OScene* s = (OScene*)sp;
    OFilm* f = new OFilm(s, p->width, p->height, p->filter_radius);
    fill_renderer(f->r, s, c, p, threads);
    return f;
}
void orc_film_destroy(void* f) { delete (OFilm*)f; }

int orc_render_sample(void* fp, uint32_t iterations) {
    OFilm* f = (OFilm*)fp;
    const auto t0 = std::chrono::steady_clock::now();
    f->r.sample(f->next_sample, iterations, f->buf, &f->ctr, nullptr);
    f->last_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    f->next_sample += iterations;
    return 0;
}
// Render only rows y0, y0+stride, ... < y1 of one batch without touching the Buffer (bounded
// timing sample of the CPU baseline).
double orc_time_rows(void* fp, uint32_t y0, uint32_t y1, uint32_t stride, uint32_t iterations, uint64_t counters[4]) {
    OFilm* f = (OFilm*)fp;
    Renderer r = f->r;
    const int nt = r.threads > 0 ? r.threads : std::max(1, (int)std::thread::hardware_concurrency());
    std::vector<Counters> ctrs(nt);
    if (stride == 0) stride = 1;
    std::atomic<uint32_t> next{0};
    std::vector<Vec3> sink(nt);
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; t++)
        pool.emplace_back([&, t] {
            for (;;) {
                const uint32_t y = y0 + next.fetch_add(1) * stride;
                if (y >= y1) break;
                for (uint32_t x = 0; x < r.width; x++) sink[t] += r.get_color(x, y, 0, iterations, &ctrs[t], nullptr);
            }
        });
    for (auto& th : pool) th.join();
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    Counters tot;
    for (auto& c : ctrs) { tot.extend_rays += c.extend_rays; tot.shadow_rays += c.shadow_rays; tot.shadow_pass += c.shadow_pass; tot.shadow_near += c.shadow_near; }
    if (counters) { counters[0] = tot.extend_rays; counters[1] = tot.shadow_rays; counters[2] = tot.shadow_pass; counters[3] = tot.shadow_near; }
    return sec;
}
double orc_film_last_seconds(void* fp) { return ((OFilm*)fp)->last_seconds; }
int orc_film_counters(void* fp, uint64_t out[4]) {
    OFilm* f = (OFilm*)fp;
    out[0] = f->ctr.extend_rays; out[1] = f->ctr.shadow_rays; out[2] = f->ctr.shadow_pass; out[3] = f->ctr.shadow_near;
    return 0;
}
int orc_film_image_rgb8(void* fp, uint8_t* out) { ((OFilm*)fp)->buf.image(out); return 0; }
int orc_film_variance(void* fp, double* out) { *out = ((OFilm*)fp)->buf.variance(); return 0; }
int orc_film_mean(void* fp, double* out) { ((OFilm*)fp)->buf.mean(out); return 0; }

// per-sample radiance (no exposure, no averaging) for listed (pixel, sample) pairs
int orc_sample_radiance(void* fp, const uint32_t* px, const uint32_t* py, const uint64_t* sample, size_t n, double* out) {
    OFilm* f = (OFilm*)fp;
    for (size_t i = 0; i < n; i++) {
        const Color c = f->r.sample_radiance(px[i], py[i], sample[i]);
        out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
    }
    return 0;
}

// dump up to cap rays (every closest-hit query) of one batch rendered serially
size_t orc_dump_rays(void* fp, uint32_t iterations, size_t cap, double* o, double* d, uint8_t* kind) {
    OFilm* f = (OFilm*)fp;
    RayLog log;
    log.cap = cap;
    Buffer tmp(f->r.width, f->r.height, 0);
    f->r.sample(0, iterations, tmp, nullptr, &log);
    const size_t n = log.kind.size();
    std::memcpy(o, log.o.data(), n * 3 * sizeof(double));
    std::memcpy(d, log.d.data(), n * 3 * sizeof(double));
    std::memcpy(kind, log.kind.data(), n);
    return n;
}

// get_closest_hit over a batch (threads > 1 splits the batch)
int orc_intersect_batch(void* sp, const double* o, const double* d, size_t n, double t_min, double* t, uint32_t* object_id,
                        uint32_t* prim_id, uint32_t* inst_id, double* normal, int threads) {
    OScene* s = (OScene*)sp;
    int nt = threads > 0 ? threads : 1;
    auto work = [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; i++) {
            const Ray ray{Vec3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), Vec3(d[3 * i], d[3 * i + 1], d[3 * i + 2])};
            HitRecord h;
            uint32_t id = 0xffffffffu;
            for (size_t k = 0; k < s->scene.objects.size(); k++)
                if (s->scene.objects[k].shape->intersect(ray, t_min, h)) id = (uint32_t)k;
            const bool hit = id != 0xffffffffu;
            if (t) t[i] = hit ? h.time : INF;
            if (object_id) object_id[i] = id;
            if (prim_id) prim_id[i] = hit ? h.prim_id : 0;
            if (inst_id) inst_id[i] = (hit && s->object_is_group[id]) ? h.inst_id : 0;
            if (normal) { normal[3 * i] = h.normal.x; normal[3 * i + 1] = h.normal.y; normal[3 * i + 2] = h.normal.z; }
        }
    };
    if (nt == 1) { work(0, n); return 0; }
    std::vector<std::thread> pool;
    for (int k = 0; k < nt; k++) pool.emplace_back(work, n * k / nt, n * (k + 1) / nt);
    for (auto& th : pool) th.join();
    return 0;
}

// known-answer helpers for tests
void orc_hex_color(uint32_t x, double out[3]) { Color c = hex_color(x); out[0] = c.x; out[1] = c.y; out[2] = c.z; }
void orc_color_bytes(const double c[3], uint8_t out[3]) { color_bytes(Vec3(c[0], c[1], c[2]), out); }
uint64_t orc_rng_u64(uint64_t seed, uint64_t domain, uint64_t a, uint64_t b, uint32_t skip) {
    Rng r(seed, domain, a, b);
    for (uint32_t i = 0; i < skip; i++) r.next_u64();
    return r.next_u64();
}
int orc_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

// photon_render (src/photon.rs:655-720)
int orc_photon_render(void* sp, const rptx_camera* c, const rptx_params* p, uint64_t photon_count, int32_t kind, int threads,
                      uint8_t* out_rgb8, double* out_mean, uint64_t counts[2], double seconds[3]) {
    OScene* s = (OScene*)sp;
    Renderer r;
    fill_renderer(r, s, c, p, threads);
    PhotonStats st;
    Buffer buf(p->width, p->height, p->filter_radius);
    if (!photon_render(r, photon_count, (PhotonRenderKind)kind, buf, &st)) { s->err = "Only found non-object lights while photon mapping"; return 1; }
    if (out_rgb8) buf.image(out_rgb8);
    if (out_mean) buf.mean(out_mean);
    if (counts) { counts[0] = st.n_surface; counts[1] = st.n_volume; }
    if (seconds) { seconds[0] = st.emit_s; seconds[1] = st.build_s; seconds[2] = st.gather_s; }
    return 0;
}


// ---- the photon passes one by one (rows = the 13 x 8-byte records of include/rptx.h) ----
struct OPhotons { PhotonList list; };
struct OPhotonMap { PhotonMap map; OScene* scene; };

static void to_row(const Photon& p, double* row) {
    row[0] = p.position.x; row[1] = p.position.y; row[2] = p.position.z;
    row[3] = p.direction.x; row[4] = p.direction.y; row[5] = p.direction.z;
    row[6] = p.power.x; row[7] = p.power.y; row[8] = p.power.z;
    row[9] = p.starting_position.x; row[10] = p.starting_position.y; row[11] = p.starting_position.z;
    const uint64_t k = p.key();
    std::memcpy(&row[12], &k, 8);
}
static Photon from_row(const double* row) {
    Photon p;
    p.position = Vec3(row[0], row[1], row[2]); p.direction = Vec3(row[3], row[4], row[5]);
    p.power = Vec3(row[6], row[7], row[8]); p.starting_position = Vec3(row[9], row[10], row[11]);
    uint64_t k;
    std::memcpy(&k, &row[12], 8);
    p.index = k >> 16; p.depth = (uint32_t)(k & 0xffff);
    return p;
}

void* orc_photons_emit(void* sp, const rptx_camera* c, const rptx_params* p, uint64_t photon_count, uint64_t first, uint64_t count,
                       int32_t kind, int threads) {
    OScene* s = (OScene*)sp;
    Renderer r;
    fill_renderer(r, s, c, p, threads);
    OPhotons* out = new OPhotons();
    if (!emit_photons(r, photon_count, (PhotonRenderKind)kind, &out->list, first, count)) {
        s->err = "Only found non-object lights while photon mapping";
        delete out;
        return nullptr;
    }
    return out;
}
void orc_photons_destroy(void* h) { delete (OPhotons*)h; }
void orc_photons_count(void* h, uint64_t counts[2]) {
    counts[0] = ((OPhotons*)h)->list.surface.size(); counts[1] = ((OPhotons*)h)->list.volume.size();
}
void orc_photons_rows(void* h, int32_t which, double* rows) {
    const auto& v = which == 0 ? ((OPhotons*)h)->list.surface : ((OPhotons*)h)->list.volume;
    for (size_t i = 0; i < v.size(); i++) to_row(v[i], rows + 13 * i);
}

void* orc_photon_map_build(void* sp, const double* surface_rows, size_t ns, const double* volume_rows, size_t nv, int32_t kind) {
    OPhotonMap* m = new OPhotonMap();
    m->scene = (OScene*)sp;
    PhotonList list;
    for (size_t i = 0; i < ns; i++) list.surface.push_back(from_row(surface_rows + 13 * i));
    for (size_t i = 0; i < nv; i++) list.volume.push_back(from_row(volume_rows + 13 * i));
    m->map.build((PhotonRenderKind)kind, std::move(list));
    return m;
}
void orc_photon_map_destroy(void* h) { delete (OPhotonMap*)h; }

// KdTree::nearests over a batch of queries
int orc_photon_map_knn(void* h, int32_t which, const double* q, size_t n, uint32_t k, uint64_t* keys, double* d2, uint32_t* counts) {
    const PointKd& kd = which == 0 ? ((OPhotonMap*)h)->map.surface_map : ((OPhotonMap*)h)->map.volume_map;
    std::vector<PointKd::Found> found;
    for (size_t i = 0; i < n; i++) {
        kd.nearests(Vec3(q[3 * i], q[3 * i + 1], q[3 * i + 2]), k, &found);
        for (uint32_t j = 0; j < k; j++) {
            keys[i * k + j] = j < found.size() ? found[j].item->key() : ~0ull;
            d2[i * k + j] = j < found.size() ? found[j].d2 : INF;
        }
        counts[i] = (uint32_t)found.size();
    }
    return 0;
}

size_t orc_photon_map_spheres(void* h, double* center, double* radius) {
    const auto& sp = ((OPhotonMap*)h)->map.spheres;
    if (center)
        for (size_t i = 0; i < sp.size(); i++) {
            center[3 * i] = sp[i].position.x; center[3 * i + 1] = sp[i].position.y; center[3 * i + 2] = sp[i].position.z;
            radius[i] = sp[i].radius;
        }
    return sp.size();
}

// PASS 3 over a prebuilt map
int orc_photon_map_gather(void* h, const rptx_camera* c, const rptx_params* p, int threads, uint8_t* out_rgb8, double* out_mean,
                          double* seconds) {
    OPhotonMap* m = (OPhotonMap*)h;
    Renderer r;
    fill_renderer(r, m->scene, c, p, threads);
    Buffer buf(p->width, p->height, p->filter_radius);
    const auto t0 = std::chrono::steady_clock::now();
    photon_gather(r, m->map, buf);
    if (seconds) *seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (out_rgb8) buf.image(out_rgb8);
    if (out_mean) buf.mean(out_mean);
    return 0;
}

}  // extern "C"
