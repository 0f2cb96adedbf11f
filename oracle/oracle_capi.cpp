// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).  PARITY UNPINNED.
// C entry points (ctypes) over the CPU restatement: KAT helpers, scene build / flatten /
// unflatten, batch closest hit (src/main.rs:44), the multi-threaded per-column renderer
// (src/main.rs:62-108) and scatter replay with injected random numbers.
#include <atomic>
#include <chrono>
#include <cstdio>
#include <thread>

#include "scene.hpp"

using namespace orc;

struct OrcScene {
    std::unique_ptr<Scene> sc;
    std::unique_ptr<FlatHolder> flat;  // for nested (builtin) scenes
};

struct OrcStats {
    uint64_t paths, rays, box_tests, prim_tests, shade_calls, unit_dir_violations;
    double seconds;
    uint32_t threads;
};

static thread_local std::string g_err;

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

// ---------------------------------------------------------------------------------------- KATs
void orc_xoshiro_next(const uint64_t s_in[4], int n, uint64_t* out) {
    SmallRng r;
    for (int i = 0; i < 4; i++) r.s[i] = s_in[i];
    for (int i = 0; i < n; i++) out[i] = r.next_u64();
}
void orc_seed_from_u64(uint64_t seed, uint64_t s_out[4]) {
    SmallRng r = SmallRng::seed_from_u64(seed);
    for (int i = 0; i < 4; i++) s_out[i] = r.s[i];
}
void orc_gen_f32(uint64_t seed, int n, float* out) {
    SmallRng r = SmallRng::seed_from_u64(seed);
    for (int i = 0; i < n; i++) out[i] = r.gen_f32();
}
void orc_gen_range(uint64_t seed, uint64_t range, int n, uint64_t* out) {
    SmallRng r = SmallRng::seed_from_u64(seed);
    for (int i = 0; i < n; i++) out[i] = r.gen_range_u64(range);
}
void orc_shuffle(uint64_t seed, uint32_t len, uint32_t* out) {
    SmallRng r = SmallRng::seed_from_u64(seed);
    for (uint32_t i = 0; i < len; i++) out[i] = i;
    r.shuffle(out, len);
}
void orc_get_uv(const float n[3], float out[2]) {
    Vec2 uv = Sphere::get_uv(v3(n));
    out[0] = uv.x;
    out[1] = uv.y;
}
static void put3(float* d, Vec3A v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
void orc_offset_hit_point(const float p[3], const float n[3], float out[3]) { put3(out, offset_hit_point(v3(p), v3(n))); }
void orc_reflect(const float v[3], const float n[3], float out[3]) { put3(out, reflect(v3(v), v3(n))); }
void orc_refract(const float v[3], const float n[3], float eta, float out[3]) { put3(out, refract(v3(v), v3(n), eta)); }
float orc_reflectance(float c, float r) { return reflectance(c, r); }
float orc_schlick_fresnel(float u) { return schlick_fresnel(u); }
float orc_fresnel_dielectric_2(float c, float eta) { return fresnel_dielectric_2(c, eta); }
float orc_gtr1(float ndh, float a) { return gtr1(ndh, a); }
float orc_gtr2(float ndh, float a) { return gtr2(ndh, a); }
float orc_smith_geo_ggx(float ndv, float a) { return smith_geo_ggx(ndv, a); }
void orc_sky(uint32_t kind, const float d[3], float out[3]) {
    Sky s;
    s.kind = (SkyKind)kind;
    put3(out, s(v3(d)));
}
int orc_checker_is_odd(const float p[3]) {
    CheckerTex c(Vec3A(1, 1, 1), Vec3A(0, 0, 0));
    return c.value(Vec2{}, v3(p)).x > 0.5f ? 1 : 0;
}
int orc_aabb_hit(const float mn[3], const float mx[3], const float o[3], const float d[3], float t_min, float t_max) {
    AABB b{v3(mn), v3(mx)};
    Ray r{v3(o), v3(d), Vec2{}, 0.0f};
    return b.hit(r, t_min, t_max) ? 1 : 0;
}
static void cam_to_rt(const Camera& c, RtCamera* out) {
    put3(out->origin, c.origin);
    put3(out->horizontal, c.horizontal);
    put3(out->vertical, c.vertical);
    put3(out->lower_left_corner, c.lower_left_corner);
    put3(out->u, c.u);
    put3(out->v, c.v);
    out->lens_radius = c.lens_radius;
    out->time0 = c.time0;
    out->time1 = c.time1;
}
static Camera cam_from_rt(const RtCamera* c) {
    Camera k;
    k.origin = v3(c->origin);
    k.horizontal = v3(c->horizontal);
    k.vertical = v3(c->vertical);
    k.lower_left_corner = v3(c->lower_left_corner);
    k.u = v3(c->u);
    k.v = v3(c->v);
    k.lens_radius = c->lens_radius;
    k.time0 = c->time0;
    k.time1 = c->time1;
    return k;
}
void orc_camera_new(const float from[3], const float at[3], const float vup[3], float vfov, float aspect, RtCamera* out) {
    Camera c(v3(from), v3(at), v3(vup), vfov, aspect);
    cam_to_rt(c, out);
}
// pinhole rays only (lens_radius == 0, time0 == time1): no random numbers consumed
void orc_camera_get_rays(const RtCamera* cam, const float* uv, size_t n, float* o, float* d) {
    Camera c = cam_from_rt(cam);
    c.lens_radius = 0.0f;
    c.time1 = c.time0;
    for (size_t i = 0; i < n; i++) {
        Ray r = c.get_ray(uv[2 * i], uv[2 * i + 1]);
        put3(o + 3 * i, r.o);
        put3(d + 3 * i, r.d);
    }
}
// SKY_COLOR of the scene (lib.rs:11; demo_scene.rs:22-35), incl. the env-map sky with its image
int orc_scene_sky(OrcScene* s, const float d[3], float out[3]) {
    if (!s) return -1;
    put3(out, s->sc->sky(v3(d)));
    return 0;
}
// First gradients / permutation entries of a freshly built Perlin table (texture.rs:60-91) drawn from
// SmallRng::seed_from_u64(seed): rand_vec[0..n_vec) as xyz triples, perm_x/perm_y/perm_z[0..n_perm).
void orc_perlin_table(uint64_t seed, uint32_t n_vec, float* rand_vec, uint32_t n_perm, uint32_t* px, uint32_t* py, uint32_t* pz) {
    SmallRng saved = RNG();
    RNG() = SmallRng::seed_from_u64(seed);
    Perlin p;
    RNG() = saved;
    for (uint32_t i = 0; i < n_vec && i < 256; i++) put3(rand_vec + 3 * i, p.rand_vec[i]);
    for (uint32_t i = 0; i < n_perm && i < 256; i++) {
        px[i] = (uint32_t)p.perm_x[i];
        py[i] = (uint32_t)p.perm_y[i];
        pz[i] = (uint32_t)p.perm_z[i];
    }
}
// Perlin noise / turbulence of texture `tex_id` (a PerlinTex) of a from-flat scene
int orc_texture_value(OrcScene* s, uint32_t tex_id, const float uv[2], const float p[3], float out[3]) {
    if (!s || tex_id >= s->sc->texs.size()) return -1;
    put3(out, s->sc->texs[tex_id]->value(Vec2{uv[0], uv[1]}, v3(p)));
    return 0;
}

// -------------------------------------------------------------------------------------- scenes
OrcScene* orc_scene_builtin(const char* name, float aspect, const char* asset_dir) {
    try {
        auto s = std::make_unique<OrcScene>();
        s->sc = build_named_scene(name, aspect, asset_dir ? asset_dir : ".");
        s->flat = flatten_scene(*s->sc);  // also assigns primitive ids in visitation order
        return s.release();
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
OrcScene* orc_scene_from_flat(const RtFlatScene* fs) {
    try {
        auto s = std::make_unique<OrcScene>();
        s->sc = scene_from_flat(*fs);
        return s.release();
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}
const RtFlatScene* orc_scene_flat(OrcScene* s) { return (s && s->flat) ? &s->flat->flat : nullptr; }
void orc_scene_free(OrcScene* s) { delete s; }

// world.hit(&r, t_min, t_max, &mut rec) over a batch (src/main.rs:44). skip_media != 0 makes
// ConstantMedium::hit return false (the deterministic parity batch excludes stochastic prims).
int orc_closest_hit(OrcScene* s, const float* o, const float* d, const float* times, size_t n, float t_min, float t_max,
                    int skip_media, RtHit* out) {
    if (!s) return -1;
    media_disabled() = skip_media != 0;
    for (size_t i = 0; i < n; i++) {
        Ray r{v3(o + 3 * i), v3(d + 3 * i), Vec2{}, times ? times[i] : 0.0f};
        HitRecord rec;
        RtHit h{};
        if (s->sc->world.hit(r, t_min, t_max, rec)) {
            h.prim = rec.prim;
            h.t = rec.t;
            put3(h.p, rec.p);
            put3(h.n, rec.norm);
            h.uv[0] = rec.uv.x;
            h.uv[1] = rec.uv.y;
            h.front_face = rec.front_face ? 1u : 0u;
        } else {
            h.prim = -1;
        }
        out[i] = h;
    }
    media_disabled() = false;
    return 0;
}

// world.hit with the media INCLUDED; every gen::<f32>() of ray i returns xi[i] (a constant injected sequence, so the
// result does not depend on the order in which a traversal meets several media).
int orc_closest_hit_xi(OrcScene* s, const float* o, const float* d, const float* xi, size_t n, float t_min, float t_max, RtHit* out) {
    if (!s) return -1;
    SmallRng saved = RNG();
    float seq[256];
    for (size_t i = 0; i < n; i++) {
        for (float& v : seq) v = xi[i];
        RNG().inject = seq;
        RNG().inject_n = 256;
        RNG().inject_i = 0;
        Ray r{v3(o + 3 * i), v3(d + 3 * i), Vec2{}, 0.0f};
        HitRecord rec;
        RtHit h{};
        if (s->sc->world.hit(r, t_min, t_max, rec)) {
            h.prim = rec.prim;
            h.t = rec.t;
            put3(h.p, rec.p);
            put3(h.n, rec.norm);
            h.uv[0] = rec.uv.x;
            h.uv[1] = rec.uv.y;
            h.front_face = rec.front_face ? 1u : 0u;
        } else {
            h.prim = -1;
        }
        out[i] = h;
    }
    RNG() = saved;
    return 0;
}

// ConstantMedium::hit (hitable.rs:536-579, through the Translate/RotateY wrappers around it) of world item
// `item` — which must be the medium with flat primitive index `flat_prim` of a from-flat scene — with ONE injected
// random number per case.  out per case (11 floats): hit, t, draws consumed, p[3], norm[3], front_face, 0
int orc_medium_cases(OrcScene* s, uint32_t flat_prim, const float* rays, const float* tmm, const float* xi, size_t n, float* out) {
    if (!s) return -1;
    const Hitable* item = nullptr;
    for (const auto& it : s->sc->world.items) {  // find the world item whose innermost object is that medium
        const Hitable* h = it.get();
        for (;;) {
            if (auto t = dynamic_cast<const Translate*>(h)) h = t->ptr.get();
            else if (auto r = dynamic_cast<const RotateY*>(h)) h = r->ptr.get();
            else break;
        }
        auto cm = dynamic_cast<const ConstantMedium*>(h);
        if (cm && cm->id == (int)flat_prim) item = it.get();
    }
    if (!item) return -2;
    SmallRng saved = RNG();
    for (size_t i = 0; i < n; i++) {
        float seq[4] = {xi[i], xi[i], xi[i], xi[i]};
        RNG().inject = seq;
        RNG().inject_n = 4;
        RNG().inject_i = 0;
        Ray r{v3(rays + 6 * i), v3(rays + 6 * i + 3), Vec2{}, 0.0f};
        HitRecord rec;
        float* q = out + 11 * i;
        for (int k = 0; k < 11; k++) q[k] = 0.0f;
        const bool hit = item->hit(r, tmm[2 * i], tmm[2 * i + 1], rec);
        q[0] = hit ? 1.0f : 0.0f;
        q[2] = (float)RNG().inject_i;
        if (hit) {
            q[1] = rec.t;
            put3(q + 3, rec.p);
            put3(q + 6, rec.norm);
            q[9] = rec.front_face ? 1.0f : 0.0f;
        }
    }
    RNG() = saved;
    return 0;
}

// The reference's pixel driver (src/main.rs:77-108): one job per image COLUMN, the worker's
// thread-local RNG reseeded with seed_base + i (reference: 95 + i), all spp camera rays of a pixel
// generated before any is traced, c = sum / spp.  Output: linear mean RGB (before powf(0.5)),
// pixel (i,j) at (j*nx+i)*3, j = 0 is the bottom row.  nthreads <= 0 -> all logical cores
// (threadpool::Builder::new().build() default, src/main.rs:73).
// column_begin/column_end/column_stride bound the jobs (a bounded, strided sample of the frame's
// columns for timing; stride 1 = every column).
int orc_render(OrcScene* s, uint32_t nx, uint32_t ny, uint32_t spp, int max_depth, int nthreads, uint64_t seed_base,
               uint32_t column_begin, uint32_t column_end, uint32_t column_stride, float* out_mean, OrcStats* stats) {
    if (!s || !out_mean) return -1;
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads <= 0) nthreads = 1;
    if (column_end > nx) column_end = nx;
    if (column_stride == 0) column_stride = 1;
    const Scene& sc = *s->sc;
    const uint32_t n_jobs = column_end > column_begin ? (column_end - column_begin + column_stride - 1) / column_stride : 0;
    std::atomic<uint32_t> next{0};
    std::atomic<uint64_t> a_rays{0}, a_box{0}, a_prim{0}, a_shade{0}, a_viol{0};
    auto t0 = std::chrono::steady_clock::now();
    auto worker = [&]() {
        Counters& c = CNT();
        c = Counters();
        while (true) {
            uint32_t job = next.fetch_add(1);
            if (job >= n_jobs) break;
            const uint32_t i = column_begin + job * column_stride;
            RNG() = SmallRng::seed_from_u64(seed_base + (uint64_t)i);
            std::vector<Ray> rays;
            rays.reserve(spp);
            for (uint32_t j = 0; j < ny; j++) {
                Vec3A col;
                rays.clear();
                for (uint32_t k = 0; k < spp; k++) {
                    float u = ((float)i + RNG().gen_f32()) / (float)nx;
                    float v = ((float)j + RNG().gen_f32()) / (float)ny;
                    rays.push_back(sc.cam.get_ray(u, v));
                }
                for (const Ray& r : rays) col += ray_color(r, sc, 0, max_depth);
                col /= (float)spp;
                float* px = out_mean + ((size_t)j * nx + i) * 3;
                px[0] = col.x;
                px[1] = col.y;
                px[2] = col.z;
            }
        }
        a_rays += c.rays;
        a_box += c.box_tests;
        a_prim += c.prim_tests;
        a_shade += c.shade_calls;
        a_viol += c.unit_dir_violations;
    };
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; t++) th.emplace_back(worker);
    for (auto& t : th) t.join();
    auto t1 = std::chrono::steady_clock::now();
    if (stats) {
        stats->paths = (uint64_t)n_jobs * ny * spp;
        stats->rays = a_rays;
        stats->box_tests = a_box;
        stats->prim_tests = a_prim;
        stats->shade_calls = a_shade;
        stats->unit_dir_violations = a_viol;
        stats->seconds = std::chrono::duration<double>(t1 - t0).count();
        stats->threads = (uint32_t)nthreads;
    }
    return 0;
}

// Reference output stage (src/main.rs:98-105,127): powf(0.5), *255.99 as u8, vertical flip.
void orc_resolve_rgb8(const float* mean, uint32_t nx, uint32_t ny, uint8_t* out) {
    for (uint32_t j = 0; j < ny; j++)
        for (uint32_t i = 0; i < nx; i++) {
            const float* px = mean + ((size_t)j * nx + i) * 3;
            Vec3A c = Vec3A(px[0], px[1], px[2]).powf(1.0f / 2.0f);
            uint8_t* o = out + ((size_t)(ny - 1 - j) * nx + i) * 3;
            o[0] = rust_f32_as_u8(c.x * 255.99f);
            o[1] = rust_f32_as_u8(c.y * 255.99f);
            o[2] = rust_f32_as_u8(c.z * 255.99f);
        }
}

// Replay Material::emitted + Material::scatter for material `mat_id` of a from-flat scene with the
// random numbers `xi` injected in place of gen::<f32>() draws.
// rec_in: p[3] norm[3] tang[3] uv[2] t front_face  (12 floats + front_face as float 0/1)
// out: emitted[3] attenuation[3] scat_o[3] scat_d[3]; returns scatter()'s bool (0/1), <0 on error;
// *n_consumed = number of injected values consumed.
int orc_scatter(OrcScene* s, uint32_t mat_id, const float ray_o[3], const float ray_d[3], const float* rec_in,
                const float* xi, size_t n_xi, float* out, uint32_t* n_consumed) {
    if (!s || mat_id >= s->sc->mats.size()) return -1;
    HitRecord rec;
    rec.p = v3(rec_in);
    rec.norm = v3(rec_in + 3);
    rec.tang = v3(rec_in + 6);
    rec.uv = Vec2{rec_in[9], rec_in[10]};
    rec.t = rec_in[11];
    rec.front_face = rec_in[12] != 0.0f;
    Ray r{v3(ray_o), v3(ray_d), Vec2{}, 0.0f};
    SmallRng saved = RNG();
    RNG().inject = xi;
    RNG().inject_n = n_xi;
    RNG().inject_i = 0;
    const Material& m = *s->sc->mats[mat_id];
    Ray scattered{Vec3A(), Vec3A(), r.s, 0.0f};
    Vec3A att(1, 1, 1);
    Vec3A em = m.emitted(rec.uv, rec.p);
    bool ok = m.scatter(r, rec, att, scattered);
    if (n_consumed) *n_consumed = (uint32_t)RNG().inject_i;
    RNG() = saved;
    put3(out, em);
    put3(out + 3, att);
    put3(out + 6, scattered.o);
    put3(out + 9, scattered.d);
    return ok ? 1 : 0;
}

}  // extern "C"
