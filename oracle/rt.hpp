// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).  PARITY UNPINNED (no reference tests
// exist); pinned only by KATs K1-K12 of SURVEY.md section 4.
//
// Structurally faithful C++17 restatement of the reference's hot path:
//   src/math.rs (all), src/hitable.rs (all), src/material.rs (all), src/pbr.rs (all),
//   src/texture.rs (all), src/camera.rs (all), src/lib.rs (RNG / sky contract),
//   src/main.rs:36-108 (ray_color + sample loop + per-column seeding).
// Same names, same operation order, same quirks: virtual dispatch, a
// shared_ptr material copy per accepted candidate hit, recursive ray_color, 1/d recomputed per
// BVH node, object-median BVH on a random axis.
//
// Oracle-only additions (clearly marked "ORACLE ADDITION"): a primitive id carried in HitRecord,
// work counters, a `flatten` visitor that emits an RtFlatScene, MovingSphere / thin-lens / shutter
// extensions (not in the reference snapshot; semantics from Shirley's RTIOW section 12 / RTNW section 2).
#pragma once
#include <algorithm>
#include <cassert>
#include <cfloat>
#include <memory>
#include <string>
#include <vector>

#include "vec.hpp"

namespace orc {

constexpr float PI = 3.14159265358979323846264338327950288f;
constexpr float FRAC_1_PI = 0.318309886183790671537767526745028724f;

// ORACLE ADDITION: per-thread work counters
struct Counters {
    uint64_t rays = 0, box_tests = 0, prim_tests = 0, shade_calls = 0;
    uint64_t unit_dir_violations = 0;  // times the reference's release-mode assert!(main.rs:39) would have fired
};
inline Counters& CNT() {
    static thread_local Counters c;
    return c;
}

// ------------------------------------------------------------------------------------------ math.rs
inline bool vec3a_near_zero(Vec3A v) {  // math.rs:8-11
    const float s = FLT_EPSILON;
    return (std::fabs(v.x) < s) && (std::fabs(v.y) < s) && (std::fabs(v.z) < s);
}
inline bool vec3a_near_one(Vec3A v) { return std::fabs(v.length() - 1.0f) < 1e-6f; }  // math.rs:13-15
inline Vec3A vec3a_random() {  // math.rs:17-25
    float x = RNG().gen_f32();
    float y = RNG().gen_f32();
    float z = RNG().gen_f32();
    return Vec3A(x, y, z);
}
inline Vec3A vec3a_random_range(float min, float max) { return vec3a_random() * (max - min) + min; }  // 27-29
inline Vec3A random_in_unit_sphere() {  // math.rs:30-37
    while (true) {
        Vec3A v = vec3a_random_range(-1.0f, 1.0f);
        if (v.length_squared() < 1.0f) return v;
    }
}
inline Vec3A random_on_unit_sphere() { return random_in_unit_sphere().normalize(); }  // 38-40
inline Vec3A random_in_hemisphere(Vec3A norm) {                                       // 43-50
    Vec3A v = random_in_unit_sphere();
    return v.dot(norm) > 0.0f ? v : -v;
}
inline Vec3A random_on_hemisphere(Vec3A norm) { return random_in_hemisphere(norm).normalize(); }  // 51-53

struct Ray {  // math.rs:55-66
    Vec3A o, d;
    Vec2 s;
    float tm = 0.0f;  // ORACLE ADDITION (extension): ray time for MovingSphere
    Vec3A at(float t) const { return o + d * t; }
};

inline Vec3A reflect(Vec3A v, Vec3A n) { return v - 2.0f * v.dot(n) * n; }  // math.rs:68-70
inline Vec3A refract(Vec3A uv, Vec3A n, float etai_over_etat) {           // math.rs:72-77
    float cos_theta = -std::fmin(uv.dot(n), 1.0f);
    Vec3A r_out_perp = etai_over_etat * (uv + cos_theta * n);
    Vec3A r_out_parallel = -std::sqrt(std::fabs(1.0f - r_out_perp.length_squared())) * n;
    return r_out_perp + r_out_parallel;
}
inline float schlick_fresnel(float u) { return powi(1.0f - u, 5); }  // math.rs:79-81
inline float reflectance(float cosine, float ref_idx) {             // math.rs:84-88
    float r0 = (1.0f - ref_idx) / (1.0f + ref_idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * schlick_fresnel(cosine);
}

struct AABB {  // math.rs:90-131
    Vec3A min, max;
    bool hit(const Ray& r, float t_min, float t_max) const {
        CNT().box_tests++;
        for (int i = 0; i < 3; i++) {
            float inv_d = 1.0f / r.d[i];
            float t0 = (min[i] - r.o[i]) * inv_d;
            float t1 = (max[i] - r.o[i]) * inv_d;
            if (inv_d < 0.0f) std::swap(t0, t1);
            t_min = std::fmax(t_min, t0);  // Rust f32::max == fmaxf (NaN-ignoring)
            t_max = std::fmin(t_max, t1);
            if (t_max <= t_min) return false;
        }
        return true;
    }
    AABB surround(const AABB& rhs) const {
        AABB o;
        o.min = Vec3A(std::fmin(min.x, rhs.min.x), std::fmin(min.y, rhs.min.y), std::fmin(min.z, rhs.min.z));
        o.max = Vec3A(std::fmax(max.x, rhs.max.x), std::fmax(max.y, rhs.max.y), std::fmax(max.z, rhs.max.z));
        return o;
    }
};

inline Vec3A smooth(Vec3A v) { return v * v * (3.0f - 2.0f * v); }  // math.rs:133-135

inline Vec3A offset_hit_point(Vec3A p, Vec3A n) {  // math.rs:137-152 (RT Gems ch. 6)
    const float ORIGIN = 1.0f / 32.0f, INT_SCALE = 256.0f, FLOAT_SCALE = 1.0f / 65536.0f;
    int32_t of_i_x = rust_f32_as_i32(n.x * INT_SCALE);
    int32_t of_i_y = rust_f32_as_i32(n.y * INT_SCALE);
    int32_t of_i_z = rust_f32_as_i32(n.z * INT_SCALE);
    // release-mode Rust integer add wraps
    float p_i_x = u2f(f2u(p.x) + (uint32_t)(p.x < 0.0f ? -of_i_x : of_i_x));
    float p_i_y = u2f(f2u(p.y) + (uint32_t)(p.y < 0.0f ? -of_i_y : of_i_y));
    float p_i_z = u2f(f2u(p.z) + (uint32_t)(p.z < 0.0f ? -of_i_z : of_i_z));
    float x = std::fabs(p.x) < ORIGIN ? p.x + n.x * FLOAT_SCALE : p_i_x;
    float y = std::fabs(p.y) < ORIGIN ? p.y + n.y * FLOAT_SCALE : p_i_y;
    float z = std::fabs(p.z) < ORIGIN ? p.z + n.z * FLOAT_SCALE : p_i_z;
    return Vec3A(x, y, z);
}
inline float lerp(float from, float to, float s) { return from + (to - from) * s; }  // math.rs:154-156

// --------------------------------------------------------------------------------------- texture.rs
struct Texture {  // texture.rs:11-13
    virtual ~Texture() = default;
    virtual Vec3A value(Vec2 uv, Vec3A p) const = 0;
    int flat_id = -1;  // ORACLE ADDITION
};
using TexPtr = std::shared_ptr<Texture>;

struct ConstantTex : Texture {  // texture.rs:15-23
    Vec3A col;
    explicit ConstantTex(Vec3A c) : col(c) {}
    Vec3A value(Vec2, Vec3A) const override { return col; }
};

struct CheckerTex : Texture {  // texture.rs:25-49
    TexPtr odd, even;
    CheckerTex(Vec3A odd_col, Vec3A even_col)
        : odd(std::make_shared<ConstantTex>(odd_col)), even(std::make_shared<ConstantTex>(even_col)) {}
    CheckerTex(TexPtr o, TexPtr e) : odd(std::move(o)), even(std::move(e)) {}
    Vec3A value(Vec2 uv, Vec3A p) const override {
        float sines = std::sin(p.x * 10.0f) * std::sin(p.y * 10.0f) * std::sin(p.z * 10.0f);
        return sines < 0.0f ? odd->value(uv, p) : even->value(uv, p);
    }
};

constexpr int PERLIN_POINT_COUNT = 256;
struct Perlin {  // texture.rs:51-147
    std::vector<Vec3A> rand_vec;
    std::vector<size_t> perm_x, perm_y, perm_z;
    struct NoInit {};
    explicit Perlin(NoInit) {}  // ORACLE ADDITION: tables filled from an RtPerlin, no RNG draws
    Perlin() {  // Default::default, texture.rs:60-91: draws from the constructing thread's RNG
        rand_vec.assign(PERLIN_POINT_COUNT, Vec3A());
        for (auto& v : rand_vec) v = vec3a_random_range(-1.0f, 1.0f);
        std::vector<size_t> p(PERLIN_POINT_COUNT);
        for (size_t i = 0; i < p.size(); i++) p[i] = i;
        RNG().shuffle(p.data(), p.size());
        perm_x = p;
        RNG().shuffle(p.data(), p.size());
        perm_y = p;
        RNG().shuffle(p.data(), p.size());
        perm_z = p;
    }
    static float trilinear_interp(const Vec3A c[2][2][2], Vec3A uvw) {  // texture.rs:93-112
        float accum = 0.0f;
        Vec3A uvw2 = smooth(uvw);
        float u = uvw2.x, v = uvw2.y, w = uvw2.z;
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++)
                for (int k = 0; k < 2; k++) {
                    Vec3A weight = uvw - Vec3A((float)i, (float)j, (float)k);
                    accum += c[i][j][k].dot(weight) * (i == 1 ? u : 1.0f - u) * (j == 1 ? v : 1.0f - v) *
                             (k == 1 ? w : 1.0f - w);
                }
        return accum;
    }
    float noise(Vec3A p) const {  // texture.rs:125-146
        int64_t i = rust_f32_as_isize(std::floor(p.x));
        int64_t j = rust_f32_as_isize(std::floor(p.y));
        int64_t k = rust_f32_as_isize(std::floor(p.z));
        Vec3A c[2][2][2];
        for (int di = 0; di < 2; di++)
            for (int dj = 0; dj < 2; dj++)
                for (int dk = 0; dk < 2; dk++) {
                    // wrapping add then rem_euclid(256) == & 255 on two's complement
                    size_t index = perm_x[(size_t)(((uint64_t)i + (uint64_t)di) & 255)] ^
                                   perm_y[(size_t)(((uint64_t)j + (uint64_t)dj) & 255)] ^
                                   perm_z[(size_t)(((uint64_t)k + (uint64_t)dk) & 255)];
                    c[di][dj][dk] = rand_vec[index];
                }
        Vec3A uvw = p - p.floor();
        return trilinear_interp(c, uvw);
    }
    float turb(Vec3A p) const {  // texture.rs:115-124
        float accum = 0.0f, w = 1.0f;
        for (int i = 0; i < 7; i++) {
            accum += w * noise(p);
            p *= 2.0f;
            w *= 0.5f;
        }
        return std::fabs(accum);
    }
};

struct PerlinTex : Texture {  // texture.rs:149-168
    Perlin perlin;
    float scale;
    explicit PerlinTex(float s) : perlin(), scale(s) {}
    PerlinTex(float s, Perlin::NoInit t) : perlin(t), scale(s) {}
    Vec3A value(Vec2, Vec3A p) const override {
        return (std::sin(10.0f * perlin.turb(p) + scale * p.z) + 1.0f) * 0.5f * Vec3A(1, 1, 1);
    }
};

struct ImageTex : Texture {  // texture.rs:169-193; img = decoded RGB8 / 255 (to_rgb32f)
    uint32_t w = 0, h = 0;
    std::vector<float> img;                 // w*h*3
    std::shared_ptr<std::vector<uint8_t>> rgb8;  // ORACLE ADDITION: the source bytes (for flatten)
    ImageTex(uint32_t w_, uint32_t h_, const uint8_t* bytes) : w(w_), h(h_) {
        rgb8 = std::make_shared<std::vector<uint8_t>>(bytes, bytes + (size_t)w * h * 3);
        img.resize((size_t)w * h * 3);
        for (size_t i = 0; i < img.size(); i++) img[i] = (float)bytes[i] / 255.0f;
    }
    Vec3A value(Vec2 uv, Vec3A) const override {
        float u = rust_clamp(uv.x, 0.0f, 1.0f);
        float v = 1.0f - rust_clamp(uv.y, 0.0f, 1.0f);
        uint32_t i = std::min(rust_f32_as_u32(u * (float)w), w - 1);
        uint32_t j = std::min(rust_f32_as_u32(v * (float)h), h - 1);
        const float* px = &img[((size_t)j * w + i) * 3];
        return Vec3A(px[0], px[1], px[2]);
    }
};

// ------------------------------------------------------------------------------------- hitable.rs (1)
struct Material;
using MatPtr = std::shared_ptr<Material>;

struct HitRecord {  // hitable.rs:13-42
    Vec3A p, norm, tang;
    float t = 0.0f;
    bool front_face = false;
    MatPtr mat;
    Vec2 uv;
    int prim = -1;  // ORACLE ADDITION: flat id of the primitive that wrote the record
    void set_face_normal(const Ray& r, Vec3A outward_normal) {
        front_face = r.d.dot(outward_normal) < 0.0f;
        norm = front_face ? outward_normal : -outward_normal;
    }
    Vec3A world_to_local(Vec3A v) const {
        Vec3A bitang = norm.cross(tang);
        return Vec3A(v.dot(tang), v.dot(bitang), v.dot(norm));
    }
    Vec3A world_to_local_with_rot(Vec3A v, float rot) const {
        Vec3A tg = std::cos(rot) * tang - std::sin(rot) * norm.cross(tang);
        Vec3A bitang = norm.cross(tg);
        return Vec3A(v.dot(tg), v.dot(bitang), v.dot(norm));
    }
};

// -------------------------------------------------------------------------------------- material.rs
struct Material {  // material.rs:10-15
    virtual ~Material() = default;
    virtual bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const = 0;
    virtual Vec3A emitted(Vec2, Vec3A) const { return Vec3A(); }
    int flat_id = -1;  // ORACLE ADDITION
};

struct Emission : Material {  // material.rs:17-29
    TexPtr emit;
    explicit Emission(TexPtr e) : emit(std::move(e)) {}
    bool scatter(const Ray&, const HitRecord&, Vec3A&, Ray&) const override { return false; }
    Vec3A emitted(Vec2 uv, Vec3A p) const override { return emit->value(uv, p); }
};

struct Diffuse : Material {  // material.rs:31-46
    TexPtr albedo;
    explicit Diffuse(TexPtr a) : albedo(std::move(a)) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A scatter_direction = rec.norm + random_in_unit_sphere().normalize();
        if (vec3a_near_zero(scatter_direction)) scatter_direction = rec.norm;
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        scattered = Ray{p, scatter_direction.normalize(), r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p);
        return true;
    }
};

struct Lambert : Material {  // material.rs:48-59
    TexPtr albedo;
    explicit Lambert(TexPtr a) : albedo(std::move(a)) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        scattered = Ray{p, random_on_hemisphere(rec.norm), r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p) * 2.0f * rec.norm.dot(scattered.d);
        return true;
    }
};

struct Metal : Material {  // material.rs:61-73
    Vec3A albedo;
    float fuzz;
    Metal(Vec3A a, float f) : albedo(a), fuzz(f) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A reflected = reflect(r_in.d, rec.norm) + fuzz * random_in_unit_sphere();
        scattered = Ray{rec.p, reflected.normalize(), r_in.s, r_in.tm};
        attenuation = albedo;
        return reflected.dot(rec.norm) > 0.0f;
    }
};

struct Dielectric : Material {  // material.rs:75-97
    float ior;
    explicit Dielectric(float i) : ior(i) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        attenuation = Vec3A(1, 1, 1);
        float ref_idx = rec.front_face ? 1.0f / ior : ior;
        float cos_theta = -std::fmin(r_in.d.dot(rec.norm), 1.0f);  // precedence quirk (Appendix A8)
        float sin_thera = std::sqrt(1.0f - cos_theta * cos_theta);
        bool cannot_refract = sin_thera * ref_idx > 1.0f;
        float rnd_num = RNG().gen_f32();
        Vec3A dir = (cannot_refract || reflectance(cos_theta, ref_idx) > rnd_num) ? reflect(r_in.d, rec.norm)
                                                                                 : refract(r_in.d, rec.norm, ref_idx);
        dir = dir.normalize();
        scattered = Ray{rec.p, dir, r_in.s, r_in.tm};
        return true;
    }
};

struct Isotropic : Material {  // material.rs:99-113
    TexPtr albedo;
    explicit Isotropic(TexPtr a) : albedo(std::move(a)) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        scattered = Ray{rec.p, random_on_unit_sphere(), r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p);
        return true;
    }
};

// ------------------------------------------------------------------------------------------- pbr.rs
struct OrenNayar : Material {  // pbr.rs:12-42
    TexPtr albedo;
    float roughness;
    OrenNayar(TexPtr a, float r) : albedo(std::move(a)), roughness(r) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float cos_i = std::fabs(rec.norm.dot(r_in.d));
        float cos_o = rec.norm.dot(dir_o);
        float sin_i = std::sqrt(1.0f - cos_i * cos_i);
        float sin_o = std::sqrt(1.0f - cos_o * cos_o);
        float max_cos = std::fmax(cos_i * cos_o + sin_i * sin_o, 0.0f);
        float r2 = roughness * roughness;
        float a = 1.0f - 0.5f * r2 / (r2 + 0.33f);
        float b = 0.45f * r2 / (r2 + 0.09f);
        float sin_alpha, tan_beta;
        if (cos_i > cos_o) {
            sin_alpha = sin_o;
            tan_beta = sin_i / cos_i;
        } else {
            sin_alpha = sin_i;
            tan_beta = sin_o / cos_o;
        }
        float w = a + b * max_cos * sin_alpha * tan_beta;
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p) * w * 2.0f * cos_o;
        return true;
    }
};

struct BurleyDiffuse : Material {  // pbr.rs:45-69
    TexPtr albedo;
    float roughness;
    BurleyDiffuse(TexPtr a, float r) : albedo(std::move(a)), roughness(r) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_i = rec.norm.dot(-r_in.d);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_o = h.dot(dir_o);
        float fl = schlick_fresnel(n_dot_o);
        float fv = schlick_fresnel(n_dot_i);
        float fd90 = 0.5f + 2.0f * h_dot_o * h_dot_o * roughness;
        float fd = lerp(1.0f, fd90, fl) * lerp(1.0f, fd90, fv);
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p) * fd * 2.0f * n_dot_o;
        return true;
    }
};

inline float gtr1(float n_dot_h, float a) {  // pbr.rs:72-79
    if (a >= 1.0f) return FRAC_1_PI;
    float a2 = a * a;
    float t = 1.0f + (a2 - 1.0f) * n_dot_h * n_dot_h;
    return (a2 - 1.0f) / (PI * std::log(a2) * t);
}
inline float gtr2(float n_dot_h, float a) {  // pbr.rs:81-85
    float a2 = a * a;
    float t = 1.0f + (a2 - 1.0f) * n_dot_h * n_dot_h;
    return a2 / (PI * t * t);
}
inline float gtr2_ansio(Vec3A h_local, float ax, float ay) {  // pbr.rs:87-89
    return 1.0f / (PI * ax * ay * powi(powi(h_local.x / ax, 2) + powi(h_local.y / ay, 2) + h_local.z * h_local.z, 2));
}
inline float smith_geo_ggx(float n_dot_v, float alpha) {  // pbr.rs:92-96
    float a = alpha * alpha;
    float b = n_dot_v * n_dot_v;
    return 1.0f / (n_dot_v + std::sqrt(a + b - a * b));
}
inline float smith_geo_ggx_aniso(Vec3A v_local, float ax, float ay) {  // pbr.rs:98-100
    return 1.0f / (v_local.z + std::sqrt(powi(v_local.x * ax, 2) + powi(v_local.y * ay, 2) + v_local.z * v_local.z));
}
inline float fresnel_dielectric(float n_dot_i, float n_dot_t, float eta) {  // pbr.rs:107-113
    float rs = (n_dot_i - eta * n_dot_t) / (n_dot_i + eta * n_dot_t);
    float rp = (eta * n_dot_i - n_dot_t) / (eta * n_dot_i + n_dot_t);
    return (rs * rs + rp * rp) / 2.0f;
}
inline float fresnel_dielectric_2(float n_dot_i, float eta) {  // pbr.rs:120-129
    float n_dot_t_sq = 1.0f - (1.0f - n_dot_i * n_dot_i) / (eta * eta);
    if (n_dot_t_sq < 0.0f) return 1.0f;
    float n_dot_t = std::sqrt(n_dot_t_sq);
    return fresnel_dielectric(std::fabs(n_dot_i), n_dot_t, eta);
}
inline float smith_masking_gtr2_2(Vec3A v_world, Vec3A n, float roughness) {  // pbr.rs:145-152
    float alpha = roughness * roughness;
    float a2 = alpha * alpha;
    float v2_z_ = v_world.dot(n);
    float v2_z = v2_z_ * v2_z_;
    float lambda = (-1.0f + std::sqrt(1.0f + a2 * (1.0f - v2_z) / v2_z)) / 2.0f;
    return 1.0f / (1.0f + lambda);
}

struct RoughPlastic : Material {  // pbr.rs:153-189
    TexPtr spec_color, diff_color;
    float roughness, eta;
    RoughPlastic(TexPtr s, TexPtr d, float r, float e)
        : spec_color(std::move(s)), diff_color(std::move(d)), roughness(r), eta(e) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_i = rec.norm.dot(-r_in.d);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_i = h.dot(-r_in.d);
        float h_dot_o = h.dot(dir_o);
        float n_dot_h = rec.norm.dot(h);
        Vec3A kd = diff_color->value(rec.uv, rec.p);
        Vec3A ks = spec_color->value(rec.uv, rec.p);
        float rough = rust_clamp(roughness, 0.01f, 1.0f);
        float f_o = fresnel_dielectric_2(h_dot_o, eta);
        float d = gtr2(n_dot_h, rough);
        float g = smith_masking_gtr2_2(-r_in.d, rec.norm, rough) * smith_masking_gtr2_2(dir_o, rec.norm, rough);
        Vec3A spec_contrib = ks * (g * f_o * d) / (4.0f * n_dot_i * n_dot_o);
        float f_i = fresnel_dielectric_2(h_dot_i, eta);
        Vec3A diff_contrib = kd * (1.0f - f_o) * (1.0f - f_i) * FRAC_1_PI;
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = (spec_contrib + diff_contrib) * n_dot_o * 2.0f * PI;
        return true;
    }
};

struct DisneyDiffuse : Material {  // pbr.rs:192-223
    TexPtr albedo;
    float roughness, subsurface;
    DisneyDiffuse(TexPtr a, float r, float s) : albedo(std::move(a)), roughness(r), subsurface(s) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_i = rec.norm.dot(-r_in.d);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_o = h.dot(dir_o);
        float fo = schlick_fresnel(n_dot_o);
        float fi = schlick_fresnel(n_dot_i);
        float fd90 = 0.5f + 2.0f * h_dot_o * h_dot_o * roughness;
        float fd = lerp(1.0f, fd90, fo) * lerp(1.0f, fd90, fi);
        float fss90 = roughness * h_dot_o * h_dot_o;
        float fss_wi = lerp(1.0f, fss90, fi);
        float fss_wo = lerp(1.0f, fss90, fo);
        float fss = 1.25f * (fss_wi * fss_wo * (1.0f / (n_dot_i + n_dot_o) - 0.5f) + 0.5f);
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = albedo->value(rec.uv, rec.p) * lerp(fd, fss, subsurface) * 2.0f * n_dot_o;
        return true;
    }
};

struct DisneyMetal : Material {  // pbr.rs:224-278
    TexPtr albedo;
    float roughness, anisotropic, rot;
    DisneyMetal(TexPtr a, float r, float an, float ro)
        : albedo(std::move(a)), roughness(r), anisotropic(an), rot(ro) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_i = rec.norm.dot(-r_in.d);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_o = h.dot(dir_o);
        float n_dot_h = rec.norm.dot(h);
        Vec3A alb = albedo->value(rec.uv, rec.p);
        Vec3A fm = alb.lerp(Vec3A(1, 1, 1), schlick_fresnel(h_dot_o));
        const float alpha_min = 0.0001f;
        float dm, gm;
        if (anisotropic > -10.0f) {
            float aspect = std::sqrt(1.0f - 0.9f * anisotropic);
            float ax = std::fmax(roughness * roughness / aspect, alpha_min);
            float ay = std::fmax(roughness * roughness * aspect, alpha_min);
            float rt = rot * 2.0f * PI;
            Vec3A h_local = rec.world_to_local_with_rot(h, rt);
            dm = gtr2_ansio(h_local, ax, ay);
            Vec3A i_local = rec.world_to_local_with_rot(-r_in.d, rt);
            Vec3A o_local = rec.world_to_local_with_rot(dir_o, rt);
            gm = smith_geo_ggx_aniso(i_local, ax, ay) * smith_geo_ggx_aniso(o_local, ax, ay);
        } else {
            float r2 = std::fmax(roughness * roughness, alpha_min);
            dm = gtr2(n_dot_h, r2);
            gm = smith_geo_ggx(n_dot_i, r2) * smith_geo_ggx(n_dot_o, r2);
        }
        Vec3A metal_w = fm * dm * gm;
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = metal_w * n_dot_o * 2.0f * PI;
        return true;
    }
};

struct DisneySheen : Material {  // pbr.rs:281-308
    TexPtr albedo;
    float tint;
    DisneySheen(TexPtr a, float t) : albedo(std::move(a)), tint(t) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_o = h.dot(dir_o);
        Vec3A alb = albedo->value(rec.uv, rec.p);
        float luminance = Vec3A(0.3f, 0.6f, 0.1f).dot(alb);
        Vec3A c_tint = luminance > 0.0f ? alb / luminance : Vec3A(1, 1, 1);
        Vec3A c_sheen = Vec3A(1, 1, 1).lerp(c_tint, tint);
        Vec3A f_sheen = c_sheen * schlick_fresnel(h_dot_o);
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = f_sheen * n_dot_o * 2.0f * PI;
        return true;
    }
};

struct DisneyClearcoat : Material {  // pbr.rs:310-335
    float clearcoat_gloss;
    explicit DisneyClearcoat(float g) : clearcoat_gloss(g) {}
    bool scatter(const Ray& r_in, const HitRecord& rec, Vec3A& attenuation, Ray& scattered) const override {
        Vec3A p = offset_hit_point(rec.p, rec.norm);
        Vec3A dir_o = random_on_hemisphere(rec.norm);
        float n_dot_i = rec.norm.dot(-r_in.d);
        float n_dot_o = rec.norm.dot(dir_o);
        Vec3A h = (dir_o - r_in.d).normalize();
        float h_dot_o = h.dot(dir_o);
        float n_dot_h = rec.norm.dot(h);
        float fc = lerp(0.4f, 1.0f, schlick_fresnel(h_dot_o));
        float dc = gtr1(n_dot_h, lerp(0.1f, 0.001f, clearcoat_gloss));
        float gc = smith_geo_ggx(n_dot_i, 0.25f) * smith_geo_ggx(n_dot_o, 0.25f);
        float cc = 0.25f * fc * dc * gc;
        scattered = Ray{p, dir_o, r_in.s, r_in.tm};
        attenuation = Vec3A::splat(cc) * n_dot_o * 2.0f * PI;
        return true;
    }
};

// ------------------------------------------------------------------------------------- hitable.rs (2)
struct Flattener;  // ORACLE ADDITION, flatten.hpp

struct Hitable {  // hitable.rs:50-54
    virtual ~Hitable() = default;
    virtual bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const = 0;
    virtual bool bbox(AABB& aabb) const = 0;
    virtual std::string memo() const = 0;
    // ORACLE ADDITION: emit primitives in the reference's visitation order
    virtual void flatten(Flattener& f, uint32_t xform_chain, bool boundary_only) const = 0;
};
using HitPtr = std::shared_ptr<Hitable>;

struct Sphere : Hitable {  // hitable.rs:56-112
    Vec3A c;
    float r;
    MatPtr mat;
    std::string name;
    mutable int id = -1;  // ORACLE ADDITION
    Sphere(Vec3A c_, float r_, MatPtr m, std::string n = "") : c(c_), r(r_), mat(std::move(m)), name(std::move(n)) {}
    static Vec2 get_uv(Vec3A n) {  // hitable.rs:65-71
        float theta = std::acos(-n.y);
        float phi = std::atan2(-n.z, n.x) + PI;
        float u = phi / (2.0f * PI);
        float v = theta / PI;
        return Vec2{u, v};
    }
    bool hit(const Ray& ray, float t_min, float t_max, HitRecord& rec) const override {
        CNT().prim_tests++;
        Vec3A oc = ray.o - c;
        float a = ray.d.length_squared();
        float half_b = oc.dot(ray.d);
        float cc = oc.length_squared() - r * r;
        float discriminant = half_b * half_b - a * cc;
        if (discriminant < 0.0f) return false;
        float sqrtd = std::sqrt(discriminant);
        float root = (-half_b - sqrtd) / a;
        if (root < t_min || t_max < root) {
            root = (-half_b + sqrtd) / a;
            if (root < t_min || t_max < root) return false;
        }
        rec.t = root;
        rec.p = ray.at(rec.t);
        Vec3A outward_normal = (rec.p - c) / r;
        rec.tang = Vec3A(0, 1, 0).cross(outward_normal).normalize();
        rec.set_face_normal(ray, outward_normal);
        rec.uv = get_uv(outward_normal);
        rec.mat = mat;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override {
        aabb.min = c - r;
        aabb.max = c + r;
        return true;
    }
    std::string memo() const override { return name; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

// EXTENSION (not in the reference snapshot): RTNW section 2 moving sphere.
struct MovingSphere : Hitable {
    Vec3A c0, c1;
    float time0, time1, r;
    MatPtr mat;
    mutable int id = -1;
    MovingSphere(Vec3A c0_, Vec3A c1_, float t0, float t1, float r_, MatPtr m)
        : c0(c0_), c1(c1_), time0(t0), time1(t1), r(r_), mat(std::move(m)) {}
    Vec3A center(float time) const { return c0 + ((time - time0) / (time1 - time0)) * (c1 - c0); }
    bool hit(const Ray& ray, float t_min, float t_max, HitRecord& rec) const override {
        CNT().prim_tests++;
        Vec3A c = center(ray.tm);
        Vec3A oc = ray.o - c;
        float a = ray.d.length_squared();
        float half_b = oc.dot(ray.d);
        float cc = oc.length_squared() - r * r;
        float discriminant = half_b * half_b - a * cc;
        if (discriminant < 0.0f) return false;
        float sqrtd = std::sqrt(discriminant);
        float root = (-half_b - sqrtd) / a;
        if (root < t_min || t_max < root) {
            root = (-half_b + sqrtd) / a;
            if (root < t_min || t_max < root) return false;
        }
        rec.t = root;
        rec.p = ray.at(rec.t);
        Vec3A outward_normal = (rec.p - c) / r;
        rec.tang = Vec3A(0, 1, 0).cross(outward_normal).normalize();
        rec.set_face_normal(ray, outward_normal);
        rec.uv = Sphere::get_uv(outward_normal);
        rec.mat = mat;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override {
        AABB b0{c0 - r, c0 + r}, b1{c1 - r, c1 + r};
        aabb = b0.surround(b1);
        return true;
    }
    std::string memo() const override { return "MovingSphere"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct HitableList : Hitable {  // hitable.rs:114-156 (type alias Vec<Arc<dyn Hitable>> + impl)
    std::vector<HitPtr> items;
    HitableList() = default;
    explicit HitableList(std::vector<HitPtr> v) : items(std::move(v)) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        HitRecord temp_rec;
        float closest_so_far = t_max;
        bool hit_anything = false;
        for (const auto& item : items) {
            bool be_hit = item->hit(r, t_min, closest_so_far, temp_rec);
            if (be_hit) {
                hit_anything = true;
                closest_so_far = temp_rec.t;
            }
        }
        if (hit_anything) rec = temp_rec;
        return hit_anything;
    }
    bool bbox(AABB& aabb) const override {
        if (items.empty()) return false;
        AABB b;
        bool first = true;
        for (const auto& o : items) {
            if (o->bbox(b)) {
                if (first) {
                    aabb = b;
                    first = false;
                } else {
                    aabb = aabb.surround(b);
                }
            } else {
                return false;
            }
        }
        return true;
    }
    std::string memo() const override { return "HitableList"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

// f32::total_cmp key
static inline int32_t total_cmp_key(float f) {
    int32_t i = f2i(f);
    i ^= (int32_t)(((uint32_t)(i >> 31)) >> 1);
    return i;
}

struct BvhNode : Hitable {  // hitable.rs:158-241
    AABB aabb;
    HitPtr left, right;
    BvhNode(std::vector<HitPtr>& objects, size_t start, size_t end) {  // hitable.rs:177-221
        size_t axis = (size_t)RNG().gen_range_u64(3);
        size_t object_span = end - start;
        assert(object_span > 0);
        if (object_span == 1) {
            left = objects[start];
            right = objects[start];
        } else {
            std::stable_sort(objects.begin() + start, objects.begin() + end, [axis](const HitPtr& a, const HitPtr& b) {
                AABB ba, bb;
                a->bbox(ba);
                b->bbox(bb);
                return total_cmp_key(ba.min[(int)axis]) < total_cmp_key(bb.min[(int)axis]);
            });
            if (object_span == 2) {
                left = objects[start];
                right = objects[start + 1];
            } else {
                size_t mid = start + object_span / 2;
                left = std::make_shared<BvhNode>(objects, start, mid);
                right = std::make_shared<BvhNode>(objects, mid, end);
            }
        }
        AABB box_a, box_b;
        left->bbox(box_a);
        right->bbox(box_b);
        aabb = box_a.surround(box_b);
    }
    std::string memo() const override { return "BvhNode"; }
    bool bbox(AABB& out) const override {
        out = aabb;
        return true;
    }
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {  // hitable.rs:232-240
        if (!aabb.hit(r, t_min, t_max)) return false;
        bool hit_left = left->hit(r, t_min, t_max, rec);
        bool hit_right = right->hit(r, t_min, hit_left ? rec.t : t_max, rec);
        return hit_left || hit_right;
    }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct XYRect : Hitable {  // hitable.rs:244-282
    Vec3A min, max;
    MatPtr mat;
    mutable int id = -1;
    XYRect(Vec3A mn, Vec3A mx, MatPtr m) : min(mn), max(mx), mat(std::move(m)) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        CNT().prim_tests++;
        float t = (min.z - r.o.z) / r.d.z;
        if (t < t_min || t > t_max) return false;
        Vec3A p = r.at(t);
        if (p.x < min.x || p.x > max.x || p.y < min.y || p.y > max.y) return false;
        Vec3A uv = (p - min) / (max - min);
        rec.uv = Vec2{uv.x, uv.y};
        rec.p = p;
        rec.t = t;
        rec.set_face_normal(r, Vec3A(0, 0, 1));
        rec.mat = mat;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override {
        aabb.min = Vec3A(min.x, min.y, min.z - 0.0001f);
        aabb.max = Vec3A(max.x, max.y, max.z + 0.0001f);
        return true;
    }
    std::string memo() const override { return "XYRect"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct XZRect : Hitable {  // hitable.rs:284-322
    Vec3A min, max;
    MatPtr mat;
    mutable int id = -1;
    XZRect(Vec3A mn, Vec3A mx, MatPtr m) : min(mn), max(mx), mat(std::move(m)) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        CNT().prim_tests++;
        float t = (min.y - r.o.y) / r.d.y;
        if (t < t_min || t > t_max) return false;
        Vec3A p = r.at(t);
        if (p.x < min.x || p.x > max.x || p.z < min.z || p.z > max.z) return false;
        Vec3A uv = (p - min) / (max - min);
        rec.uv = Vec2{uv.x, uv.z};
        rec.p = p;
        rec.t = t;
        rec.set_face_normal(r, Vec3A(0, 1, 0));
        rec.mat = mat;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override {
        aabb.min = Vec3A(min.x, min.y - 0.0001f, min.z);
        aabb.max = Vec3A(max.x, max.y + 0.0001f, max.z);
        return true;
    }
    std::string memo() const override { return "XZRect"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct YZRect : Hitable {  // hitable.rs:324-362
    Vec3A min, max;
    MatPtr mat;
    mutable int id = -1;
    YZRect(Vec3A mn, Vec3A mx, MatPtr m) : min(mn), max(mx), mat(std::move(m)) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        CNT().prim_tests++;
        float t = (min.x - r.o.x) / r.d.x;
        if (t < t_min || t > t_max) return false;
        Vec3A p = r.at(t);
        if (p.z < min.z || p.z > max.z || p.y < min.y || p.y > max.y) return false;
        Vec3A uv = (p - min) / (max - min);
        rec.uv = Vec2{uv.y, uv.z};
        rec.p = p;
        rec.t = t;
        rec.set_face_normal(r, Vec3A(1, 0, 0));
        rec.mat = mat;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override {
        aabb.min = Vec3A(min.x - 0.0001f, min.y, min.z);
        aabb.max = Vec3A(max.x + 0.0001f, max.y, max.z);
        return true;
    }
    std::string memo() const override { return "YZRect"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct GBox : Hitable {  // hitable.rs:364-402
    AABB aabb;
    HitableList sides;
    GBox(Vec3A min, Vec3A max, MatPtr mat) {
        sides.items = {
            std::make_shared<XYRect>(Vec3A(min.x, min.y, min.z), Vec3A(max.x, max.y, min.z), mat),
            std::make_shared<XYRect>(Vec3A(min.x, min.y, max.z), Vec3A(max.x, max.y, max.z), mat),
            std::make_shared<XZRect>(Vec3A(min.x, min.y, min.z), Vec3A(max.x, min.y, max.z), mat),
            std::make_shared<XZRect>(Vec3A(min.x, max.y, min.z), Vec3A(max.x, max.y, max.z), mat),
            std::make_shared<YZRect>(Vec3A(min.x, min.y, min.z), Vec3A(min.x, max.y, max.z), mat),
            std::make_shared<YZRect>(Vec3A(max.x, min.y, min.z), Vec3A(max.x, max.y, max.z), mat),
        };
        aabb = AABB{min, max};
    }
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        return sides.hit(r, t_min, t_max, rec);
    }
    bool bbox(AABB& out) const override {
        out = aabb;
        return true;
    }
    std::string memo() const override { return "GBox"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct Translate : Hitable {  // hitable.rs:404-436
    Vec3A offset;
    HitPtr ptr;
    Translate(Vec3A off, HitPtr p) : offset(off), ptr(std::move(p)) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        Ray moved_r{r.o - offset, r.d, r.s, r.tm};
        if (ptr->hit(moved_r, t_min, t_max, rec)) {
            rec.p += offset;
            return true;
        }
        return false;
    }
    bool bbox(AABB& aabb) const override {
        AABB g;
        if (ptr->bbox(g)) {
            aabb = AABB{g.min + offset, g.max + offset};
            return true;
        }
        return false;
    }
    std::string memo() const override { return "Translate"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

struct RotateY : Hitable {  // hitable.rs:438-520
    HitPtr ptr;
    float angle, sin_theta, cos_theta;
    bool has_box;
    AABB aabb;
    RotateY(HitPtr p, float angle_) : ptr(std::move(p)), angle(angle_) {
        float radians = angle * (PI / 180.0f);  // f32::to_radians
        sin_theta = std::sin(radians);
        cos_theta = std::cos(radians);
        AABB b;
        has_box = ptr->bbox(b);
        Vec3A mn(INFINITY, INFINITY, INFINITY), mx(-INFINITY, -INFINITY, -INFINITY);
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++)
                for (int k = 0; k < 2; k++) {
                    float x = i == 0 ? b.min.x : b.max.x;
                    float y = j == 0 ? b.min.y : b.max.y;
                    float z = k == 0 ? b.min.z : b.max.z;
                    float new_x = cos_theta * x + sin_theta * z;
                    float new_z = -sin_theta * x + cos_theta * z;
                    Vec3A tester(new_x, y, new_z);
                    for (int c = 0; c < 3; c++) {
                        mn[c] = std::fmin(mn[c], tester[c]);
                        mx[c] = std::fmax(mx[c], tester[c]);
                    }
                }
        aabb = AABB{mn, mx};
    }
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        Vec3A o = r.o, d = r.d;
        o.x = cos_theta * r.o.x - sin_theta * r.o.z;
        o.z = sin_theta * r.o.x + cos_theta * r.o.z;
        d.x = cos_theta * r.d.x - sin_theta * r.d.z;
        d.z = sin_theta * r.d.x + cos_theta * r.d.z;
        Ray rot_r{o, d, r.s, r.tm};
        if (ptr->hit(rot_r, t_min, t_max, rec)) {
            Vec3A p = rec.p, n = rec.norm;
            p.x = cos_theta * rec.p.x + sin_theta * rec.p.z;
            p.z = -sin_theta * rec.p.x + cos_theta * rec.p.z;
            n.x = cos_theta * rec.norm.x + sin_theta * rec.norm.z;
            n.z = -sin_theta * rec.norm.x + cos_theta * rec.norm.z;
            rec.p = p;
            rec.set_face_normal(rot_r, n);  // quirk A6: object-space ray vs rotated, already flipped normal
            return true;
        }
        return false;
    }
    bool bbox(AABB& out) const override {
        out = aabb;
        return has_box;
    }
    std::string memo() const override { return "RotateY"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

inline bool& media_disabled() {
    static thread_local bool v = false;
    return v;
}
struct ConstantMedium : Hitable {  // hitable.rs:523-588
    HitPtr boundary;
    MatPtr phase_fn;
    float neg_inv_density;
    mutable int id = -1;
    ConstantMedium(HitPtr b, float density, TexPtr tex)
        : boundary(std::move(b)), phase_fn(std::make_shared<Isotropic>(std::move(tex))), neg_inv_density(-1.0f / density) {}
    bool hit(const Ray& r, float t_min, float t_max, HitRecord& rec) const override {
        if (media_disabled()) return false;  // ORACLE ADDITION: deterministic closest-hit batches skip media
        CNT().prim_tests++;
        HitRecord rec_1;
        if (!boundary->hit(r, -INFINITY, INFINITY, rec_1)) return false;
        HitRecord rec_2;
        if (!boundary->hit(r, rec_1.t + 0.0001f, INFINITY, rec_2)) return false;
        if (rec_1.t < t_min) rec_1.t = t_min;
        if (rec_2.t > t_max) rec_2.t = t_max;
        if (rec_1.t >= rec_2.t) return false;
        if (rec_1.t < 0.0f) rec_1.t = 0.0f;
        float ray_len = r.d.length();
        float dist_inside_boundary = (rec_2.t - rec_1.t) * ray_len;
        float hit_dist = neg_inv_density * std::log(RNG().gen_f32());
        if (hit_dist > dist_inside_boundary) return false;
        rec.t = rec_1.t + hit_dist / ray_len;
        rec.p = r.at(rec.t);
        rec.norm = Vec3A(1, 0, 0);
        rec.front_face = true;
        rec.mat = phase_fn;
        rec.prim = id;
        return true;
    }
    bool bbox(AABB& aabb) const override { return boundary->bbox(aabb); }
    std::string memo() const override { return "ConstantMedium"; }
    void flatten(Flattener& f, uint32_t xf, bool bo) const override;
};

// ---------------------------------------------------------------------------------------- camera.rs
struct Camera {  // camera.rs:6-46 (+ thin-lens / shutter EXTENSION fields)
    Vec3A origin, horizontal, vertical, lower_left_corner;
    Vec3A u, v;
    float lens_radius = 0.0f, time0 = 0.0f, time1 = 0.0f;
    Camera() = default;
    Camera(Vec3A lookfrom, Vec3A lookat, Vec3A vup, float vfov, float aspect_ratio) {
        origin = lookfrom;
        float theta = vfov * (PI / 180.0f);
        float viewport_height = std::tan(theta / 2.0f) * 2.0f;
        float viewport_width = viewport_height * aspect_ratio;
        Vec3A w = (lookfrom - lookat).normalize();
        u = vup.cross(w).normalize();
        v = w.cross(u);
        horizontal = viewport_width * u;
        vertical = viewport_height * v;
        lower_left_corner = origin - horizontal / 2.0f - vertical / 2.0f - w;
    }
    // EXTENSION ctor (RTIOW section 12): focus_dist scales the viewport, aperture sets lens radius.
    Camera(Vec3A lookfrom, Vec3A lookat, Vec3A vup, float vfov, float aspect_ratio, float aperture, float focus_dist,
           float t0, float t1) {
        origin = lookfrom;
        float theta = vfov * (PI / 180.0f);
        float viewport_height = std::tan(theta / 2.0f) * 2.0f;
        float viewport_width = viewport_height * aspect_ratio;
        Vec3A w = (lookfrom - lookat).normalize();
        u = vup.cross(w).normalize();
        v = w.cross(u);
        horizontal = focus_dist * viewport_width * u;
        vertical = focus_dist * viewport_height * v;
        lower_left_corner = origin - horizontal / 2.0f - vertical / 2.0f - focus_dist * w;
        lens_radius = aperture / 2.0f;
        time0 = t0;
        time1 = t1;
    }
    Ray get_ray(float s, float t) const {  // camera.rs:40-46
        if (lens_radius == 0.0f && time0 == time1) {
            return Ray{origin, (lower_left_corner + s * horizontal + t * vertical - origin).normalize(), Vec2{s, t}, time0};
        }
        // EXTENSION path: lens sample (rejection in the unit disk, x then y), then shutter time.
        Vec3A offset;
        if (lens_radius != 0.0f) {
            float px, py;
            while (true) {
                px = RNG().gen_f32() * 2.0f - 1.0f;
                py = RNG().gen_f32() * 2.0f - 1.0f;
                if (px * px + py * py < 1.0f) break;
            }
            float rx = lens_radius * px, ry = lens_radius * py;
            offset = u * rx + v * ry;
        }
        float tm = time0;
        if (time0 != time1) tm = time0 + RNG().gen_f32() * (time1 - time0);
        Vec3A o = origin + offset;
        return Ray{o, (lower_left_corner + s * horizontal + t * vertical - origin - offset).normalize(), Vec2{s, t}, tm};
    }
};

// ------------------------------------------------------------------ lib.rs / demo_scene.rs:19-35: sky
enum class SkyKind { Gradient = 0, Black = 1, EnvMap = 2, Constant = 3 };
struct Sky {
    SkyKind kind = SkyKind::Gradient;
    std::shared_ptr<ImageTex> env;  // ENV_TEX, demo_scene.rs:19
    Vec3A color;                    // EXTENSION: constant sky for furnace tests
    Vec3A operator()(Vec3A d) const {
        switch (kind) {
            case SkyKind::Gradient: {  // demo_scene.rs:28-31
                float t = d.y * 0.5f + 0.5f;
                return Vec3A(1, 1, 1).lerp(Vec3A(0.5f, 0.7f, 1.0f), t);
            }
            case SkyKind::Black:
                return Vec3A();
            case SkyKind::EnvMap: {  // demo_scene.rs:22-26
                Vec2 uv = Sphere::get_uv(d);
                Vec3A c = env->value(Vec2{1.0f - uv.x, uv.y}, Vec3A());
                return c * c;
            }
            case SkyKind::Constant:
                return color;
        }
        return Vec3A();
    }
};

// -------------------------------------------------------------------------------- main.rs:36-60
struct Scene {
    HitableList world;  // what the scene function returns (a 1-element list holding the BvhNode, or a raw list)
    Camera cam;
    Sky sky;
    std::vector<MatPtr> mats;  // ORACLE ADDITION: by flat id (scenes rebuilt from an RtFlatScene only)
    std::vector<TexPtr> texs;
};

inline Vec3A ray_color(const Ray& r, const Scene& sc, int depth, int max_depth) {
    // main.rs:39 `assert!(vec3a_near_one(r.d))` panics the worker thread in the reference (the rest of
    // that column stays black).  The oracle counts the event and drops the path instead of aborting.
    if (!vec3a_near_one(r.d)) {
        CNT().unit_dir_violations++;
        return Vec3A();
    }
    if (depth > max_depth) return Vec3A();
    HitRecord rec;
    CNT().rays++;
    if (sc.world.hit(r, 1e-3f, FLT_MAX, rec)) {
        CNT().shade_calls++;
        Ray scattered{Vec3A(), Vec3A(), r.s, r.tm};
        Vec3A attenuation(1, 1, 1);
        Vec3A ret = rec.mat->emitted(rec.uv, rec.p);
        if (rec.mat->scatter(r, rec, attenuation, scattered)) {
            ret += attenuation * ray_color(scattered, sc, depth + 1, max_depth);
        }
        return ret;
    }
    return sc.sky(r.d);
}

}  // namespace orc
