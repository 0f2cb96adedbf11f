// Host-side scene construction API of the B200 backend.
//
// Mirrors the reference's constructor surface so that src/demo_scene.rs can be re-expressed 1:1:
//   Sphere{c,r,mat}            hitable.rs:56-62      XYRect/XZRect/YZRect{min,max,mat}  hitable.rs:244-362
//   GBox::new(min,max,mat)     hitable.rs:369-383    Translate{offset,ptr}              hitable.rs:404-407
//   RotateY::new(ptr,angle)    hitable.rs:447-478    ConstantMedium::new(b,density,tex) hitable.rs:529-533
//   BvhNode::new(list,s,e)     hitable.rs:177-221    HitableList = Vec<Arc<dyn Hitable>> hitable.rs:114
//   Emission/Diffuse/Lambert/Metal/Dielectric/Isotropic   material.rs:17-113
//   OrenNayar/BurleyDiffuse/RoughPlastic/Disney*          pbr.rs:12-335
//   ConstantTex/CheckerTex/PerlinTex/ImageTex             texture.rs:15-193
//   Camera::new(lookfrom,lookat,vup,vfov,aspect)          camera.rs:14-39
// Unlike the reference these objects carry NO hit/scatter code: they are a description that
// `flatten()` turns into the RtFlatScene POD arrays (include/rtgpu.h) consumed by the CUDA kernels.
// What they do reproduce exactly is everything that shapes the flattened result: bounding boxes,
// BvhNode::new's random-axis stable sorts (they define the primitive visitation order, i.e. the
// tie-break rule), and every random draw of the scene builders (rand 0.8.5 SmallRng restated).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "../../include/rtgpu.h"

namespace rtb {

// ---- float3 with glam Vec3A operation order (dot = (x*x+y*y)+z*z, normalize = v / sqrt(dot)) ----
struct V3 {
    float x = 0, y = 0, z = 0;
    V3() = default;
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : i == 1 ? y : z; }
};
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline V3 operator+(V3 a, float s) { return {a.x + s, a.y + s, a.z + s}; }
inline V3 operator-(V3 a, float s) { return {a.x - s, a.y - s, a.z - s}; }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline V3 normalize(V3 a) {
    float l = std::sqrt(dot(a, a));
    return {a.x / l, a.y / l, a.z / l};
}

// ---- rand 0.8.5 SmallRng (Xoshiro256++ / SplitMix64 seeding), used only to BUILD scenes ----------
class HostRng {
public:
    explicit HostRng(uint64_t seed = 1995) { reseed(seed); }
    void reseed(uint64_t state) {
        for (auto& w : s_) {
            state += 0x9e3779b97f4a7c15ULL;
            uint64_t z = state;
            z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
            z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
            w = z ^ (z >> 31);
        }
    }
    uint64_t next_u64() {
        uint64_t r = rotl(s_[0] + s_[3], 23) + s_[0];
        uint64_t t = s_[1] << 17;
        s_[2] ^= s_[0];
        s_[3] ^= s_[1];
        s_[1] ^= s_[2];
        s_[0] ^= s_[3];
        s_[2] ^= t;
        s_[3] = rotl(s_[3], 45);
        return r;
    }
    uint32_t next_u32() { return (uint32_t)(next_u64() >> 32); }
    float gen_f32() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
    uint64_t below_u64(uint64_t range) {  // gen_range(0..range), usize path
        uint64_t zone = (range << __builtin_clzll(range)) - 1;
        for (;;) {
            unsigned __int128 m = (unsigned __int128)next_u64() * range;
            if ((uint64_t)m <= zone) return (uint64_t)(m >> 64);
        }
    }
    uint32_t below_u32(uint32_t range) {  // gen_range(0..range), u32 path (shuffle)
        uint32_t zone = (range << __builtin_clz(range)) - 1;
        for (;;) {
            uint64_t m = (uint64_t)next_u32() * range;
            if ((uint32_t)m <= zone) return (uint32_t)(m >> 32);
        }
    }
    V3 vec3_range(float lo, float hi) {  // math.rs:17-29 vec3a_random_range
        float x = gen_f32(), y = gen_f32(), z = gen_f32();
        return V3(x, y, z) * (hi - lo) + lo;
    }

private:
    static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
    uint64_t s_[4];
};
// The reference's thread-local RNG of the thread that builds the scene (lib.rs:7-9, seed 1995).
HostRng& main_rng();

// ---- textures ------------------------------------------------------------------------------------
struct TextureNode {
    uint32_t kind = RT_TEX_CONSTANT;
    V3 col;
    std::shared_ptr<TextureNode> odd, even;
    float scale = 0;
    std::shared_ptr<RtPerlin> perlin;
    uint32_t img_w = 0, img_h = 0;
    std::shared_ptr<std::vector<uint8_t>> img;
};
using Tex = std::shared_ptr<TextureNode>;
Tex ConstantTex(V3 col);
Tex CheckerTex(V3 odd_col, V3 even_col);  // texture.rs:31-38 CheckerTex::new
Tex CheckerTex(Tex odd, Tex even);
Tex PerlinTex(float scale);  // texture.rs:154-160: consumes main_rng() exactly like Perlin::default
Tex ImageTex(uint32_t w, uint32_t h, const uint8_t* rgb8);
Tex ImageTexFromPpm(const std::string& path);  // decoded texel array (see tools/decode_assets.py)

// ---- materials -----------------------------------------------------------------------------------
struct MaterialNode {
    uint32_t kind = RT_MAT_DIFFUSE;
    Tex tex0, tex1;
    float p[4] = {0, 0, 0, 0};
};
using Mat = std::shared_ptr<MaterialNode>;
Mat Emission(Tex emit);
Mat Diffuse(Tex albedo);
Mat Lambert(Tex albedo);
Mat Metal(V3 albedo, float fuzz);
Mat Dielectric(float ior);
Mat Isotropic(Tex albedo);
Mat OrenNayar(Tex albedo, float roughness);
Mat BurleyDiffuse(Tex albedo, float roughness);
Mat RoughPlastic(Tex spec_color, Tex diff_color, float roughness, float eta);
Mat DisneyDiffuse(Tex albedo, float roughness, float subsurface);
Mat DisneyMetal(Tex albedo, float roughness, float anisotropic, float rot);
Mat DisneySheen(Tex albedo, float tint);
Mat DisneyClearcoat(float clearcoat_gloss);

// ---- hitables ------------------------------------------------------------------------------------
struct Box3 {
    V3 min, max;
};
enum class HKind { Sphere, MovingSphere, XYRect, XZRect, YZRect, GBox, Translate, RotateY, Medium, Bvh, List };
struct HitableNode {
    HKind kind;
    V3 a, b;  // sphere: a = centre (b = centre1 for moving); rect/box: a = min, b = max; translate: a = offset
    float r = 0, time0 = 0, time1 = 0;
    float angle = 0, sin_theta = 0, cos_theta = 1;
    float neg_inv_density = 0;
    Mat mat;
    std::vector<std::shared_ptr<HitableNode>> kids;  // wrapper: 1 child; GBox: 6 sides; Bvh/List: items in visitation order
    Box3 box;
    bool has_box = true;
};
using Hit = std::shared_ptr<HitableNode>;
using HitableList = std::vector<Hit>;

Hit Sphere(V3 c, float r, Mat mat);
Hit MovingSphere(V3 c0, V3 c1, float time0, float time1, float r, Mat mat);  // EXTENSION
Hit XYRect(V3 min, V3 max, Mat mat);
Hit XZRect(V3 min, V3 max, Mat mat);
Hit YZRect(V3 min, V3 max, Mat mat);
Hit GBox(V3 min, V3 max, Mat mat);
Hit Translate(V3 offset, Hit ptr);
Hit RotateY(Hit ptr, float angle_degrees);
Hit ConstantMedium(Hit boundary, float density, Tex phase_albedo);
// BvhNode::new(objects, start, end): sorts objects[start..end) in place exactly like the reference
// (random axis from main_rng() at EVERY recursion level, stable sort by bbox.min[axis] with
// total_cmp, split at start + span/2) and returns a node whose visitation order is the final order.
Hit BvhNode(HitableList& objects, size_t start, size_t end);

// ---- camera --------------------------------------------------------------------------------------
RtCamera Camera(V3 lookfrom, V3 lookat, V3 vup, float vfov, float aspect_ratio);
// EXTENSION (RTIOW section 12 thin lens + RTNW section 2 shutter)
RtCamera CameraLens(V3 lookfrom, V3 lookat, V3 vup, float vfov, float aspect_ratio, float aperture, float focus_dist,
                    float time0, float time1);

// ---- flattened scene that owns its arrays --------------------------------------------------------
struct FlatSceneOwner {
    std::vector<RtPrim> prims;
    std::vector<RtXformOp> ops;
    std::vector<RtXformChain> chains;
    std::vector<RtMaterial> materials;
    std::vector<RtTexture> textures;
    std::vector<RtPerlin> perlin;
    std::vector<RtImage> images;
    std::vector<std::shared_ptr<std::vector<uint8_t>>> image_store;
    RtFlatScene flat{};
};
struct SkySpec {
    uint32_t kind = RT_SKY_GRADIENT;
    Tex env;  // RT_SKY_ENVMAP: an ImageTex
    V3 color;
};
// world: what a scene function returns (`Vec<Arc<dyn Hitable>>`, demo_scene.rs:37).
std::unique_ptr<FlatSceneOwner> flatten(const HitableList& world, const RtCamera& cam, const SkySpec& sky);

// ---- the reference's demo scenes + BASELINE.json config scenes ----------------------------------
std::unique_ptr<FlatSceneOwner> build_named_scene(const std::string& name, float aspect_ratio, const std::string& asset_dir);

}  // namespace rtb
