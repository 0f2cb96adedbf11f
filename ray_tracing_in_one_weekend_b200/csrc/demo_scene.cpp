// The reference's scene functions (src/demo_scene.rs) re-expressed on the host API of
// scene_api.hpp, plus the BASELINE.json config scenes that have no counterpart in the reference
// snapshot composed through the same API.
//
// Every scene function has the reference's shape `fn(aspect_ratio) -> (Vec<Arc<dyn Hitable>>, Camera)`
// and sets the sky like `SKY_COLOR.set(..)`; here the triple is flattened straight away.
#include <stdexcept>

#include "scene_api.hpp"

namespace rtb {
namespace {

struct SceneOut {
    HitableList world;
    RtCamera cam;
    SkySpec sky;
};

HitableList build_bvh(HitableList& world) {  // demo_scene.rs:223-227
    return HitableList{BvhNode(world, 0, world.size())};
}
Tex rgb(float r, float g, float b) { return ConstantTex(V3(r, g, b)); }

enum class Cover { Verbatim, Textured, Defocus };

// demo_scene.rs:37-86 `sphere_scene` (the "cover"): 4 + 23x23 spheres, local SmallRng(95).
// Textured = BASELINE config 2: checker ground (the commented line demo_scene.rs:40), an
// earth-mapped Diffuse sphere and a Perlin Diffuse sphere.  Defocus = thin-lens EXTENSION.
SceneOut sphere_scene(float aspect, const std::string& assets, Cover variant) {
    SceneOut s;
    s.sky.kind = RT_SKY_GRADIENT;
    Tex checker = variant == Cover::Textured ? CheckerTex(V3(0.2f, 0.3f, 0.1f), V3(0.9f, 0.9f, 0.9f)) : nullptr;
    Tex perlin = PerlinTex(4.0f);
    Tex earth_map = ImageTexFromPpm(assets + "/earthmap.ppm");
    Mat material_ground = Diffuse(variant == Cover::Textured ? checker : perlin);
    Mat material_1 = variant == Cover::Textured ? Diffuse(earth_map) : Emission(earth_map);
    Mat material_2 = Dielectric(1.5f);
    Mat material_3 = Metal(V3(0.8f, 0.6f, 0.2f), 0.0f);
    HitableList world = {
        Sphere(V3(1.0f, -1000.0f, -1.0f), 1000.0f, material_ground),
        Sphere(V3(0.0f, 1.0f, 3.0f), 1.0f, material_1),
        Sphere(V3(-4.0f, 1.0f, 0.0f), 1.0f, variant == Cover::Textured ? Diffuse(perlin) : material_2),
        Sphere(V3(4.0f, 1.0f, 0.0f), 1.0f, material_3),
    };
    HostRng rng(95);
    for (int a = -11; a <= 11; a++) {
        for (int b = -11; b <= 11; b++) {
            float choose_mat = rng.gen_f32();
            float cx = (float)a + 0.9f * rng.gen_f32();
            float cz = (float)b + 0.9f * rng.gen_f32();
            Mat mat;
            if (choose_mat < 0.8f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                float b0 = rng.gen_f32(), b1 = rng.gen_f32(), b2 = rng.gen_f32();
                mat = Diffuse(ConstantTex(V3(a0, a1, a2) * V3(b0, b1, b2)));
            } else if (choose_mat < 0.95f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                V3 albedo = V3(a0, a1, a2) * 0.5f + 0.5f;
                float fuzz = rng.gen_f32();
                mat = Metal(albedo, fuzz);
            } else {
                mat = material_2;
            }
            world.push_back(Sphere(V3(cx, 0.2f, cz), 0.2f, mat));
        }
    }
    if (variant == Cover::Defocus)
        s.cam = CameraLens(V3(13, 2, 3), V3(0, 0, 0), V3(0, 1, 0), 20.0f, aspect, 0.1f, 10.0f, 0.0f, 0.0f);
    else
        s.cam = Camera(V3(13, 2, 3), V3(0, 0, 0), V3(0, 1, 0), 20.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

// BASELINE config 5 (EXTENSION): the cover generator on a grid x grid lattice squeezed into the
// same [-11.5, 11.5]^2 footprint; Diffuse spheres move during the shutter interval [0,1).
SceneOut cover_motion(float aspect, const std::string& assets, int grid) {
    SceneOut s;
    s.sky.kind = RT_SKY_GRADIENT;
    Tex perlin = PerlinTex(4.0f);
    Tex earth_map = ImageTexFromPpm(assets + "/earthmap.ppm");
    Mat material_2 = Dielectric(1.5f);
    HitableList world = {
        Sphere(V3(1.0f, -1000.0f, -1.0f), 1000.0f, Diffuse(perlin)),
        Sphere(V3(0.0f, 1.0f, 3.0f), 1.0f, Emission(earth_map)),
        Sphere(V3(-4.0f, 1.0f, 0.0f), 1.0f, material_2),
        Sphere(V3(4.0f, 1.0f, 0.0f), 1.0f, Metal(V3(0.8f, 0.6f, 0.2f), 0.0f)),
    };
    HostRng rng(95);
    const float cell = 23.0f / (float)grid;
    const float rad = 0.2f * cell;
    for (int ia = 0; ia < grid; ia++) {
        for (int ib = 0; ib < grid; ib++) {
            float a = -11.0f + (float)ia * cell;
            float b = -11.0f + (float)ib * cell;
            float choose_mat = rng.gen_f32();
            float cx = a + 0.9f * cell * rng.gen_f32();
            float cz = b + 0.9f * cell * rng.gen_f32();
            V3 center(cx, rad, cz);
            if (choose_mat < 0.8f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                float b0 = rng.gen_f32(), b1 = rng.gen_f32(), b2 = rng.gen_f32();
                Mat mat = Diffuse(ConstantTex(V3(a0, a1, a2) * V3(b0, b1, b2)));
                V3 c1 = center + V3(0.0f, 0.5f * rng.gen_f32() * cell, 0.0f);
                world.push_back(MovingSphere(center, c1, 0.0f, 1.0f, rad, mat));
            } else if (choose_mat < 0.95f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                V3 albedo = V3(a0, a1, a2) * 0.5f + 0.5f;
                float fuzz = rng.gen_f32();
                world.push_back(Sphere(center, rad, Metal(albedo, fuzz)));
            } else {
                world.push_back(Sphere(center, rad, material_2));
            }
        }
    }
    s.cam = CameraLens(V3(13, 2, 3), V3(0, 0, 0), V3(0, 1, 0), 20.0f, aspect, 0.0f, 1.0f, 0.0f, 1.0f);
    s.world = build_bvh(world);
    return s;
}

SceneOut simple_light_scene(float aspect) {  // demo_scene.rs:88-110
    SceneOut s;
    s.sky.kind = RT_SKY_BLACK;
    Mat mat_perlin = Diffuse(PerlinTex(4.0f));
    Mat light = Emission(rgb(4, 4, 4));
    HitableList world = {
        Sphere(V3(1.0f, -1000.0f, -1.0f), 1000.0f, mat_perlin),
        Sphere(V3(0.0f, 2.0f, 0.0f), 2.0f, mat_perlin),
        Sphere(V3(0.0f, 6.5f, 0.0f), 2.0f, light),
        XYRect(V3(3, 1, -2), V3(5, 3, -2), light),
    };
    s.cam = Camera(V3(26, 3, 6), V3(0, 0, 0), V3(0, 1, 0), 20.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

// demo_scene.rs:112-148.  solid=false is the shipped smoke variant (the two transformed boxes
// only bound ConstantMedium volumes); solid=true pushes the boxes themselves (BASELINE config 3).
SceneOut cornell_box(float aspect, bool solid) {
    SceneOut s;
    s.sky.kind = RT_SKY_BLACK;
    Mat red = Diffuse(rgb(0.65f, 0.05f, 0.05f));
    Mat white = Diffuse(rgb(0.73f, 0.73f, 0.73f));
    Mat green = Diffuse(rgb(0.12f, 0.45f, 0.15f));
    Mat light = Emission(rgb(7, 7, 7));
    Hit box_1 = Translate(V3(265, 0, 295), RotateY(GBox(V3(0, 0, 0), V3(165, 330, 165), white), 15.0f));
    Hit box_2 = Translate(V3(130, 0, 65), RotateY(GBox(V3(0, 0, 0), V3(165, 165, 165), white), -18.0f));
    HitableList world = {
        XZRect(V3(113, 554, 127), V3(443, 554, 432), light),
        XYRect(V3(0, 0, 555), V3(555, 555, 555), white),
        XZRect(V3(0, 0, 0), V3(555, 0, 555), white),
        XZRect(V3(0, 555, 0), V3(555, 555, 555), white),
        YZRect(V3(0, 0, 0), V3(0, 555, 555), red),
        YZRect(V3(555, 0, 0), V3(555, 555, 555), green),
        solid ? box_1 : ConstantMedium(box_1, 0.01f, rgb(0, 0, 0)),
        solid ? box_2 : ConstantMedium(box_2, 0.01f, rgb(1, 1, 1)),
    };
    s.cam = Camera(V3(278, 278, -800), V3(278, 278, 0), V3(0, 1, 0), 40.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

SceneOut final_scene(float aspect, const std::string& assets) {  // demo_scene.rs:150-221
    SceneOut s;
    s.sky.kind = RT_SKY_BLACK;
    Mat ground = Diffuse(rgb(0.48f, 0.83f, 0.53f));
    Mat white = Diffuse(rgb(0.73f, 0.73f, 0.73f));
    Mat brown = BurleyDiffuse(rgb(0.7f, 0.3f, 0.1f), 0.9f);
    Mat light = Emission(rgb(7, 7, 7));
    Mat dielectric = Dielectric(1.5f);
    Mat metal = Metal(V3(0.8f, 0.8f, 0.9f), 1.0f);
    Hit earth = Sphere(V3(400, 200, 400), 100.0f, Diffuse(ImageTexFromPpm(assets + "/earthmap.ppm")));
    Hit perlin_sphere = Sphere(V3(220, 280, 300), 80.0f, Diffuse(PerlinTex(0.1f)));
    Hit brown_sphere = Sphere(V3(400, 400, 200), 50.0f, brown);
    Hit dielectric_sphere = Sphere(V3(260, 150, 45), 50.0f, dielectric);
    Hit metal_sphere = Sphere(V3(0, 150, 145), 50.0f, metal);
    Hit boundary = Sphere(V3(360, 150, 145), 50.0f, dielectric);
    Hit medium = ConstantMedium(boundary, 0.2f, rgb(0.2f, 0.4f, 0.9f));
    HitableList spheres;
    for (int i = 0; i < 1000; i++) spheres.push_back(Sphere(main_rng().vec3_range(0.0f, 165.0f), 10.0f, white));
    Hit sph = Translate(V3(-100, 270, 395), RotateY(BvhNode(spheres, 0, 1000), 15.0f));
    HitableList boxes;
    for (int i = 0; i < 20; i++) {
        for (int j = 0; j < 20; j++) {
            float w = 100.0f;
            float x0 = -1000.0f + (float)i * w, z0 = -1000.0f + (float)j * w, y0 = 0.0f;
            float x1 = x0 + w;
            float y1 = main_rng().gen_f32() * 100.0f + 1.0f;
            float z1 = z0 + w;
            boxes.push_back(GBox(V3(x0, y0, z0), V3(x1, y1, z1), ground));
        }
    }
    Hit bx = BvhNode(boxes, 0, 400);
    HitableList world = {
        XZRect(V3(123, 544, 147), V3(423, 554, 412), light),
        sph, earth, perlin_sphere, brown_sphere, dielectric_sphere, metal_sphere, bx, boundary, medium,
    };
    s.cam = Camera(V3(478, 278, -600), V3(278, 278, 0), V3(0, 1, 0), 40.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

SceneOut test_sphere(float aspect) {  // demo_scene.rs:229-244: what the shipped main() renders; raw list, no BVH
    SceneOut s;
    s.sky.kind = RT_SKY_GRADIENT;
    Mat ground = Lambert(rgb(0.5f, 0.5f, 0.5f));
    s.world = {Sphere(V3(0.0f, -100.5f, -1.0f), 100.0f, ground), Sphere(V3(0.0f, 0.0f, -1.0f), 0.5f, ground)};
    s.cam = Camera(V3(0, 0, 0), V3(0, 0, -1), V3(0, 1, 0), 90.0f, aspect);
    return s;
}

// BASELINE config 4: every pbr.rs material under the newport_loft environment (tex_sky_color,
// demo_scene.rs:21-26, is dead code in the reference; no shipped scene uses it).
SceneOut pbr_ibl(float aspect, const std::string& assets) {
    SceneOut s;
    s.sky.kind = RT_SKY_ENVMAP;
    s.sky.env = ImageTexFromPpm(assets + "/newport_loft.ppm");
    Mat row[8] = {
        OrenNayar(rgb(0.8f, 0.3f, 0.3f), 0.5f),
        BurleyDiffuse(rgb(0.3f, 0.8f, 0.3f), 0.6f),
        RoughPlastic(rgb(1, 1, 1), rgb(0.2f, 0.3f, 0.8f), 0.2f, 1.5f),
        DisneyDiffuse(rgb(0.8f, 0.7f, 0.2f), 0.5f, 0.5f),
        DisneyMetal(rgb(0.95f, 0.64f, 0.54f), 0.3f, 0.0f, 0.0f),
        DisneyMetal(rgb(0.9f, 0.9f, 0.9f), 0.4f, 0.8f, 0.25f),
        DisneySheen(rgb(0.6f, 0.2f, 0.7f), 0.5f),
        DisneyClearcoat(0.8f),
    };
    HitableList world = {Sphere(V3(0.0f, -100.5f, -1.0f), 100.0f, RoughPlastic(rgb(1, 1, 1), rgb(0.5f, 0.5f, 0.5f), 0.3f, 1.5f))};
    for (int k = 0; k < 8; k++) world.push_back(Sphere(V3(((float)k - 3.5f) * 1.2f, 0.0f, -1.0f), 0.5f, row[k]));
    s.cam = Camera(V3(0, 1.5f, 7), V3(0, 0, -1), V3(0, 1, 0), 40.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

// Furnace scenes for the statistical KAT K10: one convex sphere under a constant white sky.
// EXTENSION scene for the positions the reference allows a ConstantMedium in (any Hitable wraps any Hitable,
// hitable.rs:404-436,438-520,523-588) beyond what its own scenes use: media UNDER Translate / RotateY wrappers, a
// boundary with wrappers of its own below such a medium, a sphere boundary.  The Cornell room of demo_scene.rs:112-148.
SceneOut smoke_instances(float aspect) {
    SceneOut s;
    s.sky.kind = RT_SKY_BLACK;
    Mat red = Diffuse(rgb(0.65f, 0.05f, 0.05f));
    Mat white = Diffuse(rgb(0.73f, 0.73f, 0.73f));
    Mat green = Diffuse(rgb(0.12f, 0.45f, 0.15f));
    Mat light = Emission(rgb(7, 7, 7));
    // Translate(ConstantMedium(Sphere))
    Hit ball = Translate(V3(400, 120, 200), ConstantMedium(Sphere(V3(0, 0, 0), 100.0f, white), 0.02f, rgb(0.9f, 0.9f, 0.9f)));
    // Translate(RotateY(ConstantMedium(GBox))): the boundary sits directly below the medium
    Hit tall = Translate(V3(265, 0, 295), RotateY(ConstantMedium(GBox(V3(0, 0, 0), V3(165, 330, 165), white), 0.01f, rgb(0, 0, 0)), 15.0f));
    // RotateY(ConstantMedium(Translate(GBox))): the boundary carries a wrapper of its own below the medium
    Hit low = RotateY(ConstantMedium(Translate(V3(60, 0, 130), GBox(V3(0, 0, 0), V3(150, 150, 150), white)), 0.015f, rgb(0.2f, 0.4f, 0.9f)), -10.0f);
    HitableList world = {
        XZRect(V3(113, 554, 127), V3(443, 554, 432), light),
        XYRect(V3(0, 0, 555), V3(555, 555, 555), white),
        XZRect(V3(0, 0, 0), V3(555, 0, 555), white),
        XZRect(V3(0, 555, 0), V3(555, 555, 555), white),
        YZRect(V3(0, 0, 0), V3(0, 555, 555), red),
        YZRect(V3(555, 0, 0), V3(555, 555, 555), green),
        ball, tall, low,
    };
    s.cam = Camera(V3(278, 278, -800), V3(278, 278, 0), V3(0, 1, 0), 40.0f, aspect);
    s.world = build_bvh(world);
    return s;
}

// Well-formed scenes the backend must REFUSE (RT_ERR_UNSUPPORTED) because the reference's picture of them depends on
// stale HitRecord fields (quirk A4) or on a stochastic boundary: used by the validation tests.
SceneOut unsupported_scene(float aspect, const std::string& which, const std::string& assets) {
    SceneOut s;
    s.sky.kind = RT_SKY_GRADIENT;
    Mat white = Diffuse(rgb(0.73f, 0.73f, 0.73f));
    if (which == "nested_medium") {
        Hit inner = ConstantMedium(Sphere(V3(0, 0, -1), 0.5f, white), 1.0f, rgb(1, 1, 1));
        s.world = {ConstantMedium(inner, 1.0f, rgb(1, 1, 1))};
    } else if (which == "disney_metal_rect") {
        s.world = {XYRect(V3(-1, -1, -2), V3(1, 1, -2), DisneyMetal(rgb(0.6f, 0.5f, 0.4f), 0.4f, 0.5f, 0.1f))};
    } else if (which == "image_isotropic_medium") {
        s.world = {ConstantMedium(Sphere(V3(0, 0, -1), 0.5f, white), 1.0f, ImageTexFromPpm(assets + "/earthmap.ppm"))};
    } else {
        throw std::runtime_error("unknown unsupported-scene " + which);
    }
    s.cam = Camera(V3(0, 0, 0), V3(0, 0, -1), V3(0, 1, 0), 90.0f, aspect);
    return s;
}

SceneOut furnace(float aspect, const std::string& which) {
    SceneOut s;
    s.sky.kind = RT_SKY_CONSTANT;
    s.sky.color = V3(1, 1, 1);
    Tex a = rgb(0.6f, 0.5f, 0.4f);
    Mat m;
    if (which == "diffuse") m = Diffuse(a);
    else if (which == "lambert") m = Lambert(a);
    else if (which == "metal") m = Metal(V3(0.6f, 0.5f, 0.4f), 0.0f);
    else if (which == "dielectric") m = Dielectric(1.5f);
    else if (which == "oren_nayar") m = OrenNayar(a, 0.5f);
    else if (which == "burley") m = BurleyDiffuse(a, 0.5f);
    else if (which == "rough_plastic") m = RoughPlastic(rgb(1, 1, 1), a, 0.3f, 1.5f);
    else if (which == "disney_diffuse") m = DisneyDiffuse(a, 0.5f, 0.5f);
    else if (which == "disney_metal") m = DisneyMetal(a, 0.4f, 0.5f, 0.1f);
    else if (which == "disney_sheen") m = DisneySheen(a, 0.5f);
    else if (which == "disney_clearcoat") m = DisneyClearcoat(0.5f);
    else throw std::runtime_error("unknown furnace material " + which);
    s.world = {Sphere(V3(0.0f, 0.0f, -1.0f), 0.5f, m)};
    s.cam = Camera(V3(0, 0, 0), V3(0, 0, -1), V3(0, 1, 0), 20.0f, aspect);
    return s;
}

}  // namespace

std::unique_ptr<FlatSceneOwner> build_named_scene(const std::string& name, float aspect, const std::string& assets) {
    main_rng().reseed(1995);  // a fresh process' main-thread generator (lib.rs:8)
    SceneOut s;
    if (name == "test_sphere") s = test_sphere(aspect);
    else if (name == "sphere_scene") s = sphere_scene(aspect, assets, Cover::Verbatim);
    else if (name == "textured") s = sphere_scene(aspect, assets, Cover::Textured);
    else if (name == "cover_defocus") s = sphere_scene(aspect, assets, Cover::Defocus);
    else if (name == "cover_100k_motion") s = cover_motion(aspect, assets, 316);
    else if (name == "cover_4k_motion") s = cover_motion(aspect, assets, 64);
    else if (name == "simple_light_scene") s = simple_light_scene(aspect);
    else if (name == "cornell_box") s = cornell_box(aspect, false);
    else if (name == "cornell_box_solid") s = cornell_box(aspect, true);
    else if (name == "final_scene") s = final_scene(aspect, assets);
    else if (name == "pbr_ibl") s = pbr_ibl(aspect, assets);
    else if (name == "smoke_instances") s = smoke_instances(aspect);
    else if (name.rfind("unsupported_", 0) == 0) s = unsupported_scene(aspect, name.substr(12), assets);
    else if (name.rfind("furnace_", 0) == 0) s = furnace(aspect, name.substr(8));
    else throw std::runtime_error("unknown scene '" + name + "'");
    return flatten(s.world, s.cam, s.sky);
}

}  // namespace rtb
