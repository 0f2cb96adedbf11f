// Host scene API: constructors, bounding boxes, BvhNode::new ordering, flattener.
// See scene_api.hpp for the reference citations.  Compiled with -ffp-contract=off so that every
// float operation is rounded individually, as in the reference's default x86-64 Rust build.
#include "scene_api.hpp"

#include <algorithm>
#include <cstdio>
#include <map>
#include <stdexcept>

namespace rtb {

HostRng& main_rng() {
    static thread_local HostRng rng(1995);
    return rng;
}

// ------------------------------------------------------------------------------------- textures
Tex ConstantTex(V3 col) {
    auto t = std::make_shared<TextureNode>();
    t->kind = RT_TEX_CONSTANT;
    t->col = col;
    return t;
}
Tex CheckerTex(Tex odd, Tex even) {
    auto t = std::make_shared<TextureNode>();
    t->kind = RT_TEX_CHECKER;
    t->odd = std::move(odd);
    t->even = std::move(even);
    return t;
}
Tex CheckerTex(V3 odd_col, V3 even_col) { return CheckerTex(ConstantTex(odd_col), ConstantTex(even_col)); }

Tex PerlinTex(float scale) {
    // texture.rs:60-91: 256 gradients in [-1,1)^3 (x,y,z draws each), then three successive
    // Fisher-Yates shuffles of the SAME index vector (perm_y continues from perm_x's order).
    auto t = std::make_shared<TextureNode>();
    t->kind = RT_TEX_PERLIN;
    t->scale = scale;
    auto tab = std::make_shared<RtPerlin>();
    HostRng& rng = main_rng();
    for (int i = 0; i < 256; i++) {
        V3 g = rng.vec3_range(-1.0f, 1.0f);
        tab->rand_vec[i][0] = g.x;
        tab->rand_vec[i][1] = g.y;
        tab->rand_vec[i][2] = g.z;
        tab->rand_vec[i][3] = 0.0f;
    }
    uint32_t p[256];
    for (uint32_t i = 0; i < 256; i++) p[i] = i;
    uint8_t* dst[3] = {tab->perm_x, tab->perm_y, tab->perm_z};
    for (auto* d : dst) {
        for (uint32_t i = 255; i >= 1; i--) std::swap(p[i], p[rng.below_u32(i + 1)]);
        for (int i = 0; i < 256; i++) d[i] = (uint8_t)p[i];
    }
    t->perlin = tab;
    return t;
}
Tex ImageTex(uint32_t w, uint32_t h, const uint8_t* rgb8) {
    auto t = std::make_shared<TextureNode>();
    t->kind = RT_TEX_IMAGE;
    t->img_w = w;
    t->img_h = h;
    t->img = std::make_shared<std::vector<uint8_t>>(rgb8, rgb8 + (size_t)w * h * 3);
    return t;
}
Tex ImageTexFromPpm(const std::string& path) {
    FILE* fh = std::fopen(path.c_str(), "rb");
    if (!fh) throw std::runtime_error("cannot open image " + path);
    char magic[3] = {0, 0, 0};
    int w = 0, h = 0, maxv = 0;
    bool ok = std::fscanf(fh, "%2s %d %d %d", magic, &w, &h, &maxv) == 4 && magic[0] == 'P' && magic[1] == '6' && maxv == 255;
    if (ok) std::fgetc(fh);
    std::vector<uint8_t> buf;
    if (ok) {
        buf.resize((size_t)w * h * 3);
        ok = std::fread(buf.data(), 1, buf.size(), fh) == buf.size();
    }
    std::fclose(fh);
    if (!ok) throw std::runtime_error("not a binary 8-bit PPM: " + path);
    return ImageTex((uint32_t)w, (uint32_t)h, buf.data());
}

// ------------------------------------------------------------------------------------ materials
static Mat make_mat(uint32_t kind, Tex t0 = nullptr, Tex t1 = nullptr, float p0 = 0, float p1 = 0, float p2 = 0, float p3 = 0) {
    auto m = std::make_shared<MaterialNode>();
    m->kind = kind;
    m->tex0 = std::move(t0);
    m->tex1 = std::move(t1);
    m->p[0] = p0; m->p[1] = p1; m->p[2] = p2; m->p[3] = p3;
    return m;
}
Mat Emission(Tex e) { return make_mat(RT_MAT_EMISSION, std::move(e)); }
Mat Diffuse(Tex a) { return make_mat(RT_MAT_DIFFUSE, std::move(a)); }
Mat Lambert(Tex a) { return make_mat(RT_MAT_LAMBERT, std::move(a)); }
Mat Metal(V3 albedo, float fuzz) { return make_mat(RT_MAT_METAL, nullptr, nullptr, albedo.x, albedo.y, albedo.z, fuzz); }
Mat Dielectric(float ior) { return make_mat(RT_MAT_DIELECTRIC, nullptr, nullptr, ior); }
Mat Isotropic(Tex a) { return make_mat(RT_MAT_ISOTROPIC, std::move(a)); }
Mat OrenNayar(Tex a, float r) { return make_mat(RT_MAT_OREN_NAYAR, std::move(a), nullptr, r); }
Mat BurleyDiffuse(Tex a, float r) { return make_mat(RT_MAT_BURLEY_DIFFUSE, std::move(a), nullptr, r); }
Mat RoughPlastic(Tex s, Tex d, float r, float eta) { return make_mat(RT_MAT_ROUGH_PLASTIC, std::move(s), std::move(d), r, eta); }
Mat DisneyDiffuse(Tex a, float r, float ss) { return make_mat(RT_MAT_DISNEY_DIFFUSE, std::move(a), nullptr, r, ss); }
Mat DisneyMetal(Tex a, float r, float an, float rot) { return make_mat(RT_MAT_DISNEY_METAL, std::move(a), nullptr, r, an, rot); }
Mat DisneySheen(Tex a, float tint) { return make_mat(RT_MAT_DISNEY_SHEEN, std::move(a), nullptr, tint); }
Mat DisneyClearcoat(float g) { return make_mat(RT_MAT_DISNEY_CLEARCOAT, nullptr, nullptr, g); }

// ------------------------------------------------------------------------------------- hitables
static Box3 surround(const Box3& a, const Box3& b) {  // math.rs:115-130
    Box3 o;
    o.min = V3(std::fmin(a.min.x, b.min.x), std::fmin(a.min.y, b.min.y), std::fmin(a.min.z, b.min.z));
    o.max = V3(std::fmax(a.max.x, b.max.x), std::fmax(a.max.y, b.max.y), std::fmax(a.max.z, b.max.z));
    return o;
}
static Hit make_hit(HKind k) {
    auto h = std::make_shared<HitableNode>();
    h->kind = k;
    return h;
}
Hit Sphere(V3 c, float r, Mat mat) {
    Hit h = make_hit(HKind::Sphere);
    h->a = c;
    h->r = r;
    h->mat = std::move(mat);
    h->box = Box3{c - r, c + r};  // hitable.rs:104-108
    return h;
}
Hit MovingSphere(V3 c0, V3 c1, float time0, float time1, float r, Mat mat) {
    Hit h = make_hit(HKind::MovingSphere);
    h->a = c0;
    h->b = c1;
    h->r = r;
    h->time0 = time0;
    h->time1 = time1;
    h->mat = std::move(mat);
    h->box = surround(Box3{c0 - r, c0 + r}, Box3{c1 - r, c1 + r});
    return h;
}
static Hit make_rect(HKind k, V3 mn, V3 mx, Mat mat) {
    Hit h = make_hit(k);
    h->a = mn;
    h->b = mx;
    h->mat = std::move(mat);
    // hitable.rs:274-278,314-318,354-358: thin axis padded by 0.0001
    V3 lo = mn, hi = mx;
    if (k == HKind::XYRect) { lo.z = mn.z - 0.0001f; hi.z = mx.z + 0.0001f; }
    if (k == HKind::XZRect) { lo.y = mn.y - 0.0001f; hi.y = mx.y + 0.0001f; }
    if (k == HKind::YZRect) { lo.x = mn.x - 0.0001f; hi.x = mx.x + 0.0001f; }
    h->box = Box3{lo, hi};
    return h;
}
Hit XYRect(V3 mn, V3 mx, Mat m) { return make_rect(HKind::XYRect, mn, mx, std::move(m)); }
Hit XZRect(V3 mn, V3 mx, Mat m) { return make_rect(HKind::XZRect, mn, mx, std::move(m)); }
Hit YZRect(V3 mn, V3 mx, Mat m) { return make_rect(HKind::YZRect, mn, mx, std::move(m)); }
Hit GBox(V3 mn, V3 mx, Mat mat) {  // hitable.rs:369-383 (side order is semantic: later side wins ties)
    Hit h = make_hit(HKind::GBox);
    h->a = mn;
    h->b = mx;
    h->kids = {
        XYRect(V3(mn.x, mn.y, mn.z), V3(mx.x, mx.y, mn.z), mat), XYRect(V3(mn.x, mn.y, mx.z), V3(mx.x, mx.y, mx.z), mat),
        XZRect(V3(mn.x, mn.y, mn.z), V3(mx.x, mn.y, mx.z), mat), XZRect(V3(mn.x, mx.y, mn.z), V3(mx.x, mx.y, mx.z), mat),
        YZRect(V3(mn.x, mn.y, mn.z), V3(mn.x, mx.y, mx.z), mat), YZRect(V3(mx.x, mn.y, mn.z), V3(mx.x, mx.y, mx.z), mat),
    };
    h->box = Box3{mn, mx};
    return h;
}
Hit Translate(V3 offset, Hit ptr) {
    Hit h = make_hit(HKind::Translate);
    h->a = offset;
    h->has_box = ptr->has_box;
    h->box = Box3{ptr->box.min + offset, ptr->box.max + offset};  // hitable.rs:420-431
    h->kids = {std::move(ptr)};
    return h;
}
Hit RotateY(Hit ptr, float angle) {  // hitable.rs:447-478
    Hit h = make_hit(HKind::RotateY);
    h->angle = angle;
    float radians = angle * (3.14159265358979323846f / 180.0f);
    h->sin_theta = std::sin(radians);
    h->cos_theta = std::cos(radians);
    h->has_box = ptr->has_box;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    const Box3& b = ptr->box;
    for (int i = 0; i < 2; i++)
        for (int j = 0; j < 2; j++)
            for (int k = 0; k < 2; k++) {
                float x = i == 0 ? b.min.x : b.max.x;
                float y = j == 0 ? b.min.y : b.max.y;
                float z = k == 0 ? b.min.z : b.max.z;
                float t[3] = {h->cos_theta * x + h->sin_theta * z, y, -h->sin_theta * x + h->cos_theta * z};
                for (int c = 0; c < 3; c++) {
                    mn[c] = std::fmin(mn[c], t[c]);
                    mx[c] = std::fmax(mx[c], t[c]);
                }
            }
    h->box = Box3{V3(mn[0], mn[1], mn[2]), V3(mx[0], mx[1], mx[2])};
    h->kids = {std::move(ptr)};
    return h;
}
Hit ConstantMedium(Hit boundary, float density, Tex phase_albedo) {  // hitable.rs:529-533
    Hit h = make_hit(HKind::Medium);
    h->neg_inv_density = -1.0f / density;
    h->mat = Isotropic(std::move(phase_albedo));
    h->has_box = boundary->has_box;
    h->box = boundary->box;
    h->kids = {std::move(boundary)};
    return h;
}

static int32_t total_cmp_key(float f) {  // f32::total_cmp
    int32_t i;
    std::memcpy(&i, &f, 4);
    i ^= (int32_t)(((uint32_t)(i >> 31)) >> 1);
    return i;
}
static void bvh_order(HitableList& objects, size_t start, size_t end) {
    // Only the side effects of BvhNode::new that matter to the flattened scene: one axis draw per
    // node (also for spans 1 and 2), the stable sort, the median split.
    int axis = (int)main_rng().below_u64(3);
    size_t span = end - start;
    if (span == 0) throw std::runtime_error("BvhNode::new on an empty range");
    if (span == 1) return;
    std::stable_sort(objects.begin() + start, objects.begin() + end,
                     [axis](const Hit& a, const Hit& b) { return total_cmp_key(a->box.min[axis]) < total_cmp_key(b->box.min[axis]); });
    if (span == 2) return;
    size_t mid = start + span / 2;
    bvh_order(objects, start, mid);
    bvh_order(objects, mid, end);
}
Hit BvhNode(HitableList& objects, size_t start, size_t end) {
    bvh_order(objects, start, end);
    Hit h = make_hit(HKind::Bvh);
    h->kids.assign(objects.begin() + start, objects.begin() + end);
    Box3 b = h->kids[0]->box;
    for (size_t i = 1; i < h->kids.size(); i++) b = surround(b, h->kids[i]->box);
    h->box = b;
    return h;
}

// --------------------------------------------------------------------------------------- camera
static void put(float* d, V3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }
RtCamera CameraLens(V3 lookfrom, V3 lookat, V3 vup, float vfov, float aspect_ratio, float aperture, float focus_dist,
                    float time0, float time1) {
    float theta = vfov * (3.14159265358979323846f / 180.0f);  // f32::to_radians
    float viewport_height = std::tan(theta / 2.0f) * 2.0f;
    float viewport_width = viewport_height * aspect_ratio;
    V3 w = normalize(lookfrom - lookat);
    V3 u = normalize(cross(vup, w));
    V3 v = cross(w, u);
    V3 horizontal, vertical, llc;
    if (focus_dist == 1.0f) {  // camera.rs:31-33 verbatim
        horizontal = viewport_width * u;
        vertical = viewport_height * v;
        llc = lookfrom - horizontal / 2.0f - vertical / 2.0f - w;
    } else {
        horizontal = focus_dist * viewport_width * u;
        vertical = focus_dist * viewport_height * v;
        llc = lookfrom - horizontal / 2.0f - vertical / 2.0f - focus_dist * w;
    }
    RtCamera c{};
    put(c.origin, lookfrom);
    put(c.horizontal, horizontal);
    put(c.vertical, vertical);
    put(c.lower_left_corner, llc);
    put(c.u, u);
    put(c.v, v);
    c.lens_radius = aperture / 2.0f;
    c.time0 = time0;
    c.time1 = time1;
    return c;
}
RtCamera Camera(V3 lookfrom, V3 lookat, V3 vup, float vfov, float aspect_ratio) {
    return CameraLens(lookfrom, lookat, vup, vfov, aspect_ratio, 0.0f, 1.0f, 0.0f, 0.0f);
}

// ------------------------------------------------------------------------------------ flattener
namespace {
struct Flat {
    FlatSceneOwner& o;
    std::map<const MaterialNode*, uint32_t> mat_ids;
    std::map<const TextureNode*, uint32_t> tex_ids;
    std::map<const void*, uint32_t> img_ids;
    uint32_t next_object = 0, cur_object = 0;
    int depth = 0;

    uint32_t image(const TextureNode* t) {
        auto f = img_ids.find(t->img.get());
        if (f != img_ids.end()) return f->second;
        RtImage im{t->img_w, t->img_h, t->img->data()};
        o.image_store.push_back(t->img);
        o.images.push_back(im);
        return img_ids[t->img.get()] = (uint32_t)o.images.size() - 1;
    }
    uint32_t tex(const Tex& t) {
        if (!t) throw std::runtime_error("material without a texture");
        auto f = tex_ids.find(t.get());
        if (f != tex_ids.end()) return f->second;
        RtTexture r{};
        r.kind = t->kind;
        switch (t->kind) {
            case RT_TEX_CONSTANT: r.p[0] = t->col.x; r.p[1] = t->col.y; r.p[2] = t->col.z; break;
            case RT_TEX_CHECKER: r.a = tex(t->odd); r.b = tex(t->even); break;
            case RT_TEX_PERLIN:
                o.perlin.push_back(*t->perlin);
                r.a = (uint32_t)o.perlin.size() - 1;
                r.p[0] = t->scale;
                break;
            case RT_TEX_IMAGE: r.a = image(t.get()); break;
        }
        o.textures.push_back(r);
        return tex_ids[t.get()] = (uint32_t)o.textures.size() - 1;
    }
    uint32_t mat(const Mat& m) {
        auto f = mat_ids.find(m.get());
        if (f != mat_ids.end()) return f->second;
        RtMaterial r{};
        r.kind = m->kind;
        bool needs_tex0 = !(m->kind == RT_MAT_METAL || m->kind == RT_MAT_DIELECTRIC || m->kind == RT_MAT_DISNEY_CLEARCOAT);
        if (needs_tex0) r.tex0 = tex(m->tex0);
        if (m->kind == RT_MAT_ROUGH_PLASTIC) r.tex1 = tex(m->tex1);
        for (int i = 0; i < 4; i++) r.p[i] = m->p[i];
        o.materials.push_back(r);
        return mat_ids[m.get()] = (uint32_t)o.materials.size() - 1;
    }
    uint32_t chain(uint32_t parent, RtXformOp op) {
        RtXformChain pc = o.chains[parent];
        RtXformChain nc{(uint32_t)o.ops.size(), pc.n_ops + 1};
        for (uint32_t i = 0; i < pc.n_ops; i++) {
            RtXformOp c = o.ops[pc.first_op + i];
            o.ops.push_back(c);
        }
        o.ops.push_back(op);
        o.chains.push_back(nc);
        return (uint32_t)o.chains.size() - 1;
    }
    uint32_t prim(uint32_t kind, const Mat& m, uint32_t xf, bool bo, std::initializer_list<float> g) {
        RtPrim p{};
        p.kind = kind;
        p.mat = mat(m);
        p.xform = xf;
        p.flags = bo ? RT_PRIM_FLAG_BOUNDARY_ONLY : 0u;
        p.object = cur_object;
        int i = 0;
        for (float v : g) p.g[i++] = v;
        o.prims.push_back(p);
        return (uint32_t)o.prims.size() - 1;
    }
    void walk(const Hit& h, uint32_t xf, bool bo) {
        const bool aggregate = h->kind == HKind::Bvh || h->kind == HKind::List;
        if (!aggregate) {
            if (depth == 0) cur_object = next_object++;
            depth++;
        }
        switch (h->kind) {
            case HKind::Sphere: prim(RT_PRIM_SPHERE, h->mat, xf, bo, {h->a.x, h->a.y, h->a.z, h->r}); break;
            case HKind::MovingSphere:
                prim(RT_PRIM_MOVING_SPHERE, h->mat, xf, bo, {h->a.x, h->a.y, h->a.z, h->r, h->b.x, h->b.y, h->b.z, h->time0, h->time1});
                break;
            case HKind::XYRect: prim(RT_PRIM_XYRECT, h->mat, xf, bo, {h->a.x, h->a.y, h->a.z, h->b.x, h->b.y, h->b.z}); break;
            case HKind::XZRect: prim(RT_PRIM_XZRECT, h->mat, xf, bo, {h->a.x, h->a.y, h->a.z, h->b.x, h->b.y, h->b.z}); break;
            case HKind::YZRect: prim(RT_PRIM_YZRECT, h->mat, xf, bo, {h->a.x, h->a.y, h->a.z, h->b.x, h->b.y, h->b.z}); break;
            case HKind::GBox:
            case HKind::Bvh:
            case HKind::List:
                for (const Hit& k : h->kids) walk(k, xf, bo);
                break;
            case HKind::Translate: {
                RtXformOp op{RT_XF_TRANSLATE, {h->a.x, h->a.y, h->a.z}};
                walk(h->kids[0], chain(xf, op), bo);
                break;
            }
            case HKind::RotateY: {
                RtXformOp op{RT_XF_ROTATE_Y, {h->sin_theta, h->cos_theta, h->angle}};
                walk(h->kids[0], chain(xf, op), bo);
                break;
            }
            case HKind::Medium: {
                // The wrappers around the medium (chain xf) wrap its boundary too (hitable.rs:404-436,486-505 over
                // 536-579): the boundary primitives carry xf followed by their own wrappers, the medium carries xf.
                // A medium as another medium's boundary flattens too; rt_scene_create answers RT_ERR_UNSUPPORTED.
                uint32_t first = (uint32_t)o.prims.size();
                walk(h->kids[0], xf, true);
                uint32_t count = (uint32_t)o.prims.size() - first;
                uint32_t id = prim(RT_PRIM_MEDIUM, h->mat, xf, bo, {h->neg_inv_density});
                o.prims[id].aux0 = first;
                o.prims[id].aux1 = count;
                break;
            }
        }
        if (!aggregate) depth--;
    }
};
}  // namespace

std::unique_ptr<FlatSceneOwner> flatten(const HitableList& world, const RtCamera& cam, const SkySpec& sky) {
    auto own = std::make_unique<FlatSceneOwner>();
    own->chains.push_back(RtXformChain{0, 0});
    Flat f{*own};
    for (const Hit& h : world) f.walk(h, 0, false);
    RtFlatScene& fs = own->flat;
    fs.sky.kind = sky.kind;
    fs.sky.image = 0;
    fs.sky.color[0] = sky.color.x; fs.sky.color[1] = sky.color.y; fs.sky.color[2] = sky.color.z;
    if (sky.kind == RT_SKY_ENVMAP) {
        if (!sky.env || sky.env->kind != RT_TEX_IMAGE) throw std::runtime_error("envmap sky needs an ImageTex");
        fs.sky.image = f.image(sky.env.get());
    }
    fs.abi_version = RT_ABI_VERSION;
    fs.n_prims = (uint32_t)own->prims.size();
    fs.n_xform_ops = (uint32_t)own->ops.size();
    fs.n_xform_chains = (uint32_t)own->chains.size();
    fs.n_materials = (uint32_t)own->materials.size();
    fs.n_textures = (uint32_t)own->textures.size();
    fs.n_perlin = (uint32_t)own->perlin.size();
    fs.n_images = (uint32_t)own->images.size();
    fs.prims = own->prims.data();
    fs.xform_ops = own->ops.data();
    fs.xform_chains = own->chains.data();
    fs.materials = own->materials.data();
    fs.textures = own->textures.data();
    fs.perlin = own->perlin.data();
    fs.images = own->images.data();
    fs.camera = cam;
    return own;
}

}  // namespace rtb
