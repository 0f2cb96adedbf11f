// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).  PARITY UNPINNED.
//
// (1) Flattener: walks the oracle's object graph in the reference's visitation order and emits an
//     RtFlatScene (include/rtgpu.h) — the same neutral POD description the product's host API
//     emits, so tests can compare the two byte-for-byte.
// (2) Unflattener: rebuilds a virtual-dispatch object graph (a flat HitableList = the reference's
//     "list" closest-hit semantics, hitable.rs:117-132) from an RtFlatScene, so that oracle and GPU
//     consume IDENTICAL inputs.
// (3) Restatement of src/demo_scene.rs (the five shipped scenes, incl. RNG draw order) and the
//     BASELINE.json config scenes composed through the same API.
#pragma once
#include <cstdio>
#include <functional>
#include <map>
#include <stdexcept>

#include "../include/rtgpu.h"
#include "rt.hpp"

namespace orc {

// ------------------------------------------------------------------------------------- flattener
struct Flattener {
    std::vector<RtPrim> prims;
    std::vector<RtXformOp> ops;
    std::vector<RtXformChain> chains;
    std::vector<RtMaterial> materials;
    std::vector<RtTexture> textures;
    std::vector<RtPerlin> perlin;
    std::vector<RtImage> images;
    std::vector<std::shared_ptr<std::vector<uint8_t>>> image_store;
    std::map<const Material*, uint32_t> mat_ids;
    std::map<const Texture*, uint32_t> tex_ids;
    std::map<const void*, uint32_t> image_ids;
    uint32_t next_object = 0;
    int object_depth = 0;
    uint32_t cur_object = 0;

    Flattener() { chains.push_back(RtXformChain{0, 0}); }

    uint32_t push_chain(uint32_t parent, const RtXformOp& op) {
        RtXformChain pc = chains[parent];
        RtXformChain nc;
        nc.first_op = (uint32_t)ops.size();
        for (uint32_t i = 0; i < pc.n_ops; i++) {
            RtXformOp c = ops[pc.first_op + i];
            ops.push_back(c);
        }
        ops.push_back(op);
        nc.n_ops = pc.n_ops + 1;
        chains.push_back(nc);
        return (uint32_t)chains.size() - 1;
    }
    uint32_t image_id(const ImageTex* it) {
        auto f = image_ids.find(it->rgb8.get());
        if (f != image_ids.end()) return f->second;
        RtImage im;
        im.width = it->w;
        im.height = it->h;
        im.rgb8 = it->rgb8->data();
        image_store.push_back(it->rgb8);
        images.push_back(im);
        uint32_t id = (uint32_t)images.size() - 1;
        image_ids[it->rgb8.get()] = id;
        return id;
    }
    uint32_t tex_id(const TexPtr& t) {
        auto f = tex_ids.find(t.get());
        if (f != tex_ids.end()) return f->second;
        RtTexture rt{};
        if (auto c = dynamic_cast<const ConstantTex*>(t.get())) {
            rt.kind = RT_TEX_CONSTANT;
            rt.p[0] = c->col.x; rt.p[1] = c->col.y; rt.p[2] = c->col.z;
        } else if (auto c = dynamic_cast<const CheckerTex*>(t.get())) {
            rt.kind = RT_TEX_CHECKER;
            rt.a = tex_id(c->odd);
            rt.b = tex_id(c->even);
        } else if (auto c = dynamic_cast<const PerlinTex*>(t.get())) {
            rt.kind = RT_TEX_PERLIN;
            RtPerlin pt{};
            for (int i = 0; i < 256; i++) {
                pt.rand_vec[i][0] = c->perlin.rand_vec[i].x;
                pt.rand_vec[i][1] = c->perlin.rand_vec[i].y;
                pt.rand_vec[i][2] = c->perlin.rand_vec[i].z;
                pt.rand_vec[i][3] = 0.0f;
                pt.perm_x[i] = (uint8_t)c->perlin.perm_x[i];
                pt.perm_y[i] = (uint8_t)c->perlin.perm_y[i];
                pt.perm_z[i] = (uint8_t)c->perlin.perm_z[i];
            }
            perlin.push_back(pt);
            rt.a = (uint32_t)perlin.size() - 1;
            rt.p[0] = c->scale;
        } else if (auto c = dynamic_cast<const ImageTex*>(t.get())) {
            rt.kind = RT_TEX_IMAGE;
            rt.a = image_id(c);
        } else {
            throw std::runtime_error("unknown texture");
        }
        textures.push_back(rt);
        uint32_t id = (uint32_t)textures.size() - 1;
        tex_ids[t.get()] = id;
        return id;
    }
    uint32_t mat_id(const MatPtr& m) {
        auto f = mat_ids.find(m.get());
        if (f != mat_ids.end()) return f->second;
        RtMaterial rm{};
        if (auto c = dynamic_cast<const Emission*>(m.get())) { rm.kind = RT_MAT_EMISSION; rm.tex0 = tex_id(c->emit); }
        else if (auto c = dynamic_cast<const Diffuse*>(m.get())) { rm.kind = RT_MAT_DIFFUSE; rm.tex0 = tex_id(c->albedo); }
        else if (auto c = dynamic_cast<const Lambert*>(m.get())) { rm.kind = RT_MAT_LAMBERT; rm.tex0 = tex_id(c->albedo); }
        else if (auto c = dynamic_cast<const Metal*>(m.get())) {
            rm.kind = RT_MAT_METAL; rm.p[0] = c->albedo.x; rm.p[1] = c->albedo.y; rm.p[2] = c->albedo.z; rm.p[3] = c->fuzz;
        }
        else if (auto c = dynamic_cast<const Dielectric*>(m.get())) { rm.kind = RT_MAT_DIELECTRIC; rm.p[0] = c->ior; }
        else if (auto c = dynamic_cast<const Isotropic*>(m.get())) { rm.kind = RT_MAT_ISOTROPIC; rm.tex0 = tex_id(c->albedo); }
        else if (auto c = dynamic_cast<const OrenNayar*>(m.get())) { rm.kind = RT_MAT_OREN_NAYAR; rm.tex0 = tex_id(c->albedo); rm.p[0] = c->roughness; }
        else if (auto c = dynamic_cast<const BurleyDiffuse*>(m.get())) { rm.kind = RT_MAT_BURLEY_DIFFUSE; rm.tex0 = tex_id(c->albedo); rm.p[0] = c->roughness; }
        else if (auto c = dynamic_cast<const RoughPlastic*>(m.get())) {
            rm.kind = RT_MAT_ROUGH_PLASTIC; rm.tex0 = tex_id(c->spec_color); rm.tex1 = tex_id(c->diff_color);
            rm.p[0] = c->roughness; rm.p[1] = c->eta;
        }
        else if (auto c = dynamic_cast<const DisneyDiffuse*>(m.get())) {
            rm.kind = RT_MAT_DISNEY_DIFFUSE; rm.tex0 = tex_id(c->albedo); rm.p[0] = c->roughness; rm.p[1] = c->subsurface;
        }
        else if (auto c = dynamic_cast<const DisneyMetal*>(m.get())) {
            rm.kind = RT_MAT_DISNEY_METAL; rm.tex0 = tex_id(c->albedo); rm.p[0] = c->roughness; rm.p[1] = c->anisotropic; rm.p[2] = c->rot;
        }
        else if (auto c = dynamic_cast<const DisneySheen*>(m.get())) { rm.kind = RT_MAT_DISNEY_SHEEN; rm.tex0 = tex_id(c->albedo); rm.p[0] = c->tint; }
        else if (auto c = dynamic_cast<const DisneyClearcoat*>(m.get())) { rm.kind = RT_MAT_DISNEY_CLEARCOAT; rm.p[0] = c->clearcoat_gloss; }
        else throw std::runtime_error("unknown material");
        materials.push_back(rm);
        uint32_t id = (uint32_t)materials.size() - 1;
        mat_ids[m.get()] = id;
        return id;
    }
    // every non-aggregate hitable met at depth 0 is one "object"
    struct ObjScope {
        Flattener& f;
        explicit ObjScope(Flattener& f_) : f(f_) {
            if (f.object_depth == 0) f.cur_object = f.next_object++;
            f.object_depth++;
        }
        ~ObjScope() { f.object_depth--; }
    };
    uint32_t push_prim(uint32_t kind, const MatPtr& m, uint32_t xf, bool bo, const float* g, int ng) {
        RtPrim p{};
        p.kind = kind;
        p.mat = m ? mat_id(m) : 0;
        p.xform = xf;
        p.flags = bo ? RT_PRIM_FLAG_BOUNDARY_ONLY : 0;
        p.object = cur_object;
        for (int i = 0; i < ng; i++) p.g[i] = g[i];
        prims.push_back(p);
        return (uint32_t)prims.size() - 1;
    }
};

inline void Sphere::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    float g[4] = {c.x, c.y, c.z, r};
    id = (int)f.push_prim(RT_PRIM_SPHERE, mat, xf, bo, g, 4);
}
inline void MovingSphere::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    float g[9] = {c0.x, c0.y, c0.z, r, c1.x, c1.y, c1.z, time0, time1};
    id = (int)f.push_prim(RT_PRIM_MOVING_SPHERE, mat, xf, bo, g, 9);
}
inline void XYRect::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    float g[6] = {min.x, min.y, min.z, max.x, max.y, max.z};
    id = (int)f.push_prim(RT_PRIM_XYRECT, mat, xf, bo, g, 6);
}
inline void XZRect::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    float g[6] = {min.x, min.y, min.z, max.x, max.y, max.z};
    id = (int)f.push_prim(RT_PRIM_XZRECT, mat, xf, bo, g, 6);
}
inline void YZRect::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    float g[6] = {min.x, min.y, min.z, max.x, max.y, max.z};
    id = (int)f.push_prim(RT_PRIM_YZRECT, mat, xf, bo, g, 6);
}
inline void HitableList::flatten(Flattener& f, uint32_t xf, bool bo) const {
    for (const auto& it : items) it->flatten(f, xf, bo);
}
inline void BvhNode::flatten(Flattener& f, uint32_t xf, bool bo) const {
    // left-to-right == the order BvhNode::hit visits; a span-1 node holds the same object twice
    // (hitable.rs:188) — emit it once.
    left->flatten(f, xf, bo);
    if (right.get() != left.get()) right->flatten(f, xf, bo);
}
inline void GBox::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    sides.flatten(f, xf, bo);
}
inline void Translate::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    RtXformOp op{};
    op.kind = RT_XF_TRANSLATE;
    op.v[0] = offset.x; op.v[1] = offset.y; op.v[2] = offset.z;
    ptr->flatten(f, f.push_chain(xf, op), bo);
}
inline void RotateY::flatten(Flattener& f, uint32_t xf, bool bo) const {
    Flattener::ObjScope s(f);
    RtXformOp op{};
    op.kind = RT_XF_ROTATE_Y;
    op.v[0] = sin_theta; op.v[1] = cos_theta; op.v[2] = angle;
    ptr->flatten(f, f.push_chain(xf, op), bo);
}
inline void ConstantMedium::flatten(Flattener& f, uint32_t xf, bool bo) const {
    // The wrappers around the medium (chain xf) wrap its boundary too: the boundary primitives carry xf followed
    // by their own wrappers, the medium record carries xf (hitable.rs:404-436,486-505 over 536-579).
    Flattener::ObjScope s(f);
    uint32_t first = (uint32_t)f.prims.size();
    boundary->flatten(f, xf, true);
    uint32_t count = (uint32_t)f.prims.size() - first;
    float g[1] = {neg_inv_density};
    id = (int)f.push_prim(RT_PRIM_MEDIUM, phase_fn, xf, bo, g, 1);
    f.prims[id].aux0 = first;
    f.prims[id].aux1 = count;
}

// A flattened scene that owns its arrays.
struct FlatHolder {
    Flattener f;
    RtFlatScene flat{};
    void finish(const Scene& sc) {
        flat.abi_version = RT_ABI_VERSION;
        flat.n_prims = (uint32_t)f.prims.size();
        flat.n_xform_ops = (uint32_t)f.ops.size();
        flat.n_xform_chains = (uint32_t)f.chains.size();
        flat.n_materials = (uint32_t)f.materials.size();
        flat.n_textures = (uint32_t)f.textures.size();
        flat.n_perlin = (uint32_t)f.perlin.size();
        flat.n_images = (uint32_t)f.images.size();
        flat.prims = f.prims.data();
        flat.xform_ops = f.ops.data();
        flat.xform_chains = f.chains.data();
        flat.materials = f.materials.data();
        flat.textures = f.textures.data();
        flat.perlin = f.perlin.data();
        flat.images = f.images.data();
        flat.sky.kind = (uint32_t)sc.sky.kind;
        flat.sky.image = 0;
        flat.sky.color[0] = sc.sky.color.x; flat.sky.color[1] = sc.sky.color.y; flat.sky.color[2] = sc.sky.color.z;
        if (sc.sky.kind == SkyKind::EnvMap) {
            flat.sky.image = f.image_id(sc.sky.env.get());
            flat.n_images = (uint32_t)f.images.size();
            flat.images = f.images.data();
        }
        const Camera& c = sc.cam;
        auto put = [](float* d, Vec3A v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; };
        put(flat.camera.origin, c.origin);
        put(flat.camera.horizontal, c.horizontal);
        put(flat.camera.vertical, c.vertical);
        put(flat.camera.lower_left_corner, c.lower_left_corner);
        put(flat.camera.u, c.u);
        put(flat.camera.v, c.v);
        flat.camera.lens_radius = c.lens_radius;
        flat.camera.time0 = c.time0;
        flat.camera.time1 = c.time1;
    }
};

inline std::unique_ptr<FlatHolder> flatten_scene(const Scene& sc) {
    auto h = std::make_unique<FlatHolder>();
    sc.world.flatten(h->f, 0, false);
    h->finish(sc);
    return h;
}

// ----------------------------------------------------------------------------------- unflattener
inline Vec3A v3(const float* p) { return Vec3A(p[0], p[1], p[2]); }

inline std::unique_ptr<Scene> scene_from_flat(const RtFlatScene& fs) {
    auto sc = std::make_unique<Scene>();
    std::vector<std::shared_ptr<ImageTex>> images;
    for (uint32_t i = 0; i < fs.n_images; i++)
        images.push_back(std::make_shared<ImageTex>(fs.images[i].width, fs.images[i].height, fs.images[i].rgb8));
    std::vector<TexPtr> tex(fs.n_textures);
    // textures may reference later/earlier textures (checker): resolve recursively
    std::function<TexPtr(uint32_t)> get_tex = [&](uint32_t i) -> TexPtr {
        if (tex[i]) return tex[i];
        const RtTexture& t = fs.textures[i];
        switch (t.kind) {
            case RT_TEX_CONSTANT: tex[i] = std::make_shared<ConstantTex>(v3(t.p)); break;
            case RT_TEX_CHECKER: tex[i] = std::make_shared<CheckerTex>(get_tex(t.a), get_tex(t.b)); break;
            case RT_TEX_PERLIN: {
                auto pt = std::make_shared<PerlinTex>(t.p[0], Perlin::NoInit{});
                const RtPerlin& rp = fs.perlin[t.a];
                pt->perlin.rand_vec.resize(256);
                pt->perlin.perm_x.resize(256); pt->perlin.perm_y.resize(256); pt->perlin.perm_z.resize(256);
                for (int k = 0; k < 256; k++) {
                    pt->perlin.rand_vec[k] = v3(rp.rand_vec[k]);
                    pt->perlin.perm_x[k] = rp.perm_x[k];
                    pt->perlin.perm_y[k] = rp.perm_y[k];
                    pt->perlin.perm_z[k] = rp.perm_z[k];
                }
                tex[i] = pt;
                break;
            }
            case RT_TEX_IMAGE: tex[i] = images.at(t.a); break;
            default: throw std::runtime_error("bad texture kind");
        }
        tex[i]->flat_id = (int)i;
        return tex[i];
    };
    std::vector<MatPtr> mats(fs.n_materials);
    for (uint32_t i = 0; i < fs.n_materials; i++) {
        const RtMaterial& m = fs.materials[i];
        switch (m.kind) {
            case RT_MAT_EMISSION: mats[i] = std::make_shared<Emission>(get_tex(m.tex0)); break;
            case RT_MAT_DIFFUSE: mats[i] = std::make_shared<Diffuse>(get_tex(m.tex0)); break;
            case RT_MAT_LAMBERT: mats[i] = std::make_shared<Lambert>(get_tex(m.tex0)); break;
            case RT_MAT_METAL: mats[i] = std::make_shared<Metal>(v3(m.p), m.p[3]); break;
            case RT_MAT_DIELECTRIC: mats[i] = std::make_shared<Dielectric>(m.p[0]); break;
            case RT_MAT_ISOTROPIC: mats[i] = std::make_shared<Isotropic>(get_tex(m.tex0)); break;
            case RT_MAT_OREN_NAYAR: mats[i] = std::make_shared<OrenNayar>(get_tex(m.tex0), m.p[0]); break;
            case RT_MAT_BURLEY_DIFFUSE: mats[i] = std::make_shared<BurleyDiffuse>(get_tex(m.tex0), m.p[0]); break;
            case RT_MAT_ROUGH_PLASTIC: mats[i] = std::make_shared<RoughPlastic>(get_tex(m.tex0), get_tex(m.tex1), m.p[0], m.p[1]); break;
            case RT_MAT_DISNEY_DIFFUSE: mats[i] = std::make_shared<DisneyDiffuse>(get_tex(m.tex0), m.p[0], m.p[1]); break;
            case RT_MAT_DISNEY_METAL: mats[i] = std::make_shared<DisneyMetal>(get_tex(m.tex0), m.p[0], m.p[1], m.p[2]); break;
            case RT_MAT_DISNEY_SHEEN: mats[i] = std::make_shared<DisneySheen>(get_tex(m.tex0), m.p[0]); break;
            case RT_MAT_DISNEY_CLEARCOAT: mats[i] = std::make_shared<DisneyClearcoat>(m.p[0]); break;
            default: throw std::runtime_error("bad material kind");
        }
        mats[i]->flat_id = (int)i;
    }
    for (uint32_t i = 0; i < fs.n_textures; i++) get_tex(i);
    sc->mats = mats;
    sc->texs = tex;
    // wraps `h` in the wrappers ops[first_op + skip .. first_op + n_ops) of chain `xf`, innermost first
    auto wrap = [&](HitPtr h, uint32_t xf, uint32_t skip) -> HitPtr {
        const RtXformChain& ch = fs.xform_chains[xf];
        for (int k = (int)ch.n_ops - 1; k >= (int)skip; k--) {
            const RtXformOp& op = fs.xform_ops[ch.first_op + k];
            if (op.kind == RT_XF_TRANSLATE) {
                h = std::make_shared<Translate>(v3(op.v), h);
            } else {
                auto r = std::make_shared<RotateY>(h, op.v[2]);
                r->sin_theta = op.v[0];  // identical inputs: take the flattened sin/cos verbatim
                r->cos_theta = op.v[1];
                h = r;
            }
        }
        return h;
    };
    auto make_geom = [&](uint32_t i, uint32_t skip) -> HitPtr {
        const RtPrim& p = fs.prims[i];
        HitPtr h;
        switch (p.kind) {
            case RT_PRIM_SPHERE: { auto s = std::make_shared<Sphere>(v3(p.g), p.g[3], mats.at(p.mat)); s->id = (int)i; h = s; break; }
            case RT_PRIM_MOVING_SPHERE: { auto s = std::make_shared<MovingSphere>(v3(p.g), v3(p.g + 4), p.g[7], p.g[8], p.g[3], mats.at(p.mat)); s->id = (int)i; h = s; break; }
            case RT_PRIM_XYRECT: { auto s = std::make_shared<XYRect>(v3(p.g), v3(p.g + 3), mats.at(p.mat)); s->id = (int)i; h = s; break; }
            case RT_PRIM_XZRECT: { auto s = std::make_shared<XZRect>(v3(p.g), v3(p.g + 3), mats.at(p.mat)); s->id = (int)i; h = s; break; }
            case RT_PRIM_YZRECT: { auto s = std::make_shared<YZRect>(v3(p.g), v3(p.g + 3), mats.at(p.mat)); s->id = (int)i; h = s; break; }
            default: throw std::runtime_error("bad prim kind");
        }
        return wrap(h, p.xform, skip);
    };
    for (uint32_t i = 0; i < fs.n_prims; i++) {
        const RtPrim& p = fs.prims[i];
        if (p.flags & RT_PRIM_FLAG_BOUNDARY_ONLY) continue;
        if (p.kind == RT_PRIM_MEDIUM) {
            // boundary chains start with the medium's own chain: those ops wrap the medium, the rest the boundary
            const uint32_t n_outer = fs.xform_chains[p.xform].n_ops;
            auto bl = std::make_shared<HitableList>();
            for (uint32_t k = 0; k < p.aux1; k++) {
                if (fs.prims[p.aux0 + k].kind == RT_PRIM_MEDIUM) throw std::runtime_error("nested medium");
                bl->items.push_back(make_geom(p.aux0 + k, n_outer));
            }
            auto iso = std::dynamic_pointer_cast<Isotropic>(mats.at(p.mat));
            if (!iso) throw std::runtime_error("medium material must be Isotropic");
            auto cm = std::make_shared<ConstantMedium>(bl, 1.0f, iso->albedo);
            cm->phase_fn = iso;
            cm->neg_inv_density = p.g[0];
            cm->id = (int)i;
            sc->world.items.push_back(wrap(cm, p.xform, 0));
        } else {
            sc->world.items.push_back(make_geom(i, 0));
        }
    }
    sc->sky.kind = (SkyKind)fs.sky.kind;
    sc->sky.color = v3(fs.sky.color);
    if (sc->sky.kind == SkyKind::EnvMap) sc->sky.env = images.at(fs.sky.image);
    const RtCamera& c = fs.camera;
    sc->cam.origin = v3(c.origin);
    sc->cam.horizontal = v3(c.horizontal);
    sc->cam.vertical = v3(c.vertical);
    sc->cam.lower_left_corner = v3(c.lower_left_corner);
    sc->cam.u = v3(c.u);
    sc->cam.v = v3(c.v);
    sc->cam.lens_radius = c.lens_radius;
    sc->cam.time0 = c.time0;
    sc->cam.time1 = c.time1;
    return sc;
}

// ------------------------------------------------------------------------------ asset loading (PPM)
inline std::shared_ptr<ImageTex> load_ppm(const std::string& path) {
    FILE* fh = std::fopen(path.c_str(), "rb");
    if (!fh) throw std::runtime_error("cannot open " + path);
    char magic[3] = {0};
    int w = 0, h = 0, mx = 0;
    if (std::fscanf(fh, "%2s %d %d %d", magic, &w, &h, &mx) != 4 || std::string(magic) != "P6" || mx != 255) {
        std::fclose(fh);
        throw std::runtime_error("bad ppm " + path);
    }
    std::fgetc(fh);
    std::vector<uint8_t> buf((size_t)w * h * 3);
    size_t got = std::fread(buf.data(), 1, buf.size(), fh);
    std::fclose(fh);
    if (got != buf.size()) throw std::runtime_error("short ppm " + path);
    return std::make_shared<ImageTex>((uint32_t)w, (uint32_t)h, buf.data());
}

// -------------------------------------------------------------------------- demo_scene.rs restated
inline HitableList build_bvh(std::vector<HitPtr>& world) {  // demo_scene.rs:223-227
    HitPtr bvh = std::make_shared<BvhNode>(world, 0, world.size());
    return HitableList({bvh});
}
template <class T, class... A>
std::shared_ptr<T> mk(A&&... a) { return std::make_shared<T>(std::forward<A>(a)...); }
inline TexPtr ctex(float r, float g, float b) { return mk<ConstantTex>(Vec3A(r, g, b)); }

// variant: 0 = sphere_scene verbatim (demo_scene.rs:37-86); 1 = cfg2 "textured" (checker ground as in the
// commented line demo_scene.rs:40, Diffuse(ImageTex earth), Diffuse(PerlinTex)); 2 = defocus camera extension
inline std::unique_ptr<Scene> sphere_scene(float aspect_ratio, const std::string& assets, int variant) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Gradient;
    TexPtr checker;
    if (variant == 1) checker = mk<CheckerTex>(Vec3A(0.2f, 0.3f, 0.1f), Vec3A(0.9f, 0.9f, 0.9f));
    TexPtr perlin = mk<PerlinTex>(4.0f);
    TexPtr earth_map = load_ppm(assets + "/earthmap.ppm");
    MatPtr material_ground = mk<Diffuse>(variant == 1 ? checker : perlin);
    MatPtr material_1 = variant == 1 ? MatPtr(mk<Diffuse>(earth_map)) : MatPtr(mk<Emission>(earth_map));
    MatPtr material_2 = mk<Dielectric>(1.5f);
    MatPtr material_3 = mk<Metal>(Vec3A(0.8f, 0.6f, 0.2f), 0.0f);
    std::vector<HitPtr> world = {
        mk<Sphere>(Vec3A(1.0f, -1000.0f, -1.0f), 1000.0f, material_ground, "Ground"),
        mk<Sphere>(Vec3A(0.0f, 1.0f, 3.0f), 1.0f, material_1, "Sphere_1"),
        mk<Sphere>(Vec3A(-4.0f, 1.0f, 0.0f), 1.0f, variant == 1 ? MatPtr(mk<Diffuse>(perlin)) : material_2, "Sphere_2"),
        mk<Sphere>(Vec3A(4.0f, 1.0f, 0.0f), 1.0f, material_3, "Sphere_3"),
    };
    SmallRng rng = SmallRng::seed_from_u64(95);
    for (int a = -11; a <= 11; a++)
        for (int b = -11; b <= 11; b++) {
            float choose_mat = rng.gen_f32();
            float cx = (float)a + 0.9f * rng.gen_f32();
            float cz = (float)b + 0.9f * rng.gen_f32();
            Vec3A center(cx, 0.2f, cz);
            MatPtr mat;
            if (choose_mat < 0.8f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                float b0 = rng.gen_f32(), b1 = rng.gen_f32(), b2 = rng.gen_f32();
                Vec3A col = Vec3A(a0, a1, a2) * Vec3A(b0, b1, b2);
                mat = mk<Diffuse>(mk<ConstantTex>(col));
            } else if (choose_mat < 0.95f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                Vec3A albedo = Vec3A(a0, a1, a2) * 0.5f + 0.5f;
                float fuzz = rng.gen_f32();
                mat = mk<Metal>(albedo, fuzz);
            } else {
                mat = material_2;
            }
            world.push_back(mk<Sphere>(center, 0.2f, mat, "Sphere"));
        }
    if (variant == 2)
        sc->cam = Camera(Vec3A(13, 2, 3), Vec3A(0, 0, 0), Vec3A(0, 1, 0), 20.0f, aspect_ratio, 0.1f, 10.0f, 0.0f, 0.0f);
    else
        sc->cam = Camera(Vec3A(13, 2, 3), Vec3A(0, 0, 0), Vec3A(0, 1, 0), 20.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

// cfg5 EXTENSION: the cover generator on a 316x316 grid in the same footprint,
// Diffuse spheres become MovingSphere, shutter [0,1).
inline std::unique_ptr<Scene> cover_100k_motion(float aspect_ratio, const std::string& assets, int grid = 316) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Gradient;
    TexPtr perlin = mk<PerlinTex>(4.0f);
    TexPtr earth_map = load_ppm(assets + "/earthmap.ppm");
    MatPtr material_ground = mk<Diffuse>(perlin);
    MatPtr material_1 = mk<Emission>(earth_map);
    MatPtr material_2 = mk<Dielectric>(1.5f);
    MatPtr material_3 = mk<Metal>(Vec3A(0.8f, 0.6f, 0.2f), 0.0f);
    std::vector<HitPtr> world = {
        mk<Sphere>(Vec3A(1.0f, -1000.0f, -1.0f), 1000.0f, material_ground, "Ground"),
        mk<Sphere>(Vec3A(0.0f, 1.0f, 3.0f), 1.0f, material_1, "Sphere_1"),
        mk<Sphere>(Vec3A(-4.0f, 1.0f, 0.0f), 1.0f, material_2, "Sphere_2"),
        mk<Sphere>(Vec3A(4.0f, 1.0f, 0.0f), 1.0f, material_3, "Sphere_3"),
    };
    SmallRng rng = SmallRng::seed_from_u64(95);
    const float cell = 23.0f / (float)grid;
    const float rad = 0.2f * cell;
    for (int ia = 0; ia < grid; ia++)
        for (int ib = 0; ib < grid; ib++) {
            float a = -11.0f + (float)ia * cell;
            float b = -11.0f + (float)ib * cell;
            float choose_mat = rng.gen_f32();
            float cx = a + 0.9f * cell * rng.gen_f32();
            float cz = b + 0.9f * cell * rng.gen_f32();
            Vec3A center(cx, rad, cz);
            if (choose_mat < 0.8f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                float b0 = rng.gen_f32(), b1 = rng.gen_f32(), b2 = rng.gen_f32();
                Vec3A col = Vec3A(a0, a1, a2) * Vec3A(b0, b1, b2);
                MatPtr mat = mk<Diffuse>(mk<ConstantTex>(col));
                Vec3A c1 = center + Vec3A(0.0f, 0.5f * rng.gen_f32() * cell, 0.0f);
                world.push_back(mk<MovingSphere>(center, c1, 0.0f, 1.0f, rad, mat));
            } else if (choose_mat < 0.95f) {
                float a0 = rng.gen_f32(), a1 = rng.gen_f32(), a2 = rng.gen_f32();
                Vec3A albedo = Vec3A(a0, a1, a2) * 0.5f + 0.5f;
                float fuzz = rng.gen_f32();
                world.push_back(mk<Sphere>(center, rad, mk<Metal>(albedo, fuzz), "Sphere"));
            } else {
                world.push_back(mk<Sphere>(center, rad, material_2, "Sphere"));
            }
        }
    sc->cam = Camera(Vec3A(13, 2, 3), Vec3A(0, 0, 0), Vec3A(0, 1, 0), 20.0f, aspect_ratio, 0.0f, 1.0f, 0.0f, 1.0f);
    sc->world = build_bvh(world);
    return sc;
}

inline std::unique_ptr<Scene> simple_light_scene(float aspect_ratio) {  // demo_scene.rs:88-110
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Black;
    TexPtr perlin = mk<PerlinTex>(4.0f);
    MatPtr mat_perlin = mk<Diffuse>(perlin);
    MatPtr material_1 = mk<Emission>(ctex(4, 4, 4));
    std::vector<HitPtr> world = {
        mk<Sphere>(Vec3A(1.0f, -1000.0f, -1.0f), 1000.0f, mat_perlin, "Ground"),
        mk<Sphere>(Vec3A(0.0f, 2.0f, 0.0f), 2.0f, mat_perlin, "Sphere_1"),
        mk<Sphere>(Vec3A(0.0f, 6.5f, 0.0f), 2.0f, material_1, "Sphere_2"),
        mk<XYRect>(Vec3A(3, 1, -2), Vec3A(5, 3, -2), material_1),
    };
    sc->cam = Camera(Vec3A(26, 3, 6), Vec3A(0, 0, 0), Vec3A(0, 1, 0), 20.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

// solid == false: verbatim smoke variant (demo_scene.rs:112-148); solid == true: cfg3 (ii), the
// rotated/translated boxes themselves are pushed instead of the media.
inline std::unique_ptr<Scene> cornell_box(float aspect_ratio, bool solid) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Black;
    MatPtr red = mk<Diffuse>(ctex(0.65f, 0.05f, 0.05f));
    MatPtr white = mk<Diffuse>(ctex(0.73f, 0.73f, 0.73f));
    MatPtr green = mk<Diffuse>(ctex(0.12f, 0.45f, 0.15f));
    MatPtr light = mk<Emission>(ctex(7, 7, 7));
    HitPtr box_1 = mk<GBox>(Vec3A(), Vec3A(165, 330, 165), white);
    box_1 = mk<RotateY>(box_1, 15.0f);
    box_1 = mk<Translate>(Vec3A(265, 0, 295), box_1);
    HitPtr mediun_1 = mk<ConstantMedium>(box_1, 0.01f, ctex(0, 0, 0));
    HitPtr box_2 = mk<GBox>(Vec3A(), Vec3A(165, 165, 165), white);
    box_2 = mk<RotateY>(box_2, -18.0f);
    box_2 = mk<Translate>(Vec3A(130, 0, 65), box_2);
    HitPtr mediun_2 = mk<ConstantMedium>(box_2, 0.01f, ctex(1, 1, 1));
    std::vector<HitPtr> world = {
        mk<XZRect>(Vec3A(113, 554, 127), Vec3A(443, 554, 432), light),
        mk<XYRect>(Vec3A(0, 0, 555), Vec3A(555, 555, 555), white),
        mk<XZRect>(Vec3A(0, 0, 0), Vec3A(555, 0, 555), white),
        mk<XZRect>(Vec3A(0, 555, 0), Vec3A(555, 555, 555), white),
        mk<YZRect>(Vec3A(0, 0, 0), Vec3A(0, 555, 555), red),
        mk<YZRect>(Vec3A(555, 0, 0), Vec3A(555, 555, 555), green),
        solid ? box_1 : mediun_1,
        solid ? box_2 : mediun_2,
    };
    sc->cam = Camera(Vec3A(278, 278, -800), Vec3A(278, 278, 0), Vec3A(0, 1, 0), 40.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

inline std::unique_ptr<Scene> final_scene(float aspect_ratio, const std::string& assets) {  // demo_scene.rs:150-221
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Black;
    MatPtr ground = mk<Diffuse>(ctex(0.48f, 0.83f, 0.53f));
    MatPtr white = mk<Diffuse>(ctex(0.73f, 0.73f, 0.73f));
    MatPtr brown = mk<BurleyDiffuse>(ctex(0.7f, 0.3f, 0.1f), 0.9f);
    MatPtr light = mk<Emission>(ctex(7, 7, 7));
    MatPtr dielectric = mk<Dielectric>(1.5f);
    MatPtr metal = mk<Metal>(Vec3A(0.8f, 0.8f, 0.9f), 1.0f);
    TexPtr earth_map = load_ppm(assets + "/earthmap.ppm");
    MatPtr earth_mat = mk<Diffuse>(earth_map);
    HitPtr earth = mk<Sphere>(Vec3A(400, 200, 400), 100.0f, earth_mat, "EarthSphere");
    TexPtr perlin = mk<PerlinTex>(0.1f);
    MatPtr mat_perlin = mk<Diffuse>(perlin);
    HitPtr perlin_sphere = mk<Sphere>(Vec3A(220, 280, 300), 80.0f, mat_perlin, "PerlinSphere");
    HitPtr brown_sphere = mk<Sphere>(Vec3A(400, 400, 200), 50.0f, brown, "BrownSphere");
    HitPtr dielectric_sphere = mk<Sphere>(Vec3A(260, 150, 45), 50.0f, dielectric, "DielectricSphere");
    HitPtr metal_sphere = mk<Sphere>(Vec3A(0, 150, 145), 50.0f, metal, "MetalSphere");
    HitPtr boundary = mk<Sphere>(Vec3A(360, 150, 145), 50.0f, dielectric, "boundary");
    HitPtr medium = mk<ConstantMedium>(boundary, 0.2f, ctex(0.2f, 0.4f, 0.9f));
    std::vector<HitPtr> spheres;
    for (int i = 0; i < 1000; i++) spheres.push_back(mk<Sphere>(vec3a_random_range(0.0f, 165.0f), 10.0f, white, "Ground"));
    HitPtr sph = mk<BvhNode>(spheres, 0, 1000);
    sph = mk<RotateY>(sph, 15.0f);
    sph = mk<Translate>(Vec3A(-100, 270, 395), sph);
    std::vector<HitPtr> boxes;
    for (int i = 0; i < 20; i++)
        for (int j = 0; j < 20; j++) {
            float w = 100.0f;
            float x0 = -1000.0f + (float)i * w;
            float z0 = -1000.0f + (float)j * w;
            float y0 = 0.0f;
            float x1 = x0 + w;
            float y1 = RNG().gen_f32() * 100.0f + 1.0f;
            float z1 = z0 + w;
            boxes.push_back(mk<GBox>(Vec3A(x0, y0, z0), Vec3A(x1, y1, z1), ground));
        }
    HitPtr bx = mk<BvhNode>(boxes, 0, 400);
    std::vector<HitPtr> world = {
        mk<XZRect>(Vec3A(123, 544, 147), Vec3A(423, 554, 412), light),
        sph, earth, perlin_sphere, brown_sphere, dielectric_sphere, metal_sphere, bx, boundary, medium,
    };
    sc->cam = Camera(Vec3A(478, 278, -600), Vec3A(278, 278, 0), Vec3A(0, 1, 0), 40.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

inline std::unique_ptr<Scene> test_sphere(float aspect_ratio) {  // demo_scene.rs:229-244
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Gradient;
    MatPtr ground = mk<Lambert>(ctex(0.5f, 0.5f, 0.5f));
    sc->world.items = {
        mk<Sphere>(Vec3A(0.0f, -100.5f, -1.0f), 100.0f, ground, "Ground"),
        mk<Sphere>(Vec3A(0.0f, 0.0f, -1.0f), 0.5f, ground, "Test"),
    };
    sc->cam = Camera(Vec3A(0, 0, 0), Vec3A(0, 0, -1), Vec3A(0, 1, 0), 90.0f, aspect_ratio);
    return sc;
}

// cfg4 `pbr_ibl` (no reference scene exists, composed through the same API):
// sky = tex_sky_color over newport_loft, RoughPlastic ground, eight pbr.rs spheres in a row.
inline std::unique_ptr<Scene> pbr_ibl(float aspect_ratio, const std::string& assets) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::EnvMap;
    sc->sky.env = load_ppm(assets + "/newport_loft.ppm");
    MatPtr ground = mk<RoughPlastic>(ctex(1, 1, 1), ctex(0.5f, 0.5f, 0.5f), 0.3f, 1.5f);
    std::vector<MatPtr> m = {
        mk<OrenNayar>(ctex(0.8f, 0.3f, 0.3f), 0.5f),
        mk<BurleyDiffuse>(ctex(0.3f, 0.8f, 0.3f), 0.6f),
        mk<RoughPlastic>(ctex(1, 1, 1), ctex(0.2f, 0.3f, 0.8f), 0.2f, 1.5f),
        mk<DisneyDiffuse>(ctex(0.8f, 0.7f, 0.2f), 0.5f, 0.5f),
        mk<DisneyMetal>(ctex(0.95f, 0.64f, 0.54f), 0.3f, 0.0f, 0.0f),
        mk<DisneyMetal>(ctex(0.9f, 0.9f, 0.9f), 0.4f, 0.8f, 0.25f),
        mk<DisneySheen>(ctex(0.6f, 0.2f, 0.7f), 0.5f),
        mk<DisneyClearcoat>(0.8f),
    };
    std::vector<HitPtr> world = {mk<Sphere>(Vec3A(0.0f, -100.5f, -1.0f), 100.0f, ground, "Ground")};
    for (int k = 0; k < 8; k++) world.push_back(mk<Sphere>(Vec3A(((float)k - 3.5f) * 1.2f, 0.0f, -1.0f), 0.5f, m[k], "pbr"));
    sc->cam = Camera(Vec3A(0, 1.5f, 7), Vec3A(0, 0, -1), Vec3A(0, 1, 0), 40.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

// K10 furnace: one convex sphere, constant white sky.
inline std::unique_ptr<Scene> furnace(float aspect_ratio, const std::string& which) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Constant;
    sc->sky.color = Vec3A(1, 1, 1);
    MatPtr m;
    TexPtr a = ctex(0.6f, 0.5f, 0.4f);
    if (which == "diffuse") m = mk<Diffuse>(a);
    else if (which == "lambert") m = mk<Lambert>(a);
    else if (which == "metal") m = mk<Metal>(Vec3A(0.6f, 0.5f, 0.4f), 0.0f);
    else if (which == "dielectric") m = mk<Dielectric>(1.5f);
    else if (which == "oren_nayar") m = mk<OrenNayar>(a, 0.5f);
    else if (which == "burley") m = mk<BurleyDiffuse>(a, 0.5f);
    else if (which == "rough_plastic") m = mk<RoughPlastic>(ctex(1, 1, 1), a, 0.3f, 1.5f);
    else if (which == "disney_diffuse") m = mk<DisneyDiffuse>(a, 0.5f, 0.5f);
    else if (which == "disney_metal") m = mk<DisneyMetal>(a, 0.4f, 0.5f, 0.1f);
    else if (which == "disney_sheen") m = mk<DisneySheen>(a, 0.5f);
    else if (which == "disney_clearcoat") m = mk<DisneyClearcoat>(0.5f);
    else throw std::runtime_error("unknown furnace material " + which);
    sc->world.items = {mk<Sphere>(Vec3A(0.0f, 0.0f, -1.0f), 0.5f, m, "Furnace")};
    sc->cam = Camera(Vec3A(0, 0, 0), Vec3A(0, 0, -1), Vec3A(0, 1, 0), 20.0f, aspect_ratio);
    return sc;
}

// EXTENSION (see csrc/demo_scene.cpp smoke_instances): media under Translate / RotateY wrappers
inline std::unique_ptr<Scene> smoke_instances(float aspect_ratio) {
    auto sc = std::make_unique<Scene>();
    sc->sky.kind = SkyKind::Black;
    MatPtr red = mk<Diffuse>(ctex(0.65f, 0.05f, 0.05f));
    MatPtr white = mk<Diffuse>(ctex(0.73f, 0.73f, 0.73f));
    MatPtr green = mk<Diffuse>(ctex(0.12f, 0.45f, 0.15f));
    MatPtr light = mk<Emission>(ctex(7, 7, 7));
    HitPtr ball = mk<Translate>(Vec3A(400, 120, 200),
                                mk<ConstantMedium>(mk<Sphere>(Vec3A(0, 0, 0), 100.0f, white), 0.02f, ctex(0.9f, 0.9f, 0.9f)));
    HitPtr tall = mk<Translate>(Vec3A(265, 0, 295),
                                mk<RotateY>(mk<ConstantMedium>(mk<GBox>(Vec3A(), Vec3A(165, 330, 165), white), 0.01f, ctex(0, 0, 0)), 15.0f));
    HitPtr low = mk<RotateY>(mk<ConstantMedium>(mk<Translate>(Vec3A(60, 0, 130), mk<GBox>(Vec3A(), Vec3A(150, 150, 150), white)), 0.015f,
                                                ctex(0.2f, 0.4f, 0.9f)), -10.0f);
    std::vector<HitPtr> world = {
        mk<XZRect>(Vec3A(113, 554, 127), Vec3A(443, 554, 432), light),
        mk<XYRect>(Vec3A(0, 0, 555), Vec3A(555, 555, 555), white),
        mk<XZRect>(Vec3A(0, 0, 0), Vec3A(555, 0, 555), white),
        mk<XZRect>(Vec3A(0, 555, 0), Vec3A(555, 555, 555), white),
        mk<YZRect>(Vec3A(0, 0, 0), Vec3A(0, 555, 555), red),
        mk<YZRect>(Vec3A(555, 0, 0), Vec3A(555, 555, 555), green),
        ball, tall, low,
    };
    sc->cam = Camera(Vec3A(278, 278, -800), Vec3A(278, 278, 0), Vec3A(0, 1, 0), 40.0f, aspect_ratio);
    sc->world = build_bvh(world);
    return sc;
}

inline std::unique_ptr<Scene> build_named_scene(const std::string& name, float aspect, const std::string& assets) {
    RNG() = SmallRng::seed_from_u64(1995);  // lib.rs:8 — the "main thread" generator of a fresh process
    if (name == "test_sphere") return test_sphere(aspect);
    if (name == "sphere_scene") return sphere_scene(aspect, assets, 0);
    if (name == "textured") return sphere_scene(aspect, assets, 1);
    if (name == "cover_defocus") return sphere_scene(aspect, assets, 2);
    if (name == "cover_100k_motion") return cover_100k_motion(aspect, assets, 316);
    if (name == "cover_4k_motion") return cover_100k_motion(aspect, assets, 64);
    if (name == "simple_light_scene") return simple_light_scene(aspect);
    if (name == "cornell_box") return cornell_box(aspect, false);
    if (name == "cornell_box_solid") return cornell_box(aspect, true);
    if (name == "final_scene") return final_scene(aspect, assets);
    if (name == "pbr_ibl") return pbr_ibl(aspect, assets);
    if (name == "smoke_instances") return smoke_instances(aspect);
    if (name.rfind("furnace_", 0) == 0) return furnace(aspect, name.substr(8));
    throw std::runtime_error("unknown scene " + name);
}

}  // namespace orc
