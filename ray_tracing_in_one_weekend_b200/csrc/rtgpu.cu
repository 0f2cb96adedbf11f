// librtgpu.so — C ABI entry points (include/rtgpu.h): scene upload, batch closest hit, render, resolve, FP32 peak
// probe.  There is no CPU fallback: without a usable CUDA device every compute entry point fails with RT_ERR_CUDA.
// Multi-GPU entry points (NCCL) live in rt_comm.cu, test-only hooks in test_hooks.cu (librtgpu_test.so).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "bvh_build.hpp"
#include "gpu_bvh.hpp"
#include "kernels.cuh"
#include "kernels_cta.cuh"
#include "rt_internal.hpp"
#include "scene_api.hpp"

using namespace rtb;

static thread_local std::string g_err;
namespace rtb {
int rt_fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
int fail_png(const std::string& msg) { return rt_fail(RT_ERR_IO, msg); }  // png_writer.cpp
}  // namespace rtb
static int fail(int code, const std::string& msg) { return rt_fail(code, msg); }

// Development knobs: only builds made with -DRT_DEV_KNOBS (tools/build_variant.py) look at the environment, and
// only inside rt_scene_create; the shipped library never calls getenv.
static int dev_knob(const char* name, int dflt) {
#ifdef RT_DEV_KNOBS
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : dflt;
#else
    (void)name;
    return dflt;
#endif
}

// ------------------------------------------------------------------------------------- buffer cache
namespace rtb {
BufCache& BufCache::get() {
    static BufCache* c = new BufCache();  // leaky on purpose (see rt_internal.hpp)
    return *c;
}
constexpr size_t CACHE_MAX_IDLE_DEV = 2ull << 30, CACHE_MAX_IDLE_PIN = 1ull << 30;
cudaError_t BufCache::take(int device, size_t bytes, void** out, size_t* got) {
    bytes = std::max<size_t>((bytes + 4095) & ~(size_t)4095, 4096);
    {
        std::lock_guard<std::mutex> lk(mu_);
        int best = -1;
        for (int i = 0; i < (int)idle_.size(); i++)  // best fit, never more than 2x what was asked for
            if (idle_[i].device == device && idle_[i].bytes >= bytes && idle_[i].bytes <= 2 * bytes &&
                (best < 0 || idle_[i].bytes < idle_[best].bytes))
                best = i;
        if (best >= 0) {
            *out = idle_[best].p;
            *got = idle_[best].bytes;
            (device < 0 ? idle_bytes_pin_ : idle_bytes_dev_) -= idle_[best].bytes;
            idle_.erase(idle_.begin() + best);
            return cudaSuccess;
        }
    }
    *got = bytes;
    if (device < 0) return cudaMallocHost(out, bytes);
    cudaError_t e = cudaMalloc(out, bytes);
    if (e == cudaErrorMemoryAllocation) {  // make room: drop what idles, try once more
        cudaGetLastError();
        trim();
        e = cudaMalloc(out, bytes);
    }
    return e;
}
void BufCache::give(int device, void* p, size_t bytes) {
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(mu_);
        size_t& idle = device < 0 ? idle_bytes_pin_ : idle_bytes_dev_;
        if (idle + bytes <= (device < 0 ? CACHE_MAX_IDLE_PIN : CACHE_MAX_IDLE_DEV)) {
            idle_.push_back(Buf{p, bytes, device});
            idle += bytes;
            return;
        }
    }
    if (device < 0) cudaFreeHost(p);
    else cudaFree(p);
}
cudaError_t BufCache::take_device(int device, size_t bytes, void** out, size_t* got) { return take(device, bytes, out, got); }
void BufCache::give_device(int device, void* p, size_t bytes) { give(device, p, bytes); }
cudaError_t BufCache::take_pinned(size_t bytes, void** out, size_t* got) { return take(-1, bytes, out, got); }
void BufCache::give_pinned(void* p, size_t bytes) { give(-1, p, bytes); }
void BufCache::trim() {
    std::vector<Buf> drop;
    {
        std::lock_guard<std::mutex> lk(mu_);
        drop.swap(idle_);
        idle_bytes_dev_ = idle_bytes_pin_ = 0;
    }
    int cur = 0;
    cudaGetDevice(&cur);
    for (const Buf& b : drop) {
        if (b.device < 0) cudaFreeHost(b.p);
        else {
            cudaSetDevice(b.device);
            cudaFree(b.p);
        }
    }
    cudaSetDevice(cur);
}
}  // namespace rtb

RtScene::~RtScene() {
    cudaSetDevice(device);
    // the arena goes back to the cache, not to the driver: nothing may still be reading it
    if (ev_last) cudaEventSynchronize(ev_last);
    if (stream) cudaStreamSynchronize(stream);
    for (auto t : texobjs) cudaDestroyTextureObject(t);
    if (gpu_nodes) cudaFree(gpu_nodes);
    for (cudaEvent_t e : ring)
        if (e) cudaEventDestroy(e);
    for (cudaEvent_t e : {ev_last, ev_stage[0], ev_stage[1]})
        if (e) cudaEventDestroy(e);
    if (stream) cudaStreamDestroy(stream);
}

namespace {

std::string validate(const RtFlatScene* d) {
    if (!d) return "desc is NULL";
    if (d->abi_version != RT_ABI_VERSION) return "abi_version mismatch";
    if (d->n_prims == 0 || !d->prims) return "scene has no primitives";
    if (d->n_xform_chains == 0 || !d->xform_chains || d->xform_chains[0].n_ops != 0) return "xform chain 0 must be the identity";
    if (d->n_prims > LEAF_FIRST_MASK) return "too many primitives";
    if (d->n_xform_chains > 0xffffu) return "too many transform chains (max 65535)";
    for (uint32_t i = 0; i < d->n_xform_chains; i++) {
        const RtXformChain& c = d->xform_chains[i];
        if (c.n_ops > (uint32_t)MAX_CHAIN) return "transform chain deeper than 4";
        if ((uint64_t)c.first_op + c.n_ops > d->n_xform_ops) return "transform chain out of range";
    }
    for (uint32_t i = 0; i < d->n_xform_ops; i++)
        if (d->xform_ops[i].kind > RT_XF_ROTATE_Y) return "bad transform op kind";
    for (uint32_t i = 0; i < d->n_textures; i++) {
        const RtTexture& t = d->textures[i];
        if (t.kind > RT_TEX_IMAGE) return "bad texture kind";
        if (t.kind == RT_TEX_CHECKER && (t.a >= d->n_textures || t.b >= d->n_textures)) return "checker child out of range";
        if (t.kind == RT_TEX_PERLIN && t.a >= d->n_perlin) return "perlin table out of range";
        if (t.kind == RT_TEX_IMAGE && t.a >= d->n_images) return "image out of range";
    }
    for (uint32_t i = 0; i < d->n_images; i++)
        if (!d->images[i].rgb8 || d->images[i].width == 0 || d->images[i].height == 0) return "empty image";
    for (uint32_t i = 0; i < d->n_materials; i++) {
        const RtMaterial& m = d->materials[i];
        if (m.kind > RT_MAT_DISNEY_CLEARCOAT) return "bad material kind";
        bool t0 = !(m.kind == RT_MAT_METAL || m.kind == RT_MAT_DIELECTRIC || m.kind == RT_MAT_DISNEY_CLEARCOAT);
        if (t0 && m.tex0 >= d->n_textures) return "material texture out of range";
        if (m.kind == RT_MAT_ROUGH_PLASTIC && m.tex1 >= d->n_textures) return "material texture out of range";
    }
    for (uint32_t i = 0; i < d->n_prims; i++) {
        const RtPrim& p = d->prims[i];
        if (p.kind > RT_PRIM_MEDIUM) return "bad primitive kind";
        if (p.mat >= d->n_materials) return "primitive material out of range";
        if (p.xform >= d->n_xform_chains) return "primitive transform out of range";
        if (p.kind == RT_PRIM_MEDIUM) {
            if (p.aux1 == 0 || (uint64_t)p.aux0 + p.aux1 > d->n_prims) return "medium boundary out of range";
            if (d->materials[p.mat].kind != RT_MAT_ISOTROPIC) return "medium material must be Isotropic";
            const RtXformChain& mc = d->xform_chains[p.xform];
            for (uint32_t k = 0; k < p.aux1; k++) {
                const RtPrim& b = d->prims[p.aux0 + k];
                if (!(b.flags & RT_PRIM_FLAG_BOUNDARY_ONLY)) return "medium boundary must be boundary-only geometry";
                // a wrapper around the medium wraps its boundary too: the boundary's chain starts with the medium's
                if (b.xform >= d->n_xform_chains) return "primitive transform out of range";
                const RtXformChain& bc = d->xform_chains[b.xform];
                bool prefix = bc.n_ops >= mc.n_ops && (uint64_t)bc.first_op + bc.n_ops <= d->n_xform_ops;
                for (uint32_t j = 0; prefix && j < mc.n_ops; j++)
                    prefix = std::memcmp(&d->xform_ops[bc.first_op + j], &d->xform_ops[mc.first_op + j], sizeof(RtXformOp)) == 0;
                if (!prefix) return "medium boundary chain must start with the medium's own transform chain";
            }
        }
    }
    if (d->sky.kind > RT_SKY_CONSTANT) return "bad sky kind";
    if (d->sky.kind == RT_SKY_ENVMAP && d->sky.image >= d->n_images) return "sky image out of range";
    return "";
}

// does texture i (transitively through checker children) contain a texture of `kind`?
bool tex_uses_kind(const RtFlatScene* d, uint32_t i, uint32_t kind, int depth = 0) {
    const RtTexture& t = d->textures[i];
    if (t.kind == kind) return true;
    if (t.kind == RT_TEX_CHECKER && depth < 8) return tex_uses_kind(d, t.a, kind, depth + 1) || tex_uses_kind(d, t.b, kind, depth + 1);
    return false;
}

// does texture i (transitively) read uv?
bool tex_needs_uv(const RtFlatScene* d, uint32_t i, int depth = 0) {
    const RtTexture& t = d->textures[i];
    if (t.kind == RT_TEX_IMAGE) return true;
    if (t.kind == RT_TEX_CHECKER && depth < 8) return tex_needs_uv(d, t.a, depth + 1) || tex_needs_uv(d, t.b, depth + 1);
    return false;
}

// Shading class of a material for the wavefront's material-sorted shade pass (rt_device.cuh CLS_*).
uint32_t material_class(const RtFlatScene* d, const RtMaterial& m) {
    if (m.kind >= RT_MAT_OREN_NAYAR) return (uint32_t)CLS_PBR;
    if (m.kind == RT_MAT_METAL) return (uint32_t)CLS_METAL;
    if (m.kind == RT_MAT_DIELECTRIC) return (uint32_t)CLS_DIELECTRIC;
    if (m.kind == RT_MAT_DIFFUSE && d->textures[m.tex0].kind == RT_TEX_CONSTANT) return (uint32_t)CLS_DIFFUSE_CONST;
    if (m.kind == RT_MAT_DIFFUSE && d->textures[m.tex0].kind == RT_TEX_PERLIN) return (uint32_t)CLS_DIFFUSE_PERLIN;
    return (uint32_t)CLS_CLASSIC;
}

// Shade queues of the CTA wavefront renderer.  The kernel has CLS_COUNT queues per block: queue 0 takes the misses
// and every shading class of the scene needs one; the queues a scene leaves free are handed to the heaviest
// (material kind, texture kind) groups of its multi-kind classes (CLS_PBR, CLS_CLASSIC), so that a warp of the shade
// phase runs ONE material kind wherever the scene allows it (pbr_ibl: seven pbr.rs kinds in one class otherwise).
// Weight of a group = surface area of the world primitives that carry it (a proxy for how often it is hit).
struct ShadeQueues {
    std::vector<uint8_t> mat_queue;  // per material
    uint32_t q_cls = 0, q_order = 0;
};
ShadeQueues assign_shade_queues(const RtFlatScene* d) {
    ShadeQueues out;
    out.mat_queue.assign(d->n_materials, 0);
    auto sig_of = [&](const RtMaterial& m) {
        const bool t0 = !(m.kind == RT_MAT_METAL || m.kind == RT_MAT_DIELECTRIC || m.kind == RT_MAT_DISNEY_CLEARCOAT);
        return material_class(d, m) | (m.kind << 8) | ((t0 ? d->textures[m.tex0].kind + 1u : 0u) << 16);
    };
    std::vector<double> mat_w(d->n_materials, 0.0);
    std::vector<char> used(d->n_materials, 0);
    for (uint32_t i = 0; i < d->n_prims; i++) {
        const RtPrim& p = d->prims[i];
        if (p.mat >= d->n_materials) continue;
        double w = 1.0;
        if (p.kind == RT_PRIM_SPHERE || p.kind == RT_PRIM_MOVING_SPHERE) {
            w = 12.566370614359172 * (double)p.g[3] * (double)p.g[3];
        } else if (p.kind == RT_PRIM_XYRECT || p.kind == RT_PRIM_XZRECT || p.kind == RT_PRIM_YZRECT) {
            w = 1.0;
            for (int k = 0; k < 3; k++) {
                const double e = std::fabs((double)p.g[3 + k] - (double)p.g[k]);
                if (e > 0.0) w *= e;
            }
        }
        mat_w[p.mat] += w;
        used[p.mat] = 1;
    }
    struct Group { uint32_t sig; double w; int queue; };
    std::vector<Group> groups;
    for (uint32_t i = 0; i < d->n_materials; i++) {
        if (!used[i]) continue;
        const uint32_t sg = sig_of(d->materials[i]);
        auto it = std::find_if(groups.begin(), groups.end(), [&](const Group& g) { return g.sig == sg; });
        if (it == groups.end()) groups.push_back(Group{sg, mat_w[i], -1});
        else it->w += mat_w[i];
    }
    // one queue per class in use ...
    int cls_queue[CLS_COUNT];
    uint8_t q_cls[8] = {(uint8_t)CLS_MISS, 0, 0, 0, 0, 0, 0, 0};
    int n_q = 1;
    for (int c = 1; c < CLS_COUNT; c++) {
        cls_queue[c] = -1;
        if (std::any_of(groups.begin(), groups.end(), [&](const Group& g) { return (int)(g.sig & 0xffu) == c; })) {
            cls_queue[c] = n_q;
            q_cls[n_q++] = (uint8_t)c;
        }
    }
    // ... and the free ones to the heaviest group of a class whose queue still mixes several groups
    const bool split = dev_knob("RT_CTA_SPLIT_QUEUES", 1) != 0;  // development builds: 0 = one queue per class
    for (; split && n_q < CLS_COUNT; n_q++) {
        Group* best = nullptr;
        for (Group& g : groups) {
            if (g.queue >= 0) continue;
            const uint32_t c = g.sig & 0xffu;
            const auto shared = std::count_if(groups.begin(), groups.end(), [&](const Group& h) { return (h.sig & 0xffu) == c && h.queue < 0; });
            if (shared >= 2 && (!best || g.w > best->w)) best = &g;
        }
        if (!best) break;
        best->queue = n_q;
        q_cls[n_q] = (uint8_t)(best->sig & 0xffu);
    }
    for (Group& g : groups)
        if (g.queue < 0) g.queue = cls_queue[g.sig & 0xffu];
    for (uint32_t i = 0; i < d->n_materials; i++) {
        if (!used[i]) {  // unreferenced material: any queue of its class (or the class's would-be queue 0 slot)
            const int c = (int)material_class(d, d->materials[i]);
            out.mat_queue[i] = (uint8_t)(cls_queue[c] >= 0 ? cls_queue[c] : 0);
            continue;
        }
        const uint32_t sg = sig_of(d->materials[i]);
        out.mat_queue[i] = (uint8_t)std::find_if(groups.begin(), groups.end(), [&](const Group& g) { return g.sig == sg; })->queue;
    }
    // The shade phase drains queues 1, 2, ..., 6 and then 0: renumber the hit queues so that the most expensive
    // class comes first (its long chunks start first) and the misses last.
    static const int rank[CLS_COUNT] = {/*MISS*/ 6, /*DIFFUSE_CONST*/ 4, /*METAL*/ 5, /*DIELECTRIC*/ 3, /*PERLIN*/ 0, /*CLASSIC*/ 2, /*PBR*/ 1};
    int by_cost[CLS_COUNT];
    for (int q = 0; q < CLS_COUNT; q++) by_cost[q] = q;
    std::stable_sort(by_cost + 1, by_cost + n_q, [&](int a, int b) { return rank[q_cls[a]] < rank[q_cls[b]]; });
    uint8_t renum[8] = {0, 1, 2, 3, 4, 5, 6, 7}, cls_sorted[8] = {(uint8_t)CLS_MISS, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 1; k < n_q; k++) {
        renum[by_cost[k]] = (uint8_t)k;
        cls_sorted[k] = q_cls[by_cost[k]];
    }
    for (auto& mq : out.mat_queue) mq = renum[mq];
    for (int q = 0; q < CLS_COUNT; q++) {
        out.q_cls |= (uint32_t)cls_sorted[q] << (4 * q);
        out.q_order |= (uint32_t)((q + 1) % CLS_COUNT) << (4 * q);  // informational: 1, 2, ..., 6, 0
    }
    return out;
}

DevPrim make_dev_prim(const RtFlatScene* d, uint32_t flat, uint32_t dev_index, std::vector<DevMotion>& motion,
                      const std::vector<uint32_t>& flat_to_dev, const std::vector<uint8_t>& mat_queue) {
    const RtPrim& p = d->prims[flat];
    DevPrim q{};
    q.mat = p.mat;
    // kind | flag (1 bit) << 8 | CTA shade queue (3 bits) << 9 | shading class << 12 | transform chain << 16
    q.kfx = (p.kind & 0xffu) | ((p.flags & 0x1u) << 8) | (((uint32_t)mat_queue[p.mat] & 7u) << 9) |
            (material_class(d, d->materials[p.mat]) << 12) | (p.xform << 16);
    q.flat = flat;
    auto bits = [](uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; };
    auto ubits = [](float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; };
    switch (p.kind) {
        case RT_PRIM_SPHERE:
            q.g0 = p.g[0]; q.g1 = p.g[1]; q.g2 = p.g[2]; q.g3 = p.g[3];
            break;
        case RT_PRIM_MOVING_SPHERE: {
            q.g0 = p.g[0]; q.g1 = p.g[1]; q.g2 = p.g[2]; q.g3 = p.g[3];
            DevMotion m{};
            m.dcx = p.g[4] - p.g[0]; m.dcy = p.g[5] - p.g[1]; m.dcz = p.g[6] - p.g[2];
            m.time0 = p.g[7];
            m.dt = p.g[8] - p.g[7];
            // the motion table runs parallel to the device primitive array (rt_device.cuh load_motion): the
            // kernels load record dev_index together with the primitive, q.x is informational
            if (motion.size() <= dev_index) motion.resize((size_t)dev_index + 1, DevMotion{0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f});
            motion[dev_index] = m;
            q.x = dev_index;
            break;
        }
        case RT_PRIM_XYRECT:  // plane z; a = x, b = y
            q.g0 = p.g[2]; q.g1 = p.g[0]; q.g2 = p.g[3]; q.g3 = p.g[1]; q.x = ubits(p.g[4]);
            break;
        case RT_PRIM_XZRECT:  // plane y; a = x, b = z
            q.g0 = p.g[1]; q.g1 = p.g[0]; q.g2 = p.g[3]; q.g3 = p.g[2]; q.x = ubits(p.g[5]);
            break;
        case RT_PRIM_YZRECT:  // plane x; a = y, b = z
            q.g0 = p.g[0]; q.g1 = p.g[1]; q.g2 = p.g[4]; q.g3 = p.g[2]; q.x = ubits(p.g[5]);
            break;
        case RT_PRIM_MEDIUM:
            q.g0 = p.g[0];
            q.g1 = bits(flat_to_dev[p.aux0]);
            q.g2 = bits(p.aux1);
            {  // g3 = 1: the boundary is <= 6 rects under one transform chain (medium_test's single-pass path)
                bool box = p.aux1 <= 6;
                for (uint32_t k = 0; k < p.aux1 && box; k++) {
                    const RtPrim& b = d->prims[p.aux0 + k];
                    box = (b.kind == RT_PRIM_XYRECT || b.kind == RT_PRIM_XZRECT || b.kind == RT_PRIM_YZRECT) &&
                          b.xform == d->prims[p.aux0].xform;
                }
                q.g3 = bits(box ? 1u : 0u);
            }
            break;
    }
    return q;
}
// What the backend cannot render the way the reference would (-> RT_ERR_UNSUPPORTED instead of a silently different
// picture).  (1) A ConstantMedium whose boundary is itself a ConstantMedium (hitable.rs:523-527 allows any Hitable;
// a stochastic boundary has no closed volume).  (2) Quirk A4: the reference's rects never write
// rec.tang (hitable.rs:262-269) and its media write neither uv nor tang (568-576), so a material that reads them
// there sees whatever an EARLIER candidate of the same traversal left behind — traversal-order dependent, not
// reproducible by any other traversal: DisneyMetal's anisotropic branch (pbr.rs:243-262) on anything but a sphere,
// and an Isotropic phase function whose texture reads uv (texture.rs:185-192) on a medium.
std::string unsupported(const RtFlatScene* d) {
    for (uint32_t i = 0; i < d->n_prims; i++) {
        const RtPrim& p = d->prims[i];
        const RtMaterial& m = d->materials[p.mat];
        if (p.kind == RT_PRIM_MEDIUM) {
            for (uint32_t k = 0; k < p.aux1; k++)
                if (d->prims[p.aux0 + k].kind == RT_PRIM_MEDIUM) return "a ConstantMedium as the boundary of a ConstantMedium is not supported";
            if (tex_needs_uv(d, m.tex0))
                return "Isotropic with a uv-reading texture on a ConstantMedium reads a stale HitRecord.uv in the reference (hitable.rs:568-576): not reproducible";
        } else if (m.kind == RT_MAT_DISNEY_METAL && m.p[1] > -10.0f && !(p.kind == RT_PRIM_SPHERE || p.kind == RT_PRIM_MOVING_SPHERE)) {
            return "DisneyMetal on a rect reads a stale HitRecord.tang in the reference (hitable.rs:262-269, pbr.rs:243-262): not reproducible; use it on spheres";
        }
    }
    return "";
}

size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

extern "C" {

const char* rt_last_error(void) { return g_err.c_str(); }

void rt_release_cached_memory(void) { BufCache::get().trim(); }

int rt_partition_samples(uint32_t sample_begin, uint32_t sample_end, int n_ranks, int rank, uint32_t* out_begin, uint32_t* out_end) {
    if (!out_begin || !out_end || n_ranks < 1 || rank < 0 || rank >= n_ranks || sample_end < sample_begin) return fail(RT_ERR_ARG, "bad partition arguments");
    const uint32_t n = sample_end - sample_begin, base = n / (uint32_t)n_ranks, extra = n % (uint32_t)n_ranks;
    const uint32_t r = (uint32_t)rank;
    *out_begin = sample_begin + r * base + std::min(r, extra);
    *out_end = *out_begin + base + (r < extra ? 1u : 0u);
    return RT_OK;
}

int rt_scene_create(const RtFlatScene* desc, int device, RtScene** out) {
    // Builder policy (results are identical either way): the host SAH tree up to 262 144 world primitives, the device
    // LBVH from there on.  The SAH tree traverses faster — cover_100k_motion (99 860 spheres) renders 1 172 against 977
    // Mpaths/s at 1080p, cover_4k_motion 2 792 against 2 380 (B200, tools/builder_ab.py) — and builds in 0.3 s at
    // 100 k primitives, which a frame of a few samples per pixel at that size already repays; a caller that rebuilds
    // its scene every frame asks for the device builder with rt_scene_create_ex(RT_SCENE_GPU_BVH).
    constexpr uint32_t AUTO_GPU_BVH_MIN_PRIMS = 262144u;
    uint32_t flags = 0;
    if (desc && desc->prims && desc->n_prims >= AUTO_GPU_BVH_MIN_PRIMS) {
        uint32_t world = 0;
        for (uint32_t i = 0; i < desc->n_prims; i++) world += (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) ? 0u : 1u;
        if (world >= AUTO_GPU_BVH_MIN_PRIMS) flags = RT_SCENE_GPU_BVH;
    }
    const int forced = dev_knob("RT_SCENE_FLAGS", -1);
    if (forced >= 0) flags = (uint32_t)forced;
    return rt_scene_create_ex(desc, device, flags, out);
}

int rt_scene_create_ex(const RtFlatScene* desc, int device, uint32_t flags, RtScene** out) {
    if (!out) return fail(RT_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (flags & ~RT_SCENE_GPU_BVH) return fail(RT_ERR_ARG, "unknown scene flag");
    std::string why = validate(desc);
    if (!why.empty()) return fail(RT_ERR_ARG, why);
    why = unsupported(desc);
    if (!why.empty()) return fail(RT_ERR_UNSUPPORTED, why);
    int n_dev = 0;
    CU_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(RT_ERR_CUDA, "no such CUDA device");
    CU_TRY(cudaSetDevice(device));

    const bool timing = dev_knob("RT_TIMING", 0) != 0;  // development builds: stage times on stderr
    auto t_prev = std::chrono::steady_clock::now();
    auto stage = [&](const char* what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[rt_scene_create] %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(now - t_prev).count());
        t_prev = now;
    };
    std::unique_ptr<RtScene> s(new RtScene());
    s->device = device;
    CU_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    for (cudaEvent_t& e : s->ring) CU_TRY(cudaEventCreate(&e));
    s->ev0 = s->ring[0];
    s->ev1 = s->ring[1];
    CU_TRY(cudaEventCreateWithFlags(&s->ev_last, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s->ev_stage[0], cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s->ev_stage[1], cudaEventDisableTiming));
    CU_TRY(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device));  // (cudaGetDeviceProperties costs milliseconds)
    int tex_align = 512, pitch_align = 32;
    CU_TRY(cudaDeviceGetAttribute(&tex_align, cudaDevAttrTextureAlignment, device));
    CU_TRY(cudaDeviceGetAttribute(&pitch_align, cudaDevAttrTexturePitchAlignment, device));
    s->knob_mode = dev_knob("RT_DEFAULT_MODE", RT_MODE_CTA_WAVEFRONT);
    s->knob_paths = dev_knob("RT_CTA_PATHS", 0);
    s->knob_unified = dev_knob("RT_CTA_UNIFIED", -1);
    s->knob_refill_camera = dev_knob("RT_CTA_REFILL_CAMERA", 24);
    s->knob_refill = dev_knob("RT_CTA_REFILL", 8);
    s->knob_blocks = dev_knob("RT_CTA_BLOCKS", 64);
    s->knob_carveout = dev_knob("RT_CTA_CARVEOUT", -1);
    s->knob_full_kernel = dev_knob("RT_CTA_FULL_KERNEL", 0);
    s->knob_units_per_warp = dev_knob("RT_UNITS_PER_WARP", 24);
    s->knob_flat = dev_knob("RT_CTA_FLAT", -1);
    stage("stream/events/attrs");

    // "Giant" primitives (rt_device.cuh giants_test): untransformed static spheres whose extent exceeds 64 x the median
    // extent of the world primitives, at most two, in scenes of more than 64 world primitives (smaller scenes take the
    // flat loop or gain nothing).  The builders see a copy of the description in which they are flagged boundary-only,
    // i.e. not part of the world; the kernels test them once per ray before the tree.
    std::vector<uint32_t> giants;
    std::vector<RtPrim> prims_for_bvh;
    RtFlatScene bvh_desc = *desc;
    uint32_t n_world_total = 0;
    {
        std::vector<float> ext;
        for (uint32_t i = 0; i < desc->n_prims; i++) {
            const RtPrim& p = desc->prims[i];
            if (p.flags & RT_PRIM_FLAG_BOUNDARY_ONLY) continue;
            n_world_total++;
            if (p.kind == RT_PRIM_SPHERE || p.kind == RT_PRIM_MOVING_SPHERE) ext.push_back(2.0f * p.g[3]);
            else if (p.kind != RT_PRIM_MEDIUM) ext.push_back(std::fmax(std::fmax(std::fabs(p.g[3] - p.g[0]), std::fabs(p.g[4] - p.g[1])), std::fabs(p.g[5] - p.g[2])));
        }
        if (n_world_total > 64u && !ext.empty() && dev_knob("RT_GIANTS", 1) != 0) {
            std::nth_element(ext.begin(), ext.begin() + ext.size() / 2, ext.end());
            const float median = ext[ext.size() / 2];
            for (uint32_t i = 0; i < desc->n_prims && giants.size() < 2; i++) {
                const RtPrim& p = desc->prims[i];
                if (!(p.flags & RT_PRIM_FLAG_BOUNDARY_ONLY) && p.kind == RT_PRIM_SPHERE && p.xform == 0 && 2.0f * p.g[3] > 64.0f * median) giants.push_back(i);
            }
        }
        if (!giants.empty()) {
            prims_for_bvh.assign(desc->prims, desc->prims + desc->n_prims);
            for (uint32_t g : giants) prims_for_bvh[g].flags |= RT_PRIM_FLAG_BOUNDARY_ONLY;
            bvh_desc.prims = prims_for_bvh.data();
        }
    }
    // acceleration structure: host SAH builder (default) or the GPU LBVH builder
    BvhBuildResult bvh;
    uint32_t builder = 0;
    float build_ms = 0.0f;
    auto b0 = std::chrono::steady_clock::now();
    if (flags & RT_SCENE_GPU_BVH) {
        GpuBvhResult g;
        std::string err;
        if (!gpu_build_bvh4(bvh_desc, s->stream, g, err)) return fail(RT_ERR_CUDA, "GPU BVH build: " + err);
        if (g.max_depth * 3u + 4u <= (uint32_t)STACK_SIZE) {
            s->gpu_nodes = g.nodes;
            bvh.leaf_order = std::move(g.leaf_order);
            bvh.n_world_prims = g.n_world_prims;
            bvh.max_depth = g.max_depth;
            bvh.max_leaf_prims = 1;
            bvh.nodes.resize(g.n_nodes);  // sizes only; the nodes already live on the device
            builder = 1;
        } else {
            cudaFree(g.nodes);  // degenerate Morton tree, too deep for the traversal stack: use the SAH builder
        }
    }
    if (builder == 0) {
        try {
            build_bvh4(bvh_desc, (uint32_t)dev_knob("RT_LEAF_MAX", 1), bvh);
        } catch (const std::exception& e) {
            return fail(RT_ERR_ARG, e.what());
        }
    }
    build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - b0).count();
    if (bvh.max_depth * 3u + 4u > (uint32_t)STACK_SIZE) return fail(RT_ERR_UNSUPPORTED, "BVH too deep for the traversal stack");
    stage("bvh build");

    // device primitive array: world primitives in BVH leaf order, then boundary-only primitives in flat order
    // ... with the giants between the two
    std::vector<uint32_t> order = bvh.leaf_order;
    const uint32_t giant_first = (uint32_t)order.size();
    for (uint32_t g : giants) order.push_back(g);
    for (uint32_t i = 0; i < desc->n_prims; i++)
        if (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) order.push_back(i);
    std::vector<uint32_t> flat_to_dev(desc->n_prims, 0xffffffffu);
    for (uint32_t k = 0; k < order.size(); k++) flat_to_dev[order[k]] = k;
    std::vector<DevMotion> motion;
    std::vector<DevPrim> prims(order.size());
    const ShadeQueues queues = assign_shade_queues(desc);
    s->q_cls = queues.q_cls;
    s->q_order = queues.q_order;
    if (timing) std::fprintf(stderr, "[rtgpu] shade queues: class per queue %07x, drain order %07x\n", s->q_cls, s->q_order);
    bool has_media = false;
    for (uint32_t k = 0; k < order.size(); k++) {
        prims[k] = make_dev_prim(desc, order[k], k, motion, flat_to_dev, queues.mat_queue);
        has_media |= desc->prims[order[k]].kind == RT_PRIM_MEDIUM;
    }
    // one record per primitive (static ones: dc = 0, dt = 1): every kernel built with F_MOTION loads record i with
    // primitive i, and the megakernel / batch entry points run the all-features build on any scene
    motion.resize(order.size(), DevMotion{0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f});
    // boundary ranges must stay contiguous on the device
    for (uint32_t i = 0; i < desc->n_prims; i++)
        if (desc->prims[i].kind == RT_PRIM_MEDIUM)
            for (uint32_t k = 1; k < desc->prims[i].aux1; k++)
                if (flat_to_dev[desc->prims[i].aux0 + k] != flat_to_dev[desc->prims[i].aux0] + k)
                    return fail(RT_ERR_ARG, "medium boundary primitives are not contiguous");

    std::vector<DevXformOp> ops(desc->n_xform_ops);
    for (uint32_t i = 0; i < desc->n_xform_ops; i++)
        ops[i] = DevXformOp{desc->xform_ops[i].kind, desc->xform_ops[i].v[0], desc->xform_ops[i].v[1], desc->xform_ops[i].v[2]};
    std::vector<DevXformChain> chains(desc->n_xform_chains);
    for (uint32_t i = 0; i < desc->n_xform_chains; i++) chains[i] = DevXformChain{desc->xform_chains[i].first_op, desc->xform_chains[i].n_ops};
    std::vector<DevMaterial> mats(desc->n_materials);
    for (uint32_t i = 0; i < desc->n_materials; i++) {
        const RtMaterial& m = desc->materials[i];
        bool t0 = !(m.kind == RT_MAT_METAL || m.kind == RT_MAT_DIELECTRIC || m.kind == RT_MAT_DISNEY_CLEARCOAT);
        uint32_t mflags = 0;
        if (t0 && tex_needs_uv(desc, m.tex0)) mflags |= 1u;
        if (m.kind == RT_MAT_ROUGH_PLASTIC && tex_needs_uv(desc, m.tex1)) mflags |= 1u;
        const uint32_t cls = material_class(desc, m);
        s->class_mask |= 1u << cls;
        mflags |= cls << 8;
        mats[i] = DevMaterial{m.kind, m.tex0, m.tex1, mflags, m.p[0], m.p[1], m.p[2], m.p[3]};
        if (cls == (uint32_t)CLS_DIFFUSE_CONST) {
            // Diffuse has no parameters of its own (material.rs:31-33): keep the ConstantTex colour in the
            // material record so the class-specialised shader needs one table lookup, not two
            const RtTexture& t = desc->textures[m.tex0];
            mats[i].p0 = t.p[0];
            mats[i].p1 = t.p[1];
            mats[i].p2 = t.p[2];
        }
    }
    std::vector<DevTexture> texs(desc->n_textures);
    for (uint32_t i = 0; i < desc->n_textures; i++) {
        const RtTexture& t = desc->textures[i];
        texs[i] = DevTexture{t.kind, t.a, t.b, 0, t.p[0], t.p[1], t.p[2], t.p[3]};
    }
    stage("host packing");

    // ---- ONE device allocation (the "arena") and ONE host-to-device copy for every table of the scene: laid out
    //      here, assembled in a pinned staging buffer, uploaded with a single cudaMemcpyAsync.
    size_t total = 0;
    auto reserve = [&](size_t bytes, size_t align) {
        total = align_up(total, align);
        const size_t off = total;
        total += std::max<size_t>(bytes, 16);
        return off;
    };
    const size_t n_grad = std::max<size_t>(1, (size_t)desc->n_perlin * 256), n_perm = std::max<size_t>(1, (size_t)desc->n_perlin * 768);
    const size_t off_scratch = reserve(24 * sizeof(unsigned long long), 256);
    const size_t off_nodes = s->gpu_nodes ? 0 : reserve(bvh.nodes.size() * sizeof(BvhNode4), 256);
    const size_t off_prims = reserve(prims.size() * sizeof(DevPrim), 256);
    const size_t off_motion = reserve(motion.size() * sizeof(DevMotion), 256);
    const size_t off_ops = reserve(ops.size() * sizeof(DevXformOp), 256);
    const size_t off_chains = reserve(chains.size() * sizeof(DevXformChain), 256);
    const size_t off_mats = reserve(mats.size() * sizeof(DevMaterial), 256);
    const size_t off_texs = reserve(texs.size() * sizeof(DevTexture), 256);
    const size_t off_grad = reserve(n_grad * sizeof(float4), (size_t)std::max(tex_align, 256));
    const size_t off_perm = reserve(n_perm, 256);
    const size_t off_imgtab = reserve(desc->n_images * sizeof(DevImage), 256);
    std::vector<uint32_t> flat_order;  // world primitives in flat (visitation) order -> device index
    for (uint32_t i = 0; i < desc->n_prims; i++)
        if (!(desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY)) flat_order.push_back(flat_to_dev[i]);
    const size_t off_flat = reserve(flat_order.size() * sizeof(uint32_t), 256);
    std::vector<size_t> off_img(desc->n_images), pitch_img(desc->n_images);
    for (uint32_t i = 0; i < desc->n_images; i++) {
        pitch_img[i] = align_up((size_t)desc->images[i].width * 4, (size_t)std::max(pitch_align, 128));
        off_img[i] = reserve(pitch_img[i] * desc->images[i].height, (size_t)std::max(tex_align, 512));
    }
    total = align_up(total, 256);
    Lease staging;
    CU_TRY(staging.pinned(total));
    CU_TRY(s->arena.device_mem(device, total));
    char* h = staging.as<char>();
    char* dbase = s->arena.as<char>();
    auto put = [&](size_t off, const void* src, size_t bytes) {
        if (bytes) std::memcpy(h + off, src, bytes);
    };
    std::memset(h + off_scratch, 0, 24 * sizeof(unsigned long long));
    if (!s->gpu_nodes) put(off_nodes, bvh.nodes.data(), bvh.nodes.size() * sizeof(BvhNode4));
    put(off_prims, prims.data(), prims.size() * sizeof(DevPrim));
    put(off_motion, motion.data(), motion.size() * sizeof(DevMotion));
    put(off_ops, ops.data(), ops.size() * sizeof(DevXformOp));
    put(off_chains, chains.data(), chains.size() * sizeof(DevXformChain));
    put(off_mats, mats.data(), mats.size() * sizeof(DevMaterial));
    put(off_texs, texs.data(), texs.size() * sizeof(DevTexture));
    put(off_flat, flat_order.data(), flat_order.size() * sizeof(uint32_t));
    {  // Perlin tables: 256 float4 gradients + 768 permutation bytes per table
        float4* grad = reinterpret_cast<float4*>(h + off_grad);
        uint8_t* perm = reinterpret_cast<uint8_t*>(h + off_perm);
        grad[0] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        perm[0] = 0;
        for (uint32_t t = 0; t < desc->n_perlin; t++) {
            const RtPerlin& p = desc->perlin[t];
            for (int k = 0; k < 256; k++) {
                grad[t * 256 + k] = make_float4(p.rand_vec[k][0], p.rand_vec[k][1], p.rand_vec[k][2], 0.0f);
                perm[t * 768 + k] = p.perm_x[k];
                perm[t * 768 + 256 + k] = p.perm_y[k];
                perm[t * 768 + 512 + k] = p.perm_z[k];
            }
        }
    }
    for (uint32_t i = 0; i < desc->n_images; i++) {  // RGB8 -> RGBA8 rows at the texture pitch
        const RtImage& im = desc->images[i];
        for (uint32_t y = 0; y < im.height; y++) {
            uchar4* row = reinterpret_cast<uchar4*>(h + off_img[i] + (size_t)y * pitch_img[i]);
            const uint8_t* src = im.rgb8 + (size_t)y * im.width * 3;
            for (uint32_t x = 0; x < im.width; x++) row[x] = make_uchar4(src[3 * x], src[3 * x + 1], src[3 * x + 2], 255);
        }
    }
    stage("staging");
    // texture objects (their handles go into the image table, which is part of the same upload)
    {
        cudaResourceDesc rd{};
        rd.resType = cudaResourceTypeLinear;
        rd.res.linear.devPtr = dbase + off_grad;
        rd.res.linear.desc = cudaCreateChannelDesc<float4>();
        rd.res.linear.sizeInBytes = n_grad * sizeof(float4);
        cudaTextureDesc td{};
        td.readMode = cudaReadModeElementType;
        cudaTextureObject_t to = 0;
        CU_TRY(cudaCreateTextureObject(&to, &rd, &td, nullptr));
        s->texobjs.push_back(to);
        s->ds.perlin_grad = to;
    }
    {
        // images: uchar4 texels in pitch-linear memory inside the arena behind point-sampled, clamped, unnormalised
        // texture objects; the texel index is computed in integer arithmetic exactly as texture.rs:186-189
        DevImage* imgs = reinterpret_cast<DevImage*>(h + off_imgtab);
        for (uint32_t i = 0; i < desc->n_images; i++) {
            const RtImage& im = desc->images[i];
            cudaResourceDesc rd{};
            rd.resType = cudaResourceTypePitch2D;
            rd.res.pitch2D.devPtr = dbase + off_img[i];
            rd.res.pitch2D.desc = cudaCreateChannelDesc<uchar4>();
            rd.res.pitch2D.width = im.width;
            rd.res.pitch2D.height = im.height;
            rd.res.pitch2D.pitchInBytes = pitch_img[i];
            cudaTextureDesc td{};
            td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
            td.filterMode = cudaFilterModePoint;
            td.readMode = cudaReadModeElementType;
            td.normalizedCoords = 0;
            cudaTextureObject_t to = 0;
            CU_TRY(cudaCreateTextureObject(&to, &rd, &td, nullptr));
            s->texobjs.push_back(to);
            imgs[i] = DevImage{to, im.width, im.height};
        }
    }
    CU_TRY(cudaMemcpyAsync(dbase, h, total, cudaMemcpyHostToDevice, s->stream));
    s->scratch = reinterpret_cast<unsigned long long*>(dbase + off_scratch);
    s->ds.nodes = s->gpu_nodes ? reinterpret_cast<const BvhNode4*>(s->gpu_nodes) : reinterpret_cast<const BvhNode4*>(dbase + off_nodes);
    s->ds.prims = reinterpret_cast<const DevPrim*>(dbase + off_prims);
    s->ds.motion = reinterpret_cast<const DevMotion*>(dbase + off_motion);
    s->ds.ops = reinterpret_cast<const DevXformOp*>(dbase + off_ops);
    s->ds.chains = reinterpret_cast<const DevXformChain*>(dbase + off_chains);
    s->ds.mats = reinterpret_cast<const DevMaterial*>(dbase + off_mats);
    s->ds.texs = reinterpret_cast<const DevTexture*>(dbase + off_texs);
    s->ds.perlin_grad_lin = reinterpret_cast<const float4*>(dbase + off_grad);
    s->ds.perlin_perm = reinterpret_cast<const uint8_t*>(dbase + off_perm);
    s->ds.images = reinterpret_cast<const DevImage*>(dbase + off_imgtab);
    s->ds.flat_order = reinterpret_cast<const uint32_t*>(dbase + off_flat);
    s->ds.n_perlin = desc->n_perlin;
    s->ds.sky = DevSky{desc->sky.kind, desc->sky.image, desc->sky.color[0], desc->sky.color[1], desc->sky.color[2]};
    const RtCamera& c = desc->camera;
    auto v3 = [](const float* p) { return V3f{p[0], p[1], p[2]}; };
    s->ds.cam = DevCamera{v3(c.origin), v3(c.horizontal), v3(c.vertical), v3(c.lower_left_corner), v3(c.u), v3(c.v), c.lens_radius, c.time0, c.time1};
    s->ds.n_world_prims = n_world_total;  // (the flat loop's count; scenes that take it have no giants)
    s->ds.n_prims = (uint32_t)prims.size();
    s->ds.has_media = has_media ? 1u : 0u;
    s->ds.giant_first = giant_first;
    s->ds.n_giants = (uint32_t)giants.size();
    s->n_world_prims = n_world_total;
    for (uint32_t i = 0; i < desc->n_prims; i++) {
        const RtPrim& p = desc->prims[i];
        if (p.xform != 0) s->features |= F_XFORM;
        if (p.kind == RT_PRIM_MEDIUM) s->features |= F_MEDIA;
        if (p.kind == RT_PRIM_MOVING_SPHERE) s->features |= F_MOTION;
        if (p.kind == RT_PRIM_XYRECT || p.kind == RT_PRIM_XZRECT || p.kind == RT_PRIM_YZRECT) s->features |= F_RECT;
    }
    s->camera = c;
    s->bvh = RtBvhInfo{(uint32_t)bvh.nodes.size(), n_world_total, bvh.max_depth, bvh.max_leaf_prims, bvh.sah_cost, (uint32_t)sizeof(BvhNode4), builder, build_ms,
                       (uint64_t)total};
    CU_TRY(cudaStreamSynchronize(s->stream));  // the staging buffer goes back to the cache when this function returns
    stage("upload + sync");
    *out = s.release();
    return RT_OK;
}

void rt_scene_destroy(RtScene* scene) { delete scene; }

int rt_flat_shade_queues(const RtFlatScene* desc, uint8_t* mat_queue, uint32_t* q_class, uint32_t* q_order) {
    if (!desc || !mat_queue || !q_class || !q_order) return fail(RT_ERR_ARG, "NULL argument");
    for (uint32_t i = 0; i < desc->n_prims; i++)
        if (desc->prims[i].mat >= desc->n_materials) return fail(RT_ERR_ARG, "primitive with a bad material index");
    const ShadeQueues q = assign_shade_queues(desc);
    std::memcpy(mat_queue, q.mat_queue.data(), q.mat_queue.size());
    *q_class = q.q_cls;
    *q_order = q.q_order;
    return RT_OK;
}

int rt_flat_validate(const RtFlatScene* desc) {
    std::string why = validate(desc);
    if (!why.empty()) return fail(RT_ERR_ARG, why);
    why = unsupported(desc);
    if (!why.empty()) return fail(RT_ERR_UNSUPPORTED, why);
    return RT_OK;
}

int rt_scene_bvh_info(const RtScene* scene, RtBvhInfo* out) {
    if (!scene || !out) return fail(RT_ERR_ARG, "NULL argument");
    *out = scene->bvh;
    return RT_OK;
}

int rt_scene_kernel_times(RtScene* s, double* out_ms, int capacity, int* out_n) {
    if (!s || !out_ms || !out_n || capacity < 0) return fail(RT_ERR_ARG, "bad argument");
    CU_TRY(cudaSetDevice(s->device));
    const unsigned long long have = std::min<unsigned long long>(s->n_renders, (unsigned long long)RtScene::KERNEL_RING);
    const int n = (int)std::min<unsigned long long>(have, (unsigned long long)capacity);
    for (int k = 0; k < n; k++) {  // oldest of the last n first
        const int slot = (int)((s->n_renders - (unsigned long long)n + (unsigned long long)k) % RtScene::KERNEL_RING);
        float ms = 0.0f;
        CU_TRY(cudaEventSynchronize(s->ring[2 * slot + 1]));
        CU_TRY(cudaEventElapsedTime(&ms, s->ring[2 * slot], s->ring[2 * slot + 1]));
        out_ms[k] = ms;
    }
    *out_n = n;
    return RT_OK;
}

int rt_closest_hit_device(RtScene* s, const float* o, const float* d, const float* times, size_t n, float t_min, float t_max,
                          RtHit* out, void* cuda_stream) {
    if (!s || !o || !d || !out) return fail(RT_ERR_ARG, "NULL argument");
    if (n == 0) return RT_OK;
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : s->stream;
    unsigned blocks = (unsigned)((n + 127) / 128);
    k_closest_hit_batch<<<blocks, 128, 0, st>>>(s->ds, o, d, times, n, t_min, t_max, out);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaEventRecord(s->ev_last, st));
    return RT_OK;
}

int rt_closest_hit(RtScene* s, const float* o, const float* d, const float* times, size_t n, float t_min, float t_max, RtHit* out) {
    if (!s || !o || !d || !out) return fail(RT_ERR_ARG, "NULL argument");
    if (n == 0) return RT_OK;
    CU_TRY(cudaSetDevice(s->device));
    // one device buffer: origins | directions | times | hits
    const size_t off_d = align_up(n * 12, 256), off_t = off_d + align_up(n * 12, 256), off_h = off_t + align_up(n * 4, 256);
    Lease buf;
    CU_TRY(buf.device_mem(s->device, off_h + n * sizeof(RtHit)));
    char* b = buf.as<char>();
    CU_TRY(cudaMemcpyAsync(b, o, n * 12, cudaMemcpyHostToDevice, s->stream));
    CU_TRY(cudaMemcpyAsync(b + off_d, d, n * 12, cudaMemcpyHostToDevice, s->stream));
    if (times) CU_TRY(cudaMemcpyAsync(b + off_t, times, n * 4, cudaMemcpyHostToDevice, s->stream));
    int rc = rt_closest_hit_device(s, reinterpret_cast<float*>(b), reinterpret_cast<float*>(b + off_d),
                                   times ? reinterpret_cast<float*>(b + off_t) : nullptr, n, t_min, t_max,
                                   reinterpret_cast<RtHit*>(b + off_h), s->stream);
    if (rc != RT_OK) {
        cudaStreamSynchronize(s->stream);
        return rc;
    }
    cudaError_t e = cudaMemcpyAsync(out, b + off_h, n * sizeof(RtHit), cudaMemcpyDeviceToHost, s->stream);
    cudaError_t e2 = cudaStreamSynchronize(s->stream);  // also before `buf` returns to the cache
    if (e != cudaSuccess || e2 != cudaSuccess) return fail(RT_ERR_CUDA, std::string("closest hit: ") + cudaGetErrorString(e != cudaSuccess ? e : e2));
    return RT_OK;
}

int rt_camera_get_rays(const RtCamera* cam, const float* uv, size_t n, float* o_xyz, float* d_xyz) {
    if (!cam || !uv || !o_xyz || !d_xyz) return fail(RT_ERR_ARG, "NULL argument");
    // Camera::get_ray (camera.rs:40-46) on the host, glam operation order (scene_api.hpp V3)
    V3 origin(cam->origin[0], cam->origin[1], cam->origin[2]);
    V3 h(cam->horizontal[0], cam->horizontal[1], cam->horizontal[2]);
    V3 v(cam->vertical[0], cam->vertical[1], cam->vertical[2]);
    V3 llc(cam->lower_left_corner[0], cam->lower_left_corner[1], cam->lower_left_corner[2]);
    for (size_t i = 0; i < n; i++) {
        V3 dir = normalize(llc + uv[2 * i] * h + uv[2 * i + 1] * v - origin);
        o_xyz[3 * i] = origin.x; o_xyz[3 * i + 1] = origin.y; o_xyz[3 * i + 2] = origin.z;
        d_xyz[3 * i] = dir.x; d_xyz[3 * i + 1] = dir.y; d_xyz[3 * i + 2] = dir.z;
    }
    return RT_OK;
}

}  // extern "C"

// Launch plan of the CTA-wavefront kernel for this scene, resolved on first use and kept in the handle.
// Scenes without media whose tree has at most this many 128-byte nodes (1 MB) trace scattered and camera rays from ONE
// refilled list (kernels_cta.cuh, traverse phase).  Measured on B200 at 1080p, separate -> unified with a camera-part
// threshold of 24 idle lanes (profiles/r02_experiments/unified_todo_list_per_scene.log): cover scene (282 nodes) +2.4 %,
// textured +4.1 %, cover_defocus +2.2 %, cover_4k_motion +0.7 %; cover_100k_motion (tens of thousands of nodes: the
// coherent camera warps' node fetches are broadcasts) -1.7 %, final_scene (media set, one-loop kernel) -1.4 %.
static constexpr uint32_t UNIFIED_TODO_MAX_NODES = 8192;

static int cta_plan(RtScene* s, bool count, const CtaPlan** out) {
    CtaPlan& p = s->cta_plan[count ? 1 : 0];
    if (!p.fn) {
        // Pool size: a compile-time constant of every kernel build (kernels_cta.cuh RT_CTA_POOL_PATHS) — 1152 paths
        // (92 KB per block, 196 KB carve-out, ~60 KB of L1) = 3.6 paths per thread of the 320-thread media builds, and
        // one warp's worth less (1120) for every build that may stage Perlin table 0 in shared memory (5.5 KB; scenes
        // with Diffuse(PerlinTex) and no media), so that two blocks stay inside the same carve-out with it (measured
        // in round 1: 1152 > 1120 > 992 > 864, staging +2..8 % on those scenes).  Builds with a run-time pool size
        // (development, RT_CTA_POOL_PATHS=0) take it from the RT_CTA_PATHS knob.
        const bool media = (s->features & F_MEDIA) != 0;
        p.stage_perlin = !media && (s->class_mask >> CLS_DIFFUSE_PERLIN & 1u) && s->ds.n_perlin > 0 ? 1u : 0u;
        // Tiny scenes (<= 32 world primitives, no moving spheres): the warp-uniform primitive loop replaces the BVH
        p.flat = (s->features & F_MOTION) == 0u && s->n_world_prims <= (uint32_t)(s->knob_flat >= 0 ? s->knob_flat : 32);
        int built_paths = 0;
        void* fn = cta_kernel_for(s->features, s->knob_full_kernel ? 0x7fu : s->class_mask, count, p.flat, nullptr, &p.threads, &built_paths);
        const int dflt_paths = p.stage_perlin ? 1120 : 1152;
        p.n_paths = built_paths > 0 ? (uint32_t)built_paths
                                    : (uint32_t)std::max(32, std::min(4096, s->knob_paths > 0 ? s->knob_paths : dflt_paths)) & ~31u;
        p.refill_idle = (uint32_t)std::max(1, std::min(32, s->knob_refill));
        p.refill_camera = (uint32_t)std::max(1, std::min(32, s->knob_refill_camera));
        // One refilled list for scattered and camera rays (small trees, no media); otherwise the camera rays keep their
        // coherent warps (UNIFIED_TODO_MAX_NODES above).
        p.unified_todo = s->knob_unified >= 0 ? (uint32_t)(s->knob_unified != 0) : (!media && s->bvh.n_nodes <= UNIFIED_TODO_MAX_NODES ? 1u : 0u);
        p.smem = cta_smem_bytes(p.n_paths, p.stage_perlin != 0u);
        CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
        if (s->knob_carveout >= 0)  // development: shared-memory carve-out in percent (the rest is L1)
            CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, s->knob_carveout));
        int bps = 0;
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fn, p.threads, p.smem));
        if (bps < 1) return fail(RT_ERR_CUDA, "CTA wavefront kernel does not fit on an SM");
        p.blocks_per_sm = std::min(bps, s->knob_blocks);
        p.fn = fn;
    }
    *out = &p;
    return RT_OK;
}

extern "C" {

int rt_render_device(RtScene* s, const RtRenderParams* rp, float* accum, void* cuda_stream, RtStats* stats) {
    if (!s || !rp || !accum) return fail(RT_ERR_ARG, "NULL argument");
    if (rp->width == 0 || rp->height == 0 || rp->sample_end < rp->sample_begin) return fail(RT_ERR_ARG, "bad render params");
    if (rp->mode != RT_MODE_AUTO && rp->mode != RT_MODE_MEGAKERNEL && rp->mode != RT_MODE_CTA_WAVEFRONT) return fail(RT_ERR_ARG, "bad render mode");
    if (rp->max_depth > DEPTH_MASK) return fail(RT_ERR_ARG, "max_depth too large");
    if (rp->flags & ~(RT_RENDER_COUNTERS | RT_RENDER_RUSSIAN_ROULETTE | RT_RENDER_COSINE_SAMPLING | RT_RENDER_LOBE_SAMPLING)) return fail(RT_ERR_ARG, "unknown render flag");
    if ((uint64_t)rp->width * rp->height >= (1ull << 31)) return fail(RT_ERR_ARG, "image too large");
    auto w0 = std::chrono::steady_clock::now();
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : s->stream;
    uint32_t mode = rp->mode == RT_MODE_AUTO ? (uint32_t)s->knob_mode : rp->mode;
    const bool count = (rp->flags & RT_RENDER_COUNTERS) != 0;
    const uint32_t n_samples = rp->sample_end - rp->sample_begin;
    const uint32_t depth_word = rp->max_depth | ((rp->flags & RT_RENDER_RUSSIAN_ROULETTE) ? DEPTH_RR_FLAG : 0u) |
                                ((rp->flags & RT_RENDER_COSINE_SAMPLING) ? DEPTH_COS_FLAG : 0u) |
                                ((rp->flags & RT_RENDER_LOBE_SAMPLING) ? DEPTH_LOBE_FLAG : 0u);
    uint32_t launches = 0;
    if (n_samples) {
        CU_TRY(cudaMemsetAsync(s->scratch, 0, 18 * sizeof(unsigned long long), st));  // statistics + both work counters
        const int slot = (int)(s->n_renders++ % RtScene::KERNEL_RING);
        s->ev0 = s->ring[2 * slot];
        s->ev1 = s->ring[2 * slot + 1];
        CU_TRY(cudaEventRecord(s->ev0, st));
        if (mode == RT_MODE_CTA_WAVEFRONT) {
            const CtaPlan* plan = nullptr;
            int rc = cta_plan(s, count, &plan);
            if (rc != RT_OK) return rc;
            CtaArgs a{};
            a.accum = accum;
            a.path_counter = s->scratch + 16;
            a.stats = s->scratch;
            a.width = rp->width;
            a.height = rp->height;
            a.s_begin = rp->sample_begin;
            a.max_depth = depth_word;
            a.per_sample = rp->width * rp->height;
            a.total_paths = (unsigned long long)a.per_sample * n_samples;
            a.stage_perlin = plan->stage_perlin;
            a.n_paths = plan->n_paths;
            a.refill_idle = plan->refill_idle;
            a.refill_camera = plan->refill_camera;
            a.unified_todo = plan->unified_todo;
            a.seed = rp->seed;
            a.q_cls = s->q_cls;
            // never more blocks than there are pools of work
            const unsigned long long want = (a.total_paths + a.n_paths - 1) / a.n_paths;
            const unsigned grid = (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>(want, (unsigned long long)s->sm_count * plan->blocks_per_sm));
            void* args[2] = {(void*)&s->ds, (void*)&a};
            CU_TRY(cudaLaunchKernel(plan->fn, dim3(grid), dim3((unsigned)plan->threads), args, plan->smem, st));
        } else {
            const bool simple = s->features == 0;
            int& bps = s->mega_blocks_per_sm[count ? 1 : 0];
            if (bps == 0) {
                const void* fn = simple ? (count ? (const void*)k_render_mega<0u, true> : (const void*)k_render_mega<0u, false>)
                                        : (count ? (const void*)k_render_mega<F_ALL, true> : (const void*)k_render_mega<F_ALL, false>);
                CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fn, MEGA_THREADS, 0));
                if (bps < 1) return fail(RT_ERR_CUDA, "render kernel does not fit on an SM");
            }
            RenderArgs a{};
            a.width = rp->width;
            a.height = rp->height;
            a.s_begin = rp->sample_begin;
            a.s_end = rp->sample_end;
            a.max_depth = depth_word;
            a.tiles_x = (rp->width + 7) / 8;
            a.n_tiles = a.tiles_x * ((rp->height + 3) / 4);
            a.seed = rp->seed;
            a.accum = accum;
            a.work_counter = reinterpret_cast<unsigned int*>(s->scratch + 17);
            a.stats = s->scratch;
            const uint32_t grid = (uint32_t)(s->sm_count * bps);
            const uint32_t resident_warps = grid * (MEGA_THREADS / 32);
            // split the sample range into chunks so that there are >= ~24 work units per resident warp
            // (bounds the tail), but never chunks shorter than 8 samples
            uint32_t want_units = resident_warps * (uint32_t)s->knob_units_per_warp;
            uint32_t n_chunks = (want_units + a.n_tiles - 1) / a.n_tiles;
            n_chunks = std::max(1u, std::min(n_chunks, std::max(1u, n_samples / 8u)));
            a.chunk_spp = std::max(1u, (n_samples + n_chunks - 1) / n_chunks);
            a.n_chunks = (n_samples + a.chunk_spp - 1) / a.chunk_spp;
            if (simple) {
                if (count) k_render_mega<0u, true><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
                else k_render_mega<0u, false><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
            } else {
                if (count) k_render_mega<F_ALL, true><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
                else k_render_mega<F_ALL, false><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
            }
            CU_TRY(cudaGetLastError());
            mode = RT_MODE_MEGAKERNEL;
        }
        launches = 1;
        CU_TRY(cudaEventRecord(s->ev1, st));
        CU_TRY(cudaEventRecord(s->ev_last, st));
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        unsigned long long h[16] = {0};
        if (n_samples) {
            CU_TRY(cudaMemcpyAsync(h, s->scratch, sizeof(h), cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaStreamSynchronize(st));
            float ms = 0.0f;
            CU_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
            stats->ms_kernel = ms;
        }
        stats->paths = (uint64_t)rp->width * rp->height * n_samples;
        stats->rays = h[0];
        stats->shade_calls = h[1];
        stats->node_visits = h[2];
        stats->box_tests = h[3];
        stats->prim_tests = h[4];
        stats->launches = launches;
        stats->mode_used = n_samples ? mode : (rp->mode == RT_MODE_AUTO ? (uint32_t)RT_MODE_CTA_WAVEFRONT : rp->mode);
#if RT_CTA_PROFILE
        for (int c = 0; c < 8; c++) stats->class_counts[c] = h[8 + c];  // development build ONLY: phase cycle counters ride in class_counts
#else
        if (mode == RT_MODE_CTA_WAVEFRONT)
            for (int q = 0; q < CLS_COUNT; q++) stats->class_counts[(s->q_cls >> (4 * q)) & 0xfu] += h[8 + q];  // queue -> class
#endif
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
    }
    return RT_OK;
}

}  // extern "C"

namespace rtb {
cudaError_t launch_divide(float* buf, size_t n, float divisor, cudaStream_t st) {
    k_divide<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(buf, n, divisor);
    return cudaGetLastError();
}
// Copies `bytes` of device memory to a host buffer on `st` and waits.  Page-locked destinations (cudaMallocHost /
// cudaHostRegister memory, e.g. a pinned torch tensor) are written by ONE direct DMA; pageable destinations go
// through two cached pinned staging buffers, the memcpy of chunk k-1 overlapping the DMA of chunk k.
int d2h_frame(RtScene* s, cudaStream_t st, const void* dev, void* host, size_t bytes) {
    cudaPointerAttributes pa{};
    const bool pinned = cudaPointerGetAttributes(&pa, host) == cudaSuccess && pa.type == cudaMemoryTypeHost;
    cudaGetLastError();  // (an unregistered pointer reports an error on old drivers)
    if (pinned) {
        CU_TRY(cudaMemcpyAsync(host, dev, bytes, cudaMemcpyDeviceToHost, st));
        CU_TRY(cudaStreamSynchronize(st));
        return RT_OK;
    }
    const size_t CH = 4u << 20;
    Lease stg;
    CU_TRY(stg.pinned(2 * CH));
    cudaEvent_t* ev = s->ev_stage;
    const size_t n_ch = (bytes + CH - 1) / CH;
    for (size_t k = 0; k <= n_ch; k++) {
        if (k < n_ch) {
            const size_t n = std::min(CH, bytes - k * CH);
            CU_TRY(cudaMemcpyAsync(stg.as<char>() + (k & 1) * CH, (const char*)dev + k * CH, n, cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaEventRecord(ev[k & 1], st));
        }
        if (k > 0) {
            const size_t j = k - 1, n = std::min(CH, bytes - j * CH);
            CU_TRY(cudaEventSynchronize(ev[j & 1]));
            std::memcpy((char*)host + j * CH, stg.as<char>() + (j & 1) * CH, n);
        }
    }
    return RT_OK;
}
}  // namespace rtb

extern "C" {

int rt_render(RtScene* s, const RtRenderParams* rp, float* out_rgb_mean, RtStats* stats) {
    if (!s || !rp || !out_rgb_mean) return fail(RT_ERR_ARG, "NULL argument");
    auto wall0 = std::chrono::steady_clock::now();
    CU_TRY(cudaSetDevice(s->device));
    const size_t nf = (size_t)rp->width * rp->height * 3;
    if (nf == 0) return fail(RT_ERR_ARG, "empty image");
    Lease accum;
    CU_TRY(accum.device_mem(s->device, nf * sizeof(float)));
    CU_TRY(cudaMemsetAsync(accum.p, 0, nf * sizeof(float), s->stream));
    RtStats local{};
    int rc = rt_render_device(s, rp, accum.as<float>(), s->stream, &local);
    if (rc == RT_OK) {
        const float spp = (float)(rp->spp_total ? rp->spp_total : (rp->sample_end - rp->sample_begin));
        cudaError_t e = launch_divide(accum.as<float>(), nf, spp, s->stream);
        if (e != cudaSuccess) rc = fail(RT_ERR_CUDA, std::string("k_divide: ") + cudaGetErrorString(e));
    }
    if (rc == RT_OK) rc = d2h_frame(s, s->stream, accum.p, out_rgb_mean, nf * sizeof(float));
    cudaStreamSynchronize(s->stream);  // before `accum` returns to the cache
    if (rc != RT_OK) return rc;
    local.launches += 1;
    local.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    if (stats) *stats = local;
    return RT_OK;
}

int rt_resolve_rgb8(RtScene* s, const float* rgb_mean, uint32_t width, uint32_t height, uint8_t* out_rgb8) {
    if (!s || !rgb_mean || !out_rgb8 || width == 0 || height == 0) return fail(RT_ERR_ARG, "bad argument");
    CU_TRY(cudaSetDevice(s->device));
    const size_t np = (size_t)width * height, off_out = align_up(np * 12, 256);
    Lease buf;
    CU_TRY(buf.device_mem(s->device, off_out + np * 3));
    int rc = RT_OK;
    cudaError_t e = cudaMemcpyAsync(buf.p, rgb_mean, np * 12, cudaMemcpyHostToDevice, s->stream);
    if (e == cudaSuccess) {
        k_resolve_rgb8<<<(unsigned)((np + 255) / 256), 256, 0, s->stream>>>(buf.as<float>(), width, height, buf.as<uint8_t>() + off_out);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_rgb8, buf.as<uint8_t>() + off_out, np * 3, cudaMemcpyDeviceToHost, s->stream);
    const cudaError_t e2 = cudaStreamSynchronize(s->stream);  // also before `buf` returns to the cache
    if (e != cudaSuccess || e2 != cudaSuccess) rc = fail(RT_ERR_CUDA, std::string("resolve: ") + cudaGetErrorString(e != cudaSuccess ? e : e2));
    return rc;
}

int rt_measure_fp32_peak(int device, double* out_ffma_per_s, double* out_sm_mhz_effective) {
    if (!out_ffma_per_s) return fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(device));
    int sms = 0;
    CU_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    Lease d;
    CU_TRY(d.device_mem(device, 4096));
    cudaEvent_t e0, e1;
    CU_TRY(cudaEventCreate(&e0));
    CU_TRY(cudaEventCreate(&e1));
    const int iters = 4096, blocks = sms * 8, threads = 256;
    double best = 0.0;
    cudaError_t err = cudaSuccess;
    for (int rep = 0; rep < 5 && err == cudaSuccess; rep++) {
        cudaEventRecord(e0);
        k_fp32_peak<<<blocks, threads>>>(d.as<float>(), iters, 1.0f);
        cudaEventRecord(e1);
        err = cudaEventSynchronize(e1);
        if (err != cudaSuccess) break;
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        double ffma = (double)blocks * threads * iters * 16.0 * 8.0;
        if (rep > 0) best = std::max(best, ffma / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (err != cudaSuccess) return fail(RT_ERR_CUDA, std::string("fp32 peak: ") + cudaGetErrorString(err));
    *out_ffma_per_s = best;
    if (out_sm_mhz_effective) *out_sm_mhz_effective = best / ((double)sms * 128.0) / 1e6;
    return RT_OK;
}

// ------------------------------------------------------------------------- host scene builders
struct RtHostScene {
    std::unique_ptr<FlatSceneOwner> own;
};

int rt_host_scene_build(const char* name, float aspect_ratio, const char* asset_dir, RtHostScene** out) {
    if (!name || !out) return fail(RT_ERR_ARG, "NULL argument");
    *out = nullptr;
    try {
        std::unique_ptr<RtHostScene> hs(new RtHostScene());
        hs->own = build_named_scene(name, aspect_ratio, asset_dir ? asset_dir : ".");
        *out = hs.release();
        return RT_OK;
    } catch (const std::exception& e) {
        std::string msg = e.what();
        return fail(msg.find("cannot open") != std::string::npos || msg.find("PPM") != std::string::npos ? RT_ERR_IO : RT_ERR_ARG, msg);
    }
}
const RtFlatScene* rt_host_scene_flat(const RtHostScene* hs) { return hs ? &hs->own->flat : nullptr; }
void rt_host_scene_destroy(RtHostScene* hs) { delete hs; }

}  // extern "C"
