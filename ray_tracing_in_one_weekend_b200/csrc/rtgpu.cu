// librtgpu.so — C ABI entry points (include/rtgpu.h): scene upload, batch closest hit, render,
// resolve, FP32 peak probe.  There is no CPU fallback: without a usable CUDA device every compute
// entry point fails with RT_ERR_CUDA.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "bvh_build.hpp"
#include "gpu_bvh.hpp"
#include "kernels.cuh"
#include "kernels_cta.cuh"
#include "scene_api.hpp"

using namespace rtb;

static thread_local std::string g_err;
static int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
namespace rtb {
int fail_png(const std::string& msg) { return fail(RT_ERR_IO, msg); }  // png_writer.cpp
}
#define CU_TRY(expr)                                                                                        \
    do {                                                                                                    \
        cudaError_t e__ = (expr);                                                                           \
        if (e__ != cudaSuccess)                                                                             \
            return fail(RT_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));                 \
    } while (0)

constexpr int WF_MAX_CHAINS = 8;
constexpr int WF_GRAPH_ITERS = 8;  // wavefront iterations per graph launch (even: the count buffers alternate)

struct RtScene {
    int device = 0;
    cudaStream_t stream = nullptr;
    DeviceScene ds{};
    std::vector<void*> allocs;
    std::vector<cudaArray_t> arrays;
    std::vector<cudaTextureObject_t> texobjs;
    RtBvhInfo bvh{};
    RtCamera camera{};
    // render scratch
    unsigned int* work_counter = nullptr;
    unsigned long long* cta_path_counter = nullptr;
    unsigned long long* stats = nullptr;
    float* accum = nullptr;
    size_t accum_floats = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int mega_blocks_per_sm = 0, mega_blocks_per_sm_cnt = 0, sm_count = 0;
    uint32_t features = 0;  // F_* bits present in the scene
    uint32_t class_mask = 0;  // shading classes (CLS_*) of the scene's materials
    // CTA wavefront: shade queue -> class (4 bits per queue) and the order the shade phase drains the queues in
    uint32_t q_cls = 0, q_order = 0;
    // wavefront pool + the CUDA graph of WF_GRAPH_ITERS iterations (built on first use)
    WfArgs wf[WF_MAX_CHAINS]{};  // independent sub-pools ("chains") whose kernels overlap on the GPU
    int wf_chains = 0;
    std::vector<void*> wf_allocs;
    uint32_t* wf_host_counts = nullptr;  // pinned
    WfFrame* wf_host_frame = nullptr;    // pinned staging copy of the per-frame parameters
    cudaGraph_t wf_graph[2] = {nullptr, nullptr};  // [0] plain, [1] with traversal counters
    cudaGraphExec_t wf_exec[2] = {nullptr, nullptr};
    uint32_t wf_graph_launches = 0;  // kernel nodes per graph launch

    ~RtScene() {
        cudaSetDevice(device);
        for (auto t : texobjs) cudaDestroyTextureObject(t);
        for (auto a : arrays) cudaFreeArray(a);
        for (auto p : allocs) cudaFree(p);
        if (accum) cudaFree(accum);
        for (int g = 0; g < 2; g++) {
            if (wf_exec[g]) cudaGraphExecDestroy(wf_exec[g]);
            if (wf_graph[g]) cudaGraphDestroy(wf_graph[g]);
        }
        for (auto p : wf_allocs) cudaFree(p);
        if (wf_host_counts) cudaFreeHost(wf_host_counts);
        if (wf_host_frame) cudaFreeHost(wf_host_frame);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {

template <class T>
cudaError_t upload(RtScene* s, const T* host, size_t n, const T** dev) {
    *dev = nullptr;
    void* p = nullptr;
    size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return e;
    s->allocs.push_back(p);
    if (n) {
        e = cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) return e;
    }
    *dev = reinterpret_cast<const T*>(p);
    return cudaSuccess;
}

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
            if (p.xform != 0 || (p.flags & RT_PRIM_FLAG_BOUNDARY_ONLY)) return "medium under a transform / as a boundary is unsupported";
            if (p.aux1 == 0 || (uint64_t)p.aux0 + p.aux1 > d->n_prims) return "medium boundary out of range";
            for (uint32_t k = 0; k < p.aux1; k++) {
                const RtPrim& b = d->prims[p.aux0 + k];
                if (!(b.flags & RT_PRIM_FLAG_BOUNDARY_ONLY) || b.kind == RT_PRIM_MEDIUM) return "medium boundary must be boundary-only geometry";
            }
            if (d->materials[p.mat].kind != RT_MAT_ISOTROPIC) return "medium material must be Isotropic";
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
int env_int(const char* name, int dflt);
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
    const bool split = env_int("RT_CTA_SPLIT_QUEUES", 1) != 0;  // development knob: 0 = one queue per class
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

int env_int(const char* name, int dflt) {
    const char* v = std::getenv(name);
    return v ? std::atoi(v) : dflt;
}

}  // namespace

extern "C" {

const char* rt_last_error(void) { return g_err.c_str(); }

int rt_scene_create(const RtFlatScene* desc, int device, RtScene** out) {
    // builder policy (results are identical either way): scenes of >= 32768 world primitives are built on the
    // device (same traversal speed as the SAH tree at that size, ~20x faster build), smaller ones on the host
    uint32_t flags = 0;
    if (std::getenv("RT_SCENE_FLAGS")) {
        flags = (uint32_t)env_int("RT_SCENE_FLAGS", 0);
    } else if (desc && desc->prims && desc->n_prims >= 32768u) {
        uint32_t world = 0;
        for (uint32_t i = 0; i < desc->n_prims; i++) world += (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) ? 0u : 1u;
        if (world >= 32768u) flags = RT_SCENE_GPU_BVH;
    }
    return rt_scene_create_ex(desc, device, flags, out);
}

int rt_scene_create_ex(const RtFlatScene* desc, int device, uint32_t flags, RtScene** out) {
    if (!out) return fail(RT_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (flags & ~RT_SCENE_GPU_BVH) return fail(RT_ERR_ARG, "unknown scene flag");
    std::string why = validate(desc);
    if (!why.empty()) return fail(RT_ERR_ARG, why);
    int n_dev = 0;
    CU_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(RT_ERR_CUDA, "no such CUDA device");
    CU_TRY(cudaSetDevice(device));

    const bool timing = std::getenv("RT_TIMING") != nullptr;  // development aid: stage times on stderr
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
    CU_TRY(cudaEventCreate(&s->ev0));
    CU_TRY(cudaEventCreate(&s->ev1));
    CU_TRY(cudaDeviceGetAttribute(&s->sm_count, cudaDevAttrMultiProcessorCount, device));  // (cudaGetDeviceProperties costs milliseconds)

    stage("stream/events/attrs");
    // acceleration structure: host SAH builder (default) or the GPU LBVH builder
    BvhBuildResult bvh;
    const BvhNode4* gpu_nodes = nullptr;
    uint32_t builder = 0;
    float build_ms = 0.0f;
    auto b0 = std::chrono::steady_clock::now();
    if (flags & RT_SCENE_GPU_BVH) {
        GpuBvhResult g;
        std::string err;
        if (!gpu_build_bvh4(*desc, s->stream, g, err)) return fail(RT_ERR_CUDA, "GPU BVH build: " + err);
        if (g.max_depth * 3u + 4u <= (uint32_t)STACK_SIZE) {
            s->allocs.push_back(g.nodes);
            gpu_nodes = g.nodes;
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
            build_bvh4(*desc, (uint32_t)env_int("RT_LEAF_MAX", 1), bvh);
        } catch (const std::exception& e) {
            return fail(RT_ERR_ARG, e.what());
        }
    }
    build_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - b0).count();
    stage("bvh build");

    // device primitive array: world primitives in BVH leaf order, then boundary-only primitives in flat order
    std::vector<uint32_t> order = bvh.leaf_order;
    for (uint32_t i = 0; i < desc->n_prims; i++)
        if (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) order.push_back(i);
    std::vector<uint32_t> flat_to_dev(desc->n_prims, 0xffffffffu);
    for (uint32_t k = 0; k < order.size(); k++) flat_to_dev[order[k]] = k;
    std::vector<DevMotion> motion;
    std::vector<DevPrim> prims(order.size());
    const ShadeQueues queues = assign_shade_queues(desc);
    s->q_cls = queues.q_cls;
    s->q_order = queues.q_order;
    if (std::getenv("RT_TIMING")) std::fprintf(stderr, "[rtgpu] shade queues: class per queue %07x, drain order %07x\n", s->q_cls, s->q_order);
    bool has_media = false;
    for (uint32_t k = 0; k < order.size(); k++) {
        prims[k] = make_dev_prim(desc, order[k], k, motion, flat_to_dev, queues.mat_queue);
        has_media |= desc->prims[order[k]].kind == RT_PRIM_MEDIUM;
    }
    // one record per primitive (static ones: dc = 0, dt = 1): every kernel built with F_MOTION loads record i with
    // primitive i, and the megakernel / wavefront / batch entry points run the all-features build on any scene
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
        uint32_t flags = 0;
        if (t0 && tex_needs_uv(desc, m.tex0)) flags |= 1u;
        if (m.kind == RT_MAT_ROUGH_PLASTIC && tex_needs_uv(desc, m.tex1)) flags |= 1u;
        const uint32_t cls = material_class(desc, m);
        s->class_mask |= 1u << cls;
        flags |= cls << 8;
        mats[i] = DevMaterial{m.kind, m.tex0, m.tex1, flags, m.p[0], m.p[1], m.p[2], m.p[3]};
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
    if (gpu_nodes) s->ds.nodes = gpu_nodes;
    else CU_TRY(upload(s.get(), bvh.nodes.data(), bvh.nodes.size(), &s->ds.nodes));
    CU_TRY(upload(s.get(), prims.data(), prims.size(), &s->ds.prims));
    CU_TRY(upload(s.get(), motion.data(), motion.size(), &s->ds.motion));
    CU_TRY(upload(s.get(), ops.data(), ops.size(), &s->ds.ops));
    CU_TRY(upload(s.get(), chains.data(), chains.size(), &s->ds.chains));
    CU_TRY(upload(s.get(), mats.data(), mats.size(), &s->ds.mats));
    CU_TRY(upload(s.get(), texs.data(), texs.size(), &s->ds.texs));

    stage("table uploads");
    // Perlin tables: gradients behind a texture object (texel fetch), permutations as bytes
    {
        std::vector<float4> grad(std::max<size_t>(1, (size_t)desc->n_perlin * 256));
        std::vector<uint8_t> perm(std::max<size_t>(1, (size_t)desc->n_perlin * 768));
        for (uint32_t t = 0; t < desc->n_perlin; t++) {
            const RtPerlin& p = desc->perlin[t];
            for (int k = 0; k < 256; k++) {
                grad[t * 256 + k] = make_float4(p.rand_vec[k][0], p.rand_vec[k][1], p.rand_vec[k][2], 0.0f);
                perm[t * 768 + k] = p.perm_x[k];
                perm[t * 768 + 256 + k] = p.perm_y[k];
                perm[t * 768 + 512 + k] = p.perm_z[k];
            }
        }
        const float4* dgrad = nullptr;
        CU_TRY(upload(s.get(), grad.data(), grad.size(), &dgrad));
        CU_TRY(upload(s.get(), perm.data(), perm.size(), &s->ds.perlin_perm));
        cudaResourceDesc rd{};
        rd.resType = cudaResourceTypeLinear;
        rd.res.linear.devPtr = const_cast<float4*>(dgrad);
        rd.res.linear.desc = cudaCreateChannelDesc<float4>();
        rd.res.linear.sizeInBytes = grad.size() * sizeof(float4);
        cudaTextureDesc td{};
        td.readMode = cudaReadModeElementType;
        cudaTextureObject_t to = 0;
        CU_TRY(cudaCreateTextureObject(&to, &rd, &td, nullptr));
        s->texobjs.push_back(to);
        s->ds.perlin_grad = to;
        s->ds.perlin_grad_lin = dgrad;
        s->ds.n_perlin = desc->n_perlin;
    }
    // images: uchar4 CUDA arrays behind point-sampled, clamped, unnormalised texture objects
    {
        std::vector<DevImage> imgs(desc->n_images);
        for (uint32_t i = 0; i < desc->n_images; i++) {
            const RtImage& im = desc->images[i];
            std::vector<uchar4> rgba((size_t)im.width * im.height);
            for (size_t k = 0; k < rgba.size(); k++) rgba[k] = make_uchar4(im.rgb8[3 * k], im.rgb8[3 * k + 1], im.rgb8[3 * k + 2], 255);
            cudaChannelFormatDesc cd = cudaCreateChannelDesc<uchar4>();
            cudaArray_t arr = nullptr;
            CU_TRY(cudaMallocArray(&arr, &cd, im.width, im.height));
            s->arrays.push_back(arr);
            CU_TRY(cudaMemcpy2DToArray(arr, 0, 0, rgba.data(), (size_t)im.width * 4, (size_t)im.width * 4, im.height, cudaMemcpyHostToDevice));
            cudaResourceDesc rd{};
            rd.resType = cudaResourceTypeArray;
            rd.res.array.array = arr;
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
        CU_TRY(upload(s.get(), imgs.data(), imgs.size(), &s->ds.images));
    }
    stage("perlin + images");
    s->ds.sky = DevSky{desc->sky.kind, desc->sky.image, desc->sky.color[0], desc->sky.color[1], desc->sky.color[2]};
    const RtCamera& c = desc->camera;
    auto v3 = [](const float* p) { return V3f{p[0], p[1], p[2]}; };
    s->ds.cam = DevCamera{v3(c.origin), v3(c.horizontal), v3(c.vertical), v3(c.lower_left_corner), v3(c.u), v3(c.v), c.lens_radius, c.time0, c.time1};
    s->ds.n_world_prims = bvh.n_world_prims;
    s->ds.n_prims = (uint32_t)prims.size();
    s->ds.has_media = has_media ? 1u : 0u;
    for (uint32_t i = 0; i < desc->n_prims; i++) {
        const RtPrim& p = desc->prims[i];
        if (p.xform != 0) s->features |= F_XFORM;
        if (p.kind == RT_PRIM_MEDIUM) s->features |= F_MEDIA;
        if (p.kind == RT_PRIM_MOVING_SPHERE) s->features |= F_MOTION;
        if (p.kind == RT_PRIM_XYRECT || p.kind == RT_PRIM_XZRECT || p.kind == RT_PRIM_YZRECT) s->features |= F_RECT;
    }
    s->camera = c;
    s->bvh = RtBvhInfo{(uint32_t)bvh.nodes.size(), bvh.n_world_prims, bvh.max_depth, bvh.max_leaf_prims, bvh.sah_cost, (uint32_t)sizeof(BvhNode4), builder, build_ms};
    if (bvh.max_depth * 3u + 4u > (uint32_t)STACK_SIZE) return fail(RT_ERR_UNSUPPORTED, "BVH too deep for the traversal stack");

    void* p = nullptr;
    CU_TRY(cudaMalloc(&p, sizeof(unsigned int)));
    s->allocs.push_back(p);
    s->work_counter = (unsigned int*)p;
    CU_TRY(cudaMalloc(&p, 16 * sizeof(unsigned long long)));
    s->allocs.push_back(p);
    s->stats = (unsigned long long*)p;
    CU_TRY(cudaMalloc(&p, sizeof(unsigned long long)));
    s->allocs.push_back(p);
    s->cta_path_counter = (unsigned long long*)p;
    // the traversal stack lives in local memory: make sure L1 is preferred over shared memory
    if (s->features == 0) {
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->mega_blocks_per_sm, k_render_mega<0u, false>, MEGA_THREADS, 0));
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->mega_blocks_per_sm_cnt, k_render_mega<0u, true>, MEGA_THREADS, 0));
    } else {
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->mega_blocks_per_sm, k_render_mega<F_ALL, false>, MEGA_THREADS, 0));
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&s->mega_blocks_per_sm_cnt, k_render_mega<F_ALL, true>, MEGA_THREADS, 0));
    }
    if (s->mega_blocks_per_sm < 1 || s->mega_blocks_per_sm_cnt < 1) return fail(RT_ERR_CUDA, "render kernel does not fit on an SM");
    CU_TRY(cudaDeviceSynchronize());
    stage("scratch + occupancy + sync");
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

int rt_scene_bvh_info(const RtScene* scene, RtBvhInfo* out) {
    if (!scene || !out) return fail(RT_ERR_ARG, "NULL argument");
    *out = scene->bvh;
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
    return RT_OK;
}

int rt_closest_hit(RtScene* s, const float* o, const float* d, const float* times, size_t n, float t_min, float t_max, RtHit* out) {
    if (!s || !o || !d || !out) return fail(RT_ERR_ARG, "NULL argument");
    if (n == 0) return RT_OK;
    CU_TRY(cudaSetDevice(s->device));
    float *d_o = nullptr, *d_d = nullptr, *d_t = nullptr;
    RtHit* d_out = nullptr;
    auto cleanup = [&]() {
        cudaFree(d_o); cudaFree(d_d); cudaFree(d_t); cudaFree(d_out);
    };
    cudaError_t e;
    if ((e = cudaMalloc(&d_o, n * 12)) != cudaSuccess || (e = cudaMalloc(&d_d, n * 12)) != cudaSuccess ||
        (e = cudaMalloc(&d_out, n * sizeof(RtHit))) != cudaSuccess || (times && (e = cudaMalloc(&d_t, n * 4)) != cudaSuccess)) {
        cleanup();
        return fail(RT_ERR_CUDA, std::string("cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_o, o, n * 12, cudaMemcpyHostToDevice, s->stream);
    cudaMemcpyAsync(d_d, d, n * 12, cudaMemcpyHostToDevice, s->stream);
    if (times) cudaMemcpyAsync(d_t, times, n * 4, cudaMemcpyHostToDevice, s->stream);
    int rc = rt_closest_hit_device(s, d_o, d_d, d_t, n, t_min, t_max, d_out, s->stream);
    if (rc == RT_OK) {
        cudaMemcpyAsync(out, d_out, n * sizeof(RtHit), cudaMemcpyDeviceToHost, s->stream);
        e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = fail(RT_ERR_CUDA, std::string("closest hit: ") + cudaGetErrorString(e));
    }
    cleanup();
    return rc;
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

// The wavefront pool is split into `wf_chains` independent sub-pools.  Each chain runs its own
// traverse -> shade -> traverse ... sequence; the chains share only the path counter, the frame
// parameters, the accumulation buffer and the statistics, so the GPU overlaps one chain's kernel tails
// and launch gaps with another chain's kernels.
static uint32_t wf_chunk() { return (uint32_t)std::max(32, std::min(4096, env_int("RT_WF_CHUNK", 128))); }

static int wf_ensure_pool(RtScene* s) {
    if (s->wf_chains) return RT_OK;
    const int chains = std::max(1, std::min(WF_MAX_CHAINS, env_int("RT_WF_CHAINS", 4)));
    uint32_t P = (uint32_t)env_int("RT_WF_POOL", 1 << 20) / (uint32_t)chains;
    P = std::max(256u * 32u, (P + 8191u) & ~8191u);
    auto alloc = [&](size_t bytes, void** out) -> cudaError_t {
        cudaError_t e = cudaMalloc(out, bytes);
        if (e == cudaSuccess) s->wf_allocs.push_back(*out);
        return e;
    };
    unsigned long long* path_counter = nullptr;
    WfFrame* frame = nullptr;
    uint32_t* counters = nullptr;  // per chain: n_class[2][WF_NCLASS], trav_counter[2] (+ pad) = 32 words
    CU_TRY(alloc(sizeof(unsigned long long), (void**)&path_counter));
    CU_TRY(alloc(sizeof(WfFrame), (void**)&frame));
    CU_TRY(alloc((size_t)chains * 32 * sizeof(uint32_t), (void**)&counters));
    for (int c = 0; c < chains; c++) {
        WfArgs& w = s->wf[c];
        CU_TRY(alloc((size_t)P * 16, (void**)&w.ray_o));
        CU_TRY(alloc((size_t)P * 16, (void**)&w.ray_d));
        CU_TRY(alloc((size_t)P * 16, (void**)&w.beta));
        CU_TRY(alloc((size_t)P * 8, (void**)&w.rng));
        CU_TRY(alloc((size_t)P * 8, (void**)&w.hit));
        CU_TRY(alloc((size_t)P * 4 * CLS_COUNT, (void**)&w.queue));
        w.n_class = counters + c * 32;
        w.trav_counter = counters + c * 32 + 2 * WF_NCLASS;
        w.path_counter = path_counter;
        w.frame = frame;
        w.stats = s->stats;
        w.P = P;
    }
    CU_TRY(cudaMallocHost((void**)&s->wf_host_counts, (size_t)chains * 32 * sizeof(uint32_t)));
    CU_TRY(cudaMallocHost((void**)&s->wf_host_frame, sizeof(WfFrame)));
    s->wf_chains = chains;
    return RT_OK;
}

extern "C++" {
template <uint32_t F>
static void* wf_shade_fn(int cls) {
    switch (cls) {
        case CLS_MISS: return (void*)k_wf_shade<F, CLS_MISS>;
        case CLS_DIFFUSE_CONST: return (void*)k_wf_shade<F, CLS_DIFFUSE_CONST>;
        case CLS_METAL: return (void*)k_wf_shade<F, CLS_METAL>;
        case CLS_DIELECTRIC: return (void*)k_wf_shade<F, CLS_DIELECTRIC>;
        case CLS_DIFFUSE_PERLIN: return (void*)k_wf_shade<F, CLS_DIFFUSE_PERLIN>;
        case CLS_CLASSIC: return (void*)k_wf_shade<F, CLS_CLASSIC>;
        default: return (void*)k_wf_shade<F, CLS_PBR>;
    }
}
}  // extern "C++"

// Builds the graph of WF_GRAPH_ITERS iterations per chain: traverse -> {one shade kernel per class present} ->
// traverse ...  Kernel arguments are per-scene constants (frame parameters are read from device memory), so the
// graph is instantiated once per scene and variant.
static int wf_ensure_graph(RtScene* s, bool count) {
    const int g = count ? 1 : 0;
    if (s->wf_exec[g]) return RT_OK;
    const bool simple = s->features == 0;
    void* trav_fn = simple ? (count ? (void*)k_wf_traverse<0u, true> : (void*)k_wf_traverse<0u, false>)
                           : (count ? (void*)k_wf_traverse<F_ALL, true> : (void*)k_wf_traverse<F_ALL, false>);
    int trav_bps = 0;
    CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trav_bps, trav_fn, WF_TRAV_THREADS, 0));
    trav_bps = std::max(1, std::min(trav_bps, env_int("RT_WF_TRAV_BLOCKS", 64)));
    const uint32_t P = s->wf[0].P;
    // persistent traverse grid: what fits on the GPU, but never more warps than there are chunks of rays
    const uint32_t trav_cap = std::max(1u, P / (wf_chunk() * WF_TRAV_WARPS));
    const dim3 trav_grid(std::min<uint32_t>(trav_cap, (uint32_t)(s->sm_count * trav_bps)));
    const uint32_t shade_cap = (P + WF_SHADE_THREADS - 1) / WF_SHADE_THREADS;
    const dim3 shade_grid(std::min<uint32_t>(shade_cap, (uint32_t)(s->sm_count * env_int("RT_WF_SHADE_BLOCKS", 8))));
    const uint32_t mask = s->class_mask | (1u << CLS_MISS);

    cudaGraph_t graph = nullptr;
    CU_TRY(cudaGraphCreate(&graph, 0));
    s->wf_graph[g] = graph;
    uint32_t n_nodes = 0;
    for (int c = 0; c < s->wf_chains; c++) {
        std::vector<cudaGraphNode_t> deps;
        for (int it = 0; it < WF_GRAPH_ITERS; it++) {
            int buf = it & 1;
            void* args[3] = {(void*)&s->ds, (void*)&s->wf[c], (void*)&buf};
            cudaKernelNodeParams kp{};
            kp.func = trav_fn;
            kp.gridDim = trav_grid;
            kp.blockDim = dim3(WF_TRAV_THREADS);
            kp.sharedMemBytes = 0;
            kp.kernelParams = args;
            cudaGraphNode_t trav = nullptr;
            CU_TRY(cudaGraphAddKernelNode(&trav, graph, deps.data(), deps.size(), &kp));
            n_nodes++;
            deps.clear();
            for (int cls = 0; cls < CLS_COUNT; cls++) {
                if (!(mask & (1u << cls))) continue;
                kp.func = simple ? wf_shade_fn<0u>(cls) : wf_shade_fn<F_ALL>(cls);
                kp.gridDim = shade_grid;
                kp.blockDim = dim3(WF_SHADE_THREADS);
                cudaGraphNode_t sh = nullptr;
                CU_TRY(cudaGraphAddKernelNode(&sh, graph, &trav, 1, &kp));
                deps.push_back(sh);
                n_nodes++;
            }
        }
    }
    CU_TRY(cudaGraphInstantiate(&s->wf_exec[g], graph, 0));
    s->wf_graph_launches = n_nodes;
    return RT_OK;
}

// Wavefront frame: generate, then graph launches of WF_GRAPH_ITERS x (traverse, shade classes) per chain until
// the pools drain.  The loop needs the live count on the host, so this mode synchronises `st` internally.
static int render_wavefront(RtScene* s, const RtRenderParams* rp, float* accum, cudaStream_t st, bool count, uint32_t* launches,
                            uint32_t* iterations) {
    int rc = wf_ensure_pool(s);
    if (rc != RT_OK) return rc;
    rc = wf_ensure_graph(s, count);
    if (rc != RT_OK) return rc;
    const int chains = s->wf_chains;
    WfFrame& f = *s->wf_host_frame;
    f.accum = accum;
    f.width = rp->width;
    f.height = rp->height;
    f.s_begin = rp->sample_begin;
    f.n_samples = rp->sample_end - rp->sample_begin;
    f.max_depth = rp->max_depth | ((rp->flags & RT_RENDER_RUSSIAN_ROULETTE) ? DEPTH_RR_FLAG : 0u) |
                  ((rp->flags & RT_RENDER_COSINE_SAMPLING) ? DEPTH_COS_FLAG : 0u);
    f.refill_threshold = (uint32_t)env_int("RT_WF_REFILL", 24);
    f.chunk = wf_chunk();
    f.seed = rp->seed;
    f.total_paths = (unsigned long long)rp->width * rp->height * f.n_samples;
    CU_TRY(cudaMemcpyAsync((void*)s->wf[0].frame, &f, sizeof(WfFrame), cudaMemcpyHostToDevice, st));
    CU_TRY(cudaMemsetAsync(s->wf[0].n_class, 0, (size_t)chains * 32 * sizeof(uint32_t), st));
    CU_TRY(cudaMemsetAsync(s->wf[0].path_counter, 0, sizeof(unsigned long long), st));
    const uint32_t P = s->wf[0].P;
    for (int c = 0; c < chains; c++) k_wf_generate<<<P / 256, 256, 0, st>>>(s->ds, s->wf[c]);
    CU_TRY(cudaGetLastError());
    *launches = (uint32_t)chains;
    const unsigned long long min_iters = f.total_paths / ((unsigned long long)P * chains);  // every path needs at least one ray
    const int last_buf = (WF_GRAPH_ITERS - 1) & 1;
    uint32_t it = 0;
    for (;;) {
        CU_TRY(cudaGraphLaunch(s->wf_exec[count ? 1 : 0], st));
        *launches += s->wf_graph_launches;
        it += WF_GRAPH_ITERS;
        if (it >= min_iters) {
            // n_class[buf][7] = number of rays the last traverse of each chain traced
            CU_TRY(cudaMemcpyAsync(s->wf_host_counts, s->wf[0].n_class, (size_t)chains * 32 * sizeof(uint32_t),
                                   cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaStreamSynchronize(st));
            uint32_t live = 0;
            for (int c = 0; c < chains; c++) live += s->wf_host_counts[c * 32 + last_buf * WF_NCLASS + 7];
            if (live == 0) break;
        }
        if (it > 100000000u) return fail(RT_ERR_CUDA, "wavefront did not drain");
    }
    *iterations = it;
    return RT_OK;
}

int rt_render_device(RtScene* s, const RtRenderParams* rp, float* accum, void* cuda_stream, RtStats* stats) {
    if (!s || !rp || !accum) return fail(RT_ERR_ARG, "NULL argument");
    if (rp->width == 0 || rp->height == 0 || rp->sample_end < rp->sample_begin) return fail(RT_ERR_ARG, "bad render params");
    if (rp->mode > RT_MODE_CTA_WAVEFRONT) return fail(RT_ERR_ARG, "bad render mode");
    if (rp->max_depth > DEPTH_MASK) return fail(RT_ERR_ARG, "max_depth too large");
    uint32_t mode = rp->mode;
    if (mode == RT_MODE_AUTO) mode = (uint32_t)env_int("RT_DEFAULT_MODE", RT_MODE_CTA_WAVEFRONT);
    if (mode == RT_MODE_CTA_WAVEFRONT && rp->sample_end > rp->sample_begin) {
        auto w0 = std::chrono::steady_clock::now();
        CU_TRY(cudaSetDevice(s->device));
        cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : s->stream;
        const bool count = (rp->flags & RT_RENDER_COUNTERS) != 0;
        if ((uint64_t)rp->width * rp->height >= (1ull << 31)) return fail(RT_ERR_ARG, "image too large");
        CU_TRY(cudaMemsetAsync(s->stats, 0, 16 * sizeof(unsigned long long), st));
        CU_TRY(cudaMemsetAsync(s->cta_path_counter, 0, sizeof(unsigned long long), st));
        CtaArgs a{};
        a.accum = accum;
        a.path_counter = s->cta_path_counter;
        a.stats = s->stats;
        a.width = rp->width;
        a.height = rp->height;
        a.s_begin = rp->sample_begin;
        a.max_depth = rp->max_depth | ((rp->flags & RT_RENDER_RUSSIAN_ROULETTE) ? DEPTH_RR_FLAG : 0u) |
                  ((rp->flags & RT_RENDER_COSINE_SAMPLING) ? DEPTH_COS_FLAG : 0u);
        a.per_sample = rp->width * rp->height;
        a.total_paths = (unsigned long long)a.per_sample * (rp->sample_end - rp->sample_begin);
        // Pool size: 1152 paths (92 KB per block, 196 KB carve-out, ~60 KB of L1) = 3 paths per thread of a 384-thread
        // block, 3.6 per thread of the 320-thread blocks of the media kernels (measured: 1152 > 1120 > 992 > 864).
        // Scenes with Diffuse(PerlinTex) (and no media) stage Perlin table 0 in shared memory (4.75 KB): one warp's
        // worth of paths less keeps the block inside the same carve-out (+2..8 % on those scenes).
        const bool media = (s->features & F_MEDIA) != 0;
        a.stage_perlin = !media && (s->class_mask >> CLS_DIFFUSE_PERLIN & 1u) && s->ds.n_perlin > 0 ? 1u : 0u;
        int threads = 0;
        // (RT_CTA_FULL_KERNEL=1, development: always the build with every shading class)
        void* fn = cta_kernel_for(s->features, env_int("RT_CTA_FULL_KERNEL", 0) ? 0x7fu : s->class_mask, count, nullptr, &threads);
        const int dflt_paths = a.stage_perlin ? 1120 : 1152;
        a.n_paths = (uint32_t)std::max(32, std::min(4096, env_int("RT_CTA_PATHS", dflt_paths))) & ~31u;
        a.refill_idle = (uint32_t)std::max(1, std::min(32, env_int("RT_CTA_REFILL", 8)));
        a.seed = rp->seed;
        a.q_cls = s->q_cls;
        const size_t smem = cta_smem_bytes(a.n_paths, a.stage_perlin != 0u);
        CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        if (std::getenv("RT_CTA_CARVEOUT"))  // development: shared-memory carve-out in percent (the rest is L1)
            CU_TRY(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, env_int("RT_CTA_CARVEOUT", 100)));
        int bps = 0;
        CU_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, fn, threads, smem));
        if (bps < 1) return fail(RT_ERR_CUDA, "CTA wavefront kernel does not fit on an SM");
        bps = std::min(bps, env_int("RT_CTA_BLOCKS", 64));
        // never more blocks than there are pools of work
        const unsigned long long want = (a.total_paths + a.n_paths - 1) / a.n_paths;
        const unsigned grid = (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>(want, (unsigned long long)s->sm_count * bps));
        CU_TRY(cudaEventRecord(s->ev0, st));
        void* args[2] = {(void*)&s->ds, (void*)&a};
        CU_TRY(cudaLaunchKernel(fn, dim3(grid), dim3((unsigned)threads), args, smem, st));
        CU_TRY(cudaEventRecord(s->ev1, st));
        if (stats) {
            std::memset(stats, 0, sizeof(*stats));
            unsigned long long h[16] = {0};
            CU_TRY(cudaMemcpyAsync(h, s->stats, sizeof(h), cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaStreamSynchronize(st));
            float ms = 0.0f;
            CU_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
            stats->ms_kernel = ms;
            stats->paths = a.total_paths;
            stats->rays = h[0];
            stats->shade_calls = h[1];
            stats->node_visits = h[2];
            stats->box_tests = h[3];
            stats->prim_tests = h[4];
            stats->launches = 1;
            stats->mode_used = RT_MODE_CTA_WAVEFRONT;
#if RT_CTA_PROFILE
            for (int c = 0; c < 8; c++) stats->class_counts[c] = h[8 + c];  // development build: phase cycle counters
#else
            for (int q = 0; q < CLS_COUNT; q++) stats->class_counts[(s->q_cls >> (4 * q)) & 0xfu] += h[8 + q];  // queue -> class
#endif
            stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
        }
        return RT_OK;
    }
    if (mode == RT_MODE_WAVEFRONT && rp->sample_end > rp->sample_begin) {
        auto w0 = std::chrono::steady_clock::now();
        CU_TRY(cudaSetDevice(s->device));
        cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : s->stream;
        CU_TRY(cudaMemsetAsync(s->stats, 0, 8 * sizeof(unsigned long long), st));
        CU_TRY(cudaEventRecord(s->ev0, st));
        uint32_t launches = 0, iters = 0;
        int rc = render_wavefront(s, rp, accum, st, (rp->flags & RT_RENDER_COUNTERS) != 0, &launches, &iters);
        if (rc != RT_OK) return rc;
        CU_TRY(cudaEventRecord(s->ev1, st));
        if (stats) {
            std::memset(stats, 0, sizeof(*stats));
            unsigned long long h[8] = {0};
            CU_TRY(cudaMemcpyAsync(h, s->stats, sizeof(h), cudaMemcpyDeviceToHost, st));
            CU_TRY(cudaStreamSynchronize(st));
            float ms = 0.0f;
            CU_TRY(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
            stats->ms_kernel = ms;
            stats->paths = (uint64_t)rp->width * rp->height * (rp->sample_end - rp->sample_begin);
            stats->rays = h[0];
            stats->shade_calls = h[1];
            stats->node_visits = h[2];
            stats->box_tests = h[3];
            stats->prim_tests = h[4];
            stats->launches = launches;
            stats->mode_used = RT_MODE_WAVEFRONT;
            stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - w0).count();
        }
        return RT_OK;
    }
    auto wall0 = std::chrono::steady_clock::now();
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t st = cuda_stream ? (cudaStream_t)cuda_stream : s->stream;
    const bool count = (rp->flags & RT_RENDER_COUNTERS) != 0;
    const uint32_t n_samples = rp->sample_end - rp->sample_begin;
    RenderArgs a{};
    a.width = rp->width;
    a.height = rp->height;
    a.s_begin = rp->sample_begin;
    a.s_end = rp->sample_end;
    a.max_depth = rp->max_depth | ((rp->flags & RT_RENDER_RUSSIAN_ROULETTE) ? DEPTH_RR_FLAG : 0u) |
                  ((rp->flags & RT_RENDER_COSINE_SAMPLING) ? DEPTH_COS_FLAG : 0u);
    a.tiles_x = (rp->width + 7) / 8;
    a.n_tiles = a.tiles_x * ((rp->height + 3) / 4);
    a.seed = rp->seed;
    a.accum = accum;
    a.work_counter = s->work_counter;
    a.stats = s->stats;
    const int bps = count ? s->mega_blocks_per_sm_cnt : s->mega_blocks_per_sm;
    const uint32_t grid = (uint32_t)(s->sm_count * bps);
    const uint32_t resident_warps = grid * (MEGA_THREADS / 32);
    // split the sample range into chunks so that there are >= ~24 work units per resident warp
    // (bounds the tail), but never chunks shorter than 8 samples
    uint32_t want_units = resident_warps * (uint32_t)env_int("RT_UNITS_PER_WARP", 24);
    uint32_t n_chunks = (want_units + a.n_tiles - 1) / a.n_tiles;
    n_chunks = std::max(1u, std::min(n_chunks, std::max(1u, n_samples / 8u)));
    a.chunk_spp = (n_samples + n_chunks - 1) / n_chunks;
    if (a.chunk_spp == 0) a.chunk_spp = 1;
    a.n_chunks = n_samples ? (n_samples + a.chunk_spp - 1) / a.chunk_spp : 0;
    uint32_t launches = 0;
    if (n_samples) {
        CU_TRY(cudaMemsetAsync(s->work_counter, 0, sizeof(unsigned int), st));
        CU_TRY(cudaMemsetAsync(s->stats, 0, 8 * sizeof(unsigned long long), st));
        if (stats) CU_TRY(cudaEventRecord(s->ev0, st));
        if (s->features == 0) {
            if (count) k_render_mega<0u, true><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
            else k_render_mega<0u, false><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
        } else {
            if (count) k_render_mega<F_ALL, true><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
            else k_render_mega<F_ALL, false><<<grid, MEGA_THREADS, 0, st>>>(s->ds, a);
        }
        CU_TRY(cudaGetLastError());
        launches = 1;
        if (stats) CU_TRY(cudaEventRecord(s->ev1, st));
    }
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        unsigned long long h[8] = {0};
        if (n_samples) {
            CU_TRY(cudaMemcpyAsync(h, s->stats, sizeof(h), cudaMemcpyDeviceToHost, st));
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
        stats->mode_used = RT_MODE_MEGAKERNEL;
        stats->ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    }
    return RT_OK;
}

int rt_render(RtScene* s, const RtRenderParams* rp, float* out_rgb_mean, RtStats* stats) {
    if (!s || !rp || !out_rgb_mean) return fail(RT_ERR_ARG, "NULL argument");
    auto wall0 = std::chrono::steady_clock::now();
    CU_TRY(cudaSetDevice(s->device));
    size_t nf = (size_t)rp->width * rp->height * 3;
    if (nf == 0) return fail(RT_ERR_ARG, "empty image");
    if (s->accum_floats < nf) {
        if (s->accum) cudaFree(s->accum);
        s->accum = nullptr;
        s->accum_floats = 0;
        CU_TRY(cudaMalloc(&s->accum, nf * sizeof(float)));
        s->accum_floats = nf;
    }
    CU_TRY(cudaMemsetAsync(s->accum, 0, nf * sizeof(float), s->stream));
    RtStats local{};
    int rc = rt_render_device(s, rp, s->accum, s->stream, &local);
    if (rc != RT_OK) return rc;
    float spp = (float)(rp->spp_total ? rp->spp_total : (rp->sample_end - rp->sample_begin));
    k_divide<<<(unsigned)((nf + 255) / 256), 256, 0, s->stream>>>(s->accum, nf, spp);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaMemcpyAsync(out_rgb_mean, s->accum, nf * sizeof(float), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    local.launches += 1;
    local.ms_total = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    if (stats) *stats = local;
    return RT_OK;
}

int rt_resolve_rgb8(RtScene* s, const float* rgb_mean, uint32_t width, uint32_t height, uint8_t* out_rgb8) {
    if (!s || !rgb_mean || !out_rgb8 || width == 0 || height == 0) return fail(RT_ERR_ARG, "bad argument");
    CU_TRY(cudaSetDevice(s->device));
    size_t np = (size_t)width * height;
    float* d_in = nullptr;
    uint8_t* d_out = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc(&d_in, np * 12)) != cudaSuccess || (e = cudaMalloc(&d_out, np * 3)) != cudaSuccess) {
        cudaFree(d_in);
        return fail(RT_ERR_CUDA, std::string("cudaMalloc: ") + cudaGetErrorString(e));
    }
    cudaMemcpyAsync(d_in, rgb_mean, np * 12, cudaMemcpyHostToDevice, s->stream);
    k_resolve_rgb8<<<(unsigned)((np + 255) / 256), 256, 0, s->stream>>>(d_in, width, height, d_out);
    cudaMemcpyAsync(out_rgb8, d_out, np * 3, cudaMemcpyDeviceToHost, s->stream);
    e = cudaStreamSynchronize(s->stream);
    cudaFree(d_in);
    cudaFree(d_out);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("resolve: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_measure_fp32_peak(int device, double* out_ffma_per_s, double* out_sm_mhz_effective) {
    if (!out_ffma_per_s) return fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CU_TRY(cudaGetDeviceProperties(&prop, device));
    float* d = nullptr;
    CU_TRY(cudaMalloc(&d, 4));
    cudaEvent_t e0, e1;
    CU_TRY(cudaEventCreate(&e0));
    CU_TRY(cudaEventCreate(&e1));
    const int iters = 4096, blocks = prop.multiProcessorCount * 8, threads = 256;
    double best = 0.0;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        k_fp32_peak<<<blocks, threads>>>(d, iters, 1.0f);
        cudaEventRecord(e1);
        cudaError_t e = cudaEventSynchronize(e1);
        if (e != cudaSuccess) {
            cudaFree(d);
            return fail(RT_ERR_CUDA, std::string("fp32 peak: ") + cudaGetErrorString(e));
        }
        float ms = 0.0f;
        cudaEventElapsedTime(&ms, e0, e1);
        double ffma = (double)blocks * threads * iters * 16.0 * 8.0;
        if (rep > 0) best = std::max(best, ffma / (ms * 1e-3));
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d);
    *out_ffma_per_s = best;
    if (out_sm_mhz_effective) *out_sm_mhz_effective = best / ((double)prop.multiProcessorCount * 128.0) / 1e6;
    return RT_OK;
}

// ----------------------------------------------------------------------- test-only entry points
// (declared in tests, not in rtgpu.h: parity of the shading / texture device code against the
// oracle with identical injected random numbers)
int rt_test_shade_cases(RtScene* s, const uint32_t* mat_ids, const float* rays, const float* recs, const float* xi,
                        uint32_t xi_stride, uint32_t n, float* out) {
    if (!s || !mat_ids || !rays || !recs || !xi || !out) return fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    uint32_t* d_m = nullptr;
    float *d_r = nullptr, *d_c = nullptr, *d_x = nullptr, *d_o = nullptr;
    CU_TRY(cudaMalloc(&d_m, n * 4));
    CU_TRY(cudaMalloc(&d_r, (size_t)n * 24));
    CU_TRY(cudaMalloc(&d_c, (size_t)n * 52));
    CU_TRY(cudaMalloc(&d_x, (size_t)n * xi_stride * 4));
    CU_TRY(cudaMalloc(&d_o, (size_t)n * 56));
    cudaMemcpy(d_m, mat_ids, n * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_r, rays, (size_t)n * 24, cudaMemcpyHostToDevice);
    cudaMemcpy(d_c, recs, (size_t)n * 52, cudaMemcpyHostToDevice);
    cudaMemcpy(d_x, xi, (size_t)n * xi_stride * 4, cudaMemcpyHostToDevice);
    k_shade_cases<<<(n + 63) / 64, 64, 0, s->stream>>>(s->ds, d_m, d_r, d_c, d_x, xi_stride, n, d_o);
    cudaError_t e = cudaStreamSynchronize(s->stream);
    cudaMemcpy(out, d_o, (size_t)n * 56, cudaMemcpyDeviceToHost);
    cudaFree(d_m); cudaFree(d_r); cudaFree(d_c); cudaFree(d_x); cudaFree(d_o);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("shade cases: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_test_texture_cases(RtScene* s, uint32_t tex, const float* uv, const float* p, uint32_t n, float* out) {
    if (!s || !uv || !p || !out) return fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    float *d_uv = nullptr, *d_p = nullptr, *d_o = nullptr;
    CU_TRY(cudaMalloc(&d_uv, (size_t)n * 8));
    CU_TRY(cudaMalloc(&d_p, (size_t)n * 12));
    CU_TRY(cudaMalloc(&d_o, (size_t)n * 12));
    cudaMemcpy(d_uv, uv, (size_t)n * 8, cudaMemcpyHostToDevice);
    cudaMemcpy(d_p, p, (size_t)n * 12, cudaMemcpyHostToDevice);
    k_texture_cases<<<(n + 127) / 128, 128, 0, s->stream>>>(s->ds, tex, d_uv, d_p, n, d_o);
    cudaError_t e = cudaStreamSynchronize(s->stream);
    cudaMemcpy(out, d_o, (size_t)n * 12, cudaMemcpyDeviceToHost);
    cudaFree(d_uv); cudaFree(d_p); cudaFree(d_o);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("texture cases: ") + cudaGetErrorString(e));
    return RT_OK;
}

// Host-only (no GPU): build the BVH of a flat scene and check its structural invariants —
// every world primitive is referenced by exactly one leaf, child counts/encodings are sane, and
// every child box encloses the boxes/leaves below it.  Returns 0 when valid.
int rt_test_bvh_validate(const RtFlatScene* desc, uint32_t leaf_max, RtBvhInfo* info) {
    std::string why = validate(desc);
    if (!why.empty()) return fail(RT_ERR_ARG, why);
    BvhBuildResult bvh;
    try {
        build_bvh4(*desc, leaf_max, bvh);
    } catch (const std::exception& e) {
        return fail(RT_ERR_ARG, e.what());
    }
    std::vector<int> seen(desc->n_prims, 0);
    struct B { float lo[3], hi[3]; };
    std::vector<B> node_box(bvh.nodes.size());
    // bottom-up: children always have larger indices than their parent (pre-order emission)
    for (int n = (int)bvh.nodes.size() - 1; n >= 0; n--) {
        const BvhNode4& nd = bvh.nodes[n];
        B acc{{INFINITY, INFINITY, INFINITY}, {-INFINITY, -INFINITY, -INFINITY}};
        uint32_t nk = nd.pad[0];
        if (nk < 1 || nk > 4) return fail(RT_ERR_ARG, "bvh: bad child count");
        for (uint32_t k = 0; k < nk; k++) {
            B cb{{nd.lox[k], nd.loy[k], nd.loz[k]}, {nd.hix[k], nd.hiy[k], nd.hiz[k]}};
            uint32_t c = nd.child[k];
            if (c & LEAF_FLAG) {
                uint32_t first = c & LEAF_FIRST_MASK, cnt = ((c >> LEAF_COUNT_SHIFT) & 0xfu) + 1u;
                if (first + cnt > bvh.leaf_order.size() || cnt > leaf_max) return fail(RT_ERR_ARG, "bvh: bad leaf range");
                for (uint32_t i = 0; i < cnt; i++) {
                    uint32_t flat = bvh.leaf_order[first + i];
                    seen[flat]++;
                    const RtPrim& p = desc->prims[flat];
                    if (p.xform == 0 && (p.kind == RT_PRIM_SPHERE)) {  // untransformed spheres: exact containment check
                        for (int a = 0; a < 3; a++)
                            if (p.g[a] - p.g[3] < cb.lo[a] || p.g[a] + p.g[3] > cb.hi[a]) return fail(RT_ERR_ARG, "bvh: sphere outside its leaf box");
                    }
                }
            } else {
                if (c <= (uint32_t)n || c >= bvh.nodes.size()) return fail(RT_ERR_ARG, "bvh: bad child index");
                for (int a = 0; a < 3; a++)
                    if (node_box[c].lo[a] < cb.lo[a] || node_box[c].hi[a] > cb.hi[a]) return fail(RT_ERR_ARG, "bvh: child box not enclosed");
            }
            for (int a = 0; a < 3; a++) {
                acc.lo[a] = std::fmin(acc.lo[a], cb.lo[a]);
                acc.hi[a] = std::fmax(acc.hi[a], cb.hi[a]);
            }
        }
        node_box[n] = acc;
    }
    for (uint32_t i = 0; i < desc->n_prims; i++) {
        int want = (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) ? 0 : 1;
        if (seen[i] != want) return fail(RT_ERR_ARG, "bvh: primitive referenced " + std::to_string(seen[i]) + " times");
    }
    if (info) *info = RtBvhInfo{(uint32_t)bvh.nodes.size(), bvh.n_world_prims, bvh.max_depth, bvh.max_leaf_prims, bvh.sah_cost, (uint32_t)sizeof(BvhNode4), 0u, 0.0f};
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
