// TEST-ONLY entry points (librtgpu_test.so = the product objects + this file; librtgpu.so does not contain them).
// Parity of the device shading / texture / medium code against the oracle with IDENTICAL injected random numbers,
// and the host BVH builder's structural invariants.  Declared in tests (ray_tracing_in_one_weekend_b200/api.py
// load_test_library), not in include/rtgpu.h.
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "bvh_build.hpp"
#include "rt_internal.hpp"

#define RT_TEST_API extern "C" __attribute__((visibility("default")))

namespace rtb {

// Shade parity: replay emitted + scatter for `n` (ray, record, material) cases with injected random
// numbers (xi_stride per case), so CUDA materials and the oracle consume identical draws.
// rec_in per case: p[3] norm[3] tang[3] uv[2] t front_face (13 floats); out per case: emitted[3]
// attenuation[3] scat_o[3] scat_d[3] ok consumed (14 floats).
__global__ void k_shade_cases(DeviceScene sc, const uint32_t* __restrict__ mat_ids, const float* __restrict__ rays,
                              const float* __restrict__ recs, const float* __restrict__ xi, uint32_t xi_stride, uint32_t n,
                              float* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r;
    r.o = mk3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    r.d = mk3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
    r.tm = 0.0f;
    const float* q = recs + 13 * (size_t)i;
    HitRec rec;
    rec.p = mk3(q[0], q[1], q[2]);
    rec.norm = mk3(q[3], q[4], q[5]);
    rec.tang = mk3(q[6], q[7], q[8]);
    rec.u = q[9];
    rec.v = q[10];
    rec.t = q[11];
    rec.front_face = q[12] != 0.0f;
    rec.mat = mat_ids[i];
    ArrayRng rng{xi + (size_t)i * xi_stride, xi_stride, 0};
    DevMaterial m = load_mat(sc, rec.mat);
    V3f em, att;
    Ray sc_ray;
    bool ok = shade(sc, m, r, rec, rng, em, att, sc_ray);
    float* o = out + 14 * (size_t)i;
    o[0] = em.x; o[1] = em.y; o[2] = em.z;
    o[3] = att.x; o[4] = att.y; o[5] = att.z;
    o[6] = sc_ray.o.x; o[7] = sc_ray.o.y; o[8] = sc_ray.o.z;
    o[9] = sc_ray.d.x; o[10] = sc_ray.d.y; o[11] = sc_ray.d.z;
    o[12] = ok ? 1.0f : 0.0f;
    o[13] = (float)rng.i;
}

// Texture / sky parity: value(uv, p) of texture `tex` (or sky(d) when tex == 0xffffffff, d = p).
__global__ void k_texture_cases(DeviceScene sc, uint32_t tex, const float* __restrict__ uv, const float* __restrict__ p,
                                uint32_t n, float* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3f pt = mk3(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
    V3f c = tex == 0xffffffffu ? sky_color(sc, pt) : texture_value(sc, tex, uv[2 * i], uv[2 * i + 1], pt);
    out[3 * i] = c.x;
    out[3 * i + 1] = c.y;
    out[3 * i + 2] = c.z;
}


// A generator that returns the same injected number on every draw: a ray that meets several media draws once per
// medium in VISITATION order (hitable.rs:564), which differs between the reference's list and a BVH; with a
// constant sequence the closest hit does not depend on that order.
struct ConstRng {
    float v;
    uint32_t n;
    RT_DEV float next() {
        n++;
        return v;
    }
};

// ConstantMedium::hit (hitable.rs:536-579) of the medium with flat index `flat_prim`, one injected number per case.
// out per case (11 floats): hit, t, draws consumed, p[3], norm[3], front_face, 0
__global__ void k_medium_cases(DeviceScene sc, uint32_t flat_prim, const float* __restrict__ rays, const float* __restrict__ tmm,
                               const float* __restrict__ xi, uint32_t n, float* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t dev = 0xffffffffu;
    for (uint32_t k = 0; k < sc.n_prims; k++)
        if (sc.prims[k].flat == flat_prim) dev = k;
    float* o = out + 11 * (size_t)i;
    for (int k = 0; k < 11; k++) o[k] = 0.0f;
    if (dev == 0xffffffffu) return;
    Ray r;
    r.o = mk3(rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]);
    r.d = mk3(rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]);
    r.tm = 0.0f;
    ConstRng rng{xi[i], 0u};
    const DevPrim p = load_prim(sc.prims, dev);
    float t = 0.0f;
    const bool hit = medium_test<F_ALL>(sc, p, r, tmm[2 * i], tmm[2 * i + 1], rng, t);
    o[0] = hit ? 1.0f : 0.0f;
    o[2] = (float)rng.n;
    if (hit) {
        const HitRec rec = fill_hit<F_ALL>(sc, r, dev, t, true);
        o[1] = rec.t;
        o[3] = rec.p.x; o[4] = rec.p.y; o[5] = rec.p.z;
        o[6] = rec.norm.x; o[7] = rec.norm.y; o[8] = rec.norm.z;
        o[9] = rec.front_face ? 1.0f : 0.0f;
    }
}

// world.hit with the media INCLUDED (main.rs:44), the same injected number for every draw of ray i.
__global__ void k_closest_hit_xi(DeviceScene sc, const float* __restrict__ o_xyz, const float* __restrict__ d_xyz, const float* __restrict__ xi,
                                 uint32_t n, float t_min, float t_max, RtHit* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r;
    r.o = mk3(o_xyz[3 * i], o_xyz[3 * i + 1], o_xyz[3 * i + 2]);
    r.d = mk3(d_xyz[3 * i], d_xyz[3 * i + 1], d_xyz[3 * i + 2]);
    r.tm = 0.0f;
    ConstRng rng{xi[i], 0u};
    ClosestHit h = closest_hit<F_ALL, false>(sc, r, t_min, t_max, rng, /*skip_media=*/false, nullptr);
    RtHit o;
    o.prim = -1;
    o.t = 0.0f;
    o.p[0] = o.p[1] = o.p[2] = 0.0f;
    o.n[0] = o.n[1] = o.n[2] = 0.0f;
    o.uv[0] = o.uv[1] = 0.0f;
    o.front_face = 0u;
    if (h.prim != 0xffffffffu) {
        HitRec rec = fill_hit<F_ALL>(sc, r, h.prim, h.t, true);
        o.prim = (int32_t)h.flat;
        o.t = rec.t;
        o.p[0] = rec.p.x; o.p[1] = rec.p.y; o.p[2] = rec.p.z;
        o.n[0] = rec.norm.x; o.n[1] = rec.norm.y; o.n[2] = rec.norm.z;
        o.uv[0] = rec.u; o.uv[1] = rec.v;
        o.front_face = rec.front_face ? 1u : 0u;
    }
    out[i] = o;
}

}  // namespace rtb

using namespace rtb;

namespace {
// Runs `launch(device pointers...)` over host arrays: every input is copied in, every output copied back.
struct DevArrays {
    std::vector<void*> ptrs;
    ~DevArrays() {
        for (void* p : ptrs) cudaFree(p);
    }
    cudaError_t in(const void* host, size_t bytes, void** dev) {
        cudaError_t e = cudaMalloc(dev, bytes ? bytes : 4);
        if (e != cudaSuccess) return e;
        ptrs.push_back(*dev);
        return bytes ? cudaMemcpy(*dev, host, bytes, cudaMemcpyHostToDevice) : cudaSuccess;
    }
};
}  // namespace

RT_TEST_API int rt_test_shade_cases(RtScene* s, const uint32_t* mat_ids, const float* rays, const float* recs, const float* xi,
                                    uint32_t xi_stride, uint32_t n, float* out) {
    if (!s || !mat_ids || !rays || !recs || !xi || !out) return rt_fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    DevArrays a;
    void *d_m, *d_r, *d_c, *d_x, *d_o;
    CU_TRY(a.in(mat_ids, (size_t)n * 4, &d_m));
    CU_TRY(a.in(rays, (size_t)n * 24, &d_r));
    CU_TRY(a.in(recs, (size_t)n * 52, &d_c));
    CU_TRY(a.in(xi, (size_t)n * xi_stride * 4, &d_x));
    CU_TRY(cudaMalloc(&d_o, (size_t)n * 56 + 4));
    a.ptrs.push_back(d_o);
    k_shade_cases<<<(n + 63) / 64, 64, 0, s->stream>>>(s->ds, (const uint32_t*)d_m, (const float*)d_r, (const float*)d_c, (const float*)d_x, xi_stride, n, (float*)d_o);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaMemcpy(out, d_o, (size_t)n * 56, cudaMemcpyDeviceToHost));
    return RT_OK;
}

RT_TEST_API int rt_test_texture_cases(RtScene* s, uint32_t tex, const float* uv, const float* p, uint32_t n, float* out) {
    if (!s || !uv || !p || !out) return rt_fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    DevArrays a;
    void *d_uv, *d_p, *d_o;
    CU_TRY(a.in(uv, (size_t)n * 8, &d_uv));
    CU_TRY(a.in(p, (size_t)n * 12, &d_p));
    CU_TRY(cudaMalloc(&d_o, (size_t)n * 12 + 4));
    a.ptrs.push_back(d_o);
    k_texture_cases<<<(n + 127) / 128, 128, 0, s->stream>>>(s->ds, tex, (const float*)d_uv, (const float*)d_p, n, (float*)d_o);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaMemcpy(out, d_o, (size_t)n * 12, cudaMemcpyDeviceToHost));
    return RT_OK;
}

RT_TEST_API int rt_test_medium_cases(RtScene* s, uint32_t flat_prim, const float* rays, const float* tmin_tmax, const float* xi, uint32_t n,
                                     float* out) {
    if (!s || !rays || !tmin_tmax || !xi || !out) return rt_fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    DevArrays a;
    void *d_r, *d_t, *d_x, *d_o;
    CU_TRY(a.in(rays, (size_t)n * 24, &d_r));
    CU_TRY(a.in(tmin_tmax, (size_t)n * 8, &d_t));
    CU_TRY(a.in(xi, (size_t)n * 4, &d_x));
    CU_TRY(cudaMalloc(&d_o, (size_t)n * 44 + 4));
    a.ptrs.push_back(d_o);
    k_medium_cases<<<(n + 63) / 64, 64, 0, s->stream>>>(s->ds, flat_prim, (const float*)d_r, (const float*)d_t, (const float*)d_x, n, (float*)d_o);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaMemcpy(out, d_o, (size_t)n * 44, cudaMemcpyDeviceToHost));
    return RT_OK;
}

RT_TEST_API int rt_test_closest_hit_xi(RtScene* s, const float* o, const float* d, const float* xi, uint32_t n, float t_min, float t_max, RtHit* out) {
    if (!s || !o || !d || !xi || !out) return rt_fail(RT_ERR_ARG, "NULL argument");
    CU_TRY(cudaSetDevice(s->device));
    DevArrays a;
    void *d_o, *d_d, *d_x, *d_h;
    CU_TRY(a.in(o, (size_t)n * 12, &d_o));
    CU_TRY(a.in(d, (size_t)n * 12, &d_d));
    CU_TRY(a.in(xi, (size_t)n * 4, &d_x));
    CU_TRY(cudaMalloc(&d_h, (size_t)n * sizeof(RtHit) + 4));
    a.ptrs.push_back(d_h);
    k_closest_hit_xi<<<(n + 127) / 128, 128, 0, s->stream>>>(s->ds, (const float*)d_o, (const float*)d_d, (const float*)d_x, n, t_min, t_max, (RtHit*)d_h);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaMemcpy(out, d_h, (size_t)n * sizeof(RtHit), cudaMemcpyDeviceToHost));
    return RT_OK;
}

// operator/(V3f, float) (rt_device.cuh: three quotients through one refined reciprocal) against the IEEE division, for
// EVERY mantissa of the divisor at one exponent: thread m divides every consecutive triple of `a_vals` by
// s = +-(1.m) * 2^(exp_s - 127) both ways and counts the components whose bits differ.
__global__ void k_div3_sweep(const float* __restrict__ a_vals, uint32_t n_a, uint32_t exp_s, uint32_t negate, unsigned long long* mism) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;
    if (m >= (1u << 23)) return;
    const float s = __uint_as_float((negate ? 0x80000000u : 0u) | (exp_s << 23) | m);
    uint32_t bad = 0;
    for (uint32_t i = 0; i + 2 < n_a; i++) {
        const V3f a = mk3(a_vals[i], a_vals[i + 1], a_vals[i + 2]);
        const V3f q = a / s;
        bad += __float_as_uint(q.x) != __float_as_uint(__fdiv_rn(a.x, s));
        bad += __float_as_uint(q.y) != __float_as_uint(__fdiv_rn(a.y, s));
        bad += __float_as_uint(q.z) != __float_as_uint(__fdiv_rn(a.z, s));
    }
    if (bad) atomicAdd(mism, (unsigned long long)bad);
}

RT_TEST_API int rt_test_div3_sweep(const float* a_vals, uint32_t n_a, uint32_t exp_s, uint32_t negate, unsigned long long* out_mismatches) {
    if (!a_vals || n_a < 3 || n_a > 4096 || exp_s < 1 || exp_s > 254 || !out_mismatches) return rt_fail(RT_ERR_ARG, "rt_test_div3_sweep: bad argument");
    float* d_a = nullptr;
    unsigned long long* d_m = nullptr;
    CU_TRY(cudaMalloc(&d_a, n_a * sizeof(float)));
    CU_TRY(cudaMalloc(&d_m, sizeof(unsigned long long)));
    CU_TRY(cudaMemcpy(d_a, a_vals, n_a * sizeof(float), cudaMemcpyHostToDevice));
    CU_TRY(cudaMemset(d_m, 0, sizeof(unsigned long long)));
    k_div3_sweep<<<(1u << 23) / 256, 256>>>(d_a, n_a, exp_s, negate, d_m);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaMemcpy(out_mismatches, d_m, sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    cudaFree(d_a);
    cudaFree(d_m);
    return RT_OK;
}

// Host-only (no GPU): build the BVH of a flat scene and check its structural invariants —
// every world primitive is referenced by exactly one leaf, child counts/encodings are sane, and
// every child box encloses the boxes/leaves below it.  Returns 0 when valid.
RT_TEST_API int rt_test_bvh_validate(const RtFlatScene* desc, uint32_t leaf_max, RtBvhInfo* info) {
    int rc = rt_flat_validate(desc);
    if (rc != RT_OK) return rc;
    BvhBuildResult bvh;
    try {
        build_bvh4(*desc, leaf_max, bvh);
    } catch (const std::exception& e) {
        return rt_fail(RT_ERR_ARG, e.what());
    }
    auto fail = [](int c, const std::string& m) { return rt_fail(c, m); };
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
    // Leaf boxes of media and of primitives under wrappers: every corner of the (boundary) geometry, carried to world
    // space with this hook's own transform code (object -> world = the wrappers undone innermost first:
    // hitable.rs:413 `rec.p += offset`, 495-498 rotation by +theta), lies inside the box its leaf stores.
    {
        std::vector<uint32_t> world;
        std::vector<Box6> ref, padded;
        compute_world_boxes(*desc, world, ref, padded);
        auto to_world = [&](uint32_t chain, float* p) {
            const RtXformChain& ch = desc->xform_chains[chain];
            for (int k = (int)ch.n_ops - 1; k >= 0; k--) {
                const RtXformOp& op = desc->xform_ops[ch.first_op + k];
                if (op.kind == RT_XF_TRANSLATE) {
                    for (int a = 0; a < 3; a++) p[a] += op.v[a];
                } else {
                    const float sn = op.v[0], cs = op.v[1], x = p[0], z = p[2];
                    p[0] = cs * x + sn * z;
                    p[2] = -sn * x + cs * z;
                }
            }
        };
        auto check_geom = [&](const RtPrim& g, const Box6& box) -> bool {
            for (int corner = 0; corner < 8; corner++) {
                float p[3];
                for (int a = 0; a < 3; a++) {
                    if (g.kind == RT_PRIM_SPHERE || g.kind == RT_PRIM_MOVING_SPHERE) p[a] = g.g[a] + ((corner >> a & 1) ? 0.577f : -0.577f) * g.g[3];
                    else p[a] = (corner >> a & 1) ? g.g[3 + a] : g.g[a];
                }
                to_world(g.xform, p);
                for (int a = 0; a < 3; a++)
                    if (p[a] < box.lo[a] - 1e-3f || p[a] > box.hi[a] + 1e-3f) return false;
            }
            return true;
        };
        for (size_t k = 0; k < world.size(); k++) {
            const RtPrim& p = desc->prims[world[k]];
            if (p.kind == RT_PRIM_MEDIUM) {
                for (uint32_t j = 0; j < p.aux1; j++)
                    if (!check_geom(desc->prims[p.aux0 + j], padded[k])) return fail(RT_ERR_ARG, "bvh: medium boundary outside the medium's box");
            } else if (!check_geom(p, padded[k])) {
                return fail(RT_ERR_ARG, "bvh: primitive outside its box");
            }
        }
    }
    for (uint32_t i = 0; i < desc->n_prims; i++) {
        int want = (desc->prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY) ? 0 : 1;
        if (seen[i] != want) return fail(RT_ERR_ARG, "bvh: primitive referenced " + std::to_string(seen[i]) + " times");
    }
    if (info) *info = RtBvhInfo{(uint32_t)bvh.nodes.size(), bvh.n_world_prims, bvh.max_depth, bvh.max_leaf_prims, bvh.sah_cost, (uint32_t)sizeof(BvhNode4), 0u, 0.0f, 0ull};
    return RT_OK;
}

