// Host-side acceleration-structure build: conservative world-space primitive bounds, binned-SAH
// top-down build emitted directly as a 4-wide BVH with 128-byte nodes (device_types.h).
//
// Replaces BvhNode::new (src/hitable.rs:177-221: random axis, object-median split, pointer-linked
// binary tree, children visited left-then-right without ordering).  The reference's tree defines
// only the primitive ORDER (tie-break rule), which the flattener already captured in the flat
// indices; the tree used for traversal is free to differ as long as culling stays conservative
// with respect to the exact primitive tests (SURVEY.md section 7 "hard parts").
#include "bvh_build.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <numeric>
#include <stdexcept>

namespace rtb {

namespace {

struct B3 {
    float lo[3] = {INFINITY, INFINITY, INFINITY};
    float hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    void grow(const B3& o) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::fmin(lo[a], o.lo[a]);
            hi[a] = std::fmax(hi[a], o.hi[a]);
        }
    }
    void grow_pt(const float* p) {
        for (int a = 0; a < 3; a++) {
            lo[a] = std::fmin(lo[a], p[a]);
            hi[a] = std::fmax(hi[a], p[a]);
        }
    }
    float area() const {
        float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
        if (!(dx >= 0) || !(dy >= 0) || !(dz >= 0)) return 0.0f;
        return 2.0f * (dx * dy + dy * dz + dz * dx);
    }
    float extent_max() const { return std::fmax(hi[0] - lo[0], std::fmax(hi[1] - lo[1], hi[2] - lo[2])); }
};

// object-space point -> world space through a chain (ops listed outermost first, so apply the
// inverse of the ray transform innermost-first: RotateY rotates points by +theta, hitable.rs:498-499)
void point_to_world(const RtFlatScene& fs, uint32_t chain, float* p) {
    const RtXformChain& ch = fs.xform_chains[chain];
    for (int k = (int)ch.n_ops - 1; k >= 0; k--) {
        const RtXformOp& op = fs.xform_ops[ch.first_op + k];
        if (op.kind == RT_XF_TRANSLATE) {
            p[0] += op.v[0];
            p[1] += op.v[1];
            p[2] += op.v[2];
        } else {
            float s = op.v[0], c = op.v[1];
            float x = c * p[0] + s * p[2];
            float z = -s * p[0] + c * p[2];
            p[0] = x;
            p[2] = z;
        }
    }
}

B3 box_to_world(const RtFlatScene& fs, uint32_t chain, const B3& b) {
    if (fs.xform_chains[chain].n_ops == 0) return b;
    B3 w;
    for (int i = 0; i < 8; i++) {
        float p[3] = {(i & 1) ? b.hi[0] : b.lo[0], (i & 2) ? b.hi[1] : b.lo[1], (i & 4) ? b.hi[2] : b.lo[2]};
        point_to_world(fs, chain, p);
        w.grow_pt(p);
    }
    return w;
}

void pad_box(B3& b, float abs_pad) {
    float maxabs = 0.0f;
    for (int a = 0; a < 3; a++) maxabs = std::fmax(maxabs, std::fmax(std::fabs(b.lo[a]), std::fabs(b.hi[a])));
    float pad = abs_pad + 1e-6f * maxabs + 1e-5f * b.extent_max();
    for (int a = 0; a < 3; a++) {
        b.lo[a] -= pad;
        b.hi[a] += pad;
    }
}

// Object-space bounds of primitive i exactly as the reference's bbox() would report them.
B3 reference_bbox(const RtFlatScene& fs, uint32_t i) {
    const RtPrim& p = fs.prims[i];
    B3 b;
    switch (p.kind) {
        case RT_PRIM_SPHERE:
            for (int a = 0; a < 3; a++) {
                b.lo[a] = p.g[a] - p.g[3];
                b.hi[a] = p.g[a] + p.g[3];
            }
            break;
        case RT_PRIM_MOVING_SPHERE:
            for (int a = 0; a < 3; a++) {
                b.lo[a] = std::fmin(p.g[a], p.g[4 + a]) - p.g[3];
                b.hi[a] = std::fmax(p.g[a], p.g[4 + a]) + p.g[3];
            }
            break;
        case RT_PRIM_XYRECT:
        case RT_PRIM_XZRECT:
        case RT_PRIM_YZRECT: {
            int thin = p.kind == RT_PRIM_XYRECT ? 2 : (p.kind == RT_PRIM_XZRECT ? 1 : 0);
            for (int a = 0; a < 3; a++) {
                b.lo[a] = p.g[a];
                b.hi[a] = p.g[3 + a];
            }
            b.lo[thin] = p.g[thin] - 0.0001f;
            b.hi[thin] = p.g[3 + thin] + 0.0001f;
            break;
        }
        case RT_PRIM_MEDIUM:
            for (uint32_t k = 0; k < p.aux1; k++) {
                B3 c = box_to_world(fs, fs.prims[p.aux0 + k].xform, reference_bbox(fs, p.aux0 + k));
                b.grow(c);
            }
            break;
        default:
            throw std::runtime_error("bvh: unknown primitive kind");
    }
    return b;
}

struct Builder {
    const std::vector<B3>& boxes;
    std::vector<float> cen;  // 3 per world prim
    std::vector<uint32_t> idx;
    std::vector<BvhNode4>& nodes;
    std::vector<uint32_t>& leaf_order;
    uint32_t leaf_max;
    uint32_t max_depth = 0, max_leaf = 0;
    double sah = 0.0;
    float root_area = 1.0f;

    struct Range {
        uint32_t begin, end;
        B3 box;
    };
    B3 bounds(uint32_t b, uint32_t e) const {
        B3 r;
        for (uint32_t i = b; i < e; i++) r.grow(boxes[idx[i]]);
        return r;
    }
    // binned SAH split of [begin,end); returns mid (begin < mid < end). Falls back to an index median.
    uint32_t split(uint32_t begin, uint32_t end, bool force_median) {
        uint32_t n = end - begin;
        if (n < 2) return begin;
        B3 cb;
        for (uint32_t i = begin; i < end; i++) cb.grow_pt(&cen[3 * idx[i]]);
        int best_axis = -1;
        float best_cost = INFINITY, best_pos = 0.0f;
#ifndef RT_BVH_BINS
#define RT_BVH_BINS 32
#endif
        constexpr int NB = RT_BVH_BINS;
        if (!force_median) {
            for (int a = 0; a < 3; a++) {
                float lo = cb.lo[a], hi = cb.hi[a];
                if (!(hi > lo)) continue;
                B3 bb[NB];
                uint32_t cnt[NB] = {0};
                float scale = (float)NB / (hi - lo);
                for (uint32_t i = begin; i < end; i++) {
                    int b = std::min(NB - 1, std::max(0, (int)((cen[3 * idx[i] + a] - lo) * scale)));
                    cnt[b]++;
                    bb[b].grow(boxes[idx[i]]);
                }
                float la[NB], ra[NB];
                uint32_t lc[NB], rc[NB];
                B3 acc;
                uint32_t c = 0;
                for (int b = 0; b < NB; b++) {
                    acc.grow(bb[b]);
                    c += cnt[b];
                    la[b] = acc.area();
                    lc[b] = c;
                }
                acc = B3();
                c = 0;
                for (int b = NB - 1; b >= 0; b--) {
                    acc.grow(bb[b]);
                    c += cnt[b];
                    ra[b] = acc.area();
                    rc[b] = c;
                }
                for (int b = 0; b < NB - 1; b++) {
                    if (lc[b] == 0 || rc[b + 1] == 0) continue;
                    float cost = la[b] * (float)lc[b] + ra[b + 1] * (float)rc[b + 1];
                    if (cost < best_cost) {
                        best_cost = cost;
                        best_axis = a;
                        best_pos = lo + (float)(b + 1) / scale;
                    }
                }
            }
        }
        if (best_axis >= 0) {
            auto mid_it = std::partition(idx.begin() + begin, idx.begin() + end,
                                         [&](uint32_t i) { return cen[3 * i + best_axis] < best_pos; });
            uint32_t mid = (uint32_t)(mid_it - idx.begin());
            if (mid > begin && mid < end) return mid;
        }
        // median on the widest centroid axis (degenerate / forced)
        int axis = 0;
        float w = -1.0f;
        for (int a = 0; a < 3; a++) {
            float e = cb.hi[a] - cb.lo[a];
            if (e > w) {
                w = e;
                axis = a;
            }
        }
        uint32_t mid = begin + n / 2;
        std::nth_element(idx.begin() + begin, idx.begin() + mid, idx.begin() + end,
                         [&](uint32_t x, uint32_t y) { return cen[3 * x + axis] < cen[3 * y + axis]; });
        return mid;
    }

    uint32_t make_leaf(const Range& r) {
        uint32_t first = (uint32_t)leaf_order.size();
        uint32_t n = r.end - r.begin;
        // inside a leaf keep ascending flat order (deterministic)
        std::sort(idx.begin() + r.begin, idx.begin() + r.end);
        for (uint32_t i = r.begin; i < r.end; i++) leaf_order.push_back(idx[i]);
        max_leaf = std::max(max_leaf, n);
        sah += (double)r.box.area() / root_area * n;
        return LEAF_FLAG | ((n - 1) << LEAF_COUNT_SHIFT) | first;
    }

    uint32_t build_node(const Range& r, uint32_t depth) {
        max_depth = std::max(max_depth, depth);
        uint32_t me = (uint32_t)nodes.size();
        nodes.emplace_back();
        std::vector<Range> kids{r};
        const bool force_median = depth > 24;
        while (kids.size() < 4) {
            // split the child with the largest area * count that still has more than one primitive
            int pick = -1;
            float best = -1.0f;
            for (size_t k = 0; k < kids.size(); k++) {
                uint32_t n = kids[k].end - kids[k].begin;
                if (n < 2) continue;
#ifdef RT_BVH_PICK_AREA
                float score = kids[k].box.area();
#else
                float score = kids[k].box.area() * (float)n;
#endif
                if (score > best) {
                    best = score;
                    pick = (int)k;
                }
            }
            if (pick < 0) break;
            Range c = kids[pick];
            uint32_t mid = split(c.begin, c.end, force_median);
            Range a{c.begin, mid, bounds(c.begin, mid)}, b{mid, c.end, bounds(mid, c.end)};
            kids[pick] = a;
            kids.insert(kids.begin() + pick + 1, b);
        }
        BvhNode4 nd;
        for (int k = 0; k < 4; k++) {
            nd.lox[k] = nd.loy[k] = nd.loz[k] = INFINITY;
            nd.hix[k] = nd.hiy[k] = nd.hiz[k] = -INFINITY;
            nd.child[k] = 0;
            nd.pad[k] = 0;
        }
        sah += (double)r.box.area() / root_area;  // one node visit
        for (size_t k = 0; k < kids.size(); k++) {
            const Range& c = kids[k];
            nd.lox[k] = c.box.lo[0]; nd.hix[k] = c.box.hi[0];
            nd.loy[k] = c.box.lo[1]; nd.hiy[k] = c.box.hi[1];
            nd.loz[k] = c.box.lo[2]; nd.hiz[k] = c.box.hi[2];
            uint32_t n = c.end - c.begin;
            nd.child[k] = (n <= leaf_max) ? make_leaf(c) : build_node(c, depth + 1);
        }
        nd.pad[0] = (uint32_t)kids.size();
        nodes[me] = nd;
        return me;
    }
};

}  // namespace

// World primitives (those without RT_PRIM_FLAG_BOUNDARY_ONLY), their reference bounds and the padded,
// conservative bounds the BVH stores.  Shared by the host SAH builder and the GPU LBVH builder (gpu_bvh.cu).
static void world_boxes(const RtFlatScene& fs, std::vector<uint32_t>& world, std::vector<B3>& ref, std::vector<B3>& boxes) {
    world.clear();
    for (uint32_t i = 0; i < fs.n_prims; i++)
        if (!(fs.prims[i].flags & RT_PRIM_FLAG_BOUNDARY_ONLY)) world.push_back(i);
    if (world.empty()) throw std::runtime_error("scene has no world primitives");

    // "compact" scene region for the sphere-test error bound: camera + boxes of primitives that are
    // not giants (extent <= 64 x median extent) — ray origins (camera, hit points) live there.
    ref.resize(world.size());
    std::vector<float> ext(world.size());
    for (size_t k = 0; k < world.size(); k++) {
        // (a medium's box is the union of its boundary boxes, which carry their full chains: already world space)
        const RtPrim& wp = fs.prims[world[k]];
        ref[k] = wp.kind == RT_PRIM_MEDIUM ? reference_bbox(fs, world[k]) : box_to_world(fs, wp.xform, reference_bbox(fs, world[k]));
        ext[k] = ref[k].extent_max();
    }
    std::vector<float> sorted_ext = ext;
    std::nth_element(sorted_ext.begin(), sorted_ext.begin() + sorted_ext.size() / 2, sorted_ext.end());
    float median_ext = sorted_ext[sorted_ext.size() / 2];
    B3 compact;
    compact.grow_pt(fs.camera.origin);
    for (size_t k = 0; k < world.size(); k++)
        if (ext[k] <= 64.0f * median_ext) compact.grow(ref[k]);

    boxes.resize(world.size());
    for (size_t k = 0; k < world.size(); k++) {
        const RtPrim& p = fs.prims[world[k]];
        B3 b = ref[k];
        float abs_pad = 0.0f;
        if (p.kind == RT_PRIM_SPHERE || p.kind == RT_PRIM_MOVING_SPHERE) {
            // The reference's float sphere test (hitable.rs:76-91) evaluates disc = half_b^2 - a*c with
            // |oc|^2-sized terms: it can report a hit for a ray that misses the exact sphere by up to
            // delta = sqrt(r^2 + E) - r, E ~ 16 ulp(|oc|^2).  Bounds must contain those hits too.
            float c[3] = {p.g[0], p.g[1], p.g[2]};
            point_to_world(fs, p.xform, c);
            float far2 = 0.0f;
            for (int corner = 0; corner < 8; corner++) {
                float d2 = 0.0f;
                for (int a = 0; a < 3; a++) {
                    float q = (corner >> a & 1) ? compact.hi[a] : compact.lo[a];
                    d2 += (q - c[a]) * (q - c[a]);
                }
                far2 = std::fmax(far2, d2);
            }
            float r = p.g[3];
            float L = std::sqrt(far2) + r;
            float E = 16.0f * 5.9604645e-8f * L * L;
            abs_pad = std::sqrt(r * r + E) - r;
        }
        pad_box(b, abs_pad);
        boxes[k] = b;
    }
}

void compute_world_boxes(const RtFlatScene& fs, std::vector<uint32_t>& world, std::vector<Box6>& ref_out, std::vector<Box6>& padded_out) {
    std::vector<B3> ref, boxes;
    world_boxes(fs, world, ref, boxes);
    ref_out.resize(world.size());
    padded_out.resize(world.size());
    for (size_t k = 0; k < world.size(); k++)
        for (int a = 0; a < 3; a++) {
            ref_out[k].lo[a] = ref[k].lo[a];
            ref_out[k].hi[a] = ref[k].hi[a];
            padded_out[k].lo[a] = boxes[k].lo[a];
            padded_out[k].hi[a] = boxes[k].hi[a];
        }
}

void build_bvh4(const RtFlatScene& fs, uint32_t leaf_max, BvhBuildResult& out) {
    out = BvhBuildResult();
    if (leaf_max < 1) leaf_max = 1;
    if (leaf_max > 16) leaf_max = 16;
    std::vector<uint32_t> world;
    std::vector<B3> ref, boxes;
    world_boxes(fs, world, ref, boxes);
    out.n_world_prims = (uint32_t)world.size();

    Builder bl{boxes, {}, {}, out.nodes, out.leaf_order, leaf_max};
    bl.cen.resize(3 * world.size());
    bl.idx.resize(world.size());
    std::iota(bl.idx.begin(), bl.idx.end(), 0u);
    for (size_t k = 0; k < world.size(); k++)
        for (int a = 0; a < 3; a++) bl.cen[3 * k + a] = 0.5f * (ref[k].lo[a] + ref[k].hi[a]);
    Builder::Range root{0, (uint32_t)world.size(), bl.bounds(0, (uint32_t)world.size())};
    bl.root_area = std::fmax(root.box.area(), 1e-30f);
    bl.build_node(root, 1);
    // leaf_order currently indexes `world`; convert to flat primitive indices
    for (auto& v : out.leaf_order) v = world[v];
    out.max_depth = bl.max_depth;
    out.max_leaf_prims = bl.max_leaf;
    out.sah_cost = (float)bl.sah;
    for (int a = 0; a < 3; a++) {
        out.root_lo[a] = root.box.lo[a];
        out.root_hi[a] = root.box.hi[a];
    }
}

}  // namespace rtb
