// Development tool (CPU only): traversal statistics of the 4-wide BVH the host builder makes for a spheres-only
// scene — node visits, box tests and primitive tests per ray of a path-traced sample of the picture — so that
// builder variants can be compared without GPU time.  The traversal is the kernel's (rt_device.cuh node_visit /
// leaf_visit / giants_test: nearest hit child first, the others pushed with their entry distance, popped entries
// re-checked against the closest hit so far); materials are simplified to diffuse / mirror / glass bounces, which is
// enough for ray statistics (the GPU's RT_RENDER_COUNTERS numbers for the cover scene: 20.1 box tests, 5.25 node
// visits, 1.54 primitive tests per ray).
//
//   g++ -O2 -std=c++17 -I. tools/bvh_sim.cpp -o /tmp/bvh_sim -Lray_tracing_in_one_weekend_b200 -lrtgpu \
//       -Wl,-rpath,$PWD/ray_tracing_in_one_weekend_b200 [-DRT_BVH_...]
//   /tmp/bvh_sim sphere_scene 480 270 4
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>

#include "../ray_tracing_in_one_weekend_b200/csrc/bvh_build.cpp"

using namespace rtb;

struct V {
    float x, y, z;
};
static V operator+(V a, V b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static V operator-(V a, V b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static V operator*(V a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static float dot(V a, V b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static V norm(V a) {
    float l = std::sqrt(dot(a, a));
    return {a.x / l, a.y / l, a.z / l};
}
static V v3(const float* p) { return {p[0], p[1], p[2]}; }

static std::mt19937_64 gen(12345);
static float rnd() { return (float)(gen() >> 40) * (1.0f / 16777216.0f); }
static V in_sphere() {
    for (;;) {
        V v{2 * rnd() - 1, 2 * rnd() - 1, 2 * rnd() - 1};
        if (dot(v, v) < 1.0f) return v;
    }
}

struct WNode {  // node of width <= 16 for the model (collapsed from the builder's 4-wide tree)
    int n = 0;
    float lo[16][3], hi[16][3];
    uint32_t child[16];
};

struct RayTrace {
    std::vector<uint8_t> runs;  // node visits before each leaf visit; the last run ends the ray
    bool last_has_leaf = false;
    bool camera = false;
};
static std::vector<RayTrace> traces;

struct Counters {
    unsigned long long rays = 0, visits = 0, boxes = 0, prims = 0, pushes = 0, pops = 0, leaf_rounds = 0;
};

static bool sphere_hit(const RtPrim& p, V o, V d, float tmin, float tmax, float& t) {
    V oc = o - V{p.g[0], p.g[1], p.g[2]};
    float a = dot(d, d), hb = dot(oc, d), c = dot(oc, oc) - p.g[3] * p.g[3];
    float disc = hb * hb - a * c;
    if (disc < 0) return false;
    float s = std::sqrt(disc);
    float r = (-hb - s) / a;
    if (r < tmin || r > tmax) {
        r = (-hb + s) / a;
        if (r < tmin || r > tmax) return false;
    }
    t = r;
    return true;
}


// ---- SIMT model of the scattered-ray part of the traverse phase (kernels_cta.cuh A2) --------------------------
// One block: `warps` warps pull entries of the iteration's todo list from a shared counter; a warp runs the
// while-while loop (node steps until every busy lane is at a leaf or done, one leaf step, pop), retires finished
// lanes and refills idle ones once `thr` lanes are idle and the list still has entries.  Costs in issue slots.
struct SimtResult {
    double node_steps = 0, node_lane_visits = 0, leaf_steps = 0, leaf_lane_visits = 0, fetch_steps = 0, fetch_lanes = 0, issue = 0, phase = 0;
    double idle_done = 0, idle_leaf = 0;  // lane-slots of node steps lost to finished lanes / to lanes waiting at a leaf
};
static SimtResult simt(const std::vector<const RayTrace*>& rays, size_t per_iter, int warps, int thr, int kstep, double Cn, double Cl, double Cf,
                       size_t n_cam = 0, int mode = 1, int midround = 0) {
    SimtResult R;
    for (size_t it0 = 0; it0 + per_iter <= rays.size(); it0 += per_iter) {
        size_t next = it0, end = it0 + per_iter - (mode == 0 ? n_cam : 0);
        size_t nextp = end, endp = it0 + per_iter;  // camera part (mode 0)
        struct Lane {
            const RayTrace* r = nullptr;
            size_t seg = 0;
            int left = 0;       // node visits left in the current run
            bool at_leaf = false, done = true;
        };
        struct Warp {
            Lane l[32];
            double clock = 0;
            bool more = true, finished = false;
        };
        std::vector<Warp> ws(warps);
        auto start_seg = [](Lane& L) {
            // position on segment L.seg
            if (L.seg >= L.r->runs.size()) { L.done = true; return; }
            L.left = L.r->runs[L.seg];
            L.at_leaf = false;
            if (L.left == 0) L.at_leaf = true;  // (cannot happen: every run has >= 1 visit) 
        };
        for (;;) {
            // the warp with the smallest clock acts next
            int wi = -1;
            for (int w = 0; w < warps; w++)
                if (!ws[w].finished && (wi < 0 || ws[w].clock < ws[wi].clock)) wi = w;
            if (wi < 0) break;
            Warp& W = ws[wi];
            // refill
            int idle = 0;
            for (auto& L : W.l) idle += L.done;
            if (W.more && idle > 0) {
                int got = 0;
                for (auto& L : W.l)
                    if (L.done && next < end) {
                        L.r = rays[next++];
                        L.seg = 0;
                        L.done = false;
                        start_seg(L);
                        got++;
                    }
                if (next >= end) W.more = false;
                if (got) {
                    R.fetch_steps++;
                    R.fetch_lanes += got;
                    W.clock += Cf;
                    R.issue += Cf;
                }
            }
            int busy = 0;
            for (auto& L : W.l) busy += !L.done;
            if (!busy) {
                // A1 (mode 0): 32 neighbouring camera rays at a time, run to completion
                if (mode == 0 && nextp < endp) {
                    int got = 0;
                    for (auto& L : W.l)
                        if (nextp < endp) {
                            L.r = rays[nextp++];
                            L.seg = 0;
                            L.done = false;
                            start_seg(L);
                            got++;
                        }
                    R.fetch_steps++;
                    R.fetch_lanes += got;
                    W.clock += Cf;
                    R.issue += Cf;
                } else {
                    W.finished = true;
                    continue;
                }
            }
            // rounds until the refill condition
            for (;;) {
                // node loop: until every busy lane is at a leaf or done (or kstep steps, the if-if variant)
                int steps = 0;
                for (;;) {
                    int in_node = 0;
                    for (auto& L : W.l) in_node += (!L.done && !L.at_leaf);
                    if (!in_node) break;
                    if (kstep && steps >= kstep) break;
                    if (midround && W.more) {  // refill check inside the node loop
                        int dn = 0;
                        for (auto& L : W.l) dn += L.done;
                        if (dn >= midround) { W.clock += 3; R.issue += 3; goto refill_now; }
                    }
                    steps++;
                    R.node_steps++;
                    R.node_lane_visits += in_node;
                    W.clock += Cn;
                    R.issue += Cn;
                    for (auto& L : W.l) {
                        if (L.done) { R.idle_done++; continue; }
                        if (L.at_leaf) { R.idle_leaf++; continue; }
                        if (--L.left == 0) {
                            bool has_leaf = L.seg + 1 < L.r->runs.size() || L.r->last_has_leaf;
                            if (has_leaf) L.at_leaf = true;
                            else L.done = true;
                        }
                    }
                }
                int at_leaf = 0;
                for (auto& L : W.l) at_leaf += (!L.done && L.at_leaf);
                if (at_leaf) {
                    R.leaf_steps++;
                    R.leaf_lane_visits += at_leaf;
                    W.clock += Cl;
                    R.issue += Cl;
                    for (auto& L : W.l)
                        if (!L.done && L.at_leaf) {
                            L.seg++;
                            L.at_leaf = false;
                            if (L.seg >= L.r->runs.size()) L.done = true;
                            else start_seg(L);
                        }
                }
                busy = 0;
                for (auto& L : W.l) busy += !L.done;
                if (!busy) break;
                if (W.more && 32 - busy >= thr) break;
            }
        refill_now:;
        }
        double ph = 0;
        for (auto& W : ws) ph = std::max(ph, W.clock);
        R.phase += ph;
    }
    return R;
}

static double envd(const char* k, double d) { return getenv(k) ? atof(getenv(k)) : d; }

int main(int argc, char** argv) {
    const char* name = argc > 1 ? argv[1] : "sphere_scene";
    int W = argc > 2 ? atoi(argv[2]) : 480, H = argc > 3 ? atoi(argv[3]) : 270, spp = argc > 4 ? atoi(argv[4]) : 4;
    RtHostScene* hs = nullptr;
    if (rt_host_scene_build(name, (float)W / H, "assets", &hs) != RT_OK) {
        fprintf(stderr, "scene: %s\n", rt_last_error());
        return 1;
    }
    const RtFlatScene* fs = rt_host_scene_flat(hs);
    // giants (rtgpu.cu): static untransformed spheres whose extent is > 64 x the median, at most two, > 64 world prims
    std::vector<RtPrim> prims(fs->prims, fs->prims + fs->n_prims);
    std::vector<uint32_t> giants;
    {
        std::vector<float> ext;
        for (auto& p : prims)
            if (!(p.flags & RT_PRIM_FLAG_BOUNDARY_ONLY)) ext.push_back(p.kind == RT_PRIM_SPHERE ? 2 * p.g[3] : 1.0f);
        if (ext.size() > 64) {
            std::nth_element(ext.begin(), ext.begin() + ext.size() / 2, ext.end());
            float med = ext[ext.size() / 2];
            for (uint32_t i = 0; i < prims.size() && giants.size() < 2; i++)
                if (prims[i].kind == RT_PRIM_SPHERE && prims[i].xform == 0 && 2 * prims[i].g[3] > 64 * med) giants.push_back(i);
        }
        for (uint32_t g : giants) prims[g].flags |= RT_PRIM_FLAG_BOUNDARY_ONLY;
    }
    RtFlatScene f2 = *fs;
    f2.prims = prims.data();
    BvhBuildResult bvh;
    build_bvh4(f2, 1, bvh);
    printf("%s: %u world prims in the tree, %zu giants, %zu nodes, depth %u, sah %.3f\n", name, bvh.n_world_prims, giants.size(),
           bvh.nodes.size(), bvh.max_depth, bvh.sah_cost);

    // width-W tree for the model: collapse the 4-wide tree (the interior child with the largest area is replaced by its
    // children while the node has room); BVH_WIDTH=4 reproduces the builder's tree
    const int width = getenv("BVH_WIDTH") ? atoi(getenv("BVH_WIDTH")) : 4;
    std::vector<WNode> wn(bvh.nodes.size());
    {
        auto area = [](const float* lo, const float* hi) {
            float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
            return 2 * (dx * dy + dy * dz + dz * dx);
        };
        for (size_t i = 0; i < bvh.nodes.size(); i++) {
            const BvhNode4& b = bvh.nodes[i];
            WNode& w = wn[i];
            for (uint32_t k = 0; k < b.pad[0]; k++) {
                w.lo[w.n][0] = b.lox[k]; w.lo[w.n][1] = b.loy[k]; w.lo[w.n][2] = b.loz[k];
                w.hi[w.n][0] = b.hix[k]; w.hi[w.n][1] = b.hiy[k]; w.hi[w.n][2] = b.hiz[k];
                w.child[w.n++] = b.child[k];
            }
        }
        if (width > 4) {
            // top-down so that a node's children are already-final 4-wide nodes when pulled up (use the original copies)
            std::vector<WNode> orig = wn;
            std::vector<uint32_t> todo{0};
            size_t used = 0;
            while (!todo.empty()) {
                uint32_t id = todo.back();
                todo.pop_back();
                used++;
                WNode w = orig[id];
                for (;;) {
                    int pick = -1;
                    float best = -1;
                    for (int k = 0; k < w.n; k++)
                        if (!(w.child[k] & LEAF_FLAG)) {
                            const WNode& ch = orig[w.child[k]];
                            if (w.n - 1 + ch.n > width) continue;
                            float a = area(w.lo[k], w.hi[k]);
                            if (a > best) best = a, pick = k;
                        }
                    if (pick < 0) break;
                    const WNode ch = orig[w.child[pick]];
                    // replace slot `pick` with the first grandchild, append the rest
                    for (int g = 0; g < ch.n; g++) {
                        int slot = g == 0 ? pick : w.n++;
                        memcpy(w.lo[slot], ch.lo[g], 12);
                        memcpy(w.hi[slot], ch.hi[g], 12);
                        w.child[slot] = ch.child[g];
                    }
                }
                wn[id] = w;
                for (int k = 0; k < w.n; k++)
                    if (!(w.child[k] & LEAF_FLAG)) todo.push_back(w.child[k]);
            }
            printf("width %d: %zu nodes reachable\n", width, used);
        }
    }
    const RtCamera& cam = fs->camera;
    Counters cnt[2];  // camera rays, scattered rays
    unsigned long long hist_visits[64] = {0};
    for (int s = 0; s < spp; s++)
        for (int y = 0; y < H; y++)
            for (int x = 0; x < W; x++) {
                float u = (x + rnd()) / W, v = (y + rnd()) / H;
                V rd;
                do rd = V{2 * rnd() - 1, 2 * rnd() - 1, 0};
                while (dot(rd, rd) >= 1);
                rd = rd * cam.lens_radius;
                V off = v3(cam.u) * rd.x + v3(cam.v) * rd.y;
                V o = v3(cam.origin) + off;
                V d = norm(v3(cam.lower_left_corner) + v3(cam.horizontal) * u + v3(cam.vertical) * v - v3(cam.origin) - off);
                for (int depth = 0; depth <= 50; depth++) {
                    Counters& c = cnt[depth ? 1 : 0];
                    c.rays++;
                    float best = 3.4e38f;
                    int bp = -1;
                    for (uint32_t g : giants) {
                        float t;
                        c.prims++;
                        if (sphere_hit(prims[g], o, d, 1e-3f, best, t)) best = t, bp = (int)g;
                    }
                    // traversal
                    float inv[3] = {1 / d.x, 1 / d.y, 1 / d.z}, oo[3] = {o.x, o.y, o.z};
                    struct E {
                        uint32_t id;
                        float t;
                    } stack[64];
                    int sp = 0;
                    uint32_t cur = 0;
                    unsigned my_visits = 0, run = 0;
                    RayTrace tr;
                    tr.camera = depth == 0;
                    const uint32_t NONE = 0x7fffffffu;
                    while (cur != NONE) {
                        while (!(cur & LEAF_FLAG) && cur != NONE) {
                            const WNode& nd = wn[cur];
                            c.visits++;
                            my_visits++;
                            run++;
                            c.boxes += nd.n;
                            float e[16];
                            for (int k = 0; k < nd.n; k++) {
                                float tn = 1e-3f, tf = best;
                                for (int a = 0; a < 3; a++) {
                                    float t0 = (nd.lo[k][a] - oo[a]) * inv[a], t1 = (nd.hi[k][a] - oo[a]) * inv[a];
                                    if (inv[a] < 0) std::swap(t0, t1);
                                    tn = std::fmax(tn, t0);
                                    tf = std::fmin(tf, t1);
                                }
                                e[k] = tn <= tf ? tn : INFINITY;
                            }
                            int nk = 0;
                            for (int k = 1; k < nd.n; k++)
                                if (e[k] < e[nk]) nk = k;
                            if (e[nk] < INFINITY) {
                                for (int k = 0; k < nd.n; k++)
                                    if (k != nk && e[k] < INFINITY) stack[sp++] = {nd.child[k], e[k]}, c.pushes++;
                                cur = nd.child[nk];
                            } else {
                                cur = NONE;
                                while (sp > 0) {
                                    E t = stack[--sp];
                                    c.pops++;
                                    if (t.t <= best) {
                                        cur = t.id;
                                        break;
                                    }
                                }
                            }
                        }
                        if (cur == NONE) break;
                        c.leaf_rounds++;
                        tr.runs.push_back((uint8_t)run);
                        run = 0;
                        tr.last_has_leaf = true;
                        uint32_t first = cur & LEAF_FIRST_MASK, n = ((cur >> LEAF_COUNT_SHIFT) & 15) + 1;
                        for (uint32_t i = 0; i < n; i++) {
                            uint32_t fi = bvh.leaf_order[first + i];
                            float t;
                            c.prims++;
                            if (sphere_hit(prims[fi], o, d, 1e-3f, best, t)) best = t, bp = (int)fi;
                        }
                        cur = NONE;
                        while (sp > 0) {
                            E t = stack[--sp];
                            c.pops++;
                            if (t.t <= best) {
                                cur = t.id;
                                break;
                            }
                        }
                    }
                    hist_visits[my_visits < 63 ? my_visits : 63]++;
                    if (run) {
                        tr.runs.push_back((uint8_t)run);
                        tr.last_has_leaf = false;
                    }
                    traces.push_back(std::move(tr));
                    if (bp < 0) break;
                    const RtPrim& p = prims[bp];
                    V hp = o + d * best;
                    V n = (hp - V{p.g[0], p.g[1], p.g[2]}) * (1.0f / p.g[3]);
                    bool front = dot(d, n) < 0;
                    V nf = front ? n : n * -1.0f;
                    const RtMaterial& m = fs->materials[p.mat];
                    V nd;
                    if (m.kind == RT_MAT_METAL) {
                        V r = d - n * (2 * dot(d, n));
                        nd = r + in_sphere() * m.p[3];
                        if (dot(nd, n) <= 0) break;
                    } else if (m.kind == RT_MAT_DIELECTRIC) {
                        float ior = m.p[0], ratio = front ? 1 / ior : ior;
                        float ct = std::fmin(dot(d * -1.0f, nf), 1.0f), st = std::sqrt(1 - ct * ct);
                        float r0 = (1 - ratio) / (1 + ratio);
                        r0 *= r0;
                        float refl = r0 + (1 - r0) * std::pow(1 - ct, 5.0f);
                        if (ratio * st > 1 || refl > rnd()) nd = d - nf * (2 * dot(d, nf));
                        else {
                            V perp = (d + nf * ct) * ratio;
                            V par = nf * -std::sqrt(std::fabs(1 - dot(perp, perp)));
                            nd = perp + par;
                        }
                    } else {
                        nd = nf + norm(in_sphere());
                        if (dot(nd, nd) < 1e-12f) nd = nf;
                        if (rnd() > 0.6f && m.kind != RT_MAT_EMISSION) { /* albedo-like absorption does not end paths here */ }
                    }
                    o = hp;
                    d = norm(nd);
                }
            }
    for (int k = 0; k < 2; k++) {
        const Counters& c = cnt[k];
        printf("%-9s rays %10llu  visits/ray %6.3f  boxes/ray %6.2f  prims/ray %5.3f  leaf rounds/ray %5.3f  pushes/ray %5.2f  pops/ray %5.2f\n",
               k ? "scattered" : "camera", c.rays, (double)c.visits / c.rays, (double)c.boxes / c.rays, (double)c.prims / c.rays,
               (double)c.leaf_rounds / c.rays, (double)c.pushes / c.rays, (double)c.pops / c.rays);
    }
    unsigned long long R = cnt[0].rays + cnt[1].rays;
    printf("all       rays %10llu  visits/ray %6.3f  boxes/ray %6.2f  prims/ray %5.3f  rays/path %.3f\n", R,
           (double)(cnt[0].visits + cnt[1].visits) / R, (double)(cnt[0].boxes + cnt[1].boxes) / R, (double)(cnt[0].prims + cnt[1].prims) / R,
           (double)R / cnt[0].rays);
    {
        unsigned long long segs = 0, zero = 0, later = 0;
        for (auto& t : traces)
            for (size_t i = 0; i < t.runs.size(); i++) {
                segs++;
                if (i > 0) later++;
                if (t.runs[i] == 0) zero++;
            }
        printf("segments %llu, after a leaf %llu, zero-length runs (a popped entry that is a leaf) %llu\n", segs, later, zero);
    }
    // SIMT model of the scattered-ray traversal (argv[5] = scattered rays per block iteration)
    const double CN = envd("COST_NODE", 118), CL = envd("COST_LEAF", 70), CF = envd("COST_FETCH", 150);
    {
        std::vector<const RayTrace*> sc;
        // the shade phase emits scattered rays class by class; approximate by path order within chunks of the pool
        for (auto& t : traces)
            if (!t.camera) sc.push_back(&t);
        size_t per_iter = argc > 5 ? (size_t)atoi(argv[5]) : 670;
        printf("\nSIMT model, %zu scattered rays, %zu per block iteration, 12 warps (costs: node 118, leaf 70, fetch 150 issue slots)\n", sc.size(), per_iter);
        printf("%-28s %10s %10s %10s %10s %10s %10s %10s\n", "policy", "node lanes", "leaf lanes", "fetch lns", "idle:done", "idle:leaf", "issue/ray", "phase/ray");
        auto show = [&](const char* nm, const SimtResult& r) {
            double slots = r.node_steps * 32;
            printf("%-28s %10.2f %10.2f %10.2f %9.1f%% %9.1f%% %10.2f %10.3f\n", nm, r.node_lane_visits / r.node_steps, r.leaf_lane_visits / r.leaf_steps,
                   r.fetch_lanes / r.fetch_steps, 100 * r.idle_done / slots, 100 * r.idle_leaf / slots, r.issue / sc.size(), r.phase / sc.size());
        };
        for (int thr : {1, 4, 8, 16, 24, 32}) {
            char nm[64];
            snprintf(nm, sizeof nm, "while-while thr %d", thr);
            show(nm, simt(sc, per_iter, 12, thr, 0, CN, CL, CF));
        }
        for (int k : {1, 2, 3}) {
            char nm[64];
            snprintf(nm, sizeof nm, "if-if k=%d thr 8", k);
            show(nm, simt(sc, per_iter, 12, 8, k, CN, CL, CF));
        }
        for (size_t n : {335, 670, 1340, 2680, 100000}) {
            char nm[64];
            snprintf(nm, sizeof nm, "thr 8, %zu rays/iter", n);
            show(nm, simt(sc, n, 12, 8, 0, CN, CL, CF));
        }
    }
    // both parts of the todo list: per iteration `per_iter` scattered rays followed by the camera rays that replace ended paths
    {
        std::vector<const RayTrace*> sc, cam;
        for (auto& t : traces) (t.camera ? cam : sc).push_back(&t);
        size_t n_sc = argc > 5 ? (size_t)atoi(argv[5]) : 670;
        size_t n_cam = (size_t)((double)n_sc * cam.size() / sc.size());
        std::vector<const RayTrace*> mix;
        size_t a = 0, b = 0;
        while (a + n_sc <= sc.size() && b + n_cam <= cam.size()) {
            for (size_t i = 0; i < n_sc; i++) mix.push_back(sc[a++]);
            for (size_t i = 0; i < n_cam; i++) mix.push_back(cam[b++]);
        }
        printf("\nboth parts: %zu scattered + %zu camera rays per block iteration\n", n_sc, n_cam);
        auto show = [&](const char* nm, const SimtResult& r) {
            double slots = r.node_steps * 32;
            printf("%-28s %10.2f %10.2f %10.2f %9.1f%% %9.1f%% %10.2f %10.3f\n", nm, r.node_lane_visits / r.node_steps, r.leaf_lane_visits / r.leaf_steps,
                   r.fetch_lanes / r.fetch_steps, 100 * r.idle_done / slots, 100 * r.idle_leaf / slots, r.issue / mix.size(), r.phase / mix.size());
        };
        show("A2 refill, then A1 coherent", simt(mix, n_sc + n_cam, 12, 8, 0, CN, CL, CF, n_cam, 0));
        show("one list, refill throughout", simt(mix, n_sc + n_cam, 12, 8, 0, CN, CL, CF, n_cam, 1));
        show("one list, thr 16", simt(mix, n_sc + n_cam, 12, 16, 0, CN, CL, CF, n_cam, 1));
        show("one list, thr 24", simt(mix, n_sc + n_cam, 12, 24, 0, CN, CL, CF, n_cam, 1));
        show("one list, mid-round refill 8", simt(mix, n_sc + n_cam, 12, 8, 0, CN, CL, CF, n_cam, 1, 8));
        show("one list, mid-round refill 16", simt(mix, n_sc + n_cam, 12, 16, 0, CN, CL, CF, n_cam, 1, 16));
        show("one list, mid-round refill 4", simt(mix, n_sc + n_cam, 12, 4, 0, CN, CL, CF, n_cam, 1, 4));
    }
    rt_host_scene_destroy(hs);
    return 0;
}
