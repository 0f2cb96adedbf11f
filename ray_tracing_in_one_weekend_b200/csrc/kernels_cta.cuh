// CTA-local wavefront renderer (sm_100a): the whole wavefront loop runs INSIDE one persistent kernel.
//
// Every thread block (448 threads for the sphere-only feature sets, 320 for the media sets, 384 otherwise; two blocks per SM) owns a pool of N paths in shared
// memory (N = 1152, or 1120 in the builds that may stage a Perlin table beside it: a compile-time constant of each build,
// RT_CTA_POOL_PATHS; 64 B of path state + 18 B of queue space per path = 92 KB per block out of Blackwell's 227 KB per SM)
// and iterates, separated by __syncthreads():
//
//   D  regenerate   one 64-bit atomicAdd per block and iteration reserves path indices for every path that
//                   ended; threads generate the new camera rays (main.rs:88-93) into the dead slots and append
//                   them to the TODO list in path-index order (32 neighbouring pixels of a row per warp)
//   A  traverse     A2: scattered rays first — warps pull todo entries from a shared-memory counter, one ray
//                   per lane, while-while BVH traversal (rt_device.cuh); idle lanes are refilled from the
//                   counter as soon as `refill_idle` lanes of the warp are idle.  A1: then the camera rays, a
//                   coherent warp of 32 at a time, run to completion (uniform chunks that fill the tail).
//                   Small trees without media (CtaArgs.unified_todo): the refill runs on through the camera rays, so
//                   the lanes a warp has left when the scattered part runs dry are topped up instead of finishing alone.
//                   A lane that finishes writes (t, prim) back and appends the path to one of the block's seven
//                   SHADE QUEUES in shared memory: queue 0 for a miss, else bits 9..11 of the primitive record.
//                   The host assigns the queues per scene (rtgpu.cu assign_shade_queues): one per shading class
//                   in use, the free ones to the heaviest material kinds of the mixed classes (pbr.rs, "classic")
//   C  shade        warps pull 32-entry chunks of the queues from a counter, most expensive class first
//                   (every queue range padded to a warp multiple, so a warp runs ONE material group):
//                   sky / emitted / scatter (material.rs, pbr.rs, texture.rs), radiance to the frame with
//                   red.global.add.f32, the scattered ray back into the slot and onto the todo list, or the
//                   slot onto the dead list
//
// Nothing but the scene (L1-resident), the accumulation atomics and one counter atomic per block-iteration
// touches global memory: no path-state traffic, no global queues, no kernel boundaries, no grid-wide tail
// until the frame itself drains.  Blocks are independent; the two resident blocks of an SM fill each other's
// barrier waits.  Replaces main.rs:38-60,84-97 with the same device functions and RNG streams as the other
// two renderers (a path's stream depends only on (seed, pixel, global sample index)).
// One translation unit per scene feature set instantiates the kernel (cta_inst.cu).
#pragma once
#include "rt_device.cuh"

namespace rtb {

// Block size per feature set (every translation unit exports its own through cta_kernel_f<F>): the media sets
// need ~100 registers per thread; at 384 threads x 2 blocks (80 registers) they carry 320-370 B of spills, at 320
// threads (96 registers, 20 warps per SM) 56 B — final_scene +15 %, cornell_box +7 % (B200, 1080p).  Every other
// set fits 80 registers and prefers the four extra warps (cornell_box_solid, cover_100k_motion: -1..-4 % at 320/352).
// FLAT: the traverse phase runs the warp-uniform primitive loop (flat_closest_hit) instead of the BVH traversal —
// the tiny-scene builds (cta_inst.cu -DRT_CTA_FLAT=1 -DRT_CTA_SUFFIX=_flat).  Without a traversal stack and slab
// constants even the media set fits 80 registers (36 B of spills): 384 threads, cornell_box +14 % over 320.
#ifndef RT_CTA_FLAT
#define RT_CTA_FLAT 0
#endif
#if !defined(RT_CTA_THREADS) && defined(RT_CTA_FLAT_THREADS) && RT_CTA_FLAT
#define RT_CTA_THREADS RT_CTA_FLAT_THREADS  // (development: block size of the tiny-scene builds only)
#endif
// Round 2, after the kernel lost 5 % of its instructions: the sphere-only sets (with or without motion) without the pbr.rs
// class take 448 threads
// (72 registers, 56 B of spills, 28 warps per SM) — cover scene +1.9 %, cover_4k_motion +1.4 %, cover_100k_motion
// +1.5 %, textured +0.6 %; 512 threads (64 registers, 240 B of spills) lose 2-7 %, and 448 costs the flat builds 1-2 % and
// the media set 15 % (profiles/r02_experiments/threads_448_sphere_sets.log).
#ifndef RT_CTA_THREADS
#if defined(RT_CTA_F) && ((RT_CTA_F) & 2) && !RT_CTA_FLAT
#define RT_CTA_THREADS 320
#elif defined(RT_CTA_F) && ((RT_CTA_F) & ~4) == 0 && !RT_CTA_FLAT && defined(RT_CTA_CLS_MASK) && (RT_CTA_CLS_MASK) == 0x3f
#define RT_CTA_THREADS 448  // (the builds without the pbr.rs class: with it the 72-register build spills 100-180 B, unmeasured)
#else
#define RT_CTA_THREADS 384
#endif
#endif
#ifndef RT_CTA_MIN_BLOCKS
#define RT_CTA_MIN_BLOCKS 2  // register cap 85 (80 used): fits traversal and the common shading classes
#endif
// Paths per block pool, a COMPILE-TIME constant of every build (exported next to the block size): the pool's arrays
// sit at multiples of N in shared memory, and with N in a register every path-state access carried its own uniform
// multiply/shift to find them (144 instructions of the spheres-only kernel, all in the per-ray fetch / retire / shade
// prologues).  1152 = 3 paths per thread; the sets that may stage the Perlin table (no media) give one warp's worth of
// paths to it so that two blocks stay inside the 196 KB carve-out (rtgpu.cu cta_plan).  0 = run-time (CtaArgs.n_paths,
// development sweeps: tools/build_variant.py -DRT_CTA_POOL_PATHS=0 + RT_CTA_PATHS).
#ifndef RT_CTA_POOL_PATHS
#if defined(RT_CTA_F) && ((RT_CTA_F) & 2)
#define RT_CTA_POOL_PATHS 1152
#else
#define RT_CTA_POOL_PATHS 1120
#endif
#endif
// Traverse phase: 1 = scattered rays and camera rays share ONE instance of the traversal loop.  The media sets inline
// a medium test into every leaf visit: one instance instead of two takes 2 100 instructions (34 KB) out of the
// kernel and its spills from 56 to 8 B — cornell_box +6.6 %, final_scene +4.9 % (the instruction cache was the
// issue: `no_instruction` 0.95 cycles per issue).  The other sets keep the two specialised loops (one loop:
// -2.2 % cover_100k_motion, -1.5 % simple_light_scene, +-0.3 % cover scene).
#ifndef RT_CTA_ONE_LOOP
#if defined(RT_CTA_F) && ((RT_CTA_F) & 2)
#define RT_CTA_ONE_LOOP 1
#else
#define RT_CTA_ONE_LOOP 0
#endif
#endif
// Shading classes compiled into the kernel (bit CLS_*).  The pbr.rs materials are 39 % of the spheres-only kernel's
// instructions (13.4 k -> 8.2 k without them, spills 20 -> 4 B): scenes that hold none run a build without that code
// (cta_inst.cu -DRT_CTA_CLS_MASK=0x3f -DRT_CTA_SUFFIX=_m3f) — less for the instruction cache to hold: cover scene
// +1.7 %, cover_4k_motion +1.9 %, textured +0.8 %.
#ifndef RT_CTA_CLS_MASK
#define RT_CTA_CLS_MASK 0x7fu
#endif
#ifndef RT_CTA_LEAF_WHILE
#define RT_CTA_LEAF_WHILE 1  // traverse phase: a popped stack entry that is a leaf is tested in the same round (cover_100k_motion +1.0 %, others +-0.2 %)
#endif
#ifndef RT_CTA_DECODE_TREE
#define RT_CTA_DECODE_TREE 1  // shade phase: chunk -> queue by a compare tree (0: counting compares + select chain)
#endif
#ifndef RT_CTA_PROFILE
#define RT_CTA_PROFILE 0  // development: phase / work cycle counters into stats[8..13] (class_counts of RtStats)
#endif
constexpr int CTA_THREADS = RT_CTA_THREADS;
constexpr int CTA_WARPS = CTA_THREADS / 32;
constexpr int CTA_MIN_BLOCKS = RT_CTA_MIN_BLOCKS;
constexpr uint32_t CTA_POOL_PATHS = RT_CTA_POOL_PATHS;
constexpr uint32_t CTA_NONE = 0xffffffffu;

struct CtaArgs {
    float* accum;
    unsigned long long* path_counter;  // next unassigned path index of this launch (zeroed before launch)
    unsigned long long* stats;         // [0] rays [1] shade calls [2..4] counters [8..14] rays per shading class
    unsigned long long total_paths;    // width * height * n_samples
    uint32_t width, height, s_begin, max_depth;
    uint32_t per_sample;   // width * height (< 2^31)
    uint32_t n_paths;      // N: paths per block pool (multiple of 32, <= 4096); builds with CTA_POOL_PATHS != 0 ignore it
    uint32_t refill_idle;  // a warp retires/refills once this many of its lanes are idle
    uint32_t refill_camera;  // ... the same once the refill has reached the camera rays (unified_todo)
    uint32_t unified_todo;   // 1: lane refill runs through the camera-ray part of the todo list as well
    uint32_t stage_perlin;  // 1: Perlin table 0 is copied into shared memory (scenes with Diffuse(PerlinTex))
    // Shade queues (rtgpu.cu assign_shade_queues): queue 0 = misses, queue q = bits 9..11 of the hit primitive's
    // record, numbered most expensive class first.  4 bits per queue: the shading class (CLS_*) it runs.
    uint32_t q_cls;
    uint64_t seed;
};

// Bytes of dynamic shared memory for a pool of n paths.
constexpr size_t CTA_PERLIN_BYTES = 256 * 16 + 768 * 2;  // 256 gradients + 768 permutation PAIRS (rt_device.cuh PerlinShared)
inline size_t cta_smem_bytes(uint32_t n, bool perlin) {
    return (((size_t)n * (64 + 2 * (CLS_COUNT + 2)) + 15) & ~(size_t)15) + (perlin ? CTA_PERLIN_BYTES : 0);
}

struct CtaPool {
    float4* o;       // o.xyz, time
    float4* d;       // d.xyz, bits(depth | alive << 31)
    float4* b;       // throughput rgb, bits(pixel)
    uint4* m;        // rng.lo, rng.hi, bits(t), prim
    uint16_t* list;  // [CLS_COUNT][N] path ids per shading class
    uint16_t* dead;  // [N] path ids to regenerate
    uint16_t* todo;  // [N] path ids to traverse next: scattered paths (in shading order), then the regenerated ones
    const float4* perlin;  // Perlin table 0 (256 gradients, then 768 permutation bytes) or nullptr
    uint32_t N;
};

// One queue entry of class CLS: the body of ray_color for a hit (main.rs:45-56) or a miss (main.rs:58).
template <uint32_t F, int CLS>
RT_DEV void cta_shade_item(const DeviceScene& sc, const CtaArgs& a, const CtaPool& P, uint32_t pid, uint32_t* s_ndead,
                           uint32_t* s_ntodo, uint32_t& n_shade) {
    const float4 rd = P.d[pid], bt = P.b[pid];
    Ray ray;
    ray.d = mk3(rd.x, rd.y, rd.z);
    ray.o = mk3(0.0f, 0.0f, 0.0f);
    ray.tm = 0.0f;
    V3f beta = mk3(bt.x, bt.y, bt.z);
    const uint32_t pixel = __float_as_uint(bt.w);
    uint32_t depth = __float_as_uint(rd.w) & DEPTH_MASK;  // (the slot's alive flag is bit 31)
    V3f add = mk3(0.0f, 0.0f, 0.0f);
    bool alive = false;
    PcgRng rng;
    rng.state = 0;
    if (CLS == CLS_MISS) {
        // main.rs:58; a path dropped by the unit-direction check arrives here with beta == 0
        if (beta.x != 0.0f || beta.y != 0.0f || beta.z != 0.0f) add = beta * sky_color(sc, ray.d);
    } else {
        const float4 ro = P.o[pid];
        const uint4 mm = P.m[pid];
        ray.o = mk3(ro.x, ro.y, ro.z);
        ray.tm = ro.w;
        rng.state = ((unsigned long long)mm.y << 32) | mm.x;
        n_shade++;
        DevPrim hp = load_prim(sc.prims, mm.w);
        DevMaterial m = load_mat(sc, hp.mat);
        const bool want_uv = (CLS == CLS_CLASSIC || CLS == CLS_PBR) && (m.reserved & 1u) != 0u;
        HitRec rec = fill_hit<F>(sc, ray, mm.w, __uint_as_float(mm.z), want_uv);
        V3f emitted, att;
        Ray scattered;
        bool ok = shade_cls<CLS>(sc, m, ray, rec, rng, emitted, att, scattered, sampling_of(a.max_depth), P.perlin);
        if (CLS == CLS_CLASSIC) add = beta * emitted;  // main.rs:47 (only Emission emits, material.rs:26-28)
        if (ok && (a.max_depth & DEPTH_RR_FLAG)) ok = russian_roulette(att, rng);
        if (ok) {                                      // main.rs:48-54
            beta = beta * att;
            ray = scattered;
            depth++;
            alive = depth <= (a.max_depth & DEPTH_MASK);  // main.rs:40-42
        }
    }
    if (add.x != 0.0f || add.y != 0.0f || add.z != 0.0f) {
        float* px_out = a.accum + (size_t)pixel * 3u;
        atomicAdd(px_out + 0, add.x);
        atomicAdd(px_out + 1, add.y);
        atomicAdd(px_out + 2, add.z);
    }
    if (alive) {
        P.o[pid] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.tm);
        P.d[pid] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(depth | WF_ALIVE));
        P.b[pid] = make_float4(beta.x, beta.y, beta.z, bt.w);
        uint2* st = reinterpret_cast<uint2*>(P.m + pid);
        *st = make_uint2((uint32_t)rng.state, (uint32_t)(rng.state >> 32));
        P.todo[atomicAdd(s_ntodo, 1u)] = (uint16_t)pid;
    } else {
        P.d[pid].w = __uint_as_float(0u);
        const uint32_t pos = atomicAdd(s_ndead, 1u);
        P.dead[pos] = (uint16_t)pid;
    }
}

// Lane 0 takes `n` entries from a block-wide shared-memory counter and every lane gets the old value.  ptxas rewrites
// every shared-memory atomic whose ADDRESS it can prove warp-uniform into its aggregation sequence (vote, find-leader,
// two popcounts, lane masks, shuffle: 17 instructions), also when ONE lane executes it and also through inline PTX:
// three such sites ran 135 M times per 64-spp frame of the cover scene, 2 % of the kernel's warp instructions
// (profiles/r02_cta_ncu_headline.md: device_atomic_functions.hpp:112 rows at 1.0 threads per instruction).  Adding
// threadIdx.y (always 0: blocks are one-dimensional) makes the address thread-dependent as far as ptxas can tell, and
// the atomic stays the single ATOMS.ADD it is.
#ifndef RT_CTA_PLAIN_TAKE
#define RT_CTA_PLAIN_TAKE 1
#endif
RT_DEV uint32_t warp_take(uint32_t* counter, uint32_t n, uint32_t lane) {
    uint32_t old = 0;
#if RT_CTA_PLAIN_TAKE
    if (lane == 0) old = atomicAdd(counter + threadIdx.y, n);
#else
    if (lane == 0) old = atomicAdd(counter, n);
#endif
    return __shfl_sync(0xffffffffu, old, 0);
}

// Traversal registers of one lane.
struct CtaLane {
    uint32_t pid;
    Ray ray;
    RaySlab rs;
    TravState ts;
    XformCache xc;
    PcgRng rng;
};

// Load the ray of pool slot `idx` into the lane and set its traversal up (main.rs:39-44).
template <uint32_t F, bool COUNT = false, bool FLAT = false>
RT_DEV void cta_fetch(const DeviceScene& sc, const CtaPool& P, uint32_t idx, CtaLane& L, TravCounters* tc) {
    const float4 rd = P.d[idx], ro = P.o[idx];
    L.pid = idx;
    L.ray.o = mk3(ro.x, ro.y, ro.z);
    L.ray.d = mk3(rd.x, rd.y, rd.z);
    L.ray.tm = ro.w;
    L.ts.init(3.402823466e+38f);
    if (F & F_XFORM) {
        L.xc.chain = 0;
        L.xc.ro = L.ray;
    }
    if (F & F_MEDIA) {
        const uint2 st = *reinterpret_cast<const uint2*>(P.m + idx);
        L.rng.state = ((unsigned long long)st.y << 32) | st.x;
    }
    if (vec3a_near_one(L.ray.d)) {
        L.rs.init(L.ray);
        if (!FLAT) giants_test<COUNT>(sc, L.ray, 1e-3f, L.ts, tc);  // (the flat loop walks every world primitive itself)
    } else {
        // main.rs:39: a non-unit direction loses the path; it retires as a miss with zero throughput
        P.b[idx] = make_float4(0.0f, 0.0f, 0.0f, P.b[idx].w);
        L.ts.cur = TRAV_NONE;
    }
}

// Traversal of the lane's ray is complete: hit record back to the slot, slot id to its class queue.
template <uint32_t F>
RT_DEV void cta_retire(const CtaPool& P, CtaLane& L, uint32_t* s_cnt) {
    uint2* hit = reinterpret_cast<uint2*>(P.m + L.pid) + 1;
    *hit = make_uint2(__float_as_uint(L.ts.best_t), L.ts.best_prim);
    if (F & F_MEDIA) {
        uint2* st = reinterpret_cast<uint2*>(P.m + L.pid);
        *st = make_uint2((uint32_t)L.rng.state, (uint32_t)(L.rng.state >> 32));
    }
    const uint32_t cls = L.ts.best_prim == 0xffffffffu ? 0u : ((L.ts.best_kfx >> 9) & 0x7u);  // shade queue
    const uint32_t pos = atomicAdd(&s_cnt[cls], 1u);
    P.list[cls * P.N + pos] = (uint16_t)L.pid;
    L.pid = CTA_NONE;
}

template <uint32_t F, bool COUNT, uint32_t MASK, bool FLAT = false>
__global__ void __launch_bounds__(CTA_THREADS, CTA_MIN_BLOCKS) k_render_cta(DeviceScene sc, CtaArgs a) {
    extern __shared__ float4 cta_smem[];
    __shared__ uint32_t s_cnt[8];  // per-class queue lengths of this iteration
    __shared__ uint32_t s_next;    // next entry of the scattered part of the todo list the traverse phase hands out
    __shared__ uint32_t s_nextp;   // next entry of the regenerated (camera ray) part
    __shared__ uint32_t s_chunk;   // next 32-entry chunk of the shade phase
    __shared__ uint32_t s_ndead;   // entries of the dead list
    __shared__ uint32_t s_ntodo;   // scattered paths appended to the todo list by the shade phase
    __shared__ uint32_t s_nsurv, s_nregen, s_r0, s_s0;

    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t N = CTA_POOL_PATHS ? CTA_POOL_PATHS : a.n_paths;
    CtaPool P;
    P.o = cta_smem;
    P.d = P.o + N;
    P.b = P.d + N;
    P.m = reinterpret_cast<uint4*>(P.b + N);
    P.list = reinterpret_cast<uint16_t*>(P.m + N);
    P.dead = P.list + (size_t)CLS_COUNT * N;
    P.todo = P.dead + N;
    P.N = N;
    P.perlin = nullptr;
    if (a.stage_perlin) {
        float4* dst = reinterpret_cast<float4*>(reinterpret_cast<char*>(cta_smem) + (((size_t)N * (64 + 2 * (CLS_COUNT + 2)) + 15) & ~(size_t)15));
        for (uint32_t i = tid; i < 256u; i += CTA_THREADS) dst[i] = __ldg(sc.perlin_grad_lin + i);
        uint16_t* pairs = reinterpret_cast<uint16_t*>(dst + 256);
        for (uint32_t i = tid; i < 768u; i += CTA_THREADS) {
            const uint32_t axis = i >> 8, k = i & 255u;
            pairs[i] = (uint16_t)(__ldg(sc.perlin_perm + axis * 256u + k) | (__ldg(sc.perlin_perm + axis * 256u + ((k + 1u) & 255u)) << 8));
        }
        P.perlin = dst;
    }

    for (uint32_t i = tid; i < N; i += CTA_THREADS) {
        P.d[i] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(0u));
        P.dead[i] = (uint16_t)i;
    }
    if (tid == 0) {
        s_ndead = N;  // every slot "dies" once before the first iteration
        s_ntodo = 0u;
    }
    uint32_t n_rays = 0, n_shade = 0;
    unsigned long long cls_total[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // thread 0, COUNT builds only
    TravCounters tc;
    uint2 stack[FLAT ? 1 : STACK_SIZE];
    if (tid == 0) {
#pragma unroll
        for (int c = 0; c < 8; c++) s_cnt[c] = 0u;
    }
    __syncthreads();

#if RT_CTA_PROFILE
    long long pf_phase[4] = {0, 0, 0, 0};  // thread 0: wall cycles of D+regen, traverse, shade (incl. barrier waits), iterations
    long long pf_work[2] = {0, 0};         // lane 0 of every warp: own work cycles in traverse, shade
    long long pf_t = clock64();
#endif
    for (;;) {
        // ================================================================== D: regenerate ended paths
        if (tid == 0) {
            const uint32_t n_dead = s_ndead;
            uint32_t n_regen = 0;
            if (n_dead) {
                const unsigned long long base = atomicAdd(a.path_counter, (unsigned long long)n_dead);
                if (base < a.total_paths) {
                    const unsigned long long left = a.total_paths - base;
                    n_regen = left < (unsigned long long)n_dead ? (uint32_t)left : n_dead;
                    s_s0 = (uint32_t)(base / a.per_sample);
                    s_r0 = (uint32_t)(base % a.per_sample);
                }
            }
            s_nregen = n_regen;
            s_nsurv = s_ntodo;
            s_ntodo = 0u;
            s_ndead = 0u;
            s_next = 0u;
            s_nextp = 0u;
            s_chunk = 0u;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                if (COUNT) cls_total[c] += s_cnt[c];
                s_cnt[c] = 0u;
            }
        }
        __syncthreads();
        const uint32_t n_regen = s_nregen, n_surv = s_nsurv;
        if (n_regen + n_surv == 0u) break;
        for (uint32_t i = tid; i < n_regen; i += CTA_THREADS) {
            const uint32_t pid = P.dead[i];
            // path index k = base + i  ->  (sample, pixel) = divmod(k, width * height), main.rs:84-93
            uint32_t r = s_r0 + i, s = s_s0;
            while (r >= a.per_sample) {
                r -= a.per_sample;
                s++;
            }
            const uint32_t py = r / a.width, px = r - py * a.width;
            PcgRng rng;
            rng.seed(a.seed, r, a.s_begin + s);
            const float u = ((float)px + rng.next()) / (float)a.width;
            const float v = ((float)py + rng.next()) / (float)a.height;
            const Ray ray = camera_get_ray(sc.cam, u, v, rng);
            P.o[pid] = make_float4(ray.o.x, ray.o.y, ray.o.z, ray.tm);
            P.d[pid] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(WF_ALIVE));
            P.b[pid] = make_float4(1.0f, 1.0f, 1.0f, __uint_as_float(r));
            uint2* st = reinterpret_cast<uint2*>(P.m + pid);
            *st = make_uint2((uint32_t)rng.state, (uint32_t)(rng.state >> 32));
            P.todo[n_surv + i] = (uint16_t)pid;  // consecutive pixels of a row sit next to each other
        }
        __syncthreads();
#if RT_CTA_PROFILE
        { long long t = clock64(); if (tid == 0) { pf_phase[0] += t - pf_t; pf_phase[3]++; } pf_t = t; }
#endif

        // ================================================================== A: traverse (one ray per lane)
        {
            CtaLane L;
            L.pid = CTA_NONE;
            L.ray.o = L.ray.d = mk3(0.0f, 0.0f, 0.0f);
            L.ray.tm = 0.0f;
            L.rs.init(L.ray);
            L.ts.init(0.0f);
            L.ts.cur = TRAV_NONE;
            L.xc.chain = 0;
            L.xc.ro = L.ray;
            L.rng.state = 0;
            if (FLAT) {
                // Tiny scenes: 32 todo entries per warp at a time (scattered and regenerated alike), every lane walks
                // the same primitive list — no refill, no stack, no divergence between node and leaf code.
                (void)stack;
                const uint32_t n_todo = n_surv + n_regen;
                for (;;) {
                    const uint32_t base = warp_take(&s_next, 32u, lane);
                    if (base >= n_todo) break;
                    bool active = base + lane < n_todo;
                    if (active) {
                        cta_fetch<F, COUNT, true>(sc, P, P.todo[base + lane], L, &tc);
                        n_rays++;
                        active = L.ts.cur != TRAV_NONE;  // main.rs:39: a non-unit direction loses the path
                    }
                    flat_closest_hit<F, COUNT>(sc, L.ray, 1e-3f, L.ts, L.rng, false, active, &tc);
                    if (base + lane < n_todo) {
                        L.ts.cur = TRAV_NONE;
                        cta_retire<F>(P, L, s_cnt);
                    }
                    __syncwarp();
                }
            } else {
            // The todo list holds the scattered rays [0, n_surv) and behind them the camera rays of the regenerated
            // paths.  Entries [0, n_list) are handed out with lane refill, the rest 32 neighbouring pixels at a time
            // to whole warps that run them to completion.  n_list = n_surv keeps the camera rays coherent (large
            // trees: their node fetches stay broadcasts); n_list = everything ("unified", small trees) tops the lanes a
            // warp has left when the scattered part runs dry up with camera rays instead of letting them finish on
            // their own — the tail is 30 % of the scattered part's lane slots at 56 scattered rays per warp and
            // iteration (tools/bvh_sim.cpp).  The host decides per scene (rtgpu.cu cta_plan).
            const uint32_t n_list = a.unified_todo ? n_surv + n_regen : n_surv;
            const uint32_t n_rest = n_surv + n_regen - n_list;
#if RT_CTA_ONE_LOOP
            // One instance of the traversal loop serves both parts of the todo list (instruction-cache footprint):
            // stage 0 = scattered rays, idle lanes refilled from the block's counter; stage 1 = camera rays, a
            // coherent chunk of 32 neighbouring pixels at a time, run to completion.  `stage` is warp-uniform.
            bool more = n_list != 0u;
            uint32_t stage = more ? 0u : 1u;
            uint32_t thr = a.refill_idle;
            for (;;) {
                if (L.pid != CTA_NONE && L.ts.cur == TRAV_NONE) cta_retire<F>(P, L, s_cnt);
                uint32_t idx = CTA_NONE;  // todo entry this lane fetches now
                if (stage == 0u) {
                    if (more) {
                        const uint32_t wm = __ballot_sync(0xffffffffu, L.pid == CTA_NONE);
                        if (wm != 0u) {
                            const uint32_t base = warp_take(&s_next, (uint32_t)__popc(wm), lane);
                            if (base + (uint32_t)__popc(wm) >= n_list) more = false;
                            if (base + (uint32_t)__popc(wm) >= n_surv) thr = a.refill_camera;
                            const uint32_t i = base + (uint32_t)__popc(wm & ((1u << lane) - 1u));
                            if (L.pid == CTA_NONE && i < n_list) idx = i;
                        }
                    }
                    if (!__any_sync(0xffffffffu, L.pid != CTA_NONE || idx != CTA_NONE)) stage = 1u;  // scattered part drained
                }
                if (stage == 1u) {  // every lane is idle here
                    const uint32_t base = warp_take(&s_nextp, 32u, lane);
                    if (base >= n_rest) break;
                    if (base + lane < n_rest) idx = n_list + base + lane;
                }
                if (idx != CTA_NONE) {
                    cta_fetch<F, COUNT>(sc, P, P.todo[idx], L, &tc);
                    n_rays++;
                }
                // ---- traverse until enough lanes are idle again (stage 0, while there is something to refill them
                //      with) or until the whole warp is done
                for (;;) {
                    while (!(L.ts.cur & LEAF_FLAG) && L.ts.cur != TRAV_NONE) node_visit<COUNT>(sc, L.rs, 1e-3f, L.ts, stack, &tc);
#if RT_CTA_LEAF_WHILE
                    while (L.ts.cur & LEAF_FLAG) {  // (a popped entry that is a leaf is tested in the same round)
#else
                    if (L.ts.cur != TRAV_NONE) {
#endif
                        leaf_visit<F, COUNT>(sc, L.ray, 1e-3f, L.ts, L.xc, L.rng, false, &tc);
                        L.ts.pop(stack);
                    }
                    const uint32_t busy = __ballot_sync(0xffffffffu, L.ts.cur != TRAV_NONE);
                    if (busy == 0u) break;
                    if (stage == 0u && more && 32u - (uint32_t)__popc(busy) >= thr) break;
                }
            }
#else
            // ---- A2: scattered rays (incoherent): idle lanes are refilled from the block's counter
            bool more = n_list != 0u;  // the todo list may still hold entries nobody has looked at (warp-uniform)
            uint32_t thr = a.refill_idle;
            for (;;) {
                if (L.pid != CTA_NONE && L.ts.cur == TRAV_NONE) cta_retire<F>(P, L, s_cnt);
                while (more) {
                    const uint32_t wm = __ballot_sync(0xffffffffu, L.pid == CTA_NONE);
                    if (wm == 0u) break;
                    const uint32_t base = warp_take(&s_next, (uint32_t)__popc(wm), lane);
                    if (base + (uint32_t)__popc(wm) >= n_list) more = false;
                    if (base + (uint32_t)__popc(wm) >= n_surv) thr = a.refill_camera;  // (the camera part of a unified list)
                    const uint32_t idx = base + (uint32_t)__popc(wm & ((1u << lane) - 1u));
                    if (L.pid == CTA_NONE && idx < n_list) {
                        cta_fetch<F, COUNT>(sc, P, P.todo[idx], L, &tc);
                        n_rays++;
                    }
                }
                if (!__any_sync(0xffffffffu, L.pid != CTA_NONE)) break;
                // ---- traverse until enough lanes are idle again (and there is something to refill them with)
                for (;;) {
                    while (!(L.ts.cur & LEAF_FLAG) && L.ts.cur != TRAV_NONE) node_visit<COUNT>(sc, L.rs, 1e-3f, L.ts, stack, &tc);
#if RT_CTA_LEAF_WHILE
                    while (L.ts.cur & LEAF_FLAG) {  // (a popped entry that is a leaf is tested in the same round)
#else
                    if (L.ts.cur != TRAV_NONE) {
#endif
                        leaf_visit<F, COUNT>(sc, L.ray, 1e-3f, L.ts, L.xc, L.rng, false, &tc);
                        L.ts.pop(stack);
                    }
                    const uint32_t busy = __ballot_sync(0xffffffffu, L.ts.cur != TRAV_NONE);
                    if (busy == 0u) break;
                    if (more && 32u - (uint32_t)__popc(busy) >= thr) break;
                }
            }
            // ---- A1: camera rays, 32 neighbouring pixels per warp: coherent, so the warp runs them to completion
            for (;;) {
                const uint32_t base = warp_take(&s_nextp, 32u, lane);
                if (base >= n_rest) break;
                if (base + lane < n_rest) {
                    cta_fetch<F, COUNT>(sc, P, P.todo[n_list + base + lane], L, &tc);
                    n_rays++;
                    while (L.ts.cur != TRAV_NONE) {
                        while (!(L.ts.cur & LEAF_FLAG) && L.ts.cur != TRAV_NONE) node_visit<COUNT>(sc, L.rs, 1e-3f, L.ts, stack, &tc);
                        if (L.ts.cur == TRAV_NONE) break;
                        leaf_visit<F, COUNT>(sc, L.ray, 1e-3f, L.ts, L.xc, L.rng, false, &tc);
                        L.ts.pop(stack);
                    }
                    cta_retire<F>(P, L, s_cnt);
                }
                __syncwarp();
            }
#endif
            }  // !FLAT
        }
#if RT_CTA_PROFILE
        if (lane == 0) pf_work[0] += clock64() - pf_t;
#endif
        __syncthreads();
#if RT_CTA_PROFILE
        { long long t = clock64(); if (tid == 0) pf_phase[1] += t - pf_t; pf_t = t; }
#endif

        // ================================================================== C: shade, class by class
        {
            // Class ranges in a common index space, each padded to a warp multiple, the most expensive class
            // first; warps pull 32-entry chunks from a counter (longest-processing-time-first keeps the tail short).
            uint32_t start[CLS_COUNT + 1];
            start[0] = 0u;
#pragma unroll
            for (int c = 0; c < CLS_COUNT; c++) start[c + 1] = start[c] + ((s_cnt[(c + 1) % CLS_COUNT] + 31u) & ~31u);  // 1..6, then 0
            for (;;) {
                const uint32_t g = warp_take(&s_chunk, 32u, lane);
                if (g >= start[CLS_COUNT]) break;
#if RT_CTA_DECODE_TREE
                // range k of the chunk (g is warp-uniform): three uniform compares down a tree instead of six + six selects
                static_assert(CLS_COUNT == 7, "the decode tree below is written for seven queues");
                uint32_t k, first;
                if (g < start[3]) {
                    if (g < start[1]) { k = 0u; first = 0u; }
                    else if (g < start[2]) { k = 1u; first = start[1]; }
                    else { k = 2u; first = start[2]; }
                } else if (g < start[5]) {
                    if (g < start[4]) { k = 3u; first = start[3]; }
                    else { k = 4u; first = start[4]; }
                } else {
                    if (g < start[6]) { k = 5u; first = start[5]; }
                    else { k = 6u; first = start[6]; }
                }
#else
                int k = 0;
#pragma unroll
                for (int c = 1; c < CLS_COUNT; c++) k += g >= start[c] ? 1 : 0;
                uint32_t first = 0;
#pragma unroll
                for (int c = 1; c < CLS_COUNT; c++) first = k == c ? start[c] : first;
#endif
                const uint32_t q = k == CLS_COUNT - 1 ? 0u : (uint32_t)k + 1u;  // queue of this chunk
                const int cls = (int)((a.q_cls >> (4 * q)) & 0xfu);  // the class it shades as
                const uint32_t idx = g - first + lane;
                if (idx < s_cnt[q]) {
                    const uint32_t pid = P.list[q * N + idx];
                    switch (cls) {
                        case CLS_MISS: cta_shade_item<F, CLS_MISS>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        case CLS_DIFFUSE_CONST: cta_shade_item<F, CLS_DIFFUSE_CONST>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        case CLS_METAL: cta_shade_item<F, CLS_METAL>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        case CLS_DIELECTRIC: cta_shade_item<F, CLS_DIELECTRIC>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        case CLS_DIFFUSE_PERLIN: cta_shade_item<F, CLS_DIFFUSE_PERLIN>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        case CLS_CLASSIC: cta_shade_item<F, CLS_CLASSIC>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade); break;
                        default:
                            if ((MASK >> CLS_PBR) & 1u) cta_shade_item<F, CLS_PBR>(sc, a, P, pid, &s_ndead, &s_ntodo, n_shade);
                            break;
                    }
                }
                __syncwarp();
            }
        }
#if RT_CTA_PROFILE
        if (lane == 0) pf_work[1] += clock64() - pf_t;
#endif
        __syncthreads();
#if RT_CTA_PROFILE
        { long long t = clock64(); if (tid == 0) pf_phase[2] += t - pf_t; pf_t = t; }
#endif
    }
#if RT_CTA_PROFILE
    if (tid == 0) {
        atomicAdd(a.stats + 8, (unsigned long long)pf_phase[0]);
        atomicAdd(a.stats + 9, (unsigned long long)pf_phase[1]);
        atomicAdd(a.stats + 10, (unsigned long long)pf_phase[2]);
        atomicAdd(a.stats + 11, (unsigned long long)pf_phase[3]);
    }
    if (lane == 0) {
        atomicAdd(a.stats + 12, (unsigned long long)pf_work[0]);
        atomicAdd(a.stats + 13, (unsigned long long)pf_work[1]);
    }
#endif

    const unsigned long long rays = warp_sum_u64(n_rays), shades = warp_sum_u64(n_shade);
    if (lane == 0) {
        atomicAdd(a.stats + 0, rays);
        atomicAdd(a.stats + 1, shades);
    }
    if (COUNT) {
        unsigned long long nv = warp_sum_u64(tc.node_visits), bt = warp_sum_u64(tc.box_tests), pt = warp_sum_u64(tc.prim_tests);
        if (lane == 0) {
            atomicAdd(a.stats + 2, nv);
            atomicAdd(a.stats + 3, bt);
            atomicAdd(a.stats + 4, pt);
        }
        if (tid == 0) {
#pragma unroll
            for (int c = 0; c < CLS_COUNT; c++) atomicAdd(a.stats + 8 + c, cls_total[c]);
        }
    }
}

// Host-side entry points, one translation unit per instantiated feature set (cta_inst.cu, -DRT_CTA_F=<F>).
// Feature sets built: 0 (spheres only), F_MOTION, F_RECT, F_RECT|F_XFORM, F_RECT|F_XFORM|F_MEDIA, F_ALL.
void* cta_kernel_f0(bool count, int* threads, int* pool_paths);  // *threads = block size the kernel was built for
void* cta_kernel_f4(bool count, int* threads, int* pool_paths);
void* cta_kernel_f8(bool count, int* threads, int* pool_paths);
void* cta_kernel_f9(bool count, int* threads, int* pool_paths);
void* cta_kernel_f11(bool count, int* threads, int* pool_paths);
void* cta_kernel_f15(bool count, int* threads, int* pool_paths);
void* cta_kernel_f0_m3f(bool count, int* threads, int* pool_paths);  // the same without the pbr.rs shading class
void* cta_kernel_f4_m3f(bool count, int* threads, int* pool_paths);
void* cta_kernel_f0_flat(bool count, int* threads, int* pool_paths);  // FLAT builds (tiny scenes): spheres; + rects; + transforms; + media
void* cta_kernel_f8_flat(bool count, int* threads, int* pool_paths);
void* cta_kernel_f9_flat(bool count, int* threads, int* pool_paths);
void* cta_kernel_f11_flat(bool count, int* threads, int* pool_paths);

// The kernel of the smallest built feature set that covers `features`.
inline void* cta_kernel_for(uint32_t features, uint32_t class_mask, bool count, bool flat, uint32_t* built, int* threads, int* pool_paths = nullptr) {
    struct E { uint32_t f; void* (*fn)(bool, int*, int*); };
    if (flat) {
        const E flat_table[] = {{0u, cta_kernel_f0_flat}, {8u, cta_kernel_f8_flat}, {9u, cta_kernel_f9_flat}, {11u, cta_kernel_f11_flat}};
        for (const E& e : flat_table)
            if ((features & ~e.f) == 0u) {
                if (built) *built = e.f;
                return e.fn(count, threads, pool_paths);
            }
    }
    if (!((class_mask >> CLS_PBR) & 1u)) {  // no pbr.rs material in the scene
        if (features == 0u) { if (built) *built = 0u; return cta_kernel_f0_m3f(count, threads, pool_paths); }
        if (features == 4u) { if (built) *built = 4u; return cta_kernel_f4_m3f(count, threads, pool_paths); }
    }
    const E table[] = {{0u, cta_kernel_f0}, {4u, cta_kernel_f4}, {8u, cta_kernel_f8}, {9u, cta_kernel_f9},
                       {11u, cta_kernel_f11}, {15u, cta_kernel_f15}};
    for (const E& e : table)
        if ((features & ~e.f) == 0u) {
            if (built) *built = e.f;
            return e.fn(count, threads, pool_paths);
        }
    if (built) *built = 15u;
    return cta_kernel_f15(count, threads, pool_paths);
}

}  // namespace rtb
