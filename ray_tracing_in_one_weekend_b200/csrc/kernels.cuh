// CUDA kernels of the backend (sm_100a).  See rt_device.cuh for the numerics contract.
#pragma once
#include "rt_device.cuh"

namespace rtb {

// ------------------------------------------------------------------------------ k_closest_hit_batch
// One thread per ray.  Replaces `world.hit(&r, t_min, t_max, &mut rec)` (main.rs:44) for a batch
// and returns the full record the reference's primitive would have written.
__global__ void __launch_bounds__(128) k_closest_hit_batch(DeviceScene sc, const float* __restrict__ o_xyz,
                                                           const float* __restrict__ d_xyz,
                                                           const float* __restrict__ times, size_t n, float t_min,
                                                           float t_max, RtHit* __restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r;
    r.o = mk3(o_xyz[3 * i], o_xyz[3 * i + 1], o_xyz[3 * i + 2]);
    r.d = mk3(d_xyz[3 * i], d_xyz[3 * i + 1], d_xyz[3 * i + 2]);
    r.tm = times ? times[i] : 0.0f;
    PcgRng rng;
    rng.state = 0;
    ClosestHit h = closest_hit<F_ALL, false>(sc, r, t_min, t_max, rng, /*skip_media=*/true, nullptr);
    RtHit o;
    if (h.prim == 0xffffffffu) {
        o.prim = -1;
        o.t = 0.0f;
        o.p[0] = o.p[1] = o.p[2] = 0.0f;
        o.n[0] = o.n[1] = o.n[2] = 0.0f;
        o.uv[0] = o.uv[1] = 0.0f;
        o.front_face = 0;
    } else {
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

// ------------------------------------------------------------------------------------ k_render_mega
// Persistent-thread megakernel: every warp pulls (8x4-pixel tile, sample chunk) work units from a
// global counter; each lane owns one pixel of the tile, regenerates a new camera path as soon as
// its current path terminates (so lanes stay busy while path lengths vary from 1 to 51 rays), and
// keeps the pixel's radiance sum in registers until the unit is finished.
// Replaces the sample loop + ray_color recursion (main.rs:38-60,84-97) iteratively:
//   L += beta * emitted;  beta *= attenuation;  on a miss L += beta * sky(d);  at most max_depth+1 rays.
struct RenderArgs {
    uint32_t width, height;
    uint32_t s_begin, s_end;  // global sample indices rendered by this launch
    uint32_t max_depth;
    uint32_t tiles_x, n_tiles;
    uint32_t chunk_spp, n_chunks;
    uint64_t seed;
    float* accum;                  // width*height*3 radiance sums
    unsigned int* work_counter;    // zeroed before launch
    unsigned long long* stats;     // [0] rays [1] shade calls [2] node visits [3] box tests [4] prim tests
};

constexpr int MEGA_THREADS = 128;

template <uint32_t F, bool COUNT>
__global__ void __launch_bounds__(MEGA_THREADS) k_render_mega(DeviceScene sc, RenderArgs a) {
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long n_rays = 0, n_shade = 0;
    TravCounters tc;
    const float nx = (float)a.width, ny = (float)a.height;
    const uint32_t n_units = a.n_tiles * a.n_chunks;

    for (;;) {
        uint32_t unit = 0;
        if (lane == 0) unit = atomicAdd(a.work_counter, 1u);
        unit = __shfl_sync(0xffffffffu, unit, 0);
        if (unit >= n_units) break;
        const uint32_t chunk = unit / a.n_tiles, tile = unit - chunk * a.n_tiles;
        const uint32_t ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
        const uint32_t px = tx * 8u + (lane & 7u), py = ty * 4u + (lane >> 3);
        const bool valid = px < a.width && py < a.height;
        const uint32_t pixel = py * a.width + px;
        uint32_t s = a.s_begin + chunk * a.chunk_spp;
        const uint32_t s_hi = min(a.s_end, s + a.chunk_spp);

        V3f sum = mk3(0.0f, 0.0f, 0.0f);
        V3f beta = mk3(1.0f, 1.0f, 1.0f);
        Ray ray;
        ray.o = ray.d = mk3(0.0f, 0.0f, 0.0f);
        ray.tm = 0.0f;
        PcgRng rng;
        rng.state = 0;
        uint32_t depth = 0;
        bool alive = false;

        for (;;) {
            if (!alive && valid && s < s_hi) {
                // main.rs:88-93: u = (i + xi) / nx, v = (j + xi) / ny, u drawn before v
                rng.seed(a.seed, pixel, s);
                float u = ((float)px + rng.next()) / nx;
                float v = ((float)py + rng.next()) / ny;
                ray = camera_get_ray(sc.cam, u, v, rng);
                beta = mk3(1.0f, 1.0f, 1.0f);
                depth = 0;
                alive = true;
                s++;
            }
            if (!__any_sync(0xffffffffu, alive)) break;
            if (alive) {
                // main.rs:39: the reference asserts |d| == 1 within 1e-6 and loses the path otherwise
                if (!vec3a_near_one(ray.d)) {
                    alive = false;
                } else {
                    n_rays++;
                    ClosestHit h = closest_hit<F, COUNT>(sc, ray, 1e-3f, 3.402823466e+38f, rng, false, &tc);
                    if (h.prim == 0xffffffffu) {
                        sum = sum + beta * sky_color(sc, ray.d);  // main.rs:58
                        alive = false;
                    } else {
                        n_shade++;
                        DevPrim hp = load_prim(sc.prims, h.prim);
                        DevMaterial m = load_mat(sc, hp.mat);
                        HitRec rec = fill_hit<F>(sc, ray, h.prim, h.t, (m.reserved & 1u) != 0u);
                        V3f emitted, att;
                        Ray scattered;
                        bool ok = shade(sc, m, ray, rec, rng, emitted, att, scattered, sampling_of(a.max_depth));
                        sum = sum + beta * emitted;  // main.rs:47
                        if (ok && (a.max_depth & DEPTH_RR_FLAG)) ok = russian_roulette(att, rng);
                        if (ok) {                    // main.rs:48-54
                            beta = beta * att;
                            ray = scattered;
                            depth++;
                            if (depth > (a.max_depth & DEPTH_MASK)) alive = false;  // main.rs:40-42
                        } else {
                            alive = false;
                        }
                    }
                }
            }
        }
        if (valid) {
            float* px_out = a.accum + (size_t)pixel * 3u;
            atomicAdd(px_out + 0, sum.x);
            atomicAdd(px_out + 1, sum.y);
            atomicAdd(px_out + 2, sum.z);
        }
    }
    n_rays = warp_sum_u64(n_rays);
    n_shade = warp_sum_u64(n_shade);
    if (lane == 0) {
        atomicAdd(a.stats + 0, n_rays);
        atomicAdd(a.stats + 1, n_shade);
    }
    if (COUNT) {
        unsigned long long nv = warp_sum_u64(tc.node_visits), bt = warp_sum_u64(tc.box_tests), pt = warp_sum_u64(tc.prim_tests);
        if (lane == 0) {
            atomicAdd(a.stats + 2, nv);
            atomicAdd(a.stats + 3, bt);
            atomicAdd(a.stats + 4, pt);
        }
    }
}

// ---------------------------------------------------------------------------------------- k_resolve
// main.rs:98-105,127: c = sum / spp; c = c.powf(0.5); (c * 255.99) as u8 (saturating, NaN -> 0);
// vertical flip so that output row 0 is the top of the picture.
RT_DEV uint8_t quantise(float c) {
    float v = powf(c, 0.5f) * 255.99f;
    if (!(v == v)) return 0;
    if (v >= 255.0f) return 255;
    if (v <= 0.0f) return 0;
    return (uint8_t)v;
}
__global__ void k_resolve_rgb8(const float* __restrict__ mean, uint32_t width, uint32_t height, uint8_t* __restrict__ out) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= width * height) return;
    uint32_t x = i % width, y = i / width;
    const float* c = mean + (size_t)i * 3u;
    uint8_t* o = out + ((size_t)(height - 1u - y) * width + x) * 3u;
    o[0] = quantise(c[0]);
    o[1] = quantise(c[1]);
    o[2] = quantise(c[2]);
}
__global__ void k_divide(float* __restrict__ buf, size_t n, float divisor) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) buf[i] = buf[i] / divisor;  // c /= spp (main.rs:98): a true division
}

// ------------------------------------------------------------------------------------ k_fp32_peak
// Issue-rate microbenchmark: 8 independent FFMA chains per thread (explicit fmaf intrinsics are not
// affected by -fmad=false).  Gives the MEASURED denominator of the FP32-lane roofline.
__global__ void __launch_bounds__(256) k_fp32_peak(float* out, int iters, float seed) {
    float a0 = seed, a1 = seed + 1.0f, a2 = seed + 2.0f, a3 = seed + 3.0f;
    float a4 = seed + 4.0f, a5 = seed + 5.0f, a6 = seed + 6.0f, a7 = seed + 7.0f;
    const float m = 0.999999f, c = 1e-7f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
            a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
        }
    }
    float s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == 12345.678f) out[0] = s;  // keep the chains alive
}

}  // namespace rtb
