/* Plain C99 caller of the drop-in boundary (include/rtgpu.h): what the reference's main() does
 * (src/main.rs:62-129) — build a scene, render it, apply the output stage, save a PNG — with the sample loop
 * and ray_color replaced by librtgpu.so.  No C++, no CUDA headers, no torch on this side of the ABI.
 *
 *   gcc -std=c99 -Iinclude examples/render_scene.c -Lray_tracing_in_one_weekend_b200 -lrtgpu \
 *       -Wl,-rpath,$PWD/ray_tracing_in_one_weekend_b200 -o render_scene
 *   ./render_scene sphere_scene 480 270 64 assets out.png
 *
 * tests/test_host_api.py compiles and links this file on the CPU; tests/test_gpu_render.py runs it on the B200. */
#include <stdio.h>
#include <stdlib.h>

#include "rtgpu.h"

#define CHECK(call)                                                                   \
    do {                                                                              \
        int rc_ = (call);                                                             \
        if (rc_ != RT_OK) {                                                           \
            fprintf(stderr, "%s failed: %d (%s)\n", #call, rc_, rt_last_error());     \
            return 1;                                                                 \
        }                                                                             \
    } while (0)

int main(int argc, char** argv) {
    const char* name = argc > 1 ? argv[1] : "sphere_scene";
    const uint32_t nx = argc > 2 ? (uint32_t)atoi(argv[2]) : 480u;
    const uint32_t ny = argc > 3 ? (uint32_t)atoi(argv[3]) : 270u;
    const uint32_t ns = argc > 4 ? (uint32_t)atoi(argv[4]) : 64u;
    const char* assets = argc > 5 ? argv[5] : "assets";
    const char* out = argc > 6 ? argv[6] : "out.png";

    /* src/main.rs:74: let (world, cam) = scene_fn(aspect)  ->  host scene API, flattened */
    RtHostScene* hs = NULL;
    CHECK(rt_host_scene_build(name, (float)nx / (float)ny, assets, &hs));
    RtScene* scene = NULL;
    CHECK(rt_scene_create(rt_host_scene_flat(hs), 0, &scene));

    /* src/main.rs:77-108: the column jobs, the sample loop and ray_color  ->  one call */
    RtRenderParams rp;
    rp.width = nx;
    rp.height = ny;
    rp.spp_total = ns;
    rp.sample_begin = 0;
    rp.sample_end = ns;
    rp.max_depth = 50; /* src/main.rs:36 */
    rp.mode = RT_MODE_AUTO;
    rp.flags = 0;
    rp.seed = 95; /* the reference seeds column i with 95 + i (src/main.rs:82) */
    float* mean = (float*)malloc((size_t)nx * ny * 3 * sizeof(float));
    uint8_t* rgb8 = (uint8_t*)malloc((size_t)nx * ny * 3);
    if (!mean || !rgb8) return 2;
    RtStats st;
    CHECK(rt_render(scene, &rp, mean, &st));

    /* src/main.rs:98-105,127-128: gamma 2, *255.99 as u8, flip, save */
    CHECK(rt_resolve_rgb8(scene, mean, nx, ny, rgb8));
    CHECK(rt_write_png(out, rgb8, nx, ny));
    printf("%s %ux%u %u spp: %llu paths, %llu rays, %.3f ms kernel, %.1f Mpaths/s -> %s\n", name, nx, ny, ns,
           (unsigned long long)st.paths, (unsigned long long)st.rays, st.ms_kernel,
           (double)st.paths / st.ms_kernel / 1e3, out);

    free(mean);
    free(rgb8);
    rt_scene_destroy(scene);
    rt_host_scene_destroy(hs);
    return 0;
}
