/* Plain C99 caller of the drop-in boundary (include/rtgpu.h): what the reference's main() does
 * (src/main.rs:62-129) — build a scene, render it, apply the output stage, save a PNG — with the sample loop
 * and ray_color replaced by librtgpu.so.  No C++, no CUDA headers, no torch on this side of the ABI.
 *
 *   gcc -std=c99 -Iinclude examples/render_scene.c -Lray_tracing_in_one_weekend_b200 -lrtgpu \
 *       -Wl,-rpath,$PWD/ray_tracing_in_one_weekend_b200 -o render_scene
 *   ./render_scene sphere_scene 480 270 64 assets out.png [n_gpus]
 *
 * With n_gpus > 1 the ONE process drives n devices: the scene is replicated (one rt_scene_create per device), the
 * samples of every pixel are partitioned, and rt_render_reduce combines the per-GPU sum buffers with one NCCL
 * reduce — what replaces the reference's mpsc gather of finished pixels (src/main.rs:72,101,114).
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
    const int n_gpus = argc > 7 ? atoi(argv[7]) : 1;
    enum { MAX_GPUS = 16 };
    if (n_gpus < 1 || n_gpus > MAX_GPUS) {
        fprintf(stderr, "n_gpus must be 1..%d\n", (int)MAX_GPUS);
        return 2;
    }

    /* src/main.rs:74: let (world, cam) = scene_fn(aspect)  ->  host scene API, flattened */
    RtHostScene* hs = NULL;
    CHECK(rt_host_scene_build(name, (float)nx / (float)ny, assets, &hs));
    RtScene* scenes[MAX_GPUS];
    int devices[MAX_GPUS];
    int g;
    for (g = 0; g < n_gpus; g++) { /* the scene is replicated: one device copy per GPU */
        devices[g] = g;
        scenes[g] = NULL;
        CHECK(rt_scene_create(rt_host_scene_flat(hs), g, &scenes[g]));
    }
    RtScene* scene = scenes[0];

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
    RtComm* comm = NULL;
    if (n_gpus == 1) {
        CHECK(rt_render(scene, &rp, mean, &st));
    } else {
        /* src/main.rs:72,101,114: workers send finished pixels through a channel  ->  one ncclReduce per frame */
        CHECK(rt_comm_create_all(devices, n_gpus, &comm));
        CHECK(rt_render_reduce(scenes, comm, &rp, mean, &st));
    }

    /* src/main.rs:98-105,127-128: gamma 2, *255.99 as u8, flip, save */
    CHECK(rt_resolve_rgb8(scene, mean, nx, ny, rgb8));
    CHECK(rt_write_png(out, rgb8, nx, ny));
    printf("%s %ux%u %u spp: %llu paths, %llu rays, %.3f ms kernel, %.1f Mpaths/s -> %s\n", name, nx, ny, ns,
           (unsigned long long)st.paths, (unsigned long long)st.rays, st.ms_kernel,
           (double)st.paths / st.ms_kernel / 1e3, out);

    free(mean);
    free(rgb8);
    rt_comm_destroy(comm);
    for (g = 0; g < n_gpus; g++) rt_scene_destroy(scenes[g]);
    rt_host_scene_destroy(hs);
    return 0;
}
