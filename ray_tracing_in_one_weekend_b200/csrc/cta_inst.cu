// One instantiation of the CTA-local wavefront renderer (kernels_cta.cuh) per translation unit:
//   nvcc -DRT_CTA_F=<feature set> -c cta_inst.cu -o cta_f<F>.o
// so the feature sets compile in parallel and each kernel carries only the code its scenes need
// (a spheres-only scene never pays registers or spills for media, transforms or rects).
#include "kernels_cta.cuh"

#ifndef RT_CTA_F
#error "compile with -DRT_CTA_F=<feature bit set>"
#endif

#define RT_CAT2(a, b) a##b
#define RT_CAT(a, b) RT_CAT2(a, b)

#ifdef RT_CTA_SUFFIX  // e.g. -DRT_CTA_CLS_MASK=0x3f -DRT_CTA_SUFFIX=_m3f: the build without the pbr.rs shading class;
                      //      -DRT_CTA_FLAT=1 -DRT_CTA_SUFFIX=_flat: the tiny-scene build (warp-uniform primitive loop)
#define RT_CTA_ENTRY RT_CAT(RT_CAT(cta_kernel_f, RT_CTA_F), RT_CTA_SUFFIX)
#else
#define RT_CTA_ENTRY RT_CAT(cta_kernel_f, RT_CTA_F)
#endif

namespace rtb {
void* RT_CTA_ENTRY(bool count, int* threads, int* pool_paths) {
    if (threads) *threads = CTA_THREADS;
    if (pool_paths) *pool_paths = (int)CTA_POOL_PATHS;
    return count ? (void*)k_render_cta<(uint32_t)RT_CTA_F, true, (uint32_t)RT_CTA_CLS_MASK, (RT_CTA_FLAT) != 0>
                 : (void*)k_render_cta<(uint32_t)RT_CTA_F, false, (uint32_t)RT_CTA_CLS_MASK, (RT_CTA_FLAT) != 0>;
}
}  // namespace rtb
