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

namespace rtb {
void* RT_CAT(cta_kernel_f, RT_CTA_F)(bool count, int* threads) {
    if (threads) *threads = CTA_THREADS;
    return count ? (void*)k_render_cta<(uint32_t)RT_CTA_F, true> : (void*)k_render_cta<(uint32_t)RT_CTA_F, false>;
}
}  // namespace rtb
