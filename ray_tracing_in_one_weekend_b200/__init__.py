"""B200-native path-tracing backend for the hot path of zhouhang95/ray_tracing_in_one_weekend.

Product = csrc/ (CUDA kernels + C ABI + host scene API) behind include/rtgpu.h; this package is the
thin ctypes mirror used by tests and bench.py.  No CPU fallback exists.
"""
from .api import (  # noqa: F401
    HostScene, Scene, Comm, RtError, RtRenderParams, RtStats, RtFlatScene, RtCamera, HIT_DTYPE, load_library,
    camera_get_rays, pixel_center_uv, flat_to_numpy, measure_fp32_peak, load_test_library, partition_samples,
    flat_validate, release_cached_memory, RT_ERR_ARG, RT_ERR_CUDA, RT_ERR_IO, RT_ERR_UNSUPPORTED, RT_ERR_NCCL, LIB_PATH, ASSET_DIR, EXPORTED_SYMBOLS,
    MODE_AUTO, MODE_MEGAKERNEL, MODE_CTA_WAVEFRONT, RENDER_COUNTERS, RENDER_RUSSIAN_ROULETTE, RENDER_COSINE_SAMPLING, RENDER_LOBE_SAMPLING, SCENE_GPU_BVH,
)
