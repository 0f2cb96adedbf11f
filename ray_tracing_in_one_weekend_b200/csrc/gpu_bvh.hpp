// GPU LBVH builder (gpu_bvh.cu): same node format as the host SAH builder, nodes left in device memory.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "../../include/rtgpu.h"
#include "device_types.h"

namespace rtb {

struct GpuBvhResult {
    BvhNode4* nodes = nullptr;         // device memory, cudaMalloc'ed; ownership passes to the caller
    uint32_t n_nodes = 0;
    uint32_t n_world_prims = 0;
    uint32_t max_depth = 0;            // BVH4 levels
    float gpu_ms = 0.0f;               // Morton codes + sort + tree + fit + collapse, CUDA events
    std::vector<uint32_t> leaf_order;  // flat primitive indices in leaf (= Morton) order
};

// Builds the tree on `stream`'s device.  Returns false with `err` set on CUDA errors or malformed input.
bool gpu_build_bvh4(const RtFlatScene& fs, cudaStream_t stream, GpuBvhResult& out, std::string& err);

}  // namespace rtb
