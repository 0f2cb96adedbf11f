#pragma once
#include <vector>

#include "../../include/rtgpu.h"
#include "device_types.h"

namespace rtb {

struct BvhBuildResult {
    std::vector<BvhNode4> nodes;       // nodes[0] is the root
    std::vector<uint32_t> leaf_order;  // flat primitive indices in leaf order (device prim i = flat leaf_order[i])
    uint32_t n_world_prims = 0;
    uint32_t max_depth = 0;
    uint32_t max_leaf_prims = 0;
    float sah_cost = 0.0f;
    float root_lo[3] = {0, 0, 0}, root_hi[3] = {0, 0, 0};
};

struct Box6 {
    float lo[3], hi[3];
};

// World primitives (flat indices) with their reference bounds (centroids come from these) and the padded
// conservative bounds every BVH must store (bvh_build.cpp).  Throws std::runtime_error on an empty world.
void compute_world_boxes(const RtFlatScene& fs, std::vector<uint32_t>& world, std::vector<Box6>& ref, std::vector<Box6>& padded);

// Builds the 4-wide SAH BVH over the world primitives (those without RT_PRIM_FLAG_BOUNDARY_ONLY).
// Throws std::runtime_error on malformed input.
void build_bvh4(const RtFlatScene& fs, uint32_t leaf_max, BvhBuildResult& out);

}  // namespace rtb
