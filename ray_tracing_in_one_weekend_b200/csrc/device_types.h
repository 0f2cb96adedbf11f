// Plain data layouts shared by the host-side builders (bvh_build.cpp, scene_upload) and the CUDA
// kernels.  Everything the traversal/shading kernels fetch is 16- or 32-byte aligned and read with
// 128-bit loads (LDG.E.128 through the read-only path).
#pragma once
#include <stdint.h>

namespace rtb {

// 4-wide BVH node, 128 bytes = 4 x 32-byte sectors, child boxes stored SoA so that one float4 load
// yields the same plane of all four children.  Empty slots have lo = +inf, hi = -inf (never hit).
struct alignas(32) BvhNode4 {
    float lox[4], hix[4];
    float loy[4], hiy[4];
    float loz[4], hiz[4];
    uint32_t child[4];  // interior: node index; leaf: LEAF_FLAG | (count-1) << LEAF_COUNT_SHIFT | first device prim
    uint32_t pad[4];
};
static_assert(sizeof(BvhNode4) == 128, "BvhNode4 must be 128 bytes");

constexpr uint32_t LEAF_FLAG = 0x80000000u;
constexpr uint32_t LEAF_COUNT_SHIFT = 27;            // 4 bits of (count-1): up to 16 prims per leaf
constexpr uint32_t LEAF_FIRST_MASK = (1u << 27) - 1;  // 134M primitives

// Device primitive record, 32 bytes = two 128-bit loads.
//   q0 (float4)  sphere / moving sphere: cx, cy, cz, r
//                rect: k (plane coordinate = min.k), a0, a1, b0   (a,b = the two in-plane axes, in
//                      the order the reference tests them / builds uv: XY -> x,y  XZ -> x,z  YZ -> y,z)
//                medium: neg_inv_density, bits(first boundary device prim), bits(boundary count), 0
//   q1.x         rect: b1 (float);  moving sphere: bits(index into DeviceScene::motion)
//   q1.y         material index
//   q1.z         kind | boundary-only flag << 8 | CTA shade queue << 9 | shading class << 12 | xform chain << 16
//   q1.w         flat index (position in RtFlatScene.prims == the reference's visitation order)
struct alignas(32) DevPrim {
    float g0, g1, g2, g3;
    uint32_t x, mat, kfx, flat;
};
static_assert(sizeof(DevPrim) == 32, "DevPrim must be 32 bytes");

// Motion record for RT_PRIM_MOVING_SPHERE: centre(t) = c0 + ((t - time0) / dt) * dc
struct alignas(16) DevMotion {
    float dcx, dcy, dcz, time0;
    float dt, pad0, pad1, pad2;
};

struct alignas(16) DevXformOp {
    uint32_t kind;
    float v0, v1, v2;  // translate: offset; rotate-y: sin, cos, angle
};
struct DevXformChain {
    uint32_t first_op, n_ops;
};

// 32-byte material / texture records (same content as RtMaterial / RtTexture)
struct alignas(32) DevMaterial {
    uint32_t kind, tex0, tex1, reserved;
    float p0, p1, p2, p3;
};
struct alignas(32) DevTexture {
    uint32_t kind, a, b, reserved;
    float p0, p1, p2, p3;
};

}  // namespace rtb
