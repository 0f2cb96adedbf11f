/*
 * rtgpu.h — C ABI of the B200-native path-tracing backend (librtgpu.so).
 *
 * This is the drop-in boundary for ONE hot path of zhouhang95/ray_tracing_in_one_weekend:
 * closest-hit ray/sphere/rect intersection through a BVH plus the per-sample integration loop.
 * The reference has no FFI seam of its own (single-crate CPU Rust binary); the seam chosen is the
 * call pair `cam.get_ray` + `ray_color(r, &world, 0)` inside the sample loop (src/main.rs:84-106)
 * and, for the deterministic batch check, `world.hit(&r, 1e-3, f32::MAX, &mut rec)`
 * (src/main.rs:44).  Every entry point below cites the reference interface it replaces.
 *
 * Conventions
 *   - plain C, POD structs, caller-owned host memory unless a name ends in `_device`;
 *   - every function returns 0 on success and a negative RtStatus on error; the message is
 *     available from rt_last_error() (thread-local).  Nothing panics/aborts across the boundary
 *     (the reference's convention is unwrap()/panic — src/main.rs:47,48,58,63,101 — which must
 *     not be mirrored across FFI);
 *   - all geometry/shading arithmetic is IEEE binary32, in the reference's operation order.
 *   - there is NO CPU fallback: a call on a machine without a usable CUDA device fails with
 *     RT_ERR_CUDA.
 *   - THREADING: a handle (RtScene, RtComm) allows ONE call in flight at a time — the per-scene statistics and
 *     work counters are shared by its renders; different handles may be used from different threads.  The
 *     reference's objects are Send + Sync and shared read-only across its workers (src/hitable.rs:50,
 *     src/material.rs:10, src/texture.rs:11); its only mutable state is the per-thread RNG (src/lib.rs:7-9), which
 *     becomes a per-path counter-seeded stream here.
 *   - the library reads no environment variables.
 */
#ifndef RTGPU_H
#define RTGPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_ABI_VERSION 2u

typedef enum RtStatus {
    RT_OK = 0,
    RT_ERR_ARG = -1,      /* bad argument / inconsistent scene description */
    RT_ERR_CUDA = -2,     /* CUDA runtime error (no device, OOM, launch failure) */
    RT_ERR_IO = -3,       /* asset file missing/unreadable (host scene builders), PNG file not writable */
    RT_ERR_UNSUPPORTED = -4, /* a scene the backend cannot render the way the reference would (see rt_scene_create) */
    RT_ERR_NCCL = -5      /* NCCL error (communicator setup, reduce) */
} RtStatus;

/* ------------------------------------------------------------------------------------------------
 * Flattened scene description ("RtFlatScene"): what the reference's trait-object graph
 * (Vec<Arc<dyn Hitable>> + Arc<dyn Material> + Arc<dyn Texture> + Camera + SKY_COLOR) becomes
 * after flattening.  Primitive ORDER is semantic: index i is the position of the primitive in the
 * reference's left-to-right visitation order (HitableList order; for a BvhNode, the order of its
 * object slice after BvhNode::new's in-place sorts — src/hitable.rs:177-221), because the
 * reference accepts a candidate when t <= closest_so_far, i.e. among exactly equal t the LATER
 * visited primitive wins (src/hitable.rs:86-91,117-132,254,294,334).
 * ---------------------------------------------------------------------------------------------- */

typedef enum RtPrimKind {
    RT_PRIM_SPHERE = 0,        /* src/hitable.rs:56-112   g = {cx,cy,cz,r} */
    RT_PRIM_XYRECT = 1,        /* src/hitable.rs:244-282  g = {min.xyz, max.xyz}; plane z = min.z */
    RT_PRIM_XZRECT = 2,        /* src/hitable.rs:284-322  plane y = min.y */
    RT_PRIM_YZRECT = 3,        /* src/hitable.rs:324-362  plane x = min.x */
    RT_PRIM_MOVING_SPHERE = 4, /* EXTENSION (not in the reference snapshot; RTNW section 2)
                                  g = {c0.xyz, r, c1.xyz, time0, time1} */
    RT_PRIM_MEDIUM = 5         /* src/hitable.rs:523-588  g = {neg_inv_density};
                                  aux0/aux1 = first/count of its boundary primitives (a HitableList, scanned in
                                  order).  `xform` = the Translate/RotateY wrappers AROUND the medium; every boundary
                                  primitive's chain starts with the same ops (a wrapper of the medium wraps its
                                  boundary too) followed by the boundary's own wrappers. */
} RtPrimKind;

#define RT_PRIM_FLAG_BOUNDARY_ONLY 1u /* only reachable through a RT_PRIM_MEDIUM, not in the world */

typedef struct RtPrim {
    uint32_t kind;    /* RtPrimKind */
    uint32_t mat;     /* material index (MEDIUM: its Isotropic phase-function material) */
    uint32_t xform;   /* transform-chain index; 0 is always the identity chain */
    uint32_t flags;
    uint32_t object;  /* index of the top-level reference object this primitive came from */
    uint32_t aux0;
    uint32_t aux1;
    uint32_t reserved;
    float g[12];
} RtPrim; /* 80 bytes */

typedef enum RtXformKind {
    RT_XF_TRANSLATE = 0, /* src/hitable.rs:404-436  v = offset */
    RT_XF_ROTATE_Y = 1   /* src/hitable.rs:438-520  v = {sin_theta, cos_theta, angle_degrees} */
} RtXformKind;

typedef struct RtXformOp {
    uint32_t kind;
    float v[3];
} RtXformOp;

/* ops [first_op, first_op+n_ops) are listed OUTERMOST wrapper first, i.e. in the order the
 * reference applies them to the incoming ray. */
typedef struct RtXformChain {
    uint32_t first_op;
    uint32_t n_ops;
} RtXformChain;

typedef enum RtMaterialKind {
    RT_MAT_EMISSION = 0,         /* src/material.rs:17-29   tex0 = emit */
    RT_MAT_DIFFUSE = 1,          /* src/material.rs:31-46   tex0 = albedo (cosine weighted) */
    RT_MAT_LAMBERT = 2,          /* src/material.rs:48-59   tex0 = albedo (uniform hemisphere) */
    RT_MAT_METAL = 3,            /* src/material.rs:61-73   p = {albedo.rgb, fuzz} */
    RT_MAT_DIELECTRIC = 4,       /* src/material.rs:75-97   p = {ior} */
    RT_MAT_ISOTROPIC = 5,        /* src/material.rs:99-113  tex0 = albedo */
    RT_MAT_OREN_NAYAR = 6,       /* src/pbr.rs:12-42        tex0, p = {roughness} */
    RT_MAT_BURLEY_DIFFUSE = 7,   /* src/pbr.rs:45-69        tex0, p = {roughness} */
    RT_MAT_ROUGH_PLASTIC = 8,    /* src/pbr.rs:153-189      tex0 = spec, tex1 = diff, p = {roughness, eta} */
    RT_MAT_DISNEY_DIFFUSE = 9,   /* src/pbr.rs:192-223      tex0, p = {roughness, subsurface} */
    RT_MAT_DISNEY_METAL = 10,    /* src/pbr.rs:224-278      tex0, p = {roughness, anisotropic, rot} */
    RT_MAT_DISNEY_SHEEN = 11,    /* src/pbr.rs:281-308      tex0, p = {tint} */
    RT_MAT_DISNEY_CLEARCOAT = 12 /* src/pbr.rs:310-335      p = {clearcoat_gloss} */
} RtMaterialKind;

typedef struct RtMaterial {
    uint32_t kind;
    uint32_t tex0;
    uint32_t tex1;
    uint32_t reserved;
    float p[4];
} RtMaterial; /* 32 bytes */

typedef enum RtTextureKind {
    RT_TEX_CONSTANT = 0, /* src/texture.rs:15-23    p = {r,g,b} */
    RT_TEX_CHECKER = 1,  /* src/texture.rs:25-49    a = odd texture, b = even texture */
    RT_TEX_PERLIN = 2,   /* src/texture.rs:149-168  a = perlin table index, p = {scale} */
    RT_TEX_IMAGE = 3     /* src/texture.rs:169-193  a = image index */
} RtTextureKind;

typedef struct RtTexture {
    uint32_t kind;
    uint32_t a;
    uint32_t b;
    uint32_t reserved;
    float p[4];
} RtTexture; /* 32 bytes */

/* src/texture.rs:51-91: 256 un-normalised gradients in [-1,1)^3 and three permutations of 0..255 */
typedef struct RtPerlin {
    float rand_vec[256][4]; /* xyz + pad */
    uint8_t perm_x[256];
    uint8_t perm_y[256];
    uint8_t perm_z[256];
} RtPerlin;

/* src/texture.rs:169-181: decoded image, 8-bit RGB, row 0 = top; value = u8 / 255 */
typedef struct RtImage {
    uint32_t width;
    uint32_t height;
    const uint8_t* rgb8; /* width*height*3 bytes */
} RtImage;

typedef enum RtSkyKind {
    RT_SKY_GRADIENT = 0, /* src/demo_scene.rs:28-31  lerp(1, (0.5,0.7,1), 0.5*d.y+0.5) */
    RT_SKY_BLACK = 1,    /* src/demo_scene.rs:33-35 */
    RT_SKY_ENVMAP = 2,   /* src/demo_scene.rs:22-26  equirect lookup, u mirrored, value squared */
    RT_SKY_CONSTANT = 3  /* EXTENSION for furnace tests: constant `color` */
} RtSkyKind;

typedef struct RtSky {
    uint32_t kind;
    uint32_t image; /* RT_SKY_ENVMAP: image index */
    float color[3];
} RtSky;

/* src/camera.rs:6-46.  u, v, lens_radius, time0, time1 are EXTENSIONS (thin lens: RTIOW section 12;
 * shutter: RTNW section 2); with lens_radius == 0 and time0 == time1 the camera draws no extra
 * random numbers and reduces exactly to the reference's pinhole Camera::get_ray. */
typedef struct RtCamera {
    float origin[3];
    float horizontal[3];
    float vertical[3];
    float lower_left_corner[3];
    float u[3];
    float v[3];
    float lens_radius;
    float time0;
    float time1;
} RtCamera;

typedef struct RtFlatScene {
    uint32_t abi_version; /* RT_ABI_VERSION */
    uint32_t n_prims;
    uint32_t n_xform_ops;
    uint32_t n_xform_chains; /* >= 1; chain 0 must be the identity */
    uint32_t n_materials;
    uint32_t n_textures;
    uint32_t n_perlin;
    uint32_t n_images;
    const RtPrim* prims;
    const RtXformOp* xform_ops;
    const RtXformChain* xform_chains;
    const RtMaterial* materials;
    const RtTexture* textures;
    const RtPerlin* perlin;
    const RtImage* images;
    RtSky sky;
    RtCamera camera;
} RtFlatScene;

/* ------------------------------------------------------------------------------------------------
 * Device scene handle
 * ---------------------------------------------------------------------------------------------- */
typedef struct RtScene RtScene; /* opaque: owns device memory, texture objects, streams */

/* Uploads `desc` to CUDA device `device` (copies everything; the caller may free `desc` on return) — every table
 * in ONE device allocation filled by ONE host-to-device copy from a pinned staging buffer — and
 * builds the 4-wide BVH (128-byte, 32-byte-aligned nodes) over the world primitives: binned SAH on the host
 * below 262144 world primitives, LBVH on the device from there on (see rt_scene_create_ex).
 * Replaces: the scene hand-off `let (world, cam) = scene_fn(aspect)` + `build_bvh`
 * (src/main.rs:74, src/demo_scene.rs:223-227, src/hitable.rs:177-221).
 * RT_ERR_ARG: malformed description.  RT_ERR_UNSUPPORTED: a well-formed scene whose picture in the reference depends
 * on STALE HitRecord fields (rects never write rec.tang, src/hitable.rs:262-269; media write
 * neither uv nor tang, 568-576) — DisneyMetal on anything but a sphere, an Isotropic phase function with a
 * uv-reading (image) texture on a medium — or a ConstantMedium whose boundary is itself a ConstantMedium. */
int rt_scene_create(const RtFlatScene* desc, int device, RtScene** out);

/* Same, with options.  RT_SCENE_GPU_BVH: build the acceleration structure ON THE DEVICE (Morton codes, radix
 * sort, binary radix tree, bottom-up fit, collapse to the same 4-wide nodes) instead of the host SAH builder:
 * ~1 ms instead of ~0.3 s at 100k primitives, for scenes that are very large or rebuilt every frame; the SAH tree
 * traverses faster (by 15-20 % on the moving-sphere cover scenes) and stays the default.  Results are identical either way (any conservative tree yields the
 * same closest hits).  Falls back to the host builder when the Morton tree is too deep for the traversal stack. */
#define RT_SCENE_GPU_BVH 1u
int rt_scene_create_ex(const RtFlatScene* desc, int device, uint32_t flags, RtScene** out);
void rt_scene_destroy(RtScene* scene);

typedef struct RtBvhInfo {
    uint32_t n_nodes;
    uint32_t n_world_prims;
    uint32_t max_depth;
    uint32_t max_leaf_prims;
    float sah_cost;       /* host SAH builder only */
    uint32_t node_bytes;
    uint32_t builder;     /* 0 = host SAH, 1 = GPU LBVH */
    float build_ms;       /* wall clock of the build inside rt_scene_create */
    uint64_t upload_bytes; /* size of the ONE host-to-device copy that carried every table of the scene */
} RtBvhInfo;
int rt_scene_bvh_info(const RtScene* scene, RtBvhInfo* out);

/* CUDA-event durations (ms) of the render kernels of the last n <= min(capacity, 32) rt_render_device /
 * rt_render_reduce* calls on this scene, oldest first — measured on the stream each kernel ran on; waits for those
 * kernels to finish.  Lets a caller time the kernel of an asynchronous frame loop without a synchronisation per frame. */
int rt_scene_kernel_times(RtScene* scene, double* out_ms, int capacity, int* out_n);

/* The checks rt_scene_create runs on a description, without a device: RT_OK, RT_ERR_ARG or RT_ERR_UNSUPPORTED. */
int rt_flat_validate(const RtFlatScene* desc);

/* Device buffers and pinned host staging buffers are kept in a process-wide cache between calls (a host that
 * creates a scene per frame does not pay the driver allocator every time); this frees whatever is idle. */
void rt_release_cached_memory(void);

/* Host-only planning step of rt_scene_create, exposed for inspection and tests (no device needed): how the
 * CTA-wavefront renderer's shade queues are assigned for this scene.  The renderer sorts the hits of an iteration
 * into 7 queues per block before shading them (the "material-sorted shade pass"): queue 0 takes the misses
 * (main.rs:58), every shading class of the scene gets one, and the queues the scene leaves free go to its heaviest
 * (material kind, texture kind) groups, so that a warp shades ONE kind of material (src/material.rs, src/pbr.rs)
 * wherever the scene allows it.  mat_queue[n_materials]: queue of every material; *q_class / *q_order: 4 bits per
 * queue — the shading class queue q runs, and the order the queues are drained in (hit queues are numbered most
 * expensive class first, so this is always 1, 2, ..., 6, 0). */
int rt_flat_shade_queues(const RtFlatScene* desc, uint8_t* mat_queue, uint32_t* q_class, uint32_t* q_order);

/* Result of one closest-hit query == the fields of the reference's HitRecord that a primitive
 * writes (src/hitable.rs:13-22).  prim = -1 on a miss (then the other fields are 0). */
typedef struct RtHit {
    int32_t prim; /* index into RtFlatScene.prims */
    float t;
    float p[3];
    float n[3]; /* face-forwarded normal (rec.norm) */
    float uv[2];
    uint32_t front_face;
} RtHit; /* 44 bytes */

/* Batch closest hit over HOST buffers. o_xyz/d_xyz are n*3 floats (AoS).
 * Replaces `world.hit(&r, t_min, t_max, &mut rec)` (src/main.rs:44 -> src/hitable.rs:117-132,232-240).
 * `times` may be NULL (ray time 0).  RT_PRIM_MEDIUM primitives are stochastic in the reference
 * (src/hitable.rs:564) and are skipped by this deterministic query. */
int rt_closest_hit(RtScene* scene, const float* o_xyz, const float* d_xyz, const float* times,
                   size_t n, float t_min, float t_max, RtHit* out);
/* Same, all pointers are device pointers on the scene's device; asynchronous on `cuda_stream`
 * (a cudaStream_t passed as void*, NULL = the scene's own non-blocking stream; the legacy default stream is
 * cudaStreamLegacy, i.e. (void*)0x1). */
int rt_closest_hit_device(RtScene* scene, const float* o_xyz, const float* d_xyz,
                          const float* times, size_t n, float t_min, float t_max, RtHit* out,
                          void* cuda_stream);

/* Same query as rt_closest_hit (HOST buffers in and out, same records), allocating NOTHING: the device staging of
 * the rays and of the hit records lives in a static device array inside the library (4.5 MiB, held by every GPU that
 * loads the library until the process exits) instead of leased / cudaMalloc'ed scratch.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * At most RT_STATIC_MAX_RAYS rays per call: a larger n is rejected with RT_ERR_ARG (split the batch, or call
 * rt_closest_hit).  *out_n_hits (optional) = rays that hit something.  Synchronous on the scene's own stream.
 * Every call on a device shares the one array: calls that overlap on the same GPU (other host threads, other scenes
 * on that device) race — serialise them. */
#define RT_STATIC_MAX_RAYS 65536u
int rt_closest_hit_static(RtScene* scene, const float* o_xyz, const float* d_xyz, const float* times,
                          size_t n, float t_min, float t_max, RtHit* out, uint64_t* out_n_hits);

/* Camera rays for pixel centres or explicit (u,v): replaces Camera::get_ray (src/camera.rs:40-46)
 * on the host side so that GPU and oracle are fed bit-identical rays. uv is n*2 floats. */
int rt_camera_get_rays(const RtCamera* cam, const float* uv, size_t n, float* o_xyz, float* d_xyz);

typedef enum RtRenderMode {
    RT_MODE_AUTO = 0,       /* = RT_MODE_CTA_WAVEFRONT, the fastest on every measured scene */
    RT_MODE_MEGAKERNEL = 1, /* persistent per-thread path loop with in-warp path regeneration; kept as the
                               independent cross-check of the default renderer (same rays, other scheduling) */
    /* 2 was the global-memory wavefront renderer of ABI 1 (1.7x slower everywhere; removed) */
    RT_MODE_CTA_WAVEFRONT = 3 /* one persistent kernel; every thread block runs the wavefront loop (regenerate,
                                 traverse, material-sorted shade, compaction) on a path pool and queues held in
                                 its shared memory */
} RtRenderMode;

typedef struct RtRenderParams {
    uint32_t width;
    uint32_t height;
    uint32_t spp_total;    /* total samples per pixel of the frame (divisor of the mean) */
    uint32_t sample_begin; /* this call renders global sample indices [sample_begin, sample_end) */
    uint32_t sample_end;   /*   of EVERY pixel — the multi-GPU partition (scene replicated) */
    uint32_t max_depth;    /* reference: MAX_DEPTH = 50 (src/main.rs:36) -> at most 51 rays/path */
    uint32_t mode;         /* RtRenderMode */
    uint32_t flags;        /* RT_RENDER_* */
    uint64_t seed;
} RtRenderParams;

#define RT_RENDER_COUNTERS 1u /* also count BVH node visits / primitive tests (slower build of the kernels) */
#define RT_RENDER_COSINE_SAMPLING 4u /* OPT-IN estimator, outside the parity contract: Lambert and every pbr.rs
                                        material draw the scattered direction cosine-weighted (pdf cos/pi) instead
                                        of uniformly on the hemisphere (src/material.rs:52-59, src/pbr.rs) and
                                        re-weight by 1/(2 cos).  Same expectation, less variance. */
#define RT_RENDER_LOBE_SAMPLING 8u /* OPT-IN estimator, outside the parity contract (takes precedence over
                                      RT_RENDER_COSINE_SAMPLING): every pbr.rs material draws the scattered direction
                                      from a 50/50 mixture of cosine-weighted and GGX-half-vector sampling of its own
                                      lobe width instead of uniformly on the hemisphere (src/pbr.rs:19,53,163,201,235,
                                      289,317) and re-weights by (1/2pi)/pdf; Lambert draws cosine-weighted.  Same
                                      expectation, bounded weights on the sharp lobes (no fireflies / NaN pixels). */
#define RT_RENDER_RUSSIAN_ROULETTE 2u /* OPT-IN estimator, outside the parity contract: after every scatter the
                                         path survives with probability min(max_element(attenuation), 1) and its
                                         throughput is divided by it (the reference has this commented out,
                                         src/main.rs:49-53).  Same expectation, fewer rays per path, other noise. */

typedef struct RtStats {
    uint64_t paths;        /* depth-0 ray_color calls == width*height*(sample_end-sample_begin) */
    uint64_t rays;         /* closest-hit queries == world.hit calls from ray_color (src/main.rs:44) */
    uint64_t node_visits;  /* only with RT_RENDER_COUNTERS */
    uint64_t box_tests;    /* only with RT_RENDER_COUNTERS */
    uint64_t prim_tests;   /* only with RT_RENDER_COUNTERS */
    uint64_t shade_calls;  /* rays that hit something */
    double ms_kernel;      /* CUDA-event time of the render kernels of this call */
    double ms_total;       /* host wall clock of the call, including copies */
    uint32_t launches;     /* kernels launched by this call */
    uint32_t mode_used;
    uint64_t class_counts[8]; /* only with RT_RENDER_COUNTERS in RT_MODE_CTA_WAVEFRONT: rays per shading class
                                 (0 miss/sky, 1 Diffuse(ConstantTex), 2 Metal, 3 Dielectric, 4 Diffuse(PerlinTex),
                                 5 other material.rs combinations, 6 pbr.rs materials) */
} RtStats;

/* Renders global samples [sample_begin, sample_end) and ADDS the radiance SUMS (not means) to the
 * device accumulation buffer `accum_sum_rgb` (width*height*3 floats, pixel (i,j) at (j*width+i)*3,
 * j = 0 is the BOTTOM row as in src/main.rs:84-105).  Asynchronous on `cuda_stream` unless
 * `stats` is non-NULL (then the call synchronises to fill it).
 * Replaces the sample loop + ray_color (src/main.rs:38-60,84-97). */
int rt_render_device(RtScene* scene, const RtRenderParams* params, float* accum_sum_rgb,
                     void* cuda_stream, RtStats* stats);

/* Host-buffer call: zero an internal accumulation buffer, render the sample range, divide
 * by spp_total (src/main.rs:98) and copy the linear float image to `out_rgb_mean`
 * (width*height*3 floats).  No gamma, no quantisation.  A page-locked `out_rgb_mean` (cudaMallocHost /
 * cudaHostRegister) is written by one direct DMA; a pageable one through a pinned staging pipeline. */
int rt_render(RtScene* scene, const RtRenderParams* params, float* out_rgb_mean, RtStats* stats);

/* Reference output stage (src/main.rs:98-105,127): c = mean; c = c.powf(0.5); (c*255.99) as u8
 * (saturating, NaN -> 0); vertical flip so row 0 is the TOP of the picture.  Host buffers. */
int rt_resolve_rgb8(RtScene* scene, const float* rgb_mean, uint32_t width, uint32_t height,
                    uint8_t* out_rgb8);

/* Writes an 8-bit RGB image (row 0 = top, as rt_resolve_rgb8 produces it) as a PNG file: what
 * image::save_buffer(file_name, &buf, nx, ny, ColorType::Rgb8) does at src/main.rs:128.
 * Host only (zlib deflate); RT_ERR_IO when the file cannot be written. */
int rt_write_png(const char* path, const uint8_t* rgb8, uint32_t width, uint32_t height);

/* ------------------------------------------------------------------------------------------------
 * Multi-GPU: scene replicated, samples of every pixel partitioned over the ranks, ONE NCCL sum-reduce of the
 * float32 accumulation buffers to rank 0 per frame (NVLink 5 / NVSwitch) — replaces the reference's mpsc gather
 * of finished pixels (src/main.rs:72,101,114); paths are independent (src/main.rs:84-106), nothing else is
 * exchanged.  Because a path's random stream depends only on (seed, pixel, global sample index), the N-GPU frame
 * equals the 1-GPU frame up to float summation order.
 * ---------------------------------------------------------------------------------------------- */
typedef struct RtComm RtComm; /* opaque: owns one NCCL communicator per local device */
#define RT_COMM_UNIQUE_ID_BYTES 128

/* Global sample indices [*out_begin, *out_end) rank `rank` of `n_ranks` renders out of [sample_begin, sample_end):
 * the ranges tile the interval, sizes differ by at most one. */
int rt_partition_samples(uint32_t sample_begin, uint32_t sample_end, int n_ranks, int rank,
                         uint32_t* out_begin, uint32_t* out_end);

/* One process per GPU: rank 0 calls rt_comm_unique_id and ferries the 128 bytes to the other ranks by whatever
 * means the launcher has (a file, MPI, torch.distributed); then EVERY rank calls rt_comm_create (collective). */
int rt_comm_unique_id(uint8_t out_id[RT_COMM_UNIQUE_ID_BYTES]);
int rt_comm_create(const uint8_t id[RT_COMM_UNIQUE_ID_BYTES], int n_ranks, int rank, int device, RtComm** out);
/* One process that drives n devices (ncclCommInitAll): local device i is rank i. */
int rt_comm_create_all(const int* devices, int n_devices, RtComm** out);
int rt_comm_info(const RtComm* comm, int* n_ranks, int* first_rank, int* n_local);
void rt_comm_destroy(RtComm* comm);

/* One frame on all ranks.  `frame` describes the WHOLE frame: [sample_begin, sample_end) is split with
 * rt_partition_samples, spp_total is the divisor.  scenes[i] (i < n_local) is the replica on the communicator's
 * local device i.  Every rank renders its share into a zeroed sum buffer, one ncclReduce(sum, f32, root 0) follows
 * on the same stream, rank 0 divides by spp_total and the process that owns rank 0 receives the linear float image
 * in `out_rgb_mean` (host memory, width*height*3; ignored on the other processes).  Collective: every process of the
 * communicator must make the call.  `stats` (optional) covers this process's ranks. */
int rt_render_reduce(RtScene* const* scenes, RtComm* comm, const RtRenderParams* frame, float* out_rgb_mean,
                     RtStats* stats);
/* Same without the host copy: accum[i] are caller-owned DEVICE buffers (width*height*3 floats on local device i,
 * overwritten); streams[i] (cudaStream_t as void*; NULL array or NULL entry = the scene's own stream) carry the
 * memset, the render kernel, the reduce and, on rank 0, the division.  Asynchronous.  On return from the stream,
 * accum of rank 0 holds the frame's mean. */
int rt_render_reduce_device(RtScene* const* scenes, RtComm* comm, const RtRenderParams* frame,
                            float* const* accum, void* const* streams);

/* FP32 issue-rate microbenchmark (dependent-free FFMA chains on every SM): returns measured
 * lane-FFMA/s — the measured denominator of the instruction-issue roofline. */
int rt_measure_fp32_peak(int device, double* out_ffma_per_s, double* out_sm_mhz_effective);

const char* rt_last_error(void);

/* ------------------------------------------------------------------------------------------------
 * Host-side scene construction (C++ API in csrc/scene_api.hpp mirrors the reference's
 * constructors: Sphere{c,r,mat}, XYRect{min,max,mat}, GBox::new, Translate{offset,ptr},
 * RotateY::new(ptr,angle), ConstantMedium::new, BvhNode::new, Camera::new, every Material and
 * Texture).  These C entry points expose the re-expressed demo scenes (src/demo_scene.rs) and the
 * BASELINE.json config scenes so that non-C++ callers (tests, bench) can obtain an RtFlatScene.
 * ---------------------------------------------------------------------------------------------- */
typedef struct RtHostScene RtHostScene;

/* name: "test_sphere" | "sphere_scene" | "simple_light_scene" | "cornell_box" | "cornell_box_solid"
 *       | "final_scene" | "textured" | "pbr_ibl" | "cover_defocus" | "cover_100k_motion" | "furnace_<mat>"
 * asset_dir: directory holding earthmap.ppm / newport_loft.ppm (decoded texel arrays). */
int rt_host_scene_build(const char* name, float aspect_ratio, const char* asset_dir,
                        RtHostScene** out);
const RtFlatScene* rt_host_scene_flat(const RtHostScene* hs);
void rt_host_scene_destroy(RtHostScene* hs);

#ifdef __cplusplus
}
#endif
#endif /* RTGPU_H */
