//! Raw FFI declarations for librtgpu.so — a line-by-line mirror of `include/rtgpu.h` (ABI version 2).
//!
//! UNCOMPILED: the image this crate was written in has no Rust toolchain.  It is not unchecked, though:
//! `tests/test_rust_bindings.py` parses every `#[repr(C)]` struct and every `extern "C"` function below and compares
//! names, field order, field types and argument lists with the C header, so the two cannot drift apart silently.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

pub const RT_ABI_VERSION: u32 = 2;

// RtStatus
pub const RT_OK: c_int = 0;
pub const RT_ERR_ARG: c_int = -1;
pub const RT_ERR_CUDA: c_int = -2;
pub const RT_ERR_IO: c_int = -3;
pub const RT_ERR_UNSUPPORTED: c_int = -4;
pub const RT_ERR_NCCL: c_int = -5;

// RtPrimKind
pub const RT_PRIM_SPHERE: u32 = 0;
pub const RT_PRIM_XYRECT: u32 = 1;
pub const RT_PRIM_XZRECT: u32 = 2;
pub const RT_PRIM_YZRECT: u32 = 3;
pub const RT_PRIM_MOVING_SPHERE: u32 = 4;
pub const RT_PRIM_MEDIUM: u32 = 5;
pub const RT_PRIM_FLAG_BOUNDARY_ONLY: u32 = 1;

// RtXformKind
pub const RT_XF_TRANSLATE: u32 = 0;
pub const RT_XF_ROTATE_Y: u32 = 1;

// RtMaterialKind
pub const RT_MAT_EMISSION: u32 = 0;
pub const RT_MAT_DIFFUSE: u32 = 1;
pub const RT_MAT_LAMBERT: u32 = 2;
pub const RT_MAT_METAL: u32 = 3;
pub const RT_MAT_DIELECTRIC: u32 = 4;
pub const RT_MAT_ISOTROPIC: u32 = 5;
pub const RT_MAT_OREN_NAYAR: u32 = 6;
pub const RT_MAT_BURLEY_DIFFUSE: u32 = 7;
pub const RT_MAT_ROUGH_PLASTIC: u32 = 8;
pub const RT_MAT_DISNEY_DIFFUSE: u32 = 9;
pub const RT_MAT_DISNEY_METAL: u32 = 10;
pub const RT_MAT_DISNEY_SHEEN: u32 = 11;
pub const RT_MAT_DISNEY_CLEARCOAT: u32 = 12;

// RtTextureKind
pub const RT_TEX_CONSTANT: u32 = 0;
pub const RT_TEX_CHECKER: u32 = 1;
pub const RT_TEX_PERLIN: u32 = 2;
pub const RT_TEX_IMAGE: u32 = 3;

// RtSkyKind
pub const RT_SKY_GRADIENT: u32 = 0;
pub const RT_SKY_BLACK: u32 = 1;
pub const RT_SKY_ENVMAP: u32 = 2;
pub const RT_SKY_CONSTANT: u32 = 3;

// RtRenderMode
pub const RT_MODE_AUTO: u32 = 0;
pub const RT_MODE_MEGAKERNEL: u32 = 1;
pub const RT_MODE_CTA_WAVEFRONT: u32 = 3;

pub const RT_SCENE_GPU_BVH: u32 = 1;
pub const RT_RENDER_COUNTERS: u32 = 1;
pub const RT_RENDER_RUSSIAN_ROULETTE: u32 = 2;
pub const RT_RENDER_COSINE_SAMPLING: u32 = 4;
pub const RT_RENDER_LOBE_SAMPLING: u32 = 8;
pub const RT_COMM_UNIQUE_ID_BYTES: usize = 128;
pub const RT_STATIC_MAX_RAYS: u32 = 65536;

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtPrim {
    pub kind: u32,
    pub mat: u32,
    pub xform: u32,
    pub flags: u32,
    pub object: u32,
    pub aux0: u32,
    pub aux1: u32,
    pub reserved: u32,
    pub g: [f32; 12],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtXformOp {
    pub kind: u32,
    pub v: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtXformChain {
    pub first_op: u32,
    pub n_ops: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtMaterial {
    pub kind: u32,
    pub tex0: u32,
    pub tex1: u32,
    pub reserved: u32,
    pub p: [f32; 4],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtTexture {
    pub kind: u32,
    pub a: u32,
    pub b: u32,
    pub reserved: u32,
    pub p: [f32; 4],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtPerlin {
    pub rand_vec: [[f32; 4]; 256],
    pub perm_x: [u8; 256],
    pub perm_y: [u8; 256],
    pub perm_z: [u8; 256],
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtImage {
    pub width: u32,
    pub height: u32,
    pub rgb8: *const u8,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtSky {
    pub kind: u32,
    pub image: u32,
    pub color: [f32; 3],
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtCamera {
    pub origin: [f32; 3],
    pub horizontal: [f32; 3],
    pub vertical: [f32; 3],
    pub lower_left_corner: [f32; 3],
    pub u: [f32; 3],
    pub v: [f32; 3],
    pub lens_radius: f32,
    pub time0: f32,
    pub time1: f32,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct RtFlatScene {
    pub abi_version: u32,
    pub n_prims: u32,
    pub n_xform_ops: u32,
    pub n_xform_chains: u32,
    pub n_materials: u32,
    pub n_textures: u32,
    pub n_perlin: u32,
    pub n_images: u32,
    pub prims: *const RtPrim,
    pub xform_ops: *const RtXformOp,
    pub xform_chains: *const RtXformChain,
    pub materials: *const RtMaterial,
    pub textures: *const RtTexture,
    pub perlin: *const RtPerlin,
    pub images: *const RtImage,
    pub sky: RtSky,
    pub camera: RtCamera,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtBvhInfo {
    pub n_nodes: u32,
    pub n_world_prims: u32,
    pub max_depth: u32,
    pub max_leaf_prims: u32,
    pub sah_cost: f32,
    pub node_bytes: u32,
    pub builder: u32,
    pub build_ms: f32,
    pub upload_bytes: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtHit {
    pub prim: i32,
    pub t: f32,
    pub p: [f32; 3],
    pub n: [f32; 3],
    pub uv: [f32; 2],
    pub front_face: u32,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtRenderParams {
    pub width: u32,
    pub height: u32,
    pub spp_total: u32,
    pub sample_begin: u32,
    pub sample_end: u32,
    pub max_depth: u32,
    pub mode: u32,
    pub flags: u32,
    pub seed: u64,
}

#[repr(C)]
#[derive(Clone, Copy, Default)]
pub struct RtStats {
    pub paths: u64,
    pub rays: u64,
    pub node_visits: u64,
    pub box_tests: u64,
    pub prim_tests: u64,
    pub shade_calls: u64,
    pub ms_kernel: f64,
    pub ms_total: f64,
    pub launches: u32,
    pub mode_used: u32,
    pub class_counts: [u64; 8],
}

/// Opaque handles (owned by the library).
pub enum RtScene {}
pub enum RtComm {}
pub enum RtHostScene {}

extern "C" {
    pub fn rt_scene_create(desc: *const RtFlatScene, device: c_int, out: *mut *mut RtScene) -> c_int;
    pub fn rt_scene_create_ex(desc: *const RtFlatScene, device: c_int, flags: u32, out: *mut *mut RtScene) -> c_int;
    pub fn rt_scene_destroy(scene: *mut RtScene);
    pub fn rt_scene_bvh_info(scene: *const RtScene, out: *mut RtBvhInfo) -> c_int;
    pub fn rt_scene_kernel_times(scene: *mut RtScene, out_ms: *mut f64, capacity: c_int, out_n: *mut c_int) -> c_int;
    pub fn rt_flat_validate(desc: *const RtFlatScene) -> c_int;
    pub fn rt_release_cached_memory();
    pub fn rt_flat_shade_queues(desc: *const RtFlatScene, mat_queue: *mut u8, q_class: *mut u32, q_order: *mut u32) -> c_int;
    pub fn rt_closest_hit(scene: *mut RtScene, o_xyz: *const f32, d_xyz: *const f32, times: *const f32, n: usize, t_min: f32, t_max: f32, out: *mut RtHit) -> c_int;
    pub fn rt_closest_hit_device(scene: *mut RtScene, o_xyz: *const f32, d_xyz: *const f32, times: *const f32, n: usize, t_min: f32, t_max: f32, out: *mut RtHit, cuda_stream: *mut c_void) -> c_int;
    pub fn rt_closest_hit_static(scene: *mut RtScene, o_xyz: *const f32, d_xyz: *const f32, times: *const f32, n: usize, t_min: f32, t_max: f32, out: *mut RtHit, out_n_hits: *mut u64) -> c_int;
    pub fn rt_camera_get_rays(cam: *const RtCamera, uv: *const f32, n: usize, o_xyz: *mut f32, d_xyz: *mut f32) -> c_int;
    pub fn rt_render_device(scene: *mut RtScene, params: *const RtRenderParams, accum_sum_rgb: *mut f32, cuda_stream: *mut c_void, stats: *mut RtStats) -> c_int;
    pub fn rt_render(scene: *mut RtScene, params: *const RtRenderParams, out_rgb_mean: *mut f32, stats: *mut RtStats) -> c_int;
    pub fn rt_resolve_rgb8(scene: *mut RtScene, rgb_mean: *const f32, width: u32, height: u32, out_rgb8: *mut u8) -> c_int;
    pub fn rt_write_png(path: *const c_char, rgb8: *const u8, width: u32, height: u32) -> c_int;
    pub fn rt_partition_samples(sample_begin: u32, sample_end: u32, n_ranks: c_int, rank: c_int, out_begin: *mut u32, out_end: *mut u32) -> c_int;
    pub fn rt_comm_unique_id(out_id: *mut u8) -> c_int;
    pub fn rt_comm_create(id: *const u8, n_ranks: c_int, rank: c_int, device: c_int, out: *mut *mut RtComm) -> c_int;
    pub fn rt_comm_create_all(devices: *const c_int, n_devices: c_int, out: *mut *mut RtComm) -> c_int;
    pub fn rt_comm_info(comm: *const RtComm, n_ranks: *mut c_int, first_rank: *mut c_int, n_local: *mut c_int) -> c_int;
    pub fn rt_comm_destroy(comm: *mut RtComm);
    pub fn rt_render_reduce(scenes: *const *mut RtScene, comm: *mut RtComm, frame: *const RtRenderParams, out_rgb_mean: *mut f32, stats: *mut RtStats) -> c_int;
    pub fn rt_render_reduce_device(scenes: *const *mut RtScene, comm: *mut RtComm, frame: *const RtRenderParams, accum: *const *mut f32, streams: *const *mut c_void) -> c_int;
    pub fn rt_measure_fp32_peak(device: c_int, out_ffma_per_s: *mut f64, out_sm_mhz_effective: *mut f64) -> c_int;
    pub fn rt_last_error() -> *const c_char;
    pub fn rt_host_scene_build(name: *const c_char, aspect_ratio: f32, asset_dir: *const c_char, out: *mut *mut RtHostScene) -> c_int;
    pub fn rt_host_scene_flat(hs: *const RtHostScene) -> *const RtFlatScene;
    pub fn rt_host_scene_destroy(hs: *mut RtHostScene);
}
