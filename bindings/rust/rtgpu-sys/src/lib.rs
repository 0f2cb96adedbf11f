//! Raw FFI declarations for librtgpu.so (mirror of include/rtgpu.h).
// UNTESTED
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_int, c_void};

#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtPrim { pub kind: u32, pub mat: u32, pub xform: u32, pub flags: u32, pub object: u32,
                    pub aux0: u32, pub aux1: u32, pub reserved: u32, pub g: [f32; 12] }
#[repr(C)] #[derive(Clone, Copy, Default)] pub struct RtXformOp { pub kind: u32, pub v: [f32; 3] }
#[repr(C)] #[derive(Clone, Copy, Default)] pub struct RtXformChain { pub first_op: u32, pub n_ops: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtMaterial { pub kind: u32, pub tex0: u32, pub tex1: u32, pub reserved: u32, pub p: [f32; 4] }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtTexture { pub kind: u32, pub a: u32, pub b: u32, pub reserved: u32, pub p: [f32; 4] }
#[repr(C)] pub struct RtPerlin { pub rand_vec: [[f32; 4]; 256], pub perm_x: [u8; 256], pub perm_y: [u8; 256], pub perm_z: [u8; 256] }
#[repr(C)] pub struct RtImage { pub width: u32, pub height: u32, pub rgb8: *const u8 }
#[repr(C)] #[derive(Clone, Copy, Default)] pub struct RtSky { pub kind: u32, pub image: u32, pub color: [f32; 3] }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtCamera { pub origin: [f32; 3], pub horizontal: [f32; 3], pub vertical: [f32; 3],
                      pub lower_left_corner: [f32; 3], pub u: [f32; 3], pub v: [f32; 3],
                      pub lens_radius: f32, pub time0: f32, pub time1: f32 }
#[repr(C)]
pub struct RtFlatScene { pub abi_version: u32, pub n_prims: u32, pub n_xform_ops: u32, pub n_xform_chains: u32,
    pub n_materials: u32, pub n_textures: u32, pub n_perlin: u32, pub n_images: u32,
    pub prims: *const RtPrim, pub xform_ops: *const RtXformOp, pub xform_chains: *const RtXformChain,
    pub materials: *const RtMaterial, pub textures: *const RtTexture, pub perlin: *const RtPerlin,
    pub images: *const RtImage, pub sky: RtSky, pub camera: RtCamera }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtHit { pub prim: i32, pub t: f32, pub p: [f32; 3], pub n: [f32; 3], pub uv: [f32; 2], pub front_face: u32 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtRenderParams { pub width: u32, pub height: u32, pub spp_total: u32, pub sample_begin: u32,
    pub sample_end: u32, pub max_depth: u32, pub mode: u32, pub flags: u32, pub seed: u64 }
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct RtStats { pub paths: u64, pub rays: u64, pub node_visits: u64, pub box_tests: u64, pub prim_tests: u64,
    pub shade_calls: u64, pub ms_kernel: f64, pub ms_total: f64, pub launches: u32, pub mode_used: u32,
    pub class_counts: [u64; 8] }
pub enum RtScene {}

extern "C" {
    pub fn rt_scene_create(desc: *const RtFlatScene, device: c_int, out: *mut *mut RtScene) -> c_int;
    pub fn rt_scene_create_ex(desc: *const RtFlatScene, device: c_int, flags: u32, out: *mut *mut RtScene) -> c_int; // RT_SCENE_GPU_BVH = 1
    pub fn rt_scene_destroy(scene: *mut RtScene);
    pub fn rt_closest_hit(scene: *mut RtScene, o_xyz: *const f32, d_xyz: *const f32, times: *const f32, n: usize,
                          t_min: f32, t_max: f32, out: *mut RtHit) -> c_int;
    pub fn rt_render(scene: *mut RtScene, params: *const RtRenderParams, out_rgb_mean: *mut f32, stats: *mut RtStats) -> c_int;
    pub fn rt_render_device(scene: *mut RtScene, params: *const RtRenderParams, accum_sum_rgb: *mut f32,
                            cuda_stream: *mut c_void, stats: *mut RtStats) -> c_int;
    pub fn rt_resolve_rgb8(scene: *mut RtScene, rgb_mean: *const f32, width: u32, height: u32, out_rgb8: *mut u8) -> c_int;
    pub fn rt_write_png(path: *const c_char, rgb8: *const u8, width: u32, height: u32) -> c_int;
    pub fn rt_last_error() -> *const c_char;
}
