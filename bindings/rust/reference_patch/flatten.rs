//! `src/flatten.rs` for zhouhang95/ray_tracing_in_one_weekend — the "new Rust module" of the north star: scenes are
//! still built through the existing Hitable / Material / Texture / Camera API (src/demo_scene.rs), this module walks
//! that object graph once and emits the POD arrays librtgpu.so consumes (`rtgpu_sys::RtFlatScene`, include/rtgpu.h).
//!
//! UNCOMPILED (no Rust toolchain where it was written).  It is the Rust twin of the C++ flattener the backend ships
//! and tests (`ray_tracing_in_one_weekend_b200/csrc/scene_api.cpp`, `Flat::walk`): same visitation order, same chain
//! construction, same de-duplication, field for field — and that C++ flattener is checked byte for byte against the
//! oracle's restatement of src/demo_scene.rs (tests/test_host_api.py::test_host_scene_equals_oracle_restatement).
//!
//! How it hooks in (bindings/rust/reference_patch/apply.sh makes these edits):
//!   * the three traits gain a supertrait each, so that `dyn Hitable` / `dyn Material` / `dyn Texture` can be flattened
//!     without a downcast:   `pub trait Hitable: Send + Sync + FlattenHitable`   (src/hitable.rs:50),
//!     `pub trait Material: Send + Sync + FlattenMaterial` (src/material.rs:10),
//!     `pub trait Texture: Send + Sync + FlattenTexture`   (src/texture.rs:11);
//!   * the private fields this module reads become `pub(crate)`: BvhNode {left, right} (src/hitable.rs:160-161),
//!     RotateY {ptr, angle, sin_theta, cos_theta} (439-442), ConstantMedium {boundary, phase_fn, neg_inv_density}
//!     (524-526), CheckerTex {odd, even} (src/texture.rs:26-27), Perlin and its four tables (53-58), PerlinTex
//!     {perlin, scale} (150-151), ImageTex {img} (171), Camera's four vectors (src/camera.rs:7-10).
//!
//! ORDER IS SEMANTIC: primitives are emitted in the order the reference VISITS them (HitableList order; a BvhNode's
//! left subtree before its right one — BvhNode::hit always tries left first, src/hitable.rs:232-240), because the
//! reference accepts a candidate when `t <= closest_so_far`: among exactly equal t the later-visited primitive wins
//! (src/hitable.rs:86-91,117-132).  The backend reproduces that rule from the flat index.
use std::collections::HashMap;
use std::sync::Arc;

use glam::Vec3A;
use rtgpu_sys::*;

use crate::camera::Camera;
use crate::hitable::*;
use crate::material::*;
use crate::pbr::*;
use crate::texture::*;

/// Owns the arrays an `RtFlatScene` points into.
#[derive(Default)]
pub struct FlatScene {
    pub prims: Vec<RtPrim>,
    pub xform_ops: Vec<RtXformOp>,
    pub xform_chains: Vec<RtXformChain>,
    pub materials: Vec<RtMaterial>,
    pub textures: Vec<RtTexture>,
    pub perlin: Vec<RtPerlin>,
    pub images: Vec<RtImage>,
    image_store: Vec<Vec<u8>>, // RGB8 bytes the RtImage records point into
    mat_ids: HashMap<usize, u32>, // Arc data pointer -> table index
    tex_ids: HashMap<usize, u32>,
    img_ids: HashMap<usize, u32>,
    next_object: u32,
    cur_object: u32,
    depth: u32,
}

fn arr3(v: Vec3A) -> [f32; 3] {
    [v.x, v.y, v.z]
}

/// Table key of a shared object: the address of the Arc's payload (two Arcs to one material share an entry).
fn key<T: ?Sized>(a: &Arc<T>) -> usize {
    Arc::as_ptr(a) as *const () as usize
}

impl FlatScene {
    pub fn new() -> Self {
        let mut s = Self::default();
        s.xform_chains.push(RtXformChain { first_op: 0, n_ops: 0 }); // chain 0 = identity
        s
    }

    /// Flattens a world (`Vec<Arc<dyn Hitable>>` as the scene functions return it, src/demo_scene.rs).
    pub fn from_world(world: &HitableList) -> Self {
        let mut s = Self::new();
        for h in world.iter() {
            h.flatten(&mut s, 0, false);
        }
        s
    }

    /// The chain `parent` followed by `op` (ops are listed OUTERMOST wrapper first, i.e. in the order the reference
    /// applies them to the incoming ray: Translate::hit / RotateY::hit transform the ray, then call the child).
    pub fn push_chain(&mut self, parent: u32, op: RtXformOp) -> u32 {
        let pc = self.xform_chains[parent as usize];
        let first_op = self.xform_ops.len() as u32;
        for i in 0..pc.n_ops {
            let c = self.xform_ops[(pc.first_op + i) as usize];
            self.xform_ops.push(c);
        }
        self.xform_ops.push(op);
        self.xform_chains.push(RtXformChain { first_op, n_ops: pc.n_ops + 1 });
        self.xform_chains.len() as u32 - 1
    }

    pub fn push_prim(&mut self, kind: u32, mat: &Arc<dyn Material>, chain: u32, boundary_only: bool, g: &[f32]) -> u32 {
        let mut p = RtPrim::default();
        p.kind = kind;
        p.mat = self.material(mat);
        p.xform = chain;
        p.flags = if boundary_only { RT_PRIM_FLAG_BOUNDARY_ONLY } else { 0 };
        p.object = self.cur_object;
        p.g[..g.len()].copy_from_slice(g);
        self.prims.push(p);
        self.prims.len() as u32 - 1
    }

    pub fn material(&mut self, m: &Arc<dyn Material>) -> u32 {
        if let Some(&id) = self.mat_ids.get(&key(m)) {
            return id;
        }
        let rec = m.flatten(self);
        self.materials.push(rec);
        let id = self.materials.len() as u32 - 1;
        self.mat_ids.insert(key(m), id);
        id
    }

    pub fn texture(&mut self, t: &Arc<dyn Texture>) -> u32 {
        if let Some(&id) = self.tex_ids.get(&key(t)) {
            return id;
        }
        let rec = t.flatten(self);
        self.textures.push(rec);
        let id = self.textures.len() as u32 - 1;
        self.tex_ids.insert(key(t), id);
        id
    }

    /// ImageTex keeps `to_rgb32f()` texels (u8 / 255, src/texture.rs:176-177); the backend wants the bytes back:
    /// (x * 255).round() inverts u8 / 255 exactly for every u8.
    pub fn image(&mut self, t: &ImageTex) -> u32 {
        let k = t as *const ImageTex as usize;
        if let Some(&id) = self.img_ids.get(&k) {
            return id;
        }
        let bytes: Vec<u8> = t.img.as_raw().iter().map(|&x| (x * 255.0).round() as u8).collect();
        self.image_store.push(bytes);
        let rgb8 = self.image_store.last().unwrap().as_ptr(); // the Vec's heap buffer does not move when image_store grows
        self.images.push(RtImage { width: t.img.width(), height: t.img.height(), rgb8 });
        let id = self.images.len() as u32 - 1;
        self.img_ids.insert(k, id);
        id
    }

    /// Every non-aggregate hitable met at nesting depth 0 is one "object" (RtPrim.object).
    fn enter(&mut self) {
        if self.depth == 0 {
            self.cur_object = self.next_object;
            self.next_object += 1;
        }
        self.depth += 1;
    }
    fn leave(&mut self) {
        self.depth -= 1;
    }

    /// The POD view handed to `rt_scene_create`.  `sky`: the reference keeps its sky in a global fn pointer
    /// (src/lib.rs:11, src/demo_scene.rs:22-35), which cannot be inspected — the caller states which one it set
    /// (`RT_SKY_GRADIENT` = sky_color, `RT_SKY_BLACK` = black_sky, `RT_SKY_ENVMAP` = tex_sky_color + `env`).
    pub fn as_rt(&mut self, cam: &Camera, sky_kind: u32, env: Option<&ImageTex>) -> RtFlatScene {
        let mut sky = RtSky { kind: sky_kind, image: 0, color: [0.0; 3] };
        if sky_kind == RT_SKY_ENVMAP {
            sky.image = self.image(env.expect("RT_SKY_ENVMAP needs ENV_TEX"));
        }
        let camera = RtCamera {
            origin: arr3(cam.origin),
            horizontal: arr3(cam.horizontal),
            vertical: arr3(cam.vertical),
            lower_left_corner: arr3(cam.lower_left_corner),
            ..RtCamera::default() // u, v, lens_radius, time0, time1: extensions the reference camera does not have
        };
        RtFlatScene {
            abi_version: RT_ABI_VERSION,
            n_prims: self.prims.len() as u32,
            n_xform_ops: self.xform_ops.len() as u32,
            n_xform_chains: self.xform_chains.len() as u32,
            n_materials: self.materials.len() as u32,
            n_textures: self.textures.len() as u32,
            n_perlin: self.perlin.len() as u32,
            n_images: self.images.len() as u32,
            prims: self.prims.as_ptr(),
            xform_ops: self.xform_ops.as_ptr(),
            xform_chains: self.xform_chains.as_ptr(),
            materials: self.materials.as_ptr(),
            textures: self.textures.as_ptr(),
            perlin: self.perlin.as_ptr(),
            images: self.images.as_ptr(),
            sky,
            camera,
        }
    }
}

// ------------------------------------------------------------------------------------------------ hitables
pub trait FlattenHitable {
    /// Emit this object's primitives in VISITATION order under transform chain `chain`; `boundary_only` marks
    /// geometry that only bounds a ConstantMedium.
    fn flatten(&self, out: &mut FlatScene, chain: u32, boundary_only: bool);
}

impl FlattenHitable for Sphere {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        out.enter();
        out.push_prim(RT_PRIM_SPHERE, &self.mat, chain, bo, &[self.c.x, self.c.y, self.c.z, self.r]);
        out.leave();
    }
}

fn rect(out: &mut FlatScene, kind: u32, min: Vec3A, max: Vec3A, mat: &Arc<dyn Material>, chain: u32, bo: bool) {
    out.enter();
    out.push_prim(kind, mat, chain, bo, &[min.x, min.y, min.z, max.x, max.y, max.z]);
    out.leave();
}
impl FlattenHitable for XYRect {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        rect(out, RT_PRIM_XYRECT, self.min, self.max, &self.mat, chain, bo);
    }
}
impl FlattenHitable for XZRect {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        rect(out, RT_PRIM_XZRECT, self.min, self.max, &self.mat, chain, bo);
    }
}
impl FlattenHitable for YZRect {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        rect(out, RT_PRIM_YZRECT, self.min, self.max, &self.mat, chain, bo);
    }
}

/// HitableList = Vec<Arc<dyn Hitable>> (src/hitable.rs:114): an aggregate, its items in order.
impl FlattenHitable for HitableList {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        for item in self.iter() {
            item.flatten(out, chain, bo);
        }
    }
}

/// Left before right == the order BvhNode::hit visits (src/hitable.rs:232-240).  A span-1 node holds the same Arc
/// twice (src/hitable.rs:188): emit it once.
impl FlattenHitable for BvhNode {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        self.left.flatten(out, chain, bo);
        if !Arc::ptr_eq(&self.left, &self.right) {
            self.right.flatten(out, chain, bo);
        }
    }
}

/// Six rects in the order GBox::new pushes them (src/hitable.rs:369-383): the later side wins ties.
impl FlattenHitable for GBox {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        out.enter();
        self.sides.flatten(out, chain, bo);
        out.leave();
    }
}

impl FlattenHitable for Translate {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        out.enter();
        let c = out.push_chain(chain, RtXformOp { kind: RT_XF_TRANSLATE, v: arr3(self.offset) });
        self.ptr.flatten(out, c, bo);
        out.leave();
    }
}

/// v = {sin_theta, cos_theta, angle}: the values RotateY::new computed (src/hitable.rs:448-451), taken verbatim so
/// that the backend rotates with the same bits.
impl FlattenHitable for RotateY {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        out.enter();
        let c = out.push_chain(chain, RtXformOp { kind: RT_XF_ROTATE_Y, v: [self.sin_theta, self.cos_theta, self.angle] });
        self.ptr.flatten(out, c, bo);
        out.leave();
    }
}

/// The boundary first (flagged boundary-only, under the SAME chain as the medium: a wrapper of the medium wraps its
/// boundary too), then the RT_PRIM_MEDIUM record with aux0/aux1 = first/count of those primitives.
impl FlattenHitable for ConstantMedium {
    fn flatten(&self, out: &mut FlatScene, chain: u32, bo: bool) {
        out.enter();
        let first = out.prims.len() as u32;
        self.boundary.flatten(out, chain, true);
        let count = out.prims.len() as u32 - first;
        let id = out.push_prim(RT_PRIM_MEDIUM, &self.phase_fn, chain, bo, &[self.neg_inv_density]) as usize;
        out.prims[id].aux0 = first;
        out.prims[id].aux1 = count;
        out.leave();
    }
}

// ----------------------------------------------------------------------------------------------- materials
pub trait FlattenMaterial {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial;
}

fn mat(kind: u32, tex0: u32, tex1: u32, p: [f32; 4]) -> RtMaterial {
    RtMaterial { kind, tex0, tex1, reserved: 0, p }
}

impl FlattenMaterial for Emission {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_EMISSION, out.texture(&self.emit), 0, [0.0; 4])
    }
}
impl FlattenMaterial for Diffuse {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DIFFUSE, out.texture(&self.albedo), 0, [0.0; 4])
    }
}
impl FlattenMaterial for Lambert {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_LAMBERT, out.texture(&self.albedo), 0, [0.0; 4])
    }
}
impl FlattenMaterial for Metal {
    fn flatten(&self, _out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_METAL, 0, 0, [self.albedo.x, self.albedo.y, self.albedo.z, self.fuzz])
    }
}
impl FlattenMaterial for Dielectric {
    fn flatten(&self, _out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DIELECTRIC, 0, 0, [self.ior, 0.0, 0.0, 0.0])
    }
}
impl FlattenMaterial for Isotropic {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_ISOTROPIC, out.texture(&self.albedo), 0, [0.0; 4])
    }
}
impl FlattenMaterial for OrenNayar {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_OREN_NAYAR, out.texture(&self.albedo), 0, [self.roughness, 0.0, 0.0, 0.0])
    }
}
impl FlattenMaterial for BurleyDiffuse {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_BURLEY_DIFFUSE, out.texture(&self.albedo), 0, [self.roughness, 0.0, 0.0, 0.0])
    }
}
impl FlattenMaterial for RoughPlastic {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        let spec = out.texture(&self.spec_color);
        let diff = out.texture(&self.diff_color);
        mat(RT_MAT_ROUGH_PLASTIC, spec, diff, [self.roughness, self.eta, 0.0, 0.0])
    }
}
impl FlattenMaterial for DisneyDiffuse {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DISNEY_DIFFUSE, out.texture(&self.albedo), 0, [self.roughness, self.subsurface, 0.0, 0.0])
    }
}
impl FlattenMaterial for DisneyMetal {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DISNEY_METAL, out.texture(&self.albedo), 0, [self.roughness, self.anisotropic, self.rot, 0.0])
    }
}
impl FlattenMaterial for DisneySheen {
    fn flatten(&self, out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DISNEY_SHEEN, out.texture(&self.albedo), 0, [self.tint, 0.0, 0.0, 0.0])
    }
}
impl FlattenMaterial for DisneyClearcoat {
    fn flatten(&self, _out: &mut FlatScene) -> RtMaterial {
        mat(RT_MAT_DISNEY_CLEARCOAT, 0, 0, [self.clearcoat_gloss, 0.0, 0.0, 0.0])
    }
}

// ------------------------------------------------------------------------------------------------ textures
pub trait FlattenTexture {
    fn flatten(&self, out: &mut FlatScene) -> RtTexture;
}

fn tex(kind: u32, a: u32, b: u32, p: [f32; 4]) -> RtTexture {
    RtTexture { kind, a, b, reserved: 0, p }
}

impl FlattenTexture for ConstantTex {
    fn flatten(&self, _out: &mut FlatScene) -> RtTexture {
        tex(RT_TEX_CONSTANT, 0, 0, [self.col.x, self.col.y, self.col.z, 0.0])
    }
}
/// a = odd child, b = even child (src/texture.rs:25-28: children are arbitrary textures).
impl FlattenTexture for CheckerTex {
    fn flatten(&self, out: &mut FlatScene) -> RtTexture {
        let odd = out.texture(&self.odd);
        let even = out.texture(&self.even);
        tex(RT_TEX_CHECKER, odd, even, [0.0; 4])
    }
}
/// a = index of this texture's Perlin table (256 un-normalised gradients, three permutations of 0..255,
/// src/texture.rs:51-91), p = {scale}.
impl FlattenTexture for PerlinTex {
    fn flatten(&self, out: &mut FlatScene) -> RtTexture {
        let mut t = RtPerlin { rand_vec: [[0.0; 4]; 256], perm_x: [0; 256], perm_y: [0; 256], perm_z: [0; 256] };
        for i in 0..256 {
            let g = self.perlin.rand_vec[i];
            t.rand_vec[i] = [g.x, g.y, g.z, 0.0];
            t.perm_x[i] = self.perlin.perm_x[i] as u8;
            t.perm_y[i] = self.perlin.perm_y[i] as u8;
            t.perm_z[i] = self.perlin.perm_z[i] as u8;
        }
        out.perlin.push(t);
        tex(RT_TEX_PERLIN, out.perlin.len() as u32 - 1, 0, [self.scale, 0.0, 0.0, 0.0])
    }
}
impl FlattenTexture for ImageTex {
    fn flatten(&self, out: &mut FlatScene) -> RtTexture {
        tex(RT_TEX_IMAGE, out.image(self), 0, [0.0; 4])
    }
}

// ---------------------------------------------------------------------------------- what main() calls
/// A scene on `n_gpus` devices of this process: one replica per device and, for n_gpus > 1, the NCCL communicator.
pub struct GpuScene {
    scenes: Vec<*mut RtScene>,
    comm: *mut RtComm,
}

fn check(rc: i32, what: &str) -> Result<(), String> {
    if rc == RT_OK {
        return Ok(());
    }
    let msg = unsafe { std::ffi::CStr::from_ptr(rt_last_error()) }.to_string_lossy().into_owned();
    Err(format!("{what} failed: {rc} ({msg})")) // an error VALUE: nothing unwinds across the FFI boundary
}

impl GpuScene {
    pub fn new(flat: &RtFlatScene, n_gpus: usize) -> Result<Self, String> {
        let mut s = GpuScene { scenes: Vec::new(), comm: std::ptr::null_mut() };
        for g in 0..n_gpus {
            let mut h: *mut RtScene = std::ptr::null_mut();
            check(unsafe { rt_scene_create(flat, g as i32, &mut h) }, "rt_scene_create")?;
            s.scenes.push(h);
        }
        if n_gpus > 1 {
            let devices: Vec<i32> = (0..n_gpus as i32).collect();
            check(unsafe { rt_comm_create_all(devices.as_ptr(), n_gpus as i32, &mut s.comm) }, "rt_comm_create_all")?;
        }
        Ok(s)
    }

    /// Replaces the column jobs, the sample loop and ray_color of src/main.rs:77-108 (and, on several GPUs, the mpsc
    /// gather of src/main.rs:72,101,114): the linear mean image, row 0 = bottom, before gamma.
    pub fn render(&mut self, nx: u32, ny: u32, spp: u32) -> Result<(Vec<f32>, RtStats), String> {
        let rp = RtRenderParams { width: nx, height: ny, spp_total: spp, sample_begin: 0, sample_end: spp, max_depth: 50,
                                  mode: RT_MODE_AUTO, flags: 0, seed: 95 };
        let mut mean = vec![0f32; (nx * ny * 3) as usize];
        let mut st = RtStats::default();
        if self.comm.is_null() {
            check(unsafe { rt_render(self.scenes[0], &rp, mean.as_mut_ptr(), &mut st) }, "rt_render")?;
        } else {
            check(unsafe { rt_render_reduce(self.scenes.as_ptr(), self.comm, &rp, mean.as_mut_ptr(), &mut st) }, "rt_render_reduce")?;
        }
        Ok((mean, st))
    }

    /// src/main.rs:98-105,127: gamma 2, `* 255.99 as u8`, vertical flip.
    pub fn resolve_rgb8(&mut self, mean: &[f32], nx: u32, ny: u32) -> Result<Vec<u8>, String> {
        let mut out = vec![0u8; (nx * ny * 3) as usize];
        check(unsafe { rt_resolve_rgb8(self.scenes[0], mean.as_ptr(), nx, ny, out.as_mut_ptr()) }, "rt_resolve_rgb8")?;
        Ok(out)
    }
}

impl Drop for GpuScene {
    fn drop(&mut self) {
        unsafe {
            rt_comm_destroy(self.comm);
            for &s in &self.scenes {
                rt_scene_destroy(s);
            }
        }
    }
}
