//! FFI binding + safe wrapper of `librt_b200.so` for addiswebb/ray_tracer.
//!
//! Drop-in for `src/core/ray_tracer.rs` (`RayTracer::new` / `update_bind_group`), the storage
//! `Texture` (`src/core/texture.rs`) and the compute pass of `Context::render`
//! (`src/core/context.rs:196-208`).  The records are the crate's own `#[repr(C)]` Pod structs;
//! they are re-declared here only so that this file stands alone — inside the crate, `use
//! crate::core::{scene::{Sphere, Mesh, Vertex}, camera::CameraUniform, context::Params}` instead.
//!
//! NOT COMPILED in the authoring environment (no rustc there): kept 1:1 with `include/rt_b200.h`.
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

// ---- records (byte-identical to the reference's; the texture extension names the padding words)
#[repr(C)] #[derive(Clone, Copy, Debug, Default, bytemuck::Pod, bytemuck::Zeroable)]
pub struct Params { pub width: u32, pub height: u32, pub number_of_bounces: i32, pub rays_per_pixel: i32, pub skybox: i32, pub frames: i32, pub accumulate: i32 } // context.rs:11-21
#[repr(C)] #[derive(Clone, Copy, Debug, Default, bytemuck::Pod, bytemuck::Zeroable)]
pub struct CameraUniform {                                        // camera.rs:9-29
    pub origin: [f32; 3], _p1: f32, pub lower_left_corner: [f32; 3], _p2: f32, pub horizontal: [f32; 3], _p3: f32,
    pub vertical: [f32; 3], _p4: f32, pub near: f32, pub far: f32, _p5: [f32; 2], pub w: [f32; 3], _p6: f32,
    pub u: [f32; 3], _p7: f32, pub v: [f32; 3], pub lens_radius: f32,
}
#[repr(C)] #[derive(Clone, Copy, Debug, Default, bytemuck::Pod, bytemuck::Zeroable)]
pub struct Sphere { pub position: [f32; 3], pub radius: f32, pub color: [f32; 4], pub emission_color: [f32; 4], pub emission_strength: f32, pub smoothness: f32, _pad: [f32; 2] } // scene.rs:11-21
#[repr(C)] #[derive(Clone, Copy, Debug, Default, bytemuck::Pod, bytemuck::Zeroable)]
pub struct Mesh {                                                 // scene.rs:23-37
    pub first: u32, pub triangles: u32, pub offset: u32,
    pub diffuse_tex: u32,                                         // upstream `_padding: f32` (0.0); extension: 1-based texture handle
    pub pos: [f32; 3], _p2: f32, pub color: [f32; 4], pub emission_color: [f32; 4], pub emission_strength: f32, pub specular: f32,
    pub normal_tex: u32, _p3: u32,                                // upstream `_padding3: [f32; 2]`
}
#[repr(C)] #[derive(Clone, Copy, Debug, Default, bytemuck::Pod, bytemuck::Zeroable)]
pub struct Vertex { pub pos: [f32; 3], pub tu: f32, pub normal: [f32; 3], pub tv: f32 }   // scene.rs:79-86; tu/tv = upstream padding
#[repr(C)] pub struct Texture { pub width: u32, pub height: u32, pub rgba8: *const u8 }
#[repr(C)] #[derive(Clone, Copy, Debug, Default)]
pub struct Counters { pub rays: u64, pub samples: u64, pub node_visits: u64, pub tri_tests: u64, pub sphere_tests: u64, pub hit_segments: u64, pub dispatches: u64, pub kernel_ms: f64, pub launches: u64 }

const _: () = {
    assert!(std::mem::size_of::<Params>() == 28 && std::mem::size_of::<CameraUniform>() == 128);
    assert!(std::mem::size_of::<Sphere>() == 64 && std::mem::size_of::<Mesh>() == 80 && std::mem::size_of::<Vertex>() == 32);
};

pub mod opt { pub const TRAVERSAL: i32 = 1; pub const SEED_MODE: i32 = 2; pub const PRIMARY_IDS: i32 = 3; pub const DETAIL_COUNTERS: i32 = 4;
              pub const FRAMES_PER_LAUNCH: i32 = 5; pub const LEAF_SIZE: i32 = 6; pub const SCHED: i32 = 7; pub const BUILDER: i32 = 8; pub const PLOC_RADIUS: i32 = 9; pub const KERNEL: i32 = 10; pub const WAVE_SLOTS: i32 = 11; pub const STAGING_MIB: i32 = 12; pub const ITEM_MODE: i32 = 13; pub const INSTANCING: i32 = 14; pub const PIPELINE: i32 = 15; pub const NODE_FORMAT: i32 = 16; pub const NODES_PACKED: i32 = 17; pub const TILE_ORDER: i32 = 18; }

#[repr(C)] pub struct RtCtx { _private: [u8; 0] }

extern "C" {
    pub fn rt_last_error(ctx: *const RtCtx) -> *const c_char;
    pub fn rt_version() -> u32;
    pub fn rt_create(out: *mut *mut RtCtx, device: c_int, width: u32, height: u32) -> c_int;
    pub fn rt_destroy(ctx: *mut RtCtx);
    pub fn rt_upload_scene(ctx: *mut RtCtx, spheres: *const Sphere, n_spheres: u32, vertices: *const Vertex, n_vertices: u32,
                           indices: *const u32, n_indices: u32, meshes: *const Mesh, n_meshes: u32) -> c_int;
    pub fn rt_update_records(ctx: *mut RtCtx, spheres: *const Sphere, n_spheres: u32, meshes: *const Mesh, n_meshes: u32) -> c_int;
    pub fn rt_upload_textures(ctx: *mut RtCtx, textures: *const Texture, n_textures: u32) -> c_int;
    pub fn rt_set_camera(ctx: *mut RtCtx, camera: *const CameraUniform) -> c_int;
    pub fn rt_set_params(ctx: *mut RtCtx, params: *const Params) -> c_int;
    pub fn rt_dispatch(ctx: *mut RtCtx) -> c_int;
    pub fn rt_render_frames(ctx: *mut RtCtx, first_frame: i32, n_frames: u32) -> c_int;
    pub fn rt_render_frames_static(ctx: *mut RtCtx, first_frame: i32, n_frames: u32) -> c_int;
    pub fn rt_sync(ctx: *mut RtCtx) -> c_int;
    pub fn rt_resize(ctx: *mut RtCtx, width: u32, height: u32) -> c_int;
    pub fn rt_clear(ctx: *mut RtCtx) -> c_int;
    pub fn rt_set_shard(ctx: *mut RtCtx, tile_rank: u32, tile_world: u32, frame_rank: u32, frame_world: u32) -> c_int;
    pub fn rt_combine_peers(ctxs: *const *mut RtCtx, n: u32, weights: *const f32) -> c_int;
    pub fn rt_combine_peers_rgba8(ctxs: *const *mut RtCtx, n: u32, weights: *const f32, rgba: *mut u8, n_bytes: usize) -> c_int;
    pub fn rt_comm_unique_id(id: *mut c_void, n_bytes: usize) -> c_int;
    pub fn rt_comm_init(ctx: *mut RtCtx, n_ranks: c_int, rank: c_int, id: *const c_void, n_bytes: usize) -> c_int;
    pub fn rt_comm_destroy(ctx: *mut RtCtx) -> c_int;
    pub fn rt_combine(ctx: *mut RtCtx, root: c_int) -> c_int;
    pub fn rt_combine_rgba8(ctx: *mut RtCtx, root: c_int, rgba: *mut u8, n_bytes: usize) -> c_int;
    pub fn rt_read_combined(ctx: *mut RtCtx, rgba: *mut f32, n_floats: usize) -> c_int;
    pub fn rt_combined_device_ptr(ctx: *mut RtCtx) -> *mut c_void;
    pub fn rt_mean_count(ctx: *const RtCtx) -> u32;
    pub fn rt_set_option(ctx: *mut RtCtx, option: c_int, value: i64) -> c_int;
    pub fn rt_get_option(ctx: *const RtCtx, option: c_int, value: *mut i64) -> c_int;
    pub fn rt_read_accum(ctx: *mut RtCtx, rgba: *mut f32, n_floats: usize) -> c_int;
    pub fn rt_write_accum(ctx: *mut RtCtx, rgba: *const f32, n_floats: usize) -> c_int;
    pub fn rt_read_rgba8(ctx: *mut RtCtx, rgba: *mut u8, n_bytes: usize) -> c_int;
    pub fn rt_read_primary_ids(ctx: *mut RtCtx, ids: *mut u32, n: usize) -> c_int;
    pub fn rt_get_counters(ctx: *mut RtCtx, out: *mut Counters) -> c_int;
    pub fn rt_reset_counters(ctx: *mut RtCtx) -> c_int;
    pub fn rt_accum_device_ptr(ctx: *mut RtCtx) -> *mut c_void;
    pub fn rt_cuda_stream(ctx: *mut RtCtx) -> *mut c_void;
    pub fn rt_device(ctx: *const RtCtx) -> c_int;
    pub fn rt_measure_fp32_peak(device: c_int, reps: c_int, tflops: *mut f64, seconds: *mut f64) -> c_int;
    pub fn rt_bvh_stats(ctx: *mut RtCtx, n_prims: *mut u32, n_nodes: *mut u32, max_depth: *mut u32) -> c_int;
}

#[derive(Debug)]
pub struct Error { pub code: i32, pub message: String }
impl std::fmt::Display for Error { fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result { write!(f, "rt_b200 error {}: {}", self.code, self.message) } }
impl std::error::Error for Error {}
pub type Result<T> = std::result::Result<T, Error>;

/// Replaces `core::ray_tracer::RayTracer` + the storage `Texture`: one CUDA device, one image.
pub struct RayTracer { ctx: *mut RtCtx, pub width: u32, pub height: u32 }
unsafe impl Send for RayTracer {}                                  // one caller thread at a time, like the wgpu queue it replaces

impl RayTracer {
    fn check(&self, rc: c_int) -> Result<()> {
        if rc == 0 { return Ok(()); }
        let message = unsafe { CStr::from_ptr(rt_last_error(self.ctx)) }.to_string_lossy().into_owned();
        Err(Error { code: rc, message })                           // a return value, never a panic across the boundary
    }
    /// was `RayTracer::new(&device, &texture, &params_buffer, &scene)` (ray_tracer.rs:12) + `Texture::new` (texture.rs:6)
    pub fn new(cuda_device: i32, width: u32, height: u32) -> Result<Self> {
        let mut ctx = std::ptr::null_mut();
        let rc = unsafe { rt_create(&mut ctx, cuda_device, width, height) };
        if rc != 0 {
            let message = unsafe { CStr::from_ptr(rt_last_error(std::ptr::null())) }.to_string_lossy().into_owned();
            return Err(Error { code: rc, message });
        }
        Ok(Self { ctx, width, height })
    }
    /// bindings 3..6 (scene.rs:478-507, ray_tracer.rs:51-94)
    pub fn upload_scene(&self, spheres: &[Sphere], vertices: &[Vertex], indices: &[u32], meshes: &[Mesh]) -> Result<()> {
        self.check(unsafe { rt_upload_scene(self.ctx, spheres.as_ptr(), spheres.len() as u32, vertices.as_ptr(), vertices.len() as u32,
                                            indices.as_ptr(), indices.len() as u32, meshes.as_ptr(), meshes.len() as u32) })
    }
    pub fn update_records(&self, spheres: &[Sphere], meshes: &[Mesh]) -> Result<()> {
        self.check(unsafe { rt_update_records(self.ctx, spheres.as_ptr(), spheres.len() as u32, meshes.as_ptr(), meshes.len() as u32) })
    }
    /// extension: RGBA8 images, row 0 first; `Mesh::diffuse_tex` / `normal_tex` index this table from 1
    pub fn upload_textures(&self, images: &[(u32, u32, &[u8])]) -> Result<()> {
        let t: Vec<Texture> = images.iter().map(|(w, h, px)| { assert_eq!(px.len(), (*w as usize) * (*h as usize) * 4); Texture { width: *w, height: *h, rgba8: px.as_ptr() } }).collect();
        self.check(unsafe { rt_upload_textures(self.ctx, t.as_ptr(), t.len() as u32) })
    }
    /// context.rs:185-186
    pub fn set_camera(&self, u: &CameraUniform) -> Result<()> { self.check(unsafe { rt_set_camera(self.ctx, u) }) }
    pub fn set_params(&self, p: &Params) -> Result<()> { self.check(unsafe { rt_set_params(self.ctx, p) }) }
    /// the compute pass, context.rs:196-208
    pub fn dispatch(&self) -> Result<()> { self.check(unsafe { rt_dispatch(self.ctx) }) }
    /// `n` redraws without input in between, fused on the device (frames = first, first+1, ...)
    pub fn render_frames(&self, first_frame: i32, n: u32) -> Result<()> { self.check(unsafe { rt_render_frames(self.ctx, first_frame, n) }) }
    /// `render_frames` without any allocation: scratch in the library's static device array (RT_STATIC_SCRATCH_BYTES).
    pub fn render_frames_static(&self, first_frame: i32, n: u32) -> Result<()> { self.check(unsafe { rt_render_frames_static(self.ctx, first_frame, n) }) }
    pub fn sync(&self) -> Result<()> { self.check(unsafe { rt_sync(self.ctx) }) }
    /// was `update_bind_group` after `Context::resize` (ray_tracer.rs:149, context.rs:126-142)
    pub fn resize(&mut self, width: u32, height: u32) -> Result<()> { self.check(unsafe { rt_resize(self.ctx, width, height) })?; self.width = width; self.height = height; Ok(()) }
    pub fn set_option(&self, option: i32, value: i64) -> Result<()> { self.check(unsafe { rt_set_option(self.ctx, option, value) }) }
    pub fn set_shard(&self, tile_rank: u32, tile_world: u32, frame_rank: u32, frame_world: u32) -> Result<()> { self.check(unsafe { rt_set_shard(self.ctx, tile_rank, tile_world, frame_rank, frame_world) }) }
    /// RGBA32F, row 0 = bottom row
    pub fn read_accum(&self) -> Result<Vec<f32>> {
        let mut v = vec![0f32; self.width as usize * self.height as usize * 4];
        self.check(unsafe { rt_read_accum(self.ctx, v.as_mut_ptr(), v.len()) })?; Ok(v)
    }
    /// what render.wgsl + the sRGB swapchain displayed: clamp, OETF, 8 bit
    pub fn read_rgba8(&self) -> Result<Vec<u8>> {
        let mut v = vec![0u8; self.width as usize * self.height as usize * 4];
        self.check(unsafe { rt_read_rgba8(self.ctx, v.as_mut_ptr(), v.len()) })?; Ok(v)
    }
    /// One process (or host thread) per GPU: join the NCCL communicator of the job. `id` = `comm_unique_id()` of one rank,
    /// handed to the others by the host (upstream is single-device, context.rs:196-208 is the one dispatch site this serves).
    pub fn comm_init(&self, n_ranks: i32, rank: i32, id: &[u8; 128]) -> Result<()> {
        self.check(unsafe { rt_comm_init(self.ctx, n_ranks, rank, id.as_ptr() as *const c_void, id.len()) })
    }
    /// Collective, asynchronous: every rank's tiles -> `root`, count-weighted, de-interleaved there (`read_combined`).
    pub fn combine(&self, root: i32) -> Result<()> { self.check(unsafe { rt_combine(self.ctx, root) }) }
    /// The combined image of the last `combine` (at the root) / `combine_peers` (on shards[0]); accum is never overwritten.
    pub fn read_combined(&self) -> Result<Vec<f32>> {
        let mut v = vec![0f32; self.width as usize * self.height as usize * 4];
        self.check(unsafe { rt_read_combined(self.ctx, v.as_mut_ptr(), v.len()) })?; Ok(v)
    }
    pub fn counters(&self) -> Result<Counters> { let mut c = Counters::default(); self.check(unsafe { rt_get_counters(self.ctx, &mut c) })?; Ok(c) }
}
impl Drop for RayTracer { fn drop(&mut self) { unsafe { rt_destroy(self.ctx) } } }

/// 128 bytes identifying a new NCCL communicator (`ncclGetUniqueId`); one rank makes it, all pass it to `comm_init`.
pub fn comm_unique_id() -> Result<[u8; 128]> {
    let mut id = [0u8; 128];
    let rc = unsafe { rt_comm_unique_id(id.as_mut_ptr() as *mut c_void, id.len()) };
    if rc != 0 {
        let message = unsafe { CStr::from_ptr(rt_last_error(std::ptr::null())) }.to_string_lossy().into_owned();
        return Err(Error { code: rc, message });
    }
    Ok(id)
}
/// Several GPUs from one process (one `RayTracer` per device, sharded with `set_shard`): combines every shard's image
/// into `shards[0]`'s combined image (`read_combined`) with one kernel over NVLink peer memory. `weights = None`: each
/// shard's share of the dispatches its pixels have seen.
pub fn combine_peers(shards: &[&RayTracer], weights: Option<&[f32]>) -> Result<()> {
    let ptrs: Vec<*mut RtCtx> = shards.iter().map(|s| s.ctx).collect();
    let w = weights.map_or(std::ptr::null(), |w| w.as_ptr());
    shards[0].check(unsafe { rt_combine_peers(ptrs.as_ptr(), ptrs.len() as u32, w) })
}
/// The same combine fused with the display transform (clamp, sRGB OETF, 8 bit), straight to host memory.
pub fn combine_peers_rgba8(shards: &[&RayTracer], weights: Option<&[f32]>) -> Result<Vec<u8>> {
    let ptrs: Vec<*mut RtCtx> = shards.iter().map(|s| s.ctx).collect();
    let w = weights.map_or(std::ptr::null(), |w| w.as_ptr());
    let mut v = vec![0u8; shards[0].width as usize * shards[0].height as usize * 4];
    shards[0].check(unsafe { rt_combine_peers_rgba8(ptrs.as_ptr(), ptrs.len() as u32, w, v.as_mut_ptr(), v.len()) })?;
    Ok(v)
}
