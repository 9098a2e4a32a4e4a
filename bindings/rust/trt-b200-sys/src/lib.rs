//! FFI declarations of `include/trt_b200.h` (ABI version 1) and a thin safe wrapper.
//!
//! The wire structs are the reference's own `#[repr(C)]` Pod types (`src/vec3.rs:5-12`,
//! `src/primitive.rs:4-34`, `src/lib.rs:23-50`); they are restated here so that the crate stands alone,
//! and are layout-identical (`const` size assertions below), so `&[Primitive]` / `&Uniforms` of the
//! reference can be passed with a pointer cast.
#![allow(non_camel_case_types)]
use std::ffi::CStr;
use std::os::raw::{c_char, c_int, c_void};

pub const TRT_ABI_VERSION: u32 = 1;
pub const TRT_OK: c_int = 0;
pub const TRT_ERR_INVALID: c_int = -1;
pub const TRT_ERR_NO_DEVICE: c_int = -2;
pub const TRT_ERR_CUDA: c_int = -3;
pub const TRT_ERR_NO_SCENE: c_int = -4;
pub const TRT_ERR_TOO_SMALL: c_int = -5;
pub const TRT_ERR_PARSE: c_int = -6;
pub const TRT_ERR_IO: c_int = -7;
pub const TRT_ERR_NOMEM: c_int = -8;
pub const TRT_CONV_AS_WRITTEN: u32 = 0;
pub const TRT_CONV_FUSED: u32 = 1;
pub const TRT_CONV_NATIVE: u32 = 2;
pub const TRT_MODE_ASCII: c_int = 0;
pub const TRT_MODE_TRUECOLOR: c_int = 1;
pub const TRT_MODE_NONE: c_int = 2;
pub const TRT_ACCEL_NONE: u32 = 0;
pub const TRT_ACCEL_GRID: u32 = 1;
pub const TRT_KEY_UP: c_int = 1001;
pub const TRT_KEY_DOWN: c_int = 1002;
pub const TRT_KEY_LEFT: c_int = 1003;
pub const TRT_KEY_RIGHT: c_int = 1004;

#[repr(C)] #[derive(Clone, Copy, Default, Debug, PartialEq)]
pub struct trt_vec3 { pub x: f32, pub y: f32, pub z: f32, pub _padding: f32 }

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_primitive {
    pub primitive_type: u32, pub _padding1: u32, pub _padding2: u32, pub _padding3: u32,
    pub color: trt_vec3, pub emission: trt_vec3,
    pub reflectivity: f32, pub _padding4: f32, pub _padding5: f32, pub _padding6: f32,
    pub sphere_center: trt_vec3, pub sphere_radius: f32, pub _sphere_padding: [f32; 3],
    pub plane_point: trt_vec3, pub plane_normal: trt_vec3,
    pub triangle_v0: trt_vec3, pub triangle_v1: trt_vec3, pub triangle_v2: trt_vec3,
    pub _triangle_padding: f32,
}

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_uniforms {
    pub width: u32, pub height: u32, pub samples_per_pixel: u32, pub max_depth: u32,
    pub seed: u32, pub frame_number: u32, pub primitive_count: u32, pub _padding1: u32,
    pub aspect_ratio: f32, pub char_aspect_ratio: f32, pub fov_rad: f32, pub _padding2: f32,
    pub camera_pos: trt_vec3, pub camera_forward: trt_vec3, pub camera_right: trt_vec3, pub camera_up: trt_vec3,
    pub grid_min: trt_vec3, pub inv_cell_size: trt_vec3, pub grid_dims: [u32; 3], pub _padding3: u32,
}

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_variance { pub variance: f32, pub sample_count: u32 }

const _: () = assert!(std::mem::size_of::<trt_vec3>() == 16);
const _: () = assert!(std::mem::size_of::<trt_primitive>() == 180);
const _: () = assert!(std::mem::size_of::<trt_uniforms>() == 160);
const _: () = assert!(std::mem::size_of::<trt_variance>() == 8);

#[repr(C)] pub struct trt_ctx { _private: [u8; 0] }
#[repr(C)] pub struct trt_scene { _private: [u8; 0] }
#[repr(C)] pub struct trt_multi { _private: [u8; 0] }

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_config {
    pub abi_version: u32, pub device: i32, pub max_width: u32, pub max_height: u32,
    pub conv: u32, pub count_work: u32,
    pub shard_index: u32, pub shard_count: u32, pub stripe_rows: u32,
    pub accel: u32,
    pub reserved: [u32; 6],
}

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_counters {
    pub pixels: u64, pub samples_base: u64, pub samples_extra: u64, pub closest_scans: u64, pub hits: u64, pub misses: u64,
    pub shadow_scans: u64, pub shadow_unoccluded: u64, pub shadow_lit: u64, pub shadow_needed: u64, pub mirror: u64,
    pub diffuse: u64, pub roulette_tests: u64, pub roulette_kills: u64, pub prim_tests: u64, pub prim_tests_min: u64,
    pub flops_written: u64, pub flops_min: u64, pub rng_draws: u64,
}

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_timings { pub trace_ms: f32, pub quantise_ms: f32, pub d2h_ms: f32, pub total_ms: f32, pub kernel_launches: u32, pub _pad: u32 }

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_frame_stats { pub samples_traced: u64, pub bounces: u64, pub pixels: u64, pub reserved: u64 }

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_camera { pub position: trt_vec3, pub yaw: f32, pub pitch: f32 }

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct trt_scene_info {
    pub width: u32, pub height: u32, pub samples_per_pixel: u32, pub max_depth: u32, pub frames_to_accumulate: u32,
    pub fov_degrees: f32, pub char_aspect_ratio: f32, pub n_spheres: u32, pub n_planes: u32, pub n_triangles: u32,
}

#[repr(C)] #[derive(Clone, Copy, Debug)]
pub struct trt_run_options {
    pub width: u32, pub height: u32, pub frames: u32, pub seed: u32, pub has_seed: u32,
    pub keys: *const c_char, pub out_path: *const c_char,
    pub device: i32, pub conv: u32, pub quiet: u32, pub batch_frames: u32, pub gpus: u32, pub accel: u32, pub pipeline: u32,
}

extern "C" {
    pub fn trt_version() -> *const c_char;
    pub fn trt_create(cfg: *const trt_config, out: *mut *mut trt_ctx) -> c_int;
    pub fn trt_destroy(ctx: *mut trt_ctx);
    pub fn trt_last_error(ctx: *const trt_ctx) -> *const c_char;
    pub fn trt_upload_scene(ctx: *mut trt_ctx, prims: *const trt_primitive, count: u32) -> c_int;
    pub fn trt_render_frame(ctx: *mut trt_ctx, u: *const trt_uniforms, mode: c_int, out: *mut u8, cap: usize, len: *mut usize) -> c_int;
    pub fn trt_render_frames(ctx: *mut trt_ctx, u: *const trt_uniforms, n: u32, mode: c_int, emit_all: c_int, out: *mut u8,
                             cap: usize, len: *mut usize, frame_len_or_null: *mut usize) -> c_int;
    pub fn trt_trace_async(ctx: *mut trt_ctx, u: *const trt_uniforms) -> c_int;
    pub fn trt_quantise_async(ctx: *mut trt_ctx, mode: c_int, len_needed_or_null: *mut usize) -> c_int;
    pub fn trt_sync(ctx: *mut trt_ctx) -> c_int;
    pub fn trt_read_stream(ctx: *mut trt_ctx, out: *mut u8, cap: usize, len: *mut usize) -> c_int;
    pub fn trt_read_pixels(ctx: *mut trt_ctx, dst4: *mut f32, n_pixels: usize) -> c_int;
    pub fn trt_read_accumulation(ctx: *mut trt_ctx, dst4: *mut f32, n_pixels: usize) -> c_int;
    pub fn trt_read_variance(ctx: *mut trt_ctx, dst: *mut trt_variance, n_pixels: usize) -> c_int;
    pub fn trt_write_accumulation(ctx: *mut trt_ctx, src4: *const f32, n_pixels: usize) -> c_int;
    pub fn trt_get_counters(ctx: *mut trt_ctx, out: *mut trt_counters) -> c_int;
    pub fn trt_get_timings(ctx: *mut trt_ctx, out: *mut trt_timings) -> c_int;
    pub fn trt_trace_kernel_name(ctx: *const trt_ctx) -> *const c_char;
    pub fn trt_get_frame_stats(ctx: *mut trt_ctx, out: *mut trt_frame_stats) -> c_int;
    pub fn trt_measure_fp32_peak(device: c_int, tflops: *mut f32, sm_count: *mut c_int, clock_khz: *mut c_int) -> c_int;
    pub fn trt_selftest_div3(device: c_int, n_cases: u64, seed: u32, distribution: c_int, mismatches: *mut u64, first_bad4: *mut f32) -> c_int;
    pub fn trt_device_buffer(ctx: *mut trt_ctx, which: c_int, dptr: *mut *mut c_void, bytes: *mut usize) -> c_int;
    pub fn trt_cuda_stream(ctx: *mut trt_ctx, stream: *mut *mut c_void) -> c_int;
    pub fn trt_shard_rows(ctx: *const trt_ctx, height: u32, rows: *mut u32, cap: u32) -> u32;
    pub fn trt_shard_rows_of(shard_index: u32, shard_count: u32, stripe_rows: u32, height: u32, rows: *mut u32, cap: u32) -> u32;
    pub fn trt_pack_shard_pixels(ctx: *mut trt_ctx, d_dst: *mut c_void, dst_bytes: usize) -> c_int;
    pub fn trt_quantise_gathered(ctx: *mut trt_ctx, d_gathered: *const c_void, block_stride_bytes: usize, width: u32, height: u32,
                                 shard_count: u32, stripe_rows: u32, mode: c_int, len_needed_or_null: *mut usize) -> c_int;
    /// allocates nothing: static device scratch, frames up to TRT_NOALLOC_MAX_PIXELS (no counterpart in the reference)
    pub fn trt_quantise_noalloc(device: c_int, d_pixels: *const c_void, width: u32, height: u32, mode: c_int, d_out: *mut c_void,
                                out_cap: usize, d_len: *mut c_void, cuda_stream: *mut c_void, err: *mut c_char, err_cap: usize) -> c_int;
    pub fn trt_alloc_host(bytes: usize, out: *mut *mut c_void) -> c_int;
    pub fn trt_free_host(p: *mut c_void);
    pub fn trt_host_register(p: *mut c_void, bytes: usize) -> c_int;
    pub fn trt_host_unregister(p: *mut c_void) -> c_int;
    pub fn trt_quantise_rows_async(ctx: *mut trt_ctx, mode: c_int) -> c_int;
    pub fn trt_row_lengths(ctx: *mut trt_ctx, out: *mut u32, cap: u32, n_rows: *mut u32) -> c_int;
    pub fn trt_read_stream_rows(ctx: *mut trt_ctx, host_base: *mut u8, cap: usize, row_off: *const u64, height: u32, wait: c_int) -> c_int;
    pub fn trt_frame_submit(ctx: *mut trt_ctx, u: *const trt_uniforms, mode: c_int) -> c_int;
    pub fn trt_frame_collect(ctx: *mut trt_ctx, out: *mut u8, cap: usize, len: *mut usize) -> c_int;
    pub fn trt_multi_create(devices: *const i32, n_devices: u32, cfg: *const trt_config, out: *mut *mut trt_multi) -> c_int;
    pub fn trt_multi_destroy(m: *mut trt_multi);
    pub fn trt_multi_last_error(m: *const trt_multi) -> *const c_char;
    pub fn trt_multi_size(m: *const trt_multi) -> u32;
    pub fn trt_multi_context(m: *mut trt_multi, i: u32) -> *mut trt_ctx;
    pub fn trt_multi_upload_scene(m: *mut trt_multi, prims: *const trt_primitive, count: u32) -> c_int;
    pub fn trt_multi_render_frame(m: *mut trt_multi, u: *const trt_uniforms, mode: c_int, out: *mut u8, cap: usize, len: *mut usize) -> c_int;
    pub fn trt_multi_read_pixels(m: *mut trt_multi, dst4: *mut f32, n_pixels: usize) -> c_int;
    pub fn trt_multi_read_variance(m: *mut trt_multi, dst: *mut trt_variance, n_pixels: usize) -> c_int;
    pub fn trt_multi_get_frame_stats(m: *mut trt_multi, out: *mut trt_frame_stats) -> c_int;
    pub fn trt_multi_get_timings(m: *mut trt_multi, out: *mut trt_timings) -> c_int;
    pub fn trt_multi_render_frames(m: *mut trt_multi, u: *const trt_uniforms, n: u32, mode: c_int, out: *mut u8, cap: usize, len: *mut usize) -> c_int;
    pub fn trt_multi_frames_context(m: *mut trt_multi) -> *mut trt_ctx;
    pub fn trt_trace_frames_async(ctx: *mut trt_ctx, u: *const trt_uniforms, n: u32) -> c_int;
    pub fn trt_frame_means(ctx: *mut trt_ctx, dptr: *mut *mut c_void, frame_stride_bytes: *mut usize) -> c_int;
    pub fn trt_blend_frames_async(ctx: *mut trt_ctx, d_means: *const c_void, u: *const trt_uniforms, n: u32) -> c_int;
    pub fn trt_camera_default(cam: *mut trt_camera);
    pub fn trt_camera_vectors(cam: *const trt_camera, forward: *mut trt_vec3, right: *mut trt_vec3, up: *mut trt_vec3);
    pub fn trt_camera_key(cam: *mut trt_camera, key: c_int) -> c_int;
    pub fn trt_scene_parse(json_text: *const c_char, json_len: usize, out: *mut *mut trt_scene, err: *mut c_char, err_cap: usize) -> c_int;
    pub fn trt_scene_load(path_or_null: *const c_char, out: *mut *mut trt_scene, err: *mut c_char, err_cap: usize) -> c_int;
    pub fn trt_scene_free(s: *mut trt_scene);
    pub fn trt_scene_get_info(s: *const trt_scene, info: *mut trt_scene_info) -> c_int;
    pub fn trt_scene_clamp(s: *mut trt_scene, term_cols: u32, term_rows: u32);
    pub fn trt_scene_set_size(s: *mut trt_scene, width: u32, height: u32);
    pub fn trt_scene_primitives(s: *const trt_scene, count: *mut u32) -> *const trt_primitive;
    pub fn trt_fill_uniforms(s: *const trt_scene, cam: *const trt_camera, seed: u32, frame_number: u32, out: *mut trt_uniforms);
    pub fn trt_run(full_color: c_int, verbose: c_int, scene_path_or_null: *const c_char, opts: *const trt_run_options) -> c_int;
}

// ---- safe wrapper: what `run()` in src/lib.rs needs in place of its wgpu objects ---------------------------

/// Error of a library call: status code and the library's message.
#[derive(Debug)]
pub struct Error { pub code: i32, pub message: String }
impl std::fmt::Display for Error {
    fn fmt(&self, f: &mut std::fmt::Formatter<'_>) -> std::fmt::Result { write!(f, "trt_b200 error {}: {}", self.code, self.message) }
}
impl std::error::Error for Error {}

/// One GPU context = device + buffers + pipeline of `src/lib.rs:248-383`.  Single owner, not `Sync`.
pub struct Tracer { ctx: *mut trt_ctx, out: Vec<u8> }

impl Tracer {
    fn err(ctx: *const trt_ctx, code: c_int) -> Error {
        let message = unsafe { CStr::from_ptr(trt_last_error(ctx)) }.to_string_lossy().into_owned();
        Error { code, message }
    }
    /// `conv`: TRT_CONV_NATIVE for production, TRT_CONV_AS_WRITTEN for the bit-exact build.
    pub fn new(device: i32, max_width: u32, max_height: u32, conv: u32) -> Result<Self, Error> {
        let cfg = trt_config { abi_version: TRT_ABI_VERSION, device, max_width, max_height, conv, ..Default::default() };
        let mut ctx = std::ptr::null_mut();
        let rc = unsafe { trt_create(&cfg, &mut ctx) };
        if rc != TRT_OK { return Err(Self::err(std::ptr::null(), rc)); }
        let cap = max_width as usize * max_height as usize * 26 + max_height as usize * 2 + 64;
        Ok(Tracer { ctx, out: vec![0u8; cap] })
    }
    /// `prims`: the reference's `Vec<Primitive>` (180-byte records, scan order), `src/lib.rs:120-154`.
    pub fn upload_scene<P: Copy>(&mut self, prims: &[P]) -> Result<(), Error> {
        assert_eq!(std::mem::size_of::<P>(), 180, "Primitive must be the 180-byte #[repr(C)] record");
        let rc = unsafe { trt_upload_scene(self.ctx, prims.as_ptr() as *const trt_primitive, prims.len() as u32) };
        if rc == TRT_OK { Ok(()) } else { Err(Self::err(self.ctx, rc)) }
    }
    /// One iteration of the frame loop, `src/lib.rs:442-545`: returns the ANSI rows block of the frame.
    pub fn render_frame<U: Copy>(&mut self, uniforms: &U, full_color: bool) -> Result<&[u8], Error> {
        assert_eq!(std::mem::size_of::<U>(), 160, "Uniforms must be the 160-byte #[repr(C)] record");
        let mode = if full_color { TRT_MODE_TRUECOLOR } else { TRT_MODE_ASCII };
        let mut len = 0usize;
        let rc = unsafe { trt_render_frame(self.ctx, uniforms as *const U as *const trt_uniforms, mode, self.out.as_mut_ptr(), self.out.len(), &mut len) };
        if rc == TRT_OK { Ok(&self.out[..len]) } else { Err(Self::err(self.ctx, rc)) }
    }
    /// `n` frames of a still camera in one GPU round trip; returns the last frame's rows block.
    pub fn render_frames<U: Copy>(&mut self, uniforms: &[U], full_color: bool) -> Result<&[u8], Error> {
        assert_eq!(std::mem::size_of::<U>(), 160);
        let mode = if full_color { TRT_MODE_TRUECOLOR } else { TRT_MODE_ASCII };
        let mut len = 0usize;
        let rc = unsafe { trt_render_frames(self.ctx, uniforms.as_ptr() as *const trt_uniforms, uniforms.len() as u32, mode, 0,
                                            self.out.as_mut_ptr(), self.out.len(), &mut len, std::ptr::null_mut()) };
        if rc == TRT_OK { Ok(&self.out[..len]) } else { Err(Self::err(self.ctx, rc)) }
    }
    pub fn frame_stats(&mut self) -> trt_frame_stats {
        let mut s = trt_frame_stats::default();
        unsafe { trt_get_frame_stats(self.ctx, &mut s) };
        s
    }
}
impl Drop for Tracer { fn drop(&mut self) { unsafe { trt_destroy(self.ctx) } } }

/// The same frame over several GPUs of this process (`trt_multi_*`): what `run()` holds instead of a `Tracer` when
/// it is given more than one device.  The rows block is byte-identical to a single-GPU frame.
pub struct MultiTracer { m: *mut trt_multi, out: *mut u8, cap: usize }

impl MultiTracer {
    fn err(m: *const trt_multi, code: c_int) -> Error {
        let message = unsafe { CStr::from_ptr(trt_multi_last_error(m)) }.to_string_lossy().into_owned();
        Error { code, message }
    }
    pub fn new(devices: &[i32], max_width: u32, max_height: u32, conv: u32) -> Result<Self, Error> {
        let cfg = trt_config { abi_version: TRT_ABI_VERSION, max_width, max_height, conv, ..Default::default() };
        let mut m = std::ptr::null_mut();
        let rc = unsafe { trt_multi_create(devices.as_ptr(), devices.len() as u32, &cfg, &mut m) };
        if rc != TRT_OK { return Err(Self::err(std::ptr::null(), rc)); }
        let cap = max_width as usize * max_height as usize * 26 + max_height as usize * 2 + 64;
        let mut out: *mut c_void = std::ptr::null_mut();
        // pinned memory: the GPUs copy their rows concurrently, each over its own PCIe link
        if unsafe { trt_alloc_host(cap, &mut out) } != TRT_OK { unsafe { trt_multi_destroy(m) }; return Err(Error { code: TRT_ERR_NOMEM, message: "pinned host buffer".into() }); }
        Ok(MultiTracer { m, out: out as *mut u8, cap })
    }
    pub fn upload_scene<P: Copy>(&mut self, prims: &[P]) -> Result<(), Error> {
        assert_eq!(std::mem::size_of::<P>(), 180, "Primitive must be the 180-byte #[repr(C)] record");
        let rc = unsafe { trt_multi_upload_scene(self.m, prims.as_ptr() as *const trt_primitive, prims.len() as u32) };
        if rc == TRT_OK { Ok(()) } else { Err(Self::err(self.m, rc)) }
    }
    pub fn render_frame<U: Copy>(&mut self, uniforms: &U, full_color: bool) -> Result<&[u8], Error> {
        assert_eq!(std::mem::size_of::<U>(), 160, "Uniforms must be the 160-byte #[repr(C)] record");
        let mode = if full_color { TRT_MODE_TRUECOLOR } else { TRT_MODE_ASCII };
        let mut len = 0usize;
        let rc = unsafe { trt_multi_render_frame(self.m, uniforms as *const U as *const trt_uniforms, mode, self.out, self.cap, &mut len) };
        if rc == TRT_OK { Ok(unsafe { std::slice::from_raw_parts(self.out, len) }) } else { Err(Self::err(self.m, rc)) }
    }
}
impl Drop for MultiTracer { fn drop(&mut self) { unsafe { trt_free_host(self.out as *mut c_void); trt_multi_destroy(self.m) } } }

/// The reference's `pub async fn run(full_color, verbose, scene_path)` (`src/lib.rs:100`) executed entirely by the
/// native library (scene parsing, camera, loop, terminal handling): for hosts that only want the entry point.
pub fn run(full_color: bool, verbose: bool, scene_path: Option<&str>) -> Result<(), Error> {
    let path = scene_path.map(|p| std::ffi::CString::new(p).expect("path contains NUL"));
    let opts = trt_run_options { width: 0, height: 0, frames: 0, seed: 0, has_seed: 0, keys: std::ptr::null(), out_path: std::ptr::null(),
                                 device: 0, conv: TRT_CONV_NATIVE, quiet: 0, batch_frames: 0, gpus: 0, accel: 0, pipeline: 0 };
    let rc = unsafe { trt_run(full_color as c_int, verbose as c_int, path.as_ref().map_or(std::ptr::null(), |p| p.as_ptr()), &opts) };
    if rc == TRT_OK { Ok(()) } else { Err(Error { code: rc, message: "trt_run failed (see stderr)".into() }) }
}
