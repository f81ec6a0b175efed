// rayca-cuda — `CudaRenderer`, a drop-in for `rayca_soft::SoftRenderer` on B200 (sm_100a).
//
// The plug-in interface is the reference's own trait
//     pub trait Draw { fn draw(&mut self, scene: &Scene, image: &mut Image); }      (rayca-soft/src/draw.rs:7-9)
// and the two constructors of `SoftRenderer` (rayca-soft/src/scene.rs:11-14,57-61), so a test swaps
// `SoftRenderer::default()` for `CudaRenderer::default()` and nothing else.  Everything `SoftRenderer::draw` does
// (rayca-soft/src/scene.rs:88-154) happens inside librayca_cuda; there is NO CPU fallback: without a CUDA device
// the constructor panics.
//
// UNCOMPILED in the authoring container (no Rust toolchain there); see INTEGRATION.md for what was checked instead.
pub mod ffi;
pub mod flat;

use std::ffi::CStr;
use std::ptr;

use rayca_model::{ColorType, Image, Model, Scene};
use rayca_soft::{Config, Draw, SoftRenderer};

pub use flat::FlatScene;

/// `rayca_soft::Config` (rayca-soft/src/config.rs:10-49) field for field; the strategy enums are `#[repr(u32)]`
/// (integrator/mod.rs:32-41, sampler/mod.rs:41-50) and the C header uses the same discriminants.
fn config_to_c(c: &Config) -> ffi::RaycaConfig {
    ffi::RaycaConfig {
        bvh: c.bvh as u32,
        light_samples: c.light_samples,
        light_stratify: c.light_stratify as u32,
        samples_per_pixel: c.samples_per_pixel,
        russian_roulette: c.russian_roulette as u32,
        direct_sampler: c.direct_sampler as u32,
        indirect_sampler: c.indirect_sampler as u32,
        integrator: c.integrator as u32,
        max_depth: c.max_depth,
        gamma: c.gamma,
    }
}

/// `Draw::draw` returns `()` and the reference panics on failure (`assert!` scene.rs:109, `todo!()`s): so does this.
fn check(rc: i32, what: &str) {
    if rc != ffi::RAYCA_OK {
        let (name, detail) = unsafe {
            (CStr::from_ptr(ffi::rayca_cuda_strerror(rc)).to_string_lossy().into_owned(),
             CStr::from_ptr(ffi::rayca_cuda_last_error()).to_string_lossy().into_owned())
        };
        panic!("rayca-cuda: {what} failed: {name} ({rc}): {detail}");
    }
}

pub struct CudaRenderer {
    pub config: Config,
    /// backend-only knobs (seed of the keyed RNG, traversal mode); not part of the reference's `Config`
    pub options: ffi::RaycaOptions,
    ctx: *mut ffi::rayca_ctx,
}

// one host thread drives the context at a time (`draw` takes `&mut self`), like `SoftRenderer`
unsafe impl Send for CudaRenderer {}

impl Default for CudaRenderer {
    fn default() -> Self {
        Self::new_with_config(Config::default())
    }
}

impl CudaRenderer {
    /// `SoftRenderer::new_with_config` (scene.rs:57-61) on every CUDA device of the box: one `draw` call fans out over
    /// all of them inside the library, as the reference fans out over the rayon pool (scene.rs:117-122).
    pub fn new_with_config(config: Config) -> Self {
        Self::new_with_devices(config, &[])
    }

    /// `devices` = CUDA device indices to use; empty = all visible devices.
    pub fn new_with_devices(config: Config, devices: &[i32]) -> Self {
        let mut ctx = ptr::null_mut();
        let rc = unsafe {
            if devices.is_empty() {
                ffi::rayca_cuda_init_devices(ptr::null(), 0, &mut ctx)
            } else {
                ffi::rayca_cuda_init_devices(devices.as_ptr(), devices.len() as i32, &mut ctx)
            }
        };
        check(rc, "rayca_cuda_init_devices");
        let options = ffi::RaycaOptions { rng_seed: 0x5EED, traversal: 0, shard_index: 0, shard_count: 0, tile_rows: 0, flags: 0, reserved: 0 };
        Self { config, options, ctx }
    }

    /// `SoftRenderer::create_default_model` (scene.rs:18-55): camera at z = 4 and two point lights — host-side scene
    /// construction, unchanged, so it is the reference's own function.
    pub fn create_default_model() -> Model {
        SoftRenderer::create_default_model()
    }
}

impl Draw for CudaRenderer {
    /// rayca-soft/src/scene.rs:88-154.  Like the reference, the acceleration structure is rebuilt on every call.
    fn draw(&mut self, scene: &Scene, image: &mut Image) {
        assert!(matches!(image.color_type, ColorType::RGBA8), "CudaRenderer draws RGBA8 images (image.rs:29-31)");
        let flat = FlatScene::from(scene);
        let desc = flat.desc();
        let cfg = config_to_c(&self.config);
        let (width, height) = (image.width(), image.height());
        let mut handle = ptr::null_mut();
        check(unsafe { ffi::rayca_cuda_scene_create(self.ctx, &desc, cfg.bvh, &mut handle, ptr::null_mut()) }, "rayca_cuda_scene_create");
        let rc = unsafe {
            ffi::rayca_cuda_render(self.ctx, handle, &cfg, &self.options, width, height, image.bytes_mut().as_mut_ptr(), ptr::null(), ptr::null_mut())
        };
        unsafe { ffi::rayca_cuda_scene_destroy(handle) };
        check(rc, "rayca_cuda_render");
    }
}

impl Drop for CudaRenderer {
    fn drop(&mut self) {
        unsafe { ffi::rayca_cuda_shutdown(self.ctx) };
    }
}
