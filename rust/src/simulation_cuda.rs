//! simulation_cuda.rs — `FluidSimulation` on top of libfluid_b200 (include/fluid_b200.h).
//!
//! Drop-in for the reference's `src/simulation.rs` behind `#[cfg(feature = "cuda")]`: same type and
//! method names (`FluidSimulation::new`, `tick`, `pub tick`), same `SimulationSettings` / `TickSettings`
//! field order (src/simulation.rs:95-122). The wgpu bind-group getters (src/simulation.rs:542-564) have no
//! CUDA meaning; the renderer reads particles through `download_particles` (sorted AoS, the same bytes the
//! reference's vertex/storage buffer holds) or the raw device pointers of `device_view`.
//!
//! NOT COMPILED in the build container (no rustc/cargo there); checked by review against the header. The same call
//! sequence is compiled and run from C (tests/c_abi_host.c, gcc) on the GPU box.
#![allow(non_camel_case_types)]
use std::ffi::{c_char, c_int, c_void, CStr};

#[repr(C)] #[derive(Clone, Copy, Default, Debug)]
pub struct ParticleInstance { pub position: [f32; 2], pub predicted_position: [f32; 2], pub velocity: [f32; 2], pub density: f32, pub grid: u32 }

#[repr(C)] #[derive(Clone, Copy)]
pub struct SimulationSettings { pub particle_count: u32, pub particle_spacing: f32, pub smoothing_radius: f32, pub size: [f32; 2], pub texture_size: [u32; 2] }

#[repr(C)] #[derive(Clone, Copy)]
pub struct TickSettings {
    pub delta: f32, pub gravity: [f32; 2], pub mass: f32, pub pressure_constant: f32, pub rest_density: f32,
    pub damping_factor: f32, pub viscosity_coefficient: f32, pub surface_tension_treshold: f32,
    pub surface_tension_coefficient: f32, pub mouse_force_radius: f32, pub mouse_force_power: f32,
    pub mouse_pos: [f32; 2], pub mouse_state: i32,
}

/// the reference's storage-buffer record (src/renderer.rs:82-90); its own `FluidObject` can be passed by pointer
#[repr(C, align(16))] #[derive(Clone, Copy, Debug)]
pub struct FluidObject { pub position: [f32; 2], pub kind: u32, pub pad: u32, pub pad2: [u32; 4] }

#[repr(C)] pub struct FluidSim { _opaque: [u8; 0] }
#[repr(C)] pub struct FluidMulti { _opaque: [u8; 0] }

/// `FluidMultiOptions` (include/fluid_b200.h); `Default` = equal bands of grid rows, default capacity and windows, transport AUTO
#[repr(C)] #[derive(Clone, Copy, Default)]
pub struct FluidMultiOptions { pub row_bounds: [u32; 9], pub capacity: u32, pub window_near: u32, pub window_far: u32, pub transport: i32, pub reserved: [u32; 4] }

extern "C" {
    fn fluid_last_error() -> *const c_char;
    fn fluid_sim_create(settings: *const SimulationSettings, device_ordinal: c_int, out: *mut *mut FluidSim) -> c_int;
    fn fluid_sim_destroy(sim: *mut FluidSim);
    fn fluid_sim_tick(sim: *mut FluidSim, tick: *const TickSettings) -> c_int;
    fn fluid_sim_sync(sim: *mut FluidSim) -> c_int;
    fn fluid_sim_tick_count(sim: *const FluidSim) -> u32;
    fn fluid_sim_download_particles(sim: *mut FluidSim, dst: *mut ParticleInstance) -> c_int;
    fn fluid_sim_upload_force_field(sim: *mut FluidSim, xy: *const f32) -> c_int;
    fn fluid_sim_download_start_indices(sim: *mut FluidSim, dst: *mut u32) -> c_int;
    fn fluid_sim_grid_w(sim: *const FluidSim) -> u32;
    fn fluid_sim_grid_h(sim: *const FluidSim) -> u32;
    fn fluid_sim_stream(sim: *mut FluidSim) -> *mut c_void;
    fn fluid_sim_field_from_objects(sim: *mut FluidSim, objects: *const FluidObject, n: u32, inv_proj: *const f32,
                                    mask_out: *mut u8, seeds_out: *mut u32, field_xy_out: *mut f32, ms_out: *mut f32) -> c_int;
    fn fluid_sim_field_from_mask(sim: *mut FluidSim, mask: *const u8, seeds_out: *mut u32, field_xy_out: *mut f32, ms_out: *mut f32) -> c_int;
    fn fluid_sim_render(sim: *mut FluidSim, inv_proj: *const f32, w: u32, h: u32, rgba_f32_out: *mut f32, rgba8_out: *mut u8, ms_out: *mut f32) -> c_int;
    fn fluid_sim_surface_tension(sim: *mut FluidSim, force_xy_out: *mut f32) -> c_int;
    fn fluid_sim_set_sort_mode(sim: *mut FluidSim, mode: c_int) -> c_int;
    fn fluid_sim_last_sort_path(sim: *const FluidSim) -> c_int;
    fn fluid_sim_upload_particles_async(sim: *mut FluidSim, src: *const ParticleInstance) -> c_int;
    fn fluid_sim_download_particles_async(sim: *mut FluidSim, dst: *mut ParticleInstance) -> c_int;
    fn fluid_sim_sync_transfers(sim: *mut FluidSim) -> c_int;
    fn fluid_sim_state_changed(sim: *mut FluidSim) -> c_int;
    fn fluid_sim_render_frame(sim: *mut FluidSim, inv_proj: *const f32, w: u32, h: u32, rgba8_out: *mut u8, srgb: c_int, wait: c_int) -> c_int;
    // the multi-GPU step: the same call per frame, every GPU of the box behind it
    fn fluid_multi_create(settings: *const SimulationSettings, n_ranks: c_int, devices: *const c_int, opt: *const FluidMultiOptions,
                          out: *mut *mut FluidMulti) -> c_int;
    fn fluid_multi_destroy(m: *mut FluidMulti);
    fn fluid_multi_tick(m: *mut FluidMulti, tick: *const TickSettings) -> c_int;
    fn fluid_multi_sync(m: *mut FluidMulti) -> c_int;
    fn fluid_multi_tick_count(m: *const FluidMulti) -> u32;
    fn fluid_multi_upload_particles(m: *mut FluidMulti, all: *const ParticleInstance, n: u32) -> c_int;
    fn fluid_multi_download_particles(m: *mut FluidMulti, dst: *mut ParticleInstance) -> c_int;
    fn fluid_multi_download_start_indices(m: *mut FluidMulti, dst: *mut u32) -> c_int;
    fn fluid_multi_upload_force_field(m: *mut FluidMulti, xy: *const f32) -> c_int;
    fn fluid_multi_particle_counts(m: *mut FluidMulti, counts: *mut u32) -> c_int;
    // jump flood in the library's static device scratch (no allocation in the call; at most 2^20 pixels; not in the reference)
    fn fluid_jfa_run_static(seeds_in: *const u32, w: u32, h: u32, seeds_out: *mut u32, device_ordinal: c_int, stream: *mut c_void, wait: c_int) -> c_int;
}

/// `fluid_jfa_run` for hosts that may not allocate device memory: `seeds` is an rg32uint image (w*h pairs), host or device memory.
pub fn jfa_run_static(seeds: &[u32], w: u32, h: u32, out: &mut [u32], device: i32) {
    assert!(seeds.len() == (w as usize) * (h as usize) * 2 && out.len() == seeds.len());
    check(unsafe { fluid_jfa_run_static(seeds.as_ptr(), w, h, out.as_mut_ptr(), device as c_int, std::ptr::null_mut(), 1) });
}

fn check(rc: c_int) {
    if rc != 0 {
        // the reference's signatures are infallible and panic inside wgpu on error; keep that behaviour
        let msg = unsafe { CStr::from_ptr(fluid_last_error()) }.to_string_lossy().into_owned();
        panic!("libfluid_b200 error {rc}: {msg}");
    }
}

pub struct FluidSimulation { handle: *mut FluidSim, settings: SimulationSettings, pub tick: u32 }
unsafe impl Send for FluidSimulation {}   // one caller thread at a time; deliberately not Sync

impl FluidSimulation {
    /// reference: `FluidSimulation::new(&wgpu::Device, SimulationSettings)`, src/simulation.rs:139
    pub fn new(device_ordinal: i32, settings: SimulationSettings) -> Self {
        let mut handle = std::ptr::null_mut();
        check(unsafe { fluid_sim_create(&settings, device_ordinal, &mut handle) });
        Self { handle, settings, tick: 0 }
    }
    /// reference: `tick(&mut self, &Queue, &mut CommandEncoder, TickSettings)`, src/simulation.rs:459 — records only
    pub fn tick(&mut self, settings: TickSettings) {
        check(unsafe { fluid_sim_tick(self.handle, &settings) });
        self.tick = unsafe { fluid_sim_tick_count(self.handle) };
    }
    /// reference: `device.poll(Wait)`, src/main.rs:79
    pub fn sync(&mut self) { check(unsafe { fluid_sim_sync(self.handle) }); }
    /// reference: `queue.write_buffer(simulation.force_field_texture(), 0, cast_slice(&field))`, src/renderer.rs:498-502
    pub fn write_force_field(&mut self, field: &[[f32; 2]]) {
        assert_eq!(field.len(), (self.settings.texture_size[0] * self.settings.texture_size[1]) as usize);
        check(unsafe { fluid_sim_upload_force_field(self.handle, field.as_ptr() as *const f32) });
    }
    /// headless stand-in for binding 0 of `simulation_bg()` (src/simulation.rs:557): sorted AoS records
    pub fn download_particles(&mut self) -> Vec<ParticleInstance> {
        let mut v = vec![ParticleInstance::default(); self.settings.particle_count as usize];
        check(unsafe { fluid_sim_download_particles(self.handle, v.as_mut_ptr()) });
        v
    }
    /// binding 1 (`start_indices`)
    pub fn download_start_indices(&mut self) -> Vec<u32> {
        let n = unsafe { fluid_sim_grid_w(self.handle) * fluid_sim_grid_h(self.handle) } as usize;
        let mut v = vec![0u32; n];
        check(unsafe { fluid_sim_download_start_indices(self.handle, v.as_mut_ptr()) });
        v
    }
    /// row f1 — replaces the object render pass, the texture readback, the CPU field thread and the `write_buffer`
    /// of src/renderer.rs:449-547: objects -> mask -> seeds -> jump flood -> force field, on the sim's stream, in
    /// the same frame. `inv_proj` = `self.projection.inverse().to_cols_array()` (src/renderer.rs:450).
    pub fn set_objects(&mut self, objects: &[FluidObject], inv_proj: &[f32; 16]) {
        use std::ptr::null_mut;
        check(unsafe { fluid_sim_field_from_objects(self.handle, objects.as_ptr(), objects.len() as u32, inv_proj.as_ptr(),
                                                    null_mut(), null_mut(), null_mut(), null_mut()) });
    }
    /// the same from a ready mask (the reference's `Image`, src/main.rs:379-399): 0 = obstacle, texture-sized
    pub fn set_obstacle_mask(&mut self, mask: &[u8]) {
        use std::ptr::null_mut;
        assert_eq!(mask.len(), (self.settings.texture_size[0] * self.settings.texture_size[1]) as usize);
        check(unsafe { fluid_sim_field_from_mask(self.handle, mask.as_ptr(), null_mut(), null_mut(), null_mut()) });
    }
    /// row f2 — headless fluid_shader.wgsl (src/renderer.rs:430-462): the blended RGBA8 image, row 0 = top
    pub fn render_rgba8(&mut self, inv_proj: &[f32; 16], w: u32, h: u32) -> Vec<u8> {
        let mut img = vec![0u8; (w * h * 4) as usize];
        check(unsafe { fluid_sim_render(self.handle, inv_proj.as_ptr(), w, h, std::ptr::null_mut(), img.as_mut_ptr(), std::ptr::null_mut()) });
        img
    }
    /// the per-frame form: blended image only, optionally sRGB-encoded as the reference's preferred surface format does
    /// on store (src/renderer.rs:137-140); `wait = false` only enqueues (pinned `img`, complete after `sync`)
    pub fn render_frame(&mut self, inv_proj: &[f32; 16], w: u32, h: u32, img: &mut [u8], srgb: bool, wait: bool) {
        assert_eq!(img.len(), (w * h * 4) as usize);
        check(unsafe { fluid_sim_render_frame(self.handle, inv_proj.as_ptr(), w, h, img.as_mut_ptr(), srgb as c_int, wait as c_int) });
    }
    /// row f4 — compute.wgsl:300-498 per sorted particle; a diagnostic, not applied (the call at :92 is commented out)
    pub fn surface_tension(&mut self) -> Vec<[f32; 2]> {
        let mut v = vec![[0f32; 2]; self.settings.particle_count as usize];
        check(unsafe { fluid_sim_surface_tension(self.handle, v.as_mut_ptr() as *mut f32) });
        v
    }
    /// reference: Renderer::restart_simulation, src/renderer.rs:870-872
    /// 0 = auto (counting sort over the cells unless they are crowded), 1 = radix passes, 2 = counting; same order either way.
    pub fn set_sort_mode(&mut self, mode: i32) { check(unsafe { fluid_sim_set_sort_mode(self.handle, mode) }); }
    pub fn last_sort_path(&self) -> i32 { unsafe { fluid_sim_last_sort_path(self.handle) } }
    /// Pipelined whole-state transfers (pinned slices; untouched until `sync_transfers`): the upload of the next state
    /// overlaps the download of the previous result.
    pub unsafe fn upload_particles_async(&mut self, src: &[ParticleInstance]) {
        assert_eq!(src.len(), self.settings.particle_count as usize);
        check(fluid_sim_upload_particles_async(self.handle, src.as_ptr()));
    }
    pub unsafe fn download_particles_async(&mut self, dst: &mut [ParticleInstance]) {
        assert_eq!(dst.len(), self.settings.particle_count as usize);
        check(fluid_sim_download_particles_async(self.handle, dst.as_mut_ptr()));
    }
    /// after writing position / velocity through raw device pointers (zero-copy consumers)
    pub fn state_changed(&mut self) { check(unsafe { fluid_sim_state_changed(self.handle) }); }
    pub fn sync_transfers(&mut self) { check(unsafe { fluid_sim_sync_transfers(self.handle) }); }
    pub fn restart(&mut self, device_ordinal: i32) { *self = Self::new(device_ordinal, self.settings); }
    pub fn cuda_stream(&mut self) -> *mut c_void { unsafe { fluid_sim_stream(self.handle) } }
}

impl Drop for FluidSimulation { fn drop(&mut self) { unsafe { fluid_sim_destroy(self.handle) } } }

/// `FluidSimulation` over several GPUs of one box: bands of grid rows, one per device, stepped by ONE call per frame like
/// the reference's `tick` (src/simulation.rs:459, called from src/renderer.rs:415-418). The library enqueues every band's
/// kernels and the tick's single NCCL exchange itself and reads nothing back; this thread never blocks in `tick`.
pub struct MultiFluidSimulation { handle: *mut FluidMulti, settings: SimulationSettings, n_ranks: usize, pub tick: u32 }
unsafe impl Send for MultiFluidSimulation {}

impl MultiFluidSimulation {
    /// `FluidSimulation::new_multi`: `devices[r]` = CUDA ordinal of rank r (one GPU per rank: NCCL over NVLink)
    pub fn new_multi(devices: &[i32], settings: SimulationSettings, options: FluidMultiOptions) -> Self {
        let mut handle = std::ptr::null_mut();
        check(unsafe { fluid_multi_create(&settings, devices.len() as c_int, devices.as_ptr(), &options, &mut handle) });
        Self { handle, settings, n_ranks: devices.len(), tick: 0 }
    }
    pub fn tick(&mut self, settings: TickSettings) {
        check(unsafe { fluid_multi_tick(self.handle, &settings) });
        self.tick = unsafe { fluid_multi_tick_count(self.handle) };
    }
    /// `device.poll(Wait)`; panics if a rank latched an overflow (an exchange window or a band's capacity was too small)
    pub fn sync(&mut self) { check(unsafe { fluid_multi_sync(self.handle) }); }
    pub fn write_force_field(&mut self, field: &[[f32; 2]]) {
        assert_eq!(field.len(), (self.settings.texture_size[0] * self.settings.texture_size[1]) as usize);
        check(unsafe { fluid_multi_upload_force_field(self.handle, field.as_ptr() as *const f32) });
    }
    pub fn upload_particles(&mut self, all: &[ParticleInstance]) {
        check(unsafe { fluid_multi_upload_particles(self.handle, all.as_ptr(), all.len() as u32) });
    }
    /// the whole domain in key order: the bytes a single `FluidSimulation` would hold
    pub fn download_particles(&mut self) -> Vec<ParticleInstance> {
        let mut v = vec![ParticleInstance::default(); self.settings.particle_count as usize];
        check(unsafe { fluid_multi_download_particles(self.handle, v.as_mut_ptr()) });
        v
    }
    pub fn download_start_indices(&mut self, cells: usize) -> Vec<u32> {
        let mut v = vec![0u32; cells];
        check(unsafe { fluid_multi_download_start_indices(self.handle, v.as_mut_ptr()) });
        v
    }
    pub fn particle_counts(&mut self) -> Vec<u32> {
        let mut v = vec![0u32; self.n_ranks];
        check(unsafe { fluid_multi_particle_counts(self.handle, v.as_mut_ptr()) });
        v
    }
}
impl Drop for MultiFluidSimulation { fn drop(&mut self) { unsafe { fluid_multi_destroy(self.handle) } } }
