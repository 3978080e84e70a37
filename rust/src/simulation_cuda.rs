//! simulation_cuda.rs — `FluidSimulation` on top of libfluid_b200 (include/fluid_b200.h).
//!
//! Drop-in for the reference's `src/simulation.rs` behind `#[cfg(feature = "cuda")]`: same type and
//! method names (`FluidSimulation::new`, `tick`, `pub tick`), same `SimulationSettings` / `TickSettings`
//! field order (src/simulation.rs:95-122). The wgpu bind-group getters (src/simulation.rs:542-564) have no
//! CUDA meaning; the renderer reads particles through `download_particles` (sorted AoS, the same bytes the
//! reference's vertex/storage buffer holds) or the raw device pointers of `device_view`.
//!
//! NOT COMPILED in the build container (no rustc/cargo there); checked only by review against the header.
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

#[repr(C)] pub struct FluidSim { _opaque: [u8; 0] }

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
    pub fn cuda_stream(&mut self) -> *mut c_void { unsafe { fluid_sim_stream(self.handle) } }
}

impl Drop for FluidSimulation { fn drop(&mut self) { unsafe { fluid_sim_destroy(self.handle) } } }
