/* c_abi_host.c — a host that is not Python walks the C ABI of include/fluid_b200.h in the order the reference's renderer
 * drives FluidSimulation (reference src/renderer.rs): construct (:156), upload the force field (:499), tick per frame
 * (:417), read the particles back (the headless stand-in for the bind groups of :457-458), reconstruct on restart (:874).
 * It is the call sequence of rust/src/simulation_cuda.rs, compiled with plain gcc (no Rust toolchain exists on this
 * image), and it also drives the multi-GPU block the same way (two bands sharing GPU 0) and checks that both give the
 * same bytes. Built and run by tests/test_c_abi_host.py on the GPU box:
 *   gcc -O1 -I include tests/c_abi_host.c -L gpu_fluid_simulation_b200 -lfluid_b200 -Wl,-rpath,... -o c_abi_host
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "fluid_b200.h"

#define CHECK(call)                                                                              \
    do {                                                                                         \
        int rc__ = (call);                                                                       \
        if (rc__ != FLUID_OK) {                                                                  \
            fprintf(stderr, "%s -> %d: %s\n", #call, rc__, fluid_last_error());                  \
            return 1;                                                                            \
        }                                                                                        \
    } while (0)

int main(void) {
    if (fluid_abi_version() != FLUID_B200_ABI_VERSION) { fprintf(stderr, "ABI version mismatch\n"); return 1; }
    if (sizeof(FluidParticle) != 32 || sizeof(FluidSimUniform) != 120) { fprintf(stderr, "struct layout\n"); return 1; }
    /* the reference's default scene, scaled down (src/main.rs:48-54) */
    FluidSimSettings s;
    memset(&s, 0, sizeof s);
    s.particle_count = 20000; s.particle_spacing = 0.1f; s.smoothing_radius = 0.2f;
    s.size[0] = 24.0f; s.size[1] = 24.0f; s.texture_size[0] = 64; s.texture_size[1] = 64;
    /* default TickSettings (src/renderer.rs:374-388) + gravity */
    FluidTickSettings t;
    memset(&t, 0, sizeof t);
    t.delta = 1.0f / 120.0f; t.gravity[1] = 9.8f; t.mass = 1.0f; t.pressure_constant = 50.0f; t.damping_factor = 0.1f;
    t.viscosity_coefficient = 25.0f; t.surface_tension_treshold = 0.1f; t.surface_tension_coefficient = 35.0f;
    t.mouse_force_radius = 5.0f; t.mouse_force_power = 150.0f;

    const size_t texels = (size_t)s.texture_size[0] * s.texture_size[1];
    float* field = (float*)calloc(texels * 2, sizeof(float));
    FluidParticle* a = (FluidParticle*)malloc(sizeof(FluidParticle) * s.particle_count);
    FluidParticle* b = (FluidParticle*)malloc(sizeof(FluidParticle) * s.particle_count);
    if (!field || !a || !b) return 1;
    field[2 * (texels / 2 + 7)] = 1.0f; /* one non-zero texel: the push-out branch (compute.wgsl:131-137) */

    FluidSim* sim = NULL;
    CHECK(fluid_sim_create(&s, 0, &sim));                      /* FluidSimulation::new, src/renderer.rs:156 */
    CHECK(fluid_sim_upload_force_field(sim, field));           /* queue.write_buffer(force_field_texture()), :498-502 */
    for (int frame = 0; frame < 3; frame++) CHECK(fluid_sim_tick(sim, &t)); /* Renderer::tick, :415-418 */
    CHECK(fluid_sim_sync(sim));                                /* device.poll(Wait), src/main.rs:79 */
    if (fluid_sim_tick_count(sim) != 3) { fprintf(stderr, "tick count\n"); return 1; }
    CHECK(fluid_sim_download_particles(sim, a));
    FluidSimUniform u;
    CHECK(fluid_sim_get_uniform(sim, &u));
    if (u.frame_time != 3 || u.particle_count != s.particle_count || u.grid_w != fluid_sim_grid_w(sim)) { fprintf(stderr, "uniform\n"); return 1; }
    for (uint32_t i = 1; i < s.particle_count; i++)
        if (a[i].grid < a[i - 1].grid) { fprintf(stderr, "particles are not in key order after a tick\n"); return 1; }
    fluid_sim_destroy(sim);                                    /* restart_simulation drops and rebuilds, :873-875 */

    /* the same frames through the multi-GPU block: two bands of grid rows sharing GPU 0 */
    FluidMultiOptions opt;
    memset(&opt, 0, sizeof opt);
    opt.transport = FLUID_TRANSPORT_LOCAL;
    const int devices[2] = {0, 0};
    FluidMulti* multi = NULL;
    CHECK(fluid_multi_create(&s, 2, devices, &opt, &multi));
    CHECK(fluid_multi_upload_force_field(multi, field));
    for (int frame = 0; frame < 3; frame++) CHECK(fluid_multi_tick(multi, &t));
    CHECK(fluid_multi_sync(multi));
    CHECK(fluid_multi_download_particles(multi, b));
    fluid_multi_destroy(multi);
    if (memcmp(a, b, sizeof(FluidParticle) * s.particle_count) != 0) { fprintf(stderr, "multi-GPU block differs from one GPU\n"); return 1; }

    /* error behaviour: the reference panics on particle_count < 2 (src/simulation.rs:323-324); the ABI returns a status */
    s.particle_count = 1;
    if (fluid_sim_create(&s, 0, &sim) != FLUID_ERR_INVALID_ARGUMENT) { fprintf(stderr, "particle_count < 2 must be rejected\n"); return 1; }
    printf("c_abi_host ok: 3 frames, %u particles, key-ordered, multi == single\n", 20000u);
    free(field); free(a); free(b);
    return 0;
}
