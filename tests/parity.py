"""Shared parity helpers: run one tick from an identical input state on the oracle and on the CUDA
library (through the C ABI) and compare with the criteria of BASELINE.json's north star:
  (1) keys, sorted keys, start_indices: bit-exact;
  (2) density / position: fp32 relative 1e-5; velocity (force-like): |a-b| <= 1e-5 * max(|ref|, sum|terms|*dt/rho)
  (3) JFA: exact.
"""
import numpy as np

RTOL = 1e-5


def run_pair(fluid, oracle, settings, tick_settings, particles, table=None, field=None, tick_count=0, n_ticks=1,
             sort_mode=None, debug_force=True, math_mode=0, gpu_sort_mode=None):
    """returns (oracle_sim, cuda_sim, extras) after n_ticks ticks from the same state."""
    O = oracle
    osim = O.OracleSim(settings, particles, sort_mode=O.SORT_STABLE if sort_mode is None else sort_mode)
    sim = fluid.FluidSimulation.new(0, settings)
    sim.math_mode = math_mode
    if gpu_sort_mode is not None:
        sim.sort_mode = gpu_sort_mode
    sim.upload_particles(particles)
    if table is not None:
        osim.table[:] = table
        sim.upload_start_indices(table)
    if field is not None:
        osim.texture[:] = field
        sim.upload_force_field(field)
    osim.tick_count = tick_count
    sim.tick_count = tick_count
    if debug_force:
        sim.enable_debug_force(True)
    extras = {}
    for k in range(n_ticks):
        if k == n_ticks - 1:
            # capture the oracle's pre-sort keys and debug force of the last tick
            u = O.make_uniform(settings, tick_settings, osim.tick_count + 1)
            tmp = osim.p.copy()
            O.predict_next_position(u, tmp)
            O.create_spatial_lookup(u, tmp)
            extras["oracle_unsorted_keys"] = tmp["grid"].copy()
            if debug_force:
                tmp_table = osim.table.copy()
                O.sort_stable(tmp) if osim.sort_mode != O.SORT_BITONIC else O.sort_bitonic(tmp)
                O.compute_start_indices(u, tmp, tmp_table)
                O.calculate_density(u, tmp, tmp_table, osim.walk_radius)
                _, accel, absum = O.move_particle(u, tmp, tmp_table, osim.texture, want_debug=True)
                extras["oracle_accel"], extras["oracle_abs"] = accel, absum
        osim.tick(tick_settings)
        sim.tick(tick_settings)
    sim.sync()
    return osim, sim, extras


def rel_err(a, b, floor=0.0):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.abs(a - b) / np.maximum(np.abs(b), floor if floor > 0 else 1e-300)


def compare_state(osim, sim, extras, tick_settings, check_bit_exact_floats=True):
    """asserts criteria (1) and (2); returns a dict of summary statistics."""
    got = sim.download_particles()
    ref = osim.p
    n = ref.shape[0]
    stats = {}
    # (1) bit-exact integer results
    assert np.array_equal(sim.download_unsorted_keys(), extras["oracle_unsorted_keys"]), "hash keys differ"
    assert np.array_equal(got["grid"], ref["grid"]), "sorted key arrays differ"
    assert np.array_equal(sim.download_start_indices(), osim.table), "start_indices tables differ"
    # (2) floats
    assert np.array_equal(got["predicted_position"], ref["predicted_position"]), "predicted positions differ"
    d_err = rel_err(got["density"], ref["density"]).max()
    stats["density_max_rel"] = float(d_err)
    assert d_err <= RTOL, f"density rel err {d_err}"
    dt = tick_settings.delta
    if "oracle_accel" in extras:
        accel, absum = sim.download_debug_force()
        a_err = np.abs(accel.astype(np.float64) - extras["oracle_accel"]) / np.maximum(
            np.maximum(np.abs(extras["oracle_accel"]), extras["oracle_abs"]), 1e-30)
        stats["accel_max_rel_to_terms"] = float(a_err.max())
        assert a_err.max() <= RTOL, f"acceleration err {a_err.max()}"
        vel_scale = np.maximum(np.abs(ref["velocity"]), extras["oracle_abs"] * dt / ref["density"][:, None])
    else:
        vel_scale = np.abs(ref["velocity"])
    v_err = np.abs(got["velocity"].astype(np.float64) - ref["velocity"]) / np.maximum(vel_scale, 1e-30)
    stats["velocity_max_rel"] = float(v_err.max())
    assert v_err.max() <= RTOL, f"velocity err {v_err.max()}"
    # position: relative to the domain scale the position lives on (|pos| can be ~0 at the centre)
    p_scale = np.maximum(np.abs(ref["position"]), vel_scale * dt)
    p_err = np.abs(got["position"].astype(np.float64) - ref["position"]) / np.maximum(p_scale, 1e-30)
    stats["position_max_rel"] = float(p_err.max())
    assert p_err.max() <= RTOL, f"position err {p_err.max()}"
    exact = (np.array_equal(got["density"], ref["density"]) and np.array_equal(got["velocity"], ref["velocity"])
             and np.array_equal(got["position"], ref["position"]))
    stats["floats_bit_exact"] = bool(exact)
    stats["n"] = int(n)
    if check_bit_exact_floats:
        # same tie order (stable) + same fp32 operation order => identical bits, stronger than the tolerance
        nd = int((got["density"] != ref["density"]).sum())
        nv = int((got["velocity"] != ref["velocity"]).any(axis=1).sum())
        npz = int((got["position"] != ref["position"]).any(axis=1).sum())
        stats["mismatch_counts"] = (nd, nv, npz)
    return stats
