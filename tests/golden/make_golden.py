"""Generates the golden fixtures under tests/golden/ from the CPU oracle.

The reference ships no tests, golden vectors or fixtures and cannot be built or run in this environment
(Rust + WGSL, needs a graphics API), so these vectors come from the oracle (oracle/fluid_oracle.cpp), not
from the reference itself: parity stays "unpinned" in the sense of the task statement. They pin the
oracle against regressions and give the GPU tests fixed inputs/outputs that travel to the GPU box.

    python tests/golden/make_golden.py        # rewrites the .npz files
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from gpu_fluid_simulation_b200 import SimulationSettings, TickSettings, scenes  # noqa: E402


def _settings_arrays(s, ts):
    return dict(
        settings=np.array([s.particle_count, s.texture_size[0], s.texture_size[1]], np.uint32),
        settings_f=np.array([s.particle_spacing, s.smoothing_radius, s.size[0], s.size[1]], np.float64),
        tick_f=np.array([ts.delta, ts.gravity[0], ts.gravity[1], ts.mass, ts.pressure_constant, ts.rest_density,
                         ts.damping_factor, ts.viscosity_coefficient, ts.mouse_force_radius, ts.mouse_force_power,
                         ts.mouse_pos[0], ts.mouse_pos[1], ts.mouse_state], np.float64))


def load_settings(z):
    n, tw, th = (int(v) for v in z["settings"])
    sp, h, sx, sy = (float(v) for v in z["settings_f"])
    t = [float(v) for v in z["tick_f"]]
    s = SimulationSettings(particle_count=n, particle_spacing=sp, smoothing_radius=h, size=(sx, sy), texture_size=(tw, th))
    ts = TickSettings(delta=t[0], gravity=(t[1], t[2]), mass=t[3], pressure_constant=t[4], rest_density=t[5],
                      damping_factor=t[6], viscosity_coefficient=t[7], mouse_force_radius=t[8], mouse_force_power=t[9],
                      mouse_pos=(t[10], t[11]), mouse_state=int(t[12]))
    return s, ts


def _run(O, s, ts, p, ticks, field=None, table=None):
    sim = O.OracleSim(s, p)
    if field is not None:
        sim.texture[:] = field
    if table is not None:
        sim.table[:] = table
    for _ in range(ticks):
        sim.tick(ts)
    out = dict(input=p, output=sim.p.copy(), table=sim.table.copy(), ticks=np.array([ticks], np.uint32))
    if field is not None:
        out["field"] = field
    if table is not None:
        out["table_in"] = table
    out.update(_settings_arrays(s, ts))
    return out


def case_uniform(O):
    s, ts, p = scenes.uniform_box(48, tex=(64, 64))
    return _run(O, s, ts, p, 3)


def case_dambreak(O):
    s, ts, p = scenes.dam_break(24, 96, tex=(64, 64))
    return _run(O, s, ts, p, 4)


def case_default_small(O):
    # the reference's sheared init lattice (coincident pairs -> RNG branch), scaled down
    s = SimulationSettings(particle_count=5000, size=(12.0, 12.0), texture_size=(64, 64))
    return _run(O, s, TickSettings(), O.init_lattice(s), 2)


def case_field_mouse(O):
    s, ts, p = scenes.uniform_box(40, tex=(32, 32))
    ts.mouse_state = -1
    ts.mouse_pos = (0.3, -0.2)
    ts.mouse_force_radius = 1.5
    rng = np.random.default_rng(5)
    field = (rng.integers(-3, 4, size=(32, 32, 2))).astype(np.float32)
    field[rng.random((32, 32)) < 0.5] = 0
    table = rng.integers(0, 3000, size=(O.grid_dims(s)[0] * O.grid_dims(s)[1]), dtype=np.uint32)
    return _run(O, s, ts, p, 2, field=field, table=table)


def case_jfa(O):
    rng = np.random.default_rng(11)
    h, w = 50, 70
    mask = np.where(rng.random((h, w)) < 0.03, 255, 0).astype(np.uint8)
    seeds = O.jfa_seed_from_mask(mask)
    out = O.jfa_run(seeds)
    return dict(mask=mask, seeds=seeds, out=out, field=O.jfa_to_force_field(out),
                pass8=O.jfa_pass(seeds, 8), sweep=O.generate_smooth_gradient_field(mask))


CASES = {
    "uniform_48": case_uniform,
    "dambreak_24x96": case_dambreak,
    "default_5000": case_default_small,
    "field_mouse_40": case_field_mouse,
    "jfa_70x50": case_jfa,
}

if __name__ == "__main__":
    from oracle import oracle as O
    O.build()
    for name, fn in CASES.items():
        data = fn(O)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **data)
        print(name, {k: v.shape for k, v in data.items()})
