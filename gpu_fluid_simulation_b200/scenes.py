"""Synthetic scenes of SURVEY.md §8(d): uniform box (U) and dam-break (D).

Inputs are generated once on the host (splitmix64, fixed seeds) and handed identically to the CPU checker
and to the CUDA library through upload_particles, so the PRNG is not part of parity.
"""
import numpy as np

from ._lib import PARTICLE_DTYPE
from .simulation import SimulationSettings, TickSettings

SEED_JITTER = 0xF1D00001
SEED_VEL = 0xF1D00002
_M = (1 << 64) - 1


def splitmix64(seed: int, count: int, start: int = 0) -> np.ndarray:
    """draws start .. start+count-1 of the splitmix64 stream as doubles in [0,1) (state_i = seed + (i+1)*golden, so any
    sub-range can be generated on its own: each rank of a multi-GPU run builds only its band)."""
    with np.errstate(over="ignore"):
        i = np.arange(start + 1, start + count + 1, dtype=np.uint64)
        z = np.uint64(seed & _M) + i * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return (z >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def _block(cols: int, rows: int, spacing: float, x0: float, y0: float, jitter: float, vel_amp: float, i0: int = 0, i1=None):
    """particles i0 .. i1-1 of the cols x rows jittered lattice (row-major)."""
    i1 = cols * rows if i1 is None else i1
    n = i1 - i0
    idx = np.arange(i0, i1, dtype=np.int64)
    ix = (idx % cols).astype(np.float64)
    iy = (idx // cols).astype(np.float64)
    r = splitmix64(SEED_JITTER, 2 * n, 2 * i0)
    px = x0 + (ix + 0.5) * spacing + (r[0::2] * 2.0 - 1.0) * jitter * spacing
    py = y0 + (iy + 0.5) * spacing + (r[1::2] * 2.0 - 1.0) * jitter * spacing
    p = np.zeros(n, dtype=PARTICLE_DTYPE)
    p["position"][:, 0] = px.astype(np.float32)
    p["position"][:, 1] = py.astype(np.float32)
    if vel_amp != 0.0:
        v = splitmix64(SEED_VEL, 2 * n, 2 * i0)
        p["velocity"][:, 0] = ((v[0::2] * 2.0 - 1.0) * vel_amp).astype(np.float32)
        p["velocity"][:, 1] = ((v[1::2] * 2.0 - 1.0) * vel_amp).astype(np.float32)
    p["predicted_position"] = p["position"]
    return p


def uniform_box(n_side: int, spacing=0.1, h=0.2, tex=(1024, 1024), lattice_rows=None):
    """Scene U: n x n jittered lattice filling size = (n*s, n*s); vel uniform in [-1,1]^2.
    lattice_rows=(r0, r1) returns only the particles of lattice rows r0..r1-1 (a contiguous range of the full array)."""
    size = (n_side * spacing, n_side * spacing)
    settings = SimulationSettings(particle_count=n_side * n_side, particle_spacing=spacing, smoothing_radius=h,
                                  size=size, texture_size=tex)
    r0, r1 = (0, n_side) if lattice_rows is None else lattice_rows
    p = _block(n_side, n_side, spacing, -size[0] / 2, -size[1] / 2, 0.25, 1.0, r0 * n_side, r1 * n_side)
    return settings, TickSettings(), p


def dam_break(cols: int, rows: int, spacing=0.1, h=0.2, tex=(1024, 1024), lattice_rows=None):
    """Scene D: cols x rows jittered block against the left wall of a box 4x as wide; gravity (0, 9.8).
    lattice_rows=(r0, r1): only the particles of lattice rows r0..r1-1 (a contiguous range of the full array)."""
    size = (cols * spacing * 4.0, rows * spacing)
    settings = SimulationSettings(particle_count=cols * rows, particle_spacing=spacing, smoothing_radius=h,
                                  size=size, texture_size=tex)
    r0, r1 = (0, rows) if lattice_rows is None else lattice_rows
    p = _block(cols, rows, spacing, -size[0] / 2, -size[1] / 2, 0.25, 0.0, r0 * cols, r1 * cols)
    return settings, TickSettings(gravity=(0.0, 9.8)), p


def band_lattice_rows(row_bounds, rank, n_lattice_rows):
    """lattice rows of scene U / D that lie in the grid rows [row_bounds[rank], row_bounds[rank+1]): with spacing = h/2 and
    jitter < spacing/2, lattice rows 2k and 2k+1 lie in grid row k+1 (grid row 0 is the empty border row)."""
    lo, hi = max(row_bounds[rank], 1), max(row_bounds[rank + 1], 1)
    r0, r1 = min(2 * (lo - 1), n_lattice_rows), min(2 * (hi - 1), n_lattice_rows)
    return r0, max(r0, r1)


def assert_no_coincident(p: np.ndarray):
    pos = np.ascontiguousarray(p["predicted_position"]).view(np.uint64).ravel()
    assert np.unique(pos).size == pos.size, "scene contains coincident positions"
