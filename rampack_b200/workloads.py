"""Synthetic configurations of BASELINE.json / SURVEY.md §8d (deterministic lattices; no files, no network)."""
import numpy as np

from . import capi

# name -> (n_cells, unit cell): hard spherocylinders l = 5, r = 0.5 on a simple cubic lattice, axes along z,
# packing fraction 0.475
SC_CONFIGS = {
    "sc_12k": ((28, 28, 16), (1.2, 1.2, 6.5)),
    "sc_100k": ((56, 56, 32), (1.2, 1.2, 6.5)),
    "sc_1m": ((120, 120, 70), (1.2, 1.2, 6.5)),
    "sc_4m": ((192, 192, 108), (1.2, 1.2, 6.5)),
}
SC_LENGTH, SC_RADIUS = 5.0, 0.5


def shapes_from(pos, rot):
    n = pos.shape[0]
    out = np.empty((n, 12))
    out[:, :3] = pos
    out[:, 3:] = rot.reshape(n, 9)
    return out


def sc_lattice(ncells, cell=(1.2, 1.2, 6.5)):
    """Returns (box 3x3, shapes N x 12) — positions at the cell centres, identity orientations."""
    nx, ny, nz = ncells
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    pos = np.stack([(ix.ravel() + 0.5) * cell[0], (iy.ravel() + 0.5) * cell[1], (iz.ravel() + 0.5) * cell[2]], axis=1)
    box = np.diag([nx * cell[0], ny * cell[1], nz * cell[2]]).astype(np.float64)
    rot = np.tile(np.eye(3), (pos.shape[0], 1, 1))
    return box, shapes_from(pos, rot)


def sc_config(name):
    ncells, cell = SC_CONFIGS[name]
    return sc_lattice(ncells, cell)


def random_rotations(rng, n):
    """Uniform rotations (normalised Gaussian quaternions)."""
    q = rng.normal(size=(n, 4))
    q /= np.linalg.norm(q, axis=1)[:, None]
    w, x, y, z = q.T
    return np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                     2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                     2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)], axis=1).reshape(n, 3, 3)


def kmer_desc(k, r=0.5, distance=1.0, soft=capi.RP_SOFT_NONE, epsilon=1.0, sigma=1.0, hard=True):
    """KMerTraits geometry (src/core/shapes/KMerTraits.cpp:10-32): k beads on the z axis, centred."""
    z0 = -distance * (k - 1) / 2
    centres = [(0.0, 0.0, z0 + i * distance) for i in range(k)]
    return capi.make_desc(hard=capi.RP_HARD_POLYSPHERE if hard else capi.RP_HARD_NONE, soft=soft, epsilon=epsilon,
                          sigma=sigma, centres=centres, centre_radius=[r] * k)


def convex_desc():
    """generic_convex(wrap(sphere r=0.5 @z=-0.5, cuboid 1x1x1 @z=+0.5 rotated by 45 degrees)), docs/shapes.md:656-659."""
    from .xc import XCBodyBuilder, make_xc_desc
    return make_xc_desc(XCBodyBuilder.from_script("sphere 0.5 move 0 0 -0.5 cuboid 1 1 1 rot 0 0 45 move 0 0 0.5 wrap"))


# ---------------------------------------------------------------------------------------------------------------------
# The configurations of BASELINE.json `configs` (C1 - C4; C5 is sc_4m above) as bench workloads.  Each entry:
#   desc       rp_interaction_desc factory
#   lattice    (cells, unit cell, random orientations?)          full-size configuration
#   cpu_cells  cells of the same lattice for the bounded CPU sample (same density, same per-move work)
#   p, moves   pressure and the rototranslation step sizes the thermalisation starts from (then adapted)
#   scaler     box-move scaler (NativeSimulation notation); the step is adapted during the thermalisation
#   relax      start from this linear compression of the lattice with overlap counting on (overlap_relaxation run,
#              sample_inputs/overlap_reduction.pyon) and time that phase separately
#   flops      per-pair operation counts of the reference arithmetic: (F_pre, F_full)
# ---------------------------------------------------------------------------------------------------------------------
def _desc_sphere():
    return capi.make_desc(hard=capi.RP_HARD_SPHERE, radius=0.5)


def _desc_sc():
    return capi.make_desc(hard=capi.RP_HARD_SPHEROCYLINDER, radius=SC_RADIUS, length=SC_LENGTH)


EXTRA_CONFIGS = {
    # C1: sample_inputs/integration.pyon ("gas" run) verbatim geometry, and a 10^6 version at liquid density
    "hs_1k_gas": dict(desc=_desc_sphere, lattice=((10, 10, 10), (10.0, 10.0, 10.0), False), cpu_cells=(10, 10, 10),
                      p=0.1, moves=(10.0, 0.1), scaler=("delta_v", 1000.0), flops=(12, 0), therm=2000),
    "hs_1m_liquid": dict(desc=_desc_sphere, lattice=((100, 100, 100), (1.1, 1.1, 1.1), False), cpu_cells=(30, 30, 30),
                         p=5.0, moves=(0.1, 0.1), scaler=("delta_v", 0.5), flops=(12, 0), therm=2000),
    # C2: hard spherocylinders L/D = 5, N = 100 352, pressure scan (isotropic / nematic / smectic side of the scan)
    "sc_100k_p0.8": dict(desc=_desc_sc, lattice=((56, 56, 32), (1.2, 1.2, 6.5), False), cpu_cells=(28, 28, 16),
                         p=0.8, moves=(0.05, 0.05), scaler=("delta_triclinic", 3e-5, False), flops=(14, 105),
                         therm=15000, observe=("nematic",)),
    "sc_100k_p1.4": dict(desc=_desc_sc, lattice=((56, 56, 32), (1.2, 1.2, 6.5), False), cpu_cells=(28, 28, 16),
                         p=1.4, moves=(0.05, 0.05), scaler=("delta_triclinic", 3e-5, False), flops=(14, 105),
                         therm=5000, observe=("nematic",)),
    "sc_100k_p2.0": dict(desc=_desc_sc, lattice=((56, 56, 32), (1.2, 1.2, 6.5), False), cpu_cells=(28, 28, 16),
                         p=2.0, moves=(0.05, 0.05), scaler=("delta_triclinic", 3e-5, False), flops=(14, 105),
                         therm=5000, observe=("nematic",)),
    # C3: 4-mers and a XenoCollide convex body, N = 2 * 10^5, started 10 % (linear) too dense with overlap counting
    # (relax_desc: the helper_shape of sample_inputs/overlap_reduction.pyon — the same beads with a
    # square_inverse_core(epsilon=2, sigma=1) potential added for the overlap_relaxation run)
    "kmer4_200k": dict(desc=lambda: kmer_desc(4), lattice=((58, 58, 60), (1.25, 1.25, 4.3), False),
                       cpu_cells=(20, 20, 20), p=20.0, moves=(0.1, 0.1), scaler=("delta_v", 1.5), flops=(13, 0),
                       relax=0.9, relax_desc=lambda: kmer_desc(4, soft=capi.RP_SOFT_SIC, epsilon=2.0, sigma=1.0),
                       therm=200),
    "convex_200k": dict(desc=convex_desc, lattice=((58, 58, 60), (1.5, 1.5, 2.3), False), cpu_cells=(20, 20, 20),
                        p=4.0, moves=(0.1, 0.1), scaler=("delta_v", 1.0), flops=(20, 400), relax=0.9, therm=200),
    # C4: soft 3-mers (energy path), N = 5 * 10^5 (1.5 * 10^6 interaction centres)
    "kmer3_wca_500k": dict(desc=lambda: kmer_desc(3, soft=capi.RP_SOFT_WCA, hard=False),
                           lattice=((80, 80, 79), (1.35, 1.35, 3.3), True), cpu_cells=(20, 20, 20), p=1.0,
                           moves=(0.1, 0.1), scaler=("delta_v", 50.0), flops=(12, 10), therm=200),
    "kmer3_lj_500k": dict(desc=lambda: kmer_desc(3, soft=capi.RP_SOFT_LJ, hard=False),
                          lattice=((80, 80, 79), (1.35, 1.35, 3.3), True), cpu_cells=(20, 20, 20), p=1.0,
                          moves=(0.1, 0.1), scaler=("delta_v", 50.0), flops=(12, 10), therm=100),
}


def extra_lattice(cfg, cells=None, seed=1234):
    ncells, cell, randomise = cfg["lattice"]
    ncells = cells or ncells
    box, shapes = sc_lattice(ncells, cell)
    if randomise:
        rot = random_rotations(np.random.default_rng(seed), shapes.shape[0])
        shapes[:, 3:] = rot.reshape(-1, 9)
    if cfg.get("relax"):
        f = cfg["relax"]
        box = box * f
        shapes[:, :3] *= f
    return box, shapes
