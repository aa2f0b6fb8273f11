"""Shared generators for the parity tests (seeded synthetic configurations, interaction descriptions)."""
import numpy as np

import rampack_b200 as rp


def random_rotations(rng, n):
    """n uniformly random rotation matrices (QR of Gaussian matrices, det fixed to +1)."""
    a = rng.normal(size=(n, 3, 3))
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.einsum("nii->ni", r))[:, None, :]
    q[np.linalg.det(q) < 0, :, 0] *= -1
    return q


def shapes_from(pos, rot):
    n = pos.shape[0]
    out = np.empty((n, 12))
    out[:, :3] = pos
    out[:, 3:] = rot.reshape(n, 9)
    return out


def random_packing(rng, n, box, aligned=False):
    """n particles uniformly random in the (triclinic) box, random orientations."""
    box = np.asarray(box, dtype=np.float64).reshape(3, 3)
    rel = rng.uniform(0, 1, size=(n, 3)) * (1 - 1e-9)
    pos = rel @ box.T
    # keep strictly inside after rounding
    relback = np.linalg.solve(box, pos.T).T
    bad = (relback < 0) | (relback >= 1)
    pos[bad.any(axis=1)] = (0.5 * np.ones(3)) @ box.T
    rot = np.tile(np.eye(3), (n, 1, 1)) if aligned else random_rotations(rng, n)
    return shapes_from(pos, rot)


def sc_lattice(ncells, cell=(1.2, 1.2, 6.5)):
    """Spherocylinders l=5, r=0.5 on a simple cubic lattice of cell 1.2 x 1.2 x 6.5, axes along z."""
    nx, ny, nz = ncells
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    pos = np.stack([(ix.ravel() + 0.5) * cell[0], (iy.ravel() + 0.5) * cell[1], (iz.ravel() + 0.5) * cell[2]], axis=1)
    box = np.diag([nx * cell[0], ny * cell[1], nz * cell[2]]).astype(np.float64)
    rot = np.tile(np.eye(3), (pos.shape[0], 1, 1))
    return box, shapes_from(pos, rot)


def sphere_lattice(ncells, a):
    nx, ny, nz = ncells
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    pos = np.stack([(ix.ravel() + 0.5) * a, (iy.ravel() + 0.5) * a, (iz.ravel() + 0.5) * a], axis=1)
    box = np.diag([nx * a, ny * a, nz * a]).astype(np.float64)
    rot = np.tile(np.eye(3), (pos.shape[0], 1, 1))
    return box, shapes_from(pos, rot)


def desc_sphere(r=0.5):
    return rp.make_desc(hard=rp.RP_HARD_SPHERE, radius=r)


def desc_sc(l=5.0, r=0.5):
    return rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=r, length=l)


def desc_kmer(k=3, r=0.5, distance=1.0, soft=rp.RP_SOFT_NONE, epsilon=1.0, sigma=1.0, hard=True):
    """KMerTraits geometry (src/core/shapes/KMerTraits.cpp:10-32): k beads on the z axis, centred."""
    z0 = -distance * (k - 1) / 2
    centres = [(0.0, 0.0, z0 + i * distance) for i in range(k)]
    return rp.make_desc(hard=rp.RP_HARD_POLYSPHERE if hard else rp.RP_HARD_NONE, soft=soft, epsilon=epsilon,
                        sigma=sigma, centres=centres, centre_radius=[r] * k)


def desc_polysc():
    """Two spherocylinders forming a V (a 2-segment polyspherocylinder)."""
    centres = [(0.0, 0.0, -0.8), (0.0, 0.3, 0.8)]
    half = [(0.0, 0.0, 0.7), (0.0, 0.35, 0.6)]
    return rp.make_desc(hard=rp.RP_HARD_POLYSPHEROCYLINDER, centres=centres, centre_radius=[0.4, 0.3],
                        centre_half_axis=half)


def random_pairs(rng, n, max_dist, n_centres=1, translation_scale=0.0):
    """n random rp_pair records: pos2 within max_dist of pos1 (after the translation), random orientations."""
    pairs = np.zeros(n, dtype=rp.PAIR_DTYPE)
    pairs["pos1"] = rng.uniform(-5, 5, size=(n, 3))
    t = rng.uniform(-1, 1, size=(n, 3)) * translation_scale
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1)[:, None]
    dist = max_dist * rng.uniform(0, 1, size=n) ** (1 / 3)
    pairs["pos2"] = pairs["pos1"] + d * dist[:, None] - t
    pairs["translation"] = t
    pairs["orientation1"] = random_rotations(rng, n).reshape(n, 9)
    pairs["orientation2"] = random_rotations(rng, n).reshape(n, 9)
    pairs["idx1"] = rng.integers(0, n_centres, size=n)
    pairs["idx2"] = rng.integers(0, n_centres, size=n)
    return pairs
