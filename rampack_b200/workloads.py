"""Synthetic configurations of BASELINE.json / SURVEY.md §8d (deterministic lattices; no files, no network)."""
import numpy as np

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
