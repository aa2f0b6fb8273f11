"""Lattice population on the device: rp_populate_lattice against the reference's Lattice +
SerialPopulator (core/lattice/Lattice.cpp:43-81, SerialPopulator.cpp:9-40) — positions, orientations and the lattice
box must be equal bit for bit — and rp_randomize_rotations against the distributions of
RotationRandomizingTransformer (core/lattice/RotationRandomizingTransformer.cpp:30-56)."""
import numpy as np
import pytest

import rampack_b200 as rp
import helpers as H
from oracle import checker

pytestmark = pytest.mark.gpu


def _molecules(rng, m):
    mol = np.zeros((m, 12))
    mol[:, :3] = rng.uniform(0.05, 0.95, size=(m, 3))
    mol[:, 3:] = H.random_rotations(rng, m).reshape(m, 9)
    return mol


@pytest.mark.parametrize("order", ["xyz", "zyx", "yxz"])
def test_populated_lattice_equals_the_reference_populator(order):
    if checker.kind() != "reference":
        pytest.skip("needs oracle/_ref (the reference's Lattice / SerialPopulator)")
    from oracle import refwrap as R
    rng = np.random.default_rng(7)
    cases = [
        (np.diag([1.2, 1.2, 6.5]), np.array([[0.5, 0.5, 0.5, 1, 0, 0, 0, 1, 0, 0, 0, 1.0]]), (7, 5, 3), None),
        (np.array([[2.0, 0.3, 0.1], [0.0, 1.7, -0.2], [0.1, 0.0, 2.4]]), _molecules(rng, 3), (4, 6, 5), None),
        (np.array([[2.0, 0.3, 0.1], [0.0, 1.7, -0.2], [0.1, 0.0, 2.4]]), _molecules(rng, 2), (5, 4, 6), 173),
    ]
    dev = rp.Device(0, seed=1)
    for cell, mol, dims, n in cases:
        total = dims[0] * dims[1] * dims[2] * mol.shape[0]
        n = total if n is None else n
        dev.populate_lattice(cell, mol, dims, order, n)
        box, shapes = dev.download()
        want_box, want = R.ref_populate_lattice(cell, mol, dims, order, n)
        assert np.array_equal(box, want_box)
        assert np.array_equal(shapes, want)
    dev.close()


def test_populate_a_million_spherocylinders_and_sweep():
    """BASELINE config M5's lattice (120 x 120 x 70 cells of 1.2 x 1.2 x 6.5) generated on the device equals the host
    construction the bench uploads, and is a valid overlap-free packing for the sweep."""
    dev = rp.Device(0, seed=3)
    dev.set_interaction(H.desc_sc())
    mol = np.array([[0.5, 0.5, 0.5, 1, 0, 0, 0, 1, 0, 0, 0, 1.0]])
    dev.populate_lattice(np.diag([1.2, 1.2, 6.5]), mol, (120, 120, 70))
    assert dev.n == 1008000
    box, shapes = dev.download()
    wbox, want = H.sc_lattice((120, 120, 70))
    assert np.array_equal(box, wbox)
    # (the host helper computes (i + 0.5) * a, the populator M * (i + 0.5): equal up to the last bit)
    assert np.allclose(shapes, want, rtol=0, atol=1e-12)
    assert dev.count_total_overlaps(False) == 0
    c = dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.05, 0.05, 0)
    assert c.attempted == dev.n and c.accepted > 0
    assert dev.count_total_overlaps(False) == 0
    dev.close()


def test_randomize_rotations_distributions():
    n = 200000
    dev = rp.Device(0, seed=5)
    mol = np.array([[0.5, 0.5, 0.5, 1, 0, 0, 0, 1, 0, 0, 0, 1.0]])
    dims = (100, 100, 20)
    tol = 5.0 / np.sqrt(n)

    def rotations():
        _, s = dev.download()
        r = s[:, 3:].reshape(n, 3, 3)
        assert np.abs(np.einsum("nij,nkj->nik", r, r) - np.eye(3)).max() < 1e-14      # orthogonal
        assert np.abs(np.linalg.det(r) - 1).max() < 1e-13
        return r

    # mode 0: uniform over SO(3) — every column is uniform on the sphere: zero mean, <x_i x_j> = delta_ij / 3
    dev.populate_lattice(np.eye(3) * 2.0, mol, dims)
    dev.randomize_rotations(1234, 0)
    r = rotations()
    for c in range(3):
        v = r[:, :, c]
        assert np.abs(v.mean(axis=0)).max() < tol
        assert np.abs(v.T @ v / n - np.eye(3) / 3).max() < tol
    # the rotation angle theta of a Haar-random rotation has density (1 - cos theta) / pi: <cos theta> = -1/2
    cos_theta = (np.trace(r, axis1=1, axis2=2) - 1) / 2
    assert abs(cos_theta.mean() + 0.5) < tol
    # a different seed gives different rotations, the same seed the same ones
    dev.populate_lattice(np.eye(3) * 2.0, mol, dims)
    dev.randomize_rotations(1234, 0)
    assert np.array_equal(rotations(), r)
    dev.populate_lattice(np.eye(3) * 2.0, mol, dims)
    dev.randomize_rotations(99, 0)
    assert not np.array_equal(rotations(), r)

    # mode 1: about a fixed axis by U(0, 2 pi): the axis is invariant, the angle uniform
    axis = np.array([1.0, 2.0, -0.5])
    dev.populate_lattice(np.eye(3) * 2.0, mol, dims)
    dev.randomize_rotations(7, 1, axis)
    r = rotations()
    a = axis / np.linalg.norm(axis)
    assert np.abs(r @ a - a).max() < 1e-14
    cos_theta = (np.trace(r, axis1=1, axis2=2) - 1) / 2
    assert abs(cos_theta.mean()) < tol and abs((cos_theta ** 2).mean() - 0.5) < tol

    # mode 2: about the particle's own body axis (start from random orientations: the body z axis must not move)
    dev.populate_lattice(np.eye(3) * 2.0, mol, dims)
    dev.randomize_rotations(11, 0)
    before = rotations()
    dev.randomize_rotations(12, 2, np.array([0.0, 0.0, 1.0]))
    after = rotations()
    assert np.abs(after[:, :, 2] - before[:, :, 2]).max() < 1e-13
    rel = np.einsum("nji,njk->nik", before, after)          # before^T after: a rotation about e_z
    assert abs(rel[:, 0, 0].mean()) < tol
    dev.close()


def test_populate_lattice_preconditions():
    dev = rp.Device(0, seed=1)
    mol = np.array([[0.5, 0.5, 0.5, 1, 0, 0, 0, 1, 0, 0, 0, 1.0]])
    with pytest.raises(rp.PreconditionError):
        dev.populate_lattice(np.eye(3), mol, (2, 2, 2), "xyz", 9)          # lattice.size() >= numOfShapes
    with pytest.raises(rp.PreconditionError):
        dev.populate_lattice(np.eye(3), mol, (2, 0, 2))
    outside = mol.copy()
    outside[0, 0] = 1.5
    with pytest.raises(rp.PreconditionError):
        dev.populate_lattice(np.eye(3), outside, (2, 2, 2))                 # areShapesWithinBox
    dev.close()
