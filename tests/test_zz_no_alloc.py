"""rp_nematic_order_static — the entry point that allocates nothing (its partial sums live in a file-scope __device__
array, rampack_b200/csrc/rp_noalloc.cu) — against rp_nematic_order, through the C ABI, on calls in a row that share
the array.  It has no counterpart in the reference; it exists for callers that cannot call cudaMalloc.

Tolerance: 1e-12 on Q and P2, the tolerance tests/test_gpu_observables.py uses for the order parameters (the two entry
points add the same terms in a different, fixed order).

The file name sorts last on purpose: a failure here hides no other test.  It
builds its own handles, so it passes alone in a fresh process as well as after the other tests."""
import numpy as np
import pytest

import rampack_b200 as rp

pytestmark = pytest.mark.gpu

MAX_PARTICLES = 4194304          # RP_NEMATIC_STATIC_MAX_PARTICLES (include/rampack_b200.h)
TOL = 1e-12


def tilted_rods(rng, n, spread, director):
    """n x 12 rows: random positions in a 50^3 box, body z axis tilted from `director` by ~N(0, spread) rad."""
    director = np.asarray(director, dtype=np.float64) / np.linalg.norm(director)
    helper = np.array([1.0, 0.0, 0.0]) if abs(director[0]) < 0.9 else np.array([0.0, 1.0, 0.0])
    e1 = np.cross(director, helper)
    e1 /= np.linalg.norm(e1)
    e2 = np.cross(director, e1)
    theta = rng.normal(0.0, spread, n)
    phi = rng.uniform(0.0, 2 * np.pi, n)
    z = (np.cos(theta)[:, None] * director + np.sin(theta)[:, None]
         * (np.cos(phi)[:, None] * e1 + np.sin(phi)[:, None] * e2))
    x = np.cross(z, director + 0.123 * e1)
    x /= np.linalg.norm(x, axis=1)[:, None]
    y = np.cross(z, x)
    rot = np.stack([x, y, z], axis=2)                       # columns = body axes
    shapes = np.empty((n, 12))
    shapes[:, :3] = rng.uniform(0.0, 50.0, (n, 3))
    shapes[:, 3:] = rot.reshape(n, 9)
    return np.diag([50.0, 50.0, 50.0]), shapes


def q_numpy(shapes, axis):
    a = shapes[:, 3:].reshape(-1, 3, 3) @ np.asarray(axis, dtype=np.float64)
    return 1.5 * (a[:, :, None] * a[:, None, :]).mean(axis=0) - 0.5 * np.eye(3)


def check_against_allocating_entry_point(dev, shapes, axis):
    q_want, p2_want = dev.nematic_order(axis)
    q_got, p2_got = dev.nematic_order_static(axis)
    assert np.allclose(q_got, q_want, rtol=0, atol=TOL)
    assert p2_got == pytest.approx(p2_want, abs=TOL)
    assert np.allclose(q_got, q_numpy(shapes, axis), rtol=0, atol=TOL)
    return p2_got


def test_static_scratch_matches_allocating_entry_point_on_calls_in_a_row():
    import torch
    rng = np.random.default_rng(20261105)
    dev, other = rp.Device(0, seed=5), rp.Device(0, seed=6)
    try:
        # call 1: 300 001 particles = 74 blocks of the static kernel, the last one ragged
        box, big = tilted_rods(rng, 300001, 0.35, (0.0, 0.0, 1.0))
        dev.upload(box, big)
        p2_big = check_against_allocating_entry_point(dev, big, (0.0, 0.0, 1.0))
        assert 0.5 < p2_big < 0.9
        # call 2, same handle, a SMALLER and different packing: rows 2..73 of the shared array still hold call 1's
        # partial sums and must not leak into the result
        box, small = tilted_rods(rng, 5000, 0.8, (1.0, 1.0, 0.0))
        dev.upload(box, small)
        p2_small = check_against_allocating_entry_point(dev, small, (0.0, 0.0, 1.0))
        assert abs(p2_small - p2_big) > 0.1
        # call 3: a second handle on the same device shares the array (calls are serialised here)
        box, third = tilted_rods(rng, 70000, 0.2, (0.0, 1.0, 0.0))
        other.upload(box, third)
        check_against_allocating_entry_point(other, third, (0.0, 0.0, 1.0))
        # and back on the first handle: identical to its previous answer, bit for bit (fixed summation order), with
        # no device memory taken by the call
        free_before = torch.cuda.mem_get_info(0)[0]
        q_a, p2_a = dev.nematic_order_static((0.0, 0.0, 1.0))
        q_b, p2_b = dev.nematic_order_static((0.0, 0.0, 1.0))
        free_after = torch.cuda.mem_get_info(0)[0]
        assert np.array_equal(q_a, q_b) and p2_a == p2_b == p2_small
        assert free_after >= free_before
    finally:
        dev.close()
        other.close()


def test_static_scratch_limit_is_enforced():
    # two differently tilted rods per cell: 128 x 128 x 128 cells = exactly the limit, one more layer is beyond it
    cell = np.diag([2.0, 2.0, 7.0])
    _, mol = tilted_rods(np.random.default_rng(3), 2, 0.4, (0.0, 0.0, 1.0))
    mol[:, :3] = [[0.25, 0.25, 0.5], [0.75, 0.75, 0.5]]
    dev = rp.Device(0, seed=7)
    try:
        dev.populate_lattice(cell, mol, (128, 128, 128))
        assert dev.n == MAX_PARTICLES
        q_want, p2_want = dev.nematic_order()
        q_got, p2_got = dev.nematic_order_static()
        assert np.allclose(q_got, q_want, rtol=0, atol=TOL) and p2_got == pytest.approx(p2_want, abs=TOL)
        assert np.allclose(q_got, q_numpy(mol, (0.0, 0.0, 1.0)), rtol=0, atol=TOL)
        dev.populate_lattice(cell, mol, (128, 128, 129))
        with pytest.raises(rp.PreconditionError) as e:
            dev.nematic_order_static()
        assert "RP_NEMATIC_STATIC_MAX_PARTICLES" in e.value.message and str(MAX_PARTICLES) in e.value.message
        dev.nematic_order()                                  # the allocating entry point still takes it
    finally:
        dev.close()
