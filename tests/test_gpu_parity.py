"""GPU parity tests: the CUDA path (through the C ABI) against the UNMODIFIED reference (oracle/_ref, strict flags).

Bar: overlap verdicts, overlap counts, cell numbers and +-inf move verdicts bit-exact; energies to 1e-10 relative
(the reference's own OpenMP reduction order is unspecified, src/core/Packing.cpp:697).
"""
import math

import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import helpers as H

pytestmark = pytest.mark.gpu

SHAPES = {
    "sphere": (H.desc_sphere, 1.1, 1),
    "spherocylinder": (H.desc_sc, 6.2, 1),
    "spherocylinder_short": (lambda: H.desc_sc(l=1.0, r=0.5), 2.2, 1),
    "kmer3": (lambda: H.desc_kmer(3), 1.1, 3),
    "polysc": (H.desc_polysc, 2.4, 2),
}


@pytest.mark.parametrize("name", list(SHAPES))
def test_pair_verdicts_bit_exact(device, name):
    make, max_dist, k = SHAPES[name]
    desc = make()
    rng = np.random.default_rng(42)
    pairs = H.random_pairs(rng, 400_000, max_dist, n_centres=k, translation_scale=20.0)
    ref = R.RefTraits(desc)
    want, _ = ref.pair_test_batch(pairs)
    device.set_interaction(desc)
    got, _ = device.pair_test_batch(pairs)
    assert 0.02 < want.mean() < 0.98          # the sample exercises both verdicts
    assert np.array_equal(got, want), f"{(got != want).sum()} verdicts differ"


@pytest.mark.parametrize("soft", [rp.RP_SOFT_LJ, rp.RP_SOFT_WCA, rp.RP_SOFT_SIC])
def test_pair_energies(device, soft):
    desc = H.desc_kmer(3, soft=soft, hard=False, epsilon=1.3, sigma=0.9)
    rng = np.random.default_rng(7)
    pairs = H.random_pairs(rng, 200_000, 3.0, n_centres=3, translation_scale=5.0)
    ref = R.RefTraits(desc)
    _, want = ref.pair_test_batch(pairs, want_verdict=False, want_energy=True)
    device.set_interaction(desc)
    _, got = device.pair_test_batch(pairs, want_verdict=False, want_energy=True)
    assert np.array_equal(got, want) or np.allclose(got, want, rtol=1e-14, atol=0)


def triclinic_box(lx, ly, lz):
    return np.array([[lx, 0.2 * ly, 0.1 * lz], [0, ly, -0.15 * lz], [0, 0, lz]])


@pytest.mark.parametrize("n,box", [(3000, np.diag([30.0, 40.0, 50.0])), (2000, triclinic_box(28, 30, 45)),
                                   (60, np.diag([9.0, 9.5, 10.0]))])
def test_cells_and_overlap_count_spherocylinders(device, n, box):
    rng = np.random.default_rng(n)
    shapes = H.random_packing(rng, n, box)
    desc = H.desc_sc(l=3.0, r=0.5)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    gi = device.grid_info()
    # Packing.cpp:868-881: no grid when 4 cells do not fit the largest height (the 60-particle case)
    assert bool(gi.has_grid) == (n != 60)
    if gi.has_grid:
        assert tuple(gi.cell_divisions) == ref.grid_divisions()
        want_cells, _ = R.ref_cell_numbers(box, gi.cell_size, shapes[:, :3])
        assert np.array_equal(device.cell_numbers(), want_cells)
    want = ref.count_total_overlaps(False)
    assert want > 0
    assert device.count_total_overlaps(False) == want
    assert device.count_total_overlaps(True) == 1 == ref.count_total_overlaps(True)


@pytest.mark.parametrize("name", ["sphere", "kmer3", "polysc"])
def test_overlap_count_other_shapes(device, name):
    make, _, _ = SHAPES[name]
    desc = make()
    rng = np.random.default_rng(5)
    box = triclinic_box(14, 15, 16)
    shapes = H.random_packing(rng, 1500, box)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    want = ref.count_total_overlaps(False)
    assert want > 0
    assert device.count_total_overlaps(False) == want


@pytest.mark.parametrize("soft,n,box", [(rp.RP_SOFT_LJ, 800, np.diag([14.0, 15.0, 16.0])),
                                        (rp.RP_SOFT_WCA, 1500, triclinic_box(14, 15, 16)),
                                        (rp.RP_SOFT_SIC, 40, np.diag([5.0, 5.0, 5.0]))])
def test_total_energy(device, soft, n, box):
    desc = H.desc_kmer(3, soft=soft, hard=False)
    rng = np.random.default_rng(11)
    shapes = H.random_packing(rng, n, box)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    want = ref.total_energy()
    got = device.total_energy()
    assert math.isfinite(want) and want != 0
    assert got == pytest.approx(want, rel=1e-10)


def run_move_sequence(device, ref, rng, n, steps, trans, counting=False):
    for step in range(steps):
        i = int(rng.integers(0, n))
        kind = int(rng.integers(0, 3))
        t = rng.uniform(-trans, trans, size=3)
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        rot = R.ref_rotation_matrix(axis, rng.uniform(-0.5, 0.5))
        want = ref.try_move(i, kind, t, rot)
        got = device.try_move(i, kind, t, rot)
        if math.isinf(want):
            assert got == want, f"step {step}: {got} != {want}"
        else:
            assert got == pytest.approx(want, rel=1e-10, abs=1e-12), f"step {step}"
        accept = (want <= 0) if not counting else (want < math.inf and rng.uniform() < 0.7)
        if accept:
            ref.accept_move()
            device.accept_move()
        if counting:
            assert device.cached_overlaps() == ref.cached_overlaps()


def test_try_move_sequence_hard(device):
    """Explicit moves (Packing::tryMove / acceptMove): +inf verdicts must agree move by move."""
    box, shapes = H.sc_lattice((8, 8, 3))
    desc = H.desc_sc()
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    rng = np.random.default_rng(3)
    run_move_sequence(device, ref, rng, shapes.shape[0], 300, 0.3)
    _, got = device.download()
    _, want = ref.get()
    assert np.array_equal(got, want)
    assert device.count_total_overlaps(False) == 0


def test_try_move_sequence_soft_and_counting(device):
    desc = H.desc_kmer(3, soft=rp.RP_SOFT_WCA)         # hard beads + WCA: a CompoundInteraction
    desc_ref = desc
    rng = np.random.default_rng(9)
    box = np.diag([12.0, 12.0, 13.0])
    shapes = H.random_packing(rng, 300, box)
    device.set_interaction(desc)
    device.upload(box, shapes)
    device.toggle_overlap_counting(True)
    try:
        ref_traits = R.RefTraits(desc_ref)
    except ValueError:
        pytest.skip("the shim builds no compound hard+soft polysphere")
    ref = R.RefPacking(ref_traits, box, shapes)
    ref.toggle_overlap_counting(True)
    assert device.cached_overlaps() == ref.cached_overlaps()
    run_move_sequence(device, ref, rng, 300, 150, 0.5, counting=True)
    device.toggle_overlap_counting(False)


def test_try_scaling(device):
    box, shapes = H.sc_lattice((8, 8, 3))
    desc = H.desc_sc()
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    for factor in (1.05, 0.97, 0.80):
        new_box = box * factor
        want = ref.try_scaling(new_box)
        got = device.try_scaling(new_box)
        assert got == want
        b, s = device.download()
        rb, rs = ref.get()
        assert np.array_equal(b, rb) and np.array_equal(s, rs)
        ref.revert_scaling()
        device.revert_scaling()
        b, s = device.download()
        assert np.array_equal(b, box) and np.array_equal(s, shapes)
    # a sheared box (triclinic move), accepted
    new_box = box.copy()
    new_box[0, 1] = 0.7
    new_box *= 1.02
    assert device.try_scaling(new_box) == ref.try_scaling(new_box) == 0
    _, s = device.download()
    _, rs = ref.get()
    assert np.array_equal(s, rs)
    assert device.count_total_overlaps(False) == 0


def test_try_scaling_soft(device):
    desc = H.desc_kmer(2, soft=rp.RP_SOFT_LJ, hard=False)
    rng = np.random.default_rng(21)
    box = np.diag([13.0, 14.0, 15.0])
    shapes = H.random_packing(rng, 500, box)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    for factor in (1.03, 0.96):
        want = ref.try_scaling(box * factor)
        got = device.try_scaling(box * factor)
        assert got == pytest.approx(want, rel=1e-9)
        ref.revert_scaling()
        device.revert_scaling()


@pytest.mark.parametrize("kind", [rp.RP_MOVE_TRANSLATION, rp.RP_MOVE_ROTATION, rp.RP_MOVE_ROTOTRANSLATION])
def test_sweep_spherocylinders_keeps_packing_valid(device, kind):
    """A bulk sweep must never create an overlap; verified with the reference's own countTotalOverlaps."""
    box, shapes = H.sc_lattice((20, 20, 6))
    n = shapes.shape[0]
    desc = H.desc_sc()
    device.set_interaction(desc)
    device.upload(box, shapes)
    accepted = 0
    for cycle in range(10):
        c = device.sweep(kind, 0.3, 0.2, cycle)
        assert c.attempted == n
        assert c.accepted + c.rejected_domain + c.rejected_overlap == n
        accepted += c.accepted
    assert 0.05 * 10 * n < accepted < 0.999 * 10 * n
    assert device.count_total_overlaps(False) == 0
    b, s = device.download()
    moved = np.any(s != shapes, axis=1).sum()
    assert moved > 0.3 * n
    # inside the box, orientation still orthonormal
    rel = np.linalg.solve(b, s[:, :3].T).T
    assert (rel >= 0).all() and (rel < 1).all()
    o = s[:, 3:].reshape(n, 3, 3)
    assert np.allclose(np.einsum("nij,nkj->nik", o, o), np.eye(3), atol=1e-12)
    ref = R.RefPacking(R.RefTraits(desc), b, s)
    assert ref.count_total_overlaps(False) == 0


def test_sweep_spheres(device):
    box, shapes = H.sphere_lattice((24, 24, 24), 1.3)
    n = shapes.shape[0]
    desc = H.desc_sphere()
    device.set_interaction(desc)
    device.upload(box, shapes)
    for cycle in range(10):
        c = device.sweep(rp.RP_MOVE_TRANSLATION, 0.2, 0.0, cycle)
        assert c.attempted == n
    assert device.count_total_overlaps(False) == 0
    b, s = device.download()
    ref = R.RefPacking(R.RefTraits(desc), b, s)
    assert ref.count_total_overlaps(False) == 0
    assert np.any(s != shapes, axis=1).sum() > 0.5 * n


def test_sweep_is_deterministic(device):
    box, shapes = H.sc_lattice((12, 12, 4))
    desc = H.desc_sc()
    results = []
    for _ in range(2):
        device.set_interaction(desc)
        device.upload(box, shapes)
        for cycle in range(3):
            device.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.2, 0.1, cycle)
        results.append(device.download()[1].copy())
    assert np.array_equal(results[0], results[1])


def test_sweep_fp32_cull_never_changes_a_verdict(device, monkeypatch):
    """The conservative FP32 cull of the sweep kernel must be invisible: with the cull disabled (every pair takes the
    exact FP64 test) the trajectories are bit-identical."""
    tric = np.array([[16 * 1.2, 3.0, 1.0], [0, 16 * 1.2, -2.0], [0, 0, 5 * 6.5]])
    for box, shapes in (H.sc_lattice((16, 16, 5)), (tric, None)):
        if shapes is None:
            _, lat = H.sc_lattice((16, 16, 5))
            rel = lat[:, :3] / np.array([16 * 1.2, 16 * 1.2, 5 * 6.5])
            shapes = lat.copy()
            shapes[:, :3] = rel @ box.T
        results = []
        for no_cull in (False, True):
            if no_cull:
                monkeypatch.setenv("RP_SWEEP_NO_CULL", "1")
            else:
                monkeypatch.delenv("RP_SWEEP_NO_CULL", raising=False)
            device.set_interaction(H.desc_sc())
            device.upload(box, shapes)
            acc = 0
            for cycle in range(6):
                acc += device.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.25, 0.15, cycle).accepted
            results.append((acc, device.download()[1].copy()))
        monkeypatch.delenv("RP_SWEEP_NO_CULL", raising=False)
        assert results[0][0] == results[1][0] > 0
        assert np.array_equal(results[0][1], results[1][1])
        assert device.count_total_overlaps(False) == 0
