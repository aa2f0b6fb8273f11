"""GPU parity of the bulk sweep against the UNMODIFIED reference, move by move.

rp_sweep is a different schedule of the same Markov chain as the reference's Simulation::performMoves (other RNG,
checkerboard cells instead of OpenMP domains), so trajectories cannot be compared directly.  What can: rp_sweep_trace
reports every trial of a sweep (particle, trial state, processing order, decision, energy difference); replaying the
trials in that order through the reference's own Packing::tryMove / acceptMove must reproduce
  * every accept / reject decision (hard: dE == +inf <=> rejected; counting mode: -inf / +inf / 0),
  * every energy difference to 1e-9 relative (soft potentials; Metropolis decisions follow from dE and the traced u),
  * the final configuration (to rounding: the device forms the trial in the image frame of the sweep grid, the
    replay hands the reference a translation vector / rotation matrix recovered from the traced trial state).
"""
import math

import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import helpers as H

pytestmark = pytest.mark.gpu


def replay_sweep(ref, trace, before, kind, temperature, box):
    """Feeds the traced trials to the reference in processing order.  Returns (trials replayed, mismatches)."""
    order = np.lexsort((trace["slot"], trace["order_key"], trace["batch"], trace["cell"], trace["colour_step"]))
    mismatches = []
    replayed = 0
    max_de_err = 0.0
    for idx in order:
        t = trace[idx]
        if t["status"] == 2:
            continue            # left its active cell: Packing.cpp:95-96 returns +inf before any pair work
        if t["status"] == 3:
            continue            # RP_MOVE_FLIP: the particle was not drawn this cycle
        i = int(t["particle"])
        old_pos, old_orient = before[i, :3], before[i, 3:].reshape(3, 3)
        translation = (t["position"] - t["image_shift"]) - old_pos if kind != rp.RP_MOVE_ROTATION else np.zeros(3)
        rotation = t["orientation"].reshape(3, 3) @ old_orient.T if kind != rp.RP_MOVE_TRANSLATION else np.eye(3)
        want = ref.try_move(i, kind, translation, rotation)
        got = float(t["dE"])
        replayed += 1
        if math.isinf(want) or math.isinf(got):
            if want != got:
                mismatches.append((i, got, want))
                continue
        else:
            err = abs(got - want) / max(1.0, abs(want))
            max_de_err = max(max_de_err, err)
            if err > 1e-9:
                mismatches.append((i, got, want))
                continue
        # the decision must follow from dE and u exactly like Simulation.cpp:483
        arg = -got / temperature
        expect_accept = True if arg > 700 else (False if arg < -700 else t["u"] <= math.exp(arg))
        if bool(t["status"] == 1) != expect_accept:
            mismatches.append((i, "decision", got, float(t["u"])))
        if t["status"] == 1:
            ref.accept_move()
    return replayed, mismatches, max_de_err


def run_replay(device, desc, box, shapes, kind, tstep, rstep, cycles, temperature=1.0, counting=False):
    device.toggle_overlap_counting(False)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    if counting:
        device.toggle_overlap_counting(True)
        ref.toggle_overlap_counting(True)
        assert device.cached_overlaps() == ref.cached_overlaps() > 0
    total, accepted = 0, 0
    for cycle in range(cycles):
        _, before = device.download()
        before = before.copy()
        c = device.sweep(kind, tstep, rstep, cycle, temperature=temperature)
        trace = device.sweep_trace()
        assert (trace["status"] == 1).sum() == c.accepted
        assert (trace["status"] == 2).sum() == c.rejected_domain
        replayed, mismatches, err = replay_sweep(ref, trace, before, kind, temperature, box)
        assert not mismatches, f"cycle {cycle}: {len(mismatches)} of {replayed} trials differ: {mismatches[:5]}"
        total += replayed
        accepted += int(c.accepted)
        _, got = device.download()
        _, want = ref.get()
        assert np.allclose(got, want, rtol=0, atol=1e-10)
        if counting:
            assert device.cached_overlaps() == ref.cached_overlaps()
    assert 0.02 * total < accepted < 0.995 * total      # both verdicts exercised
    if counting:
        device.toggle_overlap_counting(False)
    return total


def test_replay_spherocylinders_fast_path(device):
    box, shapes = H.sc_lattice((12, 12, 4))
    run_replay(device, H.desc_sc(), box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.2, 4)


def test_replay_spherocylinders_of_length_zero(device):
    """SpherocylinderTraits admits length == 0 (SpherocylinderTraits.cpp:15): the FP32 culls must not turn the missing
    axis into NaN distances (a NaN compares false and the pair would be dropped)."""
    box, shapes = H.sc_lattice((10, 10, 10), cell=(1.15, 1.15, 1.15))
    run_replay(device, H.desc_sc(l=0.0), box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.2, 3)
    assert device.count_total_overlaps(False) == 0


def test_replay_spherocylinders_generic_kernel(device, monkeypatch):
    monkeypatch.setenv("RP_SWEEP_GENERIC", "1")
    box, shapes = H.sc_lattice((12, 12, 4))
    run_replay(device, H.desc_sc(), box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.2, 3)


def test_generic_kernel_reproduces_fast_path_bit_for_bit(device, monkeypatch):
    """Same Philox streams, same processing order, same arithmetic: the two kernels must produce identical
    trajectories (cells hold < 64 particles here; above that the fast path resolves a cell in batches of 64, which
    is a different — equally valid — processing order)."""
    box, shapes = H.sc_lattice((24, 24, 8))
    results = []
    for generic in (False, True):
        if generic:
            monkeypatch.setenv("RP_SWEEP_GENERIC", "1")
        else:
            monkeypatch.delenv("RP_SWEEP_GENERIC", raising=False)
        device.set_interaction(H.desc_sc())
        device.upload(box, shapes)
        acc = 0
        for cycle in range(5):
            acc += device.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.25, 0.15, cycle).accepted
        results.append((acc, device.download()[1].copy()))
    assert results[0][0] == results[1][0] > 0
    assert np.array_equal(results[0][1], results[1][1])


def kmer_lattice(n_cells, k, spacing, rng):
    """k-mers (beads r = 0.5, distance 1) on a cubic lattice with random orientations and a small jitter."""
    nx, ny, nz = n_cells
    ix, iy, iz = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    pos = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1) * spacing + 0.5 * spacing
    box = np.diag([nx * spacing, ny * spacing, nz * spacing]).astype(np.float64)
    rot = H.random_rotations(rng, pos.shape[0])
    return box, H.shapes_from(pos, rot)


def test_replay_hard_kmers(device):
    rng = np.random.default_rng(17)
    box, shapes = kmer_lattice((9, 9, 9), 3, 3.05, rng)
    desc = H.desc_kmer(3)
    assert R.RefPacking(R.RefTraits(desc), box, shapes).count_total_overlaps(False) == 0
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.6, 0.5, 3)


def test_replay_hard_polyspherocylinders_triclinic(device):
    rng = np.random.default_rng(23)
    box, shapes = kmer_lattice((8, 8, 8), 2, 4.3, rng)
    tric = box.copy()
    tric[0, 1], tric[0, 2], tric[1, 2] = 2.0, -1.5, 1.0
    rel = shapes[:, :3] / np.diag(box)
    shapes[:, :3] = rel @ tric.T
    desc = H.desc_polysc()
    assert R.RefPacking(R.RefTraits(desc), tric, shapes).count_total_overlaps(False) == 0
    run_replay(device, desc, tric, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.8, 0.6, 3)


@pytest.mark.parametrize("kind", [rp.RP_MOVE_TRANSLATION, rp.RP_MOVE_ROTATION])
def test_replay_wca_kmers_single_move_kinds(device, kind):
    rng = np.random.default_rng(31)
    box, shapes = kmer_lattice((8, 8, 8), 3, 2.2, rng)
    desc = H.desc_kmer(3, soft=rp.RP_SOFT_WCA, hard=False)
    run_replay(device, desc, box, shapes, kind, 0.3, 0.4, 2, temperature=1.3)


def test_replay_wca_kmers(device):
    rng = np.random.default_rng(37)
    box, shapes = kmer_lattice((8, 8, 8), 3, 2.2, rng)
    desc = H.desc_kmer(3, soft=rp.RP_SOFT_WCA, hard=False)
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.4, 3, temperature=0.8)


def test_replay_lennard_jones_uses_the_reference_pair_set(device):
    """LJ has no cutoff in the reference (LennardJonesInteraction.cpp:8-13): the energy difference depends on which
    centres share neighbouring grid cells.  Agreement of dE to 1e-9 proves the sweep sums over the same pair set."""
    rng = np.random.default_rng(41)
    box, shapes = kmer_lattice((12, 12, 12), 2, 1.9, rng)
    desc = H.desc_kmer(2, soft=rp.RP_SOFT_LJ, hard=False)
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.15, 0.2, 2, temperature=1.0)


def test_replay_compound_hard_plus_wca(device):
    rng = np.random.default_rng(43)
    box, shapes = kmer_lattice((8, 8, 8), 3, 3.1, rng)
    desc = H.desc_kmer(3, soft=rp.RP_SOFT_WCA, sigma=1.4)
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.4, 0.4, 3, temperature=1.0)


def test_replay_overlap_counting_relaxation(device):
    """Overlap relaxation (CasinoMode.cpp:288-289): counting mode with a hard k-mer + square-inverse-core helper;
    moves that reduce the overlap count are always accepted, numOverlaps stays in step with the reference."""
    rng = np.random.default_rng(47)
    box, shapes = kmer_lattice((8, 8, 8), 3, 2.0, rng)          # compressed: many overlaps
    desc = H.desc_kmer(3, soft=rp.RP_SOFT_SIC, epsilon=2.0, sigma=1.0)
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.3, 3, temperature=1.0, counting=True)


def test_replay_flip_sampler(device):
    """FlipSampler (src/core/move_samplers/FlipSampler.cpp:18-41): a half-turn about the flip axis of the shape plus
    the translation that keeps the geometric origin in place; every particle attempts it with probability
    1 / flip_every.  The traced trials are replayed through the reference like any rototranslation."""
    desc = H.desc_kmer(3)
    box = np.diag([15.0, 15.0, 15.0])
    rng = np.random.default_rng(12)
    shapes = H.random_packing(rng, 450, box)
    device.toggle_overlap_counting(False)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    device.toggle_overlap_counting(True)        # random start: overlaps are counted, not forbidden
    ref.toggle_overlap_counting(True)
    axis, origin, every = np.array([1.0, 0.0, 0.0]), np.array([0.0, 0.0, 0.4]), 3
    attempted = accepted = 0
    for cycle in range(4):
        _, before = device.download()
        before = before.copy()
        c = device.sweep(rp.RP_MOVE_FLIP, 0.0, 0.0, cycle, flip_axis=axis, flip_origin=origin, flip_every=every)
        trace = device.sweep_trace()
        drawn = trace[trace["status"] != 3]
        assert c.attempted == drawn.shape[0]
        for t in drawn[:60]:
            i = int(t["particle"])
            o_old, o_new = before[i, 3:].reshape(3, 3), t["orientation"].reshape(3, 3)
            rot = o_new @ o_old.T
            world_axis = o_old @ axis
            assert np.allclose(rot @ world_axis, world_axis, atol=1e-12) and abs(np.trace(rot) + 1) < 1e-12
            new_pos = t["position"] - t["image_shift"]
            assert np.allclose(new_pos + o_new @ origin, before[i, :3] + o_old @ origin, atol=1e-12)
        replayed, mismatches, _ = replay_sweep(ref, trace, before, rp.RP_MOVE_ROTOTRANSLATION, 1.0, box)
        assert not mismatches, mismatches[:5]
        attempted += int(c.attempted)
        accepted += int(c.accepted)
        assert device.cached_overlaps() == ref.cached_overlaps()
    assert 0.25 * 4 * 450 < attempted < 0.42 * 4 * 450      # about a third of the particles is drawn per cycle
    assert accepted > 0
    device.toggle_overlap_counting(False)
