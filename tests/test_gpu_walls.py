"""GPU parity of hard walls (Packing::toggleWalls, countParticleWallOverlaps, Interaction::overlapWithWall —
src/core/Packing.cpp:561-580, 1075-1142 and the shapes' overlapWithWall) against the UNMODIFIED reference:
overlap counts, explicit moves (hard and counting mode) and sweep replay with walls on."""
import zlib

import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import helpers as H
import xc_cases
from test_gpu_parity import run_move_sequence, triclinic_box
from test_gpu_sweep_replay import replay_sweep

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not R.have_ref(), reason="walls are checked against oracle/_ref")]


@pytest.fixture()
def walled(device):
    yield device
    device.toggle_walls([0, 0, 0])
    device.toggle_overlap_counting(False)


SHAPES = {
    "sphere": H.desc_sphere,
    "spherocylinder": H.desc_sc,
    "kmer3": lambda: H.desc_kmer(3),
    "polysc": H.desc_polysc,
    # XenoCollide: the wall test is a support-point query (XenoCollideTraits.h:170-186)
    "xc_wrap_sphere_cuboid": lambda: xc_cases.CASES["wrap_sphere_cuboid"][0]()[0],
    "xc_polyhedral_wedge": lambda: xc_cases.CASES["polyhedral_wedge"][0]()[0],
    "xc_smooth_wedge_2centres": lambda: xc_cases.CASES["smooth_wedge_2centres"][0]()[0],
}


@pytest.mark.parametrize("walls", [(1, 0, 0), (0, 1, 1), (1, 1, 1)])
@pytest.mark.parametrize("name", list(SHAPES))
@pytest.mark.parametrize("box", [np.diag([14.0, 15.0, 16.0]), triclinic_box(14, 15, 16)], ids=["ortho", "triclinic"])
def test_overlap_count_with_walls(walled, name, walls, box):
    desc = SHAPES[name]()
    rng = np.random.default_rng(zlib.crc32(f"{name}{walls}".encode()))
    shapes = H.random_packing(rng, 400, box)
    walled.set_interaction(desc)
    walled.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    base = ref.count_total_overlaps(False)
    assert walled.count_total_overlaps(False) == base
    walled.toggle_walls(walls)
    ref.toggle_walls(walls)
    want = ref.count_total_overlaps(False)
    assert want > base                          # the sample has particles sticking through the walls
    assert walled.count_total_overlaps(False) == want
    assert walled.count_total_overlaps(True) == ref.count_total_overlaps(True)


def test_moves_against_walls_hard(walled):
    box, shapes = H.sc_lattice((8, 8, 3))
    desc = H.desc_sc()
    walled.set_interaction(desc)
    walled.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    walled.toggle_walls((1, 1, 1))
    ref.toggle_walls((1, 1, 1))
    assert walled.count_total_overlaps(False) == ref.count_total_overlaps(False) == 0
    run_move_sequence(walled, ref, np.random.default_rng(5), shapes.shape[0], 400, 0.4)
    _, got = walled.download()
    _, want = ref.get()
    assert np.array_equal(got, want)
    assert walled.count_total_overlaps(False) == 0


def test_moves_against_walls_counting_mode(walled):
    desc = H.desc_kmer(3)
    box = np.diag([9.0, 9.0, 9.0])
    rng = np.random.default_rng(8)
    shapes = H.random_packing(rng, 150, box)
    walled.set_interaction(desc)
    walled.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    walled.toggle_walls((1, 0, 1))
    ref.toggle_walls((1, 0, 1))
    walled.toggle_overlap_counting(True)
    ref.toggle_overlap_counting(True)
    assert walled.cached_overlaps() == ref.cached_overlaps() > 0
    run_move_sequence(walled, ref, rng, shapes.shape[0], 300, 0.5, counting=True)
    assert walled.cached_overlaps() == ref.cached_overlaps()


def sweep_with_walls(device, desc, box, shapes, walls, kind, tstep, rstep, cycles, counting=False):
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    device.toggle_walls(walls)
    ref.toggle_walls(walls)
    if counting:
        device.toggle_overlap_counting(True)
        ref.toggle_overlap_counting(True)
        assert device.cached_overlaps() == ref.cached_overlaps() > 0
    total = accepted = 0
    for cycle in range(cycles):
        _, before = device.download()
        before = before.copy()
        c = device.sweep(kind, tstep, rstep, cycle)
        trace = device.sweep_trace()
        replayed, mismatches, _ = replay_sweep(ref, trace, before, kind, 1.0, box)
        assert not mismatches, f"cycle {cycle}: {len(mismatches)} of {replayed} trials differ: {mismatches[:5]}"
        total += replayed
        accepted += int(c.accepted)
        _, got = device.download()
        _, want = ref.get()
        assert np.allclose(got, want, rtol=0, atol=1e-10)
        assert device.count_total_overlaps(False) == ref.count_total_overlaps(False)
        if counting:
            assert device.cached_overlaps() == ref.cached_overlaps()
    assert 0.02 * total < accepted < 0.995 * total
    return total


def test_sweep_replay_spherocylinders_between_walls(walled):
    box, shapes = H.sc_lattice((20, 20, 4))
    n = sweep_with_walls(walled, H.desc_sc(), box, shapes, (1, 1, 1), rp.RP_MOVE_ROTOTRANSLATION, 0.3, 0.2, 4)
    assert n > 4000
    # the walls did reject something: particles of the outer layers cannot leave the box
    _, s = walled.download()
    assert s[:, :3].min() > 0.5 - 1e-12


def test_sweep_replay_kmers_between_walls_counting(walled):
    box = np.diag([16.0, 16.0, 16.0])
    rng = np.random.default_rng(4)
    shapes = H.random_packing(rng, 700, box)
    sweep_with_walls(walled, H.desc_kmer(3), box, shapes, (0, 1, 1), rp.RP_MOVE_ROTOTRANSLATION, 0.4, 0.3, 3,
                     counting=True)
