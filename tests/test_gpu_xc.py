"""GPU parity of the XenoCollide path (generic convex shapes, smooth / polyhedral wedges) against the reference's own
geometry classes and XenoCollide<>::Intersect, fed the SAME rp_xc_program (oracle/ref_harness.cpp builds the
reference objects from it with their public constructors).  Bar: verdicts and overlap counts bit-exact."""
import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import helpers as H
import xc_cases
from test_gpu_sweep_replay import run_replay, kmer_lattice

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(xc_cases.CASES))
def test_xc_pair_verdicts_bit_exact(device, name):
    make, max_dist = xc_cases.CASES[name]
    desc, k = make()
    rng = np.random.default_rng(101)
    pairs = H.random_pairs(rng, 300_000, max_dist, n_centres=k, translation_scale=15.0)
    want, _ = R.RefTraits(desc).pair_test_batch(pairs)
    device.set_interaction(desc)
    got, _ = device.pair_test_batch(pairs)
    assert 0.05 < want.mean() < 0.95, f"sample does not exercise both verdicts: {want.mean():.3f}"
    diff = np.flatnonzero(got != want)
    assert diff.size == 0, f"{diff.size} of {pairs.shape[0]} verdicts differ, first at {diff[:5]}"


@pytest.mark.parametrize("name", ["wrap_sphere_cuboid", "ellipsoid", "sum_disk_segment", "smooth_wedge_2centres"])
def test_xc_overlap_count(device, name):
    make, _ = xc_cases.CASES[name]
    desc, _ = make()
    rng = np.random.default_rng(7)
    box = np.array([[14.0, 2.8, 1.4], [0, 15.0, -2.2], [0, 0, 16.0]])
    shapes = H.random_packing(rng, 1200, box)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    want = ref.count_total_overlaps(False)
    assert want > 0
    assert device.count_total_overlaps(False) == want
    assert device.count_total_overlaps(True) == 1


@pytest.mark.parametrize("name,spacing", [("wrap_sphere_cuboid", 2.8), ("smooth_wedge_2centres", 2.9)])
def test_xc_sweep_replay(device, name, spacing):
    make, _ = xc_cases.CASES[name]
    desc, _ = make()
    rng = np.random.default_rng(59)
    box, shapes = kmer_lattice((9, 9, 9), 1, spacing, rng)
    assert R.RefPacking(R.RefTraits(desc), box, shapes).count_total_overlaps(False) == 0
    run_replay(device, desc, box, shapes, rp.RP_MOVE_ROTOTRANSLATION, 0.5, 0.5, 3)
