"""north_star's pair check inside `-m gpu`: >= 10^7 spherocylinder pair tests drawn from
snapshots of the N = 100 352 packing (BASELINE config 2) that the REFERENCE's own Simulation advanced, device verdicts
(rp_pair_test_batch through the C ABI) against the reference's overlapBetween — bit-exact, with the pairs within 1e-12
relative of contact listed.  The full 10^8-pair run is tools/pairs_1e8.py (profiles/r01_pairs_1e8.json); this is the
same code on fewer snapshots so that the driver's GPU run sees it."""
import os
import sys

import numpy as np
import pytest

import rampack_b200 as rp
from rampack_b200 import workloads as W
from oracle import checker

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


def test_ten_million_snapshot_pairs_bit_exact():
    if checker.kind() != "reference":
        pytest.skip("needs oracle/_ref (the compiled reference)")
    import pairs_1e8 as P
    from oracle import refwrap as R
    threads = min(16, len(os.sched_getaffinity(0)))
    desc = rp.make_desc(hard=rp.RP_HARD_SPHEROCYLINDER, radius=W.SC_RADIUS, length=W.SC_LENGTH)
    strict = R.RefTraits(desc)
    native = R.RefTraits(desc, native=True) if R.have_ref(native=True) else None
    dev = rp.Device(0, seed=2024)
    dev.set_interaction(desc)
    box, shapes = W.sc_config("sc_100k")
    dev.upload(box, shapes)
    for c in range(150):                                   # melt the lattice on the device
        dev.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.2, 0.15, c)
    box, shapes = dev.download()
    rng = np.random.default_rng(77)
    contact2 = (2 * W.SC_RADIUS) ** 2
    total = mismatches = overlapping = near_contact = 0
    snapshots = 0
    while total < 1.0e7:
        # the next snapshot is written by the reference's own Simulation (one NVT cycle on the host)
        dom = (1, 1, max(1, min(8, threads)))
        _, box, shapes = R.ref_run_simulation(native or strict, box, shapes, 500 + snapshots, 1.0, 1.0, 0.2, 0.15, 0,
                                              0.0, dom, dom[2], 1, 0)
        snapshots += 1
        ij = P.neighbour_pairs(box, shapes, W.SC_LENGTH + 2 * W.SC_RADIUS)
        pairs = P.build_pairs(box, shapes, ij, rng)
        got, _ = dev.pair_test_batch(pairs)
        want = P.threaded_verdicts(strict, pairs, threads)
        mismatches += int((got != want).sum())
        overlapping += int(want.sum())
        p1 = pairs["pos1"]; p2 = pairs["pos2"] + pairs["translation"]
        h1 = pairs["orientation1"][:, [2, 5, 8]] * (W.SC_LENGTH / 2)
        h2 = pairs["orientation2"][:, [2, 5, 8]] * (W.SC_LENGTH / 2)
        d2 = P.seg_dist2(p1, h1, p2, h2)
        near_contact += int((np.abs(d2 - contact2) <= 1e-12 * contact2).sum())
        total += pairs.shape[0]
    print(f"{total} pairs from {snapshots} reference-written snapshots: {mismatches} mismatches, {overlapping} "
          f"overlapping, {near_contact} within 1e-12 of contact")
    assert total >= 1.0e7 and overlapping > 1.0e5
    assert mismatches == 0
    dev.close()
