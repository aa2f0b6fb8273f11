"""The reference's overlap-relaxation validation case on the device (test/validation_tests/SimulationTest.cpp:322-358:
"Simulation: overlap reduction for hard sphere liquid"): 216 hard spheres on a 6 x 6 x 6 lattice in a box of 5^3 — the
spheres overlap — relaxed in overlap-counting mode with a square-inverse-core helper interaction
(CompoundShapeTraits(sphere, sphere + SquareInverseCoreInteraction(1, 1))) at T = 1, p = 5: the run must end with no
overlaps WITHOUT over-expanding the packing (equilibrium density 0.754; the reference demands > 0.75)."""
import numpy as np
import pytest

import helpers as H
import rampack_b200 as rp

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed", [1234, 99])
def test_overlap_relaxation_of_a_too_dense_sphere_lattice(seed):
    dev = rp.Device(0, seed=seed)
    compound = rp.make_desc(hard=rp.RP_HARD_SPHERE, radius=0.5, soft=rp.RP_SOFT_SIC, epsilon=1.0, sigma=1.0)
    dev.set_interaction(compound)
    box, shapes = H.sphere_lattice((6, 6, 6), 5.0 / 6.0)
    dev.upload(box, shapes)
    dev.toggle_overlap_counting(True)
    start = dev.cached_overlaps()
    assert start > 216                       # every sphere overlaps its six lattice neighbours
    # Simulation::relaxOverlaps (Simulation.cpp:218-323): cycles with step adaptation until the count reaches zero
    sim = rp.NativeSimulation(dev, [(rp.RP_MOVE_ROTOTRANSLATION, 0.1, 0.1)], 1.0, 5.0, ("delta_v", 1.0), seed=seed)
    cycles = 0
    while dev.cached_overlaps() > 0 and cycles < 60000:
        sim.run(20, adjust_step_sizes=True)
        cycles += 20
    left = dev.cached_overlaps()
    b, s = dev.download()
    density = s.shape[0] / abs(np.linalg.det(b))
    print(f"seed {seed}: {start} -> {left} overlaps in {cycles} cycles, density {density:.4f}")
    assert left == 0
    dev.toggle_overlap_counting(False)
    assert dev.count_total_overlaps(False) == 0
    # the oracle agrees that the relaxed packing is overlap-free for the plain hard sphere
    from oracle import checker
    assert checker.count_total_overlaps(H.desc_sphere(), b, s) == 0
    assert density > 0.75
    sim.close()
    dev.close()
