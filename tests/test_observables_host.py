"""The numpy observables the statistical parity test averages (tests/stat_cases.py) against the reference's own
observable classes run on the same packing through oracle/_ref (NumberDensity, NematicOrder, SmecticOrder,
PairDensityCorrelation + RadialEnumerator).  CPU test: pins the restatements, no device involved."""
import numpy as np
import pytest

from oracle import refwrap as R
import helpers as H
import stat_cases as S

pytestmark = pytest.mark.skipif(not R.have_ref(), reason="needs oracle/_ref (the compiled reference)")


def fluid(n=900, seed=4):
    rng = np.random.default_rng(seed)
    box = np.diag([24.0, 22.0, 26.0])
    return box, H.random_packing(rng, n, box)


def test_pair_density_in_shell_is_the_reference_g_of_r():
    box, shapes = fluid()
    ref = R.RefPacking(R.RefTraits(H.desc_sc()), box, shapes)
    n_bins, max_r = 50, 10.0
    g, counts = ref.pair_density(max_r, n_bins)
    step = max_r / n_bins
    for lo, hi in ((5, 15), (10, 12), (0, 50)):
        r_lo, r_hi = lo * step, hi * step
        shell = 4.0 / 3.0 * np.pi * (r_hi ** 3 - r_lo ** 3)
        rho = shapes.shape[0] / abs(np.linalg.det(box))
        want = 2.0 * counts[lo:hi].sum() / (shapes.shape[0] * rho * shell)      # pairs i < j, both directions
        got = S.pair_density_in_shell(box, shapes, r_lo, r_hi)
        assert got == pytest.approx(want, rel=1e-12)
    # the reference's own normalisation of a single bin (PairDensityCorrelation.cpp:12-16)
    k = 20
    shell = 4.0 / 3.0 * np.pi * (((k + 1) * step) ** 3 - (k * step) ** 3)
    assert g[k] == pytest.approx(2.0 * counts[k] / (shapes.shape[0] * (shapes.shape[0] / abs(np.linalg.det(box))) * shell),
                                 rel=1e-12)


def test_order_parameters_are_the_reference_ones():
    rng = np.random.default_rng(9)
    box, shapes = H.sc_lattice((8, 8, 4))
    shapes[:, :3] = np.mod(shapes[:, :3] + rng.normal(scale=0.08, size=(shapes.shape[0], 3)), np.diag(box))
    tilt = H.random_rotations(rng, shapes.shape[0])
    u, _, vt = np.linalg.svd(np.eye(3)[None] * 0.85 + 0.15 * tilt)
    shapes[:, 3:] = (u @ vt).reshape(-1, 9)
    ref = R.RefPacking(R.RefTraits(H.desc_sc()), box, shapes)
    want = ref.observables((0, 0, 8))
    assert S.nematic_p2(shapes) == pytest.approx(want["p2"], abs=1e-12)
    # SmecticOrder divides the structure amplitude by the volume (SmecticOrder.cpp:31), the numpy helper by N
    volume = abs(np.linalg.det(box))
    assert S.smectic_tau(box, shapes) == pytest.approx(want["tau"] * volume / shapes.shape[0], rel=1e-11)
    assert S.number_density(box, shapes) == pytest.approx(shapes.shape[0] / volume, rel=1e-15)
