"""Device-side order parameters against the reference's own observable classes (NematicOrder, SmecticOrder —
src/core/observables/NematicOrder.cpp:9-55, SmecticOrder.cpp:14-62) evaluated on the same packing.  Bar: 1e-12
(the reference adds particle by particle, the device in a fixed block tree: same terms, different rounding)."""
import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import helpers as H
from test_gpu_parity import triclinic_box

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not R.have_ref(), reason="checked against oracle/_ref")]


def states():
    rng = np.random.default_rng(21)
    box, lattice = H.sc_lattice((10, 10, 4))
    # layered and aligned, slightly perturbed: P2 close to 1, a strong (0, 0, 4) density wave
    smectic = lattice.copy()
    smectic[:, :3] += rng.normal(scale=0.05, size=(lattice.shape[0], 3))
    smectic[:, :3] = np.mod(smectic[:, :3], np.diag(box))
    tilt = H.random_rotations(rng, lattice.shape[0])
    small = np.eye(3)[None] * 0.9 + 0.1 * tilt
    u, _, vt = np.linalg.svd(small)
    smectic[:, 3:] = (u @ vt).reshape(-1, 9)
    tbox = triclinic_box(22, 25, 30)
    return {"smectic": (box, smectic), "isotropic_triclinic": (tbox, H.random_packing(rng, 3000, tbox)),
            "aligned_random_positions": (tbox, H.random_packing(rng, 2500, tbox, aligned=True))}


@pytest.mark.parametrize("name", ["smectic", "isotropic_triclinic", "aligned_random_positions"])
def test_order_parameters(device, name):
    box, shapes = states()[name]
    desc = H.desc_sc()
    device.toggle_overlap_counting(False)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    hkl_range = (3, 3, 6)
    want = ref.observables(hkl_range)
    if name != "aligned_random_positions":          # perfectly aligned: the reference's Assert(p > 0) is at its edge
        q, p2 = device.nematic_order((0.0, 0.0, 1.0))
        assert np.allclose(q, want["q"], rtol=0, atol=1e-13)
        assert p2 == pytest.approx(want["p2"], abs=1e-12)
    tau, hkl, k = device.smectic_order(hkl_range)
    assert tau == pytest.approx(want["tau"], rel=1e-11, abs=1e-15)
    if name == "smectic":
        assert hkl == want["hkl"] == (0, 0, 4) and p2 > 0.9
        assert np.allclose(k, want["k"], rtol=1e-14, atol=1e-14)


def test_order_parameters_at_scale_are_deterministic(device):
    box, shapes = H.sc_lattice((60, 60, 30))         # 108 000 particles
    device.set_interaction(H.desc_sc())
    device.upload(box, shapes)
    for cycle in range(3):
        device.sweep(rp.RP_MOVE_ROTOTRANSLATION, 0.2, 0.1, cycle)
    a = (device.nematic_order(), device.smectic_order((1, 1, 32)))
    b = (device.nematic_order(), device.smectic_order((1, 1, 32)))
    assert np.array_equal(a[0][0], b[0][0]) and a[0][1] == b[0][1] and a[1][0] == b[1][0]
    assert a[0][1] > 0.95 and a[1][1] == (0, 0, 30)
    _, s = device.download()
    axes = s[:, 3:].reshape(-1, 3, 3)[:, :, 2]
    qn = 1.5 * (axes[:, :, None] * axes[:, None, :]).mean(axis=0) - 0.5 * np.eye(3)
    assert np.allclose(a[0][0], qn, rtol=0, atol=1e-13)


def test_observable_preconditions(device):
    box, shapes = H.sc_lattice((8, 8, 3))
    device.set_interaction(H.desc_sc())
    device.upload(box, shapes)
    with pytest.raises(rp.PreconditionError):
        device.smectic_order((0, 0, 0))


@pytest.mark.parametrize("name,max_r,n_bins", [("lattice", 4.0, 40), ("smectic", 5.9, 59),
                                               ("isotropic_triclinic", 9.0, 90), ("aligned_random_positions", 3.0, 7)])
def test_pair_density_correlation(device, name, max_r, n_bins):
    """rp_pair_density_correlation against one snapshot of the reference's PairDensityCorrelation + RadialEnumerator
    (PairDensityCorrelation.cpp:8-41, RadialEnumerator.cpp:12-50).  Pair counts per bin are integers: equal.  The
    lattice puts thousands of pairs on bin boundaries (spacing 1.2 = 12 bins of 0.1): which side they fall on is decided
    by the last bit of sqrt and of the division by the bin width — the device follows the reference's arithmetic."""
    if name == "lattice":
        box, shapes = H.sc_lattice((8, 8, 4))
    else:
        box, shapes = states()[name]
    desc = H.desc_sc()
    device.toggle_overlap_counting(False)
    device.set_interaction(desc)
    device.upload(box, shapes)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    want_g, want_counts = ref.pair_density(max_r, n_bins)
    counts, g = device.pair_density_correlation(max_r, n_bins)
    assert want_counts.sum() > 1000
    assert np.array_equal(counts, want_counts.astype(np.uint64))
    assert np.allclose(g, want_g, rtol=1e-12, atol=0)
    # beyond half a box height the minimum image is not the nearest one: refused, the caller keeps the host path
    from rampack_b200 import capi
    with pytest.raises(capi.PreconditionError, match="half of the smallest box height"):
        device.pair_density_correlation(0.51 * float(np.min(np.linalg.norm(box, axis=0))) + 1.0, 10)


def test_pair_density_correlation_at_scale(device):
    """N = 108 000: the reference's all-pairs loop would take minutes; hold the device to a cell-list count in numpy of
    the pairs inside max_r and to the g -> 1 limit of an ideal gas of points."""
    rng = np.random.default_rng(8)
    box = np.diag([90.0, 80.0, 70.0])
    n = 108_000
    shapes = np.zeros((n, 12))
    shapes[:, :3] = rng.uniform(0, 1, size=(n, 3)) * np.diag(box)
    shapes[:, 3:] = np.eye(3).ravel()
    device.toggle_overlap_counting(False)
    device.set_interaction(H.desc_sphere(0.01))
    device.upload(box, shapes)
    counts, g = device.pair_density_correlation(5.0, 25)
    from scipy.spatial import cKDTree
    tree = cKDTree(shapes[:, :3], boxsize=np.diag(box))
    assert int(counts.sum()) == (tree.count_neighbors(tree, 5.0) - n) // 2
    assert np.all(np.abs(g[5:] - 1.0) < 0.05)


def test_pair_density_cell_list_equals_all_pairs(device):
    """The cell-list form of g(r) (boxes at least three cells of max_r high along every axis) visits the same pairs as
    the all-pairs form (RP_GR_ALL_PAIRS=1 forces it) and bins them with the same FP64 arithmetic: equal counts, also
    in a sheared box, with a focal point off the mass centre, and with particles sitting exactly on cell boundaries."""
    import os
    rng = np.random.default_rng(21)
    box = np.array([[44.0, 6.0, -3.0], [0.0, 41.0, 5.0], [0.0, 0.0, 47.0]])
    n = 30_000
    rel = rng.uniform(0, 1, size=(n, 3))
    rel[:200] = np.round(rel[:200] * 6) / 6 % 1.0            # on cell boundaries of a 6 x 6 x 6 grid
    rel = np.clip(rel, 1e-12, 1 - 1e-12)                     # (strictly inside the box after the round trip)
    shapes = np.zeros((n, 12))
    shapes[:, :3] = rel @ box.T
    shapes[:, 3:] = H.random_rotations(rng, n).reshape(n, 9)
    device.toggle_overlap_counting(False)
    device.set_interaction(H.desc_sphere(0.01))
    device.upload(box, shapes)
    for focal in (None, (0.0, 0.3, 1.1)):
        counts, g = device.pair_density_correlation(6.5, 65, focal)
        os.environ["RP_GR_ALL_PAIRS"] = "1"
        try:
            counts_all, g_all = device.pair_density_correlation(6.5, 65, focal)
        finally:
            os.environ.pop("RP_GR_ALL_PAIRS", None)
        assert counts.sum() > 100_000
        assert np.array_equal(counts, counts_all)
        assert np.array_equal(g, g_all)
