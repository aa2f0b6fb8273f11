"""CPU tests (no GPU): the plain-C oracle port is pinned against
  (1) the reference's own known-answer cases (tests/golden/known_answers.json, transcribed from its unit tests),
  (2) outputs of the UNMODIFIED reference on seeded inputs (tests/golden/ref_vectors.npz, made by make_golden.py),
  (3) the unmodified reference compiled here (oracle/_ref), live, when that library exists.
"""
import math

import numpy as np
import pytest

import rampack_b200 as rp
from oracle import refwrap as R
import golden_cases as G
import helpers as H


def port_packing(desc, box, shapes):
    return R.PortPacking(R.PortTraits(desc), box, shapes)


@pytest.mark.parametrize("case", [c for c in G.known_answers() if c["kind"] in ("pair", "shape_pair")],
                         ids=lambda c: c["source"].split()[0])
def test_known_answer_overlaps(case):
    desc = G.desc_from_json(case["shape"])
    t = R.PortTraits(desc)
    o1, o2 = np.array(case["o1"]).reshape(3, 3), np.array(case["o2"]).reshape(3, 3)
    pos1, pos2 = np.array(case["pos1"], dtype=float), np.array(case["pos2"], dtype=float)
    centres = case["shape"].get("centres") or [[0, 0, 0]]
    for swap in (False, True):       # the reference tests assert both argument orders
        a = (pos1, o1) if not swap else (pos2, o2)
        b = (pos2, o2) if not swap else (pos1, o1)
        got = False
        for i, c1 in enumerate(centres):
            for j, c2 in enumerate(centres):
                p1 = a[0] + a[1] @ np.array(c1, dtype=float)
                p2 = b[0] + b[1] @ np.array(c2, dtype=float)
                pair = G.make_pair(p1, a[1], i, p2, b[1], j)
                v, _ = t.pair_test_batch(pair, pbc_box=case["pbc_box"])
                got = got or bool(v[0])
                if case["pbc_box"] is not None:      # the HardcodedTranslation route must agree with true PBC
                    pair_t = G.make_pair(p1, a[1], i, p2, b[1], j, G.min_image_translation(case["pbc_box"], p1, p2))
                    assert bool(t.pair_test_batch(pair_t)[0][0]) == bool(v[0])
        assert got == case["overlap"], case["source"]


@pytest.mark.parametrize("case", [c for c in G.known_answers() if c["kind"] == "energy"],
                         ids=lambda c: c["source"].split("/")[1].split(".")[0])
def test_known_answer_energies(case):
    desc = G.desc_from_json(case["shape"])
    pair = G.make_pair(case["pos1"], np.eye(3), 0, case["pos2"], np.eye(3), 0)
    _, e = R.PortTraits(desc).pair_test_batch(pair, want_verdict=False, want_energy=True)
    assert abs(e[0] - case["energy"]) <= max(case["margin"], 1e-12 * abs(case["energy"]))


@pytest.mark.parametrize("case", [c for c in G.known_answers() if c["kind"] in ("scaling", "translation")],
                         ids=lambda c: c["source"].split("(")[1][:-1].replace(" ", "_"))
def test_known_answer_packing(case):
    desc = G.desc_from_json(case["shape"])
    box = np.array(case["box"]).reshape(3, 3)
    shapes = H.shapes_from(np.array(case["positions"], dtype=float), np.tile(np.eye(3), (3, 1, 1)))
    pk = port_packing(desc, box, shapes)
    want = math.inf if case["dE"] == "inf" else case["dE"]
    if case["kind"] == "scaling":
        assert pk.try_scaling(box * case["factor"]) == want
        if case["positions_after"] is not None:
            assert np.allclose(pk.get()[1][:, :3], case["positions_after"], atol=1e-9)
        pk.revert_scaling()
        assert np.array_equal(pk.get()[1], shapes) and np.array_equal(pk.get()[0], box)
    else:
        assert pk.try_move(case["particle"], rp.RP_MOVE_TRANSLATION, case["translation"]) == want
        assert np.array_equal(pk.get()[1], shapes)       # nothing changes before accept
        if case["positions_after"] is not None:
            pk.accept_move()
            assert np.allclose(pk.get()[1][:, :3], case["positions_after"], atol=1e-9)


@pytest.mark.parametrize("name", list(G.SHAPES))
def test_golden_pair_verdicts(name):
    pairs = G.golden_pairs(name)
    got, _ = R.PortTraits(G.SHAPES[name]()).pair_test_batch(pairs)
    want = G.vectors()["verdict_" + name]
    assert 0.02 < want.mean() < 0.98
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name", list(G.SOFT))
def test_golden_pair_energies(name):
    desc = H.desc_kmer(3, soft=G.SOFT[name], hard=False, epsilon=1.3, sigma=0.9)
    pairs = G.golden_pairs(name, energy=True)
    _, got = R.PortTraits(desc).pair_test_batch(pairs, want_verdict=False, want_energy=True)
    assert np.array_equal(got, G.vectors()["energy_" + name])


@pytest.mark.parametrize("name", list(G.PACKINGS))
def test_golden_packings(name):
    v = G.vectors()
    desc = G.PACKINGS[name]()
    box, shapes = v[f"pk_{name}_box"], v[f"pk_{name}_shapes"]
    pk = port_packing(desc, box, shapes)
    div = pk.grid_divisions()
    want_div = tuple(int(x) for x in v[f"pk_{name}_div"])
    assert (div or (0, 0, 0)) == want_div
    assert ("nogrid" in name) == (div is None)
    assert pk.count_total_overlaps(False) == int(v[f"pk_{name}_overlaps"][0])
    assert pk.count_total_overlaps(True) == min(1, int(v[f"pk_{name}_overlaps"][0]))
    assert G.close(pk.total_energy(), float(v[f"pk_{name}_energy"][0]))
    assert pk.average_neighbours() == pytest.approx(float(v[f"pk_{name}_avg_neigh"][0]), rel=1e-12)
    if f"pk_{name}_cells" in v:
        cells, pdiv = R.port_cell_numbers(box, float(v[f"pk_{name}_cellsize"][0]), shapes[:, :3])
        assert np.array_equal(cells.astype(np.int64), v[f"pk_{name}_cells"])
        assert pdiv == want_div
    for row, want in zip(v[f"pk_{name}_moves"], v[f"pk_{name}_moves_dE"]):
        i, kind, acc = int(row[0]), int(row[1]), bool(row[2])
        got = pk.try_move(i, kind, row[3:6], row[6:15].reshape(3, 3))
        assert G.close(got, float(want), rel=1e-9), (name, got, want)
        if acc:
            pk.accept_move()
    assert np.array_equal(pk.get()[1], v[f"pk_{name}_after_moves"])
    for factor, want in v[f"pk_{name}_scalings"]:
        got = pk.try_scaling(pk.get()[0] * factor)
        assert G.close(got, float(want), rel=1e-9), (name, factor, got, want)
        pk.revert_scaling()


def test_golden_geometry():
    v = G.vectors()
    box = v["geom_box"]
    h, vol, inv = R.port_box_info(box)
    assert np.array_equal(h, v["geom_heights"]) and vol == v["geom_volume"][0] and np.array_equal(inv, v["geom_inverse"])
    for p, want in zip(v["geom_pbc_points"], v["geom_pbc_corrections"]):
        assert np.array_equal(R.port_pbc_correction(box, p), want)
    for a, g, want in zip(v["geom_rot_axes"], v["geom_rot_angles"], v["geom_rot_matrices"]):
        assert np.array_equal(R.port_rotation_matrix(a, g), want)


def test_ranges():
    for make in list(G.SHAPES.values()) + list(G.PACKINGS.values()):
        d = make()
        t = R.PortTraits(d)
        assert t.range_radius > 0 and t.total_range_radius >= t.range_radius
        if R.have_ref():
            r = R.RefTraits(d)
            assert t.range_radius == r.range_radius and t.total_range_radius == r.total_range_radius


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
def test_port_against_live_reference_counting_mode():
    """Overlap-counting semantics (-inf / +inf / 0 and the cached count) against the unmodified reference."""
    desc = H.desc_kmer(3)
    rng = np.random.default_rng(9)
    box = np.diag([12.0, 12.0, 13.0])
    shapes = H.random_packing(rng, 300, box)
    ref = R.RefPacking(R.RefTraits(desc), box, shapes)
    port = port_packing(desc, box, shapes)
    ref.toggle_overlap_counting(True)
    port.toggle_overlap_counting(True)
    assert ref.cached_overlaps() == port.cached_overlaps() > 0
    seen = set()
    for step in range(200):
        i = int(rng.integers(0, 300))
        t = rng.uniform(-0.5, 0.5, size=3)
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        rot = R.ref_rotation_matrix(axis, rng.uniform(-0.5, 0.5))
        want, got = ref.try_move(i, 2, t, rot), port.try_move(i, 2, t, rot)
        assert want == got
        seen.add(want)
        if want < math.inf and rng.uniform() < 0.7:
            ref.accept_move()
            port.accept_move()
        assert ref.cached_overlaps() == port.cached_overlaps()
    assert {-math.inf, math.inf, 0.0} <= seen
    for factor in (0.95, 1.05):
        assert ref.try_scaling(box * factor) == port.try_scaling(box * factor)
        assert ref.cached_overlaps() == port.cached_overlaps()
        ref.revert_scaling()
        port.revert_scaling()


def test_port_cycle_loop_keeps_packing_valid():
    box, shapes = H.sc_lattice((6, 6, 3))
    desc = H.desc_sc()
    pk = port_packing(desc, box, shapes)
    res = pk.run_cycles(seed=1, cycles=3, trans_step=0.3, rot_step=0.2, box_moves=True, pressure=1.2, box_step=0.05)
    assert res.moves_attempted == 3 * shapes.shape[0] and 0 < res.moves_accepted < res.moves_attempted
    assert res.scalings_attempted == 3
    assert pk.count_total_overlaps(False) == 0
