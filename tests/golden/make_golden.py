"""Generates tests/golden/*.npz and known_answers.json.  Run HERE (where /root/reference is mounted and oracle/_ref is
built):  python tests/golden/make_golden.py

Two kinds of fixtures:
 1. known_answers.json — known-answer cases TRANSCRIBED from the reference's own unit tests (file:line cited per
    case, paths relative to /root/reference/test/unit_tests/core); the expected values are the ones the reference's
    tests assert, not anything computed here.
 2. ref_*.npz — seeded inputs with the outputs of the UNMODIFIED reference (oracle/_ref/librampack_ref.so, built
    with -O3 -ffp-contract=off from the sources under /root/reference by oracle/Makefile).
The GPU box has no /root/reference: tests read only these committed files.
"""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import helpers as H  # noqa: E402
import rampack_b200 as rp  # noqa: E402  (struct layouts only; no device is touched)
from oracle import refwrap as R  # noqa: E402


def euler_rotation(ax, ay, az):
    """Matrix<3,3>::rotation(ax, ay, az) (src/geometry/Matrix.tpp:405-435)."""
    sx, cx, sy, cy, sz, cz = math.sin(ax), math.cos(ax), math.sin(ay), math.cos(ay), math.sin(az), math.cos(az)
    return np.array([[cy * cz, sx * sy * cz - cx * sz, cx * sy * cz + sx * sz],
                     [cy * sz, sx * sy * sz + cx * cz, cx * sy * sz - sx * cz],
                     [-sy, sx * cy, cx * cy]])


def known_answers():
    cases = []
    corr = euler_rotation(0, math.pi / 2, 0)
    sc_rots = {
        "sphere-sphere": ([0.92669949443125, -0.3009522885099318, 0.2250683608628711,
                           0.3235062902233933, 0.943614995716347, -0.07024542721869537,
                           -0.191237358292679, 0.1379074323590797, 0.971807497858173],
                          [0.1017126282787795, -0.2093565370521866, 0.972535028491077,
                           0.5191108031629524, 0.845122866944617, 0.127637431056981,
                           -0.848633322046771, 0.4918711011645504, 0.1946389081120092],
                          [1, 2, 3], [-3.5, 2.9, -0.5]),
        "sphere-cylinder": ([0.936848795202308, -0.03157560239884583, 0.34830635403497,
                             -0.03157560239884583, 0.984212198800577, 0.174153177017485,
                             -0.34830635403497, -0.174153177017485, 0.921060994002885],
                            [0.07918504528890558, 0.3698006820181268, -0.925730621823391,
                             -0.2008438095940726, 0.915521563787973, 0.348542723904781,
                             0.976417683550607, 0.1583278933673811, 0.1467677942585273],
                            [1, 2, 3], [-0.5, 4, -2]),
        "cylinder-cylinder": ([0.921060994002885, 0.1446263415347036, 0.361565853836759,
                               -0.1446263415347036, 0.989111861241777, -0.02722034689555686,
                               -0.361565853836759, -0.02722034689555686, 0.931949132761108],
                              [0.4161880560663263, 0.5570946854720225, -0.7186327388914037,
                               -0.902545539870646, 0.1570999152673587, -0.4009129145869377,
                               -0.1104493116652922, 0.815453939865286, 0.5681864320017205],
                              [1, 2, 3], [-0.5, 1.7, -0.8]),
    }
    for name, (r1, r2, t1, t2) in sc_rots.items():
        o1 = (np.array(r1).reshape(3, 3) @ corr).reshape(9).tolist()
        o2 = (np.array(r2).reshape(3, 3) @ corr).reshape(9).tolist()
        for scale, want in ((0.95, True), (1.05, False)):
            cases.append({"source": "shapes/SpherocylinderTraitsTest.cpp:14-101 (" + name + ")", "kind": "pair",
                          "shape": {"hard": "spherocylinder", "length": 3, "radius": 2},
                          "pos1": (np.array(t1) * scale).tolist(), "o1": o1,
                          "pos2": (np.array(t2) * scale).tolist(), "o2": o2, "pbc_box": None, "overlap": want})
    eye = np.eye(3).reshape(9).tolist()
    box100 = np.diag([100.0] * 3).reshape(9).tolist()
    for z, want in ((96.6, True), (96.4, False)):
        cases.append({"source": "shapes/SpherocylinderTraitsTest.cpp:103-122 (boundary conditions)", "kind": "pair",
                      "shape": {"hard": "spherocylinder", "length": 3, "radius": 2},
                      "pos1": [50, 50, z], "o1": eye, "pos2": [50, 50, 3.5], "o2": eye, "pbc_box": box100,
                      "overlap": want})
    # PolysphereTraitsTest.cpp:19-50: dimer {0,0,0} r=0.5 and {3,0,0} r=1 in a periodic 10-box
    box10 = np.diag([10.0] * 3).reshape(9).tolist()
    rot_pi = euler_rotation(0, math.pi, 0).reshape(9).tolist()
    for x, want in ((1.01, True), (0.99, False)):
        cases.append({"source": "shapes/PolysphereTraitsTest.cpp:19-50", "kind": "shape_pair",
                      "shape": {"hard": "polysphere", "centres": [[0, 0, 0], [3, 0, 0]], "radii": [0.5, 1]},
                      "pos1": [x, 5, 5], "o1": eye, "pos2": [9, 5, 5], "o2": rot_pi, "pbc_box": box10,
                      "overlap": want})
    # central potentials
    cases.append({"source": "interactions/LennardJonesInteractionTest.cpp:10-16", "kind": "energy",
                  "shape": {"soft": "lj", "epsilon": 3, "sigma": 2}, "pos1": [1, 0, 0], "pos2": [5, 0, 0],
                  "energy": -0.1845703125, "margin": 1e-12})
    cases.append({"source": "interactions/RepulsiveLennardJonesInteractionTest.cpp:14-18", "kind": "energy",
                  "shape": {"soft": "wca", "epsilon": 3, "sigma": 2}, "pos1": [1, 0, 0], "pos2": [3, 0, 0],
                  "energy": 3.0, "margin": 1e-12})
    cases.append({"source": "interactions/RepulsiveLennardJonesInteractionTest.cpp:20-24", "kind": "energy",
                  "shape": {"soft": "wca", "epsilon": 3, "sigma": 2}, "pos1": [1, 0, 0], "pos2": [5, 0, 0],
                  "energy": 0.0, "margin": 1e-12})
    cases.append({"source": "interactions/SquareInverseCoreInteractionTest.cpp:10-17", "kind": "energy",
                  "shape": {"soft": "sic", "epsilon": 3, "sigma": 2}, "pos1": [1, 0, 0], "pos2": [2, 0, 0],
                  "energy": 9.0, "margin": 1e-12})
    cases.append({"source": "interactions/SquareInverseCoreInteractionTest.cpp:10-17", "kind": "energy",
                  "shape": {"soft": "sic", "epsilon": 3, "sigma": 2}, "pos1": [1, 0, 0], "pos2": [4, 0, 0],
                  "energy": 0.0, "margin": 0.0})
    # PackingTest.cpp:86-160: three hard spheres r = 0.25 in a periodic 5-box
    pk = {"shape": {"hard": "sphere", "radius": 0.25}, "box": np.diag([5.0] * 3).reshape(9).tolist(),
          "positions": [[0.5, 0.5, 0.5], [4.5, 0.5, 0.5], [2.5, 2.5, 4.0]]}
    cases.append(dict(pk, source="PackingTest.cpp:101-109 (hard core upwards)", kind="scaling", factor=1.1, dE=0.0,
                      positions_after=[[0.55, 0.55, 0.55], [4.95, 0.55, 0.55], [2.75, 2.75, 4.4]]))
    cases.append(dict(pk, source="PackingTest.cpp:111-120 (downward without overlapping)", kind="scaling",
                      factor=0.51, dE=0.0,
                      positions_after=[[0.255, 0.255, 0.255], [2.295, 0.255, 0.255], [1.275, 1.275, 2.04]]))
    cases.append(dict(pk, source="PackingTest.cpp:122-137 (downward with overlapping)", kind="scaling", factor=0.49,
                      dE="inf", positions_after=None))
    cases.append(dict(pk, source="PackingTest.cpp:152-161 (translating, non-overlapping)", kind="translation",
                      particle=2, translation=[2, 2.45, -3.5], dE=0.0,
                      positions_after=[[0.5, 0.5, 0.5], [4.5, 0.5, 0.5], [4.5, 4.95, 0.5]]))
    cases.append(dict(pk, source="PackingTest.cpp:163-167 (translating, overlapping)", kind="translation",
                      particle=2, translation=[2, 2.55, -3.5], dE="inf", positions_after=None))
    return cases


def triclinic_box(lx, ly, lz):
    return np.array([[lx, 0.2 * ly, 0.1 * lz], [0, ly, -0.15 * lz], [0, 0, lz]])


SHAPES = {
    "sphere": (H.desc_sphere, 1.1, 1),
    "spherocylinder": (H.desc_sc, 6.2, 1),
    "spherocylinder_short": (lambda: H.desc_sc(l=1.0, r=0.5), 2.2, 1),
    "kmer3": (lambda: H.desc_kmer(3), 1.1, 3),
    "polysc": (H.desc_polysc, 2.4, 2),
}


def ref_vectors():
    out = {}
    # pair verdicts
    for name, (make, max_dist, k) in SHAPES.items():
        rng = np.random.default_rng(1000 + len(name))
        pairs = H.random_pairs(rng, 1000, max_dist, n_centres=k, translation_scale=20.0)
        verdict, _ = R.RefTraits(make()).pair_test_batch(pairs)
        out[f"pairs_{name}"] = np.frombuffer(pairs.tobytes(), dtype=np.uint8)
        out[f"verdict_{name}"] = verdict
    for soft_name, soft in (("lj", rp.RP_SOFT_LJ), ("wca", rp.RP_SOFT_WCA), ("sic", rp.RP_SOFT_SIC)):
        rng = np.random.default_rng(7)
        desc = H.desc_kmer(3, soft=soft, hard=False, epsilon=1.3, sigma=0.9)
        pairs = H.random_pairs(rng, 1000, 3.0, n_centres=3, translation_scale=5.0)
        _, energy = R.RefTraits(desc).pair_test_batch(pairs, want_verdict=False, want_energy=True)
        out[f"epairs_{soft_name}"] = np.frombuffer(pairs.tobytes(), dtype=np.uint8)
        out[f"energy_{soft_name}"] = energy
    # packings: grid geometry, cell numbers, overlap counts, energies
    configs = [("sc_ortho", H.desc_sc(l=3.0, r=0.5), 1200, np.diag([30.0, 40.0, 50.0])),
               ("sc_tric", H.desc_sc(l=3.0, r=0.5), 900, triclinic_box(28, 30, 45)),
               ("sc_nogrid", H.desc_sc(l=3.0, r=0.5), 60, np.diag([9.0, 9.5, 10.0])),
               ("sphere_tric", H.desc_sphere(), 700, triclinic_box(10, 11, 12)),
               ("kmer3_tric", H.desc_kmer(3), 600, triclinic_box(11, 12, 13)),
               ("polysc_tric", H.desc_polysc(), 600, triclinic_box(11, 12, 13)),
               ("lj_ortho", H.desc_kmer(3, soft=rp.RP_SOFT_LJ, hard=False), 500, np.diag([14.0, 15.0, 16.0])),
               ("wca_tric", H.desc_kmer(3, soft=rp.RP_SOFT_WCA, hard=False), 700, triclinic_box(12, 13, 14)),
               ("sic_nogrid", H.desc_kmer(3, soft=rp.RP_SOFT_SIC, hard=False, sigma=1.5), 40, np.diag([5.0, 5.0, 5.0]))]
    for name, desc, n, box in configs:
        rng = np.random.default_rng(sum(map(ord, name)))
        shapes = H.random_packing(rng, n, box)
        traits = R.RefTraits(desc)
        pk = R.RefPacking(traits, box, shapes)
        out[f"pk_{name}_box"] = box
        out[f"pk_{name}_shapes"] = shapes
        # without a neighbour grid getNeighbourGridCellDivisions() dereferences an empty optional (Packing.h:235-237);
        # the no-grid state shows as getAverageNumberOfNeighbours() == centres - 1 (Packing.cpp:1001-1006)
        no_grid = pk.average_neighbours() == n * max(desc.n_centres, 1) - 1
        assert no_grid == ("nogrid" in name)
        div = (0, 0, 0) if no_grid else pk.grid_divisions()
        out[f"pk_{name}_div"] = np.array(div, dtype=np.int64)
        out[f"pk_{name}_overlaps"] = np.array([pk.count_total_overlaps(False)], dtype=np.int64)
        out[f"pk_{name}_energy"] = np.array([pk.total_energy()])
        out[f"pk_{name}_avg_neigh"] = np.array([pk.average_neighbours()])
        if div[0] > 0 and desc.n_centres == 0:
            vol = abs(np.linalg.det(box))
            cell = max(traits.range_radius, np.cbrt(vol / n) / 5)
            cells, _ = R.ref_cell_numbers(box, cell, shapes[:, :3])
            out[f"pk_{name}_cells"] = cells.astype(np.int64)
            out[f"pk_{name}_cellsize"] = np.array([cell])
        # a sequence of explicit moves with the reference's verdicts (accepted when dE <= 0)
        mrng = np.random.default_rng(5)
        moves, des = [], []
        for step in range(120):
            i = int(mrng.integers(0, n))
            kind = int(mrng.integers(0, 3))
            t = mrng.uniform(-0.4, 0.4, size=3)
            axis = mrng.normal(size=3)
            axis /= np.linalg.norm(axis)
            rot = R.ref_rotation_matrix(axis, mrng.uniform(-0.5, 0.5))
            dE = pk.try_move(i, kind, t, rot)
            acc = dE <= 0
            if acc:
                pk.accept_move()
            moves.append(np.concatenate([[i, kind, acc], t, rot.reshape(9)]))
            des.append(dE)
        out[f"pk_{name}_moves"] = np.array(moves)
        out[f"pk_{name}_moves_dE"] = np.array(des)
        out[f"pk_{name}_after_moves"] = pk.get()[1]
        # box moves
        sc_res = []
        for factor in (1.04, 0.97, 0.85):
            dE = pk.try_scaling(pk.get()[0] * factor)
            sc_res.append([factor, dE])
            pk.revert_scaling()
        out[f"pk_{name}_scalings"] = np.array(sc_res)
    # geometry helpers
    rng = np.random.default_rng(3)
    box = triclinic_box(7, 8, 9)
    h, vol, inv = R.ref_box_info(box)
    out["geom_box"] = box
    out["geom_heights"] = h
    out["geom_volume"] = np.array([vol])
    out["geom_inverse"] = inv
    pts = rng.uniform(-20, 20, size=(200, 3))
    out["geom_pbc_points"] = pts
    out["geom_pbc_corrections"] = np.array([R.ref_pbc_correction(box, p) for p in pts])
    axes = rng.normal(size=(50, 3))
    axes /= np.linalg.norm(axes, axis=1)[:, None]
    angles = rng.uniform(-3, 3, size=50)
    out["geom_rot_axes"] = axes
    out["geom_rot_angles"] = angles
    out["geom_rot_matrices"] = np.array([R.ref_rotation_matrix(a, g) for a, g in zip(axes, angles)])
    return out


def main():
    with open(os.path.join(HERE, "known_answers.json"), "w") as f:
        json.dump(known_answers(), f, indent=1)
    vec = ref_vectors()
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **vec)
    print("wrote", len(vec), "arrays,", os.path.getsize(os.path.join(HERE, "ref_vectors.npz")) // 1024, "KiB")


if __name__ == "__main__":
    main()
