"""Runs the committed golden fixtures (tests/golden/) against any backend exposing the small surface below.
Backends: the plain-C oracle port (CPU, always), the compiled reference (CPU, where oracle/_ref exists) and the CUDA
path through the C ABI (GPU tests)."""
import json
import math
import os

import numpy as np

import rampack_b200 as rp
import helpers as H

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_vec = None


def vectors():
    global _vec
    if _vec is None:
        _vec = np.load(os.path.join(GOLDEN, "ref_vectors.npz"))
    return _vec


def known_answers():
    with open(os.path.join(GOLDEN, "known_answers.json")) as f:
        return json.load(f)


def desc_from_json(s):
    hard = {"sphere": rp.RP_HARD_SPHERE, "spherocylinder": rp.RP_HARD_SPHEROCYLINDER,
            "polysphere": rp.RP_HARD_POLYSPHERE, None: rp.RP_HARD_NONE}[s.get("hard")]
    soft = {"lj": rp.RP_SOFT_LJ, "wca": rp.RP_SOFT_WCA, "sic": rp.RP_SOFT_SIC, None: rp.RP_SOFT_NONE}[s.get("soft")]
    return rp.make_desc(hard=hard, soft=soft, radius=s.get("radius", 0.5 if soft else 0.0),
                        length=s.get("length", 0.0), epsilon=s.get("epsilon", 0.0), sigma=s.get("sigma", 0.0),
                        centres=s.get("centres"), centre_radius=s.get("radii"))


SHAPES = {
    "sphere": H.desc_sphere,
    "spherocylinder": H.desc_sc,
    "spherocylinder_short": lambda: H.desc_sc(l=1.0, r=0.5),
    "kmer3": lambda: H.desc_kmer(3),
    "polysc": H.desc_polysc,
}
SOFT = {"lj": rp.RP_SOFT_LJ, "wca": rp.RP_SOFT_WCA, "sic": rp.RP_SOFT_SIC}


def triclinic_box(lx, ly, lz):
    return np.array([[lx, 0.2 * ly, 0.1 * lz], [0, ly, -0.15 * lz], [0, 0, lz]])


PACKINGS = {
    "sc_ortho": lambda: H.desc_sc(l=3.0, r=0.5),
    "sc_tric": lambda: H.desc_sc(l=3.0, r=0.5),
    "sc_nogrid": lambda: H.desc_sc(l=3.0, r=0.5),
    "sphere_tric": H.desc_sphere,
    "kmer3_tric": lambda: H.desc_kmer(3),
    "polysc_tric": H.desc_polysc,
    "lj_ortho": lambda: H.desc_kmer(3, soft=rp.RP_SOFT_LJ, hard=False),
    "wca_tric": lambda: H.desc_kmer(3, soft=rp.RP_SOFT_WCA, hard=False),
    "sic_nogrid": lambda: H.desc_kmer(3, soft=rp.RP_SOFT_SIC, hard=False, sigma=1.5),
}


def golden_pairs(name, energy=False):
    v = vectors()
    raw = v[("epairs_" if energy else "pairs_") + name]
    return np.frombuffer(raw.tobytes(), dtype=rp.PAIR_DTYPE)


def make_pair(pos1, o1, idx1, pos2, o2, idx2, translation=(0, 0, 0)):
    p = np.zeros(1, dtype=rp.PAIR_DTYPE)
    p["pos1"], p["orientation1"], p["idx1"] = pos1, np.asarray(o1).reshape(9), idx1
    p["pos2"], p["orientation2"], p["idx2"] = pos2, np.asarray(o2).reshape(9), idx2
    p["translation"] = translation
    return p


def min_image_translation(box, p1, p2):
    """What BoundaryConditions::getTranslation returns for a pair in an orthorhombic periodic box (the known-answer
    PBC cases only use cubes): the lattice vector bringing p2 next to p1."""
    box = np.asarray(box).reshape(3, 3)
    rel = np.linalg.solve(box, np.asarray(p2, dtype=float) - np.asarray(p1, dtype=float))
    return -(box @ np.round(rel))


def close(a, b, rel=1e-10, abs_=1e-12):
    if math.isinf(a) or math.isinf(b):
        return a == b
    return abs(a - b) <= max(abs_, rel * max(abs(a), abs(b)))
