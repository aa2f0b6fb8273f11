"""TEST INFRASTRUCTURE ONLY — picks the strongest checker available where we are: the UNMODIFIED reference compiled
into oracle/_ref/ (strict flags) when that file exists, else the plain-C restatement oracle/liboracle_port.so.
Used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs; never by rampack_b200/."""
from . import refwrap as R


def kind():
    return "reference" if R.have_ref() else "port"


def traits(desc):
    return R.RefTraits(desc) if R.have_ref() else R.PortTraits(desc)


def packing(desc, box, shapes):
    t = traits(desc)
    return (R.RefPacking if R.have_ref() else R.PortPacking)(t, box, shapes)


def pair_verdicts(desc, pairs):
    return traits(desc).pair_test_batch(pairs)[0]


def pair_energies(desc, pairs):
    return traits(desc).pair_test_batch(pairs, want_verdict=False, want_energy=True)[1]


def count_total_overlaps(desc, box, shapes, early_exit=False):
    return packing(desc, box, shapes).count_total_overlaps(early_exit)


def total_energy(desc, box, shapes):
    return packing(desc, box, shapes).total_energy()
