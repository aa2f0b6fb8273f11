"""CPU tests of the XenoCollide host side: the XCBodyBuilder mirror emits well-formed programs that the reference
harness (the reference's own geometry classes) accepts; no device involved."""
import pytest

import rampack_b200 as rp
from rampack_b200 import capi
from rampack_b200.xc import XCBodyBuilder
from oracle import refwrap as R
import xc_cases


def test_builder_stack_semantics():
    b = XCBodyBuilder().process("sphere 0.5 move 0 0 -0.5 cuboid 1 1 1 rot 0 0 45 move 0 0 0.5 wrap")
    p = b.release()
    assert p.n_nodes == 3 and p.n_entries == 2
    assert [p.nodes[i].kind for i in range(3)] == [capi.RP_XC_SPHERE, capi.RP_XC_CUBOID, capi.RP_XC_MAX]
    # wrap pops from the back: entry 0 is the former top of the stack (the cuboid)   XCBodyBuilder.cpp:207-218
    assert p.entries[0].node == 1 and p.entries[1].node == 0
    assert list(p.entries[0].pos) == [0.0, 0.0, 0.5] and list(p.entries[1].pos) == [0.0, 0.0, -0.5]
    assert p.entries[0].rot[0] == pytest.approx(2 ** -0.5) and p.entries[1].rot[0] == 1.0
    assert p.root[0] == 2
    with pytest.raises(ValueError):
        XCBodyBuilder().process("sphere 1 sphere 1").release()      # two shapes left: forgot sum / wrap
    with pytest.raises(ValueError):
        XCBodyBuilder().process("frobnicate 1")


@pytest.mark.skipif(not R.have_ref(), reason="oracle/_ref not built")
@pytest.mark.parametrize("name", list(xc_cases.CASES))
def test_reference_accepts_every_program(name):
    make, max_dist = xc_cases.CASES[name]
    desc, k = make()
    ref = R.RefTraits(desc)
    # XenoCollideTraits::getRangeRadius = 2 * max circumsphere radius (XenoCollideTraits.h:188-202)
    assert 0 < ref.range_radius < 2 * max_dist
    assert ref.total_range_radius >= ref.range_radius
