"""CPU tests of the drop-in boundary: librampack_b200.so builds, loads and exports exactly what include/*.h declares;
no compute entry point is called (there is no GPU here) except to check that they fail loudly without a device."""
import ctypes as C
import os
import re

import pytest

import rampack_b200 as rp
from rampack_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "rampack_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rp_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = C.CDLL(capi.library_path()) if os.path.exists(capi.library_path()) else capi.load()
    declared = header_symbols()
    assert len(declared) >= 24
    for name in declared:
        assert hasattr(lib, name), f"{name} is declared in include/rampack_b200.h but not exported"
    assert sorted(capi.SYMBOLS) == declared, "rampack_b200/capi.py SYMBOLS and the header disagree"


def test_struct_layouts_match_header():
    # sizes implied by include/rampack_b200.h (all members naturally aligned)
    assert C.sizeof(capi.InteractionDesc) == 8 + 4 * 8 + 8 + 16 * 3 * 8 + 16 * 8 + 16 * 3 * 8 + 8
    assert C.sizeof(capi.XcNode) == 16 + 6 * 8 and C.sizeof(capi.XcEntry) == 8 + 12 * 8
    assert C.sizeof(capi.XcProgram) == 8 + 32 * 64 + 32 * 104 + 16 * 4
    assert C.sizeof(capi.Config) == 24
    assert C.sizeof(capi.Move) == 16 + 12 * 8
    assert C.sizeof(capi.SweepParams) == 8 + 4 * 8 + 7 * 8          # + flip_axis[3], flip_origin[3], flip_every
    assert C.sizeof(capi.SweepCounters) == 7 * 8
    assert capi.PAIR_DTYPE.itemsize == 29 * 8
    assert C.sizeof(capi.GridInfo) == 4 * 4 + 8 + 8 + 8


def test_version_and_no_cpu_fallback():
    lib = capi.load()
    assert b"sm_100a" in lib.rp_version()
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present; the no-device path cannot be exercised")
    with pytest.raises(rp.RpError) as e:
        rp.Device(device=0)
    assert e.value.status == capi.RP_ERR_CUDA
    assert "no CPU fallback" in e.value.message


def test_product_path_does_not_touch_the_oracle():
    """rampack_b200/ must not import, link or load anything under oracle/ (oracle = checker only)."""
    pkg = os.path.join(ROOT, "rampack_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle" not in text.replace("test oracle", ""), f"{f} mentions oracle/"


def test_cycle_path_keeps_off_the_copy_engines():
    """The NpT cycle (sweep, grid builds, box move, host loop) must not issue small cudaMemcpyAsync readbacks or
    cudaMemsetAsync: they queue on the DMA engines behind the 97 MB transfers of other packings on the device and stall
    the cycle by 1-2 ms each (profiles/r02_e2e_pipelining.txt).  Results go through readbackSync, counters
    through zeroAsync, the saved state of a box move through copyAsync.  rp_sweep_trace (a test hook that downloads the
    whole trial record) is the one allowed exception."""
    csrc = os.path.join(ROOT, "rampack_b200", "csrc")
    for name in ("rp_grid.cu", "rp_sim.cu", "rp_stats.cu", "rp_sweep.cu", "rp_recount.cu"):
        text = open(os.path.join(csrc, name)).read()
        if name == "rp_sweep.cu":
            text = text[:text.index("sweepTrace(")] if "sweepTrace(" in text else text
        if name == "rp_recount.cu":
            # particle energy fluctuations and the pair-test batch move arrays whose size grows with N: bulk copies
            text = re.sub(r"cudaMemcpyAsync\((e\.data\(\)|dPairs|verdict \+ off|energy \+ off)[^;]*;", "", text)
        assert "cudaMemsetAsync" not in text, f"{name}: cudaMemsetAsync on the cycle path (use zeroAsync)"
        for m in re.finditer(r"cudaMemcpyAsync\(([^;]*);", text):
            assert False, f"{name}: cudaMemcpyAsync on the cycle path: {m.group(0)[:90]}"
