"""The drop-in at work (SURVEY §8b', north_star: `rampack casino -i file.pyon` on the device).

host/_build holds two link variants of the SAME reference object files (host/Makefile):
  *_cpu   with the reference's own src/core/Packing.cpp
  *_b200  with host/Packing_b200.cpp (the reference's Packing interface over the C ABI) + the Simulation hooks
The binaries are built in the container where /root/reference is mounted and travel to the GPU box prebuilt.
"""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "host", "_build")
pytestmark = pytest.mark.gpu


def _need(*names):
    for n in names:
        if not os.path.exists(os.path.join(BUILD, n)):
            pytest.skip(f"host/_build/{n} is not built (needs /root/reference at build time)")


def _lines(binary):
    out = subprocess.run([os.path.join(BUILD, binary)], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    res = {}
    for line in out.stdout.splitlines():
        key, _, value = line.partition(" ")
        res[key] = value
    return res


def test_packing_interface_matches_the_reference_packing():
    """host/facade_check.cpp: explicit moves / accepts, scaling + revert, overlap counting, walls, energies,
    renormalisation, store / restore through the reference's Packing interface — both link variants must agree."""
    _need("facade_check_cpu", "facade_check_b200")
    cpu, dev = _lines("facade_check_cpu"), _lines("facade_check_b200")
    assert "EXCEPTION" not in dev, dev.get("EXCEPTION")
    assert cpu.keys() == dev.keys()
    energy_keys = ("energy", "dE", "fluctuations")
    bad = []
    for key, want in cpu.items():
        got = dev[key]
        if any(e in key for e in energy_keys):
            w, g = float(want), float(got)
            if abs(w - g) > 1e-10 * max(1.0, abs(w)):
                bad.append((key, want, got))
        elif "checksum" in key or "first_x" in key or "last_z" in key or "volume" in key or "fraction" in key:
            # positions after identical accepted moves: the arithmetic of Shape::translate / rotate and of the
            # rescale is the reference's, operation by operation
            if float(want) != float(got):
                bad.append((key, want, got))
        elif got != want:
            bad.append((key, want, got))
    assert not bad, bad[:20]


def _run_casino(binary, workdir):
    os.makedirs(workdir, exist_ok=True)
    pyon = os.path.join(ROOT, "tests", "data", "facade_integration.pyon")
    out = subprocess.run([os.path.join(BUILD, binary), "casino", "-i", pyon], cwd=workdir, capture_output=True,
                         text=True, timeout=1200)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    header, values = open(os.path.join(workdir, "out.txt")).read().split("\n")[:2]
    res = dict(zip(header.split(), map(float, values.split())))
    res["log"] = out.stdout
    return res


def test_rampack_casino_runs_on_the_device(tmp_path):
    """`rampack casino -i integration.pyon` (the reference's sample input with trimmed cycle counts): the reference's
    front end, Simulation, observables and RAMSNAP writer on top of the device Packing against the unmodified CPU
    program.  The two are different Markov chains of the same ensemble: densities agree within the errors."""
    _need("rampack_cpu", "rampack_b200")
    cpu = _run_casino("rampack_cpu", str(tmp_path / "cpu"))
    dev = _run_casino("rampack_b200", str(tmp_path / "b200"))
    print(f"\n  rho: device {dev['rho']:.6g} +- {dev['drho']:.2g}, reference {cpu['rho']:.6g} +- {cpu['drho']:.2g}")
    print(f"  Z:   device {dev['Z']:.6g} +- {dev['dZ']:.2g}, reference {cpu['Z']:.6g} +- {cpu['dZ']:.2g}")
    # the reference's own error is the naive standard error of 120 correlated samples (utils/Quantity.cpp:55-60);
    # block averaging of the same series gives ~2x that, so the bar is 2 combined standard errors with that factor
    sigma = 2.0 * (cpu["drho"] ** 2 + dev["drho"] ** 2) ** 0.5
    assert abs(cpu["rho"] - dev["rho"]) < 2 * sigma
    assert os.path.exists(tmp_path / "b200" / "packing_gas.ramsnap")
    assert "Integration completed" in dev["log"]
