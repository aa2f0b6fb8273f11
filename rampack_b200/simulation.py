"""Host-side cycle loop over the C ABI: the reference's Simulation::performCycle (src/core/Simulation.cpp:325-363)
with the particle moves delegated to rp_sweep (one bulk call per sampler and cycle) and the box move to
rp_try_scaling / rp_revert_scaling.  Host logic only (one box proposal and one Metropolis test per cycle); every
particle-level operation runs on the device.
"""
import math

import numpy as np

from . import capi


class TriclinicDeltaScaler:
    """TriclinicDeltaScaler::updateBox (src/core/volume_scalers/TriclinicDeltaScaler.cpp:7-25)."""

    def __init__(self, step, independent=False):
        self.step = step
        self.independent = independent

    def update_box(self, box, rng):
        dim = np.array(box, dtype=np.float64).reshape(3, 3).copy()
        if not self.independent:
            dim += rng.uniform(-self.step, self.step, size=(3, 3))
        else:
            side = int(rng.integers(0, 3))
            dim[:, side] += rng.uniform(-self.step, self.step, size=3)
        return dim


class DeltaVolumeScaler:
    """TriclinicAdapter(DeltaVolumeScaler) (src/core/volume_scalers/DeltaVolumeScaler.cpp, TriclinicAdapter.h:27-33):
    isotropic scaling by cbrt(1 + dV / V), dV uniform in (-step, step)."""

    def __init__(self, step):
        self.step = step

    def update_box(self, box, rng):
        dim = np.array(box, dtype=np.float64).reshape(3, 3)
        v = abs(np.linalg.det(dim))
        dv = rng.uniform(-self.step, self.step)
        return dim * np.cbrt((v + dv) / v) if v + dv > 0 else dim.copy()


class Counter:
    def __init__(self):
        self.attempted = 0
        self.accepted = 0


class Simulation:
    """NpT / NVT cycle loop.  `samplers` is a list of (rp_move_kind, translation_step, rotation_step)."""

    def __init__(self, device, box, samplers, temperature=1.0, pressure=1.0, box_scaler=None, seed=1234,
                 soft=False):
        self.dev = device
        self.box = np.array(box, dtype=np.float64).reshape(3, 3)
        self.samplers = samplers
        self.temperature = temperature
        self.pressure = pressure
        self.box_scaler = box_scaler
        self.rng = np.random.default_rng(seed)      # box proposals and their Metropolis draws (mts[0] in the reference)
        self.soft = soft
        self.cycle = 0
        self.moves = Counter()
        self.scalings = Counter()
        self.scalings_evaluated = 0
        self.sweep_kernel_ms = 0.0
        self.sweep_grid_ms = 0.0
        self.scaling_ms = 0.0

    def _heights(self):
        b = self.box
        s1, s2, s3 = b[:, 0], b[:, 1], b[:, 2]
        vol = abs(np.dot(np.cross(s1, s2), s3))
        return [vol / np.linalg.norm(np.cross(s2, s3)), vol / np.linalg.norm(np.cross(s3, s1)),
                vol / np.linalg.norm(np.cross(s1, s2))]

    def perform_moves(self):
        # MoveSampler step clamps: translation <= min height / 2 (TranslationSampler.cpp:20-37), rotation <= pi
        max_t = min(self._heights()) / 2
        for p, (kind, ts, rs) in enumerate(self.samplers):
            c = self.dev.sweep(kind, min(ts, max_t), min(rs, math.pi), self.cycle, self.temperature, pass_=p)
            self.moves.attempted += c.attempted
            self.moves.accepted += c.accepted
            k = self.dev.last_kernel_ms()
            self.sweep_kernel_ms += k["sweep"]
            self.sweep_grid_ms += k["sweep_grid"]

    def try_scaling(self):
        """Simulation::tryScaling (src/core/Simulation.cpp:493-531)."""
        old = self.box
        new = self.box_scaler.update_box(old, self.rng)
        old_v, new_v = abs(np.linalg.det(old)), abs(np.linalg.det(new))
        if new_v == 0:
            raise capi.PreconditionError(capi.RP_ERR_PRECONDITION, "newBox.getVolume() != 0")
        delta_v, factor, n = new_v - old_v, new_v / old_v, float(self.dev.n)
        t, p = self.temperature, self.pressure
        if self.soft:
            dE = self.dev.try_scaling(new)
            self.scalings_evaluated += 1
            self.scaling_ms += self.dev.last_kernel_ms()["scaling"]
            exponent = n * math.log(factor) - dE / t - p * delta_v / t
            if self.rng.uniform() <= math.exp(min(exponent, 700)):
                self.box = new
                return True
            self.dev.revert_scaling()
            return False
        exponent = n * math.log(factor) - p * delta_v / t
        if self.rng.uniform() > math.exp(min(exponent, 700)):
            return False
        dE = self.dev.try_scaling(new)
        self.scalings_evaluated += 1
        self.scaling_ms += self.dev.last_kernel_ms()["scaling"]
        if dE > 0:
            self.dev.revert_scaling()
            return False
        self.box = new
        return True

    def perform_cycle(self):
        self.perform_moves()
        if self.box_scaler is not None:
            self.scalings.attempted += 1
            if self.try_scaling():
                self.scalings.accepted += 1
        self.cycle += 1
