"""ctypes binding of the C ABI in include/rampack_b200.h (librampack_b200.so).

This is plumbing for tests, bench.py and the Python mirror of the reference's `Packing` (rampack_b200/packing.py).
There is no CPU fallback: if the shared library is missing it is built with nvcc, and if that fails the import
raises.
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

RP_MAX_CENTRES = 16

RP_OK, RP_ERR_PRECONDITION, RP_ERR_CUDA, RP_ERR_STATE, RP_ERR_UNSUPPORTED, RP_ERR_CAPACITY = range(6)
(RP_HARD_NONE, RP_HARD_SPHERE, RP_HARD_SPHEROCYLINDER, RP_HARD_POLYSPHERE, RP_HARD_POLYSPHEROCYLINDER,
 RP_HARD_XENOCOLLIDE) = range(6)
# rp_xc_kind
(RP_XC_POINT, RP_XC_SEGMENT, RP_XC_RECTANGLE, RP_XC_CUBOID, RP_XC_DISK, RP_XC_SPHERE, RP_XC_ELLIPSE, RP_XC_ELLIPSOID,
 RP_XC_FOOTBALL, RP_XC_SAUCER, RP_XC_SMOOTH_WEDGE, RP_XC_POLYHEDRAL_WEDGE) = range(12)
RP_XC_SUM, RP_XC_DIFF, RP_XC_MAX = 16, 17, 18
RP_XC_MAX_NODES = RP_XC_MAX_ENTRIES = 32
RP_SOFT_NONE, RP_SOFT_LJ, RP_SOFT_WCA, RP_SOFT_SIC = range(4)
RP_MOVE_TRANSLATION, RP_MOVE_ROTATION, RP_MOVE_ROTOTRANSLATION, RP_MOVE_FLIP = range(4)


class XcNode(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_entries", C.c_int32), ("first_entry", C.c_int32), ("reserved", C.c_int32),
                ("params", C.c_double * 6)]


class XcEntry(C.Structure):
    _fields_ = [("node", C.c_int32), ("reserved", C.c_int32), ("pos", C.c_double * 3), ("rot", C.c_double * 9)]


class XcProgram(C.Structure):
    _fields_ = [("n_nodes", C.c_int32), ("n_entries", C.c_int32), ("nodes", XcNode * RP_XC_MAX_NODES),
                ("entries", XcEntry * RP_XC_MAX_ENTRIES), ("root", C.c_int32 * RP_MAX_CENTRES)]


class InteractionDesc(C.Structure):
    _fields_ = [("hard_kind", C.c_int32), ("soft_kind", C.c_int32),
                ("radius", C.c_double), ("length", C.c_double), ("epsilon", C.c_double), ("sigma", C.c_double),
                ("n_centres", C.c_int32), ("reserved", C.c_int32),
                ("centres", (C.c_double * 3) * RP_MAX_CENTRES),
                ("centre_radius", C.c_double * RP_MAX_CENTRES),
                ("centre_half_axis", (C.c_double * 3) * RP_MAX_CENTRES),
                ("xc", C.POINTER(XcProgram))]


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("reserved", C.c_int32), ("seed", C.c_uint64), ("cuda_stream", C.c_void_p)]


class GridInfo(C.Structure):
    _fields_ = [("has_grid", C.c_int32), ("cell_divisions", C.c_int32 * 3), ("cell_size", C.c_double),
                ("num_centres", C.c_uint64), ("max_cell_population", C.c_uint32), ("reserved", C.c_uint32)]


class Move(C.Structure):
    _fields_ = [("particle", C.c_uint64), ("kind", C.c_int32), ("reserved", C.c_int32),
                ("translation", C.c_double * 3), ("rotation", C.c_double * 9)]


class SweepParams(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pass_", C.c_int32), ("translation_step", C.c_double),
                ("rotation_step", C.c_double), ("temperature", C.c_double), ("cycle", C.c_uint64),
                ("flip_axis", C.c_double * 3), ("flip_origin", C.c_double * 3), ("flip_every", C.c_uint64)]


class SweepCounters(C.Structure):
    _fields_ = [("attempted", C.c_uint64), ("accepted", C.c_uint64), ("rejected_domain", C.c_uint64),
                ("rejected_overlap", C.c_uint64), ("rejected_energy", C.c_uint64), ("sub_sweeps", C.c_uint64), ("kernel_launches", C.c_uint64)]


RP_SCALER_NONE, RP_SCALER_DELTA_V, RP_SCALER_LINEAR, RP_SCALER_LOG, RP_SCALER_TRICLINIC_DELTA = range(5)
RP_SIM_MAX_SAMPLERS = 4


class SimSampler(C.Structure):
    _fields_ = [("kind", C.c_int32), ("reserved", C.c_int32), ("translation_step", C.c_double),
                ("rotation_step", C.c_double), ("flip_axis", C.c_double * 3), ("flip_origin", C.c_double * 3),
                ("flip_every", C.c_uint64)]


class SimConfig(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("temperature", C.c_double), ("pressure", C.c_double),
                ("n_samplers", C.c_int32), ("scaler_kind", C.c_int32), ("samplers", SimSampler * RP_SIM_MAX_SAMPLERS),
                ("scaler_step", C.c_double), ("n_direction_groups", C.c_int32), ("direction_groups", C.c_int32 * 3),
                ("independent", C.c_int32), ("rotation_matrix_fix_every", C.c_int32)]


class SimStats(C.Structure):
    _fields_ = [("cycles", C.c_uint64), ("moves_attempted", C.c_uint64 * RP_SIM_MAX_SAMPLERS),
                ("moves_accepted", C.c_uint64 * RP_SIM_MAX_SAMPLERS), ("scalings_attempted", C.c_uint64),
                ("scalings_evaluated", C.c_uint64), ("scalings_accepted", C.c_uint64),
                ("translation_step", C.c_double * RP_SIM_MAX_SAMPLERS),
                ("rotation_step", C.c_double * RP_SIM_MAX_SAMPLERS), ("scaler_step", C.c_double),
                ("box", C.c_double * 9), ("sweep_kernel_ms", C.c_double), ("sweep_grid_ms", C.c_double),
                ("scaling_ms", C.c_double), ("cells_skipped", C.c_uint64)]


# rp_pair as a numpy structured dtype (same layout as the C struct: all members are 8 bytes wide)
PAIR_DTYPE = np.dtype([("pos1", "<f8", 3), ("orientation1", "<f8", 9), ("idx1", "<u8"),
                       ("pos2", "<f8", 3), ("orientation2", "<f8", 9), ("idx2", "<u8"),
                       ("translation", "<f8", 3)])
assert PAIR_DTYPE.itemsize == 8 * 29
# rp_trial_trace
TRACE_DTYPE = np.dtype([("particle", "<u8"), ("colour_step", "<i4"), ("cell", "<i4"), ("order_key", "<u4"),
                        ("slot", "<i4"), ("status", "<i4"), ("batch", "<i4"), ("position", "<f8", 3),
                        ("orientation", "<f8", 9), ("image_shift", "<f8", 3), ("u", "<f8"), ("dE", "<f8")])
assert TRACE_DTYPE.itemsize == 8 + 6 * 4 + 8 * 17

# every symbol include/rampack_b200.h declares (tests/test_abi.py checks the exports against the header)
SYMBOLS = ["rp_populate_lattice", "rp_randomize_rotations", "rp_create", "rp_destroy", "rp_last_error", "rp_version", "rp_set_interaction", "rp_upload",
           "rp_download", "rp_size", "rp_rebuild_grid", "rp_grid_info_get", "rp_cell_numbers",
           "rp_count_total_overlaps", "rp_total_energy", "rp_try_move", "rp_accept_move", "rp_try_scaling",
           "rp_revert_scaling", "rp_sweep", "rp_sweep_trace", "rp_toggle_overlap_counting", "rp_cached_overlaps",
           "rp_pair_test_batch", "rp_comm_unique_id", "rp_comm_init", "rp_comm_destroy", "rp_comm_stats", "rp_shard_plan",
           "rp_neighbour_stats", "rp_measure_fp64_peak", "rp_last_kernel_ms", "rp_launch_count", "rp_synchronize",
           "rp_renormalize_orientations", "rp_toggle_walls", "rp_nematic_order", "rp_smectic_order",
           "rp_pair_density_correlation", "rp_sim_create", "rp_sim_destroy", "rp_sim_run", "rp_sim_stats_get", "rp_sim_propose_box",
           "rp_download_range", "rp_count_wall_overlaps", "rp_particle_energy_fluctuations", "rp_snapshot_async",
           "rp_snapshot_wait", "rp_snapshot_done", "rp_upload_slice", "rp_nematic_order_static"]

_lib = None


def library_path():
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "librampack_b200.so")


def load():
    """Loads (building first if needed) librampack_b200.so and declares the prototypes."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if not os.path.exists(path):
        _build.build()
    lib = C.CDLL(path)
    H = C.c_void_p
    dp = C.POINTER(C.c_double)
    lib.rp_create.argtypes = [C.POINTER(Config), C.POINTER(H)]
    lib.rp_destroy.argtypes = [H]
    lib.rp_destroy.restype = None
    lib.rp_last_error.argtypes = [H]
    lib.rp_last_error.restype = C.c_char_p
    lib.rp_version.restype = C.c_char_p
    lib.rp_toggle_walls.argtypes = [H, C.POINTER(C.c_int32)]
    lib.rp_nematic_order.argtypes = [H, dp, dp, dp]
    lib.rp_nematic_order_static.argtypes = [H, dp, dp, dp]
    lib.rp_smectic_order.argtypes = [H, C.POINTER(C.c_int32), dp, dp, C.POINTER(C.c_int32), dp]
    lib.rp_pair_density_correlation.argtypes = [H, C.c_double, C.c_int32, dp, C.c_void_p, dp]
    lib.rp_renormalize_orientations.argtypes = [H, C.c_int, C.POINTER(C.c_uint64)]
    lib.rp_sim_create.argtypes = [H, C.POINTER(SimConfig), C.POINTER(H)]
    lib.rp_sim_destroy.argtypes = [H]
    lib.rp_sim_destroy.restype = None
    lib.rp_sim_run.argtypes = [H, C.c_uint64, C.c_int]
    lib.rp_sim_stats_get.argtypes = [H, C.POINTER(SimStats)]
    lib.rp_sim_propose_box.argtypes = [H, dp, dp]
    lib.rp_set_interaction.argtypes = [H, C.POINTER(InteractionDesc)]
    lib.rp_upload.argtypes = [H, dp, dp, C.c_uint64]
    lib.rp_populate_lattice.argtypes = [H, dp, dp, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_int), C.c_uint64]
    lib.rp_randomize_rotations.argtypes = [H, C.c_uint64, C.c_int, dp]
    lib.rp_download.argtypes = [H, dp, dp, C.c_uint64]
    lib.rp_size.argtypes = [H, C.POINTER(C.c_uint64)]
    lib.rp_download_range.argtypes = [H, C.c_uint64, C.c_uint64, dp]
    lib.rp_snapshot_async.argtypes = [H, dp, dp, C.c_uint64]
    lib.rp_upload_slice.argtypes = [H, dp, dp, C.c_uint64, C.c_uint64, C.c_uint64]
    lib.rp_snapshot_wait.argtypes = [H]
    lib.rp_snapshot_done.argtypes = [H, C.POINTER(C.c_int32)]
    lib.rp_count_wall_overlaps.argtypes = [H, C.c_int, C.POINTER(C.c_uint64)]
    lib.rp_particle_energy_fluctuations.argtypes = [H, dp]
    lib.rp_rebuild_grid.argtypes = [H]
    lib.rp_grid_info_get.argtypes = [H, C.POINTER(GridInfo)]
    lib.rp_cell_numbers.argtypes = [H, C.POINTER(C.c_uint64)]
    lib.rp_count_total_overlaps.argtypes = [H, C.c_int, C.POINTER(C.c_uint64)]
    lib.rp_total_energy.argtypes = [H, dp]
    lib.rp_try_move.argtypes = [H, C.POINTER(Move), dp]
    lib.rp_accept_move.argtypes = [H]
    lib.rp_try_scaling.argtypes = [H, dp, dp]
    lib.rp_revert_scaling.argtypes = [H]
    lib.rp_sweep.argtypes = [H, C.POINTER(SweepParams), C.POINTER(SweepCounters)]
    lib.rp_sweep_trace.argtypes = [H, C.c_void_p, C.c_uint64]
    lib.rp_toggle_overlap_counting.argtypes = [H, C.c_int]
    lib.rp_comm_unique_id.argtypes = [C.c_void_p]
    lib.rp_comm_init.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32]
    lib.rp_comm_destroy.argtypes = [H]
    lib.rp_shard_plan.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
    lib.rp_comm_stats.argtypes = [H, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.rp_cached_overlaps.argtypes = [H, C.POINTER(C.c_uint64)]
    lib.rp_pair_test_batch.argtypes = [H, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.rp_last_kernel_ms.argtypes = [H, dp, dp, dp, dp]
    lib.rp_neighbour_stats.argtypes = [H, dp, dp]
    lib.rp_measure_fp64_peak.argtypes = [H, dp, dp]
    lib.rp_launch_count.argtypes = [H, C.POINTER(C.c_uint64)]
    lib.rp_synchronize.argtypes = [H]
    for name in SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("rp_destroy", "rp_last_error", "rp_version"):
            fn.restype = C.c_int
    _lib = lib
    return lib


def shard_plan(n_layers_z, n_ranks, rank, colour):
    """rp_shard_plan as a dict (host logic of the sharded sweep; needs no device)."""
    lib = load()
    out = (C.c_int32 * 6)()
    st = lib.rp_shard_plan(n_layers_z, n_ranks, rank, colour, out)
    if st != RP_OK:
        raise RpError(st, lib.rp_last_error(None).decode())
    keys = ("a_lo", "a_hi", "send_layer", "recv_layer", "peer_send", "peer_recv")
    return dict(zip(keys, out))


class RpError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"rp_status {status}: {message}")
        self.status = status
        self.message = message


class PreconditionError(RpError):
    """The reference's PreconditionException (src/utils/Exceptions.h:19-66)."""


def dptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def make_desc(hard=RP_HARD_NONE, soft=RP_SOFT_NONE, radius=0.0, length=0.0, epsilon=0.0, sigma=0.0, centres=None,
              centre_radius=None, centre_half_axis=None, xc=None):
    d = InteractionDesc()
    if xc is not None:
        d.xc = C.pointer(xc)
        d._xc_keepalive = xc        # the desc only holds a pointer: keep the program alive with it
    d.hard_kind, d.soft_kind = hard, soft
    d.radius, d.length, d.epsilon, d.sigma = radius, length, epsilon, sigma
    centres = [] if centres is None else centres
    d.n_centres = len(centres)
    for i, c in enumerate(centres):
        for k in range(3):
            d.centres[i][k] = float(c[k])
        d.centre_radius[i] = float(centre_radius[i]) if centre_radius is not None else 0.0
        if centre_half_axis is not None:
            for k in range(3):
                d.centre_half_axis[i][k] = float(centre_half_axis[i][k])
    return d


class Device:
    """One rp_handle.  Methods are 1:1 with the C entry points; arrays are numpy float64."""

    def __init__(self, device=0, seed=1234, cuda_stream=None):
        self.lib = load()
        self.h = C.c_void_p()
        cfg = Config(device, 0, seed, cuda_stream)
        st = self.lib.rp_create(C.byref(cfg), C.byref(self.h))
        if st != RP_OK:
            raise RpError(st, self.lib.rp_last_error(None).decode())
        self.n = 0

    def close(self):
        if self.h:
            self.lib.rp_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != RP_OK:
            msg = self.lib.rp_last_error(self.h).decode()
            raise (PreconditionError if st == RP_ERR_PRECONDITION else RpError)(st, msg)

    def set_interaction(self, desc):
        self._check(self.lib.rp_set_interaction(self.h, C.byref(desc)))

    def upload(self, box, shapes):
        box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        shapes = np.ascontiguousarray(shapes, dtype=np.float64).reshape(-1, 12)
        self._check(self.lib.rp_upload(self.h, dptr(box), dptr(shapes), shapes.shape[0]))
        self.n = shapes.shape[0]

    def populate_lattice(self, cell_box, cell_molecules, n_cells, axis_order="xyz", n_shapes=None):
        """Lattice(UnitCell(cell_box, molecules), n_cells) + SerialPopulator(axis_order) on the device
        (rp_populate_lattice); cell_molecules: m x 12 (relative position in the cell, orientation)."""
        cell_box = np.ascontiguousarray(cell_box, dtype=np.float64).reshape(9)
        mol = np.ascontiguousarray(cell_molecules, dtype=np.float64).reshape(-1, 12)
        dims = (C.c_uint32 * 3)(*[int(x) for x in n_cells])
        order = (C.c_int * 3)(*["xyz".index(ch) for ch in axis_order])
        if n_shapes is None:
            n_shapes = int(n_cells[0]) * int(n_cells[1]) * int(n_cells[2]) * mol.shape[0]
        self._check(self.lib.rp_populate_lattice(self.h, dptr(cell_box), dptr(mol), mol.shape[0], dims, order,
                                                 n_shapes))
        self.n = n_shapes

    def randomize_rotations(self, seed, mode=0, axis=None):
        """RotationRandomizingTransformer on the device: mode 0 random, 1 about `axis`, 2 about orientation * `axis`."""
        a = np.ascontiguousarray(axis if axis is not None else (0.0, 0.0, 1.0), dtype=np.float64).reshape(3)
        self._check(self.lib.rp_randomize_rotations(self.h, seed, mode, dptr(a)))

    @staticmethod
    def slice_of(n, n_ranks, rank):
        """(first, count) of a rank's slice in the standard partition rp_upload_slice expects."""
        a, b = n * rank // n_ranks, n * (rank + 1) // n_ranks
        return a, b - a

    def upload_slice(self, box, shapes_slice, first, n):
        """Collective sharded upload: this rank's rows [first, first + len(shapes_slice)) of the n-particle packing."""
        box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        shapes_slice = np.ascontiguousarray(shapes_slice, dtype=np.float64).reshape(-1, 12)
        self._check(self.lib.rp_upload_slice(self.h, dptr(box), dptr(shapes_slice), first, shapes_slice.shape[0], n))
        self.n = n

    def download(self, out=None):
        box = np.empty(9)
        shapes = out if out is not None else np.empty((self.n, 12))
        self._check(self.lib.rp_download(self.h, dptr(box), dptr(shapes), self.n))
        return box.reshape(3, 3), shapes

    def download_range(self, first, count, out=None):
        out = np.empty((count, 12), dtype=np.float64) if out is None else out
        self._check(self.lib.rp_download_range(self.h, first, count, out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def snapshot_async(self, out):
        """Start an asynchronous snapshot into `out` (N x 12 float64, ideally pinned host memory); returns the box.
        `out` is valid after snapshot_wait() (or once snapshot_done() is True); later cycles may be queued meanwhile."""
        assert out.dtype == np.float64 and out.flags["C_CONTIGUOUS"] and out.size == self.n * 12
        box = np.empty(9)
        self._check(self.lib.rp_snapshot_async(self.h, dptr(box), dptr(out), self.n))
        return box.reshape(3, 3)

    def snapshot_wait(self):
        self._check(self.lib.rp_snapshot_wait(self.h))

    def snapshot_done(self):
        v = C.c_int32()
        self._check(self.lib.rp_snapshot_done(self.h, C.byref(v)))
        return bool(v.value)

    def count_wall_overlaps(self, early_exit=False):
        c = C.c_uint64(0)
        self._check(self.lib.rp_count_wall_overlaps(self.h, 1 if early_exit else 0, C.byref(c)))
        return int(c.value)

    def particle_energy_fluctuations(self):
        v = C.c_double(0)
        self._check(self.lib.rp_particle_energy_fluctuations(self.h, C.byref(v)))
        return float(v.value)

    def rebuild_grid(self):
        self._check(self.lib.rp_rebuild_grid(self.h))

    def grid_info(self):
        gi = GridInfo()
        self._check(self.lib.rp_grid_info_get(self.h, C.byref(gi)))
        return gi

    def cell_numbers(self):
        gi = self.grid_info()
        out = np.empty(gi.num_centres, dtype=np.uint64)
        self._check(self.lib.rp_cell_numbers(self.h, out.ctypes.data_as(C.POINTER(C.c_uint64))))
        return out

    def count_total_overlaps(self, early_exit=False):
        v = C.c_uint64()
        self._check(self.lib.rp_count_total_overlaps(self.h, int(early_exit), C.byref(v)))
        return v.value

    def total_energy(self):
        v = C.c_double()
        self._check(self.lib.rp_total_energy(self.h, C.byref(v)))
        return v.value

    def try_move(self, particle, kind, translation=(0, 0, 0), rotation=None):
        m = Move()
        m.particle, m.kind = particle, kind
        for k in range(3):
            m.translation[k] = float(translation[k])
        rot = np.eye(3) if rotation is None else np.asarray(rotation, dtype=np.float64)
        for k in range(9):
            m.rotation[k] = float(rot.reshape(9)[k])
        v = C.c_double()
        self._check(self.lib.rp_try_move(self.h, C.byref(m), C.byref(v)))
        return v.value

    def accept_move(self):
        self._check(self.lib.rp_accept_move(self.h))

    def try_scaling(self, new_box):
        nb = np.ascontiguousarray(new_box, dtype=np.float64).reshape(9)
        v = C.c_double()
        self._check(self.lib.rp_try_scaling(self.h, dptr(nb), C.byref(v)))
        return v.value

    def revert_scaling(self):
        self._check(self.lib.rp_revert_scaling(self.h))

    def sweep(self, kind, translation_step, rotation_step, cycle, temperature=1.0, pass_=0, flip_axis=(0, 0, 0),
              flip_origin=(0, 0, 0), flip_every=0):
        p = SweepParams(kind, pass_, translation_step, rotation_step, temperature, cycle)
        p.flip_axis[:] = flip_axis
        p.flip_origin[:] = flip_origin
        p.flip_every = flip_every
        c = SweepCounters()
        self._check(self.lib.rp_sweep(self.h, C.byref(p), C.byref(c)))
        return c

    def sweep_trace(self):
        """The trials of the last sweep() call, one TRACE_DTYPE entry per particle (parity harness)."""
        out = np.zeros(self.n, dtype=TRACE_DTYPE)
        self._check(self.lib.rp_sweep_trace(self.h, out.ctypes.data, out.shape[0]))
        return out

    @staticmethod
    def comm_unique_id():
        """128 bytes (ncclUniqueId) created on rank 0; distribute them to all ranks, then comm_init()."""
        lib = load()
        buf = (C.c_uint8 * 128)()
        st = lib.rp_comm_unique_id(buf)
        if st != RP_OK:
            raise RpError(st, lib.rp_last_error(None).decode())
        return bytes(buf)

    def comm_init(self, unique_id, rank, n_ranks):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(self.lib.rp_comm_init(self.h, buf, rank, n_ranks))

    def comm_destroy(self):
        self._check(self.lib.rp_comm_destroy(self.h))

    def comm_stats(self):
        a, b = C.c_uint64(), C.c_uint64()
        self._check(self.lib.rp_comm_stats(self.h, C.byref(a), C.byref(b)))
        return {"halo_bytes": a.value, "gather_bytes": b.value}

    def nematic_order(self, primary_axis=(0.0, 0.0, 1.0)):
        """(Q tensor 3x3, P2) — NematicOrder of the resident packing."""
        axis = np.ascontiguousarray(primary_axis, dtype=np.float64)
        q, p2 = np.empty(9), C.c_double()
        self._check(self.lib.rp_nematic_order(self.h, dptr(axis), dptr(q), C.byref(p2)))
        return q.reshape(3, 3), p2.value

    def nematic_order_static(self, primary_axis=(0.0, 0.0, 1.0)):
        """nematic_order through rp_nematic_order_static: no device allocation, at most
        RP_NEMATIC_STATIC_MAX_PARTICLES (4 194 304) particles, calls on one device must not overlap."""
        axis = np.ascontiguousarray(primary_axis, dtype=np.float64)
        q, p2 = np.empty(9), C.c_double()
        self._check(self.lib.rp_nematic_order_static(self.h, dptr(axis), dptr(q), C.byref(p2)))
        return q.reshape(3, 3), p2.value

    def smectic_order(self, hkl_range=(5, 5, 5), focal_point=None):
        """(|tau|, (h, k, l), k vector) — SmecticOrder of the resident packing."""
        rng = (C.c_int32 * 3)(*hkl_range)
        focal = np.ascontiguousarray(focal_point if focal_point is not None else (0, 0, 0), dtype=np.float64)
        tau, hkl, k = C.c_double(), (C.c_int32 * 3)(), np.empty(3)
        self._check(self.lib.rp_smectic_order(self.h, rng, dptr(focal), C.byref(tau), hkl, dptr(k)))
        return tau.value, tuple(hkl), k

    def pair_density_correlation(self, max_r, n_bins, focal_point=None):
        """(counts, g) — one snapshot of PairDensityCorrelation / RadialEnumerator of the resident packing."""
        focal = np.ascontiguousarray(focal_point if focal_point is not None else (0, 0, 0), dtype=np.float64)
        counts, g = np.zeros(n_bins, dtype=np.uint64), np.zeros(n_bins)
        self._check(self.lib.rp_pair_density_correlation(self.h, C.c_double(max_r), C.c_int32(n_bins), dptr(focal),
                                                         counts.ctypes.data_as(C.c_void_p), dptr(g)))
        return counts, g

    def toggle_walls(self, walls):
        self._check(self.lib.rp_toggle_walls(self.h, (C.c_int32 * 3)(*[int(w) for w in walls])))

    def renormalize_orientations(self, allow_overlaps=False):
        rej = C.c_uint64()
        self._check(self.lib.rp_renormalize_orientations(self.h, int(allow_overlaps), C.byref(rej)))
        return rej.value

    def toggle_overlap_counting(self, on):
        self._check(self.lib.rp_toggle_overlap_counting(self.h, int(on)))

    def cached_overlaps(self):
        v = C.c_uint64()
        self._check(self.lib.rp_cached_overlaps(self.h, C.byref(v)))
        return v.value

    def pair_test_batch(self, pairs, want_verdict=True, want_energy=False):
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = pairs.shape[0]
        verdict = np.zeros(n, dtype=np.uint8) if want_verdict else None
        energy = np.zeros(n, dtype=np.float64) if want_energy else None
        self._check(self.lib.rp_pair_test_batch(self.h, pairs.ctypes.data, n,
                                                verdict.ctypes.data if want_verdict else None,
                                                energy.ctypes.data if want_energy else None))
        return verdict, energy

    def last_kernel_ms(self):
        a, b, c, d = C.c_double(), C.c_double(), C.c_double(), C.c_double()
        self._check(self.lib.rp_last_kernel_ms(self.h, C.byref(a), C.byref(b), C.byref(c), C.byref(d)))
        return {"sweep": a.value, "sweep_grid": b.value, "rebuild": c.value, "scaling": d.value}

    def neighbour_stats(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.rp_neighbour_stats(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def measure_fp64_peak(self):
        a, b = C.c_double(), C.c_double()
        self._check(self.lib.rp_measure_fp64_peak(self.h, C.byref(a), C.byref(b)))
        return {"dfma_gflops": a.value, "dmul_dadd_gflops": b.value}

    def launch_count(self):
        v = C.c_uint64()
        self._check(self.lib.rp_launch_count(self.h, C.byref(v)))
        return v.value

    def synchronize(self):
        self._check(self.lib.rp_synchronize(self.h))


class NativeSimulation:
    """rp_sim: the reference's Simulation cycle loop (performCycle / tryScaling / evaluateCounters,
    src/core/Simulation.cpp:325-363, 493-607) running in C++ inside the library.

    samplers: list of (rp_move_kind, translation_step, rotation_step); scaler: None, ("delta_v", step),
    ("linear" | "log", step, groups, independent) with groups = list of axis bit masks (1 = x, 2 = y, 4 = z), or
    ("delta_triclinic", step, independent).  `device` may be None for the host-only parts (propose_box)."""

    def __init__(self, device, samplers, temperature=1.0, pressure=1.0, scaler=None, seed=1234,
                 rotation_matrix_fix_every=0):
        self.lib = load()
        self.device = device
        cfg = SimConfig()
        cfg.rotation_matrix_fix_every = rotation_matrix_fix_every
        cfg.seed, cfg.temperature, cfg.pressure = seed, temperature, pressure
        cfg.n_samplers = len(samplers)
        for i, smp in enumerate(samplers):
            kind, ts, rs = smp[:3]
            cfg.samplers[i].kind, cfg.samplers[i].translation_step, cfg.samplers[i].rotation_step = kind, ts, rs
            if kind == RP_MOVE_FLIP:      # (RP_MOVE_FLIP, 0, 0, flip_every, flip_axis, flip_origin)
                cfg.samplers[i].flip_every = smp[3]
                cfg.samplers[i].flip_axis[:] = smp[4]
                cfg.samplers[i].flip_origin[:] = smp[5] if len(smp) > 5 else (0, 0, 0)
        if scaler is None:
            cfg.scaler_kind = RP_SCALER_NONE
        else:
            name = scaler[0]
            cfg.scaler_kind = {"delta_v": RP_SCALER_DELTA_V, "linear": RP_SCALER_LINEAR, "log": RP_SCALER_LOG,
                               "delta_triclinic": RP_SCALER_TRICLINIC_DELTA}[name]
            cfg.scaler_step = scaler[1]
            if name in ("linear", "log"):
                groups = list(scaler[2]) if len(scaler) > 2 else [7]
                cfg.n_direction_groups = len(groups)
                for i, g in enumerate(groups):
                    cfg.direction_groups[i] = g
                cfg.independent = int(scaler[3]) if len(scaler) > 3 else 0
            elif name == "delta_triclinic":
                cfg.independent = int(scaler[2]) if len(scaler) > 2 else 1
        self.s = C.c_void_p()
        st = self.lib.rp_sim_create(device.h if device is not None else None, C.byref(cfg), C.byref(self.s))
        if st != RP_OK:
            raise (PreconditionError if st == RP_ERR_PRECONDITION else RpError)(st, self.lib.rp_last_error(None).decode())

    def close(self):
        if self.s:
            self.lib.rp_sim_destroy(self.s)
            self.s = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def run(self, cycles, adjust_step_sizes=False):
        st = self.lib.rp_sim_run(self.s, int(cycles), int(adjust_step_sizes))
        if st != RP_OK:
            self.device._check(st)

    def stats(self):
        out = SimStats()
        self.lib.rp_sim_stats_get(self.s, C.byref(out))
        return out

    def propose_box(self, box):
        box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        out = np.empty(9)
        st = self.lib.rp_sim_propose_box(self.s, dptr(box), dptr(out))
        if st != RP_OK:
            raise RpError(st, self.lib.rp_last_error(None).decode())
        return out.reshape(3, 3)
