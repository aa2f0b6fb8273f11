"""TEST INFRASTRUCTURE ONLY — ctypes access to oracle/_ref/librampack_ref*.so (the UNMODIFIED reference classes behind
the extern "C" shim oracle/ref_harness.cpp) and to oracle/liboracle_port.so (the plain-C restatement).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module.
It deliberately does not import anything from rampack_b200/: the struct layouts are restated here from
include/rampack_b200.h.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
RP_MAX_CENTRES = 16


class InteractionDesc(C.Structure):
    _fields_ = [("hard_kind", C.c_int32), ("soft_kind", C.c_int32),
                ("radius", C.c_double), ("length", C.c_double), ("epsilon", C.c_double), ("sigma", C.c_double),
                ("n_centres", C.c_int32), ("reserved", C.c_int32),
                ("centres", (C.c_double * 3) * RP_MAX_CENTRES),
                ("centre_radius", C.c_double * RP_MAX_CENTRES),
                ("centre_half_axis", (C.c_double * 3) * RP_MAX_CENTRES),
                ("xc", C.c_void_p)]      # const rp_xc_program * (include/rampack_b200.h)


class Move(C.Structure):
    _fields_ = [("particle", C.c_uint64), ("kind", C.c_int32), ("reserved", C.c_int32),
                ("translation", C.c_double * 3), ("rotation", C.c_double * 9)]


class SimResult(C.Structure):
    _fields_ = [("move_seconds", C.c_double), ("scaling_seconds", C.c_double),
                ("domain_decomposition_seconds", C.c_double), ("total_seconds", C.c_double),
                ("moves_attempted", C.c_uint64), ("moves_accepted", C.c_uint64),
                ("scalings_attempted", C.c_uint64), ("scalings_accepted", C.c_uint64),
                ("density_mean", C.c_double), ("density_error", C.c_double), ("ng_rebuild_seconds", C.c_double),
                ("cycles", C.c_uint64), ("threads", C.c_int), ("domains", C.c_int * 3)]


PAIR_DTYPE = np.dtype([("pos1", "<f8", 3), ("orientation1", "<f8", 9), ("idx1", "<u8"),
                       ("pos2", "<f8", 3), ("orientation2", "<f8", 9), ("idx2", "<u8"),
                       ("translation", "<f8", 3)])


def copy_desc(src):
    """Copies any ctypes struct with the rp_interaction_desc layout into this module's InteractionDesc."""
    assert C.sizeof(src) == C.sizeof(InteractionDesc), "rp_interaction_desc layouts disagree"
    d = InteractionDesc()
    C.memmove(C.byref(d), C.byref(src), C.sizeof(InteractionDesc))
    # the desc only holds a POINTER to the XenoCollide program: keep the pointee alive with the copy
    d._xc_keepalive = getattr(src, "_xc_keepalive", None)
    return d


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def ref_path(native=False):
    return os.path.join(HERE, "_ref", "librampack_ref_native.so" if native else "librampack_ref.so")


def ensure_built():
    """Builds oracle artefacts that can be built where we are (the port always; _ref only where /root/reference is)."""
    subprocess.check_call(["make", "-C", HERE, "-j8", "all"], stdout=subprocess.DEVNULL)


_libs = {}


def load_ref(native=False):
    key = "native" if native else "strict"
    if key in _libs:
        return _libs[key]
    path = ref_path(native)
    if not os.path.exists(path):
        raise FileNotFoundError(f"{path} missing: run `make -C oracle ref` where /root/reference is mounted")
    lib = C.CDLL(path)
    V = C.c_void_p
    dp = C.POINTER(C.c_double)
    u64p = C.POINTER(C.c_uint64)
    lib.rpr_traits_create.argtypes = [C.POINTER(InteractionDesc)]
    lib.rpr_traits_create.restype = V
    lib.rpr_traits_destroy.argtypes = [V]
    lib.rpr_traits_destroy.restype = None
    lib.rpr_range_radius.argtypes = [V]
    lib.rpr_range_radius.restype = C.c_double
    lib.rpr_total_range_radius.argtypes = [V]
    lib.rpr_total_range_radius.restype = C.c_double
    lib.rpr_pair_test_batch.argtypes = [V, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.rpr_packing_create.argtypes = [V, dp, dp, C.c_uint64, C.c_int, C.c_int]
    lib.rpr_packing_create.restype = V
    lib.rpr_packing_ok.argtypes = [V]
    lib.rpr_packing_error.argtypes = [V]
    lib.rpr_packing_error.restype = C.c_char_p
    lib.rpr_packing_destroy.argtypes = [V]
    lib.rpr_packing_destroy.restype = None
    lib.rpr_packing_get.argtypes = [V, dp, dp]
    lib.rpr_packing_get.restype = None
    lib.rpr_grid_divisions.argtypes = [V, u64p]
    lib.rpr_average_neighbours.argtypes = [V]
    lib.rpr_average_neighbours.restype = C.c_double
    lib.rpr_count_total_overlaps.argtypes = [V, V, C.c_int]
    lib.rpr_count_total_overlaps.restype = C.c_uint64
    lib.rpr_total_energy.argtypes = [V, V]
    lib.rpr_total_energy.restype = C.c_double
    lib.rpr_try_move.argtypes = [V, V, C.POINTER(Move), dp]
    lib.rpr_accept_move.argtypes = [V]
    lib.rpr_accept_move.restype = None
    lib.rpr_try_scaling.argtypes = [V, V, dp, dp]
    lib.rpr_revert_scaling.argtypes = [V]
    lib.rpr_revert_scaling.restype = None
    lib.rpr_toggle_overlap_counting.argtypes = [V, V, C.c_int]
    lib.rpr_toggle_overlap_counting.restype = None
    lib.rpr_observables.argtypes = [V, V, C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int)]
    lib.rpr_toggle_walls.argtypes = [V, C.POINTER(C.c_int)]
    lib.rpr_toggle_walls.restype = None
    lib.rpr_cached_overlaps.argtypes = [V]
    lib.rpr_cached_overlaps.restype = C.c_uint64
    lib.rpr_cell_numbers.argtypes = [dp, C.c_double, dp, C.c_uint64, u64p, u64p]
    lib.rpr_pbc_correction.argtypes = [dp, dp, dp]
    lib.rpr_pbc_correction.restype = None
    lib.rpr_box_info.argtypes = [dp, dp, dp, dp]
    lib.rpr_box_info.restype = None
    lib.rpr_rotation_matrix.argtypes = [dp, C.c_double, dp]
    lib.rpr_rotation_matrix.restype = None
    lib.rpr_run_simulation.argtypes = [V, dp, dp, C.c_uint64, C.c_uint64, C.c_double, C.c_double, C.c_double,
                                       C.c_double, C.c_int, C.c_double, C.POINTER(C.c_int), C.c_int, C.c_uint64,
                                       C.c_uint64, C.c_uint64, dp, dp, C.POINTER(SimResult), C.c_char_p, C.c_size_t]
    lib.rpr_omp_max_threads.restype = C.c_int
    _libs[key] = lib
    return lib


class RefTraits:
    """A reference ShapeTraits/Interaction object built from an rp_interaction_desc."""

    def __init__(self, desc, native=False):
        self.lib = load_ref(native)
        self.desc = copy_desc(desc)
        self.ptr = self.lib.rpr_traits_create(C.byref(self.desc))
        if not self.ptr:
            raise ValueError("the reference rejected this interaction description")

    def __del__(self):
        if getattr(self, "ptr", None):
            self.lib.rpr_traits_destroy(self.ptr)
            self.ptr = None

    @property
    def range_radius(self):
        return self.lib.rpr_range_radius(self.ptr)

    @property
    def total_range_radius(self):
        return self.lib.rpr_total_range_radius(self.ptr)

    def pair_test_batch(self, pairs, want_verdict=True, want_energy=False):
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = pairs.shape[0]
        verdict = np.zeros(n, dtype=np.uint8) if want_verdict else None
        energy = np.zeros(n, dtype=np.float64) if want_energy else None
        self.lib.rpr_pair_test_batch(self.ptr, pairs.ctypes.data, n,
                                     verdict.ctypes.data if want_verdict else None,
                                     energy.ctypes.data if want_energy else None)
        return verdict, energy


class RefPacking:
    """The reference's Packing (PeriodicBoundaryConditions) driven through the shim."""

    def __init__(self, traits, box, shapes, move_threads=1, scaling_threads=1):
        self.lib = traits.lib
        self.traits = traits
        box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        shapes = np.ascontiguousarray(shapes, dtype=np.float64).reshape(-1, 12)
        self.n = shapes.shape[0]
        self.ptr = self.lib.rpr_packing_create(traits.ptr, _dp(box), _dp(shapes), self.n, move_threads,
                                               scaling_threads)
        if not self.lib.rpr_packing_ok(self.ptr):
            msg = self.lib.rpr_packing_error(self.ptr).decode()
            self.lib.rpr_packing_destroy(self.ptr)
            self.ptr = None
            raise ValueError(msg)

    def __del__(self):
        if getattr(self, "ptr", None):
            self.lib.rpr_packing_destroy(self.ptr)
            self.ptr = None

    def get(self):
        box = np.empty(9)
        shapes = np.empty((self.n, 12))
        self.lib.rpr_packing_get(self.ptr, _dp(box), _dp(shapes))
        return box.reshape(3, 3), shapes

    def grid_divisions(self):
        out = (C.c_uint64 * 3)()
        self.lib.rpr_grid_divisions(self.ptr, out)
        return tuple(out)

    def average_neighbours(self):
        return self.lib.rpr_average_neighbours(self.ptr)

    def count_total_overlaps(self, early_exit=False):
        return self.lib.rpr_count_total_overlaps(self.ptr, self.traits.ptr, int(early_exit))

    def total_energy(self):
        return self.lib.rpr_total_energy(self.ptr, self.traits.ptr)

    def try_move(self, particle, kind, translation=(0, 0, 0), rotation=None):
        m = Move()
        m.particle, m.kind = particle, kind
        for k in range(3):
            m.translation[k] = float(translation[k])
        rot = np.eye(3) if rotation is None else np.asarray(rotation, dtype=np.float64)
        for k in range(9):
            m.rotation[k] = float(rot.reshape(9)[k])
        v = C.c_double()
        if self.lib.rpr_try_move(self.ptr, self.traits.ptr, C.byref(m), C.byref(v)) != 0:
            raise ValueError(self.lib.rpr_packing_error(self.ptr).decode())
        return v.value

    def accept_move(self):
        self.lib.rpr_accept_move(self.ptr)

    def try_scaling(self, new_box):
        nb = np.ascontiguousarray(new_box, dtype=np.float64).reshape(9)
        v = C.c_double()
        if self.lib.rpr_try_scaling(self.ptr, self.traits.ptr, _dp(nb), C.byref(v)) != 0:
            raise ValueError(self.lib.rpr_packing_error(self.ptr).decode())
        return v.value

    def revert_scaling(self):
        self.lib.rpr_revert_scaling(self.ptr)

    def toggle_overlap_counting(self, on):
        self.lib.rpr_toggle_overlap_counting(self.ptr, self.traits.ptr, int(on))

    def observables(self, hkl_range=(5, 5, 5)):
        """The reference's NematicOrder / SmecticOrder: dict(p2, q (3x3), tau, k (3), hkl)."""
        out, hkl = (C.c_double * 11)(), (C.c_int * 3)()
        st = self.lib.rpr_observables(self.ptr, self.traits.ptr, (C.c_int * 3)(*hkl_range), out, hkl)
        if st != 0:
            raise RuntimeError("reference observables failed")
        o = list(out)
        q = np.array([[o[1], o[2], o[3]], [o[2], o[4], o[5]], [o[3], o[5], o[6]]])
        return {"p2": o[0], "q": q, "tau": o[7], "k": np.array(o[8:11]), "hkl": tuple(hkl)}

    def pair_density(self, max_r, n_bins):
        """(g, counts) of one PairDensityCorrelation / RadialEnumerator snapshot of the reference."""
        g, counts = np.zeros(n_bins), np.zeros(n_bins)
        self.lib.rpr_pair_density.restype = C.c_int
        rc = self.lib.rpr_pair_density(self.ptr, self.traits.ptr, C.c_double(max_r),
                                       C.c_uint64(n_bins), _dp(g), _dp(counts))
        if rc != 0:
            raise RuntimeError(f"rpr_pair_density failed ({rc})")
        return g, counts

    def toggle_walls(self, walls):
        self.lib.rpr_toggle_walls(self.ptr, (C.c_int * 3)(*[int(w) for w in walls]))

    def cached_overlaps(self):
        return self.lib.rpr_cached_overlaps(self.ptr)


def ref_cell_numbers(box, cell_size, positions, native=False):
    lib = load_ref(native)
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    positions = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
    out = np.empty(positions.shape[0], dtype=np.uint64)
    div = (C.c_uint64 * 3)()
    if lib.rpr_cell_numbers(_dp(box), cell_size, _dp(positions), positions.shape[0],
                            out.ctypes.data_as(C.POINTER(C.c_uint64)), div) != 0:
        raise ValueError("NeighbourGrid construction or positionToCellNo failed")
    return out, tuple(div)


def ref_box_info(box, native=False):
    lib = load_ref(native)
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    heights, vol, inv = np.empty(3), C.c_double(), np.empty(9)
    lib.rpr_box_info(_dp(box), _dp(heights), C.byref(vol), _dp(inv))
    return heights, vol.value, inv.reshape(3, 3)


def ref_pbc_correction(box, pos, native=False):
    lib = load_ref(native)
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    pos = np.ascontiguousarray(pos, dtype=np.float64).reshape(3)
    out = np.empty(3)
    lib.rpr_pbc_correction(_dp(box), _dp(pos), _dp(out))
    return out


def ref_rotation_matrix(axis, angle, native=False):
    lib = load_ref(native)
    axis = np.ascontiguousarray(axis, dtype=np.float64).reshape(3)
    out = np.empty(9)
    lib.rpr_rotation_matrix(_dp(axis), angle, _dp(out))
    return out.reshape(3, 3)


def ref_run_simulation(traits, box, shapes, seed, temperature, pressure, trans_step, rot_step, box_scaler, box_step,
                       domains, threads, therm_cycles, avg_cycles, avg_every=1):
    """Simulation::integrate of the reference (rototranslation sampler; box_scaler 0 none / 1 delta_v /
    2 delta_triclinic).  Returns (SimResult, final box, final shapes)."""
    lib = traits.lib
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    shapes = np.ascontiguousarray(shapes, dtype=np.float64).reshape(-1, 12)
    n = shapes.shape[0]
    res = SimResult()
    box_out, shapes_out = np.empty(9), np.empty((n, 12))
    dd = (C.c_int * 3)(*domains)
    err = C.create_string_buffer(1024)
    st = lib.rpr_run_simulation(traits.ptr, _dp(box), _dp(shapes), n, seed, temperature, pressure, trans_step,
                                rot_step, box_scaler, box_step, dd, threads, therm_cycles, avg_cycles, avg_every,
                                _dp(box_out), _dp(shapes_out), C.byref(res), err, 1024)
    if st != 0:
        raise RuntimeError("reference simulation failed: " + err.value.decode())
    return res, box_out.reshape(3, 3), shapes_out


def ref_propose_boxes(kind, step, independent, groups, seed, old_box, count, native=False):
    """`count` proposals of the reference's box scaler classes from std::mt19937(seed), each from old_box.
    kind: "delta_v" | "linear" | "log" | "delta_triclinic"; groups: axis bit masks (linear / log)."""
    lib = load_ref(native)
    code = {"delta_v": 1, "linear": 2, "log": 3, "delta_triclinic": 4}[kind]
    g = (C.c_int * max(1, len(groups)))(*groups) if groups else (C.c_int * 1)(7)
    old_box = np.ascontiguousarray(old_box, dtype=np.float64).reshape(9)
    out = np.empty((count, 9))
    err = C.create_string_buffer(1024)
    lib.rpr_propose_boxes.argtypes = [C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int), C.c_int, C.c_uint64,
                                      C.POINTER(C.c_double), C.c_int, C.POINTER(C.c_double), C.c_char_p, C.c_size_t]
    st = lib.rpr_propose_boxes(code, step, int(independent), g, len(groups) if groups else 0, seed, _dp(old_box),
                               count, _dp(out), err, 1024)
    if st != 0:
        raise RuntimeError("reference scaler failed: " + err.value.decode())
    return out.reshape(count, 3, 3)


def ref_populate_lattice(cell_box, molecules, dims, axis_order, n, native=False):
    """Lattice + SerialPopulator of the reference: (lattice box, n x 12 shapes)."""
    lib = load_ref(native)
    cell_box = np.ascontiguousarray(cell_box, dtype=np.float64).reshape(9)
    molecules = np.ascontiguousarray(molecules, dtype=np.float64).reshape(-1, 12)
    box, shapes = np.empty(9), np.empty((n, 12))
    d = (C.c_int * 3)(*[int(x) for x in dims])
    err = C.create_string_buffer(1024)
    lib.rpr_populate_lattice.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_uint64,
                                         C.POINTER(C.c_int), C.c_char_p, C.c_uint64, C.POINTER(C.c_double),
                                         C.POINTER(C.c_double), C.c_char_p, C.c_size_t]
    st = lib.rpr_populate_lattice(_dp(cell_box), _dp(molecules), molecules.shape[0], d, axis_order.encode(), n,
                                  _dp(box), _dp(shapes), err, 1024)
    if st != 0:
        raise RuntimeError("reference lattice failed: " + err.value.decode())
    return box.reshape(3, 3), shapes


# ---------------------------------------------------------------------------------------------------------------------
# the plain-C restatement (oracle/rp_oracle.c -> liboracle_port.so): same Python surface as RefTraits / RefPacking
# ---------------------------------------------------------------------------------------------------------------------
class PortRunResult(C.Structure):
    _fields_ = [("moves_attempted", C.c_uint64), ("moves_accepted", C.c_uint64), ("scalings_attempted", C.c_uint64),
                ("scalings_accepted", C.c_uint64), ("move_seconds", C.c_double), ("scaling_seconds", C.c_double)]


def port_path():
    return os.path.join(HERE, "liboracle_port.so")


def load_port():
    if "port" in _libs:
        return _libs["port"]
    path = port_path()
    if not os.path.exists(path):
        subprocess.check_call(["make", "-C", HERE, "port"], stdout=subprocess.DEVNULL)
    lib = C.CDLL(path)
    V = C.c_void_p
    dp = C.POINTER(C.c_double)
    u64p = C.POINTER(C.c_uint64)
    D = C.POINTER(InteractionDesc)
    lib.rpo_pair_test_batch.argtypes = [D, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.rpo_pair_test_batch_pbc.argtypes = [D, dp, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    lib.rpo_box_info.argtypes = [dp, dp, dp, dp]
    lib.rpo_box_info.restype = None
    lib.rpo_pbc_correction.argtypes = [dp, dp, dp]
    lib.rpo_pbc_correction.restype = None
    lib.rpo_cell_numbers.argtypes = [dp, C.c_double, dp, C.c_uint64, u64p, u64p]
    lib.rpo_rotation_matrix.argtypes = [dp, C.c_double, dp]
    lib.rpo_rotation_matrix.restype = None
    lib.rpo_range_radius.argtypes = [D]
    lib.rpo_range_radius.restype = C.c_double
    lib.rpo_total_range_radius.argtypes = [D]
    lib.rpo_total_range_radius.restype = C.c_double
    lib.rpo_packing_create.argtypes = [D, dp, dp, C.c_uint64, C.c_char_p, C.c_size_t]
    lib.rpo_packing_create.restype = V
    lib.rpo_packing_destroy.argtypes = [V]
    lib.rpo_packing_destroy.restype = None
    lib.rpo_packing_get.argtypes = [V, dp, dp]
    lib.rpo_packing_get.restype = None
    lib.rpo_grid_divisions.argtypes = [V, u64p]
    lib.rpo_average_neighbours.argtypes = [V]
    lib.rpo_average_neighbours.restype = C.c_double
    lib.rpo_count_total_overlaps.argtypes = [V, C.c_int]
    lib.rpo_count_total_overlaps.restype = C.c_uint64
    lib.rpo_total_energy.argtypes = [V]
    lib.rpo_total_energy.restype = C.c_double
    lib.rpo_try_move.argtypes = [V, C.POINTER(Move), dp]
    lib.rpo_accept_move.argtypes = [V]
    lib.rpo_accept_move.restype = None
    lib.rpo_try_scaling.argtypes = [V, dp, dp]
    lib.rpo_revert_scaling.argtypes = [V]
    lib.rpo_revert_scaling.restype = None
    lib.rpo_toggle_overlap_counting.argtypes = [V, C.c_int]
    lib.rpo_toggle_overlap_counting.restype = None
    lib.rpo_cached_overlaps.argtypes = [V]
    lib.rpo_cached_overlaps.restype = C.c_uint64
    lib.rpo_run_cycles.argtypes = [V, C.c_uint64, C.c_uint64, C.c_double, C.c_double, C.c_double, C.c_double,
                                   C.c_int, C.c_double, C.POINTER(PortRunResult)]
    _libs["port"] = lib
    return lib


class PortTraits:
    def __init__(self, desc):
        self.lib = load_port()
        self.desc = copy_desc(desc)

    @property
    def range_radius(self):
        return self.lib.rpo_range_radius(C.byref(self.desc))

    @property
    def total_range_radius(self):
        return self.lib.rpo_total_range_radius(C.byref(self.desc))

    def pair_test_batch(self, pairs, want_verdict=True, want_energy=False, pbc_box=None):
        pairs = np.ascontiguousarray(pairs, dtype=PAIR_DTYPE)
        n = pairs.shape[0]
        verdict = np.zeros(n, dtype=np.uint8) if want_verdict else None
        energy = np.zeros(n, dtype=np.float64) if want_energy else None
        vp = verdict.ctypes.data if want_verdict else None
        ep = energy.ctypes.data if want_energy else None
        if pbc_box is None:
            self.lib.rpo_pair_test_batch(C.byref(self.desc), pairs.ctypes.data, n, vp, ep)
        else:
            box = np.ascontiguousarray(pbc_box, dtype=np.float64).reshape(9)
            self.lib.rpo_pair_test_batch_pbc(C.byref(self.desc), _dp(box), pairs.ctypes.data, n, vp, ep)
        return verdict, energy


class PortPacking:
    def __init__(self, traits, box, shapes, move_threads=1, scaling_threads=1):
        self.lib = traits.lib
        self.traits = traits
        box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
        shapes = np.ascontiguousarray(shapes, dtype=np.float64).reshape(-1, 12)
        self.n = shapes.shape[0]
        err = C.create_string_buffer(256)
        self.ptr = self.lib.rpo_packing_create(C.byref(traits.desc), _dp(box), _dp(shapes), self.n, err, 256)
        if not self.ptr:
            raise ValueError(err.value.decode())

    def __del__(self):
        if getattr(self, "ptr", None):
            self.lib.rpo_packing_destroy(self.ptr)
            self.ptr = None

    def get(self):
        box = np.empty(9)
        shapes = np.empty((self.n, 12))
        self.lib.rpo_packing_get(self.ptr, _dp(box), _dp(shapes))
        return box.reshape(3, 3), shapes

    def grid_divisions(self):
        out = (C.c_uint64 * 3)()
        return tuple(out) if self.lib.rpo_grid_divisions(self.ptr, out) else None

    def average_neighbours(self):
        return self.lib.rpo_average_neighbours(self.ptr)

    def count_total_overlaps(self, early_exit=False):
        return self.lib.rpo_count_total_overlaps(self.ptr, int(early_exit))

    def total_energy(self):
        return self.lib.rpo_total_energy(self.ptr)

    def try_move(self, particle, kind, translation=(0, 0, 0), rotation=None):
        m = Move()
        m.particle, m.kind = particle, kind
        for k in range(3):
            m.translation[k] = float(translation[k])
        rot = np.eye(3) if rotation is None else np.asarray(rotation, dtype=np.float64)
        for k in range(9):
            m.rotation[k] = float(rot.reshape(9)[k])
        v = C.c_double()
        if self.lib.rpo_try_move(self.ptr, C.byref(m), C.byref(v)) != 0:
            raise ValueError("particleIdx < size()")
        return v.value

    def accept_move(self):
        self.lib.rpo_accept_move(self.ptr)

    def try_scaling(self, new_box):
        nb = np.ascontiguousarray(new_box, dtype=np.float64).reshape(9)
        v = C.c_double()
        if self.lib.rpo_try_scaling(self.ptr, _dp(nb), C.byref(v)) != 0:
            raise ValueError("newBox.getVolume() != 0")
        return v.value

    def revert_scaling(self):
        self.lib.rpo_revert_scaling(self.ptr)

    def toggle_overlap_counting(self, on):
        self.lib.rpo_toggle_overlap_counting(self.ptr, int(on))

    def cached_overlaps(self):
        return self.lib.rpo_cached_overlaps(self.ptr)

    def run_cycles(self, seed, cycles, trans_step, rot_step, temperature=1.0, pressure=1.0, box_moves=False,
                   box_step=0.01):
        res = PortRunResult()
        self.lib.rpo_run_cycles(self.ptr, seed, cycles, trans_step, rot_step, temperature, pressure, int(box_moves),
                                box_step, C.byref(res))
        return res


def port_cell_numbers(box, cell_size, positions):
    lib = load_port()
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    positions = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
    out = np.empty(positions.shape[0], dtype=np.uint64)
    div = (C.c_uint64 * 3)()
    if lib.rpo_cell_numbers(_dp(box), cell_size, _dp(positions), positions.shape[0],
                            out.ctypes.data_as(C.POINTER(C.c_uint64)), div) != 0:
        raise ValueError("NeighbourGrid construction or positionToCellNo failed")
    return out, tuple(div)


def port_box_info(box):
    lib = load_port()
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    heights, vol, inv = np.empty(3), C.c_double(), np.empty(9)
    lib.rpo_box_info(_dp(box), _dp(heights), C.byref(vol), _dp(inv))
    return heights, vol.value, inv.reshape(3, 3)


def port_pbc_correction(box, pos):
    lib = load_port()
    box = np.ascontiguousarray(box, dtype=np.float64).reshape(9)
    pos = np.ascontiguousarray(pos, dtype=np.float64).reshape(3)
    out = np.empty(3)
    lib.rpo_pbc_correction(_dp(box), _dp(pos), _dp(out))
    return out


def port_rotation_matrix(axis, angle):
    lib = load_port()
    axis = np.ascontiguousarray(axis, dtype=np.float64).reshape(3)
    out = np.empty(9)
    lib.rpo_rotation_matrix(_dp(axis), angle, _dp(out))
    return out.reshape(3, 3)


def have_ref(native=False):
    return os.path.exists(ref_path(native))
