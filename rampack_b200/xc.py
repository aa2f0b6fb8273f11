"""Host mirror of the reference's XCBodyBuilder (src/geometry/xenocollide/XCBodyBuilder.cpp:50-294): the same stack
machine and command names, emitting the flat rp_xc_program of include/rampack_b200.h instead of a tree of heap objects.

    prog = XCBodyBuilder.from_script("sphere 0.5 move 0 0 -0.5 cuboid 1 1 1 rot 0 0 45 move 0 0 0.5 wrap")

Semantics kept from the reference:
  * every stack element is {geometry, pos, orientation}; `move` adds to pos, `rot x y z` (degrees) left-multiplies
    the orientation by Matrix<3,3>::rotation(ax, ay, az) (src/geometry/Matrix.tpp:411-434);
  * `sum n` / `wrap n` pop n elements from the BACK of the stack, so entry 0 is the former top of the stack
    (XCBodyBuilder.cpp:185-196, 207-218); `diff` pops shape2 then shape1 and builds XCDiff(shape1, shape2) (:76-86);
  * the finished geometry is the one remaining stack element's geometry; its own pos / orientation are dropped
    (releaseCollideGeometry, :37-48).

What the tests establish: the device and the reference harness are fed the SAME program, so verdict parity proves
device MPR == reference MPR on identical geometry.  It does not by itself prove this builder emits bit-identical
rotation entries to the C++ XCBodyBuilder for a given script (sin/cos of libm vs numpy may differ in the last ulp).
"""
import math

import numpy as np

from . import capi


class _Shape:
    def __init__(self, node):
        self.node = node
        self.pos = np.zeros(3)
        self.orientation = np.eye(3)

    def copy(self):
        s = _Shape(self.node)
        s.pos = self.pos.copy()
        s.orientation = self.orientation.copy()
        return s


def rotation_matrix(ax, ay, az):
    """Matrix<3,3>::rotation(ax, ay, az), radians, the reference's expressions (Matrix.tpp:411-434)."""
    sx, sy, sz = math.sin(ax), math.sin(ay), math.sin(az)
    cx, cy, cz = math.cos(ax), math.cos(ay), math.cos(az)
    return np.array([[cy * cz, sx * sy * cz - cx * sz, cx * sy * cz + sx * sz],
                     [cy * sz, sx * sy * sz + cx * cz, cx * sy * sz - sx * cz],
                     [-sy, sx * cy, cx * cy]])


class XCBodyBuilder:
    def __init__(self):
        self.nodes = []       # (kind, params, entries) with entries = [(node, pos, rot)]
        self.stack = []

    # ---- primitives (argument conventions of XCBodyBuilder.cpp: full lengths in, half lengths stored) ----
    def _push(self, kind, params):
        self.nodes.append((kind, list(params), []))
        self.stack.append(_Shape(len(self.nodes) - 1))

    def point(self, x, y, z): self._push(capi.RP_XC_POINT, [x, y, z])
    def segment(self, length): self._push(capi.RP_XC_SEGMENT, [length / 2])
    def rectangle(self, sx, sy): self._push(capi.RP_XC_RECTANGLE, [sx / 2, sy / 2])
    def cuboid(self, sx, sy, sz): self._push(capi.RP_XC_CUBOID, [sx / 2, sy / 2, sz / 2])
    def disk(self, radius): self._push(capi.RP_XC_DISK, [radius])
    def sphere(self, radius): self._push(capi.RP_XC_SPHERE, [radius])
    def ellipse(self, ax, ay): self._push(capi.RP_XC_ELLIPSE, [ax, ay])
    def ellipsoid(self, ax, ay, az): self._push(capi.RP_XC_ELLIPSOID, [ax, ay, az])
    def football(self, length, radius): self._push(capi.RP_XC_FOOTBALL, [length / 2, radius])
    def saucer(self, radius, thickness): self._push(capi.RP_XC_SAUCER, [radius, thickness / 2])
    # not XCBodyBuilder commands: the collide geometries of smooth_wedge / polyhedral_wedge
    def smooth_wedge(self, R, r, l): self._push(capi.RP_XC_SMOOTH_WEDGE, [R, r, l])
    def polyhedral_wedge(self, ax_bottom, ay_bottom, ax_top, ay_top, length):
        self._push(capi.RP_XC_POLYHEDRAL_WEDGE, [ax_bottom, ay_bottom, ax_top, ay_top, length])

    # ---- stack manipulation ----
    def move(self, x, y, z):
        self._need(1, "move")
        self.stack[-1].pos = self.stack[-1].pos + np.array([x, y, z], dtype=np.float64)

    def rot(self, x, y, z):
        self._need(1, "rotate")
        r = rotation_matrix(x * math.pi / 180, y * math.pi / 180, z * math.pi / 180)
        self.stack[-1].orientation = r @ self.stack[-1].orientation

    def dup(self, n=1):
        self._need(n, "dup")
        self.stack.extend(s.copy() for s in self.stack[-n:])

    def swap(self):
        self._need(2, "swap")
        self.stack[-1], self.stack[-2] = self.stack[-2], self.stack[-1]

    def pop(self):
        self._need(1, "pop")
        self.stack.pop()

    def _need(self, n, what):
        if len(self.stack) < n:
            raise ValueError(f"XCBodyBuilder: shape stack contains < {n} shapes; cannot {what}")

    def _operation(self, kind, shapes):
        entries = [(s.node, s.pos.copy(), s.orientation.copy()) for s in shapes]
        self.nodes.append((kind, [0.0] * 6, entries))
        self.stack.append(_Shape(len(self.nodes) - 1))

    def sum(self, n=2):
        self._need(max(n, 2), "compute Minkowski sum")
        self._operation(capi.RP_XC_SUM, [self.stack.pop() for _ in range(n)])

    def wrap(self, n=2):
        self._need(max(n, 2), "compute convex hull")
        self._operation(capi.RP_XC_MAX, [self.stack.pop() for _ in range(n)])

    def diff(self):
        self._need(2, "compute Minkowski difference")
        shape2 = self.stack.pop()
        shape1 = self.stack.pop()
        self._operation(capi.RP_XC_DIFF, [shape1, shape2])

    # ---- script front end (XCBodyBuilder::processCommand, :220-294) ----
    def process(self, script):
        tokens = script.split()
        arity = {"point": 3, "segment": 1, "rectangle": 2, "cuboid": 3, "disk": 1, "sphere": 1, "ellipse": 2,
                 "ellipsoid": 3, "football": 2, "saucer": 2, "move": 3, "rot": 3, "swap": 0, "pop": 0, "diff": 0}
        i = 0
        while i < len(tokens):
            cmd = tokens[i]
            i += 1
            if cmd in ("sum", "wrap", "dup"):
                n = 2 if cmd != "dup" else 1
                if i < len(tokens) and tokens[i].isdigit():
                    n = int(tokens[i])
                    i += 1
                getattr(self, cmd)(n)
            elif cmd in arity:
                args = [float(t) for t in tokens[i:i + arity[cmd]]]
                if len(args) != arity[cmd]:
                    raise ValueError(f"Malformed command arguments: {cmd}")
                i += arity[cmd]
                getattr(self, cmd)(*args)
            else:
                raise ValueError(f"Unknown XCBodyBuilder command: {cmd}")
        return self

    @classmethod
    def from_script(cls, script):
        return cls().process(script).release()

    def release(self, roots=None):
        """The rp_xc_program.  roots: node per interaction centre; default: the single remaining stack element."""
        if roots is None:
            if len(self.stack) != 1:
                raise ValueError("XCBodyBuilder: shape stack must hold exactly 1 shape; did you forget sum, diff or wrap?")
            roots = [self.stack[-1].node]
        return make_program(self.nodes, roots)


def make_program(nodes, roots):
    p = capi.XcProgram()
    if len(nodes) > capi.RP_XC_MAX_NODES:
        raise ValueError("geometry has more than RP_XC_MAX_NODES nodes")
    p.n_nodes = len(nodes)
    e = 0
    for i, (kind, params, entries) in enumerate(nodes):
        n = p.nodes[i]
        n.kind, n.n_entries, n.first_entry = kind, len(entries), e
        for q, v in enumerate(params[:6]):
            n.params[q] = float(v)
        for child, pos, rot in entries:
            if e >= capi.RP_XC_MAX_ENTRIES:
                raise ValueError("geometry has more than RP_XC_MAX_ENTRIES child entries")
            ent = p.entries[e]
            ent.node = child
            for d in range(3):
                ent.pos[d] = float(pos[d])
            for q in range(9):
                ent.rot[q] = float(np.asarray(rot).reshape(9)[q])
            e += 1
    p.n_entries = e
    for c, r in enumerate(roots):
        p.root[c] = r
    return p


def make_xc_desc(program, centres=None):
    """rp_interaction_desc for RP_HARD_XENOCOLLIDE (one geometry, or one per interaction centre)."""
    return capi.make_desc(hard=capi.RP_HARD_XENOCOLLIDE, centres=centres,
                          centre_radius=[0.0] * len(centres) if centres else None, xc=program)
