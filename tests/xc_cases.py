"""XenoCollide geometries used by the parity tests (seeded, shared by CPU and GPU tests)."""
from rampack_b200.xc import XCBodyBuilder, make_xc_desc


def _smooth_wedge_multi():
    """smooth_wedge(R=0.5, r=0.25, l=2, subdivisions=2): one CollideGeometry per interaction centre
    (SmoothWedgeTraits.cpp:36-64: alphas, r0/r1/l01 and centreZ exactly as written there)."""
    R, r, l, sub = 0.5, 0.25, 2.0, 2
    A = (l - r + R) / (l + r - R)
    alphas = [0.0]
    for _ in range(sub):
        alphas.append(A * alphas[-1] + 1)
    alphas = [a / alphas[-1] for a in alphas]
    beg_z = (-l + R - r) / 2
    b = XCBodyBuilder()
    centres, roots = [], []
    for i in range(sub):
        a0, a1 = alphas[i], alphas[i + 1]
        r0 = a0 * r + (1 - a0) * R
        r1 = a1 * r + (1 - a1) * R
        l01 = (a1 - a0) * l
        b.smooth_wedge(r0, r1, l01)
        roots.append(b.stack[-1].node)
        centres.append((0.0, 0.0, beg_z + l * (a0 + a1) / 2 - (r0 - r1) / 2))
    return make_xc_desc(b.release(roots), centres), 2


def _single(script=None, build=None):
    def make():
        if script is not None:
            return make_xc_desc(XCBodyBuilder.from_script(script)), 1
        b = XCBodyBuilder()
        build(b)
        return make_xc_desc(b.release()), 1
    return make


# name -> (factory returning (desc, n_centres), max pair distance for the random-pair sample)
CASES = {
    # docs/shapes.md:656-659: wrap of a sphere and a cuboid rotated by 45 degrees
    "wrap_sphere_cuboid": (_single("sphere 0.5 move 0 0 -0.5 cuboid 1 1 1 rot 0 0 45 move 0 0 0.5 wrap"), 3.2),
    "ellipsoid": (_single("ellipsoid 0.5 0.8 1.4"), 3.0),
    "football": (_single("football 3 0.5"), 3.2),
    "saucer": (_single("saucer 1.2 0.6"), 2.6),
    "cuboid": (_single("cuboid 1 1.5 2.5"), 3.4),
    "sum_disk_segment": (_single("disk 0.7 segment 1.5 rot 90 0 0 move 0.1 0 0 sum"), 3.4),      # a rounded slab
    "diff_ellipsoid_sphere": (_single("ellipsoid 1 0.6 0.8 sphere 0.2 move 0.1 0 0 diff"), 3.0),
    "wrap3_nested": (_single("point 0 0 1.2 rectangle 1 0.6 ellipse 0.9 0.4 rot 0 30 0 move 0 0 -0.8 wrap 3 "
                             "sphere 0.15 sum"), 3.6),
    "smooth_wedge": (_single(build=lambda b: b.smooth_wedge(0.5, 0.25, 2.0)), 3.0),
    "polyhedral_wedge": (_single(build=lambda b: b.polyhedral_wedge(1.0, 0.8, 0.3, 0.5, 2.0)), 2.8),
    "smooth_wedge_2centres": (_smooth_wedge_multi, 2.2),
}
