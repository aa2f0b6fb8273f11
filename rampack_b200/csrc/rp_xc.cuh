// rp_xc.cuh — XenoCollide (Minkowski Portal Refinement) overlap test for generic convex shapes on the device.
//
// Replaces, in the reference's operation order (every FP64 op a single rounded IEEE op, -fmad=false):
//   XenoCollide<G>::Intersect                       geometry/xenocollide/XenoCollide.h:59-168
//   XenoCollideTraits::overlapBetween               core/shapes/XenoCollideTraits.h:147-171
//   support functions of the primitives             geometry/xenocollide/XCPrimitives.cpp:34-163, XCPrimitives.h:44,145
//   XCSum / XCDiff / XCMax                          geometry/xenocollide/XCOperations.cpp:38-132
//   SmoothWedgeTraits::CollideGeometry              core/shapes/SmoothWedgeTraits.h:44-50
//   PolyhedralWedgeTraits::CollideGeometry          core/shapes/PolyhedralWedgeTraits.h:39-57
//   is_vector_zero                                  geometry/xenocollide/XCUtils.h:34-37
//
// The reference's geometry is a tree of heap objects behind virtual calls; here it is a flat "program" (nodes +
// child entries) evaluated by a depth-bounded template recursion: one non-inlined function per depth.
#pragma once

#include "rp_math.cuh"
#include "../../include/rampack_b200.h"

namespace rp {

struct XcNode {
    int kind, nEntries, firstEntry, pad;
    double p[6];
    double circum, in;       // getCircumsphereRadius / getInsphereRadius
    double centre[3];        // getCenter
};
struct XcEntry {
    int node, pad;
    double pos[3];
    M33 rot;
};
struct XcProgram {
    int nNodes, nEntries;
    XcNode nodes[RP_XC_MAX_NODES];
    XcEntry entries[RP_XC_MAX_ENTRIES];
    int root[RP_MAX_CENTRES];
};

constexpr int XC_MAX_DEPTH = 3;

// The geometry tree is evaluated by a depth-bounded recursion.  These functions must NOT be inlined: with
// __forceinline__ every level copies the whole primitive switch into each of its child call sites (code size
// grows like 5^depth per call site and the build does not finish).  One function per depth keeps it linear.
// `inline` here is the LINKAGE keyword (the header is included by several translation units), not an inlining hint
#define RP_HD_NOINLINE inline __host__ __device__ __noinline__

RP_HD bool xcIsZero(V3 v) {
    const double ZERO = 9.094947e-13;
    return fabs(v.x) < ZERO && fabs(v.y) < ZERO && fabs(v.z) < ZERO;
}
RP_HD V3 xcDiv(V3 v, double s) { return v3(v.x / s, v.y / s, v.z / s); }
RP_HD V3 xcNormalized(V3 v) { return xcDiv(v, sqrt(norm2(v))); }
// rot.transpose() * n (Matrix.tpp:104-113 on the materialised transpose)
RP_HD V3 mulT(const M33 &m, V3 v) {
    return v3((m.a[0] * v.x + m.a[3] * v.y) + m.a[6] * v.z,
              (m.a[1] * v.x + m.a[4] * v.y) + m.a[7] * v.z,
              (m.a[2] * v.x + m.a[5] * v.y) + m.a[8] * v.z);
}

template <int DEPTH>
RP_HD_NOINLINE V3 xcSupport(const XcProgram &P, int nodeIdx, V3 n);

// rot * geometry.getSupportPoint(rot.transpose() * n) + pos of a child entry
template <int DEPTH>
RP_HD_NOINLINE V3 xcEntrySupport(const XcProgram &P, const XcEntry &e, V3 n) {
    V3 s = xcSupport<DEPTH>(P, e.node, mulT(e.rot, n));
    return mul(e.rot, s) + v3(e.pos[0], e.pos[1], e.pos[2]);
}

template <int DEPTH>
RP_HD_NOINLINE V3 xcSupport(const XcProgram &P, int nodeIdx, V3 n) {
    const XcNode &N = P.nodes[nodeIdx];
    const double *p = N.p;
    switch (N.kind) {
        case RP_XC_POINT:
            return v3(p[0], p[1], p[2]);
        case RP_XC_SEGMENT:
            return v3(0, 0, n.z < 0 ? -p[0] : p[0]);
        case RP_XC_RECTANGLE:
            return v3(n.x < 0 ? -p[0] : p[0], n.y < 0 ? -p[1] : p[1], 0);
        case RP_XC_CUBOID:
            return v3(n.x < 0 ? -p[0] : p[0], n.y < 0 ? -p[1] : p[1], n.z < 0 ? -p[2] : p[2]);
        case RP_XC_DISK: {
            V3 n2 = v3(n.x, n.y, 0);
            if (xcIsZero(n2)) return v3(0, 0, 0);
            return p[0] * xcNormalized(n2);
        }
        case RP_XC_SPHERE:
            return p[0] * xcNormalized(n);
        case RP_XC_ELLIPSE: {
            V3 n2 = v3(p[0] * n.x, p[1] * n.y, 0 * n.z);
            if (xcIsZero(n2)) return v3(0, 0, 0);
            V3 u = xcNormalized(n2);
            return v3(u.x * p[0], u.y * p[1], u.z * 0);
        }
        case RP_XC_ELLIPSOID: {
            V3 n2 = xcNormalized(v3(n.x * p[0], n.y * p[1], n.z * p[2]));
            return v3(n2.x * p[0], n2.y * p[1], n2.z * p[2]);
        }
        case RP_XC_FOOTBALL: {
            // p[0] = halfLength, p[1] = radius; radius of curvature r2 = 0.5f * (h*h/r1 + r1)
            double r1 = p[1], h = p[0];
            double r2 = 0.5 * (h * h / r1 + r1);
            V3 n3 = xcNormalized(n);
            if (n3.z * r2 < -h) return v3(0, 0, -h);
            if (n3.z * r2 > h) return v3(0, 0, h);
            V3 n2 = xcNormalized(v3(n.x, n.y, 0));
            return (-n2) * (r2 - r1) + n3 * r2;
        }
        case RP_XC_SAUCER: {
            // p[0] = radius, p[1] = halfThickness
            double t = p[1], h = p[0];
            double r2 = 0.5 * (h * h / t + t);
            V3 n3 = xcNormalized(n);
            V3 n4 = v3(n.x, n.y, 0);
            if (!xcIsZero(n4)) {
                n4 = xcNormalized(n4);
                double threshold = n3.x * n3.x + n3.y * n3.y;
                if (threshold * r2 * r2 > h * h) return h * n4;
            }
            V3 n2 = v3(0, 0, 1);
            if (n.z < 0) n2 = -n2;
            return (-n2) * (r2 - t) + n3 * r2;
        }
        case RP_XC_SMOOTH_WEDGE: {
            // p = {R, r, l, Rminusr, Rpos, rpos}
            V3 nn = xcNormalized(n);
            if (p[3] > nn.z * p[2]) return p[0] * nn + v3(0, 0, p[4]);
            return p[1] * nn + v3(0, 0, p[5]);
        }
        case RP_XC_POLYHEDRAL_WEDGE: {
            // p = {vertexUp[3], vertexDown[3]}
            V3 up = v3(p[0], p[1], p[2]), down = v3(p[3], p[4], p[5]);
            if (n.x < 0) { up.x = -up.x; down.x = -down.x; }
            if (n.y < 0) { up.y = -up.y; down.y = -down.y; }
            return dot(up - down, n) > 0 ? up : down;
        }
        default:
            break;
    }
    if constexpr (DEPTH > 0) {
        const XcEntry *e = P.entries + N.firstEntry;
        if (N.kind == RP_XC_SUM) {
            V3 r = v3(0, 0, 0);
            for (int i = 0; i < N.nEntries; i++) r = r + xcEntrySupport<DEPTH - 1>(P, e[i], n);
            return r;
        }
        if (N.kind == RP_XC_DIFF) {
            V3 p1 = xcEntrySupport<DEPTH - 1>(P, e[0], n);
            // rot2 * geom2.getSupportPoint(rot2.transpose() * -n) + pos2
            V3 p2 = xcEntrySupport<DEPTH - 1>(P, e[1], -n);
            return p1 - p2;
        }
        if (N.kind == RP_XC_MAX) {
            V3 v = xcEntrySupport<DEPTH - 1>(P, e[0], n);
            double maxProjection = dot(v, n);
            for (int i = 1; i < N.nEntries; i++) {
                V3 v1 = xcEntrySupport<DEPTH - 1>(P, e[i], n);
                double projection = dot(v1, n);
                if (projection > maxProjection) { v = v1; maxProjection = projection; }
            }
            return v;
        }
    }
    return v3(0, 0, 0);
}

// XenoCollide::TransformSupportVert
RP_HD V3 xcWorldSupport(const XcProgram &P, int root, const M33 &rot, V3 pos, V3 n) {
    V3 local = xcSupport<XC_MAX_DEPTH>(P, root, mulT(rot, n));
    return mul(rot, local) + pos;
}

// XenoCollide<G>::Intersect; *supportCalls (optional) counts support-pair evaluations (roofline accounting)
RP_HD_NOINLINE bool xcIntersect(const XcProgram &P, int g1, const M33 &rot1, V3 pos1, int g2, const M33 &rot2, V3 pos2,
                       double boundaryTolerance, int *supportCalls)
{
    const XcNode &N1 = P.nodes[g1], &N2 = P.nodes[g2];
    V3 c1 = v3(N1.centre[0], N1.centre[1], N1.centre[2]), c2 = v3(N2.centre[0], N2.centre[1], N2.centre[2]);
    V3 v0 = (mul(rot2, c2) + pos2) - (mul(rot1, c1) + pos1);
    if (xcIsZero(v0)) return true;
    int calls = 0;
    auto support = [&](V3 n) {
        calls++;
        return xcWorldSupport(P, g2, rot2, pos2, n) - xcWorldSupport(P, g1, rot1, pos1, -n);
    };
    bool result = false;
    V3 n = -v0;
    V3 v1 = support(n);
    V3 v2, v3_;
    bool done = false;
    if (dot(v1, n) <= 0) { done = true; }
    if (!done) {
        n = cross(v1, v0);
        if (xcIsZero(n)) { result = true; done = true; }
    }
    if (!done) {
        v2 = support(n);
        if (dot(v2, n) <= 0) done = true;
    }
    if (!done) {
        n = cross(v1 - v0, v2 - v0);
        if (dot(n, v0) > 0) {
            V3 tmp = v1; v1 = v2; v2 = tmp;
            n = -n;
        }
        // phase one: find a valid portal; phase two: refine it (the iteration cap only guards against a hang the
        // reference would share; it is far above anything observed: tail at 17 steps)
        for (int guard = 0; guard < 100000 && !done; guard++) {
            v3_ = support(n);
            if (dot(v3_, n) <= 0) { done = true; break; }
            if (dot(cross(v1, v3_), v0) < 0) {
                v2 = v3_;
                n = cross(v1 - v0, v3_ - v0);
                continue;
            }
            if (dot(cross(v3_, v2), v0) < 0) {
                v1 = v3_;
                n = cross(v3_ - v0, v2 - v0);
                continue;
            }
            for (int guard2 = 0; guard2 < 100000; guard2++) {
                n = cross(v2 - v1, v3_ - v1);
                if (dot(n, v1) >= 0) { result = true; break; }
                V3 v4 = support(n);
                n = xcNormalized(n);
                if (-dot(v4, n) >= 0 || dot(v4 - v3_, n) <= boundaryTolerance) break;
                V3 cr = cross(v4, v0);
                if (dot(v1, cr) > 0) {
                    if (dot(v2, cr) > 0) v1 = v4;
                    else v3_ = v4;
                } else {
                    if (dot(v3_, cr) > 0) v2 = v4;
                    else v1 = v4;
                }
            }
            done = true;
        }
    }
    if (supportCalls) *supportCalls += calls;
    return result;
}

// XenoCollideTraits::overlapBetween with pos2bc = pos2 + translation already applied by the caller
RP_HD bool xcOverlap(const XcProgram &P, V3 pos1, const M33 &o1, int c1, V3 pos2bc, const M33 &o2, int c2,
                     int *supportCalls)
{
    const int g1 = P.root[c1], g2 = P.root[c2];
    double rCircumsphere = P.nodes[g1].circum + P.nodes[g2].circum;
    double rInsphere = P.nodes[g1].in + P.nodes[g2].in;
    double dist2 = norm2(pos2bc - pos1);
    if (dist2 > rCircumsphere * rCircumsphere) return false;
    if (dist2 < rInsphere * rInsphere) return true;
    return xcIntersect(P, g1, o1, pos1, g2, o2, pos2bc, 1.0e-12, supportCalls);
}

}  // namespace rp
