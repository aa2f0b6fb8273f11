// rp_xc_setup.cu — host side of the XenoCollide program: validation and the derived quantities the reference
// computes in the constructors of its geometry classes (circumsphere / insphere radius, centre), with the same
// expressions so that the pre-filter thresholds of XenoCollideTraits::overlapBetween compare identical doubles.
//   geometry/xenocollide/XCPrimitives.h:34-237, XCPrimitives.cpp:34-118
//   geometry/xenocollide/XCOperations.cpp:50-73 (XCSum), 75-99 (XCDiff), 101-185 (XCMax)
//   core/shapes/SmoothWedgeTraits.cpp:98-106, core/shapes/PolyhedralWedgeTraits.cpp:62-78
#include "rp_internal.cuh"

namespace rp {

static double vnorm(V3 v) { return std::sqrt(norm2(v)); }
static V3 vec(const double *p) { return v3(p[0], p[1], p[2]); }

// XCMax::calculateInsphereRadius2 (XCOperations.cpp:155-185)
static double maxInsphere2(V3 pos1, double insphere1, V3 pos2, double insphere2) {
    V3 diff = pos2 - pos1;
    double diffNorm = vnorm(diff);
    V3 normalDiff = v3(diff.x / diffNorm, diff.y / diffNorm, diff.z / diffNorm);
    if (diffNorm / std::max(insphere1, insphere2) < 1e-14)
        return std::min(insphere1, insphere2);
    double originT = -dot(pos1, normalDiff) / diffNorm;
    if (originT < 0) return 0;
    if (originT > 1) return 0;
    double proj = dot(pos1, normalDiff);
    double originOffset = std::sqrt(std::max(0.0, norm2(pos1) - proj * proj));
    double interpolatedR = (1 - originT) * insphere1 + originT * insphere2;
    return std::max(0.0, interpolatedR - originOffset);
}

void buildXcProgram(const rp_xc_program &src, int nRoots, XcProgram &dst) {
    std::memset(&dst, 0, sizeof(dst));
    RP_EXPECT(src.n_nodes >= 1 && src.n_nodes <= RP_XC_MAX_NODES, "1 <= xc.n_nodes <= RP_XC_MAX_NODES");
    RP_EXPECT(src.n_entries >= 0 && src.n_entries <= RP_XC_MAX_ENTRIES, "0 <= xc.n_entries <= RP_XC_MAX_ENTRIES");
    dst.nNodes = src.n_nodes;
    dst.nEntries = src.n_entries;
    for (int e = 0; e < src.n_entries; e++) {
        XcEntry &E = dst.entries[e];
        E.node = src.entries[e].node;
        for (int d = 0; d < 3; d++) E.pos[d] = src.entries[e].pos[d];
        for (int q = 0; q < 9; q++) E.rot.a[q] = src.entries[e].rot[q];
    }
    int depth[RP_XC_MAX_NODES];
    for (int i = 0; i < src.n_nodes; i++) {
        const rp_xc_node &S = src.nodes[i];
        XcNode &N = dst.nodes[i];
        N.kind = S.kind;
        N.nEntries = 0;
        N.firstEntry = 0;
        const double *a = S.params;
        for (int q = 0; q < 6; q++) N.p[q] = a[q];
        depth[i] = 0;
        switch (S.kind) {
            case RP_XC_POINT:
                N.circum = 0; N.in = 0;
                for (int d = 0; d < 3; d++) N.centre[d] = a[d];
                break;
            case RP_XC_SEGMENT:
                RP_EXPECT(a[0] > 0, "halfLength > 0");
                N.circum = a[0]; N.in = 0;
                break;
            case RP_XC_RECTANGLE:
                RP_EXPECT(a[0] > 0 && a[1] > 0, "halfSideX > 0 && halfSideY > 0");
                N.circum = vnorm(v3(a[0], a[1], 0)); N.in = 0;
                break;
            case RP_XC_CUBOID:
                RP_EXPECT(a[0] > 0 && a[1] > 0 && a[2] > 0, "cuboid half sides > 0");
                N.circum = vnorm(v3(a[0], a[1], a[2]));
                N.in = std::min(a[0], std::min(a[1], a[2]));
                break;
            case RP_XC_DISK:
                RP_EXPECT(a[0] > 0, "radius > 0");
                N.circum = a[0]; N.in = 0;
                break;
            case RP_XC_SPHERE:
                RP_EXPECT(a[0] > 0, "radius > 0");
                N.circum = a[0]; N.in = a[0];
                break;
            case RP_XC_ELLIPSE:
                RP_EXPECT(a[0] > 0 && a[1] > 0, "semiAxisX > 0 && semiAxisY > 0");
                N.circum = std::max(a[0], std::max(a[1], 0.0)); N.in = 0;
                break;
            case RP_XC_ELLIPSOID:
                RP_EXPECT(a[0] > 0 && a[1] > 0 && a[2] > 0, "ellipsoid semi-axes > 0");
                N.circum = std::max(a[0], std::max(a[1], a[2]));
                N.in = std::min(a[0], std::min(a[1], a[2]));
                break;
            case RP_XC_FOOTBALL:
                RP_EXPECT(a[1] > 0 && a[0] >= a[1], "radius > 0 && halfLength >= radius");
                N.circum = a[0]; N.in = a[1];
                break;
            case RP_XC_SAUCER:
                RP_EXPECT(a[0] > 0 && a[1] <= a[0], "radius > 0 && halfThickness <= radius");
                N.circum = a[0]; N.in = a[1];
                break;
            case RP_XC_SMOOTH_WEDGE: {
                double R = a[0], r = a[1], l = a[2];
                RP_EXPECT(R > 0 && r > 0 && l > 0, "R > 0 && r > 0 && l > 0");
                double Rminusr = R - r;
                N.p[3] = Rminusr;
                N.p[4] = (-l + R - r) / 2;
                N.p[5] = (l + R - r) / 2;
                N.circum = (l + R + r) / 2;
                N.in = 0.5 * (R + r + Rminusr * Rminusr / l);
                break;
            }
            case RP_XC_POLYHEDRAL_WEDGE: {
                double axBottom = a[0], ayBottom = a[1], axTop = a[2], ayTop = a[3], length = a[4];
                RP_EXPECT((axTop > 0 && ayBottom > 0) || (ayTop > 0 && axBottom > 0), "wedge must not be flat");
                RP_EXPECT(length > 0, "length > 0");
                V3 up = v3(axTop / 2, ayTop / 2, length / 2), down = v3(axBottom / 2, ayBottom / 2, -length / 2);
                N.p[0] = up.x; N.p[1] = up.y; N.p[2] = up.z; N.p[3] = down.x; N.p[4] = down.y; N.p[5] = down.z;
                N.circum = std::max(vnorm(up), vnorm(down));
                double irUp = std::min(axTop, ayTop), irDown = std::min(axBottom, ayBottom);
                N.in = std::min(std::min(irUp, irDown), length) / 2;
                break;
            }
            case RP_XC_SUM:
            case RP_XC_DIFF:
            case RP_XC_MAX: {
                int ne = S.n_entries, fe = S.first_entry;
                RP_EXPECT(ne >= (S.kind == RP_XC_DIFF ? 2 : 1) && (S.kind != RP_XC_DIFF || ne == 2),
                          "operation entry count");
                RP_EXPECT(fe >= 0 && fe + ne <= src.n_entries, "operation entries out of range");
                N.nEntries = ne;
                N.firstEntry = fe;
                const XcEntry *E = dst.entries + fe;
                for (int e = 0; e < ne; e++) {
                    RP_EXPECT(E[e].node >= 0 && E[e].node < i, "xc entries must refer to earlier nodes");
                    depth[i] = std::max(depth[i], depth[E[e].node] + 1);
                }
                RP_EXPECT(depth[i] <= XC_MAX_DEPTH, "xc geometry nests deeper than XC_MAX_DEPTH operations");
                auto childCentre = [&](const XcEntry &e) {
                    return mul(e.rot, vec(dst.nodes[e.node].centre)) + vec(e.pos);
                };
                if (S.kind == RP_XC_SUM) {
                    V3 centerSum = v3(0, 0, 0), c = v3(0, 0, 0);
                    double circumSum = 0, inSum = 0;
                    for (int e = 0; e < ne; e++) {
                        centerSum = centerSum + vec(E[e].pos);
                        circumSum += dst.nodes[E[e].node].circum;
                        inSum += dst.nodes[E[e].node].in;
                        c = c + childCentre(E[e]);
                    }
                    double centerDisplacement = vnorm(centerSum);
                    N.circum = centerDisplacement + circumSum;
                    N.in = std::max(0., inSum - centerDisplacement);
                    N.centre[0] = c.x; N.centre[1] = c.y; N.centre[2] = c.z;
                } else if (S.kind == RP_XC_DIFF) {
                    const XcNode &g1 = dst.nodes[E[0].node], &g2 = dst.nodes[E[1].node];
                    double centerDisplacement = vnorm(vec(E[0].pos) - vec(E[1].pos));
                    N.circum = centerDisplacement + g1.circum + g2.circum;
                    N.in = std::max(0., g1.in + g2.in - centerDisplacement);
                    V3 c = childCentre(E[0]) - childCentre(E[1]);
                    N.centre[0] = c.x; N.centre[1] = c.y; N.centre[2] = c.z;
                } else {
                    double circum = vnorm(vec(E[0].pos)) + dst.nodes[E[0].node].circum;
                    V3 c = v3(0, 0, 0);
                    for (int e = 0; e < ne; e++) {
                        circum = std::max(circum, vnorm(vec(E[e].pos)) + dst.nodes[E[e].node].circum);
                        c = c + childCentre(E[e]);
                    }
                    N.circum = circum;
                    if (ne == 1) N.in = std::max(0.0, dst.nodes[E[0].node].in - vnorm(vec(E[0].pos)));
                    else if (ne == 2) N.in = maxInsphere2(vec(E[0].pos), dst.nodes[E[0].node].in, vec(E[1].pos),
                                                          dst.nodes[E[1].node].in);
                    else N.in = 0;
                    double cnt = static_cast<double>(ne);
                    N.centre[0] = c.x / cnt; N.centre[1] = c.y / cnt; N.centre[2] = c.z / cnt;
                }
                break;
            }
            default:
                throw rp_error(RP_ERR_UNSUPPORTED, "unknown rp_xc_kind");
        }
    }
    for (int r = 0; r < std::max(nRoots, 1); r++) {
        RP_EXPECT(src.root[r] >= 0 && src.root[r] < src.n_nodes, "xc.root out of range");
        dst.root[r] = src.root[r];
    }
}

}  // namespace rp
