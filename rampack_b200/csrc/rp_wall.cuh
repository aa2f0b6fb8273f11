// rp_wall.cuh — hard walls on the box faces: Packing::countParticleWallOverlaps (core/Packing.cpp:1086-1142) and
// Interaction::overlapWithWall of every supported shape (SphereTraits.cpp:67-75, SpherocylinderTraits.cpp:56-73,
// PolysphereTraits.cpp:129-136, PolyspherocylinderTraits.cpp:152-170, XenoCollideTraits.h:170-186).
#pragma once

#include "rp_pair.cuh"
#include "rp_xc.cuh"

namespace rp {

struct Walls {
    int mask = 0;            // bit i: Packing::hasWall[i]
    V3 normal[3];            // (side[i+1] ^ side[i+2]).normalized()
    V3 furtherOrigin;        // side[0] + side[1] + side[2]
    double halfTotalRange;   // interaction.getTotalRangeRadius() / 2
};

inline __host__ Walls makeWalls(const Box &box, int mask, double totalRange) {
    Walls w;
    w.mask = mask;
    V3 s[3] = {relToAbs(box, v3(1, 0, 0)), relToAbs(box, v3(0, 1, 0)), relToAbs(box, v3(0, 0, 1))};
    for (int i = 0; i < 3; i++) {
        V3 c = cross(s[(i + 1) % 3], s[(i + 2) % 3]);
        double nrm = std::sqrt(norm2(c));
        w.normal[i] = v3(c.x / nrm, c.y / nrm, c.z / nrm);
    }
    w.furtherOrigin = (v3(0, 0, 0) + s[0]) + s[1] + s[2];       // std::accumulate over the sides
    w.halfTotalRange = totalRange / 2;
    return w;
}

// Interaction::overlapWithWall(pos, orientation, idx, wallOrigin, wallVector)
RP_HD bool overlapWithWall(const Inter &I, V3 pos, const M33 &orientation, int idx, V3 wallOrigin, V3 wallVector) {
    switch (I.hard) {
        case RP_HARD_SPHERE:
            return dot(wallVector, pos - wallOrigin) < I.radius;
        case RP_HARD_POLYSPHERE:
            return dot(wallVector, pos - wallOrigin) < I.cradius[idx];
        case RP_HARD_SPHEROCYLINDER: {
            V3 halfAxis = column(orientation, 2) * I.scHalfPlus;       // orientation.column(2) * (length / 2)
            if (dot(wallVector, (pos + halfAxis) - wallOrigin) < I.radius) return true;
            return dot(wallVector, (pos - halfAxis) - wallOrigin) < I.radius;
        }
        case RP_HARD_POLYSPHEROCYLINDER: {
            V3 halfAxis = mul(orientation, v3(I.chalf[idx][0], I.chalf[idx][1], I.chalf[idx][2]));
            if (dot(wallVector, (pos - halfAxis) - wallOrigin) < I.cradius[idx]) return true;
            return dot(wallVector, (pos + halfAxis) - wallOrigin) < I.cradius[idx];
        }
        case RP_HARD_XENOCOLLIDE: {
            const XcProgram &P = *I.xc;
            V3 normalVector = mulT(orientation, wallVector);
            V3 supportPoint = xcSupport<XC_MAX_DEPTH>(P, P.root[idx], v3(-normalVector.x, -normalVector.y, -normalVector.z));
            V3 origin = mulT(orientation, wallOrigin - pos);
            double distanceSupport = -dot(supportPoint, normalVector);
            double distanceWall = -dot(origin, normalVector);
            return !(distanceWall > distanceSupport);
        }
        default:
            return false;     // CentralInteraction::hasWallPart() == false
    }
}

// Packing::countParticleWallOverlaps for a particle at `pos` with orientation `orient`; crel(c) = rotated centre
// offset (Packing::interactionCentres).  No early exit: the callers compare with 0 or need the count.
template <typename CentreFn>
RP_HD unsigned int particleWallOverlaps(const Inter &I, const Walls &W, V3 pos, const M33 &orient, CentreFn crel) {
    unsigned int count = 0;
    for (int i = 0; i < 3; i++) {
        if (!((W.mask >> i) & 1)) continue;
        V3 wallVector = W.normal[i];
        V3 origin, vec;
        if (dot(pos, wallVector) < W.halfTotalRange) { origin = v3(0, 0, 0); vec = wallVector; }
        else if (dot(W.furtherOrigin - pos, wallVector) < W.halfTotalRange) {
            origin = W.furtherOrigin;
            vec = v3(-wallVector.x, -wallVector.y, -wallVector.z);
        } else continue;
        if (I.k == 0) count += overlapWithWall(I, pos, orient, 0, origin, vec) ? 1u : 0u;
        else for (int c = 0; c < I.k; c++) count += overlapWithWall(I, pos + crel(c), orient, c, origin, vec) ? 1u : 0u;
    }
    return count;
}

}  // namespace rp
