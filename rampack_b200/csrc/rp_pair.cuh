// rp_pair.cuh — device/host pair functions: Interaction::overlapBetween / calculateEnergyBetween for the shape
// classes the device path supports, in the reference's operation order (citations under the reference's src/).
#pragma once

#include "rp_math.cuh"
#include "rp_xc.cuh"
#include "../../include/rampack_b200.h"

namespace rp {

// Flattened Interaction parameters.  All thresholds are precomputed on the host with the same expressions the
// reference evaluates per call, so the per-pair comparisons see identical bits.
struct Inter {
    int hard;     // rp_hard_kind
    int soft;     // rp_soft_kind
    int k;        // Interaction::getInteractionCentres().size(); 0 = single centre in the origin
    int kk;       // max(k, 1): centres per particle in the neighbour grid
    double range;       // Interaction::getRangeRadius()
    double totalRange;  // Interaction::getTotalRangeRadius() (core/Interaction.cpp:55-70)
    // sphere (core/shapes/SphereTraits.cpp:56-65): std::pow(2*r, 2)
    double sphereThr2;
    double radius;      // sphere / spherocylinder radius (wall tests)
    // spherocylinder (core/shapes/SpherocylinderTraits.cpp:28-46): 4*r*r, std::pow(2*r + l, 2), 0.5*(+-1)*l
    double scInner2, scOuter2, scHalfMinus, scHalfPlus;
    // central potentials (core/interactions/*.cpp)
    double sigma2;        // sigma*sigma (LJ, WCA) or std::pow(sigma, 2) (SIC) — the same double
    double fourEps;       // 4*epsilon
    double eps;
    double wcaCut2;       // (sigma * 2^(1/6))^2
    // per-centre data
    double cpos[RP_MAX_CENTRES][3];
    double cradius[RP_MAX_CENTRES];      // polysphere bead radius / polyspherocylinder radius
    double chalf[RP_MAX_CENTRES][3];     // polyspherocylinder body-frame half axis
    double ccircum[RP_MAX_CENTRES];      // radius + |halfAxis| (core/shapes/PolyspherocylinderTraits.cpp:78-84)
    bool convex;                         // Interaction::isConvex()
    const XcProgram *xc;                 // RP_HARD_XENOCOLLIDE: device copy of the collide geometries
};

// "aux" vector carried beside each interaction centre in the sorted pair records:
//   spherocylinder      -> orientation.column(2)                     (SpherocylinderTraits.cpp:24-26)
//   polyspherocylinder  -> orientation * halfAxis_c                  (PolyspherocylinderTraits.cpp:128-129)
//   others              -> unused
RP_HD V3 pairAux(const Inter &I, const M33 &orientation, int centre) {
    if (I.hard == RP_HARD_SPHEROCYLINDER)
        return column(orientation, 2);
    if (I.hard == RP_HARD_POLYSPHEROCYLINDER)
        return mul(orientation, v3(I.chalf[centre][0], I.chalf[centre][1], I.chalf[centre][2]));
    return v3(0, 0, 0);
}

// Spherocylinder pre-filter on the centre distance; returns 1 = overlap, 0 = disjoint, 2 = needs the segment test
RP_HD int scPrefilter(const Inter &I, double distance2) {
    if (distance2 < I.scInner2) return 1;
    if (distance2 >= I.scOuter2) return 0;
    return 2;
}
// Cap centres: pos + axis * (0.5 * beginOrEnd * length), then the segment distance (SpherocylinderTraits.cpp:36-46)
RP_HD bool scFull(const Inter &I, V3 pos1, V3 axis1, V3 pos2bc, V3 axis2) {
    double d2 = segmentDistance2(pos1 + axis1 * I.scHalfMinus, pos1 + axis1 * I.scHalfPlus,
                                 pos2bc + axis2 * I.scHalfMinus, pos2bc + axis2 * I.scHalfPlus);
    return d2 < I.scInner2;
}

// Interaction::overlapBetween for (centre c1 at pos1 with aux1) vs (centre c2 at pos2 with aux2)
// (HARDK >= 0: the hard kind is a compile-time constant of the calling kernel and the switch folds away)
template <int HARDK = -1, typename BC>
RP_HD bool overlapBetween(const Inter &I, V3 pos1, V3 aux1, int c1, V3 pos2, V3 aux2, int c2, const BC &bc) {
    switch (HARDK >= 0 ? HARDK : I.hard) {
        case RP_HARD_SPHERE:
            return bc.distance2(pos1, pos2) < I.sphereThr2;
        case RP_HARD_SPHEROCYLINDER: {
            V3 pos2bc = pos2 + bc.translation(pos1, pos2);
            double distance2 = norm2(pos2bc - pos1);
            int pre = scPrefilter(I, distance2);
            if (pre != 2) return pre == 1;
            return scFull(I, pos1, aux1, pos2bc, aux2);
        }
        case RP_HARD_POLYSPHERE: {
            // core/shapes/PolysphereTraits.cpp:97-106
            double r = I.cradius[c1] + I.cradius[c2];
            return bc.distance2(pos1, pos2) < r * r;
        }
        case RP_HARD_POLYSPHEROCYLINDER: {
            // core/shapes/PolyspherocylinderTraits.cpp:109-132
            V3 pos2bc = pos2 + bc.translation(pos1, pos2);
            double distance2 = norm2(pos2bc - pos1);
            double insphereR = I.cradius[c1] + I.cradius[c2];
            double insphereR2 = insphereR * insphereR;
            if (distance2 < insphereR2) return true;
            double circumsphereR = I.ccircum[c1] + I.ccircum[c2];
            double circumsphereR2 = circumsphereR * circumsphereR;
            if (distance2 >= circumsphereR2) return false;
            return segmentDistance2(pos1 + aux1, pos1 - aux1, pos2bc + aux2, pos2bc - aux2) < insphereR2;
        }
        default:
            return false;
    }
}

// Interaction::overlapBetween for shapes whose pair test needs the whole orientation (XenoCollide); getO1 / getO2
// return the orientation matrices and are only called when the sphere pre-filters do not decide
// (core/shapes/XenoCollideTraits.h:147-171)
template <typename BC, typename O1, typename O2>
RP_HD bool overlapBetweenOriented(const Inter &I, V3 pos1, O1 &&getO1, int c1, V3 pos2, O2 &&getO2, int c2,
                                  const BC &bc, int *supportCalls = nullptr)
{
    const XcProgram &P = *I.xc;
    const int g1 = P.root[c1], g2 = P.root[c2];
    double rCircumsphere = P.nodes[g1].circum + P.nodes[g2].circum;
    double rInsphere = P.nodes[g1].in + P.nodes[g2].in;
    V3 pos2bc = pos2 + bc.translation(pos1, pos2);
    double dist2 = norm2(pos2bc - pos1);
    if (dist2 > rCircumsphere * rCircumsphere) return false;
    if (dist2 < rInsphere * rInsphere) return true;
    M33 o1 = getO1(), o2 = getO2();
    return xcIntersect(P, g1, o1, pos1, g2, o2, pos2bc, 1.0e-12, supportCalls);
}
RP_HD bool needsOrientation(const Inter &I) { return I.hard == RP_HARD_XENOCOLLIDE; }

// CentralInteraction::calculateEnergyForDistance2 (core/interactions/*.cpp)
RP_HD double energyForDistance2(const Inter &I, double distance2) {
    switch (I.soft) {
        case RP_SOFT_LJ: {
            double x2 = I.sigma2 / distance2;
            double x6 = x2 * x2 * x2;
            double x12 = x6 * x6;
            return I.fourEps * (x12 - x6);
        }
        case RP_SOFT_WCA: {
            if (distance2 >= I.wcaCut2) return 0;
            double x2 = I.sigma2 / distance2;
            double x6 = x2 * x2 * x2;
            double x12 = x6 * x6;
            return I.fourEps * (x12 - x6) + I.eps;
        }
        case RP_SOFT_SIC: {
            double v = I.sigma2 / distance2 - 1;
            return I.eps * (0.0 > v ? 0.0 : v);    // std::max(0.0, v)
        }
        default:
            return 0;
    }
}
// CentralInteraction::calculateEnergyBetween (core/interactions/CentralInteraction.h:58-67)
template <typename BC>
RP_HD double energyBetween(const Inter &I, V3 pos1, V3 pos2, const BC &bc) {
    return energyForDistance2(I, bc.distance2(pos1, pos2));
}

}  // namespace rp
