// rp_math.cuh — FP64 3-vector / 3x3 / box / boundary-condition / pair arithmetic of the Monte Carlo sweep.
//
// Every function replicates the *operation order* of the reference arithmetic it cites, so that with
// `nvcc -fmad=false` (device) and `-ffp-contract=off` (host) each operation is one correctly rounded IEEE-754
// double operation on both sides and overlap verdicts are bit-identical to the reference built without FMA
// contraction.  Citations are file:line under the reference's src/.
//
// One deliberate deviation: the reference's dot product starts from `prod = 0` (geometry/Vector.tpp:53-60); the
// leading `0 + a0*b0` is dropped here.  It can only change the sign of an exact zero, which no comparison, square,
// sum with a non-zero term or guarded division on this path can observe.
#pragma once

#include <cmath>
#include <cstdint>

#define RP_HD __host__ __device__ __forceinline__

namespace rp {

struct V3 { double x, y, z; };

RP_HD V3 v3(double x, double y, double z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
RP_HD V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
RP_HD V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
RP_HD V3 operator*(V3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
RP_HD V3 operator*(double s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
RP_HD V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
// geometry/Vector.tpp:53-60 — (a0*b0 + a1*b1) + a2*b2
RP_HD double dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
RP_HD double norm2(V3 a) { return dot(a, a); }
// geometry/Vector.tpp:105-112
RP_HD V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }

// Row-major 3x3 (geometry/Matrix.h:177)
struct M33 { double a[9]; };

// geometry/Matrix.tpp:104-113 — zero-initialised accumulation in k order
RP_HD V3 mul(const M33 &m, V3 v) {
    return v3((m.a[0] * v.x + m.a[1] * v.y) + m.a[2] * v.z,
              (m.a[3] * v.x + m.a[4] * v.y) + m.a[5] * v.z,
              (m.a[6] * v.x + m.a[7] * v.y) + m.a[8] * v.z);
}
RP_HD M33 mul(const M33 &p, const M33 &q) {
    M33 r;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            r.a[3 * i + j] = (p.a[3 * i] * q.a[j] + p.a[3 * i + 1] * q.a[3 + j]) + p.a[3 * i + 2] * q.a[6 + j];
    return r;
}
RP_HD V3 column(const M33 &m, int c) { return v3(m.a[c], m.a[3 + c], m.a[6 + c]); }

// geometry/Matrix.tpp:352-359 (det) and :258-272, 298-340 (cofactor inverse: +-minor(j,i) / det)
RP_HD double det(const M33 &m) {
    const double *a = m.a;
    return a[0] * a[4] * a[8] + a[1] * a[5] * a[6] + a[2] * a[3] * a[7]
           - a[2] * a[4] * a[6] - a[1] * a[3] * a[8] - a[0] * a[5] * a[7];
}
RP_HD double minor2(const M33 &m, int row, int col) {
    double s[4];
    int k = 0;
    for (int i = 0; i < 3; i++) {
        if (i == row) continue;
        for (int j = 0; j < 3; j++) {
            if (j == col) continue;
            s[k++] = m.a[3 * i + j];
        }
    }
    return s[0] * s[3] - s[1] * s[2];
}
RP_HD M33 inverse(const M33 &m) {
    double d = det(m);
    M33 out;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            out.a[3 * i + j] = ((i + j) % 2 == 0) ? (minor2(m, j, i) / d) : (-minor2(m, j, i) / d);
    return out;
}

// Box: dimensions (columns = box vectors) + cached inverse (core/TriclinicBox.h:20-23, 61-70)
struct Box {
    M33 m;
    M33 inv;
};
RP_HD V3 absToRel(const Box &b, V3 p) { return mul(b.inv, p); }
RP_HD V3 relToAbs(const Box &b, V3 s) { return mul(b.m, s); }

// core/TriclinicBox.cpp:27-51
inline __host__ double boxVolume(const Box &b) {
    V3 s1 = relToAbs(b, v3(1, 0, 0)), s2 = relToAbs(b, v3(0, 1, 0)), s3 = relToAbs(b, v3(0, 0, 1));
    return std::abs(dot(cross(s1, s2), s3));
}
inline __host__ void boxHeights(const Box &b, double h[3]) {
    V3 s1 = relToAbs(b, v3(1, 0, 0)), s2 = relToAbs(b, v3(0, 1, 0)), s3 = relToAbs(b, v3(0, 0, 1));
    double vol = boxVolume(b);
    h[0] = vol / std::sqrt(norm2(cross(s2, s3)));
    h[1] = vol / std::sqrt(norm2(cross(s3, s1)));
    h[2] = vol / std::sqrt(norm2(cross(s1, s2)));
}

// core/PeriodicBoundaryConditions.cpp:11-22 — wrap into [0,1)^3 in relative coordinates, mapped back with M
RP_HD V3 pbcCorrection(const Box &b, V3 pos) {
    V3 rel = absToRel(b, pos);
    double r[3] = {rel.x, rel.y, rel.z};
    double c[3] = {0, 0, 0};
    for (int i = 0; i < 3; i++) {
        while (r[i] + c[i] < 0) c[i] += 1;
        while (r[i] + c[i] >= 1) c[i] -= 1;
    }
    return relToAbs(b, v3(c[0], c[1], c[2]));
}
// core/PeriodicBoundaryConditions.cpp:24-36 — minimum-image translation of pos2 towards pos1
RP_HD V3 pbcTranslation(const Box &b, V3 p1, V3 p2) {
    V3 r1 = absToRel(b, p1), r2 = absToRel(b, p2);
    double a1[3] = {r1.x, r1.y, r1.z}, a2[3] = {r2.x, r2.y, r2.z};
    double t[3] = {0, 0, 0};
    for (int i = 0; i < 3; i++) {
        while (a2[i] + t[i] - a1[i] > 0.5) t[i] -= 1;
        while (a2[i] + t[i] - a1[i] < -0.5) t[i] += 1;
    }
    return relToAbs(b, v3(t[0], t[1], t[2]));
}
// core/PeriodicBoundaryConditions.cpp:38-50 — minimum-image distance squared
RP_HD double pbcDistance2(const Box &b, V3 p1, V3 p2) {
    V3 r1 = absToRel(b, p1), r2 = absToRel(b, p2);
    double a1[3] = {r1.x, r1.y, r1.z}, a2[3] = {r2.x, r2.y, r2.z};
    double d[3];
    for (int i = 0; i < 3; i++) {
        double c = fabs(a2[i] - a1[i]);
        while (c > 0.5) c -= 1;
        d[i] = c;
    }
    return norm2(relToAbs(b, v3(d[0], d[1], d[2])));
}

// The two BoundaryConditions flavours the pair functions see:
//   BcFixed — HardcodedTranslation of a neighbour cell (core/Packing.cpp:23-43)
//   BcPeriodic — the real PeriodicBoundaryConditions, used when there is no neighbour grid (Packing.cpp:411-422)
struct BcFixed {
    V3 t;
    RP_HD V3 translation(V3, V3) const { return t; }
    RP_HD double distance2(V3 p1, V3 p2) const { return norm2(p2 + t - p1); }
};
struct BcPeriodic {
    const Box *box;
    RP_HD V3 translation(V3 p1, V3 p2) const { return pbcTranslation(*box, p1, p2); }
    RP_HD double distance2(V3 p1, V3 p2) const { return pbcDistance2(*box, p1, p2); }
};

// geometry/SegmentDistanceCalculator.h:22-85 — softSurfer / Dan Sunday segment-segment distance squared.
// Same arithmetic and the same EPSILON as the reference; its if/else ladder only SELECTS among already computed
// values (sN, sD, tN, tD), so it is written here with selects: identical results bit for bit, but straight-line code
// the GPU can software-pipeline (two independent pair tests interleave in the sweep kernel).  The two divisions are
// evaluated unconditionally and discarded by the same |x| < EPSILON tests the reference applies first.
RP_HD double segmentDistance2(V3 s11, V3 s12, V3 s21, V3 s22) {
    const double EPS = 0.0000000001;
    V3 u = s12 - s11;
    V3 v = s22 - s21;
    V3 w = s11 - s21;
    double a = dot(u, u), b = dot(u, v), c = dot(v, v), d = dot(u, w), e = dot(v, w);
    double D = a * c - b * b;
    // first stage (reference lines 36-57)
    const bool small = D < EPS;
    const double sN0 = (b * e - c * d), tN0 = (a * e - b * d);
    const bool lt = !small && sN0 < 0.0;
    const bool gt = !small && !lt && sN0 > D;
    double sN = (small || lt) ? 0.0 : (gt ? D : sN0);
    double sD = small ? 1.0 : D;
    double tN = (small || lt) ? e : (gt ? e + b : tN0);
    double tD = (small || lt || gt) ? c : D;
    // second stage (reference lines 59-79)
    const bool tneg = tN < 0.0;
    const bool tbig = !tneg && tN > tD;
    const double y = tneg ? -d : (-d + b);
    const bool ylt = y < 0.0;
    const bool ygt = !ylt && y > a;
    if (tneg || tbig) {
        sN = ylt ? 0.0 : (ygt ? sD : y);
        sD = (ylt || ygt) ? sD : a;
    }
    tN = tneg ? 0.0 : (tbig ? tD : tN);
    double sc = (fabs(sN) < EPS ? 0.0 : sN / sD);
    double tc = (fabs(tN) < EPS ? 0.0 : tN / tD);
    V3 dP = w + (sc * u) - (tc * v);
    return norm2(dP);
}

}  // namespace rp
