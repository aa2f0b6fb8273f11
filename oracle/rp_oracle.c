/*
 * rp_oracle.c — TEST INFRASTRUCTURE ONLY.  Plain-C, single-threaded CPU restatement of the reference's Monte Carlo hot
 * path.  Compiled with -ffp-contract=off so that every FP64 operation is one correctly rounded IEEE operation in the
 * reference's operation order.  Never linked or loaded by the product path (rampack_b200/).
 *
 * Parity is pinned (tests/test_oracle.py): the reference's own known-answer vectors in tests/golden/ and, where
 * /root/reference is mounted, the unmodified reference compiled into oracle/_ref/.
 *
 * Citations are file:line under /root/reference/src.  The neighbour grid below keeps the reference's data structure
 * (ghost "reflected" cells + singly linked lists), which is deliberately NOT what the CUDA path does.
 */
#define _GNU_SOURCE
#include "rp_oracle.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* ------------------------------------------------------------------------------------------------------------------
 * geometry/Vector.tpp, geometry/Matrix.tpp
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct { double v[3]; } vec3;

static vec3 V(double x, double y, double z) { vec3 r = {{x, y, z}}; return r; }
static vec3 vadd(vec3 a, vec3 b) { return V(a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2]); }
static vec3 vsub(vec3 a, vec3 b) { return V(a.v[0] - b.v[0], a.v[1] - b.v[1], a.v[2] - b.v[2]); }
static vec3 vscale(vec3 a, double s) { return V(a.v[0] * s, a.v[1] * s, a.v[2] * s); }
/* Vector.tpp:53-60: prod = 0; prod += a_i * b_i */
static double vdot(vec3 a, vec3 b) {
    double prod = 0;
    for (int i = 0; i < 3; i++) prod += a.v[i] * b.v[i];
    return prod;
}
static double vnorm2(vec3 a) { return vdot(a, a); }
/* Vector.tpp:105-112 */
static vec3 vcross(vec3 a, vec3 b) {
    return V(a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2], a.v[0] * b.v[1] - a.v[1] * b.v[0]);
}
/* Matrix.tpp:104-113: zero-initialised accumulation in k order */
static vec3 mvec(const double m[9], vec3 x) {
    vec3 r;
    for (int i = 0; i < 3; i++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += m[3 * i + k] * x.v[k];
        r.v[i] = s;
    }
    return r;
}
static void mmul(const double a[9], const double b[9], double out[9]) {
    double r[9];
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) {
            double s = 0;
            for (int k = 0; k < 3; k++) s += a[3 * i + k] * b[3 * k + j];
            r[3 * i + j] = s;
        }
    memcpy(out, r, sizeof(r));
}
static vec3 mcolumn(const double m[9], int c) { return V(m[c], m[3 + c], m[6 + c]); }
/* Matrix.tpp:352-359 */
static double mdet(const double a[9]) {
    return a[0] * a[4] * a[8] + a[1] * a[5] * a[6] + a[2] * a[3] * a[7]
           - a[2] * a[4] * a[6] - a[1] * a[3] * a[8] - a[0] * a[5] * a[7];
}
/* Matrix.tpp:258-272, 298-340: inverse(i, j) = (-1)^(i+j) * minor(j, i) / det */
static double mminor(const double m[9], int row, int col) {
    double s[4];
    int k = 0;
    for (int i = 0; i < 3; i++) {
        if (i == row) continue;
        for (int j = 0; j < 3; j++) {
            if (j == col) continue;
            s[k++] = m[3 * i + j];
        }
    }
    return s[0] * s[3] - s[1] * s[2];
}
static void minverse(const double m[9], double out[9]) {
    double d = mdet(m);
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            out[3 * i + j] = ((i + j) % 2 == 0) ? (mminor(m, j, i) / d) : (-mminor(m, j, i) / d);
}

/* core/TriclinicBox.h:20-23, 61-70 */
typedef struct { double m[9], inv[9]; } box_t;
static void box_init(box_t *b, const double m[9]) { memcpy(b->m, m, 9 * sizeof(double)); minverse(m, b->inv); }
static vec3 abs_to_rel(const box_t *b, vec3 p) { return mvec(b->inv, p); }
static vec3 rel_to_abs(const box_t *b, vec3 s) { return mvec(b->m, s); }
/* core/TriclinicBox.cpp:27-51 */
static double box_volume(const box_t *b) {
    vec3 s1 = rel_to_abs(b, V(1, 0, 0)), s2 = rel_to_abs(b, V(0, 1, 0)), s3 = rel_to_abs(b, V(0, 0, 1));
    return fabs(vdot(vcross(s1, s2), s3));
}
static void box_heights(const box_t *b, double h[3]) {
    vec3 s1 = rel_to_abs(b, V(1, 0, 0)), s2 = rel_to_abs(b, V(0, 1, 0)), s3 = rel_to_abs(b, V(0, 0, 1));
    double vol = box_volume(b);
    h[0] = vol / sqrt(vnorm2(vcross(s2, s3)));
    h[1] = vol / sqrt(vnorm2(vcross(s3, s1)));
    h[2] = vol / sqrt(vnorm2(vcross(s1, s2)));
}

/* ------------------------------------------------------------------------------------------------------------------
 * boundary conditions: PeriodicBoundaryConditions.cpp:11-50 and HardcodedTranslation (Packing.cpp:23-43)
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct { int periodic; vec3 t; const box_t *box; } bc_t;

static vec3 pbc_correction(const box_t *b, vec3 pos) {
    vec3 rel = abs_to_rel(b, pos), c = V(0, 0, 0);
    for (int i = 0; i < 3; i++) {
        while (rel.v[i] + c.v[i] < 0) c.v[i] += 1;
        while (rel.v[i] + c.v[i] >= 1) c.v[i] -= 1;
    }
    return rel_to_abs(b, c);
}
static vec3 bc_translation(const bc_t *bc, vec3 p1, vec3 p2) {
    if (!bc->periodic) return bc->t;
    vec3 r1 = abs_to_rel(bc->box, p1), r2 = abs_to_rel(bc->box, p2), t = V(0, 0, 0);
    for (int i = 0; i < 3; i++) {
        while (r2.v[i] + t.v[i] - r1.v[i] > 0.5) t.v[i] -= 1;
        while (r2.v[i] + t.v[i] - r1.v[i] < -0.5) t.v[i] += 1;
    }
    return rel_to_abs(bc->box, t);
}
static double bc_distance2(const bc_t *bc, vec3 p1, vec3 p2) {
    if (!bc->periodic) return vnorm2(vsub(vadd(p2, bc->t), p1));
    vec3 r1 = abs_to_rel(bc->box, p1), r2 = abs_to_rel(bc->box, p2), d;
    for (int i = 0; i < 3; i++) {
        double c = fabs(r2.v[i] - r1.v[i]);
        while (c > 0.5) c -= 1;
        d.v[i] = c;
    }
    return vnorm2(rel_to_abs(bc->box, d));
}

/* ------------------------------------------------------------------------------------------------------------------
 * geometry/SegmentDistanceCalculator.h:22-85
 * ---------------------------------------------------------------------------------------------------------------- */
static double segment_distance2(vec3 s11, vec3 s12, vec3 s21, vec3 s22) {
    const double EPSILON = 0.0000000001;
    vec3 u = vsub(s12, s11), v = vsub(s22, s21), w = vsub(s11, s21);
    double a = vdot(u, u), b = vdot(u, v), c = vdot(v, v), d = vdot(u, w), e = vdot(v, w);
    double D = a * c - b * b;
    double sc, sN, sD = D, tc, tN, tD = D;
    if (D < EPSILON) {
        sN = 0.0; sD = 1.0; tN = e; tD = c;
    } else {
        sN = (b * e - c * d);
        tN = (a * e - b * d);
        if (sN < 0.0) { sN = 0.0; tN = e; tD = c; }
        else if (sN > sD) { sN = sD; tN = e + b; tD = c; }
    }
    if (tN < 0.0) {
        tN = 0.0;
        if (-d < 0.0) sN = 0.0;
        else if (-d > a) sN = sD;
        else { sN = -d; sD = a; }
    } else if (tN > tD) {
        tN = tD;
        if ((-d + b) < 0.0) sN = 0;
        else if ((-d + b) > a) sN = sD;
        else { sN = (-d + b); sD = a; }
    }
    sc = (fabs(sN) < EPSILON ? 0.0 : sN / sD);
    tc = (fabs(tN) < EPSILON ? 0.0 : tN / tD);
    vec3 dP = vsub(vadd(w, vscale(u, sc)), vscale(v, tc));
    return vnorm2(dP);
}

/* ------------------------------------------------------------------------------------------------------------------
 * Interaction (core/Interaction.h:28-130) for the shape classes of rp_interaction_desc
 * ---------------------------------------------------------------------------------------------------------------- */
static int inter_k(const rp_interaction_desc *d) { return d->n_centres; }
static int inter_kk(const rp_interaction_desc *d) { return d->n_centres > 0 ? d->n_centres : 1; }

static double hard_range(const rp_interaction_desc *d) {
    switch (d->hard_kind) {
        case RP_HARD_SPHERE: return 2 * d->radius;                          /* SphereTraits.h */
        case RP_HARD_SPHEROCYLINDER: return 2 * d->radius + d->length;      /* SpherocylinderTraits.h:81 */
        case RP_HARD_POLYSPHERE: {                                          /* PolysphereTraits.cpp:116-121 */
            double m = 0;
            for (int i = 0; i < d->n_centres; i++) if (d->centre_radius[i] > m) m = d->centre_radius[i];
            return 2 * m;
        }
        case RP_HARD_POLYSPHEROCYLINDER: {                                  /* PolyspherocylinderTraits.cpp:143-150 */
            double m = 0;
            for (int i = 0; i < d->n_centres; i++) {
                vec3 h = V(d->centre_half_axis[i][0], d->centre_half_axis[i][1], d->centre_half_axis[i][2]);
                double c = d->centre_radius[i] + sqrt(vnorm2(h));
                if (c > m) m = c;
            }
            return 2 * m;
        }
        default: return 0;
    }
}
static double soft_range(const rp_interaction_desc *d) {
    switch (d->soft_kind) {
        case RP_SOFT_LJ: return 3 * d->sigma;                               /* LennardJonesInteraction.h:26 */
        case RP_SOFT_WCA: return d->sigma * pow(2., 1. / 6);                /* RepulsiveLennardJonesInteraction.h */
        case RP_SOFT_SIC: return d->sigma;                                  /* SquareInverseCoreInteraction.h */
        default: return 0;
    }
}
/* CompoundInteraction.cpp:5-66: max of the ranges */
double rpo_range_radius(const rp_interaction_desc *d) {
    double h = hard_range(d), s = soft_range(d);
    return h > s ? h : s;
}
/* Interaction.cpp:55-70 */
double rpo_total_range_radius(const rp_interaction_desc *d) {
    double range = rpo_range_radius(d);
    if (d->n_centres == 0) return range;
    double maxRadius2 = 0;
    for (int i = 0; i < d->n_centres; i++) {
        double r2 = vnorm2(V(d->centres[i][0], d->centres[i][1], d->centres[i][2]));
        if (r2 > maxRadius2) maxRadius2 = r2;
    }
    return 2 * sqrt(maxRadius2) + range;
}
static int is_convex(const rp_interaction_desc *d) {
    return d->hard_kind == RP_HARD_SPHERE || d->hard_kind == RP_HARD_SPHEROCYLINDER;
}

static int overlap_between(const rp_interaction_desc *d, vec3 pos1, const double *o1, int idx1, vec3 pos2,
                           const double *o2, int idx2, const bc_t *bc)
{
    switch (d->hard_kind) {
        case RP_HARD_SPHERE:                                                /* SphereTraits.cpp:56-65 */
            return bc_distance2(bc, pos1, pos2) < pow(2 * d->radius, 2);
        case RP_HARD_SPHEROCYLINDER: {                                      /* SpherocylinderTraits.cpp:24-46 */
            vec3 pos2bc = vadd(pos2, bc_translation(bc, pos1, pos2));
            double distance2 = vnorm2(vsub(pos2bc, pos1));
            if (distance2 < 4 * d->radius * d->radius) return 1;
            else if (distance2 >= pow(2 * d->radius + d->length, 2)) return 0;
            vec3 a1 = mcolumn(o1, 2), a2 = mcolumn(o2, 2);
            vec3 c11 = vadd(pos1, vscale(a1, 0.5 * -1 * d->length)), c12 = vadd(pos1, vscale(a1, 0.5 * 1 * d->length));
            vec3 c21 = vadd(pos2bc, vscale(a2, 0.5 * -1 * d->length));
            vec3 c22 = vadd(pos2bc, vscale(a2, 0.5 * 1 * d->length));
            return segment_distance2(c11, c12, c21, c22) < 4 * d->radius * d->radius;
        }
        case RP_HARD_POLYSPHERE: {                                          /* PolysphereTraits.cpp:97-106 */
            double r = d->centre_radius[idx1] + d->centre_radius[idx2];
            return bc_distance2(bc, pos1, pos2) < r * r;
        }
        case RP_HARD_POLYSPHEROCYLINDER: {                                  /* PolyspherocylinderTraits.cpp:109-132 */
            vec3 h1 = V(d->centre_half_axis[idx1][0], d->centre_half_axis[idx1][1], d->centre_half_axis[idx1][2]);
            vec3 h2 = V(d->centre_half_axis[idx2][0], d->centre_half_axis[idx2][1], d->centre_half_axis[idx2][2]);
            double circ1 = d->centre_radius[idx1] + sqrt(vnorm2(h1)), circ2 = d->centre_radius[idx2] + sqrt(vnorm2(h2));
            vec3 pos2bc = vadd(pos2, bc_translation(bc, pos1, pos2));
            double distance2 = vnorm2(vsub(pos2bc, pos1));
            double insphereR = d->centre_radius[idx1] + d->centre_radius[idx2];
            double insphereR2 = insphereR * insphereR;
            if (distance2 < insphereR2) return 1;
            double circumsphereR = circ1 + circ2;
            double circumsphereR2 = circumsphereR * circumsphereR;
            if (distance2 >= circumsphereR2) return 0;
            vec3 ha1 = mvec(o1, h1), ha2 = mvec(o2, h2);
            return segment_distance2(vadd(pos1, ha1), vsub(pos1, ha1), vadd(pos2bc, ha2), vsub(pos2bc, ha2))
                   < insphereR2;
        }
        default:
            return 0;
    }
}

/* CentralInteraction.h:58-67 + LennardJonesInteraction.cpp:8-13, RepulsiveLennardJonesInteraction.cpp:8-16,
 * SquareInverseCoreInteraction.cpp:13-15 */
static double energy_for_distance2(const rp_interaction_desc *d, double distance2) {
    switch (d->soft_kind) {
        case RP_SOFT_LJ: {
            double x2 = d->sigma * d->sigma / distance2;
            double x6 = x2 * x2 * x2;
            double x12 = x6 * x6;
            return 4 * d->epsilon * (x12 - x6);
        }
        case RP_SOFT_WCA: {
            double sigmaTimesTwoToOneSixth = d->sigma * pow(2., 1. / 6);
            if (distance2 >= sigmaTimesTwoToOneSixth * sigmaTimesTwoToOneSixth) return 0;
            double x2 = d->sigma * d->sigma / distance2;
            double x6 = x2 * x2 * x2;
            double x12 = x6 * x6;
            return 4 * d->epsilon * (x12 - x6) + d->epsilon;
        }
        case RP_SOFT_SIC: {
            double v = pow(d->sigma, 2) / distance2 - 1;
            return d->epsilon * (0.0 > v ? 0.0 : v);
        }
        default:
            return 0;
    }
}
static double energy_between(const rp_interaction_desc *d, vec3 pos1, vec3 pos2, const bc_t *bc) {
    return energy_for_distance2(d, bc_distance2(bc, pos1, pos2));
}

static int pair_batch(const rp_interaction_desc *desc, const box_t *box, const rp_pair *pairs, uint64_t n,
                      uint8_t *verdict, double *energy)
{
    for (uint64_t i = 0; i < n; i++) {
        const rp_pair *p = &pairs[i];
        bc_t bc;
        bc.periodic = box != NULL;
        bc.box = box;
        bc.t = V(p->translation[0], p->translation[1], p->translation[2]);
        vec3 p1 = V(p->pos1[0], p->pos1[1], p->pos1[2]), p2 = V(p->pos2[0], p->pos2[1], p->pos2[2]);
        if (verdict)
            verdict[i] = desc->hard_kind != RP_HARD_NONE
                         && overlap_between(desc, p1, p->orientation1, (int) p->idx1, p2, p->orientation2,
                                            (int) p->idx2, &bc);
        if (energy)
            energy[i] = desc->soft_kind != RP_SOFT_NONE ? energy_between(desc, p1, p2, &bc) : 0;
    }
    return 0;
}
int rpo_pair_test_batch(const rp_interaction_desc *desc, const rp_pair *pairs, uint64_t n, uint8_t *verdict,
                        double *energy)
{
    return pair_batch(desc, NULL, pairs, n, verdict, energy);
}
int rpo_pair_test_batch_pbc(const rp_interaction_desc *desc, const double boxm[9], const rp_pair *pairs, uint64_t n,
                            uint8_t *verdict, double *energy)
{
    box_t box;
    box_init(&box, boxm);
    return pair_batch(desc, &box, pairs, n, verdict, energy);
}

void rpo_box_info(const double boxm[9], double heights[3], double *volume, double inverse[9]) {
    box_t b;
    box_init(&b, boxm);
    box_heights(&b, heights);
    *volume = box_volume(&b);
    memcpy(inverse, b.inv, sizeof(b.inv));
}
void rpo_pbc_correction(const double boxm[9], const double pos[3], double out[3]) {
    box_t b;
    box_init(&b, boxm);
    vec3 c = pbc_correction(&b, V(pos[0], pos[1], pos[2]));
    memcpy(out, c.v, sizeof(c.v));
}
/* geometry/Matrix.tpp:437-449: identity + K*sina + ((1 - cosa)*K)*K */
void rpo_rotation_matrix(const double axis[3], double angle, double out[9]) {
    double sina = sin(angle), cosa = cos(angle);
    double K[9] = {0, -axis[2], axis[1], axis[2], 0, -axis[0], -axis[1], axis[0], 0};
    double sK[9], KK[9];
    for (int i = 0; i < 9; i++) sK[i] = (1 - cosa) * K[i];
    mmul(sK, K, KK);
    for (int i = 0; i < 9; i++) out[i] = ((i % 4 == 0) ? 1.0 : 0.0) + K[i] * sina + KK[i];
}

/* ------------------------------------------------------------------------------------------------------------------
 * NeighbourGrid (core/NeighbourGrid.h:30-44, NeighbourGrid.cpp) — ghost cells + linked lists, as in the reference
 * ---------------------------------------------------------------------------------------------------------------- */
#define LIST_END ((size_t) -1)

typedef struct {
    box_t box;
    size_t div[3];          /* cellDivisions incl. the 2 ghost layers */
    double relCell[3];
    size_t numCells;
    size_t *heads, *succ, *reflected, *transIdx;
    size_t numItems;
    vec3 trans[27];
    long offsets[27];
    int numOffsets;
    long posOffsets[13];
    int numPosOffsets;
} grid_t;

static size_t grid_coords_to_cell(const grid_t *g, const size_t c[3]) {
    size_t r = 0;
    for (int i = 2; i >= 0; i--) r = g->div[i] * r + c[i];
    return r;
}
static void grid_cell_to_coords(const grid_t *g, size_t cellNo, size_t c[3]) {
    for (int i = 0; i < 3; i++) { c[i] = cellNo % g->div[i]; cellNo /= g->div[i]; }
}
/* NeighbourGrid.cpp:13-34; returns -1 through *ok when the position is outside the box beyond EPSILON */
static size_t grid_position_to_cell(const grid_t *g, vec3 pos, int *ok) {
    size_t result = 0;
    vec3 rel = abs_to_rel(&g->box, pos);
    for (int i = 2; i >= 0; i--) {
        const double EPSILON = DBL_EPSILON * 10;
        double r = rel.v[i];
        if (r < 0) {
            if (!(r > -EPSILON) && ok) *ok = 0;
            r = 0;
        } else if (r >= 1) {
            if (!(r < 1 + EPSILON) && ok) *ok = 0;
            r = 1 - EPSILON;
        }
        size_t coord = (size_t) ((int) (r / g->relCell[i]) + 1);
        result = g->div[i] * result + coord;
    }
    return result;
}
static int cmp_long(const void *a, const void *b) {
    long x = *(const long *) a, y = *(const long *) b;
    return (x > y) - (x < y);
}
static int uniq_long(long *a, int n) {
    qsort(a, n, sizeof(long), cmp_long);
    int m = 0;
    for (int i = 0; i < n; i++) if (m == 0 || a[m - 1] != a[i]) a[m++] = a[i];
    return m;
}
/* NeighbourGrid ctor + setupSizes + getReflectedCellData + fillNeighbouringCellsOffsets
 * (NeighbourGrid.cpp:80-100, 113-165, 178-217) */
static int grid_init(grid_t *g, const box_t *box, double cellSize, size_t numItems) {
    memset(g, 0, sizeof(*g));
    g->box = *box;
    double h[3];
    box_heights(box, h);
    g->numCells = 1;
    for (int i = 0; i < 3; i++) {
        g->div[i] = (size_t) floor(h[i] / cellSize) + 2;
        if (g->div[i] < 3) return 1;      /* "Neighbour grid cell too big" */
        g->relCell[i] = 1 / (double) (g->div[i] - 2);
        g->numCells *= g->div[i];
    }
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            for (int k = 0; k < 3; k++)
                g->trans[(i * 3 + j) * 3 + k] = rel_to_abs(box, V((double) i - 1, (double) j - 1., (double) k - 1.));
    g->heads = malloc(g->numCells * sizeof(size_t));
    g->reflected = malloc(g->numCells * sizeof(size_t));
    g->transIdx = malloc(g->numCells * sizeof(size_t));
    g->succ = malloc((numItems + 1) * sizeof(size_t));
    g->numItems = numItems;
    for (size_t c = 0; c < g->numCells; c++) g->heads[c] = LIST_END;
    for (size_t i = 0; i < numItems; i++) g->succ[i] = LIST_END;
    for (size_t c = 0; c < g->numCells; c++) {
        size_t co[3], t[3] = {1, 1, 1};
        grid_cell_to_coords(g, c, co);
        for (int i = 0; i < 3; i++) {
            if (co[i] == 0) { co[i] = g->div[i] - 2; t[i] = 0; }
            else if (co[i] == g->div[i] - 1) { co[i] = 1; t[i] = 2; }
        }
        g->reflected[c] = grid_coords_to_cell(g, co);
        g->transIdx[c] = (t[0] * 3 + t[1]) * 3 + t[2];
    }
    size_t test[3];
    for (int i = 0; i < 3; i++) test[i] = g->div[i] / 2;
    long testNo = (long) grid_coords_to_cell(g, test);
    g->numOffsets = g->numPosOffsets = 0;
    for (int n2 = 0; n2 < 3; n2++)
        for (int n1 = 0; n1 < 3; n1++)
            for (int n0 = 0; n0 < 3; n0++) {
                size_t co[3] = {test[0] + n0 - 1, test[1] + n1 - 1, test[2] + n2 - 1};
                long off = (long) grid_coords_to_cell(g, co) - testNo;
                g->offsets[g->numOffsets++] = off;
                if (9 * n0 + 3 * n1 + n2 > 13) g->posOffsets[g->numPosOffsets++] = off;
            }
    g->numOffsets = uniq_long(g->offsets, g->numOffsets);
    g->numPosOffsets = uniq_long(g->posOffsets, g->numPosOffsets);
    return 0;
}
static void grid_free(grid_t *g) {
    free(g->heads); free(g->succ); free(g->reflected); free(g->transIdx);
    memset(g, 0, sizeof(*g));
}
/* NeighbourGrid.cpp:219-258 */
static void grid_add(grid_t *g, size_t idx, vec3 pos) {
    size_t c = grid_position_to_cell(g, pos, NULL);
    g->succ[idx] = g->heads[c];
    g->heads[c] = idx;
}
static void grid_remove(grid_t *g, size_t idx, vec3 pos) {
    size_t c = grid_position_to_cell(g, pos, NULL);
    size_t head = g->heads[c];
    if (head == idx) {
        g->heads[c] = g->succ[idx];
        g->succ[idx] = LIST_END;
    } else {
        while (head != LIST_END) {
            if (g->succ[head] == idx) {
                g->succ[head] = g->succ[idx];
                g->succ[idx] = LIST_END;
                break;
            }
            head = g->succ[head];
        }
    }
}

int rpo_cell_numbers(const double boxm[9], double cell_size, const double *positions, uint64_t n, uint64_t *out,
                     uint64_t divisions_out[3])
{
    box_t b;
    box_init(&b, boxm);
    grid_t g;
    if (grid_init(&g, &b, cell_size, 0)) return 1;
    int ok = 1;
    for (int i = 0; i < 3; i++) divisions_out[i] = g.div[i] - 2;
    for (uint64_t i = 0; i < n; i++)
        out[i] = grid_position_to_cell(&g, V(positions[3 * i], positions[3 * i + 1], positions[3 * i + 2]), &ok);
    grid_free(&g);
    return ok ? 0 : 1;
}

/* ------------------------------------------------------------------------------------------------------------------
 * Packing (core/Packing.h:32-69, Packing.cpp)
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct { vec3 pos; double o[9]; } shape_t;

struct rpo_packing {
    rp_interaction_desc desc;
    box_t box, lastBox;
    size_t n;                 /* size(); one temp slot lives at index n */
    shape_t *shapes, *lastShapes;
    vec3 *crel, *cabs, *lastCabs;      /* interactionCentres / absoluteInteractionCentres, (n + 1) * k entries */
    double interactionRange;
    int hasGrid, tempHasGrid;
    grid_t grid, tempGrid;
    size_t lastAltered;
    long lastMoveOverlapDelta;
    int overlapCounting;
    size_t numOverlaps, lastScalingNumOverlaps;
};

static vec3 item_pos(const rpo_packing *p, size_t idx) {
    return inter_k(&p->desc) == 0 ? p->shapes[idx].pos : p->cabs[idx];
}

/* Packing.cpp:1041-1047 */
static void recalc_abs_centres(rpo_packing *p, size_t particle) {
    int k = inter_k(&p->desc);
    for (int c = 0; c < k; c++) {
        size_t idx = particle * k + c;
        vec3 pos = vadd(p->shapes[particle].pos, p->crel[idx]);
        p->cabs[idx] = vadd(pos, pbc_correction(&p->box, pos));
    }
}

/* Packing.cpp:858-929 */
static void rebuild_grid(rpo_packing *p) {
    double cellSize = p->interactionRange;
    double minCellSize = cbrt(box_volume(&p->box) / p->n) / 5;
    if (p->interactionRange < minCellSize) cellSize = minCellSize;
    double h[3];
    box_heights(&p->box, h);
    double maxH = fmax(h[0], fmax(h[1], h[2])), minH = fmin(h[0], fmin(h[1], h[2]));
    if (p->hasGrid) grid_free(&p->grid);
    p->hasGrid = 0;
    if (cellSize * 4 > maxH) return;
    static const double CELL_SIZE_EPSILON = 1 + 1e-12;
    if (cellSize * CELL_SIZE_EPSILON > minH) return;
    size_t total = p->n * inter_kk(&p->desc);
    if (grid_init(&p->grid, &p->box, cellSize, total + inter_kk(&p->desc))) return;
    p->hasGrid = 1;
    for (size_t i = 0; i < total; i++) grid_add(&p->grid, i, item_pos(p, i));
}

/* iteration over the neighbouring cells of a cell number: NeighboursView (NeighbourGrid.h:154-223) */
#define FOR_NEIGHBOUR_CELLS(g, cellNo, offs, nOffs, realCell, tr) \
    for (int o_ = 0; o_ < (nOffs); o_++) { \
        size_t ghost_ = (size_t) ((long) (cellNo) + (offs)[o_]); \
        size_t realCell = (g)->reflected[ghost_]; \
        vec3 tr = (g)->trans[(g)->transIdx[ghost_]];
#define END_FOR_NEIGHBOUR_CELLS }

/* Packing.cpp:373-429, 582-647 */
static size_t count_particle_overlaps(const rpo_packing *p, size_t orig, size_t tmp, int earlyExit) {
    const rp_interaction_desc *d = &p->desc;
    int k = inter_k(d), kk = inter_kk(d);
    size_t count = 0;
    if (p->hasGrid) {
        const grid_t *g = &p->grid;
        for (int c1 = 0; c1 < kk; c1++) {
            vec3 pos1 = k == 0 ? p->shapes[tmp].pos : p->cabs[tmp * k + c1];
            size_t cellNo = grid_position_to_cell(g, pos1, NULL);
            FOR_NEIGHBOUR_CELLS(g, cellNo, g->offsets, g->numOffsets, cell, tr)
                bc_t bc = {0, tr, NULL};
                for (size_t idx2 = g->heads[cell]; idx2 != LIST_END; idx2 = g->succ[idx2]) {
                    size_t j = idx2 / kk;
                    if (j == orig) continue;
                    if (overlap_between(d, pos1, p->shapes[tmp].o, c1, item_pos(p, idx2), p->shapes[j].o,
                                        (int) (idx2 % kk), &bc)) {
                        if (earlyExit) return 1;
                        count++;
                    }
                }
            END_FOR_NEIGHBOUR_CELLS
        }
    } else {
        bc_t bc = {1, V(0, 0, 0), &p->box};
        for (size_t j = 0; j < p->n; j++) {
            if (j == orig) continue;
            for (int c1 = 0; c1 < kk; c1++)
                for (int c2 = 0; c2 < kk; c2++) {
                    vec3 pos1 = k == 0 ? p->shapes[tmp].pos : p->cabs[tmp * k + c1];
                    vec3 pos2 = k == 0 ? p->shapes[j].pos : p->cabs[j * k + c2];
                    if (overlap_between(d, pos1, p->shapes[tmp].o, c1, pos2, p->shapes[j].o, c2, &bc)) {
                        if (earlyExit) return 1;
                        count++;
                    }
                }
        }
    }
    return count;
}

/* Packing.cpp:649-688, 711-764 */
static double particle_energy(const rpo_packing *p, size_t orig, size_t tmp) {
    const rp_interaction_desc *d = &p->desc;
    if (d->soft_kind == RP_SOFT_NONE) return 0;
    int k = inter_k(d), kk = inter_kk(d);
    double energy = 0;
    if (p->hasGrid) {
        const grid_t *g = &p->grid;
        for (int c1 = 0; c1 < kk; c1++) {
            double centreEnergy = 0;
            vec3 pos1 = k == 0 ? p->shapes[tmp].pos : p->cabs[tmp * k + c1];
            size_t cellNo = grid_position_to_cell(g, pos1, NULL);
            FOR_NEIGHBOUR_CELLS(g, cellNo, g->offsets, g->numOffsets, cell, tr)
                bc_t bc = {0, tr, NULL};
                for (size_t idx2 = g->heads[cell]; idx2 != LIST_END; idx2 = g->succ[idx2]) {
                    if (idx2 / kk == orig) continue;
                    centreEnergy += energy_between(d, pos1, item_pos(p, idx2), &bc);
                }
            END_FOR_NEIGHBOUR_CELLS
            if (k == 0) energy = centreEnergy; else energy += centreEnergy;
        }
    } else {
        bc_t bc = {1, V(0, 0, 0), &p->box};
        for (size_t j = 0; j < p->n; j++) {
            if (j == orig) continue;
            double e = 0;
            for (int c1 = 0; c1 < kk; c1++)
                for (int c2 = 0; c2 < kk; c2++) {
                    vec3 pos1 = k == 0 ? p->shapes[tmp].pos : p->cabs[tmp * k + c1];
                    vec3 pos2 = k == 0 ? p->shapes[j].pos : p->cabs[j * k + c2];
                    e += energy_between(d, pos1, pos2, &bc);
                }
            energy += e;
        }
    }
    return energy;
}

/* Packing.cpp:431-559 (overlaps) and 690-709, 766-839 (energy): half-shell enumeration */
static void total_recount(const rpo_packing *p, int earlyExit, int wantEnergy, size_t *overlapsOut, double *energyOut) {
    const rp_interaction_desc *d = &p->desc;
    int k = inter_k(d), kk = inter_kk(d);
    size_t overlaps = 0;
    double energy = 0;
    if (p->hasGrid) {
        const grid_t *g = &p->grid;
        for (size_t i = 0; i < g->div[0] - 2; i++)
            for (size_t j = 0; j < g->div[1] - 2; j++)
                for (size_t l = 0; l < g->div[2] - 2; l++) {
                    size_t co[3] = {i + 1, j + 1, l + 1};
                    size_t cellNo = grid_coords_to_cell(g, co);
                    double cellEnergy = 0;
                    for (size_t idx1 = g->heads[cellNo]; idx1 != LIST_END; idx1 = g->succ[idx1]) {
                        size_t p1 = idx1 / kk;
                        int c1 = (int) (idx1 % kk);
                        vec3 pos1 = item_pos(p, idx1);
                        bc_t none = {0, V(0, 0, 0), NULL};
                        for (size_t idx2 = g->succ[idx1]; idx2 != LIST_END; idx2 = g->succ[idx2]) {
                            size_t p2 = idx2 / kk;
                            if (k != 0 && p1 == p2) continue;
                            if (wantEnergy) cellEnergy += energy_between(d, pos1, item_pos(p, idx2), &none);
                            else if (overlap_between(d, pos1, p->shapes[p1].o, c1, item_pos(p, idx2), p->shapes[p2].o,
                                                     (int) (idx2 % kk), &none)) {
                                if (earlyExit) { *overlapsOut = 1; return; }
                                overlaps++;
                            }
                        }
                        FOR_NEIGHBOUR_CELLS(g, cellNo, g->posOffsets, g->numPosOffsets, cell, tr)
                            bc_t bc = {0, tr, NULL};
                            for (size_t idx2 = g->heads[cell]; idx2 != LIST_END; idx2 = g->succ[idx2]) {
                                size_t p2 = idx2 / kk;
                                if (k != 0 && p1 == p2) continue;
                                if (wantEnergy) cellEnergy += energy_between(d, pos1, item_pos(p, idx2), &bc);
                                else if (overlap_between(d, pos1, p->shapes[p1].o, c1, item_pos(p, idx2),
                                                         p->shapes[p2].o, (int) (idx2 % kk), &bc)) {
                                    if (earlyExit) { *overlapsOut = 1; return; }
                                    overlaps++;
                                }
                            }
                        END_FOR_NEIGHBOUR_CELLS
                    }
                    energy += cellEnergy;
                }
    } else {
        bc_t bc = {1, V(0, 0, 0), &p->box};
        for (size_t i = 0; i < p->n; i++)
            for (size_t j = i + 1; j < p->n; j++) {
                double e = 0;
                for (int c1 = 0; c1 < kk; c1++)
                    for (int c2 = 0; c2 < kk; c2++) {
                        vec3 pos1 = k == 0 ? p->shapes[i].pos : p->cabs[i * k + c1];
                        vec3 pos2 = k == 0 ? p->shapes[j].pos : p->cabs[j * k + c2];
                        if (wantEnergy) e += energy_between(d, pos1, pos2, &bc);
                        else if (overlap_between(d, pos1, p->shapes[i].o, c1, pos2, p->shapes[j].o, c2, &bc)) {
                            if (earlyExit) { *overlapsOut = 1; return; }
                            overlaps++;
                        }
                    }
                energy += e;
            }
    }
    if (overlapsOut) *overlapsOut = overlaps;
    if (energyOut) *energyOut = energy;
}

uint64_t rpo_count_total_overlaps(const rpo_packing *p, int early_exit) {
    if (p->desc.hard_kind == RP_HARD_NONE) return 0;
    size_t overlaps = 0;
    total_recount(p, early_exit, 0, &overlaps, NULL);
    return overlaps;
}
double rpo_total_energy(const rpo_packing *p) {
    if (p->desc.soft_kind == RP_SOFT_NONE) return 0;
    double e = 0;
    total_recount(p, 0, 1, NULL, &e);
    return e;
}

/* Packing.cpp:1000-1033 */
double rpo_average_neighbours(const rpo_packing *p) {
    int kk = inter_kk(&p->desc);
    size_t total = p->n * kk;
    if (!p->hasGrid) return (double) (total - 1);
    const grid_t *g = &p->grid;
    size_t num = 0;
    for (size_t idx1 = 0; idx1 < total; idx1++) {
        size_t cellNo = grid_position_to_cell(g, item_pos(p, idx1), NULL);
        FOR_NEIGHBOUR_CELLS(g, cellNo, g->offsets, g->numOffsets, cell, tr)
            (void) tr;
            for (size_t idx2 = g->heads[cell]; idx2 != LIST_END; idx2 = g->succ[idx2])
                if (idx2 / kk != idx1 / kk) num++;
        END_FOR_NEIGHBOUR_CELLS
    }
    return (double) num / (double) total;
}

static void set_error(char *err, size_t len, const char *msg) {
    if (err && len) { strncpy(err, msg, len - 1); err[len - 1] = 0; }
}

/* Packing::Packing / reset / setupForInteraction (Packing.cpp:46-82, 931-950) */
rpo_packing *rpo_packing_create(const rp_interaction_desc *desc, const double boxm[9], const double *shapes,
                                uint64_t n, char *err, size_t err_len)
{
    rpo_packing *p = calloc(1, sizeof(*p));
    p->desc = *desc;
    box_init(&p->box, boxm);
    if (box_volume(&p->box) == 0) { set_error(err, err_len, "newBox.getVolume() != 0"); free(p); return NULL; }
    if (!(rpo_range_radius(desc) > 0)) { set_error(err, err_len, "getRangeRadius() > 0"); free(p); return NULL; }
    if (n == 0) { set_error(err, err_len, "!newShapes.empty()"); free(p); return NULL; }
    for (uint64_t i = 0; i < n; i++) {      /* Packing::areShapesWithinBox (Packing.cpp:1067-1073) */
        vec3 rel = abs_to_rel(&p->box, V(shapes[12 * i], shapes[12 * i + 1], shapes[12 * i + 2]));
        for (int c = 0; c < 3; c++)
            if (rel.v[c] < 0 || rel.v[c] >= 1) {
                set_error(err, err_len, "areShapesWithinBox(newShapes, newBox)");
                free(p);
                return NULL;
            }
    }
    p->n = n;
    p->shapes = calloc(n + 1, sizeof(shape_t));
    p->lastShapes = calloc(n + 1, sizeof(shape_t));
    for (uint64_t i = 0; i < n; i++) {
        p->shapes[i].pos = V(shapes[12 * i], shapes[12 * i + 1], shapes[12 * i + 2]);
        memcpy(p->shapes[i].o, shapes + 12 * i + 3, 9 * sizeof(double));
    }
    p->interactionRange = rpo_range_radius(desc);
    int k = inter_k(desc);
    if (k > 0) {
        p->crel = calloc((n + 1) * k, sizeof(vec3));
        p->cabs = calloc((n + 1) * k, sizeof(vec3));
        p->lastCabs = calloc((n + 1) * k, sizeof(vec3));
        for (uint64_t i = 0; i < n; i++) {
            for (int c = 0; c < k; c++)
                p->crel[i * k + c] = mvec(p->shapes[i].o, V(desc->centres[c][0], desc->centres[c][1],
                                                            desc->centres[c][2]));
            recalc_abs_centres(p, i);
        }
    }
    rebuild_grid(p);
    return p;
}

void rpo_packing_destroy(rpo_packing *p) {
    if (!p) return;
    if (p->hasGrid) grid_free(&p->grid);
    if (p->tempHasGrid) grid_free(&p->tempGrid);
    free(p->shapes); free(p->lastShapes); free(p->crel); free(p->cabs); free(p->lastCabs);
    free(p);
}

void rpo_packing_get(const rpo_packing *p, double boxm[9], double *shapes) {
    memcpy(boxm, p->box.m, 9 * sizeof(double));
    for (size_t i = 0; i < p->n; i++) {
        memcpy(shapes + 12 * i, p->shapes[i].pos.v, 3 * sizeof(double));
        memcpy(shapes + 12 * i + 3, p->shapes[i].o, 9 * sizeof(double));
    }
}

int rpo_grid_divisions(const rpo_packing *p, uint64_t out[3]) {
    if (!p->hasGrid) return 0;
    for (int i = 0; i < 3; i++) out[i] = p->grid.div[i] - 2;
    return 1;
}

/* Packing::tryTranslation / tryRotation / tryMove + calculateMoveOverlapEnergy (Packing.cpp:84-164, 304-323) */
int rpo_try_move(rpo_packing *p, const rp_move *m, double *dE) {
    if (m->particle >= p->n) return 1;
    size_t i = m->particle, tmp = p->n;
    int k = inter_k(&p->desc);
    p->lastAltered = i;
    p->shapes[tmp] = p->shapes[i];
    int translated = m->kind == RP_MOVE_TRANSLATION || m->kind == RP_MOVE_ROTOTRANSLATION;
    int rotated = m->kind == RP_MOVE_ROTATION || m->kind == RP_MOVE_ROTOTRANSLATION;
    if (translated) {       /* Shape::translate (core/Shape.cpp:9-12) */
        p->shapes[tmp].pos = vadd(p->shapes[tmp].pos, V(m->translation[0], m->translation[1], m->translation[2]));
        p->shapes[tmp].pos = vadd(p->shapes[tmp].pos, pbc_correction(&p->box, p->shapes[tmp].pos));
    }
    if (rotated)            /* Shape::rotate (core/Shape.cpp:14-16) */
        mmul(m->rotation, p->shapes[tmp].o, p->shapes[tmp].o);
    for (int c = 0; c < k; c++) {
        p->crel[tmp * k + c] = p->crel[i * k + c];
        if (rotated) p->crel[tmp * k + c] = mvec(m->rotation, p->crel[tmp * k + c]);
    }
    if (k > 0) recalc_abs_centres(p, tmp);

    if (p->desc.hard_kind != RP_HARD_NONE) {
        if (p->overlapCounting) {
            size_t initial = count_particle_overlaps(p, i, i, 0);
            size_t final = count_particle_overlaps(p, i, tmp, 0);
            p->lastMoveOverlapDelta = (long) final - (long) initial;
            if (p->lastMoveOverlapDelta < 0) { *dE = -INFINITY; return 0; }
            if (p->lastMoveOverlapDelta > 0) { *dE = INFINITY; return 0; }
        } else if (count_particle_overlaps(p, i, tmp, 1) > 0) {
            *dE = INFINITY;
            return 0;
        }
    }
    double initialEnergy = particle_energy(p, i, i);
    double finalEnergy = particle_energy(p, i, tmp);
    *dE = finalEnergy - initialEnergy;
    return 0;
}

/* Packing::acceptMove (Packing.cpp:278-302) */
void rpo_accept_move(rpo_packing *p) {
    size_t i = p->lastAltered, tmp = p->n;
    int k = inter_k(&p->desc), kk = inter_kk(&p->desc);
    if (p->hasGrid)
        for (int c = 0; c < kk; c++) grid_remove(&p->grid, i * kk + c, item_pos(p, i * kk + c));
    p->shapes[i] = p->shapes[tmp];
    for (int c = 0; c < k; c++) {
        p->crel[i * k + c] = p->crel[tmp * k + c];
        p->cabs[i * k + c] = p->cabs[tmp * k + c];
    }
    if (p->hasGrid)
        for (int c = 0; c < kk; c++) grid_add(&p->grid, i * kk + c, item_pos(p, i * kk + c));
    if (p->overlapCounting) p->numOverlaps = (size_t) ((long) p->numOverlaps + p->lastMoveOverlapDelta);
}

/* Packing::isBoxUpscaled (Packing.cpp:1155-1176) */
static int is_box_upscaled(const box_t *oldBox, const box_t *newBox) {
    const double EPSILON = 100 * DBL_EPSILON;
    for (int i = 0; i < 3; i++) {
        vec3 e = V(i == 0, i == 1, i == 2);
        vec3 oldSide = rel_to_abs(oldBox, e), newSide = rel_to_abs(newBox, e);
        double d = vdot(oldSide, newSide);
        double norm2Old = vnorm2(oldSide), norm2New = vnorm2(newSide);
        if (fabs(d * d / norm2Old / norm2New - 1) > EPSILON) return 0;
        if (d * d / norm2Old / norm2Old < 1 + EPSILON) return 0;
    }
    return 1;
}

/* Packing::tryScaling (Packing.cpp:172-208) */
int rpo_try_scaling(rpo_packing *p, const double new_box[9], double *dE) {
    box_t nb;
    box_init(&nb, new_box);
    if (box_volume(&nb) == 0) return 1;
    int k = inter_k(&p->desc);
    p->lastBox = p->box;
    memcpy(p->lastShapes, p->shapes, (p->n + 1) * sizeof(shape_t));
    if (k > 0) memcpy(p->lastCabs, p->cabs, (p->n + 1) * k * sizeof(vec3));
    double initialEnergy = rpo_total_energy(p);
    p->lastScalingNumOverlaps = p->numOverlaps;
    p->box = nb;
    for (size_t i = 0; i < p->n; i++)
        p->shapes[i].pos = rel_to_abs(&p->box, abs_to_rel(&p->lastBox, p->shapes[i].pos));
    /* std::swap(neighbourGrid, tempNeighbourGrid) */
    grid_t tg = p->grid; p->grid = p->tempGrid; p->tempGrid = tg;
    int th = p->hasGrid; p->hasGrid = p->tempHasGrid; p->tempHasGrid = th;
    if (k > 0) for (size_t i = 0; i < p->n; i++) recalc_abs_centres(p, i);
    rebuild_grid(p);
    if (p->desc.hard_kind != RP_HARD_NONE) {
        if (p->overlapCounting) {
            p->numOverlaps = rpo_count_total_overlaps(p, 0);
            long delta = (long) p->numOverlaps - (long) p->lastScalingNumOverlaps;
            if (delta < 0) { *dE = -INFINITY; return 0; }
            if (delta > 0) { *dE = INFINITY; return 0; }
        } else {
            /* CompoundInteraction.cpp:24-31: a single convex hard part keeps the compound convex */
            int cannotOverlap = is_convex(&p->desc) && is_box_upscaled(&p->lastBox, &p->box);
            if (!cannotOverlap && rpo_count_total_overlaps(p, 1) > 0) { *dE = INFINITY; return 0; }
        }
    }
    *dE = rpo_total_energy(p) - initialEnergy;
    return 0;
}

/* Packing::revertScaling (Packing.cpp:363-371) */
void rpo_revert_scaling(rpo_packing *p) {
    int k = inter_k(&p->desc);
    p->box = p->lastBox;
    memcpy(p->shapes, p->lastShapes, (p->n + 1) * sizeof(shape_t));
    if (k > 0) memcpy(p->cabs, p->lastCabs, (p->n + 1) * k * sizeof(vec3));
    grid_t tg = p->grid; p->grid = p->tempGrid; p->tempGrid = tg;
    int th = p->hasGrid; p->hasGrid = p->tempHasGrid; p->tempHasGrid = th;
    p->numOverlaps = p->lastScalingNumOverlaps;
}

void rpo_toggle_overlap_counting(rpo_packing *p, int on) {
    p->overlapCounting = on != 0;
    if (p->overlapCounting) p->numOverlaps = rpo_count_total_overlaps(p, 0);
}
uint64_t rpo_cached_overlaps(const rpo_packing *p) { return p->numOverlaps; }

/* ------------------------------------------------------------------------------------------------------------------
 * cycle loop with one domain: Simulation::performCycle / tryMove / tryScaling (core/Simulation.cpp:325-531),
 * RototranslationSampler::sampleMove (core/move_samplers/RototranslationSampler.cpp:22-49),
 * TriclinicDeltaScaler::updateBox with independent sides (core/volume_scalers/TriclinicDeltaScaler.cpp:7-25)
 * ---------------------------------------------------------------------------------------------------------------- */
typedef struct { uint64_t s[4]; } rng_t;
static uint64_t splitmix(uint64_t *x) {
    uint64_t z = (*x += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t rng_next(rng_t *r) {     /* xoshiro256** */
    uint64_t res = rotl(r->s[1] * 5, 7) * 9, t = r->s[1] << 17;
    r->s[2] ^= r->s[0]; r->s[3] ^= r->s[1]; r->s[1] ^= r->s[2]; r->s[0] ^= r->s[3];
    r->s[2] ^= t; r->s[3] = rotl(r->s[3], 45);
    return res;
}
static double rng_uniform(rng_t *r, double a, double b) {
    return a + (b - a) * ((double) (rng_next(r) >> 11) * (1.0 / 9007199254740992.0));
}
static double now_seconds(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

int rpo_run_cycles(rpo_packing *p, uint64_t seed, uint64_t cycles, double trans_step, double rot_step,
                   double temperature, double pressure, int box_moves, double box_step, rpo_run_result *res)
{
    rng_t rng;
    uint64_t sm = seed;
    for (int i = 0; i < 4; i++) rng.s[i] = splitmix(&sm);
    memset(res, 0, sizeof(*res));
    for (uint64_t cycle = 0; cycle < cycles; cycle++) {
        double t0 = now_seconds();
        double h[3];
        box_heights(&p->box, h);
        double maxStep = fmin(h[0], fmin(h[1], h[2])) / 2;
        double ts = fmin(trans_step, maxStep), rs = fmin(rot_step, M_PI);
        for (size_t a = 0; a < p->n; a++) {
            rp_move m;
            memset(&m, 0, sizeof(m));
            m.kind = RP_MOVE_ROTOTRANSLATION;
            for (int c = 0; c < 3; c++) m.translation[c] = rng_uniform(&rng, -ts, ts);
            vec3 axis;
            do {
                axis = V(rng_uniform(&rng, -1, 1), rng_uniform(&rng, -1, 1), rng_uniform(&rng, -1, 1));
            } while (vnorm2(axis) > 1);
            double angle = rng_uniform(&rng, -rs, rs);
            double nrm = sqrt(vnorm2(axis));
            double ax[3] = {axis.v[0] / nrm, axis.v[1] / nrm, axis.v[2] / nrm};
            rpo_rotation_matrix(ax, angle, m.rotation);
            m.particle = rng_next(&rng) % p->n;
            double dE;
            rpo_try_move(p, &m, &dE);
            res->moves_attempted++;
            if (rng_uniform(&rng, 0, 1) <= exp(-dE / temperature)) {     /* Simulation.cpp:483 */
                rpo_accept_move(p);
                res->moves_accepted++;
            }
        }
        double t1 = now_seconds();
        res->move_seconds += t1 - t0;
        if (box_moves) {
            double nb[9];
            memcpy(nb, p->box.m, sizeof(nb));
            int side = (int) (rng_next(&rng) % 3);
            for (int i = 0; i < 3; i++) nb[3 * i + side] += rng_uniform(&rng, -box_step, box_step);
            box_t newBox;
            box_init(&newBox, nb);
            double oldV = fabs(box_volume(&p->box)), newV = fabs(box_volume(&newBox));
            double deltaV = newV - oldV, factor = newV / oldV, N = (double) p->n;
            res->scalings_attempted++;
            if (p->desc.soft_kind != RP_SOFT_NONE || p->overlapCounting) {
                double dE;
                rpo_try_scaling(p, nb, &dE);
                double exponent = N * log(factor) - dE / temperature - pressure * deltaV / temperature;
                if (rng_uniform(&rng, 0, 1) <= exp(exponent)) res->scalings_accepted++;
                else rpo_revert_scaling(p);
            } else {
                double exponent = N * log(factor) - pressure * deltaV / temperature;
                double dE;
                if (rng_uniform(&rng, 0, 1) > exp(exponent)) {
                } else if (rpo_try_scaling(p, nb, &dE), dE > 0) {
                    rpo_revert_scaling(p);
                } else {
                    res->scalings_accepted++;
                }
            }
            res->scaling_seconds += now_seconds() - t1;
        }
    }
    return 0;
}
