// rp_sweep_generic.cuh — the checkerboard sub-sweep for EVERY interaction the device supports (multi-centre shapes,
// soft potentials with the Metropolis rule, compound hard + soft, overlap-counting mode); included by rp_sweep.cu.
// Hard spheres / spherocylinders outside counting mode take the specialised k_sweep (rp_sweep_kernel2.cuh) instead.
//
// Same concurrency rule as k_sweep (random grid origin, cells >= totalRange wide, 8 parity colours, one CTA per
// active cell, moves leaving the cell rejected: DomainDecomposition.cpp:50-75, Packing.cpp:95-96), but
//   * a record carries the whole particle: mass centre, (orientation), the k interaction centres in the image frame
//     (Packing::absoluteInteractionCentres without the wrap), aux vectors, (reference-grid cell of every centre);
//   * phase 1 (parallel): one warp per trial scans the FROZEN candidates (the 26 cells around the active cell,
//     read from the cell-sorted records through L2) and accumulates, per trial, the overlap counts and energies of
//     the old and of the new state — Packing::countParticleOverlaps / calculateParticleEnergy
//     (Packing.cpp:373-429, 649-688) restricted to frozen neighbours; lanes reduce in a fixed order (deterministic);
//   * phase 2 (sequential, random order): one warp walks the trials of the cell; the pair terms with the OTHER
//     particles of the cell are evaluated against their CURRENT state (shared memory), then the reference's
//     decision is taken: Packing::calculateMoveOverlapEnergy (Packing.cpp:304-323) -> +-inf / energy difference,
//     Metropolis u <= exp(-dE/T) (Simulation.cpp:483); an accepted trial replaces the current state at once.
//   This is exactly a sequential Monte Carlo pass through the cell.
//
// Un-truncated Lennard-Jones (LennardJonesInteraction.cpp:8-13 has no cutoff test): the reference sums over whatever
// lies in the 27 neighbour-grid cells around a centre, so the pair set is defined by its grid.  The records carry the
// (extended) reference-grid cell of every centre and a pair contributes iff the two cells are adjacent — the same
// pair set as Packing::calculateInteractionCentreEnergyWithNG; the sweep cells are made wide enough to see all of it.
#pragma once

#include "rp_internal.cuh"
#include "rp_philox.cuh"
#include "rp_wall.cuh"

namespace rp {

struct GenLayout {
    int recw;        // doubles per record (even: 16-byte multiples)
    int offOrient;   // 9 doubles, -1 = absent
    int offCentre;   // kk x 3 doubles; k == 0: 0 (the centre is the mass centre)
    int offAux;      // kk x 3 doubles, -1 = absent
    int offCell;     // kk x {int x, y, z, 0} (two doubles per centre, 16-byte aligned), -1 = absent
    int kk;
};

struct GenArgs {
    int kind;
    double transStep, rotStep, temperature;
    unsigned long long seed;
    unsigned int cycle, pass;
    int colour[3], nActive[3];
    int capOwn, capOld;      // shared-memory capacities: particles of the active cell / frozen candidates around it
    float cull32Thr2;        // FP32 cull of mass-centre distances (conservative: cull range + rounding); LJ: unused
    int cullBox;             // un-truncated LJ: the cull points hold RELATIVE coordinates and the cull is the box
    float cullBoxHalf[3];    //   |d rel_i| < cullBoxHalf[i] (two reference cells + the molecules' extent + rounding)
    int counting;            // Packing::overlapCounting
    int matrices;            // in-cell pair terms as matrices (counting / soft / > 64 particles per cell) or bit masks
    // RP_MOVE_FLIP (FlipSampler.cpp:18-41): body-frame flip axis, geometric origin, attempt probability 1 / flipEvery
    double flipAxis[3], flipOrigin[3], flipProb;
    int flipOriginZero;
    double cullRange2;       // (totalRange * (1 + 1e-9))^2: mass centres farther apart cannot interact; LJ: +inf
    int refN[3];             // LJ: real cells of the reference-geometry grid
    double refRelCell[3];
    size_t cap, ccap;
    GenLayout L;
};

struct GenTrialBuffers {
    double *rec;       // recw doubles per slot: the trial state as a record
    double *orient;    // 9 per slot: new orientation
    double *crel;      // 3 * k per slot: new Packing::interactionCentres of the particle
    double *u;         // Metropolis variate
    double *dE;        // written by k_gen_sweep: the value Packing::tryMove would have returned (rp_sweep_trace)
    uint2 *meta;       // x: 0 alive / 2 left the cell, y: order key
    int2 *wall;        // wall overlaps of the {old, new} state (Packing::countParticleWallOverlaps); nullptr: no walls
    float4 *cull;      // FP32 cull point of the trial state (mass centre relative to the centre of the own cell)
};

__device__ __forceinline__ V3 ld3(const double *p) { return v3(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(double *p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// extended reference-grid cell of a position given in the image frame (NeighbourGrid::positionToCellNo without the
// wrap: floor(rel / relativeCellSize), may be < 0 or >= n)
__device__ __forceinline__ int4 refCellOf(const Box &box, const double refRelCell[3], V3 p) {
    V3 rel = absToRel(box, p);
    return make_int4(static_cast<int>(floor(rel.x / refRelCell[0])), static_cast<int>(floor(rel.y / refRelCell[1])),
                     static_cast<int>(floor(rel.z / refRelCell[2])), 0);
}
// a = cell of the first centre minus the lattice translation's (gx n0, gy n1, gz n2); b = cell of the second centre
__device__ __forceinline__ bool refCellsAdjacent(const int4 &a, const int4 &b) {
    return static_cast<unsigned int>(a.x - b.x + 1) <= 2u && static_cast<unsigned int>(a.y - b.y + 1) <= 2u
           && static_cast<unsigned int>(a.z - b.z + 1) <= 2u;
}

// fills the derived parts of a record from {mass centre p, orientation o, centre offsets crel[k]}
__device__ __forceinline__ void fillRecord(const Inter &I, const GenArgs &A, const Box &box, double *r, V3 p,
                                           const M33 &o, const V3 *crel)
{
    const GenLayout &L = A.L;
    st3(r, p);
    if (L.offOrient >= 0) for (int q = 0; q < 9; q++) r[L.offOrient + q] = o.a[q];
    for (int c = 0; c < I.k; c++) st3(r + L.offCentre + 3 * c, p + crel[c]);
    if (L.offAux >= 0) for (int c = 0; c < L.kk; c++) st3(r + L.offAux + 3 * c, pairAux(I, o, c));
    if (L.offCell >= 0)
        for (int c = 0; c < L.kk; c++)
            reinterpret_cast<int4 *>(r + L.offCell)[c] = refCellOf(box, A.refRelCell, ld3(r + L.offCentre + 3 * c));
}

// FP32 cull point of a particle state: mass centre relative to the centre of the particle's own sweep cell (box cull
// of the un-truncated LJ potential: in relative coordinates)
__device__ __forceinline__ float4 genCullPoint(const GenArgs &A, const Box &box, const V3 &p, const V3 &centre) {
    V3 d = p - centre;
    if (A.cullBox) d = absToRel(box, d);
    return make_float4(static_cast<float>(d.x), static_cast<float>(d.y), static_cast<float>(d.z), 0.f);
}

// one warp per cell: items by ascending particle index + the generic records (image frame) + the FP32 cull points
__global__ void k_gen_records(int numCells, GenArgs A, SweepGrid sg, Box box, const Inter *__restrict__ interPtr,
                              const int *__restrict__ cellStart, const int *__restrict__ itemsTmp,
                              const int *__restrict__ keys, const double *__restrict__ pos,
                              const double *__restrict__ orient, const double *__restrict__ crelAll,
                              int *__restrict__ items, double *__restrict__ rec, float4 *__restrict__ cull)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();
    int cell = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (cell >= numCells) return;
    int start = cellStart[cell], cnt = cellStart[cell + 1] - start;
    const V3 centre = sweepCellCentre(sg, box, cell % sg.n[0], (cell / sg.n[0]) % sg.n[1], cell / (sg.n[0] * sg.n[1]));
    for (int i = lane; i < cnt; i += 32) {
        int id = itemsTmp[start + i];
        int rank = 0;
        for (int j = 0; j < cnt; j++) rank += (itemsTmp[start + j] < id);
        int slot = start + rank;
        items[slot] = id;
        int f = (keys[id] >> 28) & 7;
        V3 p = v3(pos[id], pos[A.cap + id], pos[2 * A.cap + id]);
        if (f) p = p + sg.imageShift[f];
        M33 o;
        for (int q = 0; q < 9; q++) o.a[q] = orient[q * A.cap + id];
        V3 crel[RP_MAX_CENTRES];
        for (int c = 0; c < I.k; c++) {
            size_t idx = static_cast<size_t>(id) * I.k + c;
            crel[c] = v3(crelAll[idx], crelAll[A.ccap + idx], crelAll[2 * A.ccap + idx]);
        }
        fillRecord(I, A, box, rec + static_cast<size_t>(slot) * A.L.recw, p, o, crel);
        cull[slot] = genCullPoint(A, box, p, centre);
    }
}

// Matrix<3, 3>::rotation(axis, angle): identity + K*sina + ((1 - cosa)*K)*K (geometry/Matrix.tpp:437-449)
__device__ __forceinline__ M33 rodrigues(V3 axis, double angle) {
    double sina = sin(angle), cosa = cos(angle);
    M33 K = {{0, -axis.z, axis.y, axis.z, 0, -axis.x, -axis.y, axis.x, 0}};
    M33 sK;
    for (int q = 0; q < 9; q++) sK.a[q] = (1 - cosa) * K.a[q];
    M33 KK = mul(sK, K);
    M33 R;
    for (int q = 0; q < 9; q++) R.a[q] = (((q % 4 == 0) ? 1.0 : 0.0) + K.a[q] * sina) + KK.a[q];
    return R;
}

// MoveSampler rotation: axis uniform in the unit ball by rejection then normalised, angle U(-a, a)
// (RototranslationSampler.cpp:35-43), Rodrigues matrix identity + K*sina + ((1 - cosa)*K)*K (Matrix.tpp:437-449)
__device__ __forceinline__ M33 drawRotation(Rng &rng, double rotStep) {
    V3 axis;
    double n2;
    do {
        axis = v3(rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1));
        n2 = norm2(axis);
    } while (n2 > 1 || n2 == 0);
    double nrm = sqrt(n2);
    axis = v3(axis.x / nrm, axis.y / nrm, axis.z / nrm);
    double angle = rng.uniform(-rotStep, rotStep);
    return rodrigues(axis, angle);
}

// trial moves of all particles (one thread per sorted slot): Shape::translate / rotate (Shape.cpp:9-16),
// prepareTemp / rotateTempInteractionCentres (Packing.cpp:339-361), ActiveDomain rule with the cell as the domain
__global__ void __launch_bounds__(128) k_gen_trials(int n, GenArgs A, SweepGrid sg, Box box,
                                                    const Inter *__restrict__ interPtr,
                                                    const int *__restrict__ items, const int *__restrict__ keys,
                                                    const double *__restrict__ rec,
                                                    const double *__restrict__ orient,
                                                    const double *__restrict__ crelAll, GenTrialBuffers T,
                                                    Walls W, const double *__restrict__ posAll)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();
    const int slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= n) return;
    const int id = items[slot];
    const int cell = keys[id] & 0x0FFFFFFF;
    const int cc[3] = {cell % sg.n[0], (cell / sg.n[0]) % sg.n[1], cell / (sg.n[0] * sg.n[1])};
    V3 p = ld3(rec + static_cast<size_t>(slot) * A.L.recw);
    const bool flip = A.kind == RP_MOVE_FLIP;
    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION || flip);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION
                             || (flip && !A.flipOriginZero));
    Rng rng;
    rng.init(A.seed, static_cast<unsigned int>(id), A.cycle, A.pass);
    unsigned int rej = 0;
    M33 O;
    for (int q = 0; q < 9; q++) O.a[q] = orient[q * A.cap + id];
    M33 R = {{1, 0, 0, 0, 1, 0, 0, 0, 1}};
    if (flip) {
        // the particle attempts a flip with probability 1 / flipEvery; the others sit this sampler out
        if (!(rng.uniform() < A.flipProb)) rej = 5;
        // shapeFlipAxis = (orientation * flipAxis).normalized(); rotation(shapeFlipAxis, M_PI) (FlipSampler.cpp:27-29)
        V3 ax = mul(O, v3(A.flipAxis[0], A.flipAxis[1], A.flipAxis[2]));
        double nrm = sqrt(norm2(ax));
        ax = v3(ax.x / nrm, ax.y / nrm, ax.z / nrm);
        R = rodrigues(ax, 3.14159265358979323846);
    }
    if (translated) {
        V3 shift;
        if (flip) {
            // translation = -rotation * shapeGeometricOrigin + shapeGeometricOrigin (FlipSampler.cpp:34-36)
            V3 g = mul(O, v3(A.flipOrigin[0], A.flipOrigin[1], A.flipOrigin[2]));
            V3 rg = mul(R, g);
            shift = v3(-rg.x, -rg.y, -rg.z) + g;
        } else {
            double tx = rng.uniform(-A.transStep, A.transStep);
            double ty = rng.uniform(-A.transStep, A.transStep);
            double tz = rng.uniform(-A.transStep, A.transStep);
            shift = v3(tx, ty, tz);
        }
        p = p + shift;
        V3 rel = absToRel(box, p);
        double q[3] = {rel.x - sg.shift[0], rel.y - sg.shift[1], rel.z - sg.shift[2]};
        for (int d = 0; d < 3; d++) {
            if (rej == 5) break;
            if (q[d] < 0 || q[d] >= 1) rej = 2;
            else if (static_cast<int>(q[d] / sg.relCell[d]) != cc[d]) rej = 2;
        }
    }
    V3 crel[RP_MAX_CENTRES];
    for (int c = 0; c < I.k; c++) {
        size_t idx = static_cast<size_t>(id) * I.k + c;
        crel[c] = v3(crelAll[idx], crelAll[A.ccap + idx], crelAll[2 * A.ccap + idx]);
    }
    int wallOld = 0;
    if (T.wall && A.counting && I.hard != RP_HARD_NONE) {
        // the state the particle is in now, on its stored (wrapped) position
        V3 cp = v3(posAll[id], posAll[A.cap + id], posAll[2 * A.cap + id]);
        wallOld = static_cast<int>(particleWallOverlaps(I, W, cp, O, [&](int c) { return crel[c]; }));
    }
    if (rotated) {
        if (!flip) R = drawRotation(rng, A.rotStep);
        O = mul(R, O);
        for (int c = 0; c < I.k; c++) crel[c] = mul(R, crel[c]);
    }
    if (T.wall) {
        int wallNew = 0;
        if (rej == 0 && I.hard != RP_HARD_NONE) {
            V3 cp = p + pbcCorrection(box, p);       // the wrapped position the write-back would store
            wallNew = static_cast<int>(particleWallOverlaps(I, W, cp, O, [&](int c) { return crel[c]; }));
        }
        T.wall[slot] = make_int2(wallOld, wallNew);
    }
    fillRecord(I, A, box, T.rec + static_cast<size_t>(slot) * A.L.recw, p, O, crel);
    T.cull[slot] = genCullPoint(A, box, p, sweepCellCentre(sg, box, cc[0], cc[1], cc[2]));
    double *o = T.orient + 9 * static_cast<size_t>(slot);
    for (int q = 0; q < 9; q++) o[q] = O.a[q];
    for (int c = 0; c < I.k; c++) st3(T.crel + (static_cast<size_t>(slot) * I.k + c) * 3, crel[c]);
    rng.ctr.w = 0x40000000u;     // separate sub-streams: processing order, Metropolis variate
    T.meta[slot] = make_uint2(rej, philox4x32(rng.ctr, rng.k0, rng.k1).x);
    rng.ctr.w = 0x40000001u;
    rng.have = 0;
    T.u[slot] = rng.uniform();
}

// Pair terms between two particles given as records: `a` (own, untranslated) and `b` (+ lattice translation t with
// integer lattice coordinates g).  Adds the number of overlapping centre pairs to *ov (stops at the first one when
// firstOnly) and the pair energies to *en.  Loops like Packing::countParticleOverlaps / calculateParticleEnergy.
// HARD / SOFT are compile-time constants of the calling kernel: one instantiation carries one pair function.
template <int HARD, bool SOFT>
__device__ __forceinline__ void pairTerms(const Inter &I, const GenArgs &A, const double *a, const double *b, V3 t,
                                          int gx, int gy, int gz, bool wantOv, bool wantEn, bool firstOnly,
                                          int *ov, double *en)
{
    const GenLayout &L = A.L;
    V3 pa = ld3(a), pb = ld3(b) + t;
    if (!(norm2(pb - pa) < A.cullRange2)) return;
    BcFixed bc{t};
    constexpr bool HAS_AUX = HARD == RP_HARD_SPHEROCYLINDER || HARD == RP_HARD_POLYSPHEROCYLINDER;
    for (int c1 = 0; c1 < L.kk; c1++) {
        V3 pos1 = ld3(a + L.offCentre + 3 * c1);
        V3 aux1 = HAS_AUX ? ld3(a + L.offAux + 3 * c1) : v3(0, 0, 0);
        int4 cell1 = make_int4(0, 0, 0, 0);
        if (SOFT && L.offCell >= 0) {
            cell1 = reinterpret_cast<const int4 *>(a + L.offCell)[c1];
            cell1.x -= gx * A.refN[0]; cell1.y -= gy * A.refN[1]; cell1.z -= gz * A.refN[2];
        }
        for (int c2 = 0; c2 < L.kk; c2++) {
            V3 pos2 = ld3(b + L.offCentre + 3 * c2);
            if (HARD != RP_HARD_NONE && wantOv) {
                bool hit;
                if (HARD == RP_HARD_XENOCOLLIDE) {
                    hit = overlapBetweenOriented(
                        I, pos1, [&] { M33 o; for (int q = 0; q < 9; q++) o.a[q] = a[L.offOrient + q]; return o; }, c1,
                        pos2, [&] { M33 o; for (int q = 0; q < 9; q++) o.a[q] = b[L.offOrient + q]; return o; }, c2, bc);
                } else {
                    V3 aux2 = HAS_AUX ? ld3(b + L.offAux + 3 * c2) : v3(0, 0, 0);
                    hit = overlapBetween<HARD>(I, pos1, aux1, c1, pos2, aux2, c2, bc);
                }
                if (hit) {
                    (*ov)++;
                    if (firstOnly) return;
                }
            }
            if (SOFT && wantEn) {
                if (L.offCell >= 0 && !refCellsAdjacent(cell1, reinterpret_cast<const int4 *>(b + L.offCell)[c2]))
                    continue;
                *en += energyBetween(I, pos1, pos2, bc);
            }
        }
    }
}

// Soft-only interactions (no hard part), phase 1: the energies of the trial (new) state `ta` and of the current (old)
// state `oa` of one particle against the frozen candidate `b`, with ONE pass over b's record: its centres (and
// reference-grid cells) are fetched from global memory four at a time before anything depends on them — the exact
// stage of the un-truncated LJ sweep is bound by the latency of exactly these loads — and serve both states.
// Same pair set and pair energies as two pairTerms calls; only the order of the additions differs.
template <bool CELLS>
__device__ __forceinline__ void softPairBoth(const Inter &I, const GenArgs &A, const double *ta, const double *oa,
                                             const double *b, V3 t, int gx, int gy, int gz, double *enN, double *enO)
{
    const GenLayout &L = A.L;
    const V3 pb = ld3(b) + t;
    const bool nearN = norm2(pb - ld3(ta)) < A.cullRange2, nearO = norm2(pb - ld3(oa)) < A.cullRange2;
    if (!nearN && !nearO) return;
    BcFixed bc{t};
    const int shift[3] = {gx * A.refN[0], gy * A.refN[1], gz * A.refN[2]};
    for (int c20 = 0; c20 < L.kk; c20 += 4) {
        V3 pos2[4];
        int4 cell2[4];
        #pragma unroll
        for (int u = 0; u < 4; u++) {
            const int c2 = min(c20 + u, L.kk - 1);
            pos2[u] = ld3(b + L.offCentre + 3 * c2);
            if (CELLS) {
                cell2[u] = reinterpret_cast<const int4 *>(b + L.offCell)[c2];
                cell2[u].x += shift[0]; cell2[u].y += shift[1]; cell2[u].z += shift[2];
            }
        }
        for (int c1 = 0; c1 < L.kk; c1++) {
            const V3 p1n = ld3(ta + L.offCentre + 3 * c1), p1o = ld3(oa + L.offCentre + 3 * c1);
            int4 c1n = make_int4(0, 0, 0, 0), c1o = c1n;
            if (CELLS) {
                c1n = reinterpret_cast<const int4 *>(ta + L.offCell)[c1];
                c1o = reinterpret_cast<const int4 *>(oa + L.offCell)[c1];
            }
            #pragma unroll
            for (int u = 0; u < 4; u++) {
                if (c20 + u >= L.kk) break;
                if (nearN && (!CELLS || refCellsAdjacent(c1n, cell2[u]))) *enN += energyBetween(I, p1n, pos2[u], bc);
                if (nearO && (!CELLS || refCellsAdjacent(c1o, cell2[u]))) *enO += energyBetween(I, p1o, pos2[u], bc);
            }
        }
    }
}

constexpr int GEN_THREADS = 128;     // default CTA; crowded cells (block-wise sequential mode) take GEN_THREADS_WIDE
constexpr int GEN_THREADS_WIDE = 512;
constexpr int GEN_XC_MINB = 4;       // XenoCollide: register cap 128 (MPR wants 255: 2 CTAs per SM, latency bound)
constexpr int GEN_MAXOWN = 64;       // trials per cell the bit masks of the hard-only resolution cover
constexpr int GEN_MAXCELL = 512;     // particles per sweep cell the kernel supports at all

// shared memory of k_gen_sweep beyond its static part, for (capOld frozen candidates, capOwn particles in the cell)
__host__ __device__ inline size_t genSmemBytes(int capOld, int capOwn, int recw, bool matrices,
                                               int threads = GEN_THREADS) {
    capOwn = (capOwn + 1) & ~1;                                             // (keeps the double arrays 8-byte aligned)
    size_t b = 16 * static_cast<size_t>(capOld + 2 * capOwn);               // FP32 cull points: candidates, trial + current states
    b += 2 * static_cast<size_t>(capOwn) * recw * 8;                        // cur, tri
    b += 3 * static_cast<size_t>(capOwn) * 8;                               // eOld, eNew, uMet
    b += 5 * static_cast<size_t>(capOwn) * 4;                               // ovOld, ovNew, rejected, seq, orderKey
    b += (threads / 32) * 64 * 4;                                           // per-warp survivor queues
    if (matrices) b += 4 * static_cast<size_t>(capOwn) * capOwn * (8 + 2);  // in-cell pair matrices
    return b + 64;
}

// ---------------------------------------------------------------------------------------------------------------
// k_gen_sweep<HARD, SOFT> — one CTA per active cell.
//   staging   the FP32 cull points (mass centre relative to the own cell's centre, 16 B) of the 26 frozen cells land
//             in shared memory by 1-D TMA bulk copies; the FP64 records stay in global memory (L2) and are fetched
//             for survivors only; the own cell's records (current and trial states) are copied to shared memory
//   phase 1   one warp per trial: FP32 distance cull of all staged candidates (32 per iteration), survivors are
//             compacted (ballot) into the warp's queue and take the exact pair terms 32 at a time — every lane busy;
//             per-trial sums reduce in a fixed order (deterministic)
//   phase 2a  pair terms inside the cell for every ordered pair (i, j): trial / current state of i against the
//             current and the trial state of j, all threads in parallel
//             hard-only: two bit masks per trial; otherwise small matrices of counts and energies
//   phase 2b  sequential pass in the cell's random order: what used to evaluate pair functions one trial after the
//             other is now a walk over the precomputed tables (the state of j a term is taken from follows `moved`)
// The decisions are those of a sequential Monte Carlo pass through the cell, as before.
// ---------------------------------------------------------------------------------------------------------------
template <int HARD, bool SOFT, int THREADS>
__global__ void __launch_bounds__(THREADS, (HARD == RP_HARD_XENOCOLLIDE ? GEN_XC_MINB : 1))
k_gen_sweep(GenArgs A, SweepGrid sg, SweepCellOffsets off, Box box, const Inter *__restrict__ interPtr,
            const int *__restrict__ cellStart, const int *__restrict__ items, double *__restrict__ rec,
            float4 *__restrict__ cull, GenTrialBuffers T, double *__restrict__ pos, double *__restrict__ orient,
            double *__restrict__ crelAll, double *__restrict__ cabsAll, unsigned long long *__restrict__ counters,
            int *__restrict__ flags)
{
    constexpr int WARPS = THREADS / 32;
    constexpr bool hard = HARD != RP_HARD_NONE, soft = SOFT;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ Inter I;
    __shared__ int cellBegin[28], cellSrc[27], cellT[27];
    __shared__ int sStats[5];
    __shared__ long long sOverlapDelta;
    __shared__ unsigned long long sBar;
    __shared__ unsigned long long ownMask[GEN_MAXOWN], newMask[GEN_MAXOWN];
    __shared__ unsigned long long sMoved[GEN_MAXCELL / 64];      // accepted trials of the cell, one bit each
    __shared__ double sRedE[2 * (THREADS / 32)];
    __shared__ int sRedO[2 * (THREADS / 32)];
    __shared__ int sAccept;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const GenLayout &L = A.L;
    const int recw = L.recw;
    for (int i = tid; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += THREADS)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];

    const bool counting = A.counting != 0 && hard;
    const bool needOld = counting || soft;
    // in-cell pair terms: 0 = bit masks (hard only, <= 64 particles per cell), 1 = matrices (counting / soft), 2 = no
    // tables: the sequential pass evaluates them against the current states with the whole CTA (cells too crowded for
    // the matrices, e.g. the wide cells of the un-truncated Lennard-Jones potential)
    const bool matrices = A.matrices == 1, blockwise = A.matrices == 2;

    const int b = blockIdx.x;
    const int a0 = b % A.nActive[0], a1 = (b / A.nActive[0]) % A.nActive[1], a2 = b / (A.nActive[0] * A.nActive[1]);
    const int cx = 2 * a0 + A.colour[0], cy = 2 * a1 + A.colour[1], cz = 2 * a2 + A.colour[2];
    const int ownCell = (cz * sg.n[1] + cy) * sg.n[0] + cx;
    const int ownStart = cellStart[ownCell], numOwn = cellStart[ownCell + 1] - ownStart;
    if (numOwn == 0) return;

    // dynamic shared memory
    float4 *cand = reinterpret_cast<float4 *>(smemRaw);
    float4 *trialPt = cand + A.capOld;
    float4 *curPt = trialPt + A.capOwn;
    double *cur = reinterpret_cast<double *>(curPt + A.capOwn);
    double *tri = cur + static_cast<size_t>(A.capOwn) * recw;
    double *eOld = tri + static_cast<size_t>(A.capOwn) * recw;
    double *eNew = eOld + A.capOwn;
    double *uMet = eNew + A.capOwn;
    int *ovOld = reinterpret_cast<int *>(uMet + A.capOwn);
    int *ovNew = ovOld + A.capOwn;
    int *rejected = ovNew + A.capOwn;
    int *seq = rejected + A.capOwn;
    unsigned int *orderKey = reinterpret_cast<unsigned int *>(seq + A.capOwn);
    int *queue = reinterpret_cast<int *>(orderKey + A.capOwn);
    // in-cell matrices, row i = trial, column j = other particle: [0] new_i/old_j [1] old_i/old_j [2] new_i/new_j
    // [3] old_i/new_j
    double *eMat = reinterpret_cast<double *>(queue + WARPS * 64);
    short *cMat = reinterpret_cast<short *>(eMat + 4 * static_cast<size_t>(A.capOwn) * A.capOwn);
    const size_t matStride = static_cast<size_t>(A.capOwn) * A.capOwn;

    // ---- cell table + staging (warp 0) ----
    if (warp == 0) {
        int cnt = 0, src = 0, tIdx = 13;
        if (lane < 27) {
            const int c2 = neighbourCell(sg.n, cx, cy, cz, lane % 3 - 1, (lane / 3) % 3 - 1, lane / 9 - 1, &tIdx);
            src = cellStart[c2];
            cnt = lane == 13 ? 0 : cellStart[c2 + 1] - src;      // the own cell is not frozen
        }
        if (lane == 0) {
            sStats[0] = sStats[1] = sStats[2] = sStats[3] = sStats[4] = 0;
            sOverlapDelta = 0;
            mbarInit(&sBar, 1);
        }
        int incl = cnt;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 26);
        const bool ok = total <= A.capOld && numOwn <= A.capOwn;
        if (lane == 0 && ok) mbarExpectTx(&sBar, static_cast<unsigned int>((total + 2 * numOwn) * 16));
        __syncwarp();
        if (lane < 27) {
            cellBegin[lane] = incl - cnt;
            cellT[lane] = tIdx;
            cellSrc[lane] = src;
            if (lane == 26) cellBegin[27] = incl;
            if (ok && cnt > 0)
                bulkCopyG2S(cand + (incl - cnt), cull + src, static_cast<unsigned int>(cnt * 16), &sBar);
        } else if (lane == 27 && ok) {
            bulkCopyG2S(trialPt, T.cull + ownStart, static_cast<unsigned int>(numOwn * 16), &sBar);
        } else if (lane == 28 && ok) {
            bulkCopyG2S(curPt, cull + ownStart, static_cast<unsigned int>(numOwn * 16), &sBar);
        }
    }
    __syncthreads();
    const int numFrozen = cellBegin[27];
    if (numFrozen > A.capOld || numOwn > A.capOwn) {
        if (tid == 0) atomicOr(&flags[3], 1);
        return;
    }
    // own cell: current and trial records, trial meta
    for (int i = tid; i < numOwn * recw; i += THREADS) {
        cur[i] = rec[static_cast<size_t>(ownStart) * recw + i];
        tri[i] = T.rec[static_cast<size_t>(ownStart) * recw + i];
    }
    for (int i = tid; i < numOwn; i += THREADS) {
        const uint2 meta = T.meta[ownStart + i];
        rejected[i] = static_cast<int>(meta.x);
        orderKey[i] = meta.y;
        uMet[i] = T.u[ownStart + i];
        if (i < GEN_MAXOWN) { ownMask[i] = 0ull; newMask[i] = 0ull; }
        if (i < GEN_MAXCELL / 64) sMoved[i] = 0ull;
    }
    if (warp == 0) mbarWait(&sBar, 0);
    __syncthreads();
    // frozen cull points move into the frame of the active cell (centre-to-centre vector incl. the periodic image)
    for (int c = warp; c < 27; c += WARPS) {
        const int begin = cellBegin[c], cnt = cellBegin[c + 1] - begin;
        const float ox = off.o[c][0], oy = off.o[c][1], oz = off.o[c][2];
        for (int j = begin + lane; j < begin + cnt; j += 32) {
            float4 r = cand[j];
            r.x += ox; r.y += oy; r.z += oz;
            cand[j] = r;
        }
    }
    __syncthreads();

    // ---- phase 1: frozen neighbours, one warp per trial ----
    {
        int *q = queue + warp * 64;
        const unsigned int ltMask = (1u << lane) - 1u;
        for (int i = warp; i < numOwn; i += WARPS) {
            int oO = 0, oN = 0;
            double enO = 0, enN = 0;
            if (rejected[i] == 0) {
                const double *ta = tri + static_cast<size_t>(i) * recw, *oa = cur + static_cast<size_t>(i) * recw;
                // a candidate passes when it is within the cull range of the trial state or (where the old state's
                // terms are needed) of the current state
                const float4 pt = trialPt[i], po = curPt[i];
                int qn = 0;
                bool stop = false;
                auto exact = [&](int jj, bool valid) {
                    if (valid) {
                        int c = 0;
                        #pragma unroll
                        for (int s = 16; s > 0; s >>= 1) if (c + s < 27 && cellBegin[c + s] <= jj) c += s;
                        const double *rb = rec + static_cast<size_t>(cellSrc[c] + (jj - cellBegin[c])) * recw;
                        const int tIdx = cellT[c];
                        const V3 t = sg.trans[tIdx];
                        const int gx = tIdx / 9 - 1, gy = (tIdx / 3) % 3 - 1, gz = tIdx % 3 - 1;
                        if (HARD == RP_HARD_NONE && SOFT) {
                            if (L.offCell >= 0) softPairBoth<true>(I, A, ta, oa, rb, t, gx, gy, gz, &enN, &enO);
                            else softPairBoth<false>(I, A, ta, oa, rb, t, gx, gy, gz, &enN, &enO);
                        } else {
                            pairTerms<HARD, SOFT>(I, A, ta, rb, t, gx, gy, gz, hard, soft, !counting, &oN, &enN);
                            if (needOld)
                                pairTerms<HARD, SOFT>(I, A, oa, rb, t, gx, gy, gz, counting, soft, false, &oO, &enO);
                        }
                    }
                    // early exit like Packing::countParticleOverlaps(earlyExit = true)
                    if (!counting && hard && __any_sync(0xffffffffu, oN > 0)) stop = true;
                };
                for (int base = 0; base < numFrozen && !stop; base += 32) {
                    const int jj = base + lane;
                    bool pass = false;
                    if (jj < numFrozen) {
                        const float4 c = cand[jj];
                        const float dx = c.x - pt.x, dy = c.y - pt.y, dz = c.z - pt.z;
                        const float ex = c.x - po.x, ey = c.y - po.y, ez = c.z - po.z;
                        if (A.cullBox) {
                            // un-truncated LJ: a centre pair contributes iff its reference-grid cells are adjacent,
                            // i.e. only if the mass centres are closer than two reference cells + the molecules'
                            // extent along every box axis (relative coordinates)
                            pass = !(fabsf(dx) >= A.cullBoxHalf[0]) && !(fabsf(dy) >= A.cullBoxHalf[1])
                                   && !(fabsf(dz) >= A.cullBoxHalf[2]);
                            pass = pass || (!(fabsf(ex) >= A.cullBoxHalf[0]) && !(fabsf(ey) >= A.cullBoxHalf[1])
                                            && !(fabsf(ez) >= A.cullBoxHalf[2]));
                        } else {
                            pass = !(fmaf(dz, dz, fmaf(dy, dy, dx * dx)) >= A.cull32Thr2);
                            if (needOld) pass = pass || !(fmaf(ez, ez, fmaf(ey, ey, ex * ex)) >= A.cull32Thr2);
                        }
                    }
                    const unsigned int m = __ballot_sync(0xffffffffu, pass);
                    if (pass) q[qn + __popc(m & ltMask)] = jj;
                    qn += __popc(m);
                    __syncwarp();
                    if (qn >= 32) {
                        qn -= 32;
                        exact(q[qn + lane], true);
                        __syncwarp();
                    }
                }
                if (qn > 0 && !stop) exact(lane < qn ? q[lane] : 0, lane < qn);
                __syncwarp();
            }
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                oO += __shfl_xor_sync(0xffffffffu, oO, o);
                oN += __shfl_xor_sync(0xffffffffu, oN, o);
                if (SOFT) {
                    enO += __shfl_xor_sync(0xffffffffu, enO, o);
                    enN += __shfl_xor_sync(0xffffffffu, enN, o);
                }
            }
            if (lane == 0) {
                if (T.wall && rejected[i] == 0) {       // + countParticleWallOverlaps (core/Packing.cpp:424-428)
                    const int2 w = T.wall[ownStart + i];
                    oO += w.x;
                    oN += w.y;
                }
                ovOld[i] = oO; ovNew[i] = oN; eOld[i] = enO; eNew[i] = enN;
                // hard, not counting: a trial that hits a frozen particle (or a wall) is rejected whatever happens
                // inside the cell (state 1; its row and column of the in-cell tables are never needed)
                if (!counting && hard && oN > 0 && rejected[i] == 0) rejected[i] = 1;
            }
        }
    }
    // processing order: rank of the random key
    for (int i = tid; i < numOwn; i += THREADS) {
        const unsigned int k = orderKey[i];
        int rank = 0;
        for (int j = 0; j < numOwn; j++) {
            const unsigned int kj = orderKey[j];
            rank += (kj < k) || (kj == k && j < i);
        }
        seq[rank] = i;
    }
    __syncthreads();

    // ---- phase 2a: pair terms inside the cell, all threads.  One work item = one ordered pair (i, j) x one variant
    // (which state of i against which state of j); consecutive items go to DIFFERENT warps (item = lane * WARPS +
    // warp): a small cell has a few dozen items, and with the expensive pair functions (XenoCollide's portal
    // refinement) one warp working through them while the others wait was most of the kernel's time ----
    {
        const V3 zero = v3(0, 0, 0);
        const int NV = matrices ? 4 : 2;
        const int items = blockwise ? 0 : numOwn * numOwn * NV;
        for (int base = 0; base < items; base += THREADS) {
            const int it = base + lane * WARPS + warp;
            if (it >= items) continue;
            const int v = it % NV, p = it / NV;
            const int i = p / numOwn, j = p - i * numOwn;
            if (i == j || rejected[i] != 0) continue;
            const double *ta = tri + static_cast<size_t>(i) * recw, *oa = cur + static_cast<size_t>(i) * recw;
            const double *ob = cur + static_cast<size_t>(j) * recw, *tb = tri + static_cast<size_t>(j) * recw;
            const bool jMayMove = rejected[j] == 0;
            int o = 0;
            double e = 0;
            if (!matrices) {       // bit masks: v = 0 trial of i / current state of j, v = 1 trial of i / trial of j
                if (v == 0) {
                    pairTerms<HARD, SOFT>(I, A, ta, ob, zero, 0, 0, 0, true, false, true, &o, &e);
                    if (o) atomicOr(&ownMask[i], 1ull << j);
                } else if (jMayMove) {
                    pairTerms<HARD, SOFT>(I, A, ta, tb, zero, 0, 0, 0, true, false, true, &o, &e);
                    if (o) atomicOr(&newMask[i], 1ull << j);
                }
            } else {               // matrices [0] new_i/old_j [1] old_i/old_j [2] new_i/new_j [3] old_i/new_j
                const double *ai = (v & 1) ? oa : ta, *bj = (v & 2) ? tb : ob;
                if (!(v & 2) || jMayMove) {
                    if (v & 1) pairTerms<HARD, SOFT>(I, A, ai, bj, zero, 0, 0, 0, counting, soft, false, &o, &e);
                    else pairTerms<HARD, SOFT>(I, A, ai, bj, zero, 0, 0, 0, hard, soft, !counting, &o, &e);
                }
                const size_t at = static_cast<size_t>(i) * A.capOwn + j;
                eMat[v * matStride + at] = e;
                cMat[v * matStride + at] = static_cast<short>(o);
            }
        }
    }
    __syncthreads();

    // ---- phase 2b: sequential pass through the cell in its random order ----
    const double INF = __longlong_as_double(0x7ff0000000000000ll);
    if (blockwise) {
        // every step: all threads evaluate the pair terms of trial i against the CURRENT states of the other particles
        // of the cell (an accepted trial replaces its current state at once), fixed-order reduction, one decision
        int nAcc = 0, nRejDomain = 0, nRejOverlap = 0, nRejEnergy = 0, nIdle = 0;
        long long overlapDelta = 0;
        const V3 zero = v3(0, 0, 0);
        for (int step = 0; step < numOwn; step++) {
            const int i = seq[step];
            const int state = rejected[i];
            if (state == 2) { nRejDomain++; continue; }
            if (state == 5) { nIdle++; continue; }
            const double *ta = tri + static_cast<size_t>(i) * recw, *oa = cur + static_cast<size_t>(i) * recw;
            int oO = 0, oN = 0;
            double enO = 0, enN = 0;
            if (state == 0) {
                for (int j = tid; j < numOwn; j += THREADS) {
                    if (j == i) continue;
                    const double *rb = cur + static_cast<size_t>(j) * recw;
                    pairTerms<HARD, SOFT>(I, A, ta, rb, zero, 0, 0, 0, hard, soft, !counting, &oN, &enN);
                    if (needOld) pairTerms<HARD, SOFT>(I, A, oa, rb, zero, 0, 0, 0, counting, soft, false, &oO, &enO);
                }
            }
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                oO += __shfl_xor_sync(0xffffffffu, oO, o);
                oN += __shfl_xor_sync(0xffffffffu, oN, o);
                enO += __shfl_xor_sync(0xffffffffu, enO, o);
                enN += __shfl_xor_sync(0xffffffffu, enN, o);
            }
            if (lane == 0) { sRedO[2 * warp] = oO; sRedO[2 * warp + 1] = oN; sRedE[2 * warp] = enO; sRedE[2 * warp + 1] = enN; }
            __syncthreads();
            if (tid == 0) {
                oO = ovOld[i]; oN = ovNew[i]; enO = 0; enN = 0;
                for (int w = 0; w < WARPS; w++) {
                    oO += sRedO[2 * w]; oN += sRedO[2 * w + 1]; enO += sRedE[2 * w]; enN += sRedE[2 * w + 1];
                }
                double dE;
                long long delta = 0;
                if (counting) {
                    delta = static_cast<long long>(oN) - static_cast<long long>(oO);
                    dE = delta < 0 ? -INF : (delta > 0 ? INF : 0.0);
                } else {
                    dE = (hard && (oN > 0 || state == 1)) ? INF : 0.0;
                }
                const bool overlapVerdict = dE != 0;
                if (!overlapVerdict && soft) dE = (eNew[i] + enN) - (eOld[i] + enO);
                const bool accept = uMet[i] <= exp(-dE / A.temperature);
                T.dE[ownStart + i] = dE;
                sAccept = accept ? 1 : 0;
                if (accept) {
                    nAcc++;
                    overlapDelta += delta;
                    sMoved[i >> 6] |= 1ull << (i & 63);
                } else if (overlapVerdict) nRejOverlap++;
                else nRejEnergy++;
            }
            __syncthreads();
            if (sAccept)
                for (int q = tid; q < recw; q += THREADS) cur[static_cast<size_t>(i) * recw + q] = ta[q];
            __syncthreads();
        }
        if (tid == 0) {
            sStats[0] = nAcc; sStats[1] = nRejDomain; sStats[2] = nRejOverlap; sStats[3] = nRejEnergy;
            sStats[4] = nIdle;
            sOverlapDelta = overlapDelta;
        }
    } else if (warp == 0) {
        int nAcc = 0, nRejDomain = 0, nRejOverlap = 0, nRejEnergy = 0, nIdle = 0;
        long long overlapDelta = 0;
        unsigned long long movedLo = 0ull, movedHi = 0ull;     // particles 0..63 / 64..127 of the cell
        for (int step = 0; step < numOwn; step++) {
            const int i = seq[step];
            const int state = rejected[i];
            if (state == 2) { nRejDomain++; continue; }
            if (state == 5) { nIdle++; continue; }      // RP_MOVE_FLIP: not drawn this cycle
            double dE;
            long long delta = 0;
            bool overlapVerdict;
            if (!matrices) {
                // hard only: blocked by a frozen particle, by an unmoved neighbour's old state or a moved one's new
                const bool blocked = state == 1
                                     || ((ownMask[i] & ~movedLo) | (newMask[i] & movedLo)) != 0ull;
                dE = blocked ? INF : 0.0;
                overlapVerdict = blocked;
            } else {
                int oO = 0, oN = 0;
                double enO = 0, enN = 0;
                for (int base = 0; base < numOwn; base += 32) {
                    const int j = base + lane;
                    if (j < numOwn && j != i) {
                        const bool mv = j < 64 ? ((movedLo >> j) & 1ull) != 0ull : ((movedHi >> (j - 64)) & 1ull) != 0ull;
                        const size_t at = static_cast<size_t>(i) * A.capOwn + j;
                        oN += cMat[(mv ? 2 : 0) * matStride + at];
                        oO += cMat[(mv ? 3 : 1) * matStride + at];
                        if (SOFT) {
                            enN += eMat[(mv ? 2 : 0) * matStride + at];
                            enO += eMat[(mv ? 3 : 1) * matStride + at];
                        }
                    }
                }
                #pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    oO += __shfl_xor_sync(0xffffffffu, oO, o);
                    oN += __shfl_xor_sync(0xffffffffu, oN, o);
                    if (SOFT) {
                        enO += __shfl_xor_sync(0xffffffffu, enO, o);
                        enN += __shfl_xor_sync(0xffffffffu, enN, o);
                    }
                }
                oO += ovOld[i]; oN += ovNew[i];
                // Packing::calculateMoveOverlapEnergy (Packing.cpp:304-323), then the energy difference
                if (counting) {
                    delta = static_cast<long long>(oN) - static_cast<long long>(oO);
                    dE = delta < 0 ? -INF : (delta > 0 ? INF : 0.0);
                } else {
                    dE = (hard && (oN > 0 || state == 1)) ? INF : 0.0;
                }
                overlapVerdict = dE != 0;
                if (!overlapVerdict && soft) dE = (eNew[i] + enN) - (eOld[i] + enO);
            }
            // Simulation::tryMove (Simulation.cpp:483): accept iff u <= exp(-dE / T)
            const bool accept = (soft || counting) ? (uMet[i] <= exp(-dE / A.temperature)) : !overlapVerdict;
            if (lane == 0) T.dE[ownStart + i] = dE;
            if (accept) {
                nAcc++;
                overlapDelta += delta;
                if (i < 64) movedLo |= 1ull << i; else movedHi |= 1ull << (i - 64);
            } else if (overlapVerdict) nRejOverlap++;
            else nRejEnergy++;
        }
        if (lane == 0) {
            sStats[0] = nAcc; sStats[1] = nRejDomain; sStats[2] = nRejOverlap; sStats[3] = nRejEnergy;
            sStats[4] = nIdle;
            sOverlapDelta = overlapDelta;
            sMoved[0] = movedLo;
            sMoved[1] = movedHi;
        }
    }
    __syncthreads();

    // ---- write-back of the accepted moves ----
    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION || A.kind == RP_MOVE_FLIP);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION
                             || (A.kind == RP_MOVE_FLIP && !A.flipOriginZero));
    for (int i = warp; i < numOwn; i += WARPS) {
        const bool acc = ((sMoved[i >> 6] >> (i & 63)) & 1ull) != 0ull;
        if (!acc) continue;
        const int slot = ownStart + i;
        const int id = items[slot];
        const double *t = tri + static_cast<size_t>(i) * recw;
        for (int q = lane; q < recw; q += 32) rec[static_cast<size_t>(slot) * recw + q] = t[q];
        if (lane == 0) {
            T.meta[slot].x = 4u;      // accepted (read back by rp_sweep_trace only)
            cull[slot] = trialPt[i];  // the cull point the later colours stage
        }
        // Shape::translate (Shape.cpp:9-12): position += bc.getCorrection(position)
        V3 p = ld3(t);
        if (translated) {
            p = p + pbcCorrection(box, p);
            if (lane == 0) { pos[id] = p.x; pos[A.cap + id] = p.y; pos[2 * A.cap + id] = p.z; }
        } else {
            p = v3(pos[id], pos[A.cap + id], pos[2 * A.cap + id]);
        }
        if (rotated && lane < 9) orient[lane * A.cap + id] = T.orient[9 * static_cast<size_t>(slot) + lane];
        if (lane < I.k) {
            // acceptTempInteractionCentres + recalculateAbsoluteInteractionCentres (Packing.cpp:339-348, 1041-1047)
            const size_t idx = static_cast<size_t>(id) * I.k + lane;
            V3 r = ld3(T.crel + (static_cast<size_t>(slot) * I.k + lane) * 3);
            crelAll[idx] = r.x; crelAll[A.ccap + idx] = r.y; crelAll[2 * A.ccap + idx] = r.z;
            V3 c = p + r;
            c = c + pbcCorrection(box, c);
            cabsAll[idx] = c.x; cabsAll[A.ccap + idx] = c.y; cabsAll[2 * A.ccap + idx] = c.z;
        }
    }
    if (tid == 0) {
        atomicAdd(&counters[0], static_cast<unsigned long long>(numOwn - sStats[4]));
        atomicAdd(&counters[1], static_cast<unsigned long long>(sStats[0]));
        atomicAdd(&counters[2], static_cast<unsigned long long>(sStats[1]));
        atomicAdd(&counters[3], static_cast<unsigned long long>(sStats[2]));
        atomicAdd(&counters[4], static_cast<unsigned long long>(sStats[3]));
        if (sOverlapDelta != 0) atomicAdd(&counters[5], static_cast<unsigned long long>(sOverlapDelta));
    }
}

}  // namespace rp
