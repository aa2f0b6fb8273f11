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
    int offCell;     // kk packed cells (3 x 21 bit, bias 2^20), -1 = absent
    int kk;
};

struct GenArgs {
    int kind;
    double transStep, rotStep, temperature;
    unsigned long long seed;
    unsigned int cycle, pass;
    int colour[3], nActive[3];
    int capOwn;
    int counting;            // Packing::overlapCounting
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
};

__device__ __forceinline__ V3 ld3(const double *p) { return v3(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(double *p, V3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }

// extended reference-grid cell of a position given in the image frame (NeighbourGrid::positionToCellNo without the
// wrap: floor(rel / relativeCellSize), may be < 0 or >= n)
__device__ __forceinline__ long long packRefCell(const Box &box, const double refRelCell[3], V3 p) {
    V3 rel = absToRel(box, p);
    long long cx = static_cast<long long>(floor(rel.x / refRelCell[0])) + (1 << 20);
    long long cy = static_cast<long long>(floor(rel.y / refRelCell[1])) + (1 << 20);
    long long cz = static_cast<long long>(floor(rel.z / refRelCell[2])) + (1 << 20);
    return cx | (cy << 21) | (cz << 42);
}
__device__ __forceinline__ bool refCellsAdjacent(long long a, long long b, int gx, int gy, int gz, const int n[3]) {
    const long long M = (1 << 21) - 1;
    long long dx = (a & M) - ((b & M) + gx * n[0]);
    long long dy = ((a >> 21) & M) - (((b >> 21) & M) + gy * n[1]);
    long long dz = ((a >> 42) & M) - (((b >> 42) & M) + gz * n[2]);
    return dx >= -1 && dx <= 1 && dy >= -1 && dy <= 1 && dz >= -1 && dz <= 1;
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
            reinterpret_cast<long long *>(r)[L.offCell + c] = packRefCell(box, A.refRelCell, ld3(r + L.offCentre + 3 * c));
}

// one warp per cell: items by ascending particle index + the generic records (image frame)
__global__ void k_gen_records(int numCells, GenArgs A, SweepGrid sg, Box box, const Inter *__restrict__ interPtr,
                              const int *__restrict__ cellStart, const int *__restrict__ itemsTmp,
                              const int *__restrict__ keys, const double *__restrict__ pos,
                              const double *__restrict__ orient, const double *__restrict__ crelAll,
                              int *__restrict__ items, double *__restrict__ rec)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();
    int cell = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (cell >= numCells) return;
    int start = cellStart[cell], cnt = cellStart[cell + 1] - start;
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
__device__ __forceinline__ void pairTerms(const Inter &I, const GenArgs &A, const double *a, const double *b, V3 t,
                                          int gx, int gy, int gz, bool wantOv, bool wantEn, bool firstOnly,
                                          int *ov, double *en)
{
    const GenLayout &L = A.L;
    V3 pa = ld3(a), pb = ld3(b) + t;
    if (!(norm2(pb - pa) < A.cullRange2)) return;
    BcFixed bc{t};
    for (int c1 = 0; c1 < L.kk; c1++) {
        V3 pos1 = ld3(a + L.offCentre + 3 * c1);
        V3 aux1 = L.offAux >= 0 ? ld3(a + L.offAux + 3 * c1) : v3(0, 0, 0);
        for (int c2 = 0; c2 < L.kk; c2++) {
            V3 pos2 = ld3(b + L.offCentre + 3 * c2);
            if (wantOv) {
                V3 aux2 = L.offAux >= 0 ? ld3(b + L.offAux + 3 * c2) : v3(0, 0, 0);
                const bool hit = needsOrientation(I)
                    ? overlapBetweenOriented(I, pos1, [&] { M33 o; for (int q = 0; q < 9; q++) o.a[q] = a[L.offOrient + q]; return o; },
                                             c1, pos2,
                                             [&] { M33 o; for (int q = 0; q < 9; q++) o.a[q] = b[L.offOrient + q]; return o; },
                                             c2, bc)
                    : overlapBetween(I, pos1, aux1, c1, pos2, aux2, c2, bc);
                if (hit) {
                    (*ov)++;
                    if (firstOnly) return;
                }
            }
            if (wantEn) {
                if (L.offCell >= 0) {
                    long long ca = reinterpret_cast<const long long *>(a)[L.offCell + c1];
                    long long cb = reinterpret_cast<const long long *>(b)[L.offCell + c2];
                    if (!refCellsAdjacent(ca, cb, gx, gy, gz, A.refN)) continue;
                }
                *en += energyBetween(I, pos1, pos2, bc);
            }
        }
    }
}

constexpr int GEN_THREADS = 128;

__global__ void __launch_bounds__(GEN_THREADS)
k_gen_sweep(GenArgs A, SweepGrid sg, Box box, const Inter *__restrict__ interPtr, const int *__restrict__ cellStart,
            const int *__restrict__ items, double *__restrict__ rec, GenTrialBuffers T, double *__restrict__ pos,
            double *__restrict__ orient, double *__restrict__ crelAll, double *__restrict__ cabsAll,
            unsigned long long *__restrict__ counters, int *__restrict__ flags)
{
    constexpr int WARPS = GEN_THREADS / 32;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ Inter I;
    __shared__ int cellBegin[28], cellSrc[27], cellT[27];
    __shared__ int sStats[5];
    __shared__ long long sOverlapDelta;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const GenLayout &L = A.L;
    const int recw = L.recw;
    for (int i = tid; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += GEN_THREADS)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];

    const int b = blockIdx.x;
    const int a0 = b % A.nActive[0], a1 = (b / A.nActive[0]) % A.nActive[1], a2 = b / (A.nActive[0] * A.nActive[1]);
    const int cx = 2 * a0 + A.colour[0], cy = 2 * a1 + A.colour[1], cz = 2 * a2 + A.colour[2];
    if (tid < 27) {
        int tIdx;
        int c2 = neighbourCell(sg.n, cx, cy, cz, tid % 3 - 1, (tid / 3) % 3 - 1, tid / 9 - 1, &tIdx);
        cellSrc[tid] = cellStart[c2];
        cellBegin[tid] = (tid == 13) ? 0 : cellStart[c2 + 1] - cellStart[c2];   // the own cell is not frozen
        cellT[tid] = tIdx;
    }
    if (tid == 0) { sStats[0] = sStats[1] = sStats[2] = sStats[3] = sStats[4] = 0; sOverlapDelta = 0; }
    __syncthreads();
    if (tid == 0) {
        int acc = 0;
        for (int c = 0; c < 27; c++) { int n = cellBegin[c]; cellBegin[c] = acc; acc += n; }
        cellBegin[27] = acc;
    }
    const int ownCell = (cz * sg.n[1] + cy) * sg.n[0] + cx;
    const int ownStart = cellStart[ownCell], numOwn = cellStart[ownCell + 1] - ownStart;
    if (numOwn == 0) return;
    if (numOwn > A.capOwn) {
        if (tid == 0) atomicOr(&flags[3], 1);
        return;
    }
    __syncthreads();
    const int numFrozen = cellBegin[27];

    // shared memory: current states and trial states of the own cell, per-trial accumulators
    double *cur = reinterpret_cast<double *>(smemRaw);
    double *tri = cur + static_cast<size_t>(A.capOwn) * recw;
    double *eOld = tri + static_cast<size_t>(A.capOwn) * recw;
    double *eNew = eOld + A.capOwn;
    double *uMet = eNew + A.capOwn;
    int *ovOld = reinterpret_cast<int *>(uMet + A.capOwn);
    int *ovNew = ovOld + A.capOwn;
    int *rejected = ovNew + A.capOwn;
    int *seq = rejected + A.capOwn;
    int *accepted = seq + A.capOwn;
    unsigned int *orderKey = reinterpret_cast<unsigned int *>(accepted + A.capOwn);

    for (int i = tid; i < numOwn * recw; i += GEN_THREADS) {
        cur[i] = rec[static_cast<size_t>(ownStart) * recw + i];
        tri[i] = T.rec[static_cast<size_t>(ownStart) * recw + i];
    }
    for (int i = tid; i < numOwn; i += GEN_THREADS) {
        const uint2 meta = T.meta[ownStart + i];
        rejected[i] = static_cast<int>(meta.x);
        orderKey[i] = meta.y;
        uMet[i] = T.u[ownStart + i];
        accepted[i] = 0;
    }
    __syncthreads();

    const bool hard = I.hard != RP_HARD_NONE, soft = I.soft != RP_SOFT_NONE;
    const bool counting = A.counting != 0 && hard;
    const bool needOld = counting || soft;

    // ---- phase 1: frozen neighbours, one warp per trial ----
    for (int i = warp; i < numOwn; i += WARPS) {
        int oO = 0, oN = 0;
        double enO = 0, enN = 0;
        if (rejected[i] == 0) {
            const double *ta = tri + static_cast<size_t>(i) * recw, *oa = cur + static_cast<size_t>(i) * recw;
            for (int base = 0; base < numFrozen; base += 32) {
                const int jj = base + lane;
                if (jj < numFrozen) {
                    int c = 0;
                    #pragma unroll
                    for (int s = 16; s > 0; s >>= 1) if (c + s < 27 && cellBegin[c + s] <= jj) c += s;
                    const double *rb = rec + static_cast<size_t>(cellSrc[c] + (jj - cellBegin[c])) * recw;
                    const int tIdx = cellT[c];
                    const V3 t = sg.trans[tIdx];
                    const int gx = tIdx / 9 - 1, gy = (tIdx / 3) % 3 - 1, gz = tIdx % 3 - 1;
                    pairTerms(I, A, ta, rb, t, gx, gy, gz, hard, soft, !counting, &oN, &enN);
                    if (needOld) pairTerms(I, A, oa, rb, t, gx, gy, gz, counting, soft, false, &oO, &enO);
                }
                // early exit like Packing::countParticleOverlaps(earlyExit = true)
                if (!counting && hard && __any_sync(0xffffffffu, oN > 0)) break;
            }
        }
        #pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            oO += __shfl_xor_sync(0xffffffffu, oO, o);
            oN += __shfl_xor_sync(0xffffffffu, oN, o);
            enO += __shfl_xor_sync(0xffffffffu, enO, o);
            enN += __shfl_xor_sync(0xffffffffu, enN, o);
        }
        if (lane == 0) {
            if (T.wall && rejected[i] == 0) {       // + countParticleWallOverlaps (core/Packing.cpp:424-428)
                const int2 w = T.wall[ownStart + i];
                oO += w.x;
                oN += w.y;
            }
            ovOld[i] = oO; ovNew[i] = oN; eOld[i] = enO; eNew[i] = enN;
        }
    }
    // processing order: rank of the random key
    for (int i = tid; i < numOwn; i += GEN_THREADS) {
        const unsigned int k = orderKey[i];
        int rank = 0;
        for (int j = 0; j < numOwn; j++) {
            const unsigned int kj = orderKey[j];
            rank += (kj < k) || (kj == k && j < i);
        }
        seq[rank] = i;
    }
    __syncthreads();

    // ---- phase 2: sequential pass through the cell (warp 0), other particles of the cell in their CURRENT state ----
    if (warp == 0) {
        int nAcc = 0, nRejDomain = 0, nRejOverlap = 0, nRejEnergy = 0, nIdle = 0;
        long long overlapDelta = 0;
        const V3 zero = v3(0, 0, 0);
        for (int step = 0; step < numOwn; step++) {
            const int i = seq[step];
            if (rejected[i] == 2) { nRejDomain++; continue; }
            if (rejected[i] == 5) { nIdle++; continue; }      // RP_MOVE_FLIP: not drawn this cycle
            const double *ta = tri + static_cast<size_t>(i) * recw, *oa = cur + static_cast<size_t>(i) * recw;
            int oO = 0, oN = 0;
            double enO = 0, enN = 0;
            const bool frozenHit = !counting && hard && ovNew[i] > 0;
            if (!frozenHit) {
                for (int base = 0; base < numOwn; base += 32) {
                    const int j = base + lane;
                    if (j < numOwn && j != i) {
                        const double *rb = cur + static_cast<size_t>(j) * recw;
                        pairTerms(I, A, ta, rb, zero, 0, 0, 0, hard, soft, !counting, &oN, &enN);
                        if (needOld) pairTerms(I, A, oa, rb, zero, 0, 0, 0, counting, soft, false, &oO, &enO);
                    }
                    if (!counting && hard && __any_sync(0xffffffffu, oN > 0)) break;
                }
            }
            #pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                oO += __shfl_xor_sync(0xffffffffu, oO, o);
                oN += __shfl_xor_sync(0xffffffffu, oN, o);
                enO += __shfl_xor_sync(0xffffffffu, enO, o);
                enN += __shfl_xor_sync(0xffffffffu, enN, o);
            }
            oO += ovOld[i]; oN += ovNew[i];
            // Packing::calculateMoveOverlapEnergy (Packing.cpp:304-323), then the energy difference
            double dE;
            long long delta = 0;
            const double INF = __longlong_as_double(0x7ff0000000000000ll);
            if (counting) {
                delta = static_cast<long long>(oN) - static_cast<long long>(oO);
                dE = delta < 0 ? -INF : (delta > 0 ? INF : 0.0);
            } else {
                dE = (hard && oN > 0) ? INF : 0.0;
            }
            const bool overlapVerdict = dE != 0;
            if (!overlapVerdict && soft) dE = (eNew[i] + enN) - (eOld[i] + enO);
            // Simulation::tryMove (Simulation.cpp:483): accept iff u <= exp(-dE / T)
            const bool accept = uMet[i] <= exp(-dE / A.temperature);
            if (lane == 0) T.dE[ownStart + i] = dE;
            if (accept) {
                nAcc++;
                overlapDelta += delta;
                for (int q = lane; q < recw; q += 32) cur[static_cast<size_t>(i) * recw + q] = ta[q];
                if (lane == 0) accepted[i] = 1;
                __syncwarp();
            } else if (overlapVerdict) nRejOverlap++;
            else nRejEnergy++;
        }
        if (lane == 0) {
            sStats[0] = nAcc; sStats[1] = nRejDomain; sStats[2] = nRejOverlap; sStats[3] = nRejEnergy;
            sStats[4] = nIdle;
            sOverlapDelta = overlapDelta;
        }
    }
    __syncthreads();

    // ---- write-back of the accepted moves ----
    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION || A.kind == RP_MOVE_FLIP);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION
                             || (A.kind == RP_MOVE_FLIP && !A.flipOriginZero));
    for (int i = warp; i < numOwn; i += WARPS) {
        if (!accepted[i]) continue;
        const int slot = ownStart + i;
        const int id = items[slot];
        const double *t = cur + static_cast<size_t>(i) * recw;
        for (int q = lane; q < recw; q += 32) rec[static_cast<size_t>(slot) * recw + q] = t[q];
        if (lane == 0) T.meta[slot].x = 4u;      // accepted (read back by rp_sweep_trace only)
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
