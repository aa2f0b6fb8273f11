// rp_sweep_kernel.cuh — the checkerboard sub-sweep kernels (north_star kernel (b)); included by rp_sweep.cu only.
//
// k_sweep_trials  one thread per particle, once per rp_sweep call: draws the trial move of every particle from its
//                 Philox stream (MoveSampler distributions), applies the active-domain rule and stores the trial
//                 record {new position, new half axis}, the new orientation and {reject flag, order key}.
// k_sweep         one CTA per ACTIVE cell of the current colour (8 launches per rp_sweep call):
//   stage   the 27 cells around the active cell (9 x-runs, contiguous in the cell-sorted order) and the trial
//           records of the own cell are pulled into shared memory with 1-D TMA bulk copies (cp.async.bulk +
//           mbarrier); a second pass adds the lattice translations (pos2 + translation, as the reference's
//           HardcodedTranslation) and makes an FP32 copy of the positions in cell-local coordinates;
//   phase A conservative FP32 cull of ALL (trial, candidate) pairs — candidates are the staged OLD states plus the
//           NEW states of the other trials of the cell; a lane owns a candidate and builds a bitmask over trials;
//           survivors are appended to one CTA-wide queue;
//   phase B the queue is processed by all threads in lock step: exact FP64 pair test (reference operation order,
//           -fmad=false); results: "hits a frozen particle" flags and two bit matrices ownMask[i] (trial i vs OLD
//           state of own j) and newMask[i] (trial i vs NEW state of own j);
//   phase C the trials of the cell are resolved sequentially in a random order from the bit matrices alone: trial i
//           is accepted iff it hits no frozen particle, no unmoved own particle (ownMask & ~moved) and no already
//           moved own particle (newMask & moved) — exactly a sequential Monte Carlo pass through the cell with the
//           reference's accept rule (Simulation.cpp:483 with dE in {0, +inf});
//   write   accepted states go back to the sorted records and to the canonical SoA arrays.
//
// The FP32 cull never decides a verdict: a pair it drops is farther apart than the pair function's range by more
// than the proven FP32 error bound (cullThr2, computed on the host), so the FP64 test would return "disjoint".
#pragma once

#include "rp_internal.cuh"
#include "rp_philox.cuh"
#include "rp_shard.cuh"
#include "rp_wall.cuh"

namespace rp {

// staging order of the 27 neighbour cells (codes (dz+1)*9 + (dy+1)*3 + (dx+1)): own cell, faces, edges, corners
__device__ const unsigned char SW_STAGE_ORDER[27] = {13, 4, 10, 12, 14, 16, 22, 1, 3, 5, 7, 9, 11, 15, 17, 19, 21, 23, 25,
                                                     0, 2, 6, 8, 18, 20, 24, 26};
constexpr int SW_MAXOWN = 64;      // trials resolved together (bit matrices are 64 bits wide)
constexpr int SW_PUSH = 4;         // queue entries a lane may push per round when a chunk's survivors do not fit
constexpr int SW_WQCAP1 = 32 * SW_PUSH + 96;    // per-warp queue of (trial, candidate) pairs that passed the first cull
constexpr int SW_WQCAP = SW_WQCAP1 + 96;        // ... followed by the queue of pairs awaiting the exact test

// the few interaction constants the sweep kernels need (rp::Inter thresholds, computed on the host with the
// reference's expressions), passed by value
struct SweepInter {
    double sphereThr2;                    // SphereTraits.cpp:56-65
    double scInner2, scOuter2, scHalf;    // SpherocylinderTraits.cpp:28-46: 4r^2, (2r + l)^2, 0.5 * l
};

struct SweepArgs {
    int kind;                   // rp_move_kind
    double transStep, rotStep;
    unsigned long long seed;
    unsigned int cycle, pass;
    int colour[3];              // parity of the active cells per axis
    int nActive[3];             // active cells per axis
    int capOld;                 // capacity of the staged old states (max 27-neighbourhood population)
    int capOwn;                 // capacity of the trial region (max cell population)
    float cullThr2;             // FP32 cull threshold (squared distance) including the FP32 error bound
    float cullHalf;             // spherocylinders: half length of the axis segment
    float cull2Thr2;            // spherocylinders: threshold of the FP32 segment-segment cull (squared distance)
    SweepInter inter;
    size_t cap;
    int aOff2;                  // sharded runs: first active-cell layer (z) of this rank's slab, else 0
    int slotBegin, slotEnd;     // sorted slots whose trials this rank generates (all of them on one GPU)
    // sharded over peer memory: this rank's slab of cell layers [trialLayerLo, trialLayerHi) — k_sweep_prepare writes
    // trials for the slab only and records for the slab and the layer on either side; -1: everything
    int trialLayerLo, trialLayerHi;
};

// device buffers of the trials of one rp_sweep call, indexed by the sorted slot
struct TrialBuffers {
    double *rec;                // 6 doubles per slot: new position (image frame), new half axis
    double *orient;             // 9 doubles per slot: new orientation (only written for rotating samplers)
    uint2 *meta;                // x: 0 alive / 2 left the cell, y: random key of the processing order
    float4 *rec32;              // 2 float4 per slot: the FP32 cull record of the new state (Rec32 below)
};

// FP32 cull record of one particle state, 32 bytes (two float4): {x, y, z, ux} {uy, uz, 0, 0} — position relative to
// the CENTRE OF THE PARTICLE'S OWN sweep cell (image frame) and the unit axis (spherocylinders).  Only the two
// conservative culls read it; the exact test reads the FP64 record.  A CTA that stages a neighbour cell adds the
// centre-to-centre vector of the two cells (SweepCellOffsets, the same for every active cell), which also carries
// the lattice translation of a periodic image.
struct SweepCellOffsets { float o[27][3]; };      // index = neighbour code (dz+1)*9 + (dy+1)*3 + (dx+1)

// centre of sweep cell (cx, cy, cz) in the image frame of the shifted origin
RP_HD V3 sweepCellCentre(const SweepGrid &sg, const Box &box, int cx, int cy, int cz) {
    return relToAbs(box, v3((cx + 0.5) * sg.relCell[0] + sg.shift[0], (cy + 0.5) * sg.relCell[1] + sg.shift[1],
                            (cz + 0.5) * sg.relCell[2] + sg.shift[2]));
}
// the one expression every producer of a cull record uses (prepare kernel, sharded apply kernels): identical inputs
// give identical records on every rank
__device__ __forceinline__ void makeRec32(const V3 &p, const V3 &h, const V3 &centre, float invHalf, float4 *out) {
    out[0] = make_float4(static_cast<float>(p.x - centre.x), static_cast<float>(p.y - centre.y),
                         static_cast<float>(p.z - centre.z), static_cast<float>(h.x) * invHalf);
    out[1] = make_float4(static_cast<float>(h.y) * invHalf, static_cast<float>(h.z) * invHalf, 0.f, 0.f);
}

// ---- 1-D TMA bulk copy + mbarrier (PTX ISA: cp.async.bulk, mbarrier) ----
__device__ __forceinline__ unsigned int smemAddr(const void *p) {
    return static_cast<unsigned int>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbarInit(unsigned long long *bar, unsigned int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smemAddr(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbarExpectTx(unsigned long long *bar, unsigned int bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkCopyG2S(void *dst, const void *src, unsigned int bytes, unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smemAddr(dst)), "l"(src), "r"(bytes), "r"(smemAddr(bar)) : "memory");
}
__device__ __forceinline__ void mbarWait(unsigned long long *bar, unsigned int phase) {
    unsigned int done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(done) : "r"(smemAddr(bar)), "r"(phase) : "memory");
        if (!done) __nanosleep(200);     // leave the issue slots to the other CTAs of the SM
    }
}

// Exact pair test of a trial record `t` = {new position, new half axis} against a staged candidate record `c`
// (position already translated).
// Spherocylinder: SpherocylinderTraits::overlapBetween (SpherocylinderTraits.cpp:28-46); the cap centres
// pos + axis * (0.5 * -+1 * l) are formed as pos -+ h with h = axis * (0.5 * l): bit-identical because
// axis * (-x) == -(axis * x) and a + (-b) == a - b in IEEE arithmetic.
template <int HARD>
__device__ __forceinline__ bool exactOverlap(const SweepInter &I, const double *__restrict__ t,
                                             const double *__restrict__ c)
{
    V3 p1 = v3(t[0], t[1], t[2]), p2 = v3(c[0], c[1], c[2]);
    double d2 = norm2(p2 - p1);
    if (HARD == RP_HARD_SPHERE) return d2 < I.sphereThr2;
    // straight-line form of: d2 < 4r^2 -> true; d2 >= (2r + l)^2 -> false; else segment distance^2 < 4r^2.
    // (after the FP32 cull practically every pair needs the segment test, so nothing is saved by branching)
    V3 h1 = v3(t[3], t[4], t[5]), h2 = v3(c[3], c[4], c[5]);
    double s = segmentDistance2(p1 - h1, p1 + h1, p2 - h2, p2 + h2);
    return (d2 < I.scInner2) || (!(d2 >= I.scOuter2) && s < I.scInner2);
}

// Second FP32 cull for spherocylinders: squared distance between the two axis segments in single precision against
// thr2 = 4r^2 + 1e-3 (l + 2r)^2.  Centred parametrisation P1(s) = c1 + s h1, P2(t) = c2 + t h2, s, t in [-1, 1]:
// s = clamp(sN / denom) from the unconstrained optimum, t = the optimum for that s, and when t had to be clamped s =
// the optimum for the clamped t.  Whatever rounding does to s (ill-conditioned for nearly parallel axes), the other
// parameter is optimal for it, so the computed squared distance exceeds the true minimum only by
// (denom / |h|^2) (s - s*)^2 relative to the clamped optimum: with absolute errors ~3e-5 in sN and ~1e-5 in denom
// that is < 3e-5 in every regime (denom above the noise: quadratic in a small ds; denom inside the noise: the
// coefficient itself is tiny; sign of sN lost: then |sN| is noise and the two ends differ by 4 |sN| / |h|^2), plus
// ~1e-5 of input rounding (coordinates relative to the cell centre).  Observed maximum over 1.6e7 random pairs
// including axes parallel to 1e-8: 4e-6 (tests/test_cull_math.py).  The margin is 0.036 for l = 5, r = 0.5 — and the
// reference's own segment distance is never BELOW the true minimum — so a pair this cull drops is disjoint for the
// FP64 test.  (Checked like the first cull: RP_SWEEP_NO_CULL=1 must give a bit-identical trajectory.)
__device__ __forceinline__ bool segmentCullPass(const double *__restrict__ c, const V3 &origin, const float4 tf,
                                                const float4 tu, float half, float thr2)
{
    const float dx = tf.x - static_cast<float>(c[0] - origin.x), dy = tf.y - static_cast<float>(c[1] - origin.y),
                dz = tf.z - static_cast<float>(c[2] - origin.z);
    const float h1x = tu.x * half, h1y = tu.y * half, h1z = tu.z * half;
    const float h2x = static_cast<float>(c[3]), h2y = static_cast<float>(c[4]), h2z = static_cast<float>(c[5]);
    const float a = half * half;                                  // |h1|^2 = |h2|^2
    const float b = fmaf(h1z, h2z, fmaf(h1y, h2y, h1x * h2x));
    const float cc = fmaf(h1z, dz, fmaf(h1y, dy, h1x * dx));
    const float f = fmaf(h2z, dz, fmaf(h2y, dy, h2x * dx));
    const float denom = fmaf(a, a, -b * b);
    const float inva = a > 0.f ? __fdividef(1.0f, a) : 0.f;      // length 0: both parameters stay 0
    float s = __fdividef(fmaf(b, f, -cc * a), fmaxf(denom, 1e-20f));    // nearly parallel: the SIGN still decides the end
    s = fminf(fmaxf(s, -1.0f), 1.0f);
    const float t0 = fmaf(b, s, f) * inva;
    const float t = fminf(fmaxf(t0, -1.0f), 1.0f);
    const float s1 = fminf(fmaxf(fmaf(b, t, -cc) * inva, -1.0f), 1.0f);
    s = (t != t0) ? s1 : s;
    const float px = fmaf(-t, h2x, fmaf(s, h1x, dx)), py = fmaf(-t, h2y, fmaf(s, h1y, dy)),
                pz = fmaf(-t, h2z, fmaf(s, h1z, dz));
    return !(fmaf(pz, pz, fmaf(py, py, px * px)) >= thr2);        // NaN passes
}

// FP32 segment-segment cull on two cull records (same formula as segmentCullPass; d = trial - candidate)
__device__ __forceinline__ bool segmentCullPass32(const float4 ca, const float4 cb, const float4 ta, const float4 tb,
                                                  float half, float thr2)
{
    const float dx = ta.x - ca.x, dy = ta.y - ca.y, dz = ta.z - ca.z;
    const float h1x = ta.w * half, h1y = tb.x * half, h1z = tb.y * half;
    const float h2x = ca.w * half, h2y = cb.x * half, h2z = cb.y * half;
    const float a = half * half;                                  // |h1|^2 = |h2|^2
    const float b = fmaf(h1z, h2z, fmaf(h1y, h2y, h1x * h2x));
    const float cc = fmaf(h1z, dz, fmaf(h1y, dy, h1x * dx));
    const float f = fmaf(h2z, dz, fmaf(h2y, dy, h2x * dx));
    const float denom = fmaf(a, a, -b * b);
    const float inva = a > 0.f ? __fdividef(1.0f, a) : 0.f;      // length 0: both parameters stay 0
    float s = __fdividef(fmaf(b, f, -cc * a), fmaxf(denom, 1e-20f));    // nearly parallel: the SIGN still decides the end
    s = fminf(fmaxf(s, -1.0f), 1.0f);
    const float t0 = fmaf(b, s, f) * inva;
    const float t = fminf(fmaxf(t0, -1.0f), 1.0f);
    const float s1 = fminf(fmaxf(fmaf(b, t, -cc) * inva, -1.0f), 1.0f);
    s = (t != t0) ? s1 : s;
    const float px = fmaf(-t, h2x, fmaf(s, h1x, dx)), py = fmaf(-t, h2y, fmaf(s, h1y, dy)),
                pz = fmaf(-t, h2z, fmaf(s, h1z, dz));
    return !(fmaf(pz, pz, fmaf(py, py, px * px)) >= thr2);        // NaN passes
}

// largest population of a cell's 27-neighbourhood: the shared-memory capacity k_sweep is launched with
static __global__ void __launch_bounds__(256) k_sweep_neighbourhood(int numCells, SweepGrid sg,
                                                             const int *__restrict__ cellStart,
                                                             int *__restrict__ maxNeigh)
{
    const int cell = blockIdx.x * blockDim.x + threadIdx.x;
    int sum = 0;
    if (cell < numCells) {
        const int cx = cell % sg.n[0], cy = (cell / sg.n[0]) % sg.n[1], cz = cell / (sg.n[0] * sg.n[1]);
        for (int c = 0; c < 27; c++) {
            int tIdx;
            const int c2 = neighbourCell(sg.n, cx, cy, cz, c / 9 - 1, (c / 3) % 3 - 1, c % 3 - 1, &tIdx);
            sum += cellStart[c2 + 1] - cellStart[c2];
        }
    }
    sum = __reduce_max_sync(0xffffffffu, sum);
    if ((threadIdx.x & 31) == 0) atomicMax(maxNeigh, sum);
}

// ---------------------------------------------------------------------------------------------------------------
// k_sweep_prepare — one thread per particle (per slot of the scattered item list), once per rp_sweep call.  Fuses what
// used to be two gather passes over the canonical arrays:
//   records  order the cell's items by ascending particle index (deterministic whatever the scatter atomics did),
//            write the 48-byte FP64 record {P', h} (image frame; h = orientation.column(2) * (l / 2)) and the 32-byte
//            FP32 cull record;
//   trials   MoveSampler::sampleMove distributions (core/move_samplers/*.cpp), Shape::translate / rotate
//            (core/Shape.cpp:9-16), ActiveDomain::isInside with the checkerboard cell as the domain
//            (Packing.cpp:95-96): trial FP64 record, trial cull record, new orientation, {reject flag, order key};
// (The shared-memory capacity of k_sweep — the largest population of a 27-neighbourhood — is k_sweep_neighbourhood's.)
// The kernel is bound by the latency of its dependent chains (Philox rounds, FP64 Rodrigues formula, the gathers):
// every lane carries a particle and the register allocation leaves room for 24 warps per SM.
// ---------------------------------------------------------------------------------------------------------------
template <int HARD>
__global__ void __launch_bounds__(256, 3) k_sweep_prepare(int n, SweepArgs A, SweepGrid sg, Box box,
                                                          double halfLength, const int *__restrict__ cellStart,
                                                          const int *__restrict__ itemsTmp,
                                                          const int *__restrict__ keys, const double *__restrict__ pos,
                                                          const double *__restrict__ orient, int *__restrict__ items,
                                                          double *__restrict__ rec, float4 *__restrict__ rec32,
                                                          TrialBuffers T, Walls W, double radius)
{
    const int scattered = blockIdx.x * blockDim.x + threadIdx.x;
    if (scattered >= n) return;
    const int particleId = itemsTmp[scattered];
    const int key = keys[particleId];
    const int cell = key & 0x0FFFFFFF, f = (key >> 28) & 7;
    // sharded over peer memory: records only where this rank's k_sweep stages them, trials only for its own slab
    bool wantTrial = true, wantRec = true;
    if (A.trialLayerLo >= 0) {
        const int cz = cell / (sg.n[0] * sg.n[1]), nz = sg.n[2];
        wantTrial = cz >= A.trialLayerLo && cz < A.trialLayerHi;
        wantRec = wantTrial || cz == (A.trialLayerLo - 1 + nz) % nz || cz == A.trialLayerHi % nz;
    }
    // the canonical state: gathers issued before anything depends on them
    V3 p = v3(0, 0, 0);
    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    M33 O;
    V3 h = v3(0, 0, 0);
    if (wantRec) {
        p = v3(pos[particleId], pos[A.cap + particleId], pos[2 * A.cap + particleId]);
        if (rotated) {
            #pragma unroll
            for (int q = 0; q < 9; q++) O.a[q] = orient[q * A.cap + particleId];
        } else if (HARD == RP_HARD_SPHEROCYLINDER) {
            // half axis h = orientation.column(2) * (0.5 * l): the cap centres are pos -+ h (SpherocylinderTraits.cpp:24-26)
            h = v3(orient[2 * A.cap + particleId], orient[5 * A.cap + particleId], orient[8 * A.cap + particleId]);
        }
    }
    const int start = cellStart[cell], cnt = cellStart[cell + 1] - start;
    const int cc[3] = {cell % sg.n[0], (cell / sg.n[0]) % sg.n[1], cell / (sg.n[0] * sg.n[1])};
    const V3 centre = sweepCellCentre(sg, box, cc[0], cc[1], cc[2]);
    const float invHalfF = (HARD == RP_HARD_SPHEROCYLINDER && A.cullHalf > 0.f) ? __fdividef(1.0f, A.cullHalf) : 0.f;
    int rank = 0;
    for (int j = 0; j < cnt; j++) rank += (itemsTmp[start + j] < particleId);
    do {
        const int slot = start + rank;
        items[slot] = particleId;
        if (!wantRec) break;
        if (f) p = p + sg.imageShift[f];
        if (HARD == RP_HARD_SPHEROCYLINDER) h = (rotated ? column(O, 2) : h) * halfLength;
        double *r = rec + 6 * static_cast<size_t>(slot);
        r[0] = p.x; r[1] = p.y; r[2] = p.z; r[3] = h.x; r[4] = h.y; r[5] = h.z;
        makeRec32(p, h, centre, invHalfF, rec32 + 2 * static_cast<size_t>(slot));
        if (slot < A.slotBegin || slot >= A.slotEnd || !wantTrial) break;          // sharded: another rank's trial

        // ---- the trial move of this particle ----
        Rng rng;
        rng.init(A.seed, static_cast<unsigned int>(particleId), A.cycle, A.pass);
        unsigned int rej = 0;
        if (translated) {
            // TranslationSampler / RototranslationSampler: 3 x U(-step, step)  (RototranslationSampler.cpp:31-33)
            double tx = rng.uniform(-A.transStep, A.transStep);
            double ty = rng.uniform(-A.transStep, A.transStep);
            double tz = rng.uniform(-A.transStep, A.transStep);
            p = p + v3(tx, ty, tz);
            // the new mass centre must stay in its cell (image frame of the shifted origin)
            V3 rel = absToRel(box, p);
            double q[3] = {rel.x - sg.shift[0], rel.y - sg.shift[1], rel.z - sg.shift[2]};
            for (int d = 0; d < 3; d++) {
                if (q[d] < 0 || q[d] >= 1) rej = 2;
                else if (static_cast<int>(q[d] / sg.relCell[d]) != cc[d]) rej = 2;
            }
        }
        if (rotated) {
            // axis uniform in the unit ball by rejection, angle U(-a, a) (RototranslationSampler.cpp:35-43);
            // Rodrigues: identity + K*sina + ((1 - cosa)*K)*K (geometry/Matrix.tpp:437-449)
            V3 axis;
            double n2;
            do {
                axis = v3(rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(-1, 1));
                n2 = norm2(axis);
            } while (n2 > 1 || n2 == 0);
            double nrm = sqrt(n2);
            axis = v3(axis.x / nrm, axis.y / nrm, axis.z / nrm);
            double angle = rng.uniform(-A.rotStep, A.rotStep);
            double sina = sin(angle), cosa = cos(angle);
            M33 K = {{0, -axis.z, axis.y, axis.z, 0, -axis.x, -axis.y, axis.x, 0}};
            M33 sK;
            for (int q = 0; q < 9; q++) sK.a[q] = (1 - cosa) * K.a[q];
            M33 KK = mul(sK, K);
            M33 R;
            for (int q = 0; q < 9; q++) R.a[q] = (((q % 4 == 0) ? 1.0 : 0.0) + K.a[q] * sina) + KK.a[q];
            M33 newOrient = mul(R, O);     // Shape::rotate (core/Shape.cpp:14-16)
            if (HARD == RP_HARD_SPHEROCYLINDER) h = column(newOrient, 2) * halfLength;
            double *o = T.orient + 9 * static_cast<size_t>(slot);
            for (int q = 0; q < 9; q++) o[q] = newOrient.a[q];
        }
        if (W.mask && rej == 0) {
            // hard walls (Packing::countParticleWallOverlaps, core/Packing.cpp:1086-1142, with the sphere / spherocylinder
            // overlapWithWall): a trial touching a wall is rejected like an overlap with a frozen particle.  The test
            // runs on the wrapped position the write-back would store.
            V3 cp = p + pbcCorrection(box, p);
            for (int i2 = 0; i2 < 3; i2++) {
                if (!((W.mask >> i2) & 1)) continue;
                V3 wv = W.normal[i2], origin;
                if (dot(cp, wv) < W.halfTotalRange) origin = v3(0, 0, 0);
                else if (dot(W.furtherOrigin - cp, wv) < W.halfTotalRange) {
                    origin = W.furtherOrigin;
                    wv = v3(-wv.x, -wv.y, -wv.z);
                } else continue;
                bool hit;
                if (HARD == RP_HARD_SPHERE) hit = dot(wv, cp - origin) < radius;
                else hit = dot(wv, (cp + h) - origin) < radius || dot(wv, (cp - h) - origin) < radius;
                if (hit) rej = 1;
            }
        }
        double *t = T.rec + 6 * static_cast<size_t>(slot);
        t[0] = p.x; t[1] = p.y; t[2] = p.z; t[3] = h.x; t[4] = h.y; t[5] = h.z;
        makeRec32(p, h, centre, invHalfF, T.rec32 + 2 * static_cast<size_t>(slot));
        rng.ctr.w = 0x40000000u;     // separate sub-stream for the processing order
        T.meta[slot] = make_uint2(rej, philox4x32(rng.ctr, rng.k0, rng.k1).x);
    } while (false);
}

}  // namespace rp

#include "rp_sweep_kernel2.cuh"
