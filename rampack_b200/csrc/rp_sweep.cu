// rp_sweep.cu — checkerboard domain-decomposed particle moves (north_star kernel (b)).
//
// Replaces, as ONE bulk call per cycle and sampler, the reference's
//   Simulation::performMovesWithDomainDivision + Simulation::tryMove      core/Simulation.cpp:389-491
//   DomainDecomposition / ActiveDomain::isInside                           core/DomainDecomposition.cpp:36-147
//   Packing::tryTranslation / tryRotation / tryMove / acceptMove           core/Packing.cpp:84-164, 278-302
//   Packing::countParticleOverlaps (early exit)                            core/Packing.cpp:373-429
//   MoveSampler::sampleMove distributions                                  core/move_samplers/*.cpp
//
// Design (new, not the reference's structure):
//   * every cycle draws a random grid origin `s` (the reference draws a random domain origin, Simulation.cpp:401-404)
//     and bins the particle mass centres into cells of width >= totalRange, an even number per axis;
//   * the 8 parity colours of the cells are swept one after another (random order); in one sub-sweep all cells of
//     one colour are ACTIVE and each is owned by one CTA; the 26 cells around an active cell are inactive, hence
//     frozen, and two active cells are at least one cell (>= totalRange) apart, so concurrent moves never interact;
//   * a move whose new mass centre leaves its cell is rejected — the reference's ActiveDomain rule (Packing.cpp:95-96);
//   * inside a CTA every particle of the cell gets one trial move, all generated up front from per-particle Philox
//     streams; ALL (trial, frozen candidate) pair tests run data-parallel (centre-distance pre-filter for every
//     pair, survivors compacted into warp queues so that the expensive full test runs with 32 busy lanes), and only
//     the interplay between trials of the same cell is resolved sequentially (random order) by one warp from
//     bitmasks of intra-cell overlaps.
//   * positions are handled in the image frame of the shifted origin: P' = p + M*f, f_i = [rel_i < s_i], so the cell
//     grid is an ordinary periodic grid over [s, s+1)^3 and the only lattice translations are the reference's 27.
#include "rp_internal.cuh"
#include "rp_sweep_kernel.cuh"
#include "rp_sweep_generic.cuh"
#include "rp_shard.cuh"
#include <cstdlib>
#include <limits>

namespace rp {

// ---------------------------------------------------------------------------------------------------------------
// sweep grid
// ---------------------------------------------------------------------------------------------------------------
// cell coordinate in the shifted frame of a relative coordinate r' = rel - s + f in [0, 1)
__device__ __forceinline__ int shiftedCoord(double r, double relCell, int n) {
    int c = static_cast<int>(r / relCell);
    return c < 0 ? 0 : (c >= n ? n - 1 : c);
}

// keys: image flag, cell in the shifted frame, histogram
__global__ void k_sweep_keys(int n, size_t cap, const double *__restrict__ pos, Box box, SweepGrid sg,
                             int *__restrict__ keys, int *__restrict__ hist)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 rel = absToRel(box, v3(pos[i], pos[cap + i], pos[2 * cap + i]));
    double r[3] = {rel.x, rel.y, rel.z};
    int c[3], f = 0;
    for (int d = 0; d < 3; d++) {
        double q = r[d] - sg.shift[d];
        if (q < 0) { q += 1; f |= 1 << d; }
        c[d] = shiftedCoord(q, sg.relCell[d], sg.n[d]);
    }
    int cell = (c[2] * sg.n[1] + c[1]) * sg.n[0] + c[0];
    keys[i] = cell | (f << 28);
    atomicAdd(&hist[cell], 1);
}

__global__ void k_sweep_scatter(int n, const int *__restrict__ keys, const int *__restrict__ cellStart,
                                int *__restrict__ fill, int *__restrict__ itemsTmp)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c = keys[i] & 0x0FFFFFFF;
    itemsTmp[cellStart[c] + atomicAdd(&fill[c], 1)] = i;
}

// {max cell population, max neighbourhood population, cells skipped, peer timeout} -> mapped pinned memory
// ... and the four counters of this sweep added to the caller's accumulators (the cycle loop reads them later)
__global__ void k_sweep_report(const int *__restrict__ flags, int *__restrict__ out,
                               const unsigned long long *__restrict__ counters, unsigned long long *__restrict__ acc) {
    if (threadIdx.x == 0) {
        out[0] = flags[1]; out[1] = flags[2]; out[2] = flags[3]; out[3] = flags[SHARD_ABORT_FLAG];
    }
    if (acc && threadIdx.x < 4) acc[threadIdx.x] += counters[threadIdx.x];
}

// ---------------------------------------------------------------------------------------------------------------
// host driver: one call = one pass of N trial moves (every particle once) with one sampler
// ---------------------------------------------------------------------------------------------------------------
// FP32 cull threshold + a bound on the FP32 error: positions are converted to cell-local coordinates of magnitude
// <= E (1.5 cells along every box vector), so a coordinate difference carries an absolute error <= 4 E 2^-24 and sums
// of squares / projections a relative one of a few 2^-24.
//   spheres:          centre distance < range
//   spherocylinders:  distance(candidate centre, trial's axis segment) < 2r + l/2 — necessary for the two axis
//                     segments to come closer than 2r (every point of the candidate's segment lies within l/2 of
//                     its centre); it implies centre distance < range as well
void cullThresholds(const Handle *h, const double relCell[3], float *thr2, float *half) {
    const Inter &I = h->inter;
    double E = 0;
    for (int i = 0; i < 3; i++)
        E += 1.5 * relCell[i] * std::sqrt(norm2(relToAbs(h->box, v3(i == 0, i == 1, i == 2))));
    double R = I.range, u = std::ldexp(1.0, -24), delta = 4 * E * u;
    double Rc = (I.hard == RP_HARD_SPHEROCYLINDER) ? std::sqrt(I.scInner2) + I.scHalfPlus : R;
    // + 64 u R^2: the sweep kernel evaluates |d|^2 + s (s - 2 d.u) (cancellation of terms up to R^2)
    double margin = 8 * Rc * (8 * delta + 8 * u * R) + 1e-5 * Rc * Rc + 64 * u * R * R;
    *thr2 = std::nextafterf(static_cast<float>(Rc * Rc + margin), INFINITY);
    *half = static_cast<float>(I.scHalfPlus);
}

// k_sweep variants: threads per CTA (= per active cell) and the minimum CTAs per SM the register allocation aims at.
// The default is picked from the number of active cells (one wave or less: more threads per cell so that a launch
// does not cost a full 4-warp CTA lifetime); RP_SWEEP_THREADS / RP_SWEEP_MINB override it (measurement hook).
// ceiling of the dynamic shared memory of the sweep kernels (they carry < 12 KB of static shared memory; B200 allows
// 227 KB per CTA); every capacity check below refuses launches above 200 KB
constexpr size_t SWEEP_SMEM_CEILING = 208 * 1024;
struct SweepVariant { int threads, minb, index; };
static const SweepVariant SWEEP_VARIANTS[] = {{128, 4, 0}, {128, 5, 1}, {128, 6, 2}, {512, 1, 3}, {512, 2, 4},
                                              {256, 2, 5}, {256, 3, 6}};

static SweepVariant chooseVariant(int activeCells) {
    // (a launch of a handful of cells — a 1000-particle gas has ONE active cell per colour — is pure CTA latency:
    // the widest CTA wins)
    int threads = activeCells <= 148 ? 512 : (activeCells <= 148 * 3 ? 256 : 128);
    int minb = threads == 512 ? 1 : (threads == 256 ? 3 : 5);
    if (const char *e = std::getenv("RP_SWEEP_THREADS")) threads = std::atoi(e);
    if (const char *e = std::getenv("RP_SWEEP_MINB")) minb = std::atoi(e);
    for (const SweepVariant &v : SWEEP_VARIANTS)
        if (v.threads == threads && v.minb == minb) return v;
    for (const SweepVariant &v : SWEEP_VARIANTS)
        if (v.threads == threads) return v;
    return SWEEP_VARIANTS[1];
}

static size_t sweepSmemBytes(int capOld, int capOwn, int threads) {
    return static_cast<size_t>(capOld + capOwn) * 32 + static_cast<size_t>(capOwn) * 48
           + static_cast<size_t>(threads / 32) * SW_WQCAP * sizeof(unsigned int);
}

template <int HARD, int THREADS, int MINB>
static void launchColour(Handle *h, int variantIndex, int blocks, size_t smem, const SweepArgs &A, const SweepGrid &sg,
                         const SweepCellOffsets &off, const TrialBuffers &T, unsigned long long *dCounters,
                         const PeerView &peers)
{
    GridStore &g = h->sgrid;
    // The attribute is per DEVICE and per kernel instantiation — shared by every handle of the process on that device
    // — and it is a ceiling, not a reservation (occupancy follows the launch's own size).  It is therefore always set
    // to the largest size any launch may ask for (SWEEP_SMEM_CEILING): a handle with small cells must never lower it
    // under a handle with crowded ones (two handles in one process, e.g. concurrent state points on separate streams).
    size_t &configured = h->smemConfigured[variantIndex + (HARD == RP_HARD_SPHERE ? 8 : 0)];
    if (configured < SWEEP_SMEM_CEILING) {
        RP_CUDA(cudaFuncSetAttribute(k_sweep<HARD, THREADS, MINB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(SWEEP_SMEM_CEILING)));
        configured = SWEEP_SMEM_CEILING;
    }
    k_sweep<HARD, THREADS, MINB><<<blocks, THREADS, smem, h->stream>>>(A, sg, off, h->box, g.cellStart, g.items, g.rec,
                                                                      sweepRec32(g), T, h->pos, h->orient, dCounters,
                                                                      h->dFlags, peers);
}

template <int HARD>
static void launchColourVariant(Handle *h, const SweepVariant &v, int blocks, size_t smem, const SweepArgs &A,
                                const SweepGrid &sg, const SweepCellOffsets &off, const TrialBuffers &T,
                                unsigned long long *dCounters, const PeerView &peers)
{
    switch (v.index) {
        case 0: launchColour<HARD, 128, 4>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        case 1: launchColour<HARD, 128, 5>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        case 2: launchColour<HARD, 128, 6>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        case 3: launchColour<HARD, 512, 1>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        case 4: launchColour<HARD, 512, 2>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        case 5: launchColour<HARD, 256, 2>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
        default: launchColour<HARD, 256, 3>(h, v.index, blocks, smem, A, sg, off, T, dCounters, peers); break;
    }
}

template <int HARD>
static void launchSweep(Handle *h, const SweepArgs &A0, const SweepGrid &sg, const SweepCellOffsets &off,
                        const int colourOrder[8], const TrialBuffers &T, unsigned long long *dCounters)
{
    const bool sharded = shardActive(h);
    const bool translated = A0.kind != RP_MOVE_ROTATION, rotated = A0.kind != RP_MOVE_TRANSLATION;
    int n = static_cast<int>(h->n);
    int activeLayers = A0.nActive[2];
    if (sharded) {
        int aLo, aHi;
        shardSlab(A0.nActive[2], h->shard->nranks, h->shard->rank, &aLo, &aHi);
        activeLayers = aHi - aLo;
        h->shard->lastHaloBytes = h->shard->lastGatherBytes = 0;
    }
    // peer-memory path: accepted moves are written into all replicas by k_sweep itself, flags order the colours
    const bool p2p = sharded && h->shard->p2p;
    PeerView peers;
    unsigned int epoch = 0;
    if (p2p) {
        peers = shardPeerView(h, sg.n[2]);
        epoch = ++h->shard->sweepEpoch;
        // nobody writes into a replica whose grid of this sweep is not built yet
        shardSignalWait(h, SHARD_FLAG_GRID, epoch, true, false);
    }
    const int blocks = A0.nActive[0] * A0.nActive[1] * activeLayers;
    const SweepVariant v = chooseVariant(blocks);
    const size_t smem = sweepSmemBytes(A0.capOld, A0.capOwn, v.threads);
    for (int c = 0; c < 8; c++) {
        SweepArgs A = A0;
        int col = colourOrder[c];
        A.colour[0] = col & 1; A.colour[1] = (col >> 1) & 1; A.colour[2] = (col >> 2) & 1;
        launchColourVariant<HARD>(h, v, blocks, smem, A, sg, off, T, dCounters, peers);
        h->launches++;
        if (p2p) {
            if (c < 7) {        // the z-neighbours read my boundary layers in the next colour
                shardSignalWait(h, SHARD_FLAG_COLOUR, epoch * 8u + c + 1, false, false);
            }
        } else if (sharded) shardHaloExchange(h, col, sg.n[0], sg.n[1], sg.n[2], translated, rotated);
    }
    if (p2p) {                  // all replicas identical again before the next grid build / box move / download
        shardSignalWait(h, SHARD_FLAG_DONE, epoch, true, true, dCounters);
        int aLo, aHi;
        shardSlab(A0.nActive[2], h->shard->nranks, h->shard->rank, &aLo, &aHi);
        shardApplyAccepts(h, aLo, aHi, n, translated, rotated, A0.nActive[2], epoch);
    } else if (sharded) {
        shardGatherCycle(h, sg.n[2], translated, rotated);
        shardAllReduceU64(h, dCounters, 4);
    }
}

static void ensureTrialBuffers(Handle *h) {
    if (h->n > h->trialCap || !h->trialRec) {
        for (void *q : {(void *) h->trialRec, (void *) h->trialOrient, h->trialMeta, h->trialRec32}) if (q) cudaFree(q);
        h->trialRec = h->trialOrient = nullptr;
        h->trialMeta = h->trialRec32 = nullptr;
        size_t cap = h->n + h->n / 8 + 64;
        RP_CUDA(cudaMalloc(&h->trialRec, cap * 6 * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->trialOrient, cap * 9 * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->trialMeta, cap * sizeof(uint2)));
        RP_CUDA(cudaMalloc(&h->trialRec32, cap * 2 * sizeof(float4)));
        h->trialCap = cap;
    }
}

static bool fastPathEligible(const Handle *h) {
    const Inter &I = h->inter;
    return I.k == 0 && I.soft == RP_SOFT_NONE && (I.hard == RP_HARD_SPHERE || I.hard == RP_HARD_SPHEROCYLINDER)
           && !h->overlapCounting && !std::getenv("RP_SWEEP_GENERIC");
}

static void sweepGeneric(Handle *h, const rp_sweep_params &p, const SweepGrid &sg, const int colourOrder[8],
                         const int refN[3], rp_sweep_counters *out, cudaEvent_t *ev);

// the maxima the previous sweeps reported (pinned ring, non-blocking): grow the staging capacities of the next one
static void harvestSweepMaxima(Handle *h, bool block) {
    for (int i = 0; i < 4; i++) {
        if (!h->sweepMaxPending[i]) continue;
        if (block) RP_CUDA(cudaEventSynchronize(h->sweepMaxEvent[i]));
        else if (cudaEventQuery(h->sweepMaxEvent[i]) != cudaSuccess) { cudaGetLastError(); continue; }
        h->sweepMaxPending[i] = false;
        const int *m = h->sweepMaxPinned + 4 * i;
        // head room: the densest neighbourhood of the next cycles is a few particles away from this one's
        h->sweepCapOwn = std::max(h->sweepCapOwn, m[0] + m[0] / 8 + 4);
        h->sweepCapOld = std::max(h->sweepCapOld, m[1] + m[1] / 16 + 16);
        h->sweepCellsSkipped += static_cast<unsigned long long>(m[2]);
        if (m[3]) throw rp_error(RP_ERR_STATE, "rp_sweep: a peer rank did not reach a synchronisation point within "
                                               "20 s (peer timeout); the state of this handle is undefined");
    }
}

void sweepAsync(Handle *h, const rp_sweep_params &p, unsigned long long *dAcc, rp_sweep_counters *out) {
    unsigned long long *dCounters = h->dCounters;
    const Inter &I = h->inter;
    const bool flip = p.kind == RP_MOVE_FLIP;
    const bool fast = fastPathEligible(h) && !flip;
    RP_EXPECT(p.kind >= RP_MOVE_TRANSLATION && p.kind <= RP_MOVE_FLIP, "unknown sampler kind");
    bool translated = p.kind != RP_MOVE_ROTATION, rotated = p.kind != RP_MOVE_TRANSLATION;
    if (flip) {
        // FlipSampler (move_samplers/FlipSampler.cpp:10, 13-16, 45-51)
        RP_EXPECT(p.flip_every > 0, "flipEvery > 0");
        RP_EXPECT(h->n > p.flip_every, "numParticles > this->flipEvery");
        RP_EXPECT(p.flip_axis[0] != 0 || p.flip_axis[1] != 0 || p.flip_axis[2] != 0, "flip axis must not be zero");
        const double EPSILON = 1e-14;
        translated = !(p.flip_origin[0] * p.flip_origin[0] + p.flip_origin[1] * p.flip_origin[1]
                       + p.flip_origin[2] * p.flip_origin[2] < EPSILON * EPSILON);
    } else {
        if (translated) RP_EXPECT(p.translation_step > 0, "translationStepSize > 0");
        if (rotated) RP_EXPECT(p.rotation_step > 0, "rotationStepSize > 0");
    }
    if (I.soft != RP_SOFT_NONE) RP_EXPECT(p.temperature > 0, "temperature > 0");

    double heights[3];
    boxHeights(h->box, heights);
    double volume = boxVolume(h->box);
    // cell width: at least totalRange (plus a rounding margin), and wide enough to hold ~16 particles
    // (the specialised kernel amortises a staged candidate over the trials of its cell: ~16 per cell is its optimum;
    // the generic kernel runs one warp per trial over all staged candidates, so its cells are as small as the range
    // allows: 7.7 -> 4.0 ms per sweep for 5e5 WCA trimers, 2.9 -> 1.9 ms for 2e5 hard 4-mers)
    double TARGET_POPULATION = fast ? 16 : 6;
    if (const char *e = std::getenv("RP_SWEEP_TARGET_POP")) TARGET_POPULATION = std::max(1.0, std::atof(e));      // (measurement hook)
    double minWidth = I.totalRange * (1 + 1e-9);
    int refN[3] = {0, 0, 0};
    if (I.soft == RP_SOFT_LJ) {
        // un-truncated potential: the pair set is "centres in adjacent cells of the reference-geometry grid"
        // (rp_sweep_generic.cuh); two such centres are at most 2 reference cells apart along every axis
        ensureGrid(h);
        if (!h->hasGrid)
            throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: Lennard-Jones needs a box with a neighbour grid");
        double cellMax = 0;
        for (int i = 0; i < 3; i++) {
            refN[i] = h->grid.n[i];
            cellMax = std::max(cellMax, heights[i] / refN[i]);
        }
        minWidth = (2 * cellMax + (I.totalRange - I.range)) * (1 + 1e-9);
    }
    double width = std::max(minWidth, std::cbrt(TARGET_POPULATION * volume / static_cast<double>(h->n)));
    SweepGrid sg;
    size_t numCells = 1;
    for (int i = 0; i < 3; i++) {
        int n = static_cast<int>(floor(heights[i] / width));
        if (n < 2 && heights[i] >= 2 * minWidth) n = 2;
        if (n < 2)
            throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: the box must be at least 2 x totalRange high along every "
                                               "axis for the checkerboard sweep");
        // even counts for the two-colour pattern per axis.  Rounding an odd count DOWN multiplies the cell population
        // by (n / (n - 1))^3 — 3.4 at n = 3, where a box of 10^3 particles collapses to 8 cells of 125, one CTA per
        // colour — so small grids round UP when the narrower cell is still admissible
        if (n & 1) n = (n <= 7 && heights[i] / (n + 1) >= minWidth) ? n + 1 : n - 1;
        sg.n[i] = n;
        sg.relCell[i] = 1 / static_cast<double>(n);
        numCells *= n;
    }
    // random grid origin and colour order from the host side of the same Philox generator.  The cycle number is a
    // 64-bit quantity: its low word is a Philox counter word, its high word is folded into the key.
    const unsigned long long sweepSeed = h->seed ^ ((p.cycle >> 32) * 0x9E3779B97F4A7C15ull);
    Rng rng;
    rng.init(sweepSeed, 0xFFFFFFFFu, static_cast<unsigned int>(p.cycle), static_cast<unsigned int>(p.pass));
    for (int i = 0; i < 3; i++) sg.shift[i] = rng.uniform();
    int colourOrder[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    for (int i = 7; i > 0; i--) {
        int j = static_cast<int>(rng.uniform() * (i + 1));
        if (j > i) j = i;
        std::swap(colourOrder[i], colourOrder[j]);
    }
    for (int f = 0; f < 8; f++)
        sg.imageShift[f] = relToAbs(h->box, v3(f & 1, (f >> 1) & 1, (f >> 2) & 1));
    computeTranslations(h->box, sg.trans);
    h->lastSweepValid = false;
    for (int c = 0; c < 8; c++) { h->lastColourOrder[c] = colourOrder[c]; h->lastImageShift[c] = sg.imageShift[c]; }
    for (int i = 0; i < 3; i++) h->lastSweepN[i] = sg.n[i];
    h->lastSweepRotated = rotated;
    h->lastSg = sg;
    h->sweepSerial++;

    GridStore &g = h->sgrid;
    ensureGridCapacity(g, numCells, h->n, 10);
    for (int i = 0; i < 3; i++) g.n[i] = sg.n[i];
    g.numCells = static_cast<int>(numCells);
    g.numItems = static_cast<int>(h->n);
    g.box = h->box;
    g.cellSize = width;

    cudaStream_t st = h->stream;
    int n = static_cast<int>(h->n);
    // timing events come from a ring: a sweep's device times are collected later without blocking (sweepTimings)
    const int tslot = static_cast<int>(h->sweepSerial % TIMING_RING);
    if (h->timing[tslot].pending) sweepTimings(h, true);
    cudaEvent_t *ev = h->timing[tslot].e;
    RP_CUDA(cudaEventRecord(ev[0], st));
    zeroAsync(h, g.cellFill, (numCells + 1) * sizeof(int), st);
    zeroAsync(h, h->dFlags, 8 * sizeof(int), st);
    zeroAsync(h, h->dCounters, 8 * sizeof(unsigned long long), st);
    k_sweep_keys<<<(n + 255) / 256, 256, 0, st>>>(n, h->cap, h->pos, h->box, sg, g.keys, g.cellFill);
    scanCells(h, g, &h->dFlags[1]);
    k_sweep_scatter<<<(n + 255) / 256, 256, 0, st>>>(n, g.keys, g.cellStart, g.cellFill, g.itemsTmp);
    if (!fast) {
        if (shardActive(h))
            throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: the sharded sweep covers hard spheres / spherocylinders "
                                               "only (other interactions: run replicas)");
        h->launches += 2;
        sweepGeneric(h, p, sg, colourOrder, refN, out, ev);
        return;
    }
    if (shardActive(h) && sg.n[2] / 2 < h->shard->nranks)
        throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: fewer active cell layers along z than ranks; use fewer GPUs "
                                           "or orient the longest box vector along z");

    SweepArgs A;
    A.kind = p.kind;
    A.transStep = p.translation_step;
    A.rotStep = p.rotation_step;
    A.seed = sweepSeed;
    A.cycle = static_cast<unsigned int>(p.cycle);
    A.pass = static_cast<unsigned int>(p.pass);
    for (int i = 0; i < 3; i++) { A.nActive[i] = sg.n[i] / 2; A.colour[i] = 0; }
    A.capOld = A.capOwn = 0;
    {
        cullThresholds(h, sg.relCell, &A.cullThr2, &A.cullHalf);
        // test hook: with the cull disabled every pair takes the exact FP64 test; the results must be bit-identical
        A.cull2Thr2 = std::nextafterf(static_cast<float>(I.scInner2 + 1e-3 * I.range * I.range), INFINITY);
        if (std::getenv("RP_SWEEP_NO_CULL")) A.cullThr2 = A.cull2Thr2 = 3.0e38f;
    }
    A.cap = h->cap;
    A.aOff2 = 0;
    A.slotBegin = 0;
    A.slotEnd = n;
    A.trialLayerLo = A.trialLayerHi = -1;
    A.inter.sphereThr2 = I.sphereThr2;
    A.inter.scInner2 = I.scInner2;
    A.inter.scOuter2 = I.scOuter2;
    A.inter.scHalf = I.scHalfPlus;
    h->lastInvHalf = (I.hard == RP_HARD_SPHEROCYLINDER && A.cullHalf > 0.f) ? 1.0f / A.cullHalf : 0.f;
    ensureTrialBuffers(h);
    TrialBuffers T;
    T.rec = h->trialRec;
    T.orient = h->trialOrient;
    T.meta = static_cast<uint2 *>(h->trialMeta);
    T.rec32 = static_cast<float4 *>(h->trialRec32);
    // centre-to-centre vectors of the 27 neighbour cells in FP32 (the frame change of the staged cull records)
    SweepCellOffsets off;
    for (int code = 0; code < 27; code++) {
        V3 o = relToAbs(h->box, v3((code % 3 - 1) * sg.relCell[0], ((code / 3) % 3 - 1) * sg.relCell[1],
                                   (code / 9 - 1) * sg.relCell[2]));
        off.o[code][0] = static_cast<float>(o.x); off.o[code][1] = static_cast<float>(o.y);
        off.o[code][2] = static_cast<float>(o.z);
    }

    // sharded: the peer mapping (collective, may synchronise) and, over NCCL, the slab's slot range
    bool p2p = false;
    if (shardActive(h)) {
        p2p = shardEnsurePeerMapping(h);
        shardQueueLayerStarts(h, sg.n[0], sg.n[1], sg.n[2], !p2p);
        if (!p2p) {
            // NCCL path: a rank generates the trials of its own slab only -> needs the layer boundaries on the host
            RP_CUDA(cudaStreamSynchronize(st));
            int aLo, aHi;
            shardSlab(sg.n[2] / 2, h->shard->nranks, h->shard->rank, &aLo, &aHi);
            A.slotBegin = h->shard->layerStart[2 * aLo];
            A.slotEnd = h->shard->layerStart[2 * aHi];
        }
        int aLo, aHi;
        shardSlab(sg.n[2] / 2, h->shard->nranks, h->shard->rank, &aLo, &aHi);
        A.aOff2 = aLo;
        if (p2p) { A.trialLayerLo = 2 * aLo; A.trialLayerHi = 2 * aHi; }
    }
    const int prepBlocks = (n + 255) / 256;
    const Walls W = makeWalls(h->box, h->wallMask, I.totalRange);
    k_sweep_neighbourhood<<<static_cast<int>((numCells + 255) / 256), 256, 0, st>>>(g.numCells, sg, g.cellStart,
                                                                                    &h->dFlags[2]);
    if (I.hard == RP_HARD_SPHERE)
        k_sweep_prepare<RP_HARD_SPHERE><<<prepBlocks, 256, 0, st>>>(n, A, sg, h->box, 0.0, g.cellStart, g.itemsTmp,
                                                                    g.keys, h->pos, h->orient, g.items, g.rec,
                                                                    sweepRec32(g), T, W, I.radius);
    else
        k_sweep_prepare<RP_HARD_SPHEROCYLINDER><<<prepBlocks, 256, 0, st>>>(n, A, sg, h->box, I.scHalfPlus,
                                                                            g.cellStart, g.itemsTmp, g.keys, h->pos,
                                                                            h->orient, g.items, g.rec,
                                                                            sweepRec32(g), T, W, I.radius);
    h->launches += 7;
    RP_CUDA(cudaEventRecord(ev[1], st));

    // staging capacities: from the maxima of the previous sweeps when there are any (no host synchronisation), else
    // from this grid (one synchronisation, first call or after the packing changed)
    harvestSweepMaxima(h, false);
    // (the synchronous form — counters returned to the caller — always sizes from its own grid: it never skips a cell)
    if (h->sweepCapOld == 0 || out != nullptr || std::getenv("RP_SWEEP_SYNC_CAPS")) {
        int flags[4];
        readbackSync(h, st, flags, h->dFlags, 4 * sizeof(int));
        RP_CUDA(cudaGetLastError());
        h->sweepCapOwn = std::max(h->sweepCapOwn, flags[1] + flags[1] / 8 + 4);
        h->sweepCapOld = std::max(h->sweepCapOld, flags[2] + flags[2] / 16 + 16);
    }
    g.maxPopulation = h->sweepCapOwn;
    if (h->sweepCapOld >= 65536)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell neighbourhood holds more than 65535 particles");
    A.capOld = h->sweepCapOld;
    A.capOwn = h->sweepCapOwn;
    if (sweepSmemBytes(A.capOld, A.capOwn, 512) > 200 * 1024)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell neighbourhood does not fit into shared memory");

    RP_CUDA(cudaEventRecord(ev[2], st));
    if (I.hard == RP_HARD_SPHERE) launchSweep<RP_HARD_SPHERE>(h, A, sg, off, colourOrder, T, dCounters);
    else launchSweep<RP_HARD_SPHEROCYLINDER>(h, A, sg, off, colourOrder, T, dCounters);
    RP_CUDA(cudaEventRecord(ev[3], st));
    h->timing[tslot].pending = true;
    // {max population, max neighbourhood, cells skipped, peer timeout} of this sweep -> pinned ring, read by a later call
    {
        const int slot = h->sweepMaxNext;
        if (h->sweepMaxPending[slot]) {       // ring full: wait for the oldest
            RP_CUDA(cudaEventSynchronize(h->sweepMaxEvent[slot]));
            harvestSweepMaxima(h, false);
        }
        k_sweep_report<<<1, 32, 0, st>>>(h->dFlags, h->sweepMaxPinned + 4 * slot, h->dCounters, dAcc);
        RP_CUDA(cudaEventRecord(h->sweepMaxEvent[slot], st));
        h->sweepMaxPending[slot] = true;
        h->sweepMaxNext = (slot + 1) % 4;
        h->launches++;
    }
    h->gridDirty = true;       // particles moved: the reference-geometry grid is rebuilt before its next use
    h->hasTrial = false;
    h->lastSweepValid = true;
    h->lastSweepGeneric = false;
    h->lastRecw = 6;
    if (out) {
        // the synchronous form: counters of THIS call
        unsigned long long counters[4];
        readbackSync(h, st, counters, dCounters, sizeof(counters));
        RP_CUDA(cudaGetLastError());
        sweepTimings(h, true);
        const unsigned long long skippedBefore = h->sweepCellsSkipped;
        harvestSweepMaxima(h, true);
        if (h->sweepCellsSkipped != skippedBefore)
            throw rp_error(RP_ERR_CAPACITY, "rp_sweep: shared-memory staging capacity exceeded (cells skipped; the "
                                            "capacity has been raised, the packing is valid)");
        out->attempted = counters[0];
        out->accepted = counters[1];
        out->rejected_domain = counters[2];
        out->rejected_overlap = counters[3];
        out->rejected_energy = 0;
        out->sub_sweeps = 8;
        out->kernel_launches = 8 + 7;
    }
}

// device times of the sweeps whose events have completed (block: of all sweeps queued so far) -> lastSweepMs /
// lastSweepGridMs (the most recent one) and the accumulators the cycle loop reports
void sweepTimings(Handle *h, bool block) {
    for (int k = 0; k < TIMING_RING; k++) {
        // oldest first: slot (sweepSerial + 1 + k) % TIMING_RING
        Handle::TimingSlot &t = h->timing[(h->sweepSerial + 1 + k) % TIMING_RING];
        if (!t.pending) continue;
        if (block) RP_CUDA(cudaEventSynchronize(t.e[3]));
        else if (cudaEventQuery(t.e[3]) != cudaSuccess) { cudaGetLastError(); continue; }
        t.pending = false;
        float msGrid = 0, msSweep = 0;
        cudaEventElapsedTime(&msGrid, t.e[0], t.e[1]);
        cudaEventElapsedTime(&msSweep, t.e[2], t.e[3]);
        h->lastSweepMs = msSweep;
        h->lastSweepGridMs = msGrid;
        h->accSweepMs += msSweep;
        h->accSweepGridMs += msGrid;
    }
}

void sweepHarvest(Handle *h, bool block) { harvestSweepMaxima(h, block); }

bool sweepIsQueued(const Handle *h, int kind) {
    return fastPathEligible(h) && kind != RP_MOVE_FLIP;
}

void sweep(Handle *h, const rp_sweep_params &p, rp_sweep_counters *out) {
    rp_sweep_counters local;
    sweepAsync(h, p, nullptr, out ? out : &local);
}

// ---------------------------------------------------------------------------------------------------------------
// generic path: records, trials and the 8 colour launches of k_gen_sweep (rp_sweep_generic.cuh); called after the
// keys / scan / scatter of the sweep grid
// ---------------------------------------------------------------------------------------------------------------
static GenLayout genLayout(const Inter &I) {
    GenLayout L;
    L.kk = I.kk;
    int w = 3;
    L.offOrient = -1;        // shapes whose pair test needs the whole orientation (XenoCollide)
    if (I.hard == RP_HARD_XENOCOLLIDE) { L.offOrient = w; w += 9; }
    if (I.k > 0) { L.offCentre = w; w += 3 * I.k; } else L.offCentre = 0;
    L.offAux = -1;
    if (I.hard == RP_HARD_SPHEROCYLINDER || I.hard == RP_HARD_POLYSPHEROCYLINDER) { L.offAux = w; w += 3 * I.kk; }
    L.offCell = -1;
    if (I.soft == RP_SOFT_LJ) { w = (w + 1) & ~1; L.offCell = w; w += 2 * I.kk; }
    L.recw = (w + 1) & ~1;
    return L;
}

static void sweepGeneric(Handle *h, const rp_sweep_params &p, const SweepGrid &sg, const int colourOrder[8],
                         const int refN[3], rp_sweep_counters *out, cudaEvent_t *ev)
{
    const Inter &I = h->inter;
    GridStore &g = h->sgrid;
    cudaStream_t st = h->stream;
    const int n = static_cast<int>(h->n);
    GenArgs A;
    std::memset(&A, 0, sizeof(A));
    A.kind = p.kind;
    A.transStep = p.translation_step;
    A.rotStep = p.rotation_step;
    A.temperature = I.soft != RP_SOFT_NONE ? p.temperature : 1.0;
    if (p.kind == RP_MOVE_FLIP) {
        for (int i = 0; i < 3; i++) { A.flipAxis[i] = p.flip_axis[i]; A.flipOrigin[i] = p.flip_origin[i]; }
        A.flipProb = 1.0 / static_cast<double>(p.flip_every);
        const double EPSILON = 1e-14;
        A.flipOriginZero = (p.flip_origin[0] * p.flip_origin[0] + p.flip_origin[1] * p.flip_origin[1]
                            + p.flip_origin[2] * p.flip_origin[2] < EPSILON * EPSILON) ? 1 : 0;
    }
    A.seed = h->seed;
    A.cycle = static_cast<unsigned int>(p.cycle);
    A.pass = static_cast<unsigned int>(p.pass);
    for (int i = 0; i < 3; i++) { A.nActive[i] = sg.n[i] / 2; A.refN[i] = refN[i]; A.refRelCell[i] = refN[i] ? 1.0 / refN[i] : 1.0; }
    A.counting = h->overlapCounting ? 1 : 0;
    double tr = I.totalRange * (1 + 1e-9);
    A.cullRange2 = I.soft == RP_SOFT_LJ ? std::numeric_limits<double>::infinity() : tr * tr;
    A.cap = h->cap;
    A.ccap = h->centreCap;
    A.L = genLayout(I);
    const size_t recw = A.L.recw;

    devEnsure(h->genRec, h->genRecCap, recw * h->n);
    devEnsure(h->genTrialRec, h->genTrialRecCap, recw * h->n);
    devEnsure(h->genTrialCrel, h->genTrialCrelCap, 3 * static_cast<size_t>(std::max(I.k, 1)) * h->n);
    devEnsure(h->genTrialU, h->genTrialUCap, 3 * h->n);
    ensureTrialBuffers(h);
    GenTrialBuffers T;
    T.rec = h->genTrialRec;
    T.orient = h->trialOrient;
    T.crel = h->genTrialCrel;
    T.u = h->genTrialU;
    T.dE = h->genTrialU + h->n;
    T.wall = h->wallMask ? reinterpret_cast<int2 *>(h->genTrialU + 2 * h->n) : nullptr;
    T.meta = static_cast<uint2 *>(h->trialMeta);

    devEnsure(h->genCull, h->genCullCap, 2 * h->n);          // float4 per slot
    devEnsure(h->genTrialCull, h->genTrialCullCap, 2 * h->n);
    float4 *cull = reinterpret_cast<float4 *>(h->genCull);
    T.cull = reinterpret_cast<float4 *>(h->genTrialCull);
    {
        // FP32 cull of mass-centre distances: the FP64 test inside pairTerms (cullRange2) stays the deciding one; the
        // margin covers the rounding of cell-relative coordinates (|x| < a few cell widths, ulp ~ 1e-6)
        const double t32 = tr * (1 + 1e-4) + 1e-3;
        A.cull32Thr2 = I.soft == RP_SOFT_LJ ? std::numeric_limits<float>::infinity()
                                            : std::nextafterf(static_cast<float>(t32 * t32), INFINITY);
        A.cullBox = I.soft == RP_SOFT_LJ ? 1 : 0;
        if (A.cullBox) {
            // adjacent reference cells: the two centres are closer than 2 cells along every box axis; a centre is
            // within (totalRange - range) / 2 of its mass centre, i.e. within |row_i(M^-1)| times that in relative
            // coordinate i; + FP32 rounding of relative coordinates of magnitude < 3 sweep cells
            const double rMol = 0.5 * (I.totalRange - I.range);
            for (int i = 0; i < 3; i++) {
                const double row = std::sqrt(h->box.inv.a[3 * i] * h->box.inv.a[3 * i]
                                             + h->box.inv.a[3 * i + 1] * h->box.inv.a[3 * i + 1]
                                             + h->box.inv.a[3 * i + 2] * h->box.inv.a[3 * i + 2]);
                const double b = (2 * A.refRelCell[i] + 2 * rMol * row) * (1 + 1e-5) + 1e-5 * sg.relCell[i];
                A.cullBoxHalf[i] = std::nextafterf(static_cast<float>(b), INFINITY);
            }
        }
    }
    SweepCellOffsets off;
    for (int code = 0; code < 27; code++) {
        V3 o = v3((code % 3 - 1) * sg.relCell[0], ((code / 3) % 3 - 1) * sg.relCell[1], (code / 9 - 1) * sg.relCell[2]);
        if (!A.cullBox) o = relToAbs(h->box, o);
        off.o[code][0] = static_cast<float>(o.x); off.o[code][1] = static_cast<float>(o.y);
        off.o[code][2] = static_cast<float>(o.z);
    }

    int recBlocks = static_cast<int>((static_cast<size_t>(g.numCells) * 32 + 255) / 256);
    k_gen_records<<<recBlocks, 256, 0, st>>>(g.numCells, A, sg, h->box, h->dInter, g.cellStart, g.itemsTmp, g.keys,
                                             h->pos, h->orient, h->crel, g.items, h->genRec, cull);
    k_sweep_neighbourhood<<<(g.numCells + 255) / 256, 256, 0, st>>>(g.numCells, sg, g.cellStart, &h->dFlags[2]);
    h->launches += 2;
    RP_CUDA(cudaEventRecord(ev[1], st));
    int flags[4];
    readbackSync(h, st, flags, h->dFlags, 4 * sizeof(int));
    g.maxPopulation = flags[1];
    A.capOwn = (std::max(g.maxPopulation, 1) + 1) & ~1;
    A.capOld = std::max(flags[2], 1);
    if (A.capOwn > GEN_MAXCELL)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: more than 512 particles in one cell of the sweep grid");
    // in-cell pair terms: bit masks (hard only), matrices while they fit, else evaluated step by step by the whole CTA
    const bool needTables = (h->overlapCounting && I.hard != RP_HARD_NONE) || I.soft != RP_SOFT_NONE
                            || A.capOwn > GEN_MAXOWN;
    A.matrices = 0;
    size_t smem = genSmemBytes(A.capOld, A.capOwn, static_cast<int>(recw), false);
    if (needTables) {
        const size_t withMatrices = genSmemBytes(A.capOld, A.capOwn, static_cast<int>(recw), true);
        if (A.capOwn <= 2 * GEN_MAXOWN && withMatrices <= 110 * 1024) { A.matrices = 1; smem = withMatrices; }
        else A.matrices = 2;
    }
    // crowded cells (block-wise sequential mode: the wide cells of the un-truncated LJ potential hold ~100 particles
    // and ~2700 frozen candidates): 16 warps per cell instead of 4
    const bool softOn = I.soft != RP_SOFT_NONE;
    const bool wide = A.matrices == 2 && softOn && I.hard != RP_HARD_XENOCOLLIDE;
    const int threads = wide ? GEN_THREADS_WIDE : GEN_THREADS;
    if (wide) smem = genSmemBytes(A.capOld, A.capOwn, static_cast<int>(recw), false, threads);
    if (smem > 200 * 1024)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell does not fit into shared memory");
    RP_CUDA(cudaEventRecord(ev[2], st));
    k_gen_trials<<<(n + 127) / 128, 128, 0, st>>>(n, A, sg, h->box, h->dInter, g.items, g.keys, h->genRec, h->orient,
                                                  h->crel, T, makeWalls(h->box, h->wallMask, I.totalRange), h->pos);
    h->launches++;
    const int blocks = A.nActive[0] * A.nActive[1] * A.nActive[2];
    for (int c = 0; c < 8; c++) {
        GenArgs Ac = A;
        int col = colourOrder[c];
        Ac.colour[0] = col & 1; Ac.colour[1] = (col >> 1) & 1; Ac.colour[2] = (col >> 2) & 1;
        // one instantiation per (hard kind, soft potential present): the pair function is a compile-time constant
        auto launch = [&](auto kernel, int slotIndex) {
            size_t &configured = h->genSmemConfigured[slotIndex];
            if (configured < SWEEP_SMEM_CEILING) {      // (a ceiling shared by all handles on the device: see launchColour)
                RP_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             static_cast<int>(SWEEP_SMEM_CEILING)));
                configured = SWEEP_SMEM_CEILING;
            }
            kernel<<<blocks, threads, smem, st>>>(Ac, sg, off, h->box, h->dInter, g.cellStart, g.items, h->genRec,
                                                  cull, T, h->pos, h->orient, h->crel, h->cabs, h->dCounters,
                                                  h->dFlags);
        };
        const int slotIndex = I.hard * 2 + (softOn ? 1 : 0) + (wide ? 12 : 0);
        constexpr int TN = GEN_THREADS, TW = GEN_THREADS_WIDE;
        switch (I.hard) {
            case RP_HARD_NONE:
                if (wide) launch(k_gen_sweep<RP_HARD_NONE, true, TW>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_NONE, true, TN>, slotIndex);
                break;
            case RP_HARD_SPHERE:
                if (wide) launch(k_gen_sweep<RP_HARD_SPHERE, true, TW>, slotIndex);
                else if (softOn) launch(k_gen_sweep<RP_HARD_SPHERE, true, TN>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_SPHERE, false, TN>, slotIndex);
                break;
            case RP_HARD_SPHEROCYLINDER:
                if (wide) launch(k_gen_sweep<RP_HARD_SPHEROCYLINDER, true, TW>, slotIndex);
                else if (softOn) launch(k_gen_sweep<RP_HARD_SPHEROCYLINDER, true, TN>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_SPHEROCYLINDER, false, TN>, slotIndex);
                break;
            case RP_HARD_POLYSPHERE:
                if (wide) launch(k_gen_sweep<RP_HARD_POLYSPHERE, true, TW>, slotIndex);
                else if (softOn) launch(k_gen_sweep<RP_HARD_POLYSPHERE, true, TN>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_POLYSPHERE, false, TN>, slotIndex);
                break;
            case RP_HARD_POLYSPHEROCYLINDER:
                if (wide) launch(k_gen_sweep<RP_HARD_POLYSPHEROCYLINDER, true, TW>, slotIndex);
                else if (softOn) launch(k_gen_sweep<RP_HARD_POLYSPHEROCYLINDER, true, TN>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_POLYSPHEROCYLINDER, false, TN>, slotIndex);
                break;
            case RP_HARD_XENOCOLLIDE:
                if (softOn) launch(k_gen_sweep<RP_HARD_XENOCOLLIDE, true, TN>, slotIndex);
                else launch(k_gen_sweep<RP_HARD_XENOCOLLIDE, false, TN>, slotIndex);
                break;
            default:
                throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: unknown hard interaction kind");
        }
        h->launches++;
    }
    RP_CUDA(cudaEventRecord(ev[3], st));
    unsigned long long counters[6];
    readbackSync(h, st, counters, h->dCounters, sizeof(counters), flags, h->dFlags, 4 * sizeof(int));
    RP_CUDA(cudaGetLastError());
    float msGrid = 0, msSweep = 0;
    cudaEventElapsedTime(&msGrid, ev[0], ev[1]);
    cudaEventElapsedTime(&msSweep, ev[2], ev[3]);
    h->lastSweepMs = msSweep;
    h->lastSweepGridMs = msGrid;
    h->accSweepMs += msSweep;
    h->accSweepGridMs += msGrid;
    h->gridDirty = true;
    h->hasTrial = false;
    if (flags[3])
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: shared-memory staging capacity exceeded (cells skipped)");
    if (h->overlapCounting)
        h->numOverlaps = static_cast<unsigned long long>(static_cast<long long>(h->numOverlaps)
                                                         + static_cast<long long>(counters[5]));
    h->lastSweepValid = true;
    h->lastSweepGeneric = true;
    h->lastRecw = A.L.recw;
    if (out) {
        out->attempted = counters[0];
        out->accepted = counters[1];
        out->rejected_domain = counters[2];
        out->rejected_overlap = counters[3];
        out->rejected_energy = counters[4];
        out->sub_sweeps = 8;
        out->kernel_launches = 8 + 7;
    }
}

// parity harness: the trials of the last rp_sweep call (device buffers are intact until the next call)
void sweepTrace(Handle *h, rp_trial_trace *out, unsigned long long n) {
    if (!h->lastSweepValid)
        throw rp_error(RP_ERR_STATE, "rp_sweep_trace: no completed rp_sweep call to report");
    cudaStream_t st = h->stream;
    GridStore &g = h->sgrid;
    const size_t recw = h->lastRecw;
    std::vector<int> items(n), keys(n);
    std::vector<uint2> meta(n);
    std::vector<double> rec(n * recw), orient(n * 9), u(2 * n, 0.0), canon(9 * n);
    RP_CUDA(cudaMemcpyAsync(items.data(), g.items, n * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaMemcpyAsync(keys.data(), g.keys, n * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaMemcpyAsync(meta.data(), h->trialMeta, n * sizeof(uint2), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaMemcpyAsync(rec.data(), h->lastSweepGeneric ? h->genTrialRec : h->trialRec,
                            n * recw * sizeof(double), cudaMemcpyDeviceToHost, st));
    const bool haveOrient = h->lastSweepGeneric || h->lastSweepRotated;
    if (haveOrient)
        RP_CUDA(cudaMemcpyAsync(orient.data(), h->trialOrient, n * 9 * sizeof(double), cudaMemcpyDeviceToHost, st));
    else
        for (int q = 0; q < 9; q++)
            RP_CUDA(cudaMemcpyAsync(canon.data() + q * n, h->orient + q * h->cap, n * sizeof(double),
                                    cudaMemcpyDeviceToHost, st));
    if (h->lastSweepGeneric)
        RP_CUDA(cudaMemcpyAsync(u.data(), h->genTrialU, 2 * n * sizeof(double), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    int stepOfColour[8];
    for (int c = 0; c < 8; c++) stepOfColour[h->lastColourOrder[c]] = c;
    const int *sn = h->lastSweepN;
    int prevCell = -1;
    unsigned long long cellFirst = 0;
    for (unsigned long long slot = 0; slot < n; slot++) {
        const int id = items[slot];
        rp_trial_trace &t = out[id];
        const int cell = keys[id] & 0x0FFFFFFF, f = (keys[id] >> 28) & 7;
        if (cell != prevCell) { prevCell = cell; cellFirst = slot; }
        const int cx = cell % sn[0], cy = (cell / sn[0]) % sn[1], cz = cell / (sn[0] * sn[1]);
        t.particle = id;
        t.colour_step = stepOfColour[(cx & 1) | ((cy & 1) << 1) | ((cz & 1) << 2)];
        t.cell = cell;
        t.order_key = meta[slot].y;
        t.slot = static_cast<int>(slot);
        t.status = meta[slot].x == 4u ? 1 : (meta[slot].x == 2u ? 2 : (meta[slot].x == 5u ? 3 : 0));
        // the specialised kernel resolves a cell holding more than SW_MAXOWN trials in batches of SW_MAXOWN slots
        t.batch = h->lastSweepGeneric ? 0 : static_cast<int>((slot - cellFirst) / SW_MAXOWN);
        for (int d = 0; d < 3; d++) t.position[d] = rec[slot * recw + d];
        for (int q = 0; q < 9; q++) t.orientation[q] = haveOrient ? orient[slot * 9 + q] : canon[q * n + id];
        const V3 s = f ? h->lastImageShift[f] : v3(0, 0, 0);
        t.image_shift[0] = s.x; t.image_shift[1] = s.y; t.image_shift[2] = s.z;
        t.u = u[slot];
        t.dE = h->lastSweepGeneric ? u[n + slot]
                                   : (t.status == 1 ? 0.0 : std::numeric_limits<double>::infinity());
    }
}

}  // namespace rp
