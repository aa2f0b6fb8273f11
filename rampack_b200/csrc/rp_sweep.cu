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

// one warp per cell: order the items by ascending particle index and write the 48-byte AoS records
// {P'x, P'y, P'z, aux x, aux y, aux z} in the image frame; also the 27-neighbourhood population maximum
template <int HARD>
__global__ void k_sweep_records(int numCells, size_t cap, const int *__restrict__ cellStart,
                                const int *__restrict__ itemsTmp, const int *__restrict__ keys,
                                const double *__restrict__ pos, const double *__restrict__ orient, SweepGrid sg,
                                double halfLength, int *__restrict__ items, double *__restrict__ rec,
                                int *__restrict__ maxNeigh)
{
    int cell = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (cell >= numCells) return;
    int start = cellStart[cell], end = cellStart[cell + 1];
    int cnt = end - start;
    for (int i = lane; i < cnt; i += 32) {
        int id = itemsTmp[start + i];
        int rank = 0;
        for (int j = 0; j < cnt; j++) rank += (itemsTmp[start + j] < id);
        int slot = start + rank;
        items[slot] = id;
        int f = (keys[id] >> 28) & 7;
        V3 p = v3(pos[id], pos[cap + id], pos[2 * cap + id]);
        if (f) p = p + sg.imageShift[f];
        double *r = rec + 6 * static_cast<size_t>(slot);
        r[0] = p.x; r[1] = p.y; r[2] = p.z;
        if (HARD == RP_HARD_SPHEROCYLINDER) {
            // half axis h = orientation.column(2) * (0.5 * l): the cap centres are pos -+ h (SpherocylinderTraits.cpp:24-26)
            r[3] = orient[2 * cap + id] * halfLength; r[4] = orient[5 * cap + id] * halfLength;
            r[5] = orient[8 * cap + id] * halfLength;
        } else {
            r[3] = 0; r[4] = 0; r[5] = 0;
        }
    }
    // population of the 27-neighbourhood (smem sizing of the sweep kernel)
    int cx = cell % sg.n[0], cy = (cell / sg.n[0]) % sg.n[1], cz = cell / (sg.n[0] * sg.n[1]);
    int sum = 0;
    if (lane < 27) {
        int tIdx;
        int c2 = neighbourCell(sg.n, cx, cy, cz, lane / 9 - 1, (lane / 3) % 3 - 1, lane % 3 - 1, &tIdx);
        sum = cellStart[c2 + 1] - cellStart[c2];
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if (lane == 0) atomicMax(maxNeigh, sum);
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

constexpr int SW_THREADS = 128;

template <int HARD>
static void launchSweep(Handle *h, const SweepArgs &A0, const SweepGrid &sg, const int colourOrder[8], size_t smem,
                        const TrialBuffers &T)
{
    static size_t configured = 0;
    if (smem > configured) {
        RP_CUDA(cudaFuncSetAttribute(k_sweep<HARD, SW_THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(std::max<size_t>(smem, 48 * 1024))));
        configured = std::max<size_t>(smem, 48 * 1024);
    }
    GridStore &g = h->sgrid;
    int n = static_cast<int>(h->n);
    const bool sharded = shardActive(h);
    const bool translated = A0.kind != RP_MOVE_ROTATION, rotated = A0.kind != RP_MOVE_TRANSLATION;
    int activeLayers = A0.nActive[2];
    if (sharded) {
        int aLo, aHi;
        shardSlab(A0.nActive[2], h->shard->nranks, h->shard->rank, &aLo, &aHi);
        activeLayers = aHi - aLo;
        h->shard->lastHaloBytes = h->shard->lastGatherBytes = 0;
    }
    // peer-memory path: accepted moves are written into all replicas by k_sweep itself, flags order the colours
    const bool p2p = sharded && shardEnsurePeerMapping(h);
    PeerView peers;
    unsigned int epoch = 0;
    if (p2p) {
        peers = shardPeerView(h, sg.n[2]);
        epoch = ++h->shard->sweepEpoch;
    }
    SweepArgs AT = A0;
    if (p2p) { AT.slotBegin = 0; AT.slotEnd = n; }       // every rank holds the trials of all slots
    int nTrials = AT.slotEnd - AT.slotBegin;
    if (nTrials > 0)
        k_sweep_trials<HARD><<<(nTrials + 127) / 128, 128, 0, h->stream>>>(
            n, AT, sg, h->box, h->inter.scHalfPlus, g.items, g.keys, g.rec, h->orient, T,
            makeWalls(h->box, h->wallMask, h->inter.totalRange), h->inter.radius);
    h->launches++;
    if (p2p) {      // nobody writes into a replica whose grid of this sweep is not built yet
        shardSignalWait(h, SHARD_FLAG_GRID, epoch, true, false);
    }
    for (int c = 0; c < 8; c++) {
        SweepArgs A = A0;
        int col = colourOrder[c];
        A.colour[0] = col & 1; A.colour[1] = (col >> 1) & 1; A.colour[2] = (col >> 2) & 1;
        int blocks = A.nActive[0] * A.nActive[1] * activeLayers;
        k_sweep<HARD, SW_THREADS><<<blocks, SW_THREADS, smem, h->stream>>>(A, sg, h->box, g.cellStart,
                                                                          g.items, g.rec, T, h->pos, h->orient,
                                                                          h->dCounters, h->dFlags, peers);
        h->launches++;
        if (p2p) {
            if (c < 7) {        // the z-neighbours read my boundary layers in the next colour
                shardSignalWait(h, SHARD_FLAG_COLOUR, epoch * 8u + c + 1, false, false);
            }
        } else if (sharded) shardHaloExchange(h, col, sg.n[0], sg.n[1], sg.n[2], translated, rotated);
    }
    if (p2p) {                  // all replicas identical again before the next grid build / box move / download
        shardSignalWait(h, SHARD_FLAG_DONE, epoch, true, true);
        shardApplyAccepts(h, A0.slotBegin, A0.slotEnd, n, translated, rotated);
    } else if (sharded) {
        shardGatherCycle(h, sg.n[2], translated, rotated);
        shardAllReduceU64(h, h->dCounters, 4);
    }
}

static void ensureTrialBuffers(Handle *h) {
    if (h->n > h->trialCap || !h->trialRec) {
        for (void *q : {(void *) h->trialRec, (void *) h->trialOrient, h->trialMeta}) if (q) cudaFree(q);
        h->trialRec = h->trialOrient = nullptr;
        h->trialMeta = nullptr;
        size_t cap = h->n + h->n / 8 + 64;
        RP_CUDA(cudaMalloc(&h->trialRec, cap * 6 * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->trialOrient, cap * 9 * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->trialMeta, cap * sizeof(uint2)));
        h->trialCap = cap;
    }
}

static bool fastPathEligible(const Handle *h) {
    const Inter &I = h->inter;
    return I.k == 0 && I.soft == RP_SOFT_NONE && (I.hard == RP_HARD_SPHERE || I.hard == RP_HARD_SPHEROCYLINDER)
           && !h->overlapCounting && !std::getenv("RP_SWEEP_GENERIC");
}

static void sweepGeneric(Handle *h, const rp_sweep_params &p, const SweepGrid &sg, const int colourOrder[8],
                         const int refN[3], rp_sweep_counters *out);

void sweep(Handle *h, const rp_sweep_params &p, rp_sweep_counters *out) {
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
    const double TARGET_POPULATION = 16;
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
        n &= ~1;
        sg.n[i] = n;
        sg.relCell[i] = 1 / static_cast<double>(n);
        numCells *= n;
    }
    // random grid origin and colour order from the host side of the same Philox generator
    Rng rng;
    rng.init(h->seed, 0xFFFFFFFFu, static_cast<unsigned int>(p.cycle), static_cast<unsigned int>(p.pass));
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

    GridStore &g = h->sgrid;
    ensureGridCapacity(g, numCells, h->n);
    for (int i = 0; i < 3; i++) g.n[i] = sg.n[i];
    g.numCells = static_cast<int>(numCells);
    g.numItems = static_cast<int>(h->n);
    g.box = h->box;
    g.cellSize = width;

    cudaStream_t st = h->stream;
    int n = static_cast<int>(h->n);
    RP_CUDA(cudaEventRecord(h->ev[0], st));
    RP_CUDA(cudaMemsetAsync(g.cellFill, 0, (numCells + 1) * sizeof(int), st));
    RP_CUDA(cudaMemsetAsync(h->dFlags, 0, 8 * sizeof(int), st));
    RP_CUDA(cudaMemsetAsync(h->dCounters, 0, 8 * sizeof(unsigned long long), st));
    k_sweep_keys<<<(n + 255) / 256, 256, 0, st>>>(n, h->cap, h->pos, h->box, sg, g.keys, g.cellFill);
    scanCells(h, g, &h->dFlags[1]);
    k_sweep_scatter<<<(n + 255) / 256, 256, 0, st>>>(n, g.keys, g.cellStart, g.cellFill, g.itemsTmp);
    if (!fast) {
        if (shardActive(h))
            throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: the sharded sweep covers hard spheres / spherocylinders "
                                               "only (other interactions: run replicas)");
        h->launches += 2;
        sweepGeneric(h, p, sg, colourOrder, refN, out);
        return;
    }
    if (shardActive(h) && sg.n[2] / 2 < h->shard->nranks)
        throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: fewer active cell layers along z than ranks; use fewer GPUs "
                                           "or orient the longest box vector along z");
    int recBlocks = static_cast<int>((numCells * 32 + 255) / 256);
    if (I.hard == RP_HARD_SPHERE)
        k_sweep_records<RP_HARD_SPHERE><<<recBlocks, 256, 0, st>>>(g.numCells, h->cap, g.cellStart, g.itemsTmp,
                                                                    g.keys, h->pos, h->orient, sg, 0.0, g.items,
                                                                    g.rec, &h->dFlags[2]);
    else
        k_sweep_records<RP_HARD_SPHEROCYLINDER><<<recBlocks, 256, 0, st>>>(g.numCells, h->cap, g.cellStart,
                                                                            g.itemsTmp, g.keys, h->pos, h->orient,
                                                                            sg, I.scHalfPlus, g.items, g.rec,
                                                                            &h->dFlags[2]);
    h->launches += 3;
    if (shardActive(h)) shardQueueLayerStarts(h, sg.n[0], sg.n[1], sg.n[2]);
    RP_CUDA(cudaEventRecord(h->ev[1], st));
    int flags[4];
    RP_CUDA(cudaMemcpyAsync(flags, h->dFlags, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    g.maxPopulation = flags[1];
    int maxNeigh = flags[2];
    if (maxNeigh >= 65536)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell neighbourhood holds more than 65535 particles");
    // staging capacities: records are 48 bytes, so any count keeps the 16-byte alignment of the bulk copies
    const int capOld = maxNeigh, capOwn = g.maxPopulation;
    const int capCand = capOld + capOwn;
    size_t smem = static_cast<size_t>(capCand) * 6 * sizeof(double)
                  + (SW_THREADS / 32) * SW_WQCAP * sizeof(unsigned int);
    if (smem > 200 * 1024)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell neighbourhood does not fit into shared memory");

    SweepArgs A;
    A.kind = p.kind;
    A.transStep = p.translation_step;
    A.rotStep = p.rotation_step;
    A.seed = h->seed;
    A.cycle = static_cast<unsigned int>(p.cycle);
    A.pass = static_cast<unsigned int>(p.pass);
    for (int i = 0; i < 3; i++) { A.nActive[i] = sg.n[i] / 2; A.colour[i] = 0; }
    A.capOld = capOld;
    A.capOwn = capOwn;
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
    if (shardActive(h)) {
        int aLo, aHi;
        shardSlab(sg.n[2] / 2, h->shard->nranks, h->shard->rank, &aLo, &aHi);
        A.aOff2 = aLo;
        A.slotBegin = h->shard->layerStart[2 * aLo];
        A.slotEnd = h->shard->layerStart[2 * aHi];
    }
    A.inter.sphereThr2 = I.sphereThr2;
    A.inter.scInner2 = I.scInner2;
    A.inter.scOuter2 = I.scOuter2;
    A.inter.scHalf = I.scHalfPlus;
    ensureTrialBuffers(h);
    TrialBuffers T;
    T.rec = h->trialRec;
    T.orient = h->trialOrient;
    T.meta = static_cast<uint2 *>(h->trialMeta);
    RP_CUDA(cudaEventRecord(h->ev[2], st));
    if (I.hard == RP_HARD_SPHERE) launchSweep<RP_HARD_SPHERE>(h, A, sg, colourOrder, smem, T);
    else launchSweep<RP_HARD_SPHEROCYLINDER>(h, A, sg, colourOrder, smem, T);
    RP_CUDA(cudaEventRecord(h->ev[3], st));
    unsigned long long counters[4];
    RP_CUDA(cudaMemcpyAsync(counters, h->dCounters, sizeof(counters), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaMemcpyAsync(flags, h->dFlags, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    float msGrid = 0, msSweep = 0;
    cudaEventElapsedTime(&msGrid, h->ev[0], h->ev[1]);
    cudaEventElapsedTime(&msSweep, h->ev[2], h->ev[3]);
    h->lastSweepMs = msSweep;
    h->lastSweepGridMs = msGrid;
    h->gridDirty = true;       // particles moved: the reference-geometry grid is rebuilt before its next use
    h->hasTrial = false;
    if (flags[3])
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: shared-memory staging capacity exceeded (cells skipped)");
    h->lastSweepValid = true;
    h->lastSweepGeneric = false;
    h->lastRecw = 6;
    if (out) {
        out->attempted = counters[0];
        out->accepted = counters[1];
        out->rejected_domain = counters[2];
        out->rejected_overlap = counters[3];
        out->rejected_energy = 0;
        out->sub_sweeps = 8;
        out->kernel_launches = 8 + 7;
    }
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
    if (I.soft == RP_SOFT_LJ) { L.offCell = w; w += I.kk; }
    L.recw = (w + 1) & ~1;
    return L;
}

static void sweepGeneric(Handle *h, const rp_sweep_params &p, const SweepGrid &sg, const int colourOrder[8],
                         const int refN[3], rp_sweep_counters *out)
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

    int recBlocks = static_cast<int>((static_cast<size_t>(g.numCells) * 32 + 255) / 256);
    k_gen_records<<<recBlocks, 256, 0, st>>>(g.numCells, A, sg, h->box, h->dInter, g.cellStart, g.itemsTmp, g.keys,
                                             h->pos, h->orient, h->crel, g.items, h->genRec);
    h->launches++;
    RP_CUDA(cudaEventRecord(h->ev[1], st));
    int flags[4];
    RP_CUDA(cudaMemcpyAsync(flags, h->dFlags, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    g.maxPopulation = flags[1];
    A.capOwn = std::max(g.maxPopulation, 1);
    size_t smem = static_cast<size_t>(A.capOwn) * (16 * recw + 48) + 64;
    if (smem > 200 * 1024)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell does not fit into shared memory");
    static size_t configured = 0;
    if (smem > configured) {
        RP_CUDA(cudaFuncSetAttribute(k_gen_sweep, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(std::max<size_t>(smem, 48 * 1024))));
        configured = std::max<size_t>(smem, 48 * 1024);
    }
    RP_CUDA(cudaEventRecord(h->ev[2], st));
    k_gen_trials<<<(n + 127) / 128, 128, 0, st>>>(n, A, sg, h->box, h->dInter, g.items, g.keys, h->genRec, h->orient,
                                                  h->crel, T, makeWalls(h->box, h->wallMask, I.totalRange), h->pos);
    h->launches++;
    for (int c = 0; c < 8; c++) {
        GenArgs Ac = A;
        int col = colourOrder[c];
        Ac.colour[0] = col & 1; Ac.colour[1] = (col >> 1) & 1; Ac.colour[2] = (col >> 2) & 1;
        int blocks = A.nActive[0] * A.nActive[1] * A.nActive[2];
        k_gen_sweep<<<blocks, GEN_THREADS, smem, st>>>(Ac, sg, h->box, h->dInter, g.cellStart, g.items, h->genRec, T,
                                                       h->pos, h->orient, h->crel, h->cabs, h->dCounters, h->dFlags);
        h->launches++;
    }
    RP_CUDA(cudaEventRecord(h->ev[3], st));
    unsigned long long counters[6];
    RP_CUDA(cudaMemcpyAsync(counters, h->dCounters, sizeof(counters), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaMemcpyAsync(flags, h->dFlags, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    float msGrid = 0, msSweep = 0;
    cudaEventElapsedTime(&msGrid, h->ev[0], h->ev[1]);
    cudaEventElapsedTime(&msSweep, h->ev[2], h->ev[3]);
    h->lastSweepMs = msSweep;
    h->lastSweepGridMs = msGrid;
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
