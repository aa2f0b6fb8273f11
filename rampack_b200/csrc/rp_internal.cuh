// rp_internal.cuh — handle layout and device views shared by the translation units of librampack_b200.so
#pragma once

#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <stdexcept>

#include "rp_pair.cuh"

namespace rp {

// ---------------------------------------------------------------------------------------------------------------
// Reference-geometry neighbour grid (Packing::neighbourGrid) over the REAL cells only: the reference's ghost layers
// are replaced by periodic wrap + the same 27 lattice translations (core/NeighbourGrid.cpp:80-100, 203-217).
// Cells are flattened x-fastest like the reference (NeighbourGrid.cpp:45-50); items of one cell are contiguous and
// ordered by ascending centre index; pair records are SoA in that order.
// ---------------------------------------------------------------------------------------------------------------
struct GridView {
    int n[3];              // real cells per axis
    int numCells;
    double relCell[3];     // 1 / n[i]  (NeighbourGrid.cpp:196)
    int numItems;
    const int *cellStart;  // [numCells + 1]
    const int *items;      // [numItems] centre index = particle * kk + centre
    const double *px, *py, *pz;   // sorted centre positions (absolute, inside the box)
    const double *ax, *ay, *az;   // sorted aux vectors (rp::pairAux)
    V3 trans[27];          // lattice translations, index (i*3 + j)*3 + k <-> relative (i-1, j-1, k-1)
};

// 0-based REAL cell coordinate of a relative coordinate — NeighbourGrid::positionToCellNo
// (core/NeighbourGrid.cpp:13-34; the reference adds 1 for its ghost layer)
RP_HD int cellCoord(double rel, double relCell, int *outside) {
    const double EPSILON = 2.220446049250313e-16 * 10;
    if (rel < 0) {
        if (!(rel > -EPSILON) && outside) *outside = 1;
        rel = 0;
    } else if (rel >= 1) {
        if (!(rel < 1 + EPSILON) && outside) *outside = 1;
        rel = 1 - EPSILON;
    }
    return static_cast<int>(rel / relCell);
}

RP_HD int cellOfPosition(const GridView &g, const Box &box, V3 pos, int *outside) {
    V3 rel = absToRel(box, pos);
    int cx = cellCoord(rel.x, g.relCell[0], outside);
    int cy = cellCoord(rel.y, g.relCell[1], outside);
    int cz = cellCoord(rel.z, g.relCell[2], outside);
    return (cz * g.n[1] + cy) * g.n[0] + cx;
}

// Neighbour cell (dx,dy,dz in -1..1) of real cell (cx,cy,cz): periodic wrap; *tIdx = index into trans[].
// Same aliasing as NeighbourGrid::getReflectedCellData (core/NeighbourGrid.cpp:80-100).
RP_HD int neighbourCell(const int n[3], int cx, int cy, int cz, int dx, int dy, int dz, int *tIdx) {
    int c[3] = {cx + dx, cy + dy, cz + dz};
    int t[3] = {1, 1, 1};
    for (int i = 0; i < 3; i++) {
        if (c[i] < 0) { c[i] = n[i] - 1; t[i] = 0; }
        else if (c[i] >= n[i]) { c[i] = 0; t[i] = 2; }
    }
    *tIdx = (t[0] * 3 + t[1]) * 3 + t[2];
    return (c[2] * n[1] + c[1]) * n[0] + c[0];
}

// Device storage shared by the two grid flavours (reference-geometry grid and sweep grid)
struct GridStore {
    int n[3] = {0, 0, 0};
    int numCells = 0;
    int numItems = 0;
    double cellSize = 0;
    int *cellStart = nullptr;     // capacity cellCap + 1
    int *cellFill = nullptr;      // capacity cellCap + 1 (histogram / fill counters)
    int *keys = nullptr;          // capacity itemCap
    int *itemsTmp = nullptr;      // capacity itemCap
    int *items = nullptr;         // capacity itemCap
    double *rec = nullptr;        // 6 * itemCap doubles: SoA (reference grid) or AoS 48-byte records (sweep grid)
    size_t cellCap = 0, itemCap = 0;
    int *blockSums = nullptr;     // scan scratch
    size_t blockSumsCap = 0;
    int maxPopulation = 0;
    bool valid = false;
    Box box;                      // box the grid was built for (translations)
};

// geometry of the sweep grid of one rp_sweep call (rp_sweep.cu)
struct SweepGrid {
    int n[3];
    double relCell[3];     // 1 / n[i]
    double shift[3];       // random relative origin s
    V3 imageShift[8];      // M * f for f = (fx, fy, fz) in {0,1}^3, index fx + 2 fy + 4 fz
    V3 trans[27];
};

constexpr int TIMING_RING = 16;
struct SweepStats {
    unsigned long long attempted, accepted, rejDomain, rejOverlap;
};

struct Handle {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool ownStream = false;
    unsigned long long seed = 0;
    std::string error;

    // interaction
    bool hasInteraction = false;
    Inter inter;            // host copy
    Inter *dInter = nullptr;
    XcProgram *xcHost = nullptr, *dXc = nullptr;   // RP_HARD_XENOCOLLIDE geometries (host copy, device copy)

    // particles, canonical (original particle order), SoA with stride `cap`
    size_t n = 0, cap = 0;
    Box box;
    Box lastBox;
    double *pos = nullptr;       // 3 * cap
    double *orient = nullptr;    // 9 * cap
    double *crel = nullptr;      // 3 * centreCap  (Packing::interactionCentres, only k > 0)
    double *cabs = nullptr;      // 3 * centreCap  (Packing::absoluteInteractionCentres, only k > 0)
    size_t centreCap = 0;
    double *lastPos = nullptr;   // tryScaling backup (positions and absolute centres; orientations do not change)
    double *lastCabs = nullptr;
    size_t lastCap = 0, lastCentreCap = 0;
    bool hasLastScaling = false;
    bool lastGridValid = false;     // tryScaling: the grid of the pre-scaling state was up to date (else revert leaves it dirty)
    double interactionRange = 0; // Packing::interactionRange

    // reference-geometry neighbour grid (Packing::neighbourGrid / tempNeighbourGrid)
    GridStore grid, tempGrid;
    bool hasGrid = false, tempHasGrid = false;
    bool gridDirty = true;
    // sweep grid (random origin, particle mass centres, even cell counts)
    GridStore sgrid;
    GridStore ogrid;      // cell list of the g(r) observable (rp_observe.cu)
    int *sweepMoved = nullptr;    // (unused)
    size_t sweepMovedCap = 0;
    // trials of one rp_sweep call, per sorted slot: record (6 doubles), new orientation (9 doubles), meta (uint2),
    // FP32 cull record of the new state (2 float4)
    double *trialRec = nullptr, *trialOrient = nullptr;
    void *trialMeta = nullptr;
    void *trialRec32 = nullptr;
    SweepGrid lastSg;             // sweep grid of the last rp_sweep call (sharded apply kernels, trace)
    float lastInvHalf = 0.f;
    // shared-memory staging capacities of k_sweep: taken from the maxima the previous calls observed (read back
    // without blocking), so that a cycle needs no host synchronisation; 0 = unknown (first call synchronises)
    int sweepCapOld = 0, sweepCapOwn = 0;
    unsigned int *mailbox = nullptr;      // pinned, device-visible: small results a kernel writes for the host (readbackSync)
    int *sweepMaxPinned = nullptr;        // pinned ring of {max cell population, max neighbourhood, cells skipped, -}
    cudaEvent_t sweepMaxEvent[4] = {nullptr, nullptr, nullptr, nullptr};
    bool sweepMaxPending[4] = {false, false, false, false};
    int sweepMaxNext = 0;
    unsigned long long sweepCellsSkipped = 0;   // cells a sweep left out because the staging capacity was too small
    unsigned long long sweepSerial = 0;         // rp_sweep calls made on this handle (timing ring)
    unsigned long long simCycles = 0;           // cycles run by rp_sim loops on this handle (Philox counter of their sweeps)
    size_t genSmemConfigured[32] = {};     // ... of the k_gen_sweep instantiations (hard kind * 2 + soft, + 12 wide)
    size_t smemConfigured[24] = {};   // cudaFuncAttributeMaxDynamicSharedMemorySize set per kernel variant
    size_t trialCap = 0;
    // what rp_sweep_trace needs to know about the last rp_sweep call
    bool lastSweepValid = false, lastSweepGeneric = false, lastSweepRotated = false;
    int lastColourOrder[8] = {0, 1, 2, 3, 4, 5, 6, 7};
    int lastSweepN[3] = {0, 0, 0};
    int lastRecw = 6;
    V3 lastImageShift[8];
    // generic sweep (rp_sweep_generic.cuh): whole-particle records in cell order and the trial buffers
    double *genRec = nullptr, *genTrialRec = nullptr, *genTrialCrel = nullptr, *genTrialU = nullptr;
    double *genCull = nullptr, *genTrialCull = nullptr;      // float4 per slot: FP32 cull points (rp_sweep_generic.cuh)
    size_t genCullCap = 0, genTrialCullCap = 0;
    size_t genRecCap = 0, genTrialRecCap = 0, genTrialCrelCap = 0, genTrialUCap = 0;

    // explicit trial move (Packing's temp slot): pos[3] orient[9] crel[16][3] cabs[16][3]
    double *tmpState = nullptr;
    bool hasTrial = false;
    unsigned long long trialParticle = 0;
    int trialKind = 0;
    long long lastMoveOverlapDelta = 0;

    int wallMask = 0;             // Packing::hasWall, bit per axis (Packing.cpp:1075-1084)
    // overlap counting (Packing::overlapCounting / numOverlaps)
    bool overlapCounting = false;
    unsigned long long numOverlaps = 0;
    unsigned long long lastScalingNumOverlaps = 0;

    // scratch
    unsigned long long *dCounters = nullptr;  // 32 counters
    int *dFlags = nullptr;                    // 16 flags
    double *dCellEnergy = nullptr;
    size_t cellEnergyCap = 0;
    double *dMoveOut = nullptr;               // 8 doubles
    void *stage = nullptr;        // pinned host staging
    size_t stageCap = 0;
    void *dStage = nullptr;       // device staging for AoS <-> SoA transposes
    size_t dStageCap = 0;
    // asynchronous snapshot (rp_snapshot_async): device-side copy + second stream
    void *dSnap = nullptr;
    size_t dSnapCap = 0;
    cudaStream_t snapStream = nullptr;
    cudaEvent_t snapReady = nullptr, snapDone = nullptr;
    bool snapPending = false;

    // spatial sharding over the GPUs of one node (rp_shard.cuh); nullptr = single GPU
    struct Shard *shard = nullptr;

    // timing of the last call
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    double lastSweepMs = 0, lastSweepGridMs = 0, lastRebuildMs = 0, lastScalingMs = 0;
    // sweeps are queued without host synchronisation: their timing events live in a ring and are read back later
    struct TimingSlot { cudaEvent_t e[4] = {nullptr, nullptr, nullptr, nullptr}; bool pending = false; };
    TimingSlot timing[16];
    double accSweepMs = 0, accSweepGridMs = 0;      // device time of all sweeps harvested so far
    unsigned long long launches = 0;
};

struct rp_error : std::runtime_error {
    int status;
    rp_error(int status, const std::string &msg) : std::runtime_error(msg), status(status) {}
};

#define RP_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    throw rp::rp_error(RP_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)
#define RP_EXPECT(cond, msg) do { if (!(cond)) \
    throw rp::rp_error(RP_ERR_PRECONDITION, std::string("Precondition (") + #cond + ") failed: " + (msg)); } while (0)

struct Handle;
// Control traffic of the cycle stays off the copy engines: a few result words read back through cudaMemcpyAsync, or a
// counter cleared through cudaMemsetAsync, queue on the same DMA engine as the 97 MB transfers of OTHER handles on the
// device (packings in flight, snapshots) and wait for them — 1-2 ms per readback when both directions are busy.
//   zeroAsync:    a kernel clears `bytes` (multiple of 4) at p
//   copyAsync:    a kernel copies `bytes` (multiple of 8) between device buffers (the saved state of a box move)
//   readbackSync: a kernel copies up to two small device arrays (multiples of 4 bytes, <= 1 KB together) into the
//                 handle's pinned mailbox, the stream is synchronised, the words are handed to dst0 / dst1
constexpr size_t MAILBOX_BYTES = 1024;
void zeroAsync(Handle *h, void *p, size_t bytes, cudaStream_t st);
void copyAsync(Handle *h, void *dst, const void *src, size_t bytes, cudaStream_t st);      // device to device
void readbackSync(Handle *h, cudaStream_t st, void *dst0, const void *dsrc0, size_t bytes0, void *dst1 = nullptr,
                  const void *dsrc1 = nullptr, size_t bytes1 = 0);

template <typename T>
inline void devEnsure(T *&p, size_t &cap, size_t need) {
    if (need <= cap && p) return;
    if (p) cudaFree(p);
    p = nullptr;
    size_t newCap = need + need / 4 + 16;
    RP_CUDA(cudaMalloc(&p, newCap * sizeof(T)));
    cap = newCap;
}

inline void computeTranslations(const Box &box, V3 trans[27]) {
    // NeighbourGrid::calculateTranslations (core/NeighbourGrid.cpp:203-217)
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++)
            for (int k = 0; k < 3; k++)
                trans[(i * 3 + j) * 3 + k] = relToAbs(box, v3(static_cast<double>(i) - 1,
                                                              static_cast<double>(j) - 1.,
                                                              static_cast<double>(k) - 1.));
}

// ---- rp_grid.cu ----
// rec holds recDoublesPerItem * itemCap doubles (6 = the FP64 records; the sweep grid asks for 10: its 32-byte FP32
// cull records follow the FP64 records in the same allocation, see sweepRec32)
void ensureGridCapacity(GridStore &g, size_t numCells, size_t numItems, size_t recDoublesPerItem = 6);
inline float4 *sweepRec32(const GridStore &g) { return reinterpret_cast<float4 *>(g.rec + 6 * g.itemCap); }
// histogram in g.cellFill -> g.cellStart (exclusive scan), g.cellFill zeroed, max population -> dMaxPop
void scanCells(Handle *h, GridStore &g, int *dMaxPop);
// keys -> items (cell order, ascending index inside a cell); needs scanCells() first
void scatterItems(Handle *h, GridStore &g);
GridView makeView(const GridStore &g);
void rebuildReferenceGrid(Handle *h);
void ensureGrid(Handle *h);
void freeGrid(GridStore &g);
// ---- rp_recount.cu ----
unsigned long long countTotalOverlaps(Handle *h, bool earlyExit);
double totalEnergy(Handle *h);
void initCentres(Handle *h);
void recalcAbsoluteCentres(Handle *h);
double tryMove(Handle *h, const rp_move &m);
void particleScan(Handle *h, bool wantOverlaps, bool wantEnergy, unsigned char *dOverlapFlag, double *dEnergy);
double particleEnergyFluctuations(Handle *h);
unsigned long long countWallOverlaps(Handle *h, bool earlyExit);
void acceptMove(Handle *h);
double tryScaling(Handle *h, const Box &newBox);
void revertScaling(Handle *h);
void pairTestBatch(Handle *h, const rp_pair *pairs, unsigned long long n, unsigned char *verdict, double *energy);
// ---- rp_xc_setup.cu ----
void buildXcProgram(const rp_xc_program &src, int nRoots, XcProgram &dst);
// ---- rp_stats.cu ----
void neighbourStats(Handle *h, double *kCand, double *kFull);
void fp64Peak(Handle *h, double *dfmaGflops, double *dmulDaddGflops);
// ---- rp_observe.cu ----
void nematicOrder(Handle *h, const double axis[3], double qOut[9], double *p2Out);
void smecticOrder(Handle *h, const int range[3], const double focal[3], double *tauAbs, int hklOut[3], double kOut[3]);
void pairDensityCorrelation(Handle *h, double maxR, int numBins, const double focal[3], unsigned long long *countsOut,
                            double *gOut);
// ---- rp_sweep.cu ----
void sweep(Handle *h, const rp_sweep_params &p, rp_sweep_counters *out);
// the same without a host synchronisation when out == nullptr (fast path): the four counters {attempted, accepted,
// rejected: left the cell, rejected: overlap} are ADDED to dAcc[0..3] on the device (dAcc may be nullptr)
void sweepAsync(Handle *h, const rp_sweep_params &p, unsigned long long *dAcc, rp_sweep_counters *out);
void sweepTimings(Handle *h, bool block);
// staging-capacity reports of the queued sweeps (block: of all of them); throws on a peer timeout
void sweepHarvest(Handle *h, bool block);
// does a sweep with this sampler run without a host synchronisation (specialised kernel)?
bool sweepIsQueued(const Handle *h, int kind);
void cullThresholds(const Handle *h, const double relCell[3], float *thr2, float *half);
void sweepTrace(Handle *h, rp_trial_trace *out, unsigned long long n);

}  // namespace rp

struct rp_handle : rp::Handle {};
