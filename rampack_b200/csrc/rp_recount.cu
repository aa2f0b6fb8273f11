// rp_recount.cu — full-system recount (north_star kernel (c)), explicit trial moves and the box move.
//
// Replaces, with the reference's pair enumeration (same (i, j, lattice image) set):
//   Packing::countTotalOverlaps / countTotalOverlapsNGCellHelper   core/Packing.cpp:431-559
//   Packing::getTotalEnergy / getTotalEnergyNGCellHelper            core/Packing.cpp:690-709, 766-839
//   Packing::tryTranslation / tryRotation / tryMove / acceptMove    core/Packing.cpp:84-164, 278-302
//   Packing::countParticleOverlaps / calculateParticleEnergy        core/Packing.cpp:373-429, 649-688
//   Packing::tryScaling / revertScaling                             core/Packing.cpp:172-208, 363-371
#include "rp_internal.cuh"
#include "rp_recount_cull.cuh"
#include "rp_shard.cuh"
#include "rp_wall.cuh"
#include <limits>
#include <cstdlib>

namespace rp {

// ---------------------------------------------------------------------------------------------------------------
// interaction centres: Packing::setupForInteraction / recalculateAbsoluteInteractionCentres (Packing.cpp:931-950,
// 1035-1047)
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_init_centres(int n, int k, size_t cap, size_t ccap, const double *__restrict__ orient,
                               const Inter *__restrict__ inter, double *__restrict__ crel)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k) return;
    int particle = idx / k, c = idx % k;
    M33 o;
    for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + particle];
    V3 r = mul(o, v3(inter->cpos[c][0], inter->cpos[c][1], inter->cpos[c][2]));
    crel[idx] = r.x; crel[ccap + idx] = r.y; crel[2 * ccap + idx] = r.z;
}

__global__ void k_recalc_abs(int n, int k, size_t cap, size_t ccap, const double *__restrict__ pos,
                             const double *__restrict__ crel, Box box, double *__restrict__ cabs)
{
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k) return;
    int particle = idx / k;
    V3 p = v3(pos[particle], pos[cap + particle], pos[2 * cap + particle])
           + v3(crel[idx], crel[ccap + idx], crel[2 * ccap + idx]);
    V3 a = p + pbcCorrection(box, p);
    cabs[idx] = a.x; cabs[ccap + idx] = a.y; cabs[2 * ccap + idx] = a.z;
}

void initCentres(Handle *h) {
    int k = h->inter.k;
    if (k == 0) return;
    int total = static_cast<int>(h->n) * k;
    k_init_centres<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(h->n), k, h->cap, h->centreCap,
                                                                h->orient, h->dInter, h->crel);
    h->launches++;
}

void recalcAbsoluteCentres(Handle *h) {
    int k = h->inter.k;
    if (k == 0) return;
    int total = static_cast<int>(h->n) * k;
    k_recalc_abs<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(h->n), k, h->cap, h->centreCap, h->pos,
                                                              h->crel, h->box, h->cabs);
    h->launches++;
}

// ---------------------------------------------------------------------------------------------------------------
// full recount with the neighbour grid: one warp per real cell; half-shell rule 9a+3b+c > 13 on the offsets
// (core/NeighbourGrid.cpp:127-131) with a = x offset
// ---------------------------------------------------------------------------------------------------------------
// Interaction::overlapBetween for any supported hard part.  Shapes whose test needs the whole orientation matrix
// (XenoCollide) fetch it from the canonical SoA arrays only when the sphere pre-filters do not decide; o1 != nullptr
// overrides the first particle's orientation (the trial state of an explicit move).
__device__ __forceinline__ M33 loadOrientSoA(const double *__restrict__ orient, size_t cap, int particle) {
    M33 o;
    for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + particle];
    return o;
}
template <typename BC>
__device__ __forceinline__ bool overlapAny(const Inter &I, const double *__restrict__ orient, size_t cap, V3 pos1,
                                           V3 aux1, const M33 *o1, int particle1, int c1, V3 pos2, V3 aux2,
                                           int particle2, int c2, const BC &bc)
{
    if (needsOrientation(I))
        return overlapBetweenOriented(
            I, pos1, [&] { return o1 ? *o1 : loadOrientSoA(orient, cap, particle1); }, c1, pos2,
            [&] { return loadOrientSoA(orient, cap, particle2); }, c2, bc);
    return overlapBetween(I, pos1, aux1, c1, pos2, aux2, c2, bc);
}

// G lanes share a cell (G = 32: a warp per cell; sparse grids — reference cells of one sphere diameter hold less than
// one centre on average — take fewer lanes per cell so that the lanes are not idle: recountGroup below)
template <bool ENERGY, int G>
__global__ void __launch_bounds__(256) k_recount_grid(GridView g, const Inter *__restrict__ interPtr, int earlyExit,
                                                      unsigned long long *__restrict__ count, int *__restrict__ found,
                                                      double *__restrict__ cellEnergy,
                                                      const double *__restrict__ orient, size_t cap)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();

    const long long gtid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    int cell = static_cast<int>(gtid / G);
    const int lane = static_cast<int>(gtid % G);
    const bool valid = cell < g.numCells;       // (no early return: the group reductions below are warp-wide)
    if (!valid) cell = 0;
    const bool skip = !valid || (!ENERGY && earlyExit && *reinterpret_cast<volatile int *>(found));

    int cx = cell % g.n[0], cy = (cell / g.n[0]) % g.n[1], cz = cell / (g.n[0] * g.n[1]);
    int start = g.cellStart[cell], end = skip ? start : g.cellStart[cell + 1];
    unsigned long long overlaps = 0;
    double energy = 0;
    int kk = I.kk;
    bool multi = I.k != 0;

    for (int i = start; i < end; i++) {
        V3 pos1 = v3(g.px[i], g.py[i], g.pz[i]);
        V3 aux1 = v3(g.ax[i], g.ay[i], g.az[i]);
        int idx1 = g.items[i];
        int particle1 = idx1 / kk, centre1 = idx1 % kk;

        // pairs within the cell (no translation)
        BcFixed none{v3(0, 0, 0)};
        for (int j = i + 1 + lane; j < end; j += G) {
            int idx2 = g.items[j];
            int particle2 = idx2 / kk, centre2 = idx2 % kk;
            if (multi && particle1 == particle2) continue;
            V3 pos2 = v3(g.px[j], g.py[j], g.pz[j]);
            if (ENERGY) {
                energy += energyBetween(I, pos1, pos2, none);
            } else {
                V3 aux2 = v3(g.ax[j], g.ay[j], g.az[j]);
                if (overlapAny(I, orient, cap, pos1, aux1, nullptr, particle1, centre1, pos2, aux2, particle2,
                               centre2, none))
                    overlaps++;
            }
        }
        // pairs with the 13 "positive" neighbour cells
        for (int code = 14; code < 27; code++) {
            int dx = code / 9 - 1, dy = (code / 3) % 3 - 1, dz = code % 3 - 1;
            int tIdx;
            int cell2 = neighbourCell(g.n, cx, cy, cz, dx, dy, dz, &tIdx);
            BcFixed bc{g.trans[tIdx]};
            int s2 = g.cellStart[cell2], e2 = g.cellStart[cell2 + 1];
            for (int j = s2 + lane; j < e2; j += G) {
                int idx2 = g.items[j];
                int particle2 = idx2 / kk, centre2 = idx2 % kk;
                if (multi && particle1 == particle2) continue;
                V3 pos2 = v3(g.px[j], g.py[j], g.pz[j]);
                if (ENERGY) {
                    energy += energyBetween(I, pos1, pos2, bc);
                } else {
                    V3 aux2 = v3(g.ax[j], g.ay[j], g.az[j]);
                    if (overlapAny(I, orient, cap, pos1, aux1, nullptr, particle1, centre1, pos2, aux2, particle2,
                                   centre2, bc))
                        overlaps++;
                }
            }
        }
        // (per lane: the other lanes of the group finish their current centre's candidates and stop at the flag)
        if (!ENERGY && earlyExit && overlaps > 0) break;
    }

    if (ENERGY) {
        #pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) energy += __shfl_xor_sync(0xffffffffu, energy, o);
        if (lane == 0 && valid) cellEnergy[cell] = energy;
    } else {
        #pragma unroll
        for (int o = G / 2; o > 0; o >>= 1) overlaps += __shfl_xor_sync(0xffffffffu, overlaps, o);
        if (lane == 0 && overlaps > 0) {
            atomicAdd(count, overlaps);
            if (earlyExit) atomicExch(found, 1);
        }
    }
}

// lanes per cell of k_recount_grid from the average cell population
static int recountGroup(const GridStore &gs) {
    const double avg = static_cast<double>(gs.numItems) / std::max(1, gs.numCells);
    return avg < 1.5 ? 1 : (avg < 6 ? 4 : (avg < 14 ? 8 : 32));
}

template <bool ENERGY>
static void launchRecountGrid(Handle *h, const GridView &g, int earlyExit, unsigned long long *count, int *found,
                              double *cellEnergy)
{
    const int G = recountGroup(h->grid);
    const long long threads = static_cast<long long>(g.numCells) * G;
    const int blocks = static_cast<int>((threads + 255) / 256);
    cudaStream_t st = h->stream;
    switch (G) {
        case 1: k_recount_grid<ENERGY, 1><<<blocks, 256, 0, st>>>(g, h->dInter, earlyExit, count, found, cellEnergy, h->orient, h->cap); break;
        case 4: k_recount_grid<ENERGY, 4><<<blocks, 256, 0, st>>>(g, h->dInter, earlyExit, count, found, cellEnergy, h->orient, h->cap); break;
        case 8: k_recount_grid<ENERGY, 8><<<blocks, 256, 0, st>>>(g, h->dInter, earlyExit, count, found, cellEnergy, h->orient, h->cap); break;
        default: k_recount_grid<ENERGY, 32><<<blocks, 256, 0, st>>>(g, h->dInter, earlyExit, count, found, cellEnergy, h->orient, h->cap); break;
    }
}

// deterministic sum of per-cell energies: fixed strided partials + fixed tree; block b sums the contiguous chunk
// [b * chunk, (b + 1) * chunk) (one block: everything)
__global__ void k_sum_doubles(int n, int chunk, const double *__restrict__ in, double *__restrict__ out) {
    __shared__ double part[1024];
    const int begin = blockIdx.x * chunk, end = min(n, begin + chunk);
    double s = 0;
    for (int i = begin + threadIdx.x; i < end; i += blockDim.x) s += in[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = part[0];
}

// in[0..n) -> in[n] (needs room for n + 1 + SUM_BLOCKS doubles): two levels when n is large, fixed order either way
constexpr int SUM_BLOCKS = 296;
static void sumDoubles(Handle *h, int n, double *in) {
    cudaStream_t st = h->stream;
    if (n > 64 * 1024) {
        const int chunk = (n + SUM_BLOCKS - 1) / SUM_BLOCKS;
        k_sum_doubles<<<SUM_BLOCKS, 1024, 0, st>>>(n, chunk, in, in + n + 1);
        k_sum_doubles<<<1, 1024, 0, st>>>(SUM_BLOCKS, SUM_BLOCKS, in + n + 1, in + n);
        h->launches += 2;
    } else {
        k_sum_doubles<<<1, 1024, 0, st>>>(n, n, in, in + n);
        h->launches++;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// full recount without the neighbour grid (core/Packing.cpp:545-556, 582-617, 703-738): all particle pairs i < j
// with the true PeriodicBoundaryConditions
// ---------------------------------------------------------------------------------------------------------------
struct ParticleView {
    int n, k, kk;
    size_t cap, ccap;
    const double *pos, *orient, *cabs;
};

__device__ __forceinline__ V3 loadPos(const ParticleView &P, int i) {
    return v3(P.pos[i], P.pos[P.cap + i], P.pos[2 * P.cap + i]);
}
__device__ __forceinline__ M33 loadOrient(const ParticleView &P, int i) {
    M33 o;
    for (int q = 0; q < 9; q++) o.a[q] = P.orient[q * P.cap + i];
    return o;
}
__device__ __forceinline__ V3 loadCabs(const ParticleView &P, int idx) {
    return v3(P.cabs[idx], P.cabs[P.ccap + idx], P.cabs[2 * P.ccap + idx]);
}

template <bool ENERGY>
__global__ void __launch_bounds__(256) k_recount_nogrid(ParticleView P, Box box, const Inter *__restrict__ interPtr,
                                                        unsigned long long *__restrict__ count,
                                                        double *__restrict__ blockEnergy)
{
    __shared__ Inter I;
    __shared__ Box sbox;
    __shared__ double red[256];
    __shared__ unsigned long long redc[256];
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    if (threadIdx.x == 0) sbox = box;
    __syncthreads();
    BcPeriodic bc{&sbox};

    long long nPairs = static_cast<long long>(P.n) * (P.n - 1) / 2;
    unsigned long long overlaps = 0;
    double energy = 0;
    for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < nPairs;
         p += static_cast<long long>(gridDim.x) * blockDim.x)
    {
        // p -> (i, j), i < j, row-major over the strict upper triangle
        long long i = static_cast<long long>(P.n - 2
                      - floor(sqrt(static_cast<double>(-8 * p + 4ll * P.n * (P.n - 1) - 7)) / 2.0 - 0.5));
        long long j = p + i + 1 - static_cast<long long>(P.n) * (P.n - 1) / 2
                      + static_cast<long long>(P.n - i) * ((P.n - i) - 1) / 2;
        // guard against sqrt rounding
        while (i > 0 && p < i * (2ll * P.n - i - 1) / 2) { i--; }
        while (p >= (i + 1) * (2ll * P.n - i - 2) / 2) { i++; }
        j = p - i * (2ll * P.n - i - 1) / 2 + i + 1;
        int pi = static_cast<int>(i), pj = static_cast<int>(j);
        if (P.k == 0) {
            V3 pos1 = loadPos(P, pi), pos2 = loadPos(P, pj);
            if (ENERGY) {
                energy += energyBetween(I, pos1, pos2, bc);
            } else {
                M33 o1 = loadOrient(P, pi);
                V3 aux1 = pairAux(I, o1, 0), aux2 = pairAux(I, loadOrient(P, pj), 0);
                if (overlapAny(I, P.orient, P.cap, pos1, aux1, &o1, pi, 0, pos2, aux2, pj, 0, bc)) overlaps++;
            }
        } else {
            M33 o1, o2;
            if (!ENERGY) { o1 = loadOrient(P, pi); o2 = loadOrient(P, pj); }
            for (int c1 = 0; c1 < P.k; c1++) {
                V3 pos1 = loadCabs(P, pi * P.k + c1);
                for (int c2 = 0; c2 < P.k; c2++) {
                    V3 pos2 = loadCabs(P, pj * P.k + c2);
                    if (ENERGY) energy += energyBetween(I, pos1, pos2, bc);
                    else if (overlapAny(I, P.orient, P.cap, pos1, pairAux(I, o1, c1), &o1, pi, c1, pos2,
                                        pairAux(I, o2, c2), pj, c2, bc))
                        overlaps++;
                }
            }
        }
    }
    red[threadIdx.x] = energy;
    redc[threadIdx.x] = overlaps;
    __syncthreads();
    for (int o = blockDim.x / 2; o > 0; o >>= 1) {
        if (threadIdx.x < o) { red[threadIdx.x] += red[threadIdx.x + o]; redc[threadIdx.x] += redc[threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        if (ENERGY) blockEnergy[blockIdx.x] = red[0];
        else if (redc[0]) atomicAdd(count, redc[0]);
    }
}

static ParticleView particleView(const Handle *h) {
    ParticleView P;
    P.n = static_cast<int>(h->n);
    P.k = h->inter.k;
    P.kk = h->inter.kk;
    P.cap = h->cap;
    P.ccap = h->centreCap;
    P.pos = h->pos;
    P.orient = h->orient;
    P.cabs = h->cabs;
    return P;
}

static void ensureCellEnergy(Handle *h, size_t need) {
    if (need > h->cellEnergyCap || !h->dCellEnergy) {
        if (h->dCellEnergy) cudaFree(h->dCellEnergy);
        h->dCellEnergy = nullptr;
        RP_CUDA(cudaMalloc(&h->dCellEnergy, (need + 16) * sizeof(double)));
        h->cellEnergyCap = need + 16;
    }
}

// Packing::countWallOverlaps (core/Packing.cpp:561-580): one thread per particle, block sum, one atomic per block
__global__ void k_count_wall_overlaps(int n, ParticleView P, const double *__restrict__ crelAll,
                                      const Inter *__restrict__ interPtr, Walls W,
                                      unsigned long long *__restrict__ counter)
{
    __shared__ Inter I;
    __shared__ unsigned int blockSum;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    if (threadIdx.x == 0) blockSum = 0;
    __syncthreads();
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned int c = 0;
    if (p < n) {
        V3 pos = loadPos(P, p);
        M33 o = loadOrient(P, p);
        c = particleWallOverlaps(I, W, pos, o, [&](int centre) {
            size_t idx = static_cast<size_t>(p) * I.k + centre;
            return v3(crelAll[idx], crelAll[P.ccap + idx], crelAll[2 * P.ccap + idx]);
        });
    }
    if (c) atomicAdd(&blockSum, c);
    __syncthreads();
    if (threadIdx.x == 0 && blockSum) atomicAdd(counter, static_cast<unsigned long long>(blockSum));
}

unsigned long long countTotalOverlaps(Handle *h, bool earlyExit) {
    if (!h->inter.hard) return 0;
    ensureGrid(h);
    cudaStream_t st = h->stream;
    zeroAsync(h, h->dCounters, sizeof(unsigned long long), st);
    zeroAsync(h, h->dFlags + 2, sizeof(int), st);
    const Inter &I = h->inter;
    const bool culled = h->hasGrid && I.k == 0 && (I.hard == RP_HARD_SPHERE || I.hard == RP_HARD_SPHEROCYLINDER)
                        && !std::getenv("RP_RECOUNT_LEGACY");
    if (culled) {
        // two-stage pair test (FP32 cull -> compacted exact FP64 tests), rp_recount_cull.cuh
        GridView g = makeView(h->grid);
        RecountArgs A;
        A.inter.sphereThr2 = I.sphereThr2;
        A.inter.scInner2 = I.scInner2;
        A.inter.scOuter2 = I.scOuter2;
        A.inter.scHalf = I.scHalfPlus;
        A.earlyExit = earlyExit ? 1 : 0;
        cullThresholds(h, g.relCell, &A.cullThr2, &A.cullHalf);
        A.cull2Thr2 = std::nextafterf(static_cast<float>(I.scInner2 + 1e-3 * I.range * I.range), INFINITY);
        A.cellBegin = 0;
        A.cellEnd = g.numCells;
        if (shardActive(h)) {
            // every pair is enumerated from exactly one cell (half-shell rule), so a split of the cells is a split
            // of the pairs; the per-rank counts are summed on the device (ncclAllReduce below)
            const long long P = h->shard->nranks, r = h->shard->rank;
            A.cellBegin = static_cast<int>(r * g.numCells / P);
            A.cellEnd = static_cast<int>((r + 1) * g.numCells / P);
        }
        int blocks = std::max(1, (A.cellEnd - A.cellBegin + RC_WARPS - 1) / RC_WARPS);
        if (I.hard == RP_HARD_SPHERE)
            k_recount_cull<RP_HARD_SPHERE><<<blocks, RC_WARPS * 32, 0, st>>>(g, h->grid.box, A, h->dCounters,
                                                                             h->dFlags + 2);
        else
            k_recount_cull<RP_HARD_SPHEROCYLINDER><<<blocks, RC_WARPS * 32, 0, st>>>(g, h->grid.box, A, h->dCounters,
                                                                                     h->dFlags + 2);
    } else if (h->hasGrid) {
        GridView g = makeView(h->grid);
        launchRecountGrid<false>(h, g, earlyExit ? 1 : 0, h->dCounters, h->dFlags + 2, nullptr);
    } else {
        long long nPairs = static_cast<long long>(h->n) * (h->n - 1) / 2;
        int blocks = static_cast<int>(std::min<long long>(1184, (nPairs + 255) / 256));
        if (blocks > 0)
            k_recount_nogrid<false><<<blocks, 256, 0, st>>>(particleView(h), h->box, h->dInter, h->dCounters, nullptr);
    }
    h->launches++;
    if (culled && shardActive(h)) shardAllReduceU64(h, h->dCounters, 1);
    if (h->wallMask) {       // + Packing::countWallOverlaps (every rank of a sharded run counts all walls itself)
        const int n = static_cast<int>(h->n);
        k_count_wall_overlaps<<<(n + 255) / 256, 256, 0, st>>>(n, particleView(h), h->crel, h->dInter,
                                                               makeWalls(h->box, h->wallMask, I.totalRange),
                                                               h->dCounters);
        h->launches++;
    }
    unsigned long long result = 0;
    readbackSync(h, st, &result, h->dCounters, sizeof(result));
    RP_CUDA(cudaGetLastError());
    if (earlyExit && result) return 1;   // Packing.cpp:543-544
    return result;
}

double totalEnergy(Handle *h) {
    if (!h->inter.soft) return 0;
    ensureGrid(h);
    cudaStream_t st = h->stream;
    int parts = 0;
    if (h->hasGrid) {
        GridView g = makeView(h->grid);
        ensureCellEnergy(h, g.numCells + 2 + SUM_BLOCKS);
        launchRecountGrid<true>(h, g, 0, nullptr, nullptr, h->dCellEnergy);
        parts = g.numCells;
    } else {
        long long nPairs = static_cast<long long>(h->n) * (h->n - 1) / 2;
        int blocks = static_cast<int>(std::min<long long>(1184, (nPairs + 255) / 256));
        if (blocks == 0) return 0;
        ensureCellEnergy(h, blocks + 2 + SUM_BLOCKS);
        k_recount_nogrid<true><<<blocks, 256, 0, st>>>(particleView(h), h->box, h->dInter, nullptr, h->dCellEnergy);
        parts = blocks;
    }
    sumDoubles(h, parts, h->dCellEnergy);
    h->launches++;
    double result = 0;
    readbackSync(h, st, &result, h->dCellEnergy + parts, sizeof(double));
    RP_CUDA(cudaGetLastError());
    return result;
}

// ---------------------------------------------------------------------------------------------------------------
// explicit trial move: one CTA.  tmpState layout (doubles): pos[3] orient[9] crel[16][3] cabs[16][3]
// ---------------------------------------------------------------------------------------------------------------
constexpr int TMP_POS = 0, TMP_ORIENT = 3, TMP_CREL = 12, TMP_CABS = 12 + 3 * RP_MAX_CENTRES;
constexpr int TMP_SIZE = 12 + 6 * RP_MAX_CENTRES;

struct TrialState {
    V3 pos;
    M33 orient;
    V3 cabs[RP_MAX_CENTRES];
};

// Overlap count and/or energy of `S` (standing for particle `orig`) against the rest of the packing, block-parallel:
// Packing::countParticleOverlaps / calculateParticleEnergy (core/Packing.cpp:373-429, 649-688)
__device__ void scanParticle(const TrialState &S, int orig, const ParticleView &P, const GridView &g, bool hasGrid,
                             const Box &box, const Inter &I, bool wantOverlaps, bool wantEnergy,
                             unsigned long long &overlaps, double &energy)
{
    overlaps = 0;
    energy = 0;
    int kk = I.kk;
    if (hasGrid) {
        for (int c1 = 0; c1 < kk; c1++) {
            V3 pos1 = (I.k == 0) ? S.pos : S.cabs[c1];
            V3 aux1 = pairAux(I, S.orient, c1);
            int cell = cellOfPosition(g, box, pos1, nullptr);
            int cx = cell % g.n[0], cy = (cell / g.n[0]) % g.n[1], cz = cell / (g.n[0] * g.n[1]);
            for (int code = 0; code < 27; code++) {
                int dx = code / 9 - 1, dy = (code / 3) % 3 - 1, dz = code % 3 - 1;
                int tIdx;
                int cell2 = neighbourCell(g.n, cx, cy, cz, dx, dy, dz, &tIdx);
                BcFixed bc{g.trans[tIdx]};
                int s2 = g.cellStart[cell2], e2 = g.cellStart[cell2 + 1];
                for (int j = s2 + threadIdx.x; j < e2; j += blockDim.x) {
                    int idx2 = g.items[j];
                    int particle2 = idx2 / kk, centre2 = idx2 % kk;
                    if (particle2 == orig) continue;
                    V3 pos2 = v3(g.px[j], g.py[j], g.pz[j]);
                    if (wantOverlaps) {
                        V3 aux2 = v3(g.ax[j], g.ay[j], g.az[j]);
                        if (overlapAny(I, P.orient, P.cap, pos1, aux1, &S.orient, orig, c1, pos2, aux2, particle2,
                                       centre2, bc))
                            overlaps++;
                    }
                    if (wantEnergy) energy += energyBetween(I, pos1, pos2, bc);
                }
            }
        }
    } else {
        BcPeriodic bc{&box};
        for (int j = threadIdx.x; j < P.n; j += blockDim.x) {
            if (j == orig) continue;
            if (I.k == 0) {
                V3 pos2 = loadPos(P, j);
                if (wantOverlaps) {
                    V3 aux2 = pairAux(I, loadOrient(P, j), 0);
                    if (overlapAny(I, P.orient, P.cap, S.pos, pairAux(I, S.orient, 0), &S.orient, orig, 0, pos2, aux2,
                                   j, 0, bc))
                        overlaps++;
                }
                if (wantEnergy) energy += energyBetween(I, S.pos, pos2, bc);
            } else {
                M33 o2;
                if (wantOverlaps) o2 = loadOrient(P, j);
                for (int c1 = 0; c1 < I.k; c1++) {
                    for (int c2 = 0; c2 < I.k; c2++) {
                        V3 pos2 = loadCabs(P, j * I.k + c2);
                        if (wantOverlaps && overlapAny(I, P.orient, P.cap, S.cabs[c1], pairAux(I, S.orient, c1),
                                                       &S.orient, orig, c1, pos2, pairAux(I, o2, c2), j, c2, bc))
                            overlaps++;
                        if (wantEnergy) energy += energyBetween(I, S.cabs[c1], pos2, bc);
                    }
                }
            }
        }
    }
}

// out[0] = overlaps(original), out[1] = overlaps(trial), out[2] = energy(original), out[3] = energy(trial)
// The original state is only scanned when `counting` (overlaps) or the interaction is soft (energy).
__global__ void __launch_bounds__(256) k_try_move(ParticleView P, const double *__restrict__ crelAll, GridView g,
                                                  int hasGrid, Box box, const Inter *__restrict__ interPtr,
                                                  int particle, int moveKind, V3 translation, M33 rotation,
                                                  int counting, Walls W, double *__restrict__ tmp,
                                                  double *__restrict__ out)
{
    __shared__ Inter I;
    __shared__ Box sbox;
    __shared__ TrialState orig, trial;
    __shared__ double red[256];
    __shared__ unsigned long long redc[256];
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    if (threadIdx.x == 0) sbox = box;
    __syncthreads();
    if (threadIdx.x == 0) {
        orig.pos = loadPos(P, particle);
        orig.orient = loadOrient(P, particle);
        trial.pos = orig.pos;
        trial.orient = orig.orient;
        // Shape::translate (core/Shape.cpp:9-12): position += translation; position += bc.getCorrection(position)
        if (moveKind == RP_MOVE_TRANSLATION || moveKind == RP_MOVE_ROTOTRANSLATION) {
            trial.pos = trial.pos + translation;
            trial.pos = trial.pos + pbcCorrection(sbox, trial.pos);
        }
        // Shape::rotate (core/Shape.cpp:14-16): orientation = rotation * orientation
        if (moveKind == RP_MOVE_ROTATION || moveKind == RP_MOVE_ROTOTRANSLATION)
            trial.orient = mul(rotation, trial.orient);
        tmp[TMP_POS] = trial.pos.x; tmp[TMP_POS + 1] = trial.pos.y; tmp[TMP_POS + 2] = trial.pos.z;
        for (int q = 0; q < 9; q++) tmp[TMP_ORIENT + q] = trial.orient.a[q];
        // interaction centres: prepareTemp / rotateTemp / recalculateAbsolute (Packing.cpp:339-361, 1041-1047)
        for (int c = 0; c < I.k; c++) {
            int idx = particle * I.k + c;
            orig.cabs[c] = loadCabs(P, idx);
            V3 rel = v3(crelAll[idx], crelAll[P.ccap + idx], crelAll[2 * P.ccap + idx]);
            if (moveKind == RP_MOVE_ROTATION || moveKind == RP_MOVE_ROTOTRANSLATION)
                rel = mul(rotation, rel);
            V3 p = trial.pos + rel;
            trial.cabs[c] = p + pbcCorrection(sbox, p);
            tmp[TMP_CREL + 3 * c] = rel.x; tmp[TMP_CREL + 3 * c + 1] = rel.y; tmp[TMP_CREL + 3 * c + 2] = rel.z;
            tmp[TMP_CABS + 3 * c] = trial.cabs[c].x; tmp[TMP_CABS + 3 * c + 1] = trial.cabs[c].y;
            tmp[TMP_CABS + 3 * c + 2] = trial.cabs[c].z;
        }
    }
    __syncthreads();

    bool hard = I.hard != RP_HARD_NONE, soft = I.soft != RP_SOFT_NONE;
    unsigned long long ovO = 0, ovT = 0;
    double enO = 0, enT = 0;
    if ((hard && counting) || soft)
        scanParticle(orig, particle, P, g, hasGrid, sbox, I, hard && counting, soft, ovO, enO);
    if (hard || soft)
        scanParticle(trial, particle, P, g, hasGrid, sbox, I, hard, soft, ovT, enT);

    if (threadIdx.x == 0 && W.mask && hard) {
        // + countParticleWallOverlaps of the scanned states (core/Packing.cpp:424-428)
        if (counting)
            ovO += particleWallOverlaps(I, W, orig.pos, orig.orient, [&](int c) {
                int idx = particle * I.k + c;
                return v3(crelAll[idx], crelAll[P.ccap + idx], crelAll[2 * P.ccap + idx]);
            });
        ovT += particleWallOverlaps(I, W, trial.pos, trial.orient, [&](int c) {
            return v3(tmp[TMP_CREL + 3 * c], tmp[TMP_CREL + 3 * c + 1], tmp[TMP_CREL + 3 * c + 2]);
        });
    }
    double vals[4] = {static_cast<double>(ovO), static_cast<double>(ovT), enO, enT};
    for (int q = 0; q < 4; q++) {
        red[threadIdx.x] = vals[q];
        __syncthreads();
        for (int o = blockDim.x / 2; o > 0; o >>= 1) {
            if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) out[q] = red[0];
        __syncthreads();
    }
    (void) redc;
}

__global__ void k_accept_move(ParticleView P, double *pos, double *orient, double *crel, double *cabs, int particle,
                              int k, const double *__restrict__ tmp)
{
    int t = threadIdx.x;
    if (t < 3) pos[t * P.cap + particle] = tmp[TMP_POS + t];
    if (t < 9) orient[t * P.cap + particle] = tmp[TMP_ORIENT + t];
    if (t < k) {
        int idx = particle * k + t;
        for (int d = 0; d < 3; d++) {
            crel[d * P.ccap + idx] = tmp[TMP_CREL + 3 * t + d];
            cabs[d * P.ccap + idx] = tmp[TMP_CABS + 3 * t + d];
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// bulk per-particle scan of the CURRENT state: overlaps / energy of every particle with the rest of the packing —
// Packing::countParticleOverlaps(i, i, ...) / calculateParticleEnergy(i, i, ...) for all i (core/Packing.cpp:373-429,
// 649-688).  Persistent CTAs of 64 threads walk the particles; used by getParticleEnergyFluctuations and by the
// per-particle rejection of renormalizeOrientations (both rare: every 10^3..10^4 cycles).
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(64) k_particle_scan(ParticleView P, GridView g, int hasGrid, Box box,
                                                      const Inter *__restrict__ interPtr, int wantOverlaps,
                                                      int wantEnergy, unsigned char *__restrict__ overlapFlag,
                                                      double *__restrict__ energyOut)
{
    __shared__ Inter I;
    __shared__ Box sbox;
    __shared__ TrialState S;
    __shared__ double red[64];
    __shared__ unsigned long long redc[64];
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    if (threadIdx.x == 0) sbox = box;
    __syncthreads();
    for (int particle = blockIdx.x; particle < P.n; particle += gridDim.x) {
        if (threadIdx.x == 0) {
            S.pos = loadPos(P, particle);
            S.orient = loadOrient(P, particle);
            for (int c = 0; c < I.k; c++) S.cabs[c] = loadCabs(P, particle * I.k + c);
        }
        __syncthreads();
        unsigned long long ov = 0;
        double en = 0;
        scanParticle(S, particle, P, g, hasGrid, sbox, I, wantOverlaps != 0, wantEnergy != 0, ov, en);
        red[threadIdx.x] = en;
        redc[threadIdx.x] = ov;
        __syncthreads();
        for (int o = 32; o > 0; o >>= 1) {
            if (threadIdx.x < o) { red[threadIdx.x] += red[threadIdx.x + o]; redc[threadIdx.x] += redc[threadIdx.x + o]; }
            __syncthreads();
        }
        if (threadIdx.x == 0) {
            if (overlapFlag) overlapFlag[particle] = redc[0] ? 1 : 0;
            if (energyOut) energyOut[particle] = red[0];
        }
        __syncthreads();
    }
}

void particleScan(Handle *h, bool wantOverlaps, bool wantEnergy, unsigned char *dOverlapFlag, double *dEnergy) {
    ensureGrid(h);
    GridView g;
    if (h->hasGrid) g = makeView(h->grid);
    else std::memset(&g, 0, sizeof(g));
    const int blocks = static_cast<int>(std::min<size_t>(h->n, 148 * 16));
    k_particle_scan<<<blocks, 64, 0, h->stream>>>(particleView(h), g, h->hasGrid ? 1 : 0, h->box, h->dInter,
                                                  wantOverlaps ? 1 : 0, wantEnergy ? 1 : 0, dOverlapFlag, dEnergy);
    h->launches++;
    RP_CUDA(cudaGetLastError());
}

// Packing::getParticleEnergyFluctuations (core/Packing.cpp:841-856): the per-particle energies come from the device,
// the two running sums are accumulated on the host in the reference's order
double particleEnergyFluctuations(Handle *h) {
    if (!h->inter.soft) return 0;
    const size_t n = h->n;
    size_t need = n * sizeof(double);
    if (need > h->dStageCap) {
        if (h->dStage) cudaFree(h->dStage);
        h->dStage = nullptr;
        RP_CUDA(cudaMalloc(&h->dStage, need + 64));
        h->dStageCap = need + 64;
    }
    double *dEnergy = static_cast<double *>(h->dStage);
    particleScan(h, false, true, nullptr, dEnergy);
    std::vector<double> e(n);
    RP_CUDA(cudaMemcpyAsync(e.data(), dEnergy, need, cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    double energySum = 0, energySum2 = 0;
    for (size_t i = 0; i < n; i++) {
        energySum += e[i];
        energySum2 += e[i] * e[i];
    }
    double N = static_cast<double>(n);
    double doubleEnergy = std::sqrt(energySum2 / (N - 1) - std::pow(energySum, 2) / N / (N - 1));
    return doubleEnergy / 2;
}

// Packing::countWallOverlaps (core/Packing.cpp:561-580)
unsigned long long countWallOverlaps(Handle *h, bool earlyExit) {
    if (!h->wallMask || !h->inter.hard) return 0;
    cudaStream_t st = h->stream;
    zeroAsync(h, h->dCounters, sizeof(unsigned long long), st);
    const int n = static_cast<int>(h->n);
    k_count_wall_overlaps<<<(n + 255) / 256, 256, 0, st>>>(n, particleView(h), h->crel, h->dInter,
                                                           makeWalls(h->box, h->wallMask, h->inter.totalRange),
                                                           h->dCounters);
    h->launches++;
    unsigned long long result = 0;
    readbackSync(h, st, &result, h->dCounters, sizeof(result));
    RP_CUDA(cudaGetLastError());
    // with early exit the reference returns the overlaps of the particles visited before the flag was seen: at
    // least 1, at most the full count — callers only test for > 0
    return (earlyExit && result) ? 1 : result;
}

// Packing::tryTranslation / tryRotation / tryMove (core/Packing.cpp:84-164) incl. calculateMoveOverlapEnergy
// (Packing.cpp:304-323)
double tryMove(Handle *h, const rp_move &m) {
    RP_EXPECT(m.particle < h->n, "particleIdx < size()");
    RP_EXPECT(h->hasInteraction, "interaction not set");
    ensureGrid(h);
    cudaStream_t st = h->stream;
    V3 tr = v3(m.translation[0], m.translation[1], m.translation[2]);
    M33 rot;
    for (int q = 0; q < 9; q++) rot.a[q] = m.rotation[q];
    GridView g;
    if (h->hasGrid) g = makeView(h->grid);
    else std::memset(&g, 0, sizeof(g));
    k_try_move<<<1, 256, 0, st>>>(particleView(h), h->crel, g, h->hasGrid ? 1 : 0, h->box, h->dInter,
                                  static_cast<int>(m.particle), m.kind, tr, rot, h->overlapCounting ? 1 : 0,
                                  makeWalls(h->box, h->wallMask, h->inter.totalRange), h->tmpState, h->dMoveOut);
    h->launches++;
    double out[4];
    readbackSync(h, st, out, h->dMoveOut, sizeof(out));
    RP_CUDA(cudaGetLastError());
    h->hasTrial = true;
    h->trialParticle = m.particle;
    h->trialKind = m.kind;
    const double INF = std::numeric_limits<double>::infinity();
    if (h->inter.hard != RP_HARD_NONE) {
        if (h->overlapCounting) {
            h->lastMoveOverlapDelta = static_cast<long long>(out[1]) - static_cast<long long>(out[0]);
            if (h->lastMoveOverlapDelta < 0) return -INF;
            if (h->lastMoveOverlapDelta > 0) return INF;
        } else if (out[1] > 0) {
            return INF;
        }
    }
    return out[3] - out[2];
}

// Packing::acceptMove / acceptTranslation / acceptRotation (core/Packing.cpp:234-302): temp slot -> particle, grid
// remove + add (here: the grid is rebuilt lazily before its next use)
void acceptMove(Handle *h) {
    RP_EXPECT(h->hasTrial, "acceptMove without a preceding tryMove");
    k_accept_move<<<1, 32, 0, h->stream>>>(particleView(h), h->pos, h->orient, h->crel, h->cabs,
                                           static_cast<int>(h->trialParticle), h->inter.k, h->tmpState);
    h->launches++;
    RP_CUDA(cudaGetLastError());
    h->gridDirty = true;
    if (h->overlapCounting)
        h->numOverlaps = static_cast<unsigned long long>(static_cast<long long>(h->numOverlaps)
                                                         + h->lastMoveOverlapDelta);
}

// ---------------------------------------------------------------------------------------------------------------
// box move: Packing::tryScaling / revertScaling (core/Packing.cpp:172-208, 363-371)
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_rescale(int n, size_t cap, Box oldBox, Box newBox, double *__restrict__ pos) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 p = v3(pos[i], pos[cap + i], pos[2 * cap + i]);
    // shape.setPosition(box.relativeToAbsolute(lastBox.absoluteToRelative(position)))  (Packing.cpp:183-184)
    V3 q = relToAbs(newBox, absToRel(oldBox, p));
    pos[i] = q.x; pos[cap + i] = q.y; pos[2 * cap + i] = q.z;
}

// Packing::isBoxUpscaled (core/Packing.cpp:1155-1176)
static bool isBoxUpscaled(const Box &oldBox, const Box &newBox) {
    const double EPSILON = 100 * std::numeric_limits<double>::epsilon();
    for (int i = 0; i < 3; i++) {
        V3 e = v3(i == 0, i == 1, i == 2);
        V3 oldSide = relToAbs(oldBox, e), newSide = relToAbs(newBox, e);
        double d = dot(oldSide, newSide);
        double norm2Old = norm2(oldSide), norm2New = norm2(newSide);
        if (std::abs(d * d / norm2Old / norm2New - 1) > EPSILON) return false;
        if (d * d / norm2Old / norm2Old < 1 + EPSILON) return false;
    }
    return true;
}

double tryScaling(Handle *h, const Box &newBox) {
    RP_EXPECT(h->hasInteraction && h->n > 0, "packing is empty");
    RP_EXPECT(boxVolume(newBox) != 0, "newBox.getVolume() != 0");
    cudaStream_t st = h->stream;
    RP_CUDA(cudaEventRecord(h->ev[0], st));
    h->lastBox = h->box;
    devEnsure(h->lastPos, h->lastCap, 3 * h->cap);
    copyAsync(h, h->lastPos, h->pos, 3 * h->cap * sizeof(double), st);
    if (h->inter.k > 0) {
        devEnsure(h->lastCabs, h->lastCentreCap, 3 * h->centreCap);
        copyAsync(h, h->lastCabs, h->cabs, 3 * h->centreCap * sizeof(double), st);
    }
    // The grid of the OLD state is only worth building if somebody reads it: the energy of a soft interaction now, or
    // whoever comes after a revert.  After a sweep it is stale and the next sweep would discard it unread, so a hard-
    // only run leaves it stale (revertScaling then marks the restored state "grid dirty": built lazily on first use).
    double initialEnergy = totalEnergy(h);     // soft interactions: builds the current grid
    h->lastGridValid = !h->gridDirty;
    h->lastScalingNumOverlaps = h->numOverlaps;

    h->box = newBox;
    int n = static_cast<int>(h->n);
    k_rescale<<<(n + 255) / 256, 256, 0, st>>>(n, h->cap, h->lastBox, h->box, h->pos);
    h->launches++;
    std::swap(h->grid, h->tempGrid);
    std::swap(h->hasGrid, h->tempHasGrid);
    recalcAbsoluteCentres(h);
    rebuildReferenceGrid(h);
    h->hasLastScaling = true;

    const double INF = std::numeric_limits<double>::infinity();
    double result = 0;
    bool decided = false;
    if (h->inter.hard != RP_HARD_NONE) {
        if (h->overlapCounting) {
            h->numOverlaps = countTotalOverlaps(h, false);
            long long delta = static_cast<long long>(h->numOverlaps)
                              - static_cast<long long>(h->lastScalingNumOverlaps);
            if (delta < 0) { result = -INF; decided = true; }
            else if (delta > 0) { result = INF; decided = true; }
        } else {
            bool cannotOverlap = h->inter.convex && isBoxUpscaled(h->lastBox, h->box);
            if (!cannotOverlap && countTotalOverlaps(h, true) > 0) { result = INF; decided = true; }
        }
    }
    if (!decided)
        result = totalEnergy(h) - initialEnergy;
    RP_CUDA(cudaEventRecord(h->ev[1], st));
    RP_CUDA(cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]);
    h->lastScalingMs = ms;
    return result;
}

void revertScaling(Handle *h) {
    RP_EXPECT(h->hasLastScaling, "revertScaling without a preceding tryScaling");
    cudaStream_t st = h->stream;
    copyAsync(h, h->pos, h->lastPos, 3 * h->cap * sizeof(double), st);
    if (h->inter.k > 0)
        copyAsync(h, h->cabs, h->lastCabs, 3 * h->centreCap * sizeof(double), st);
    h->box = h->lastBox;
    std::swap(h->grid, h->tempGrid);
    std::swap(h->hasGrid, h->tempHasGrid);
    h->gridDirty = !h->lastGridValid;      // the swapped-back grid describes the restored state if it was built
    h->numOverlaps = h->lastScalingNumOverlaps;
    h->hasLastScaling = false;
}

// ---------------------------------------------------------------------------------------------------------------
// parity harness: Interaction::overlapBetween / calculateEnergyBetween on explicit pairs
// ---------------------------------------------------------------------------------------------------------------
__global__ void k_pair_batch(unsigned long long n, const rp_pair *__restrict__ pairs,
                             const Inter *__restrict__ interPtr, unsigned char *__restrict__ verdict,
                             double *__restrict__ energy)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();
    for (unsigned long long i = static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<unsigned long long>(gridDim.x) * blockDim.x)
    {
        const rp_pair &p = pairs[i];
        M33 o1, o2;
        for (int q = 0; q < 9; q++) { o1.a[q] = p.orientation1[q]; o2.a[q] = p.orientation2[q]; }
        V3 pos1 = v3(p.pos1[0], p.pos1[1], p.pos1[2]), pos2 = v3(p.pos2[0], p.pos2[1], p.pos2[2]);
        BcFixed bc{v3(p.translation[0], p.translation[1], p.translation[2])};
        int c1 = static_cast<int>(p.idx1), c2 = static_cast<int>(p.idx2);
        if (verdict)
            verdict[i] = (I.hard != RP_HARD_NONE)
                         && (needsOrientation(I)
                             ? overlapBetweenOriented(I, pos1, [&] { return o1; }, c1, pos2, [&] { return o2; }, c2, bc)
                             : overlapBetween(I, pos1, pairAux(I, o1, c1), c1, pos2, pairAux(I, o2, c2), c2, bc));
        if (energy)
            energy[i] = (I.soft != RP_SOFT_NONE) ? energyBetween(I, pos1, pos2, bc) : 0.0;
    }
}

void pairTestBatch(Handle *h, const rp_pair *pairs, unsigned long long n, unsigned char *verdict, double *energy) {
    RP_EXPECT(h->hasInteraction, "interaction not set");
    if (n == 0) return;
    cudaStream_t st = h->stream;
    const unsigned long long CHUNK = 1ull << 20;
    size_t need = CHUNK * (sizeof(rp_pair) + sizeof(double) + 1) + 64;
    if (need > h->dStageCap) {
        if (h->dStage) cudaFree(h->dStage);
        h->dStage = nullptr;
        RP_CUDA(cudaMalloc(&h->dStage, need));
        h->dStageCap = need;
    }
    auto *dPairs = static_cast<rp_pair *>(h->dStage);
    auto *dEnergy = reinterpret_cast<double *>(dPairs + CHUNK);
    auto *dVerdict = reinterpret_cast<unsigned char *>(dEnergy + CHUNK);
    for (unsigned long long off = 0; off < n; off += CHUNK) {
        unsigned long long m = std::min(CHUNK, n - off);
        RP_CUDA(cudaMemcpyAsync(dPairs, pairs + off, m * sizeof(rp_pair), cudaMemcpyHostToDevice, st));
        int blocks = static_cast<int>(std::min<unsigned long long>(148 * 8, (m + 255) / 256));
        k_pair_batch<<<blocks, 256, 0, st>>>(m, dPairs, h->dInter, verdict ? dVerdict : nullptr,
                                             energy ? dEnergy : nullptr);
        h->launches++;
        if (verdict) RP_CUDA(cudaMemcpyAsync(verdict + off, dVerdict, m, cudaMemcpyDeviceToHost, st));
        if (energy) RP_CUDA(cudaMemcpyAsync(energy + off, dEnergy, m * sizeof(double), cudaMemcpyDeviceToHost, st));
        RP_CUDA(cudaStreamSynchronize(st));
    }
    RP_CUDA(cudaGetLastError());
}

}  // namespace rp
