// rp_grid.cu — device-resident NeighbourGrid rebuild (north_star kernel (a)).
//
// Replaces Packing::rebuildNeighbourGrid + addInteractionCentresToNeighbourGrid (core/Packing.cpp:858-929) and the
// linked-list NeighbourGrid::add (core/NeighbourGrid.cpp:219-236): cell index per interaction centre
// (positionToCellNo, NeighbourGrid.cpp:13-34), then a counting sort on the cell index -> cellStart[] + items[] and
// coalesced SoA pair records (centre position + aux vector) in cell order.
//
// Pipeline (all on the handle's stream, no host synchronisation except reading the max cell population):
//   k_cell_keys     : key[i] = cell of centre i, histogram[key]++            reads 24 B, writes 4 B per centre
//   k_scan_*        : exclusive scan of the histogram -> cellStart            4 B per cell
//   k_scatter       : itemsTmp[cellStart[key] + fill[key]++] = i              4 B + 4 B per centre
//   k_cell_order    : per cell, rank items by ascending index (deterministic order regardless of atomics)
//   k_gather_records: sorted SoA pair records                                 reads 24(+72) B, writes 48 B per centre
#include "rp_internal.cuh"

namespace rp {

__global__ void k_cell_keys(int numItems, int k, size_t cap, size_t ccap, const double *__restrict__ pos,
                            const double *__restrict__ cabs, Box box, GridView g, int *__restrict__ keys,
                            int *__restrict__ hist, int *__restrict__ flags)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numItems) return;
    V3 p;
    if (k == 0) p = v3(pos[i], pos[cap + i], pos[2 * cap + i]);
    else p = v3(cabs[i], cabs[ccap + i], cabs[2 * ccap + i]);
    int outside = 0;
    int c = cellOfPosition(g, box, p, &outside);
    if (outside) atomicOr(&flags[0], 1);
    keys[i] = c;
    atomicAdd(&hist[c], 1);
}

// --- exclusive scan over numCells ints: per-block reduce, scan of block sums, per-block scan + offset ---
constexpr int SCAN_BLOCK = 1024;

__device__ __forceinline__ int blockExclusiveScan(int v, int *total) {
    __shared__ int warpSums[32];
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
    for (int o = 1; o < 32; o <<= 1) {
        int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int w = (lane < (blockDim.x >> 5)) ? warpSums[lane] : 0;
        int wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        warpSums[lane] = wi - w;
        if (lane == 31) *total = wi;
    }
    __syncthreads();
    int res = incl - v + warpSums[warp];
    __syncthreads();
    return res;
}

__global__ void k_scan_block_sums(int n, const int *__restrict__ in, int *__restrict__ blockSums) {
    __shared__ int total;
    int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    int v = (i < n) ? in[i] : 0;
    blockExclusiveScan(v, &total);
    if (threadIdx.x == 0) blockSums[blockIdx.x] = total;
}

// single block: exclusive scan of the block sums in place (loops if more than SCAN_BLOCK of them)
__global__ void k_scan_sums(int numBlocks, int *__restrict__ blockSums) {
    __shared__ int total;
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < numBlocks; base += SCAN_BLOCK) {
        int i = base + threadIdx.x;
        int v = (i < numBlocks) ? blockSums[i] : 0;
        int ex = blockExclusiveScan(v, &total);
        if (i < numBlocks) blockSums[i] = ex + carry;
        __syncthreads();
        if (threadIdx.x == 0) carry += total;
        __syncthreads();
    }
}

// `in` and `fill` alias (the histogram becomes the fill counter array), hence no __restrict__ on them
__global__ void k_scan_final(int n, const int *in, const int *__restrict__ blockSums,
                             int *__restrict__ out, int *fill, int *__restrict__ maxPop)
{
    __shared__ int total;
    int i = blockIdx.x * SCAN_BLOCK + threadIdx.x;
    int v = (i < n) ? in[i] : 0;
    int ex = blockExclusiveScan(v, &total) + blockSums[blockIdx.x];
    if (i < n) {
        out[i] = ex;
        fill[i] = 0;
        if (i == n - 1) out[n] = ex + v;
    }
    // max population (block max -> one atomic)
    int m = v;
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(maxPop, m);
}

__global__ void k_scatter(int numItems, const int *__restrict__ keys, const int *__restrict__ cellStart,
                          int *__restrict__ fill, int *__restrict__ itemsTmp)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= numItems) return;
    int c = keys[i];
    int slot = cellStart[c] + atomicAdd(&fill[c], 1);
    itemsTmp[slot] = i;
}

// one warp per cell: order the cell's items by ascending index (rank by counting; cells hold tens of items)
__global__ void k_cell_order(int numCells, const int *__restrict__ cellStart, const int *__restrict__ itemsTmp,
                             int *__restrict__ items)
{
    int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (warp >= numCells) return;
    int start = cellStart[warp], end = cellStart[warp + 1];
    int n = end - start;
    for (int i = lane; i < n; i += 32) {
        int id = itemsTmp[start + i];
        int rank = 0;
        for (int j = 0; j < n; j++) rank += (itemsTmp[start + j] < id);
        items[start + rank] = id;
    }
}

__global__ void k_gather_records(int numItems, int kk, int k, size_t cap, size_t ccap, size_t icap,
                                 const int *__restrict__ items, const double *__restrict__ pos,
                                 const double *__restrict__ orient, const double *__restrict__ cabs,
                                 const Inter *__restrict__ inter, double *__restrict__ rec)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= numItems) return;
    int idx = items[s];
    int particle = idx / kk, centre = idx % kk;
    V3 p;
    if (k == 0) p = v3(pos[particle], pos[cap + particle], pos[2 * cap + particle]);
    else p = v3(cabs[idx], cabs[ccap + idx], cabs[2 * ccap + idx]);
    V3 aux = v3(0, 0, 0);
    int hard = inter->hard;
    if (hard == RP_HARD_SPHEROCYLINDER) {
        aux = v3(orient[2 * cap + particle], orient[5 * cap + particle], orient[8 * cap + particle]);
    } else if (hard == RP_HARD_POLYSPHEROCYLINDER) {
        M33 o;
        for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + particle];
        aux = pairAux(*inter, o, centre);
    }
    rec[s] = p.x; rec[icap + s] = p.y; rec[2 * icap + s] = p.z;
    rec[3 * icap + s] = aux.x; rec[4 * icap + s] = aux.y; rec[5 * icap + s] = aux.z;
}

GridView makeView(const GridStore &g) {
    GridView v;
    for (int i = 0; i < 3; i++) {
        v.n[i] = g.n[i];
        v.relCell[i] = 1 / static_cast<double>(g.n[i]);
    }
    v.numCells = g.numCells;
    v.numItems = g.numItems;
    v.cellStart = g.cellStart;
    v.items = g.items;
    v.px = g.rec; v.py = g.rec + g.itemCap; v.pz = g.rec + 2 * g.itemCap;
    v.ax = g.rec + 3 * g.itemCap; v.ay = g.rec + 4 * g.itemCap; v.az = g.rec + 5 * g.itemCap;
    computeTranslations(g.box, v.trans);
    return v;
}

void freeGrid(GridStore &g) {
    for (int **p : {&g.cellStart, &g.cellFill, &g.keys, &g.itemsTmp, &g.items, &g.blockSums}) {
        if (*p) cudaFree(*p);
        *p = nullptr;
    }
    if (g.rec) cudaFree(g.rec);
    g.rec = nullptr;
    g.cellCap = g.itemCap = g.blockSumsCap = 0;
    g.valid = false;
}

void ensureGridCapacity(GridStore &g, size_t numCells, size_t numItems, size_t recDoublesPerItem) {
    RP_EXPECT(numCells < (1ull << 31) && numItems < (1ull << 31), "grid too large for 32-bit indices");
    if (numCells + 1 > g.cellCap || !g.cellStart) {
        if (g.cellStart) cudaFree(g.cellStart);
        if (g.cellFill) cudaFree(g.cellFill);
        g.cellStart = g.cellFill = nullptr;
        size_t newCap = numCells + numCells / 4 + 16;
        RP_CUDA(cudaMalloc(&g.cellStart, (newCap + 1) * sizeof(int)));
        RP_CUDA(cudaMalloc(&g.cellFill, (newCap + 1) * sizeof(int)));
        g.cellCap = newCap;
    }
    if (numItems > g.itemCap || !g.keys) {
        for (int **p : {&g.keys, &g.itemsTmp, &g.items}) { if (*p) cudaFree(*p); *p = nullptr; }
        if (g.rec) cudaFree(g.rec);
        g.rec = nullptr;
        size_t newCap = numItems + numItems / 4 + 16;
        RP_CUDA(cudaMalloc(&g.keys, newCap * sizeof(int)));
        RP_CUDA(cudaMalloc(&g.itemsTmp, newCap * sizeof(int)));
        RP_CUDA(cudaMalloc(&g.items, newCap * sizeof(int)));
        RP_CUDA(cudaMalloc(&g.rec, recDoublesPerItem * newCap * sizeof(double)));
        g.itemCap = newCap;
    }
    size_t numBlocks = (numCells + SCAN_BLOCK - 1) / SCAN_BLOCK;
    if (numBlocks > g.blockSumsCap || !g.blockSums) {
        if (g.blockSums) cudaFree(g.blockSums);
        g.blockSums = nullptr;
        RP_CUDA(cudaMalloc(&g.blockSums, (numBlocks + 16) * sizeof(int)));
        g.blockSumsCap = numBlocks + 16;
    }
}

void scanCells(Handle *h, GridStore &g, int *dMaxPop) {
    cudaStream_t st = h->stream;
    int numBlocks = (g.numCells + SCAN_BLOCK - 1) / SCAN_BLOCK;
    k_scan_block_sums<<<numBlocks, SCAN_BLOCK, 0, st>>>(g.numCells, g.cellFill, g.blockSums);
    k_scan_sums<<<1, SCAN_BLOCK, 0, st>>>(numBlocks, g.blockSums);
    k_scan_final<<<numBlocks, SCAN_BLOCK, 0, st>>>(g.numCells, g.cellFill, g.blockSums, g.cellStart, g.cellFill,
                                                   dMaxPop);
    h->launches += 3;
}

void scatterItems(Handle *h, GridStore &g) {
    cudaStream_t st = h->stream;
    int threads = 256;
    int blocks = (g.numItems + threads - 1) / threads;
    k_scatter<<<blocks, threads, 0, st>>>(g.numItems, g.keys, g.cellStart, g.cellFill, g.itemsTmp);
    int orderBlocks = static_cast<int>((static_cast<size_t>(g.numCells) * 32 + threads - 1) / threads);
    k_cell_order<<<orderBlocks, threads, 0, st>>>(g.numCells, g.cellStart, g.itemsTmp, g.items);
    h->launches += 2;
}

// Packing::rebuildNeighbourGrid cell-size policy (core/Packing.cpp:858-899) + NeighbourGrid::setupSizes
// (core/NeighbourGrid.cpp:178-201)
void rebuildReferenceGrid(Handle *h) {
    RP_EXPECT(h->hasInteraction && h->n > 0, "rebuildNeighbourGrid needs an interaction and particles");
    GridStore &g = h->grid;
    double cellSize = h->interactionRange;
    double volume = boxVolume(h->box);
    double minCellSize = std::cbrt(volume / static_cast<double>(h->n)) / 5;
    if (h->interactionRange < minCellSize)
        cellSize = minCellSize;

    double heights[3];
    boxHeights(h->box, heights);
    double maxH = std::max(heights[0], std::max(heights[1], heights[2]));
    double minH = std::min(heights[0], std::min(heights[1], heights[2]));
    g.cellSize = cellSize;
    h->gridDirty = false;
    h->lastRebuildMs = 0;
    if (cellSize * 4 > maxH) { h->hasGrid = false; g.valid = false; return; }
    static constexpr double CELL_SIZE_EPSILON = 1 + 1e-12;
    if (cellSize * CELL_SIZE_EPSILON > minH) { h->hasGrid = false; g.valid = false; return; }

    size_t numCells = 1;
    for (int i = 0; i < 3; i++) {
        // cellDivisions = floor(height / cellSize) + 2 including ghosts; real cells = floor(height / cellSize)
        g.n[i] = static_cast<int>(static_cast<std::size_t>(floor(heights[i] / cellSize)));
        RP_EXPECT(g.n[i] >= 1, "Neighbour grid cell too big");
        numCells *= g.n[i];
    }
    const Inter &I = h->inter;
    size_t numItems = h->n * I.kk;
    ensureGridCapacity(g, numCells, numItems);
    g.numCells = static_cast<int>(numCells);
    g.numItems = static_cast<int>(numItems);
    g.box = h->box;
    GridView view = makeView(g);

    cudaStream_t st = h->stream;
    RP_CUDA(cudaEventRecord(h->ev[2], st));
    zeroAsync(h, g.cellFill, (numCells + 1) * sizeof(int), st);
    zeroAsync(h, h->dFlags, 2 * sizeof(int), st);
    int threads = 256;
    int blocks = static_cast<int>((numItems + threads - 1) / threads);
    k_cell_keys<<<blocks, threads, 0, st>>>(g.numItems, I.k, h->cap, h->centreCap, h->pos, h->cabs, h->box, view,
                                            g.keys, g.cellFill, h->dFlags);
    scanCells(h, g, &h->dFlags[1]);
    scatterItems(h, g);
    k_gather_records<<<blocks, threads, 0, st>>>(g.numItems, I.kk, I.k, h->cap, h->centreCap, g.itemCap, g.items,
                                                 h->pos, h->orient, h->cabs, h->dInter, g.rec);
    h->launches += 2;
    RP_CUDA(cudaGetLastError());
    RP_CUDA(cudaEventRecord(h->ev[3], st));
    int flags[2];
    readbackSync(h, st, flags, h->dFlags, 2 * sizeof(int));
    float ms = 0;
    cudaEventElapsedTime(&ms, h->ev[2], h->ev[3]);
    h->lastRebuildMs = ms;
    g.maxPopulation = flags[1];
    g.valid = true;
    h->hasGrid = true;
    RP_EXPECT(!(flags[0] & 1), "a position lies outside the box (NeighbourGrid::positionToCellNo)");
}

void ensureGrid(Handle *h) {
    if (h->gridDirty) rebuildReferenceGrid(h);
}

}  // namespace rp
