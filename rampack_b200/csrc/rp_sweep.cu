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

namespace rp {

// ---------------------------------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) — counter-based, one independent stream per (particle, cycle, pass)
// ---------------------------------------------------------------------------------------------------------------
struct U4 { unsigned int x, y, z, w; };

RP_HD unsigned int mulhi32(unsigned int a, unsigned int b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return static_cast<unsigned int>((static_cast<unsigned long long>(a) * b) >> 32);
#endif
}

RP_HD U4 philox4x32(U4 c, unsigned int k0, unsigned int k1) {
    const unsigned int M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    #pragma unroll
    for (int r = 0; r < 10; r++) {
        unsigned int hi0 = mulhi32(M0, c.x), lo0 = M0 * c.x;
        unsigned int hi1 = mulhi32(M1, c.z), lo1 = M1 * c.z;
        U4 n;
        n.x = hi1 ^ c.y ^ k0; n.y = lo1; n.z = hi0 ^ c.w ^ k1; n.w = lo0;
        c = n;
        k0 += W0; k1 += W1;
    }
    return c;
}

struct Rng {
    U4 ctr;
    unsigned int k0, k1;
    U4 buf;
    int have;    // doubles left in buf (0..2)
    RP_HD void init(unsigned long long seed, unsigned int a, unsigned int b, unsigned int c) {
        k0 = static_cast<unsigned int>(seed); k1 = static_cast<unsigned int>(seed >> 32);
        ctr.x = a; ctr.y = b; ctr.z = c; ctr.w = 0;
        have = 0;
    }
    // uniform in [0, 1) with 53 random bits
    RP_HD double uniform() {
        if (have == 0) { buf = philox4x32(ctr, k0, k1); ctr.w++; have = 2; }
        unsigned int hi = (have == 2) ? buf.x : buf.z, lo = (have == 2) ? buf.y : buf.w;
        have--;
        unsigned long long bits = (static_cast<unsigned long long>(hi) << 32) | lo;
        return static_cast<double>(bits >> 11) * (1.0 / 9007199254740992.0);
    }
    // std::uniform_real_distribution<double>(a, b) semantics: a + (b - a) * u
    RP_HD double uniform(double a, double b) { return a + (b - a) * uniform(); }
};

// ---------------------------------------------------------------------------------------------------------------
// sweep grid
// ---------------------------------------------------------------------------------------------------------------
struct SweepGrid {
    int n[3];
    double relCell[3];     // 1 / n[i]
    double shift[3];       // random relative origin s
    V3 imageShift[8];      // M * f for f = (fx, fy, fz) in {0,1}^3, index fx + 2 fy + 4 fz
    V3 trans[27];
};

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
                                int *__restrict__ items, double *__restrict__ rec, int *__restrict__ maxNeigh)
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
            r[3] = orient[2 * cap + id]; r[4] = orient[5 * cap + id]; r[5] = orient[8 * cap + id];
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
// pair test on an already translated candidate (pos2t = pos2 + translation): same arithmetic as
// rp::overlapBetween with BcFixed
// ---------------------------------------------------------------------------------------------------------------
template <int HARD>
__device__ __forceinline__ int pairPrefilter(const Inter &I, double d2) {
    if (HARD == RP_HARD_SPHERE) return d2 < I.sphereThr2 ? 1 : 0;
    return scPrefilter(I, d2);
}

constexpr int SW_THREADS = 256;
constexpr int SW_WARPS = SW_THREADS / 32;
constexpr int SW_MAXOWN = 64;
constexpr int SW_QCAP = 64;

struct SweepArgs {
    int kind;                   // rp_move_kind
    double transStep, rotStep;
    unsigned long long seed;
    unsigned int cycle, pass;
    int colour[3];              // parity of the active cells per axis
    int nActive[3];             // active cells per axis
    int capCand;                // candidate capacity of the shared-memory staging area
    size_t cap;
};

struct TrialRec { double px, py, pz, ax, ay, az; };

template <int HARD>
__global__ void __launch_bounds__(SW_THREADS) k_sweep(SweepArgs A, SweepGrid sg, Box box,
                                                      const Inter *__restrict__ interPtr,
                                                      const int *__restrict__ cellStart,
                                                      const int *__restrict__ items, double *__restrict__ rec,
                                                      double *__restrict__ pos, double *__restrict__ orient,
                                                      unsigned long long *__restrict__ counters,
                                                      int *__restrict__ flags)
{
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ Inter I;
    __shared__ TrialRec trial[SW_MAXOWN];
    __shared__ unsigned long long ownMask[SW_MAXOWN];
    __shared__ unsigned int orderKey[SW_MAXOWN];
    __shared__ int order[SW_MAXOWN];
    __shared__ int rejected[SW_MAXOWN];     // 0 = alive, 1 = overlap with a frozen particle, 2 = left the cell
    __shared__ int accepted[SW_MAXOWN];
    __shared__ int movedList[SW_MAXOWN];
    __shared__ unsigned int queue[SW_WARPS][SW_QCAP];
    __shared__ int cellBegin[28], cellSrc[27], cellT[27];
    __shared__ int sOwnBegin, sNumCand, sNumOwn;

    double *cand = reinterpret_cast<double *>(smemRaw);     // 6 doubles per candidate, translation applied

    int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];

    // active cell of this CTA
    int b = blockIdx.x;
    int a0 = b % A.nActive[0], a1 = (b / A.nActive[0]) % A.nActive[1], a2 = b / (A.nActive[0] * A.nActive[1]);
    int cx = 2 * a0 + A.colour[0], cy = 2 * a1 + A.colour[1], cz = 2 * a2 + A.colour[2];
    int ownCell = (cz * sg.n[1] + cy) * sg.n[0] + cx;

    if (tid < 27) {
        int tIdx;
        int c2 = neighbourCell(sg.n, cx, cy, cz, tid / 9 - 1, (tid / 3) % 3 - 1, tid % 3 - 1, &tIdx);
        cellSrc[tid] = cellStart[c2];
        cellBegin[tid] = cellStart[c2 + 1] - cellStart[c2];     // count, turned into an offset below
        cellT[tid] = tIdx;
    }
    __syncthreads();
    if (tid == 0) {
        int off = 0;
        for (int c = 0; c < 27; c++) { int cnt = cellBegin[c]; cellBegin[c] = off; off += cnt; }
        cellBegin[27] = off;
        sNumCand = off;
        sOwnBegin = cellBegin[13];
        sNumOwn = cellBegin[14] - cellBegin[13];
    }
    __syncthreads();
    const int numCand = sNumCand, ownBeginAll = sOwnBegin, numOwnAll = sNumOwn;
    if (numOwnAll == 0) return;
    if (numCand > A.capCand) {
        if (tid == 0) atomicOr(&flags[3], 1);
        return;
    }
    const int ownStartAll = cellStart[ownCell];

    // ---- stage the 27 cells into shared memory, lattice translation applied (pos2 + translation) ----
    for (int c = 0; c < 27; c++) {
        int cnt6 = (cellBegin[c + 1] - cellBegin[c]) * 6;
        const double *src = rec + 6 * static_cast<size_t>(cellSrc[c]);
        double *dst = cand + 6 * cellBegin[c];
        V3 t = sg.trans[cellT[c]];
        for (int i = tid; i < cnt6; i += SW_THREADS) {
            int e = i % 6;
            double v = src[i];
            if (e == 0) v += t.x; else if (e == 1) v += t.y; else if (e == 2) v += t.z;
            dst[i] = v;
        }
    }

    // A cell with more than SW_MAXOWN particles is processed in batches of SW_MAXOWN trials; the particles of the
    // other batches are frozen candidates meanwhile (their accepted states are patched into shared memory).
    for (int batch0 = 0; batch0 < numOwnAll; batch0 += SW_MAXOWN) {
    const int numOwn = min(SW_MAXOWN, numOwnAll - batch0);
    const int ownBegin = ownBeginAll + batch0, ownStart = ownStartAll + batch0;

    // ---- trial moves: thread i < numOwn owns particle i of the batch ----
    M33 newOrient;
    V3 newPos = v3(0, 0, 0);
    int particleId = -1;
    bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    if (tid < numOwn) {
        int slot = ownStart + tid;
        particleId = items[slot];
        const double *r = rec + 6 * static_cast<size_t>(slot);
        V3 p = v3(r[0], r[1], r[2]);
        V3 aux = v3(r[3], r[4], r[5]);
        Rng rng;
        rng.init(A.seed, static_cast<unsigned int>(particleId), A.cycle, A.pass);
        int rej = 0;
        if (translated) {
            // TranslationSampler / RototranslationSampler: 3 x U(-step, step)  (RototranslationSampler.cpp:31-33)
            double tx = rng.uniform(-A.transStep, A.transStep);
            double ty = rng.uniform(-A.transStep, A.transStep);
            double tz = rng.uniform(-A.transStep, A.transStep);
            p = p + v3(tx, ty, tz);
            // ActiveDomain::isInside on the mass centre (Packing.cpp:95-96): the new position must stay in the cell
            V3 rel = absToRel(box, p);
            double q[3] = {rel.x - sg.shift[0], rel.y - sg.shift[1], rel.z - sg.shift[2]};
            int cc[3] = {cx, cy, cz};
            for (int d = 0; d < 3; d++) {
                if (q[d] < 0 || q[d] >= 1) rej = 2;
                else if (static_cast<int>(q[d] / sg.relCell[d]) != cc[d]) rej = 2;
            }
        }
        if (rotated) {
            // axis uniform in the unit ball by rejection, angle U(-a, a), Rodrigues (RototranslationSampler.cpp:35-43,
            // geometry/Matrix.tpp:437-449)
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
            M33 KK = mul(K, K);
            M33 R;
            for (int q = 0; q < 9; q++) R.a[q] = ((q % 4 == 0) ? 1.0 : 0.0) + K.a[q] * sina + (1 - cosa) * KK.a[q];
            M33 O;
            for (int q = 0; q < 9; q++) O.a[q] = orient[q * A.cap + particleId];
            newOrient = mul(R, O);     // Shape::rotate (core/Shape.cpp:14-16)
            if (HARD == RP_HARD_SPHEROCYLINDER) aux = column(newOrient, 2);
        }
        newPos = p;
        trial[tid].px = p.x; trial[tid].py = p.y; trial[tid].pz = p.z;
        trial[tid].ax = aux.x; trial[tid].ay = aux.y; trial[tid].az = aux.z;
        rejected[tid] = rej;
        accepted[tid] = 0;
        ownMask[tid] = 0ull;
        rng.ctr.w = 0x40000000u;     // separate sub-stream for the processing order
        rng.have = 0;
        orderKey[tid] = philox4x32(rng.ctr, rng.k0, rng.k1).x;
    }
    __syncthreads();

    // ---- phase A: centre-distance pre-filter of ALL (trial, candidate) pairs.  A lane owns up to two candidates
    // and collects, per candidate, the bitmask of trials it has to be tested against: a branch-free loop over
    // the trials (broadcast reads) the compiler can unroll and software-pipeline.
    // ---- phase B: the masks are expanded trial by trial into the warp queue (ballot + popc), and whenever 32
    // (trial, candidate) entries are queued the full test runs with every lane busy. ----
    {
        volatile int *vrej = rejected;
        unsigned int *q = queue[warp];
        int qn = 0;
        const double thr = (HARD == RP_HARD_SPHERE) ? I.sphereThr2 : I.scOuter2;
        unsigned long long aliveMask = 0;
        for (int i = 0; i < numOwn; i++) if (rejected[i] == 0) aliveMask |= 1ull << i;

        for (int base = warp * 64; base < numCand; base += SW_WARPS * 64) {
            int j0 = base + lane, j1 = base + 32 + lane;
            bool v0 = j0 < numCand, v1 = j1 < numCand;
            // far-away dummies for the invalid slots keep the loop branch-free
            double c0x = 1e300, c0y = 0, c0z = 0, c1x = 1e300, c1y = 0, c1z = 0;
            if (v0) { c0x = cand[6 * j0]; c0y = cand[6 * j0 + 1]; c0z = cand[6 * j0 + 2]; }
            if (v1) { c1x = cand[6 * j1]; c1y = cand[6 * j1 + 1]; c1z = cand[6 * j1 + 2]; }
            unsigned long long m0 = 0, m1 = 0;
            #pragma unroll 4
            for (int i = 0; i < numOwn; i++) {
                double tx = trial[i].px, ty = trial[i].py, tz = trial[i].pz;
                double ax = c0x - tx, ay = c0y - ty, az = c0z - tz;
                double bx = c1x - tx, by = c1y - ty, bz = c1z - tz;
                double da = (ax * ax + ay * ay) + az * az;
                double db = (bx * bx + by * by) + bz * bz;
                m0 |= static_cast<unsigned long long>(da < thr) << i;
                m1 |= static_cast<unsigned long long>(db < thr) << i;
            }
            // never test a particle against itself; skip trials that already left their cell
            int o0 = j0 - ownBegin, o1 = j1 - ownBegin;
            bool own0 = o0 >= 0 && o0 < numOwn, own1 = o1 >= 0 && o1 < numOwn;
            if (own0) m0 &= ~(1ull << o0);
            if (own1) m1 &= ~(1ull << o1);
            m0 &= aliveMask;
            m1 &= aliveMask;

            if (HARD == RP_HARD_SPHERE) {
                // the pre-filter IS the sphere test (SphereTraits.cpp:56-65)
                while (m0) {
                    int i = __ffsll(static_cast<long long>(m0)) - 1;
                    m0 &= m0 - 1;
                    if (own0) atomicOr(&ownMask[i], 1ull << o0); else vrej[i] = 1;
                }
                while (m1) {
                    int i = __ffsll(static_cast<long long>(m1)) - 1;
                    m1 &= m1 - 1;
                    if (own1) atomicOr(&ownMask[i], 1ull << o1); else vrej[i] = 1;
                }
                continue;
            }

            #pragma unroll 1
            for (int half = 0; half < 2; half++) {
                unsigned long long m = half ? m1 : m0;
                int j = half ? j1 : j0;
                unsigned long long any = m;
                for (int o = 16; o > 0; o >>= 1) any |= __shfl_xor_sync(0xffffffffu, any, o);
                while (any) {
                    int i = __ffsll(static_cast<long long>(any)) - 1;
                    any &= any - 1;
                    if (vrej[i]) continue;                       // early exit: the trial already hit a frozen particle
                    bool push = (m >> i) & 1;
                    unsigned int ballot = __ballot_sync(0xffffffffu, push);
                    if (push) q[qn + __popc(ballot & ((1u << lane) - 1))] = (static_cast<unsigned int>(i) << 16) | j;
                    qn += __popc(ballot);
                    __syncwarp();
                    if (qn >= 32) {
                        unsigned int e = q[qn - 32 + lane];
                        qn -= 32;
                        int ti = e >> 16, cj = e & 0xffff;
                        if (!vrej[ti]) {
                            const double *c = cand + 6 * cj;
                            V3 p1 = v3(trial[ti].px, trial[ti].py, trial[ti].pz), p2 = v3(c[0], c[1], c[2]);
                            int pre = scPrefilter(I, norm2(p2 - p1));
                            bool ov = pre == 1;
                            if (pre == 2)
                                ov = scFull(I, p1, v3(trial[ti].ax, trial[ti].ay, trial[ti].az), p2,
                                            v3(c[3], c[4], c[5]));
                            if (ov) {
                                int o = cj - ownBegin;
                                if (o >= 0 && o < numOwn) atomicOr(&ownMask[ti], 1ull << o);
                                else vrej[ti] = 1;
                            }
                        }
                        __syncwarp();
                    }
                }
            }
        }
        if (HARD != RP_HARD_SPHERE) {
            __syncwarp();
            if (lane < qn) {
                unsigned int e = q[lane];
                int ti = e >> 16, cj = e & 0xffff;
                if (!vrej[ti]) {
                    const double *c = cand + 6 * cj;
                    V3 p1 = v3(trial[ti].px, trial[ti].py, trial[ti].pz), p2 = v3(c[0], c[1], c[2]);
                    int pre = scPrefilter(I, norm2(p2 - p1));
                    bool ov = pre == 1;
                    if (pre == 2)
                        ov = scFull(I, p1, v3(trial[ti].ax, trial[ti].ay, trial[ti].az), p2, v3(c[3], c[4], c[5]));
                    if (ov) {
                        int o = cj - ownBegin;
                        if (o >= 0 && o < numOwn) atomicOr(&ownMask[ti], 1ull << o);
                        else vrej[ti] = 1;
                    }
                }
            }
        }
    }
    __syncthreads();

    // ---- phase C: sequential resolution inside the cell, random order, warp 0 ----
    if (tid < numOwn) {
        unsigned int k = orderKey[tid];
        int rank = 0;
        for (int j = 0; j < numOwn; j++) {
            unsigned int kj = orderKey[j];
            rank += (kj < k) || (kj == k && j < tid);
        }
        order[rank] = tid;
    }
    __syncthreads();
    int nAcc = 0, nRejDomain = 0, nRejOverlap = 0;
    if (warp == 0) {
        unsigned long long moved = 0;
        int nMoved = 0;
        for (int step = 0; step < numOwn; step++) {
            int i = order[step];
            int r = rejected[i];
            if (r == 2) { nRejDomain++; continue; }
            if (r == 1) { nRejOverlap++; continue; }
            bool rej = (ownMask[i] & ~moved) != 0;
            if (!rej && nMoved > 0) {
                V3 p1 = v3(trial[i].px, trial[i].py, trial[i].pz), a1 = v3(trial[i].ax, trial[i].ay, trial[i].az);
                bool ov = false;
                for (int m0 = 0; m0 < nMoved; m0 += 32) {
                    int m = m0 + lane;
                    if (m < nMoved) {
                        int bI = movedList[m];
                        V3 p2 = v3(trial[bI].px, trial[bI].py, trial[bI].pz);
                        int pre = pairPrefilter<HARD>(I, norm2(p2 - p1));
                        if (pre == 1) ov = true;
                        else if (pre == 2 && scFull(I, p1, a1, p2, v3(trial[bI].ax, trial[bI].ay, trial[bI].az)))
                            ov = true;
                    }
                }
                rej = __any_sync(0xffffffffu, ov);
            }
            if (rej) { nRejOverlap++; continue; }
            // hard interactions only: Metropolis u <= exp(-0/T) always accepts (Simulation.cpp:483)
            if (lane == 0) { accepted[i] = 1; movedList[nMoved] = i; }
            moved |= 1ull << i;
            nMoved++;
            nAcc++;
            __syncwarp();
        }
    }
    __syncthreads();

    // ---- write-back of the accepted moves: sorted record (image frame) and canonical arrays (wrapped) ----
    if (tid < numOwn && accepted[tid]) {
        double *r = rec + 6 * static_cast<size_t>(ownStart + tid);
        r[0] = trial[tid].px; r[1] = trial[tid].py; r[2] = trial[tid].pz;
        r[3] = trial[tid].ax; r[4] = trial[tid].ay; r[5] = trial[tid].az;
        double *cs = cand + 6 * (ownBegin + tid);      // own cell: no lattice translation
        cs[0] = trial[tid].px; cs[1] = trial[tid].py; cs[2] = trial[tid].pz;
        cs[3] = trial[tid].ax; cs[4] = trial[tid].ay; cs[5] = trial[tid].az;
        if (translated) {
            // Shape::translate (core/Shape.cpp:9-12): position += bc.getCorrection(position)
            V3 p = newPos + pbcCorrection(box, newPos);
            pos[particleId] = p.x; pos[A.cap + particleId] = p.y; pos[2 * A.cap + particleId] = p.z;
        }
        if (rotated)
            for (int q = 0; q < 9; q++) orient[q * A.cap + particleId] = newOrient.a[q];
    }
    if (tid == 0) {
        atomicAdd(&counters[0], static_cast<unsigned long long>(numOwn));
        atomicAdd(&counters[1], static_cast<unsigned long long>(nAcc));
        atomicAdd(&counters[2], static_cast<unsigned long long>(nRejDomain));
        atomicAdd(&counters[3], static_cast<unsigned long long>(nRejOverlap));
    }
    __syncthreads();
    }   // batches
}

// ---------------------------------------------------------------------------------------------------------------
// host driver: one call = one pass of N trial moves (every particle once) with one sampler
// ---------------------------------------------------------------------------------------------------------------
template <int HARD>
static void launchSweep(Handle *h, const SweepArgs &A0, const SweepGrid &sg, const int colourOrder[8], size_t smem) {
    static size_t configured = 0;
    if (smem > configured) {
        RP_CUDA(cudaFuncSetAttribute(k_sweep<HARD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     static_cast<int>(std::max<size_t>(smem, 48 * 1024))));
        configured = std::max<size_t>(smem, 48 * 1024);
    }
    GridStore &g = h->sgrid;
    for (int c = 0; c < 8; c++) {
        SweepArgs A = A0;
        int col = colourOrder[c];
        A.colour[0] = col & 1; A.colour[1] = (col >> 1) & 1; A.colour[2] = (col >> 2) & 1;
        int blocks = A.nActive[0] * A.nActive[1] * A.nActive[2];
        k_sweep<HARD><<<blocks, SW_THREADS, smem, h->stream>>>(A, sg, h->box, h->dInter, g.cellStart, g.items, g.rec,
                                                              h->pos, h->orient, h->dCounters, h->dFlags);
        h->launches++;
    }
}

void sweep(Handle *h, const rp_sweep_params &p, rp_sweep_counters *out) {
    const Inter &I = h->inter;
    if (I.k != 0 || I.soft != RP_SOFT_NONE || !(I.hard == RP_HARD_SPHERE || I.hard == RP_HARD_SPHEROCYLINDER))
        throw rp_error(RP_ERR_UNSUPPORTED, "rp_sweep: only hard spheres and hard spherocylinders are implemented");
    RP_EXPECT(p.kind >= RP_MOVE_TRANSLATION && p.kind <= RP_MOVE_ROTOTRANSLATION, "unknown sampler kind");
    bool translated = p.kind != RP_MOVE_ROTATION, rotated = p.kind != RP_MOVE_TRANSLATION;
    if (translated) RP_EXPECT(p.translation_step > 0, "translationStepSize > 0");
    if (rotated) RP_EXPECT(p.rotation_step > 0, "rotationStepSize > 0");

    double heights[3];
    boxHeights(h->box, heights);
    double volume = boxVolume(h->box);
    // cell width: at least totalRange (plus a rounding margin), and wide enough to hold ~16 particles
    const double TARGET_POPULATION = 16;
    double width = std::max(I.totalRange * (1 + 1e-9), std::cbrt(TARGET_POPULATION * volume / static_cast<double>(h->n)));
    SweepGrid sg;
    size_t numCells = 1;
    for (int i = 0; i < 3; i++) {
        int n = static_cast<int>(floor(heights[i] / width));
        if (n < 2 && heights[i] >= 2 * I.totalRange * (1 + 1e-9)) n = 2;
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
    int recBlocks = static_cast<int>((numCells * 32 + 255) / 256);
    if (I.hard == RP_HARD_SPHERE)
        k_sweep_records<RP_HARD_SPHERE><<<recBlocks, 256, 0, st>>>(g.numCells, h->cap, g.cellStart, g.itemsTmp,
                                                                    g.keys, h->pos, h->orient, sg, g.items, g.rec,
                                                                    &h->dFlags[2]);
    else
        k_sweep_records<RP_HARD_SPHEROCYLINDER><<<recBlocks, 256, 0, st>>>(g.numCells, h->cap, g.cellStart,
                                                                            g.itemsTmp, g.keys, h->pos, h->orient,
                                                                            sg, g.items, g.rec, &h->dFlags[2]);
    h->launches += 3;
    RP_CUDA(cudaEventRecord(h->ev[1], st));
    int flags[4];
    RP_CUDA(cudaMemcpyAsync(flags, h->dFlags, 4 * sizeof(int), cudaMemcpyDeviceToHost, st));
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    g.maxPopulation = flags[1];
    int maxNeigh = flags[2];
    if (maxNeigh >= 65536)
        throw rp_error(RP_ERR_CAPACITY, "rp_sweep: a cell neighbourhood holds more than 65535 particles");
    size_t smem = static_cast<size_t>(maxNeigh) * 6 * sizeof(double);
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
    A.capCand = maxNeigh;
    A.cap = h->cap;
    RP_CUDA(cudaEventRecord(h->ev[2], st));
    if (I.hard == RP_HARD_SPHERE) launchSweep<RP_HARD_SPHERE>(h, A, sg, colourOrder, smem);
    else launchSweep<RP_HARD_SPHEROCYLINDER>(h, A, sg, colourOrder, smem);
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
    if (out) {
        out->attempted = counters[0];
        out->accepted = counters[1];
        out->rejected_domain = counters[2];
        out->rejected_overlap = counters[3];
        out->sub_sweeps = 8;
        out->kernel_launches = 8 + 6;
    }
}

}  // namespace rp
