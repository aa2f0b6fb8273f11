// rp_observe.cu — device-side order parameters: the observables a large run samples every few
// hundred cycles, evaluated where the packing lives instead of downloading 96 bytes per particle.
//   rp_nematic_order  NematicOrder::calculate      (core/observables/NematicOrder.cpp:9-28, eigenvalues :30-55)
//   rp_smectic_order  SmecticOrder::calculate      (core/observables/SmecticOrder.cpp:14-62), constant shape function
//   rp_pair_density_correlation  PairDensityCorrelation::addSnapshot with RadialEnumerator
//                     (core/observables/correlation/PairDensityCorrelation.cpp:8-41, RadialEnumerator.cpp:12-50)
// Sums are formed per block in a fixed tree order and finished on the host in block order: results do not depend on
// scheduling.  The reference adds particle by particle; agreement is to rounding (tests: 1e-12).
#include "rp_internal.cuh"
#include <cmath>
#include <cstdlib>
#include <algorithm>
#include <vector>

namespace rp {

constexpr int OBS_THREADS = 256, OBS_BLOCKS = 296;     // 2 CTAs per SM

static void ensureCellEnergyBuffer(Handle *h, size_t need) { devEnsure(h->dCellEnergy, h->cellEnergyCap, need); }

template <int M>
__device__ __forceinline__ void blockReduce(double (&v)[M], double *smem, double *out) {
    for (int m = 0; m < M; m++) {
        smem[threadIdx.x] = v[m];
        __syncthreads();
        for (int o = OBS_THREADS / 2; o > 0; o >>= 1) {
            if (threadIdx.x < o) smem[threadIdx.x] += smem[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) out[m] = smem[0];
        __syncthreads();
    }
}

// partial[block][6] = sum of a_i a_j (xx xy xz yy yz zz), a = orientation * bodyAxis
__global__ void __launch_bounds__(OBS_THREADS) k_q_partial(int n, size_t cap, const double *__restrict__ orient,
                                                           V3 bodyAxis, double *__restrict__ partial)
{
    __shared__ double smem[OBS_THREADS];
    double s[6] = {0, 0, 0, 0, 0, 0};
    for (int p = blockIdx.x * OBS_THREADS + threadIdx.x; p < n; p += gridDim.x * OBS_THREADS) {
        M33 o;
        for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + p];
        V3 a = mul(o, bodyAxis);
        s[0] += a.x * a.x; s[1] += a.x * a.y; s[2] += a.x * a.z;
        s[3] += a.y * a.y; s[4] += a.y * a.z; s[5] += a.z * a.z;
    }
    blockReduce<6>(s, smem, partial + 6 * blockIdx.x);
}

// partial[k][block][2] = sum over the block's particles of exp(i kvec . (pos + orientation * focal))
__global__ void __launch_bounds__(OBS_THREADS) k_smectic_partial(int n, size_t cap, const double *__restrict__ pos,
                                                                 const double *__restrict__ orient, V3 focal,
                                                                 int useFocal, const double *__restrict__ kvecs,
                                                                 double *__restrict__ partial)
{
    __shared__ double smem[OBS_THREADS];
    const int kIdx = blockIdx.y;
    const V3 k = v3(kvecs[3 * kIdx], kvecs[3 * kIdx + 1], kvecs[3 * kIdx + 2]);
    double s[2] = {0, 0};
    for (int p = blockIdx.x * OBS_THREADS + threadIdx.x; p < n; p += gridDim.x * OBS_THREADS) {
        V3 r = v3(pos[p], pos[cap + p], pos[2 * cap + p]);
        if (useFocal) {
            M33 o;
            for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + p];
            r = r + mul(o, focal);
        }
        double sn, cs;
        sincos(dot(k, r), &sn, &cs);
        s[0] += cs;
        s[1] += sn;
    }
    blockReduce<2>(s, smem, partial + 2 * (static_cast<size_t>(kIdx) * gridDim.x + blockIdx.x));
}

// NematicOrder::calculateEigenvalues (NematicOrder.cpp:30-55): Cardano, arXiv:1306.6291
static void eigenvalues(const double Q[9], double out[3]) {
    double A11 = Q[0], A12 = Q[1], A13 = Q[2], A22 = Q[4], A23 = Q[5], A33 = Q[8];
    double b = A11 + A22 + A33;
    double c = A11 * A22 + A11 * A33 + A22 * A33 - A12 * A12 - A13 * A13 - A23 * A23;
    double d = A11 * A23 * A23 + A22 * A13 * A13 + A33 * A12 * A12 - A11 * A22 * A33 - 2 * A12 * A13 * A23;
    double p = b * b - 3 * c;
    if (!(p > 0)) throw rp_error(RP_ERR_STATE, "NematicOrder: Assert(p > 0) failed (isotropic Q tensor)");
    double sqrt_p = std::sqrt(p);
    double q = 2 * b * b * b - 9 * b * c - 27 * d;
    double ratio = q / 2 / (sqrt_p * sqrt_p * sqrt_p);
    if (std::abs(ratio) > 1) ratio = ratio < 0 ? -1 : 1;
    double Delta = std::acos(ratio);
    out[0] = 1. / 3 * (b + 2 * sqrt_p * std::cos(Delta / 3));
    out[1] = 1. / 3 * (b + 2 * sqrt_p * std::cos((Delta + 2 * M_PI) / 3));
    out[2] = 1. / 3 * (b + 2 * sqrt_p * std::cos((Delta - 2 * M_PI) / 3));
}

void nematicOrder(Handle *h, const double axis[3], double qOut[9], double *p2Out) {
    const int n = static_cast<int>(h->n);
    const int blocks = std::min(OBS_BLOCKS, (n + OBS_THREADS - 1) / OBS_THREADS);
    ensureCellEnergyBuffer(h, static_cast<size_t>(blocks) * 6);
    k_q_partial<<<blocks, OBS_THREADS, 0, h->stream>>>(n, h->cap, h->orient, v3(axis[0], axis[1], axis[2]),
                                                       h->dCellEnergy);
    h->launches++;
    std::vector<double> part(static_cast<size_t>(blocks) * 6);
    RP_CUDA(cudaMemcpyAsync(part.data(), h->dCellEnergy, part.size() * sizeof(double), cudaMemcpyDeviceToHost,
                            h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_CUDA(cudaGetLastError());
    double s[6] = {0, 0, 0, 0, 0, 0};
    for (int b = 0; b < blocks; b++) for (int m = 0; m < 6; m++) s[m] += part[6 * b + m];
    const double N = static_cast<double>(h->n);
    double Q[9] = {s[0] / N, s[1] / N, s[2] / N, s[1] / N, s[3] / N, s[4] / N, s[2] / N, s[4] / N, s[5] / N};
    // QTensor = 0.5 * (3. * QTensor - identity)
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Q[3 * i + j] = 0.5 * (3. * Q[3 * i + j] - (i == j ? 1.0 : 0.0));
    double ev[3];
    eigenvalues(Q, ev);
    double p2 = ev[0];
    for (int i = 1; i < 3; i++) if (std::abs(p2) < std::abs(ev[i])) p2 = ev[i];      // highest magnitude
    if (qOut) for (int i = 0; i < 9; i++) qOut[i] = Q[i];
    if (p2Out) *p2Out = p2;
}

void smecticOrder(Handle *h, const int range[3], const double focal[3], double *tauAbs, int hklOut[3],
                  double kOut[3]) {
    // the k vectors in the reference's iteration order (equivalent k and -k are not visited twice)
    std::vector<int> hkl;
    for (int cx = 0; cx <= range[0]; cx++)
        for (int cy = (cx == 0 ? 0 : -range[1]); cy <= range[1]; cy++)
            for (int cz = ((cx == 0 && cy == 0) ? 0 : -range[2]); cz <= range[2]; cz++) {
                if (cx == 0 && cy == 0 && cz == 0) continue;
                hkl.push_back(cx); hkl.push_back(cy); hkl.push_back(cz);
            }
    const int numK = static_cast<int>(hkl.size() / 3);
    std::vector<double> kvecs(3 * static_cast<size_t>(numK));
    for (int v = 0; v < numK; v++) {
        // kTauVector = dimInv.transpose() * (2 pi n)
        V3 kt = v3(2 * M_PI * hkl[3 * v], 2 * M_PI * hkl[3 * v + 1], 2 * M_PI * hkl[3 * v + 2]);
        const M33 &inv = h->box.inv;
        kvecs[3 * v] = (inv.a[0] * kt.x + inv.a[3] * kt.y) + inv.a[6] * kt.z;
        kvecs[3 * v + 1] = (inv.a[1] * kt.x + inv.a[4] * kt.y) + inv.a[7] * kt.z;
        kvecs[3 * v + 2] = (inv.a[2] * kt.x + inv.a[5] * kt.y) + inv.a[8] * kt.z;
    }
    const int n = static_cast<int>(h->n);
    const int blocks = std::max(1, std::min(64, (n + OBS_THREADS * 8 - 1) / (OBS_THREADS * 8)));
    const size_t need = 3 * static_cast<size_t>(numK) + 2 * static_cast<size_t>(numK) * blocks;
    ensureCellEnergyBuffer(h, need);
    double *dK = h->dCellEnergy, *dPart = h->dCellEnergy + 3 * static_cast<size_t>(numK);
    RP_CUDA(cudaMemcpyAsync(dK, kvecs.data(), kvecs.size() * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    const bool useFocal = focal && (focal[0] != 0 || focal[1] != 0 || focal[2] != 0);
    k_smectic_partial<<<dim3(blocks, numK), OBS_THREADS, 0, h->stream>>>(
        n, h->cap, h->pos, h->orient, useFocal ? v3(focal[0], focal[1], focal[2]) : v3(0, 0, 0), useFocal ? 1 : 0,
        dK, dPart);
    h->launches++;
    std::vector<double> part(2 * static_cast<size_t>(numK) * blocks);
    RP_CUDA(cudaMemcpyAsync(part.data(), dPart, part.size() * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_CUDA(cudaGetLastError());
    const double volume = boxVolume(h->box);
    double best = 0;
    int bestV = -1;
    for (int v = 0; v < numK; v++) {
        double re = 0, im = 0;
        for (int b = 0; b < blocks; b++) {
            re += part[2 * (static_cast<size_t>(v) * blocks + b)];
            im += part[2 * (static_cast<size_t>(v) * blocks + b) + 1];
        }
        re /= volume; im /= volume;                      // tau_ /= volume (SmecticOrder.cpp:31)
        double mag = std::hypot(re, im);
        if (mag > best) { best = mag; bestV = v; }
    }
    if (tauAbs) *tauAbs = best;
    for (int i = 0; i < 3; i++) {
        if (hklOut) hklOut[i] = bestV >= 0 ? hkl[3 * bestV + i] : 0;
        if (kOut) kOut[i] = bestV >= 0 ? kvecs[3 * bestV + i] : 0;
    }
}


// ---------------------------------------------------------------------------------------------------------------
// g(r): the reference visits ALL pairs i < j (RadialEnumerator.cpp:21-31) — so does this kernel, tile by tile, but a
// pair reaches the reference's FP64 arithmetic (getTranslation, norm, bin = floor(d / step)) only if a conservative
// FP32 minimum-image distance in relative coordinates says it may lie inside maxR.  Counts are integers: equal to the
// reference's bin for bin.
// ---------------------------------------------------------------------------------------------------------------
constexpr int PD_TILE = 256, PD_MAXBINS = 4096;

struct PdArgs {
    float m[9];            // box matrix in FP32 (row-major)
    float thr2;            // (maxR + FP32 error bound)^2
    double maxR, step;
    int numBins;
    V3 focal;
    int useFocal;
};

__device__ __forceinline__ V3 focalPoint(int p, size_t cap, const double *__restrict__ pos,
                                         const double *__restrict__ orient, const PdArgs &A)
{
    V3 r = v3(pos[p], pos[cap + p], pos[2 * cap + p]);
    if (A.useFocal) {
        M33 o;
        for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + p];
        r = r + mul(o, A.focal);
    }
    return r;
}

__global__ void __launch_bounds__(PD_TILE) k_pair_density(int n, size_t cap, const double *__restrict__ pos,
                                                          const double *__restrict__ orient, Box box, PdArgs A,
                                                          unsigned long long *__restrict__ hist)
{
    if (blockIdx.x < blockIdx.y) return;                   // tile pairs (I <= J) only
    __shared__ double sx[PD_TILE], sy[PD_TILE], sz[PD_TILE];
    __shared__ float4 sRel[PD_TILE];
    __shared__ unsigned int sHist[PD_MAXBINS];
    const int t = threadIdx.x;
    for (int b = t; b < A.numBins; b += PD_TILE) sHist[b] = 0u;
    const int i = blockIdx.y * PD_TILE + t, jLoad = blockIdx.x * PD_TILE + t;
    V3 pi = v3(0, 0, 0);
    float rix = 0.f, riy = 0.f, riz = 0.f;
    if (i < n) {
        pi = focalPoint(i, cap, pos, orient, A);
        const V3 r = absToRel(box, pi);
        rix = static_cast<float>(r.x); riy = static_cast<float>(r.y); riz = static_cast<float>(r.z);
    }
    if (jLoad < n) {
        const V3 pj = focalPoint(jLoad, cap, pos, orient, A);
        const V3 r = absToRel(box, pj);
        sx[t] = pj.x; sy[t] = pj.y; sz[t] = pj.z;
        sRel[t] = make_float4(static_cast<float>(r.x), static_cast<float>(r.y), static_cast<float>(r.z), 0.f);
    } else {
        sRel[t] = make_float4(1e18f, 0.f, 0.f, 0.f);
    }
    __syncthreads();
    const int jBase = blockIdx.x * PD_TILE;
    if (i < n) {
        #pragma unroll 4
        for (int jj = 0; jj < PD_TILE; jj++) {
            const float4 rj = sRel[jj];
            float dx = rj.x - rix, dy = rj.y - riy, dz = rj.z - riz;
            dx -= rintf(dx); dy -= rintf(dy); dz -= rintf(dz);
            const float ax = fmaf(A.m[2], dz, fmaf(A.m[1], dy, A.m[0] * dx));
            const float ay = fmaf(A.m[5], dz, fmaf(A.m[4], dy, A.m[3] * dx));
            const float az = fmaf(A.m[8], dz, fmaf(A.m[7], dy, A.m[6] * dx));
            const float d2 = fmaf(az, az, fmaf(ay, ay, ax * ax));
            const int j = jBase + jj;
            if (d2 < A.thr2 && j > i && j < n) {
                // RadialEnumerator.cpp:24-29, PairDensityCorrelation.cpp:29-40, Histogram.tpp:61-69
                const V3 pj = v3(sx[jj], sy[jj], sz[jj]);
                const V3 p2 = pj + pbcTranslation(box, pi, pj);
                const double distance = sqrt(norm2(p2 - pi));
                if (!(distance > A.maxR)) {
                    long long bin = static_cast<long long>(floor(distance / A.step));
                    if (bin >= A.numBins) bin = A.numBins - 1;
                    atomicAdd(&sHist[bin], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int b = t; b < A.numBins; b += PD_TILE)
        if (sHist[b]) atomicAdd(&hist[b], static_cast<unsigned long long>(sHist[b]));
}

// ---- cell-list form: same pairs, same FP64 arithmetic for the survivors, O(N K) candidates ----
// Focal points are binned into cells at least (maxR + margin) wide (relative coordinates: triclinic-safe, at least 3
// cells per axis so that the 27 neighbour cells are distinct); every unordered pair within maxR lies in the same or in
// adjacent cells and is visited once: own cell j > i (sorted order), 13 "positive" neighbour cells all (i, j).
struct PdCellArgs { int n[3]; };

__global__ void k_pd_keys(int n, size_t cap, const double *__restrict__ pos, const double *__restrict__ orient, Box box,
                          PdArgs A, PdCellArgs C, int *__restrict__ keys, int *__restrict__ hist)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 r = absToRel(box, focalPoint(i, cap, pos, orient, A));
    const double q[3] = {r.x - floor(r.x), r.y - floor(r.y), r.z - floor(r.z)};
    int c[3];
    for (int d = 0; d < 3; d++) {
        int v = static_cast<int>(q[d] * C.n[d]);
        c[d] = v < 0 ? 0 : (v >= C.n[d] ? C.n[d] - 1 : v);
    }
    const int cell = (c[2] * C.n[1] + c[1]) * C.n[0] + c[0];
    keys[i] = cell;
    atomicAdd(&hist[cell], 1);
}

// sorted records: {x, y, z} FP64 focal point + FP32 relative coordinates (packed into two more doubles)
__global__ void k_pd_records(int n, size_t cap, const double *__restrict__ pos, const double *__restrict__ orient,
                             Box box, PdArgs A, const int *__restrict__ items, double *__restrict__ rec)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const V3 p = focalPoint(items[s], cap, pos, orient, A);
    const V3 r = absToRel(box, p);
    double *o = rec + 5 * static_cast<size_t>(s);
    o[0] = p.x; o[1] = p.y; o[2] = p.z;
    float *f = reinterpret_cast<float *>(o + 3);
    f[0] = static_cast<float>(r.x); f[1] = static_cast<float>(r.y); f[2] = static_cast<float>(r.z); f[3] = 0.f;
}

constexpr int PDC_THREADS = 128;
__global__ void __launch_bounds__(PDC_THREADS) k_pair_density_cells(PdCellArgs C, const int *__restrict__ cellStart,
                                                                    const double *__restrict__ rec, Box box, PdArgs A,
                                                                    unsigned long long *__restrict__ hist)
{
    __shared__ unsigned int sHist[PD_MAXBINS];
    const int t = threadIdx.x;
    for (int b = t; b < A.numBins; b += PDC_THREADS) sHist[b] = 0u;
    __syncthreads();
    const int cell = blockIdx.x;
    const int cx = cell % C.n[0], cy = (cell / C.n[0]) % C.n[1], cz = cell / (C.n[0] * C.n[1]);
    const int s1 = cellStart[cell], n1 = cellStart[cell + 1] - s1;
    for (int code = 13; code < 27 && n1 > 0; code++) {
        const int dx = code % 3 - 1, dy = (code / 3) % 3 - 1, dz = code / 9 - 1;
        const int c2 = (((cz + dz + C.n[2]) % C.n[2]) * C.n[1] + (cy + dy + C.n[1]) % C.n[1]) * C.n[0]
                       + (cx + dx + C.n[0]) % C.n[0];
        const int s2 = cellStart[c2], n2 = cellStart[c2 + 1] - s2;
        const long long pairs = static_cast<long long>(n1) * n2;
        for (long long p = t; p < pairs; p += PDC_THREADS) {
            const int a = static_cast<int>(p / n2), b = static_cast<int>(p - static_cast<long long>(a) * n2);
            if (code == 13 && b <= a) continue;
            const double *ra = rec + 5 * static_cast<size_t>(s1 + a), *rb = rec + 5 * static_cast<size_t>(s2 + b);
            const float *fa = reinterpret_cast<const float *>(ra + 3), *fb = reinterpret_cast<const float *>(rb + 3);
            float ddx = fb[0] - fa[0], ddy = fb[1] - fa[1], ddz = fb[2] - fa[2];
            ddx -= rintf(ddx); ddy -= rintf(ddy); ddz -= rintf(ddz);
            const float ax = fmaf(A.m[2], ddz, fmaf(A.m[1], ddy, A.m[0] * ddx));
            const float ay = fmaf(A.m[5], ddz, fmaf(A.m[4], ddy, A.m[3] * ddx));
            const float az = fmaf(A.m[8], ddz, fmaf(A.m[7], ddy, A.m[6] * ddx));
            if (fmaf(az, az, fmaf(ay, ay, ax * ax)) < A.thr2) {
                // RadialEnumerator.cpp:24-29, PairDensityCorrelation.cpp:29-40, Histogram.tpp:61-69
                const V3 pi = v3(ra[0], ra[1], ra[2]), pj = v3(rb[0], rb[1], rb[2]);
                const V3 p2 = pj + pbcTranslation(box, pi, pj);
                const double distance = sqrt(norm2(p2 - pi));
                if (!(distance > A.maxR)) {
                    long long bin = static_cast<long long>(floor(distance / A.step));
                    if (bin >= A.numBins) bin = A.numBins - 1;
                    atomicAdd(&sHist[bin], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int b = t; b < A.numBins; b += PDC_THREADS)
        if (sHist[b]) atomicAdd(&hist[b], static_cast<unsigned long long>(sHist[b]));
}

void pairDensityCorrelation(Handle *h, double maxR, int numBins, const double focal[3], unsigned long long *countsOut,
                            double *gOut) {
    const int n = static_cast<int>(h->n);
    ensureCellEnergyBuffer(h, static_cast<size_t>(numBins) + 2);
    unsigned long long *dHist = reinterpret_cast<unsigned long long *>(h->dCellEnergy);
    RP_CUDA(cudaMemsetAsync(dHist, 0, sizeof(unsigned long long) * numBins, h->stream));
    PdArgs A;
    double len = 0;
    for (int q = 0; q < 9; q++) { A.m[q] = static_cast<float>(h->box.m.a[q]); len += std::abs(h->box.m.a[q]); }
    // FP32 error of the minimum-image distance: relative coordinates carry 2^-24, the matrix product multiplies by at
    // most the sum of the |entries|; ten times that, plus the same relative error on maxR
    const double thr = maxR * (1 + 1e-5) + 1e-5 * len;
    A.thr2 = std::nextafterf(static_cast<float>(thr * thr), INFINITY);
    A.maxR = maxR;
    A.step = (maxR - 0.0) / static_cast<double>(numBins);            // Histogram.tpp:91
    A.numBins = numBins;
    A.useFocal = focal && (focal[0] != 0 || focal[1] != 0 || focal[2] != 0);
    A.focal = A.useFocal ? v3(focal[0], focal[1], focal[2]) : v3(0, 0, 0);
    // cell list when every box height holds at least three cells of (maxR + margin), else all pairs, tile by tile
    double heights[3];
    boxHeights(h->box, heights);
    PdCellArgs C;
    size_t numCells = 1;
    for (int i = 0; i < 3; i++) {
        C.n[i] = static_cast<int>(std::min(512.0, std::floor(heights[i] / (thr * (1 + 1e-6)))));
        numCells *= static_cast<size_t>(std::max(C.n[i], 0));
    }
    const bool cells = C.n[0] >= 3 && C.n[1] >= 3 && C.n[2] >= 3 && !std::getenv("RP_GR_ALL_PAIRS")
                       && numCells <= 4 * static_cast<size_t>(n) + 64;
    if (cells) {
        GridStore &g = h->ogrid;
        ensureGridCapacity(g, numCells, h->n, 5);
        g.numCells = static_cast<int>(numCells);
        g.numItems = n;
        RP_CUDA(cudaMemsetAsync(g.cellFill, 0, (numCells + 1) * sizeof(int), h->stream));
        k_pd_keys<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, h->pos, h->orient, h->box, A, C, g.keys, g.cellFill);
        scanCells(h, g, reinterpret_cast<int *>(dHist + numBins));
        scatterItems(h, g);
        k_pd_records<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, h->pos, h->orient, h->box, A, g.items, g.rec);
        k_pair_density_cells<<<static_cast<int>(numCells), PDC_THREADS, 0, h->stream>>>(C, g.cellStart, g.rec, h->box,
                                                                                         A, dHist);
        h->launches += 3;
    } else {
        const int tiles = (n + PD_TILE - 1) / PD_TILE;
        k_pair_density<<<dim3(tiles, tiles), PD_TILE, 0, h->stream>>>(n, h->cap, h->pos, h->orient, h->box, A, dHist);
        h->launches++;
    }
    std::vector<unsigned long long> counts(numBins);
    RP_CUDA(cudaMemcpyAsync(counts.data(), dHist, sizeof(unsigned long long) * numBins, cudaMemcpyDeviceToHost,
                            h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_CUDA(cudaGetLastError());
    // PairDensityCorrelation.cpp:12-16: bin *= 2 / (expected molecules in the shell) / N, expected from the number
    // density (RadialEnumerator.cpp:34-50, in the reference's operation order)
    const double numberDensity = static_cast<double>(h->n) / std::abs(boxVolume(h->box));
    auto inBall = [numberDensity](double r) { return 4. / 3 * M_PI * r * r * r * numberDensity; };
    for (int b = 0; b < numBins; b++) {
        if (countsOut) countsOut[b] = counts[b];
        if (gOut) {
            // Histogram::getBinDividers (Histogram.tpp:141-151) with min = 0
            const double nb = static_cast<double>(numBins);
            const double lo = 0.0 * ((nb - b) / nb) + maxR * (static_cast<double>(b) / nb);
            const double hi = 0.0 * ((nb - (b + 1)) / nb) + maxR * (static_cast<double>(b + 1) / nb);
            double factor = inBall(hi) - inBall(lo);
            factor = 2. / factor / static_cast<double>(h->n);
            gOut[b] = static_cast<double>(counts[b]) * factor;
        }
    }
}

}  // namespace rp
