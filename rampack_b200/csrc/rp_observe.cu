// rp_observe.cu — device-side order parameters: the observables a large run samples every few
// hundred cycles, evaluated where the packing lives instead of downloading 96 bytes per particle.
//   rp_nematic_order  NematicOrder::calculate      (core/observables/NematicOrder.cpp:9-28, eigenvalues :30-55)
//   rp_smectic_order  SmecticOrder::calculate      (core/observables/SmecticOrder.cpp:14-62), constant shape function
// Sums are formed per block in a fixed tree order and finished on the host in block order: results do not depend on
// scheduling.  The reference adds particle by particle; agreement is to rounding (tests: 1e-12).
#include "rp_internal.cuh"
#include <cmath>
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

}  // namespace rp
