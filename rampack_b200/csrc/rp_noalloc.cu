// rp_noalloc.cu — rp_nematic_order_static: the nematic order parameter without any device allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is a variant of
// rp_nematic_order (rp_observe.cu, k_q_partial) whose per-block partial sums live in a file-scope __device__ array
// instead of the handle's growable scratch buffer:
//   - qScratch below: (NA_MAX_BLOCKS + 1) x 6 doubles = 49 200 bytes, held by every GPU that loads the library until
//     the process exits;
//   - one block per NA_CHUNK = 4096 particles, so the limit is NA_MAX_BLOCKS x NA_CHUNK = 4 194 304 particles
//     (RP_NEMATIC_STATIC_MAX_PARTICLES in include/rampack_b200.h); larger packings are rejected with
//     RP_ERR_PRECONDITION and the caller keeps rp_nematic_order;
//   - no cudaMalloc / cudaFree / allocating library call on the path: two kernel launches, one 48-byte
//     cudaMemcpyAsync into a stack buffer, one stream synchronisation;
//   - every call on a device shares the one array.  Nothing in it has to start at zero: each slot a launch reads was
//     written by the same call's first kernel, so the previous call's contents are simply overwritten.  Calls that
//     OVERLAP on the GPU (two handles on different streams of the same device, or two host threads) race on it;
//     callers serialise them.
// In host code the array's name is a host-side stand-in, not a device address: the address the kernels get comes from
// cudaGetSymbolAddress, once per device, with that device current (scratchFor below).
//
// Summation order: fixed tree inside a block, blocks in index order in the finishing kernel — independent of
// scheduling, but not the order of rp_nematic_order (grid-stride over at most 296 blocks, host sum), so the two agree
// to rounding (tests/test_zz_no_alloc.py: 1e-12, the tolerance of the observables tests).
#include "rp_internal.cuh"
#include <cmath>
#include <mutex>

namespace rp {

extern thread_local std::string tlsError;

constexpr int NA_THREADS = 256;
constexpr int NA_CHUNK = 4096;                 // particles per block
constexpr int NA_MAX_BLOCKS = 1024;            // -> RP_NEMATIC_STATIC_MAX_PARTICLES = 1024 * 4096
constexpr int NA_MAX_DEVICES = 64;
static_assert(static_cast<unsigned long long>(NA_MAX_BLOCKS) * NA_CHUNK == RP_NEMATIC_STATIC_MAX_PARTICLES,
              "the header's limit must be the scratch array's capacity");

// rows 0 .. NA_MAX_BLOCKS-1: per-block sums (xx xy xz yy yz zz); row NA_MAX_BLOCKS: their total
static __device__ __align__(256) double qScratch[(NA_MAX_BLOCKS + 1) * 6];

static std::mutex scratchMutex;
static double *scratchAddr[NA_MAX_DEVICES];     // device address of qScratch per device ordinal, nullptr = not asked yet

// the device address of qScratch on `device`, which must be the current device
static double *scratchFor(int device) {
    if (device < 0 || device >= NA_MAX_DEVICES)
        throw rp_error(RP_ERR_UNSUPPORTED, "rp_nematic_order_static: device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(scratchMutex);
    if (!scratchAddr[device]) {
        double *p = nullptr;
        RP_CUDA(cudaGetSymbolAddress(reinterpret_cast<void **>(&p), qScratch));
        if (!p) throw rp_error(RP_ERR_CUDA, "rp_nematic_order_static: cudaGetSymbolAddress returned a null address");
        scratchAddr[device] = p;
    }
    return scratchAddr[device];
}

// partial[block][6] = sum over the block's chunk of a_i a_j (xx xy xz yy yz zz), a = orientation * bodyAxis
__global__ void __launch_bounds__(NA_THREADS) k_q_partial_static(int n, size_t cap, const double *__restrict__ orient,
                                                                  V3 bodyAxis, double *__restrict__ partial)
{
    __shared__ double smem[NA_THREADS];
    const int first = blockIdx.x * NA_CHUNK;
    const int last = min(n, first + NA_CHUNK);
    double s[6] = {0, 0, 0, 0, 0, 0};
    for (int p = first + threadIdx.x; p < last; p += NA_THREADS) {
        M33 o;
        for (int q = 0; q < 9; q++) o.a[q] = orient[q * cap + p];
        V3 a = mul(o, bodyAxis);
        s[0] += a.x * a.x; s[1] += a.x * a.y; s[2] += a.x * a.z;
        s[3] += a.y * a.y; s[4] += a.y * a.z; s[5] += a.z * a.z;
    }
    for (int m = 0; m < 6; m++) {
        smem[threadIdx.x] = s[m];
        __syncthreads();
        for (int o = NA_THREADS / 2; o > 0; o >>= 1) {
            if (threadIdx.x < o) smem[threadIdx.x] += smem[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0) partial[6 * blockIdx.x + m] = smem[0];
        __syncthreads();
    }
}

// total[6] = partial rows 0 .. blocks-1 added in row order (one thread per component: a fixed order)
__global__ void k_q_finish_static(int blocks, const double *__restrict__ partial, double *__restrict__ total) {
    const int m = threadIdx.x;
    if (m >= 6) return;
    double s = 0;
    for (int b = 0; b < blocks; b++) s += partial[6 * b + m];
    total[m] = s;
}

// NematicOrder::calculateEigenvalues (core/observables/NematicOrder.cpp:30-55): Cardano's formula for the symmetric Q
static void eigenvaluesOfSymmetric(const double Q[9], double out[3]) {
    const double a11 = Q[0], a12 = Q[1], a13 = Q[2], a22 = Q[4], a23 = Q[5], a33 = Q[8];
    const double b = a11 + a22 + a33;
    const double c = a11 * a22 + a11 * a33 + a22 * a33 - a12 * a12 - a13 * a13 - a23 * a23;
    const double d = a11 * a23 * a23 + a22 * a13 * a13 + a33 * a12 * a12 - a11 * a22 * a33 - 2 * a12 * a13 * a23;
    const double p = b * b - 3 * c;
    if (!(p > 0)) throw rp_error(RP_ERR_STATE, "NematicOrder: Assert(p > 0) failed (isotropic Q tensor)");
    const double sqrtP = std::sqrt(p);
    const double q = 2 * b * b * b - 9 * b * c - 27 * d;
    double ratio = q / 2 / (sqrtP * sqrtP * sqrtP);
    if (std::abs(ratio) > 1) ratio = ratio < 0 ? -1 : 1;
    const double delta = std::acos(ratio);
    out[0] = 1. / 3 * (b + 2 * sqrtP * std::cos(delta / 3));
    out[1] = 1. / 3 * (b + 2 * sqrtP * std::cos((delta + 2 * M_PI) / 3));
    out[2] = 1. / 3 * (b + 2 * sqrtP * std::cos((delta - 2 * M_PI) / 3));
}

static void nematicOrderStatic(Handle *h, const double axis[3], double qOut[9], double *p2Out) {
    if (h->n > RP_NEMATIC_STATIC_MAX_PARTICLES)
        throw rp_error(RP_ERR_PRECONDITION,
                       "rp_nematic_order_static: " + std::to_string(h->n) + " particles exceed the static scratch "
                       "array's limit of " + std::to_string(RP_NEMATIC_STATIC_MAX_PARTICLES) +
                       " (RP_NEMATIC_STATIC_MAX_PARTICLES); use rp_nematic_order");
    const int n = static_cast<int>(h->n);
    const int blocks = (n + NA_CHUNK - 1) / NA_CHUNK;
    double *scratch = scratchFor(h->device);
    double *total = scratch + 6 * static_cast<size_t>(NA_MAX_BLOCKS);
    k_q_partial_static<<<blocks, NA_THREADS, 0, h->stream>>>(n, h->cap, h->orient, v3(axis[0], axis[1], axis[2]),
                                                             scratch);
    k_q_finish_static<<<1, 32, 0, h->stream>>>(blocks, scratch, total);
    h->launches += 2;
    double s[6];
    RP_CUDA(cudaMemcpyAsync(s, total, sizeof(s), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_CUDA(cudaGetLastError());
    const double N = static_cast<double>(h->n);
    double Q[9] = {s[0] / N, s[1] / N, s[2] / N, s[1] / N, s[3] / N, s[4] / N, s[2] / N, s[4] / N, s[5] / N};
    // QTensor = 0.5 * (3. * QTensor - identity)   (NematicOrder.cpp:9-28)
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) Q[3 * i + j] = 0.5 * (3. * Q[3 * i + j] - (i == j ? 1.0 : 0.0));
    double ev[3];
    eigenvaluesOfSymmetric(Q, ev);
    double p2 = ev[0];
    for (int i = 1; i < 3; i++) if (std::abs(p2) < std::abs(ev[i])) p2 = ev[i];      // highest magnitude
    if (qOut) for (int i = 0; i < 9; i++) qOut[i] = Q[i];
    if (p2Out) *p2Out = p2;
}

}  // namespace rp

extern "C" rp_status rp_nematic_order_static(rp_handle *h, const double primary_axis[3], double q_out[9],
                                             double *p2_out) {
    if (!h) { rp::tlsError = "null handle"; return RP_ERR_PRECONDITION; }
    try {
        if (cudaSetDevice(h->device) != cudaSuccess) throw rp::rp_error(RP_ERR_CUDA, "cudaSetDevice failed");
        RP_EXPECT(h->n > 0 && primary_axis != nullptr, "packing must be set, primary_axis != NULL");
        rp::nematicOrderStatic(h, primary_axis, q_out, p2_out);
    } catch (const rp::rp_error &e) {
        h->error = e.what();
        return static_cast<rp_status>(e.status);
    } catch (const std::exception &e) {
        h->error = e.what();
        return RP_ERR_STATE;
    }
    return RP_OK;
}
