// rp_api.cu — the C ABI of librampack_b200.so (include/rampack_b200.h): handle lifetime, state upload/download and
// thin forwarding to the kernels.  No exception crosses the boundary: every entry point returns an rp_status.
#include "rp_internal.cuh"
#include "rp_philox.cuh"
#include <cstring>
#include "rp_shard.cuh"
#include <limits>

namespace rp {

thread_local std::string tlsError;

// Flatten an rp_interaction_desc into the device parameter block.  Every threshold is computed with the expression
// the reference evaluates per call (citations in rp_pair.cuh) so that the compared doubles are bit-identical.
static void fillInter(const rp_interaction_desc &d, Inter &I, const XcProgram *xc) {
    std::memset(&I, 0, sizeof(I));
    I.hard = d.hard_kind;
    I.soft = d.soft_kind;
    RP_EXPECT(d.n_centres >= 0 && d.n_centres <= RP_MAX_CENTRES, "0 <= n_centres <= RP_MAX_CENTRES");
    RP_EXPECT(I.hard != RP_HARD_NONE || I.soft != RP_SOFT_NONE, "interaction has neither hard nor soft part");
    I.k = d.n_centres;
    I.kk = d.n_centres > 0 ? d.n_centres : 1;
    double hardRange = 0, softRange = 0;
    bool hardConvex = false;
    switch (d.hard_kind) {
        case RP_HARD_NONE:
            break;
        case RP_HARD_SPHERE:
            RP_EXPECT(d.radius > 0, "radius > 0");
            RP_EXPECT(d.n_centres == 0, "sphere has no interaction centres");
            I.sphereThr2 = (2 * d.radius) * (2 * d.radius);      // std::pow(2*r, 2)
            I.radius = d.radius;
            hardRange = 2 * d.radius;
            hardConvex = true;
            break;
        case RP_HARD_SPHEROCYLINDER:
            RP_EXPECT(d.radius > 0 && d.length >= 0, "radius > 0 && length >= 0");
            RP_EXPECT(d.n_centres == 0, "spherocylinder has no interaction centres");
            I.scInner2 = 4 * d.radius * d.radius;
            I.radius = d.radius;
            I.scOuter2 = (2 * d.radius + d.length) * (2 * d.radius + d.length);
            I.scHalfMinus = 0.5 * -1 * d.length;
            I.scHalfPlus = 0.5 * 1 * d.length;
            hardRange = 2 * d.radius + d.length;
            hardConvex = true;
            break;
        case RP_HARD_POLYSPHERE: {
            RP_EXPECT(d.n_centres >= 1, "polysphere needs at least one centre");
            double maxR = 0;
            for (int i = 0; i < d.n_centres; i++) {
                RP_EXPECT(d.centre_radius[i] > 0, "radius > 0");
                if (d.centre_radius[i] > maxR) maxR = d.centre_radius[i];
            }
            hardRange = 2 * maxR;
            break;
        }
        case RP_HARD_POLYSPHEROCYLINDER: {
            RP_EXPECT(d.n_centres >= 1, "polyspherocylinder needs at least one centre");
            double maxC = 0;
            for (int i = 0; i < d.n_centres; i++) {
                RP_EXPECT(d.centre_radius[i] > 0, "radius > 0");
                V3 ha = v3(d.centre_half_axis[i][0], d.centre_half_axis[i][1], d.centre_half_axis[i][2]);
                I.ccircum[i] = d.centre_radius[i] + std::sqrt(norm2(ha));
                if (I.ccircum[i] > maxC) maxC = I.ccircum[i];
            }
            hardRange = 2 * maxC;
            break;
        }
        case RP_HARD_XENOCOLLIDE: {
            // XenoCollideTraits::getRangeRadius (XenoCollideTraits.h:188-202): 2 * max circumsphere radius
            RP_EXPECT(xc != nullptr, "RP_HARD_XENOCOLLIDE needs desc.xc");
            double maxRadius = 0;
            for (int i = 0; i < std::max(d.n_centres, 1); i++)
                maxRadius = std::max(maxRadius, xc->nodes[xc->root[i]].circum);
            hardRange = 2 * maxRadius;
            hardConvex = true;
            break;
        }
        default:
            throw rp_error(RP_ERR_UNSUPPORTED, "hard interaction kind not implemented on the device");
    }
    switch (d.soft_kind) {
        case RP_SOFT_NONE:
            break;
        case RP_SOFT_LJ:
            RP_EXPECT(d.epsilon > 0 && d.sigma > 0, "epsilon > 0 && sigma > 0");
            softRange = 3 * d.sigma;
            break;
        case RP_SOFT_WCA: {
            RP_EXPECT(d.epsilon > 0 && d.sigma > 0, "epsilon > 0 && sigma > 0");
            double s = d.sigma * std::pow(2., 1. / 6);
            I.wcaCut2 = s * s;
            softRange = s;
            break;
        }
        case RP_SOFT_SIC:
            RP_EXPECT(d.epsilon != 0 && d.sigma > 0, "epsilon != 0 && sigma > 0");
            softRange = d.sigma;
            break;
        default:
            throw rp_error(RP_ERR_UNSUPPORTED, "soft interaction kind not implemented on the device");
    }
    I.sigma2 = d.sigma * d.sigma;
    I.fourEps = 4 * d.epsilon;
    I.eps = d.epsilon;
    for (int i = 0; i < d.n_centres; i++) {
        for (int c = 0; c < 3; c++) {
            I.cpos[i][c] = d.centres[i][c];
            I.chalf[i][c] = d.centre_half_axis[i][c];
        }
        I.cradius[i] = d.centre_radius[i];
    }
    // CompoundInteraction (core/interactions/CompoundInteraction.cpp:5-66): max of the ranges; convex iff the only
    // hard part is convex
    I.range = std::max(hardRange, softRange);
    I.convex = hardConvex;
    // Interaction::getTotalRangeRadius (core/Interaction.cpp:55-70)
    double maxRadius2 = 0;
    for (int i = 0; i < d.n_centres; i++) {
        double r2 = norm2(v3(d.centres[i][0], d.centres[i][1], d.centres[i][2]));
        if (r2 > maxRadius2) maxRadius2 = r2;
    }
    I.totalRange = d.n_centres == 0 ? I.range : 2 * std::sqrt(maxRadius2) + I.range;
}

Box makeBox(const double m[9]) {
    Box b;
    for (int i = 0; i < 9; i++) b.m.a[i] = m[i];
    b.inv = inverse(b.m);
    return b;
}

// ---------------------------------------------------------------------------------------------------------------
// lattice population on the device
// ---------------------------------------------------------------------------------------------------------------
// SerialPopulator::populateLattice over a regular Lattice (core/lattice/SerialPopulator.cpp:9-40, Lattice.cpp:61-81):
// shape s = molecule (s % m) of the (s / m)-th cell in the populator's axis order; position =
// cellBox.relativeToAbsolute({i, j, k} + relative position) — the same two operations in the same order, so the
// result equals the reference's bit for bit.  flag[0] |= 1 when a shape ends up outside the lattice box.
struct LatticeArgs {
    Box cell, lattice;
    int dims[3], order[3];
    int m;
};
__global__ void k_populate_lattice(int n, size_t cap, LatticeArgs L, const double *__restrict__ molecules,
                                   double *__restrict__ pos, double *__restrict__ orient, int *__restrict__ flag)
{
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    int q = s / L.m, j = s - q * L.m;
    int i[3];
    i[L.order[2]] = q % L.dims[L.order[2]]; q /= L.dims[L.order[2]];
    i[L.order[1]] = q % L.dims[L.order[1]]; q /= L.dims[L.order[1]];
    i[L.order[0]] = q;
    const double *mol = molecules + 12 * static_cast<size_t>(j);
    V3 rel = v3(static_cast<double>(i[0]), static_cast<double>(i[1]), static_cast<double>(i[2])) + v3(mol[0], mol[1], mol[2]);
    V3 p = relToAbs(L.cell, rel);
    pos[s] = p.x; pos[cap + s] = p.y; pos[2 * cap + s] = p.z;
    #pragma unroll
    for (int c = 0; c < 9; c++) orient[c * cap + s] = mol[3 + c];
    V3 r = absToRel(L.lattice, p);
    if (r.x < 0 || r.x >= 1 || r.y < 0 || r.y >= 1 || r.z < 0 || r.z >= 1) atomicOr(flag, 1);
}

// RotationRandomizingTransformer::rotateRandomly (core/lattice/RotationRandomizingTransformer.cpp:30-56) for every
// particle, from per-particle Philox streams instead of one sequential std::mt19937 (same distributions):
//   mode 0  uniformly random rotation: Matrix<3,3>::rotation(U(0, 2pi), asin(U(-1, 1)), U(0, 2pi)) (Tait-Bryan angles,
//           geometry/Matrix.tpp:411-436)
//   mode 1  rotation about the fixed (normalised) axis by U(0, 2pi) (Matrix.tpp:437-449)
//   mode 2  rotation about the particle's own axis orientation * bodyAxis (ShapeGeometry::getAxis) by U(0, 2pi)
// followed by Shape::rotate (core/Shape.cpp:14-16): orientation = rotation * orientation.
__global__ void k_randomize_rotations(int n, size_t cap, unsigned long long seed, int mode, V3 axis,
                                      double *__restrict__ orient)
{
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    Rng rng;
    rng.init(seed, static_cast<unsigned int>(p), 0x4C415454u, 0u);
    const double TWO_PI = 2 * 3.14159265358979323846;
    M33 O;
    #pragma unroll
    for (int q = 0; q < 9; q++) O.a[q] = orient[q * cap + p];
    M33 R;
    if (mode == 0) {
        const double ax = rng.uniform(0, TWO_PI), ay = asin(rng.uniform(-1, 1)), az = rng.uniform(0, TWO_PI);
        const double sx = sin(ax), sy = sin(ay), sz = sin(az), cx = cos(ax), cy = cos(ay), cz = cos(az);
        R.a[0] = cy * cz; R.a[1] = sx * sy * cz - cx * sz; R.a[2] = cx * sy * cz + sx * sz;
        R.a[3] = cy * sz; R.a[4] = sx * sy * sz + cx * cz; R.a[5] = cx * sy * sz - sx * cz;
        R.a[6] = -sy;     R.a[7] = sx * cy;                R.a[8] = cx * cy;
    } else {
        V3 a = axis;
        if (mode == 2) a = mul(O, axis);
        const double nrm = sqrt(norm2(a));
        a = v3(a.x / nrm, a.y / nrm, a.z / nrm);
        const double angle = rng.uniform(0, TWO_PI);
        const double sina = sin(angle), cosa = cos(angle);
        M33 K = {{0, -a.z, a.y, a.z, 0, -a.x, -a.y, a.x, 0}};
        M33 sK;
        for (int q = 0; q < 9; q++) sK.a[q] = (1 - cosa) * K.a[q];
        M33 KK = mul(sK, K);
        for (int q = 0; q < 9; q++) R.a[q] = (((q % 4 == 0) ? 1.0 : 0.0) + K.a[q] * sina) + KK.a[q];
    }
    M33 N = mul(R, O);
    #pragma unroll
    for (int q = 0; q < 9; q++) orient[q * cap + p] = N.a[q];
}

// AoS (N x 12, reference Shape layout) <-> SoA with stride cap
// Packing::areShapesWithinBox (core/Packing.cpp:1067-1073) on the staged upload: flag[0] |= 1 when a relative
// coordinate lies outside [0, 1)
__global__ void k_check_within_box(int n, Box box, const double *__restrict__ aos, int *__restrict__ flag) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    V3 rel = absToRel(box, v3(aos[12 * static_cast<size_t>(p)], aos[12 * static_cast<size_t>(p) + 1],
                              aos[12 * static_cast<size_t>(p) + 2]));
    if (rel.x < 0 || rel.x >= 1 || rel.y < 0 || rel.y >= 1 || rel.z < 0 || rel.z >= 1) atomicOr(flag, 1);
}
// (the unpack is skipped when the check failed: a rejected upload leaves the previous state untouched)
__global__ void k_unpack_shapes(int n, size_t cap, const double *__restrict__ aos, double *__restrict__ pos,
                                double *__restrict__ orient, const int *__restrict__ flag)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 12 || *flag) return;
    int p = i / 12, c = i % 12;
    double v = aos[i];
    if (c < 3) pos[c * cap + p] = v;
    else orient[(c - 3) * cap + p] = v;
}
__global__ void k_pack_shapes(int n, size_t cap, const double *__restrict__ pos, const double *__restrict__ orient,
                              double *__restrict__ aos)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * 12) return;
    int p = i / 12, c = i % 12;
    aos[i] = (c < 3) ? pos[c * cap + p] : orient[(c - 3) * cap + p];
}

static void ensureStage(Handle *h, size_t bytes) {
    if (bytes > h->dStageCap) {
        if (h->dStage) cudaFree(h->dStage);
        h->dStage = nullptr;
        RP_CUDA(cudaMalloc(&h->dStage, bytes + 64));
        h->dStageCap = bytes + 64;
    }
}

// Packing::fixRotationMatrix (core/Packing.cpp:1214-1219): three iterations of R <- 1.5 R - 0.5 R R^T R, in the
// reference's operator order ((0.5 R) R^T) R.  backup != nullptr keeps the old matrix (9 doubles per particle).
__global__ void k_fix_orientations(int n, size_t cap, double *__restrict__ orient, double *__restrict__ backup) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    M33 R;
    for (int q = 0; q < 9; q++) R.a[q] = orient[q * cap + p];
    if (backup) for (int q = 0; q < 9; q++) backup[9 * static_cast<size_t>(p) + q] = R.a[q];
    for (int it = 0; it < 3; it++) {
        M33 half, Rt;
        for (int q = 0; q < 9; q++) half.a[q] = 0.5 * R.a[q];
        for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) Rt.a[3 * i + j] = R.a[3 * j + i];
        M33 corr = mul(mul(half, Rt), R);
        for (int q = 0; q < 9; q++) R.a[q] = 1.5 * R.a[q] - corr.a[q];
    }
    for (int q = 0; q < 9; q++) orient[q * cap + p] = R.a[q];
}
// withdraws the fix of every particle (flags == nullptr) or of the flagged ones; counts the withdrawals
__global__ void k_restore_orientations(int n, size_t cap, double *__restrict__ orient,
                                       const double *__restrict__ backup, const unsigned char *__restrict__ flags,
                                       unsigned char *__restrict__ withdrawn, unsigned long long *__restrict__ count) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    if (flags && (!flags[p] || withdrawn[p])) return;
    for (int q = 0; q < 9; q++) orient[q * cap + p] = backup[9 * static_cast<size_t>(p) + q];
    if (flags) { withdrawn[p] = 1; atomicAdd(count, 1ull); }
}

static void setupForInteraction(Handle *h) {
    // Packing::setupForInteraction (core/Packing.cpp:931-950)
    h->interactionRange = h->inter.range;
    if (h->n == 0) return;
    if (h->inter.k > 0) {
        size_t need = h->cap * h->inter.k;
        if (need > h->centreCap || !h->crel) {
            if (h->crel) cudaFree(h->crel);
            if (h->cabs) cudaFree(h->cabs);
            h->crel = h->cabs = nullptr;
            RP_CUDA(cudaMalloc(&h->crel, 3 * need * sizeof(double)));
            RP_CUDA(cudaMalloc(&h->cabs, 3 * need * sizeof(double)));
            h->centreCap = need;
        }
        initCentres(h);
        recalcAbsoluteCentres(h);
    }
    h->gridDirty = true;
    h->hasTrial = false;
    h->hasLastScaling = false;
    rebuildReferenceGrid(h);
    if (h->overlapCounting)
        h->numOverlaps = countTotalOverlaps(h, false);
}

__global__ void k_zero_words(unsigned int *__restrict__ p, size_t words) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < words;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        p[i] = 0u;
}

__global__ void k_mailbox(unsigned int *__restrict__ box, const unsigned int *__restrict__ a, int wordsA,
                          const unsigned int *__restrict__ b, int wordsB) {
    for (int i = threadIdx.x; i < wordsA; i += blockDim.x) box[i] = a[i];
    for (int i = threadIdx.x; i < wordsB; i += blockDim.x) box[wordsA + i] = b[i];
    __threadfence_system();
}

__global__ void k_copy_words(unsigned long long *__restrict__ dst, const unsigned long long *__restrict__ src,
                             size_t words) {
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < words;
         i += static_cast<size_t>(gridDim.x) * blockDim.x)
        dst[i] = src[i];
}

void copyAsync(Handle *h, void *dst, const void *src, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return;
    if ((bytes & 7) || (reinterpret_cast<uintptr_t>(dst) & 7) || (reinterpret_cast<uintptr_t>(src) & 7))
        throw rp_error(RP_ERR_STATE, "copyAsync: not a whole number of aligned 8-byte words");
    const size_t words = bytes / 8;
    const int blocks = static_cast<int>(std::min<size_t>((words + 255) / 256, 1184));
    k_copy_words<<<blocks, 256, 0, st>>>(static_cast<unsigned long long *>(dst),
                                         static_cast<const unsigned long long *>(src), words);
    h->launches++;
}

void zeroAsync(Handle *h, void *p, size_t bytes, cudaStream_t st) {
    if (bytes == 0) return;
    if ((bytes & 3) || (reinterpret_cast<uintptr_t>(p) & 3))
        throw rp_error(RP_ERR_STATE, "zeroAsync: not a whole number of aligned words");
    const size_t words = bytes / 4;
    const int blocks = static_cast<int>(std::min<size_t>((words + 255) / 256, 592));
    k_zero_words<<<blocks, 256, 0, st>>>(static_cast<unsigned int *>(p), words);
    h->launches++;
}

void readbackSync(Handle *h, cudaStream_t st, void *dst0, const void *dsrc0, size_t bytes0, void *dst1,
                  const void *dsrc1, size_t bytes1) {
    if ((bytes0 & 3) || (bytes1 & 3) || bytes0 + bytes1 > MAILBOX_BYTES || !h->mailbox)
        throw rp_error(RP_ERR_STATE, "readbackSync: the mailbox takes at most 1 KB of whole words");
    k_mailbox<<<1, 64, 0, st>>>(h->mailbox, static_cast<const unsigned int *>(dsrc0), static_cast<int>(bytes0 / 4),
                                static_cast<const unsigned int *>(dsrc1), static_cast<int>(bytes1 / 4));
    h->launches++;
    RP_CUDA(cudaStreamSynchronize(st));
    RP_CUDA(cudaGetLastError());
    std::memcpy(dst0, h->mailbox, bytes0);
    if (bytes1) std::memcpy(dst1, reinterpret_cast<const char *>(h->mailbox) + bytes0, bytes1);
}

}  // namespace rp

using namespace rp;

#define RP_API_BEGIN(h) if (!(h)) { rp::tlsError = "null handle"; return RP_ERR_PRECONDITION; } \
    try { if (cudaSetDevice((h)->device) != cudaSuccess) throw rp::rp_error(RP_ERR_CUDA, "cudaSetDevice failed");
#define RP_API_END(h) } catch (const rp::rp_error &e) { (h)->error = e.what(); return static_cast<rp_status>(e.status); } \
    catch (const std::exception &e) { (h)->error = e.what(); return RP_ERR_STATE; } return RP_OK;

extern "C" {

const char *rp_version(void) { return "rampack_b200 0.1.0 (sm_100a)"; }

const char *rp_last_error(const rp_handle *h) { return h ? h->error.c_str() : rp::tlsError.c_str(); }

rp_status rp_create(const rp_config *config, rp_handle **out) {
    if (!config || !out) { rp::tlsError = "null argument"; return RP_ERR_PRECONDITION; }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        rp::tlsError = std::string("no usable CUDA device (there is no CPU fallback): ")
                       + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return RP_ERR_CUDA;
    }
    if (config->device < 0 || config->device >= count) { rp::tlsError = "bad device ordinal"; return RP_ERR_CUDA; }
    auto *h = new rp_handle;
    try {
        h->device = config->device;
        h->seed = config->seed;
        RP_CUDA(cudaSetDevice(h->device));
        if (config->cuda_stream) {
            h->stream = static_cast<cudaStream_t>(config->cuda_stream);
        } else {
            RP_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
            h->ownStream = true;
        }
        RP_CUDA(cudaMalloc(&h->dInter, sizeof(Inter)));
        RP_CUDA(cudaMalloc(&h->dCounters, 32 * sizeof(unsigned long long)));
        RP_CUDA(cudaMalloc(&h->dFlags, 16 * sizeof(int)));
        RP_CUDA(cudaMemset(h->dFlags, 0, 16 * sizeof(int)));
        RP_CUDA(cudaMallocHost(&h->sweepMaxPinned, 4 * 4 * sizeof(int)));
        RP_CUDA(cudaMallocHost(&h->mailbox, MAILBOX_BYTES));
        for (auto &ev : h->sweepMaxEvent) RP_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        for (auto &t : h->timing) for (auto &ev : t.e) RP_CUDA(cudaEventCreate(&ev));
        RP_CUDA(cudaMalloc(&h->dMoveOut, 8 * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->tmpState, (12 + 6 * RP_MAX_CENTRES) * sizeof(double)));
        for (auto &ev : h->ev) RP_CUDA(cudaEventCreate(&ev));
    } catch (const std::exception &ex) {
        rp::tlsError = ex.what();
        delete h;
        return RP_ERR_CUDA;
    }
    *out = h;
    return RP_OK;
}

void rp_destroy(rp_handle *h) {
    if (!h) return;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    for (void *p : {(void *) h->dInter, (void *) h->pos, (void *) h->orient, (void *) h->crel, (void *) h->cabs,
                    (void *) h->lastPos, (void *) h->lastCabs, (void *) h->tmpState, (void *) h->dCounters,
                    (void *) h->dFlags, (void *) h->dCellEnergy, (void *) h->dMoveOut, h->dStage,
                    (void *) h->sweepMoved, (void *) h->trialRec, (void *) h->trialOrient, h->trialMeta, h->trialRec32,
                    (void *) h->genRec, (void *) h->genTrialRec, (void *) h->genTrialCrel, (void *) h->genTrialU,
                    (void *) h->genCull, (void *) h->genTrialCull})
        if (p) cudaFree(p);
    shardDestroy(h);
    if (h->dXc) cudaFree(h->dXc);
    delete h->xcHost;
    if (h->stage) cudaFreeHost(h->stage);
    freeGrid(h->grid);
    freeGrid(h->tempGrid);
    freeGrid(h->sgrid);
    freeGrid(h->ogrid);
    for (auto &ev : h->ev) if (ev) cudaEventDestroy(ev);
    for (auto &ev : h->sweepMaxEvent) if (ev) cudaEventDestroy(ev);
    for (auto &t : h->timing) for (auto &ev : t.e) if (ev) cudaEventDestroy(ev);
    if (h->sweepMaxPinned) cudaFreeHost(h->sweepMaxPinned);
    if (h->mailbox) cudaFreeHost(h->mailbox);
    if (h->snapStream) {
        cudaStreamSynchronize(h->snapStream);
        cudaStreamDestroy(h->snapStream);
        cudaEventDestroy(h->snapReady);
        cudaEventDestroy(h->snapDone);
    }
    if (h->dSnap) cudaFree(h->dSnap);
    if (h->ownStream && h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

rp_status rp_set_interaction(rp_handle *h, const rp_interaction_desc *desc) {
    RP_API_BEGIN(h)
    RP_EXPECT(desc != nullptr, "desc != NULL");
    Inter I;
    if (desc->hard_kind == RP_HARD_XENOCOLLIDE) {
        RP_EXPECT(desc->xc != nullptr, "RP_HARD_XENOCOLLIDE needs desc.xc");
        if (!h->xcHost) h->xcHost = new XcProgram;
        buildXcProgram(*desc->xc, desc->n_centres, *h->xcHost);
        if (!h->dXc) RP_CUDA(cudaMalloc(&h->dXc, sizeof(XcProgram)));
        RP_CUDA(cudaMemcpyAsync(h->dXc, h->xcHost, sizeof(XcProgram), cudaMemcpyHostToDevice, h->stream));
    }
    fillInter(*desc, I, h->xcHost);
    I.xc = h->dXc;
    RP_EXPECT(I.range > 0, "interaction.getRangeRadius() > 0");
    h->inter = I;
    RP_CUDA(cudaMemcpyAsync(h->dInter, &h->inter, sizeof(Inter), cudaMemcpyHostToDevice, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    h->hasInteraction = true;
    setupForInteraction(h);
    RP_API_END(h)
}

rp_status rp_upload(rp_handle *h, const double box[9], const double *shapes, uint64_t n) {
    RP_API_BEGIN(h)
    RP_EXPECT(box && shapes, "box and shapes must not be NULL");
    RP_EXPECT(n > 0, "!shapes.empty()");
    RP_EXPECT(n < (1ull << 27), "at most 2^27 particles per handle");
    Box b = makeBox(box);
    RP_EXPECT(boxVolume(b) != 0, "newBox.getVolume() != 0");
    if (n > h->cap || !h->pos) {
        if (h->pos) cudaFree(h->pos);
        if (h->orient) cudaFree(h->orient);
        h->pos = h->orient = nullptr;
        size_t cap = n + 16;
        RP_CUDA(cudaMalloc(&h->pos, 3 * cap * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->orient, 9 * cap * sizeof(double)));
        h->cap = cap;
        h->n = 0;          // the previous state went with the old buffers
    }
    ensureStage(h, n * 12 * sizeof(double));
    RP_CUDA(cudaMemcpyAsync(h->dStage, shapes, n * 12 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    zeroAsync(h, h->dFlags, sizeof(int), h->stream);
    k_check_within_box<<<static_cast<int>((n + 255) / 256), 256, 0, h->stream>>>(
        static_cast<int>(n), b, static_cast<const double *>(h->dStage), h->dFlags);
    int total = static_cast<int>(n * 12);
    k_unpack_shapes<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(n), h->cap,
                                                                static_cast<const double *>(h->dStage), h->pos,
                                                                h->orient, h->dFlags);
    h->launches += 2;
    int outside = 0;
    readbackSync(h, h->stream, &outside, h->dFlags, sizeof(int));
    RP_EXPECT(!outside, "areShapesWithinBox(shapes, box)");
    h->n = n;
    h->box = b;
    h->gridDirty = true;
    h->hasTrial = false;
    h->hasLastScaling = false;
    h->sweepCapOld = h->sweepCapOwn = 0;      // a new packing: the next queued sweep sizes its staging from its own grid
    if (h->hasInteraction) setupForInteraction(h);
    RP_API_END(h)
}

rp_status rp_upload_slice(rp_handle *h, const double box[9], const double *shapes, uint64_t first, uint64_t count,
                          uint64_t n) {
    RP_API_BEGIN(h)
    RP_EXPECT(box != nullptr, "box must not be NULL");
    RP_EXPECT(n > 0 && n < (1ull << 27), "0 < n < 2^27");
    const int P = shardActive(h) ? h->shard->nranks : 1, r = shardActive(h) ? h->shard->rank : 0;
    unsigned long long wantFirst = 0, wantCount = n;
    if (P > 1) shardSlice(n, P, r, &wantFirst, &wantCount);
    RP_EXPECT(first == wantFirst && count == wantCount,
              "slice must be [n * rank / n_ranks, n * (rank + 1) / n_ranks) of the packing");
    RP_EXPECT(shapes != nullptr || count == 0, "shapes must not be NULL");
    Box b = makeBox(box);
    RP_EXPECT(boxVolume(b) != 0, "newBox.getVolume() != 0");
    if (n > h->cap || !h->pos) {
        if (h->pos) cudaFree(h->pos);
        if (h->orient) cudaFree(h->orient);
        h->pos = h->orient = nullptr;
        size_t cap = n + 16;
        RP_CUDA(cudaMalloc(&h->pos, 3 * cap * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->orient, 9 * cap * sizeof(double)));
        h->cap = cap;
        h->n = 0;
    }
    zeroAsync(h, h->dFlags, sizeof(int), h->stream);
    if (count > 0) {
        ensureStage(h, count * 12 * sizeof(double));
        RP_CUDA(cudaMemcpyAsync(h->dStage, shapes, count * 12 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
        k_check_within_box<<<static_cast<int>((count + 255) / 256), 256, 0, h->stream>>>(
            static_cast<int>(count), b, static_cast<const double *>(h->dStage), h->dFlags);
        int total = static_cast<int>(count * 12);
        k_unpack_shapes<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(count), h->cap,
                                                                    static_cast<const double *>(h->dStage),
                                                                    h->pos + first, h->orient + first, h->dFlags);
        h->launches += 2;
    }
    if (P > 1) {
        shardAllReduceMaxInt(h, h->dFlags);       // a shape outside the box on any rank fails the call on every rank
        shardAllGatherState(h, n);
    }
    int outside = 0;
    readbackSync(h, h->stream, &outside, h->dFlags, sizeof(int));
    RP_EXPECT(!outside, "areShapesWithinBox(shapes, box)");
    h->n = n;
    h->box = b;
    h->gridDirty = true;
    h->hasTrial = false;
    h->hasLastScaling = false;
    h->sweepCapOld = h->sweepCapOwn = 0;
    if (h->hasInteraction) setupForInteraction(h);
    RP_API_END(h)
}

rp_status rp_populate_lattice(rp_handle *h, const double cell_box[9], const double *cell_molecules, uint64_t m,
                              const uint32_t n_cells[3], const int axis_order[3], uint64_t n_shapes) {
    RP_API_BEGIN(h)
    RP_EXPECT(cell_box && cell_molecules && n_cells && axis_order, "arguments must not be NULL");
    RP_EXPECT(m > 0, "!molecules.empty()");
    RP_EXPECT(n_cells[0] > 0 && n_cells[1] > 0 && n_cells[2] > 0, "dimensions must be > 0");
    bool seen[3] = {false, false, false};
    for (int i = 0; i < 3; i++) {
        RP_EXPECT(axis_order[i] >= 0 && axis_order[i] < 3 && !seen[axis_order[i]], "axis order must be a permutation of xyz");
        seen[axis_order[i]] = true;
    }
    const unsigned long long size = static_cast<unsigned long long>(n_cells[0]) * n_cells[1] * n_cells[2] * m;
    RP_EXPECT(n_shapes > 0 && size >= n_shapes, "lattice.size() >= numOfShapes");
    RP_EXPECT(n_shapes < (1ull << 27), "at most 2^27 particles per handle");
    LatticeArgs L;
    L.cell = makeBox(cell_box);
    RP_EXPECT(boxVolume(L.cell) != 0, "cell box volume != 0");
    // Lattice::getLatticeBox (Lattice.cpp:43-48): every side (column) of the cell box times the number of cells
    double lb[9];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++) lb[3 * r + c] = cell_box[3 * r + c] * static_cast<double>(n_cells[c]);
    L.lattice = makeBox(lb);
    for (int i = 0; i < 3; i++) { L.dims[i] = static_cast<int>(n_cells[i]); L.order[i] = axis_order[i]; }
    L.m = static_cast<int>(m);
    const uint64_t n = n_shapes;
    if (n > h->cap || !h->pos) {
        if (h->pos) cudaFree(h->pos);
        if (h->orient) cudaFree(h->orient);
        h->pos = h->orient = nullptr;
        size_t cap = n + 16;
        RP_CUDA(cudaMalloc(&h->pos, 3 * cap * sizeof(double)));
        RP_CUDA(cudaMalloc(&h->orient, 9 * cap * sizeof(double)));
        h->cap = cap;
    }
    h->n = 0;          // the previous state is overwritten in place: a failed call leaves an empty packing
    ensureStage(h, m * 12 * sizeof(double));
    RP_CUDA(cudaMemcpyAsync(h->dStage, cell_molecules, m * 12 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    zeroAsync(h, h->dFlags, sizeof(int), h->stream);
    k_populate_lattice<<<static_cast<int>((n + 255) / 256), 256, 0, h->stream>>>(
        static_cast<int>(n), h->cap, L, static_cast<const double *>(h->dStage), h->pos, h->orient, h->dFlags);
    h->launches++;
    int outside = 0;
    readbackSync(h, h->stream, &outside, h->dFlags, sizeof(int));
    RP_EXPECT(!outside, "areShapesWithinBox(shapes, box)");
    h->n = n;
    h->box = L.lattice;
    h->gridDirty = true;
    h->hasTrial = false;
    h->hasLastScaling = false;
    h->sweepCapOld = h->sweepCapOwn = 0;
    if (h->hasInteraction) setupForInteraction(h);
    RP_API_END(h)
}

rp_status rp_randomize_rotations(rp_handle *h, uint64_t seed, int mode, const double axis[3]) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->n > 0, "the packing must not be empty");
    RP_EXPECT(mode >= 0 && mode <= 2, "mode: 0 random, 1 fixed axis, 2 body axis");
    V3 a = v3(0, 0, 1);
    if (mode != 0) {
        RP_EXPECT(axis != nullptr, "axis must not be NULL");
        a = v3(axis[0], axis[1], axis[2]);
        RP_EXPECT(norm2(a) > 1e-20, "vectorAxis.norm2() > 1e-20");
    }
    const int n = static_cast<int>(h->n);
    k_randomize_rotations<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, seed, mode, a, h->orient);
    h->launches++;
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_CUDA(cudaGetLastError());
    h->gridDirty = true;
    h->hasTrial = false;
    h->hasLastScaling = false;
    if (h->hasInteraction) setupForInteraction(h);
    RP_API_END(h)
}

rp_status rp_download(rp_handle *h, double box_out[9], double *shapes_out, uint64_t n) {
    RP_API_BEGIN(h)
    RP_EXPECT(n == h->n, "n == size()");
    if (box_out) for (int i = 0; i < 9; i++) box_out[i] = h->box.m.a[i];
    if (shapes_out && n > 0) {
        ensureStage(h, n * 12 * sizeof(double));
        int total = static_cast<int>(n * 12);
        k_pack_shapes<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(n), h->cap, h->pos, h->orient,
                                                                  static_cast<double *>(h->dStage));
        h->launches++;
        RP_CUDA(cudaMemcpyAsync(shapes_out, h->dStage, n * 12 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
        RP_CUDA(cudaStreamSynchronize(h->stream));
        RP_CUDA(cudaGetLastError());
    }
    RP_API_END(h)
}

rp_status rp_download_range(rp_handle *h, uint64_t first, uint64_t count, double *shapes_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(first <= h->n && count <= h->n - first, "first + count <= size()");
    RP_EXPECT(shapes_out != nullptr || count == 0, "shapes_out != NULL");
    if (count > 0) {
        ensureStage(h, count * 12 * sizeof(double));
        int total = static_cast<int>(count * 12);
        k_pack_shapes<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(count), h->cap, h->pos + first,
                                                                  h->orient + first, static_cast<double *>(h->dStage));
        h->launches++;
        RP_CUDA(cudaMemcpyAsync(shapes_out, h->dStage, count * 12 * sizeof(double), cudaMemcpyDeviceToHost,
                                h->stream));
        RP_CUDA(cudaStreamSynchronize(h->stream));
        RP_CUDA(cudaGetLastError());
    }
    RP_API_END(h)
}

rp_status rp_snapshot_async(rp_handle *h, double box_out[9], double *shapes_out, uint64_t n) {
    RP_API_BEGIN(h)
    RP_EXPECT(n == h->n, "n == size()");
    RP_EXPECT(shapes_out != nullptr || n == 0, "shapes_out != NULL");
    if (box_out) for (int i = 0; i < 9; i++) box_out[i] = h->box.m.a[i];
    if (n > 0) {
        if (!h->snapStream) {
            RP_CUDA(cudaStreamCreateWithFlags(&h->snapStream, cudaStreamNonBlocking));
            RP_CUDA(cudaEventCreateWithFlags(&h->snapReady, cudaEventDisableTiming));
            RP_CUDA(cudaEventCreateWithFlags(&h->snapDone, cudaEventDisableTiming));
        }
        const size_t bytes = n * 12 * sizeof(double);
        // the previous snapshot still reads the device-side copy: the pack below must come after it
        if (h->snapPending) RP_CUDA(cudaStreamWaitEvent(h->stream, h->snapDone, 0));
        if (bytes > h->dSnapCap) {
            if (h->snapPending) RP_CUDA(cudaEventSynchronize(h->snapDone));
            if (h->dSnap) cudaFree(h->dSnap);
            h->dSnap = nullptr;
            RP_CUDA(cudaMalloc(&h->dSnap, bytes + 64));
            h->dSnapCap = bytes + 64;
        }
        const int total = static_cast<int>(n * 12);
        k_pack_shapes<<<(total + 255) / 256, 256, 0, h->stream>>>(static_cast<int>(n), h->cap, h->pos, h->orient,
                                                                  static_cast<double *>(h->dSnap));
        h->launches++;
        RP_CUDA(cudaEventRecord(h->snapReady, h->stream));
        RP_CUDA(cudaStreamWaitEvent(h->snapStream, h->snapReady, 0));
        RP_CUDA(cudaMemcpyAsync(shapes_out, h->dSnap, bytes, cudaMemcpyDeviceToHost, h->snapStream));
        RP_CUDA(cudaEventRecord(h->snapDone, h->snapStream));
        h->snapPending = true;
    }
    RP_API_END(h)
}

rp_status rp_snapshot_wait(rp_handle *h) {
    RP_API_BEGIN(h)
    if (h->snapPending) {
        RP_CUDA(cudaEventSynchronize(h->snapDone));
        h->snapPending = false;
    }
    RP_API_END(h)
}

rp_status rp_snapshot_done(rp_handle *h, int32_t *done_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(done_out != nullptr, "done_out != NULL");
    *done_out = 1;
    if (h->snapPending) {
        cudaError_t e = cudaEventQuery(h->snapDone);
        if (e == cudaErrorNotReady) *done_out = 0;
        else {
            RP_CUDA(e);
            h->snapPending = false;
        }
    }
    RP_API_END(h)
}

rp_status rp_count_wall_overlaps(rp_handle *h, int early_exit, uint64_t *count_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && count_out != nullptr, "interaction must be set, count_out != NULL");
    *count_out = h->n > 0 ? countWallOverlaps(h, early_exit != 0) : 0;
    RP_API_END(h)
}

rp_status rp_particle_energy_fluctuations(rp_handle *h, double *fluctuations_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && h->n > 0 && fluctuations_out != nullptr, "packing and interaction must be set");
    *fluctuations_out = particleEnergyFluctuations(h);
    RP_API_END(h)
}

rp_status rp_nematic_order(rp_handle *h, const double primary_axis[3], double q_out[9], double *p2_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->n > 0 && primary_axis != nullptr, "packing must be set, primary_axis != NULL");
    nematicOrder(h, primary_axis, q_out, p2_out);
    RP_API_END(h)
}

rp_status rp_smectic_order(rp_handle *h, const int32_t hkl_range[3], const double focal_point[3], double *tau_out,
                           int32_t hkl_out[3], double k_out[3]) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->n > 0 && hkl_range != nullptr, "packing must be set, hkl_range != NULL");
    // SmecticOrder::SmecticOrder (core/observables/SmecticOrder.cpp:70-71)
    RP_EXPECT(hkl_range[0] >= 0 && hkl_range[1] >= 0 && hkl_range[2] >= 0, "hklRange[i] >= 0");
    RP_EXPECT(hkl_range[0] != 0 || hkl_range[1] != 0 || hkl_range[2] != 0, "any hklRange[i] != 0");
    RP_EXPECT(hkl_range[0] <= 32 && hkl_range[1] <= 32 && hkl_range[2] <= 32, "hklRange[i] <= 32");
    int range[3] = {hkl_range[0], hkl_range[1], hkl_range[2]}, hkl[3];
    smecticOrder(h, range, focal_point, tau_out, hkl, k_out);
    if (hkl_out) for (int i = 0; i < 3; i++) hkl_out[i] = hkl[i];
    RP_API_END(h)
}

rp_status rp_pair_density_correlation(rp_handle *h, double max_r, int32_t n_bins, const double focal_point[3],
                                      uint64_t *counts_out, double *g_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->n > 0, "packing must be set");
    // HistogramBuilder::HistogramBuilder (core/observables/HistogramBuilder.tpp): max > min, numBins > 0
    RP_EXPECT(max_r > 0, "maxR > 0");
    RP_EXPECT(n_bins > 0 && n_bins <= 4096, "0 < numBins <= 4096");
    {
        // Not a precondition of the reference, which applies its relative-coordinate minimum image at any distance;
        // beyond half a box height that image is no longer the nearest one and, at a relative separation of exactly
        // 1/2, depends on the last bit — the FP32 cull of the device kernel only promises the reference's counts below.
        double heights[3];
        boxHeights(h->box, heights);
        const double minHeight = std::min(heights[0], std::min(heights[1], heights[2]));
        RP_EXPECT(max_r <= 0.5 * minHeight * (1 - 1e-6), "maxR <= half of the smallest box height");
    }
    static_assert(sizeof(uint64_t) == sizeof(unsigned long long), "uint64_t");
    pairDensityCorrelation(h, max_r, n_bins, focal_point, reinterpret_cast<unsigned long long *>(counts_out), g_out);
    RP_API_END(h)
}

rp_status rp_toggle_walls(rp_handle *h, const int32_t walls[3]) {
    RP_API_BEGIN(h)
    RP_EXPECT(walls != nullptr, "walls != NULL");
    h->wallMask = (walls[0] ? 1 : 0) | (walls[1] ? 2 : 0) | (walls[2] ? 4 : 0);
    RP_API_END(h)
}

rp_status rp_renormalize_orientations(rp_handle *h, int allow_overlaps, uint64_t *rejected_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && h->n > 0, "packing and interaction must be set");
    if (rejected_out) *rejected_out = 0;
    if (allow_overlaps) {
        // the reference resets the packing from a copy taken BEFORE the fix (core/Packing.cpp:1179-1190): the fixed
        // matrices are discarded, only the caches are rebuilt — parity first
        setupForInteraction(h);
    } else {
        const int n = static_cast<int>(h->n);
        ensureStage(h, h->n * 9 * sizeof(double));
        double *backup = static_cast<double *>(h->dStage);
        k_fix_orientations<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, h->orient, backup);
        h->launches++;
        initCentres(h);
        recalcAbsoluteCentres(h);
        h->gridDirty = true;
        h->hasTrial = false;
        // the reference walks the particles one by one and rejects a fix that creates an overlap
        // (core/Packing.cpp:1195-1211).  Here all fixes are applied at once and checked by one early-exit recount (the
        // fixes are O(1e-15): it practically always passes).  If it does not, the fix of every particle that takes
        // part in an overlap is withdrawn — per-particle rejection like the reference's, except that of two
        // neighbours whose fixes only collide with each other both are rejected where the reference would keep the
        // first — and the check is repeated; after three rounds every fix is withdrawn.
        unsigned long long rejected = 0;
        if (h->inter.hard != RP_HARD_NONE && countTotalOverlaps(h, true) > 0) {
            unsigned char *flags = nullptr, *withdrawn = nullptr;
            RP_CUDA(cudaMalloc(&flags, 2 * h->n));
            withdrawn = flags + h->n;
            RP_CUDA(cudaMemsetAsync(withdrawn, 0, h->n, h->stream));
            bool clean = false;
            for (int round = 0; round < 3 && !clean; round++) {
                particleScan(h, true, false, flags, nullptr);
                RP_CUDA(cudaMemsetAsync(h->dCounters + 8, 0, sizeof(unsigned long long), h->stream));
                k_restore_orientations<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, h->orient, backup, flags,
                                                                              withdrawn, h->dCounters + 8);
                h->launches++;
                unsigned long long now = 0;
                RP_CUDA(cudaMemcpyAsync(&now, h->dCounters + 8, sizeof(now), cudaMemcpyDeviceToHost, h->stream));
                RP_CUDA(cudaStreamSynchronize(h->stream));
                rejected += now;
                initCentres(h);
                recalcAbsoluteCentres(h);
                h->gridDirty = true;
                clean = countTotalOverlaps(h, true) == 0;
            }
            cudaFree(flags);
            if (!clean) {
                k_restore_orientations<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->cap, h->orient, backup, nullptr,
                                                                              nullptr, nullptr);
                h->launches++;
                initCentres(h);
                recalcAbsoluteCentres(h);
                h->gridDirty = true;
                rejected = h->n;
            }
        }
        if (rejected_out) *rejected_out = rejected;
        RP_CUDA(cudaStreamSynchronize(h->stream));
    }
    RP_API_END(h)
}

rp_status rp_size(const rp_handle *h, uint64_t *n_out) {
    if (!h || !n_out) return RP_ERR_PRECONDITION;
    *n_out = h->n;
    return RP_OK;
}

rp_status rp_rebuild_grid(rp_handle *h) {
    RP_API_BEGIN(h)
    rebuildReferenceGrid(h);
    RP_API_END(h)
}

rp_status rp_grid_info_get(rp_handle *h, rp_grid_info *out) {
    RP_API_BEGIN(h)
    RP_EXPECT(out != nullptr, "out != NULL");
    ensureGrid(h);
    std::memset(out, 0, sizeof(*out));
    out->has_grid = h->hasGrid ? 1 : 0;
    out->cell_size = h->grid.cellSize;
    out->num_centres = h->n * h->inter.kk;
    if (h->hasGrid) {
        for (int i = 0; i < 3; i++) out->cell_divisions[i] = h->grid.n[i];
        out->max_cell_population = h->grid.maxPopulation;
    }
    RP_API_END(h)
}

rp_status rp_cell_numbers(rp_handle *h, uint64_t *cell_no_out) {
    RP_API_BEGIN(h)
    ensureGrid(h);
    RP_EXPECT(h->hasGrid, "the packing has no neighbour grid (Packing.cpp:868-881)");
    size_t m = h->grid.numItems;
    std::vector<int> keys(m);
    RP_CUDA(cudaMemcpyAsync(keys.data(), h->grid.keys, m * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    // reference numbering incl. ghost layers: coordinates + 1, divisions + 2 (NeighbourGrid.cpp:13-34)
    const int *n = h->grid.n;
    for (size_t i = 0; i < m; i++) {
        int c = keys[i];
        uint64_t cx = c % n[0], cy = (c / n[0]) % n[1], cz = c / (n[0] * n[1]);
        cell_no_out[i] = ((cz + 1) * (n[1] + 2) + (cy + 1)) * (n[0] + 2) + (cx + 1);
    }
    RP_API_END(h)
}

rp_status rp_count_total_overlaps(rp_handle *h, int early_exit, uint64_t *count_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && h->n > 0 && count_out, "packing and interaction must be set");
    *count_out = countTotalOverlaps(h, early_exit != 0);
    RP_API_END(h)
}

rp_status rp_total_energy(rp_handle *h, double *energy_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && h->n > 0 && energy_out, "packing and interaction must be set");
    *energy_out = totalEnergy(h);
    RP_API_END(h)
}

rp_status rp_try_move(rp_handle *h, const rp_move *move, double *dE_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(move && dE_out, "move and dE_out must not be NULL");
    *dE_out = tryMove(h, *move);
    RP_API_END(h)
}

rp_status rp_accept_move(rp_handle *h) {
    RP_API_BEGIN(h)
    acceptMove(h);
    RP_API_END(h)
}

rp_status rp_try_scaling(rp_handle *h, const double new_box[9], double *dE_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(new_box && dE_out, "new_box and dE_out must not be NULL");
    *dE_out = tryScaling(h, makeBox(new_box));
    RP_API_END(h)
}

rp_status rp_revert_scaling(rp_handle *h) {
    RP_API_BEGIN(h)
    revertScaling(h);
    RP_API_END(h)
}

rp_status rp_sweep(rp_handle *h, const rp_sweep_params *params, rp_sweep_counters *counters_out) {
    RP_API_BEGIN(h)
    RP_EXPECT(params != nullptr, "params != NULL");
    RP_EXPECT(h->hasInteraction && h->n > 0, "packing and interaction must be set");
    sweep(h, *params, counters_out);
    RP_API_END(h)
}

rp_status rp_sweep_trace(rp_handle *h, rp_trial_trace *out, uint64_t n) {
    RP_API_BEGIN(h)
    RP_EXPECT(out != nullptr && n == h->n, "out != NULL && n == size()");
    sweepTrace(h, out, n);
    RP_API_END(h)
}

rp_status rp_comm_unique_id(uint8_t id_out[128]) {
    try {
        if (!id_out) throw rp::rp_error(RP_ERR_PRECONDITION, "id_out != NULL");
        shardUniqueId(id_out);
    } catch (const rp::rp_error &e) { rp::tlsError = e.what(); return static_cast<rp_status>(e.status); }
    return RP_OK;
}

rp_status rp_comm_init(rp_handle *h, const uint8_t id[128], int32_t rank, int32_t n_ranks) {
    RP_API_BEGIN(h)
    RP_EXPECT(id != nullptr && n_ranks <= 64, "id != NULL && n_ranks <= 64");
    shardInit(h, id, rank, n_ranks);
    RP_API_END(h)
}

rp_status rp_comm_destroy(rp_handle *h) {
    RP_API_BEGIN(h)
    shardDestroy(h);
    RP_API_END(h)
}

rp_status rp_shard_plan(int32_t n_layers_z, int32_t n_ranks, int32_t rank, int32_t colour, int32_t out[6]) {
    if (!out || n_ranks < 1 || rank < 0 || rank >= n_ranks || n_layers_z < 2 * n_ranks || (n_layers_z & 1)) {
        rp::tlsError = "rp_shard_plan: need an even n_layers_z >= 2 * n_ranks and 0 <= rank < n_ranks";
        return RP_ERR_PRECONDITION;
    }
    HaloPlan p = shardHaloPlan(n_layers_z, n_ranks, rank, colour);
    out[0] = p.aLo; out[1] = p.aHi; out[2] = p.sendLayer; out[3] = p.recvLayer; out[4] = p.peerSend;
    out[5] = p.peerRecv;
    return RP_OK;
}

rp_status rp_comm_stats(rp_handle *h, uint64_t *halo_bytes, uint64_t *gather_bytes) {
    RP_API_BEGIN(h)
    if (halo_bytes) *halo_bytes = h->shard ? h->shard->lastHaloBytes : 0;
    if (gather_bytes) *gather_bytes = h->shard ? h->shard->lastGatherBytes : 0;
    RP_API_END(h)
}

rp_status rp_toggle_overlap_counting(rp_handle *h, int count_overlaps) {
    RP_API_BEGIN(h)
    // Packing::toggleOverlapCounting (core/Packing.cpp:1049-1053)
    h->overlapCounting = count_overlaps != 0;
    if (h->overlapCounting && h->hasInteraction && h->n > 0)
        h->numOverlaps = countTotalOverlaps(h, false);
    RP_API_END(h)
}

rp_status rp_cached_overlaps(rp_handle *h, uint64_t *count_out) {
    RP_API_BEGIN(h)
    // Packing::getCachedNumberOfOverlaps (core/Packing.cpp:1055-1060)
    if (!h->overlapCounting)
        throw rp_error(RP_ERR_STATE,
                       "Packing: overlap counting is toggled false; number of overlaps is not cached");
    *count_out = h->numOverlaps;
    RP_API_END(h)
}

rp_status rp_pair_test_batch(rp_handle *h, const rp_pair *pairs, uint64_t n, uint8_t *verdict_out,
                             double *energy_out)
{
    RP_API_BEGIN(h)
    pairTestBatch(h, pairs, n, verdict_out, energy_out);
    RP_API_END(h)
}

rp_status rp_last_kernel_ms(rp_handle *h, double *sweep_ms, double *sweep_grid_ms, double *rebuild_ms,
                            double *scaling_ms)
{
    RP_API_BEGIN(h)
    sweepTimings(h, true);
    if (sweep_ms) *sweep_ms = h->lastSweepMs;
    if (sweep_grid_ms) *sweep_grid_ms = h->lastSweepGridMs;
    if (rebuild_ms) *rebuild_ms = h->lastRebuildMs;
    if (scaling_ms) *scaling_ms = h->lastScalingMs;
    RP_API_END(h)
}

rp_status rp_neighbour_stats(rp_handle *h, double *candidates_per_centre, double *within_range_per_centre) {
    RP_API_BEGIN(h)
    RP_EXPECT(h->hasInteraction && h->n > 0 && candidates_per_centre && within_range_per_centre,
              "packing and interaction must be set");
    neighbourStats(h, candidates_per_centre, within_range_per_centre);
    RP_API_END(h)
}

rp_status rp_measure_fp64_peak(rp_handle *h, double *dfma_gflops, double *dmul_dadd_gflops) {
    RP_API_BEGIN(h)
    RP_EXPECT(dfma_gflops && dmul_dadd_gflops, "output pointers must not be NULL");
    fp64Peak(h, dfma_gflops, dmul_dadd_gflops);
    RP_API_END(h)
}

rp_status rp_launch_count(rp_handle *h, uint64_t *launches_out) {
    RP_API_BEGIN(h)
    *launches_out = h->launches;
    RP_API_END(h)
}

rp_status rp_synchronize(rp_handle *h) {
    RP_API_BEGIN(h)
    RP_CUDA(cudaStreamSynchronize(h->stream));
    RP_API_END(h)
}

}  // extern "C"
