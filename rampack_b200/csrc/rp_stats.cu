// rp_stats.cu — measurement helpers of the C ABI (not on the Monte Carlo path itself):
//   * neighbourStats: candidates per interaction centre in the reference-geometry grid, i.e. what
//     Packing::getAverageNumberOfNeighbours (core/Packing.cpp:1000-1033) reports ("Average neighbours per centre"),
//     plus how many of them pass the centre-distance pre-filter of the pair function.  bench.py turns these into the
//     algorithmic flop count of a trial move (W_move = K_cand * F_pre + K_full * F_full).
//   * fp64Peak: the FP64 vector-pipe ceiling of the device, (a) fused DFMA, (b) separate DMUL + DADD — the latter is
//     what a kernel compiled with -fmad=false (bit-exact verdicts) can reach.
#include "rp_internal.cuh"

namespace rp {

// one warp per real cell, like k_recount_grid, but over all 27 neighbour cells and only counting
__global__ void __launch_bounds__(256) k_neighbour_stats(GridView g, const Inter *__restrict__ interPtr,
                                                         unsigned long long *__restrict__ out)
{
    __shared__ Inter I;
    for (int i = threadIdx.x; i < static_cast<int>(sizeof(Inter) / sizeof(int)); i += blockDim.x)
        reinterpret_cast<int *>(&I)[i] = reinterpret_cast<const int *>(interPtr)[i];
    __syncthreads();
    int cell = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    int lane = threadIdx.x & 31;
    if (cell >= g.numCells) return;
    int cx = cell % g.n[0], cy = (cell / g.n[0]) % g.n[1], cz = cell / (g.n[0] * g.n[1]);
    int start = g.cellStart[cell], end = g.cellStart[cell + 1];
    int kk = I.kk;
    double thr = I.range * I.range;
    unsigned long long cand = 0, full = 0;
    for (int code = 0; code < 27; code++) {
        int tIdx;
        int cell2 = neighbourCell(g.n, cx, cy, cz, code / 9 - 1, (code / 3) % 3 - 1, code % 3 - 1, &tIdx);
        V3 t = g.trans[tIdx];
        int s2 = g.cellStart[cell2], e2 = g.cellStart[cell2 + 1];
        for (int i = start; i < end; i++) {
            V3 p1 = v3(g.px[i], g.py[i], g.pz[i]);
            int particle1 = g.items[i] / kk;
            for (int j = s2 + lane; j < e2; j += 32) {
                if (g.items[j] / kk == particle1) continue;
                cand++;
                V3 p2 = v3(g.px[j], g.py[j], g.pz[j]);
                if (norm2(p2 + t - p1) < thr) full++;
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        cand += __shfl_xor_sync(0xffffffffu, cand, o);
        full += __shfl_xor_sync(0xffffffffu, full, o);
    }
    if (lane == 0) {
        atomicAdd(&out[0], cand);
        atomicAdd(&out[1], full);
    }
}

void neighbourStats(Handle *h, double *kCand, double *kFull) {
    ensureGrid(h);
    RP_EXPECT(h->hasGrid, "neighbour statistics need a neighbour grid");
    cudaStream_t st = h->stream;
    zeroAsync(h, h->dCounters, 2 * sizeof(unsigned long long), st);
    GridView g = makeView(h->grid);
    int blocks = (g.numCells * 32 + 255) / 256;
    k_neighbour_stats<<<blocks, 256, 0, st>>>(g, h->dInter, h->dCounters);
    h->launches++;
    unsigned long long out[2];
    readbackSync(h, st, out, h->dCounters, sizeof(out));
    RP_CUDA(cudaGetLastError());
    double centres = static_cast<double>(h->n) * h->inter.kk;
    *kCand = static_cast<double>(out[0]) / centres;
    *kFull = static_cast<double>(out[1]) / centres;
}

template <int MODE>
__global__ void __launch_bounds__(256) k_fp64_peak(double *out, int iters, double a, double b) {
    double x[8];
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; it++) {
        #pragma unroll
        for (int i = 0; i < 8; i++) {
            if (MODE == 0) x[i] = __fma_rn(x[i], a, b);
            else { x[i] = __dmul_rn(x[i], a); x[i] = __dadd_rn(x[i], b); }
        }
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += x[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// GFLOP/s, 2 flop per DFMA and 1 per DMUL / DADD; best of 3
void fp64Peak(Handle *h, double *dfmaGflops, double *dmulDaddGflops) {
    int sms = 0;
    RP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device));
    int blocks = sms * 8, threads = 256, iters = 8000;
    double *out = nullptr;
    RP_CUDA(cudaMalloc(&out, sizeof(double) * blocks * threads));
    cudaStream_t st = h->stream;
    double best[2] = {0, 0};
    for (int mode = 0; mode < 2; mode++) {
        for (int rep = 0; rep < 4; rep++) {
            RP_CUDA(cudaEventRecord(h->ev[0], st));
            if (mode == 0) k_fp64_peak<0><<<blocks, threads, 0, st>>>(out, iters, 1.0000001, 1e-9);
            else k_fp64_peak<1><<<blocks, threads, 0, st>>>(out, iters, 1.0000001, 1e-9);
            RP_CUDA(cudaEventRecord(h->ev[1], st));
            RP_CUDA(cudaStreamSynchronize(st));
            float ms = 0;
            cudaEventElapsedTime(&ms, h->ev[0], h->ev[1]);
            double flops = 2.0 * blocks * threads * static_cast<double>(iters) * 8;
            if (rep > 0) best[mode] = std::max(best[mode], flops / ms * 1e-6);
            h->launches++;
        }
    }
    cudaFree(out);
    RP_CUDA(cudaGetLastError());
    *dfmaGflops = best[0];
    *dmulDaddGflops = best[1];
}

}  // namespace rp
