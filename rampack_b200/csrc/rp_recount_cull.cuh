// rp_recount_cull.cuh — full-system overlap recount for single-centre hard spheres / spherocylinders with the same
// staged pair test as the sweep kernel: conservative FP32 cull (warp-wide, branch-free), survivors compacted into
// a per-warp queue, (spherocylinders) FP32 segment-segment cull 64 pairs at a time, its survivors compacted again,
// exact FP64 test (reference operation order) 64 pairs at a time.  Included by rp_recount.cu.
//
// Pair enumeration = the reference's (core/Packing.cpp:431-559): every real cell x {own cell i < j, 13 "positive"
// neighbour cells (9a + 3b + c > 13, core/NeighbourGrid.cpp:127-131)}; one warp per cell.
#pragma once

#include "rp_internal.cuh"
#include "rp_sweep_kernel.cuh"

namespace rp {

constexpr int RC_WARPS = 4;
constexpr int RC_QCAP = 96;       // first-cull survivors: < 64 waiting + 32 pushed per round
constexpr int RC_Q2CAP = 128;     // second-cull survivors: < 64 waiting + 64 per round

struct RecountArgs {
    SweepInter inter;
    float cullThr2, cullHalf, cull2Thr2;
    int earlyExit;
    int cellBegin, cellEnd;     // cells this launch covers (sharded runs: this rank's share)
};

template <int HARD>
__global__ void __launch_bounds__(RC_WARPS * 32, 4)
k_recount_cull(GridView g, Box box, RecountArgs A, unsigned long long *__restrict__ count, int *__restrict__ found)
{
    __shared__ float4 sOwnF[RC_WARPS][SW_MAXOWN], sOwnU[RC_WARPS][SW_MAXOWN];
    __shared__ double sOwn6[RC_WARPS][SW_MAXOWN][6];
    __shared__ unsigned long long sQueue[RC_WARPS][RC_QCAP], sQueue2[RC_WARPS][RC_Q2CAP];
    __shared__ int sNbBegin[RC_WARPS][16], sNbStart[RC_WARPS][16], sNbT[RC_WARPS][16];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cell = A.cellBegin + blockIdx.x * RC_WARPS + warp;
    if (cell >= A.cellEnd) return;
    if (A.earlyExit && *reinterpret_cast<volatile int *>(found)) return;
    const SweepInter &I = A.inter;
    float4 *ownF = sOwnF[warp], *ownU = sOwnU[warp];
    double (*own6)[6] = sOwn6[warp];
    unsigned long long *q = sQueue[warp], *q2 = sQueue2[warp];
    int *nbBegin = sNbBegin[warp], *nbStart = sNbStart[warp], *nbT = sNbT[warp];

    const int cx = cell % g.n[0], cy = (cell / g.n[0]) % g.n[1], cz = cell / (g.n[0] * g.n[1]);
    const int start = g.cellStart[cell], nOwnAll = g.cellStart[cell + 1] - start;
    if (nOwnAll == 0) return;

    // the 14 candidate cells: own cell (i < j only) and the 13 positive neighbours
    int myCnt = 0;
    if (lane < 14) {
        const int code = 13 + lane;
        int tIdx;
        const int c2 = neighbourCell(g.n, cx, cy, cz, code / 9 - 1, (code / 3) % 3 - 1, code % 3 - 1, &tIdx);
        nbStart[lane] = g.cellStart[c2];
        myCnt = g.cellStart[c2 + 1] - g.cellStart[c2];
        nbT[lane] = tIdx;
    }
    int incl = myCnt;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    if (lane < 14) nbBegin[lane] = incl - myCnt;
    if (lane == 13) nbBegin[14] = incl;
    const int total = __shfl_sync(0xffffffffu, incl, 13);
    __syncwarp();

    const V3 origin = relToAbs(box, v3((cx + 0.5) * g.relCell[0], (cy + 0.5) * g.relCell[1], (cz + 0.5) * g.relCell[2]));
    const float thrF = A.cullThr2, halfF = A.cullHalf;
    unsigned long long overlaps = 0;
    int qn = 0, q2n = 0;
    const unsigned int ltMask = (1u << lane) - 1u;

    // exact tests of two queue entries per lane; entry = own index (6 bits) | candidate cell slot k (4 bits) << 8 |
    // sorted index of the candidate << 16
    auto exact2 = [&](unsigned long long e0, unsigned long long e1, bool v0, bool v1) {
        double c0[6], c1[6];
        const unsigned long long es[2] = {e0, e1};
        #pragma unroll
        for (int w = 0; w < 2; w++) {
            double *c = w == 0 ? c0 : c1;
            const int k = static_cast<int>((es[w] >> 8) & 0xf);
            const int j = static_cast<int>(es[w] >> 16);
            const V3 t = g.trans[nbT[k]];
            c[0] = g.px[j] + t.x; c[1] = g.py[j] + t.y; c[2] = g.pz[j] + t.z;
            if (HARD == RP_HARD_SPHEROCYLINDER) {
                c[3] = g.ax[j] * I.scHalf; c[4] = g.ay[j] * I.scHalf; c[5] = g.az[j] * I.scHalf;
            } else { c[3] = 0; c[4] = 0; c[5] = 0; }
        }
        const bool r0 = exactOverlap<HARD>(I, own6[e0 & 0x3f], c0);
        const bool r1 = exactOverlap<HARD>(I, own6[e1 & 0x3f], c1);
        overlaps += (v0 && r0) + (v1 && r1);
    };

    // second cull (segmentCullPass, rp_sweep_kernel.cuh) of two queue entries per lane; survivors to q2 by ballot
    auto segmentCull2 = [&](unsigned long long e0, unsigned long long e1, bool v0, bool v1) {
        if (HARD == RP_HARD_SPHEROCYLINDER) {
            const unsigned long long es[2] = {e0, e1};
            bool pass[2] = {v0, v1};
            #pragma unroll
            for (int w = 0; w < 2; w++) {
                if (!pass[w]) continue;
                const int k = static_cast<int>((es[w] >> 8) & 0xf);
                const int j = static_cast<int>(es[w] >> 16);
                const int i = static_cast<int>(es[w] & 0x3f);
                const V3 t = g.trans[nbT[k]];
                const double c[6] = {g.px[j] + t.x, g.py[j] + t.y, g.pz[j] + t.z,
                                     g.ax[j] * I.scHalf, g.ay[j] * I.scHalf, g.az[j] * I.scHalf};
                pass[w] = segmentCullPass(c, origin, ownF[i], ownU[i], halfF, A.cull2Thr2);
            }
            v0 = pass[0]; v1 = pass[1];
        }
        const unsigned int b0 = __ballot_sync(0xffffffffu, v0), b1 = __ballot_sync(0xffffffffu, v1);
        if (v0) q2[q2n + __popc(b0 & ltMask)] = e0;
        if (v1) q2[q2n + __popc(b0) + __popc(b1 & ltMask)] = e1;
        q2n += __popc(b0) + __popc(b1);
        __syncwarp();
        while (q2n >= 64) {
            q2n -= 64;
            exact2(q2[q2n + lane], q2[q2n + 32 + lane], true, true);
        }
        __syncwarp();
    };

    for (int batch0 = 0; batch0 < nOwnAll; batch0 += SW_MAXOWN) {
        const int nOwn = min(SW_MAXOWN, nOwnAll - batch0);
        __syncwarp();
        for (int i = lane; i < nOwn; i += 32) {
            const int s = start + batch0 + i;
            const V3 p = v3(g.px[s], g.py[s], g.pz[s]);
            V3 h = v3(0, 0, 0);
            if (HARD == RP_HARD_SPHEROCYLINDER) h = v3(g.ax[s], g.ay[s], g.az[s]) * I.scHalf;
            own6[i][0] = p.x; own6[i][1] = p.y; own6[i][2] = p.z; own6[i][3] = h.x; own6[i][4] = h.y; own6[i][5] = h.z;
            ownF[i] = make_float4(static_cast<float>(p.x - origin.x), static_cast<float>(p.y - origin.y),
                                  static_cast<float>(p.z - origin.z), 0.f);
            ownU[i] = make_float4(static_cast<float>(g.ax[s]), static_cast<float>(g.ay[s]),
                                  static_cast<float>(g.az[s]), 0.f);
        }
        __syncwarp();

        // candidates: flat index x over the 14 cells; own-cell candidates only pair with own particles before them
        for (int x0 = batch0; x0 < total; x0 += 32) {     // own-cell candidates before the batch pair with nobody
            const int x = x0 + lane;
            const bool valid = x < total;
            int k = 0;
            if (valid) { while (x >= nbBegin[k + 1]) k++; }
            float cxf = 1e18f, cyf = 0.f, czf = 0.f;
            int j = 0;
            if (valid) {
                j = nbStart[k] + (x - nbBegin[k]);
                const V3 t = g.trans[nbT[k]];
                cxf = static_cast<float>((g.px[j] + t.x) - origin.x);
                cyf = static_cast<float>((g.py[j] + t.y) - origin.y);
                czf = static_cast<float>((g.pz[j] + t.z) - origin.z);
            }
            unsigned int mlo = 0, mhi = 0;
            #pragma unroll
            for (int half = 0; half < 2; half++) {
                const int i0 = half * 32, i1 = min(nOwn, i0 + 32);
                unsigned int bit = 1, mm = 0;
                #pragma unroll 4
                for (int i = i0; i < i1; i++) {
                    const float4 tf = ownF[i];
                    float dx = cxf - tf.x, dy = cyf - tf.y, dz = czf - tf.z;
                    float d2;
                    if (HARD == RP_HARD_SPHEROCYLINDER) {
                        const float4 tu = ownU[i];
                        float s = fmaf(dz, tu.z, fmaf(dy, tu.y, dx * tu.x));
                        s = fminf(fmaxf(s, -halfF), halfF);
                        float rx = fmaf(-s, tu.x, dx), ry = fmaf(-s, tu.y, dy), rz = fmaf(-s, tu.z, dz);
                        d2 = fmaf(rz, rz, fmaf(ry, ry, rx * rx));
                    } else {
                        d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
                    }
                    if (d2 < thrF) mm |= bit;
                    bit <<= 1;
                }
                if (half == 0) mlo = mm; else mhi = mm;
            }
            unsigned long long m = (static_cast<unsigned long long>(mhi) << 32) | mlo;
            if (valid && k == 0) {
                // own cell: candidate at local index x pairs with own particles i (global local index batch0 + i) < x
                const int lim = x - batch0;          // own indices 0 .. lim-1 of this batch
                m &= lim >= 64 ? ~0ull : (lim <= 0 ? 0ull : ((1ull << lim) - 1ull));
            }
            if (!valid) m = 0ull;

            // push: one entry per lane and round, queue slots from a ballot
            #pragma unroll 1
            for (int half = 0; half < 2; half++) {
                unsigned int mm = half ? static_cast<unsigned int>(m >> 32) : static_cast<unsigned int>(m);
                unsigned int has;
                while ((has = __ballot_sync(0xffffffffu, mm != 0u)) != 0u) {
                    if (mm != 0u) {
                        const int i = __ffs(static_cast<int>(mm)) - 1 + 32 * half;
                        mm &= mm - 1u;
                        q[qn + __popc(has & ltMask)] = static_cast<unsigned long long>(i)
                                                       | (static_cast<unsigned long long>(k) << 8)
                                                       | (static_cast<unsigned long long>(j) << 16);
                    }
                    qn += __popc(has);
                    __syncwarp();
                    if (qn >= 64) {
                        qn -= 64;
                        segmentCull2(q[qn + lane], q[qn + 32 + lane], true, true);
                    }
                    __syncwarp();
                }
            }
        }
        // drain before the own records are replaced by the next batch
        if (qn > 0) {
            const bool v0 = lane < qn, v1 = lane + 32 < qn;
            segmentCull2(v0 ? q[lane] : 0ull, v1 ? q[lane + 32] : 0ull, v0, v1);
            qn = 0;
        }
        if (q2n > 0) {
            const bool v0 = lane < q2n, v1 = lane + 32 < q2n;
            exact2(v0 ? q2[lane] : 0ull, v1 ? q2[lane + 32] : 0ull, v0, v1);
            q2n = 0;
        }
        if (A.earlyExit && __any_sync(0xffffffffu, overlaps > 0)) break;
    }

    for (int o = 16; o > 0; o >>= 1) overlaps += __shfl_xor_sync(0xffffffffu, overlaps, o);
    if (lane == 0 && overlaps > 0) {
        atomicAdd(count, overlaps);
        if (A.earlyExit) atomicExch(found, 1);
    }
}

}  // namespace rp
