// rp_sweep_kernel2.cuh — k_sweep, the per-colour sub-sweep kernel (see rp_sweep_kernel.cuh for the overview)
#pragma once

namespace rp {

template <int HARD, int THREADS>
__global__ void __launch_bounds__(THREADS, THREADS == 256 ? 3 : 4)
k_sweep(SweepArgs A, SweepGrid sg, Box box, const int *__restrict__ cellStart,
        const int *__restrict__ items, double *__restrict__ rec, TrialBuffers T, double *__restrict__ pos,
        double *__restrict__ orient, unsigned long long *__restrict__ counters, int *__restrict__ flags,
        PeerView peers)
{
    constexpr int WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ float4 trialF[SW_MAXOWN], trialU[SW_MAXOWN];
    // the same FP32 trial data for the packed (f32x2) cull loop, two trials per entry: (-x0, -x1, -y0, -y1),
    // (-z0, -z1, ux0, ux1), (uy0, uy1, uz0, uz1)
    __shared__ float4 trialP[SW_MAXOWN / 2 + 1][3];
    __shared__ unsigned long long ownMask[SW_MAXOWN], newMask[SW_MAXOWN], seqOwn[SW_MAXOWN], seqNew[SW_MAXOWN];
    __shared__ unsigned int orderKey[SW_MAXOWN];
    __shared__ unsigned long long seqBit[SW_MAXOWN];
    __shared__ int rejected[SW_MAXOWN];     // 0 = alive, 1 = overlap with a frozen particle, 2 = left the cell
    __shared__ unsigned long long sAccepted, sRejMask, sBar;
    __shared__ int sDomain;
    __shared__ int cellBegin[28], cellT[27];
    __shared__ int sStats[3];
    __shared__ int qCount;

    // dynamic shared memory: FP64 records (old states, then the trial = new states of the own cell), the warp queues
    const int capCand = A.capOld + A.capOwn;
    double *cand = reinterpret_cast<double *>(smemRaw);
    unsigned int *queue = reinterpret_cast<unsigned int *>(cand + 6 * static_cast<size_t>(capCand));

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const SweepInter &I = A.inter;

    // active cell of this CTA
    const int b = blockIdx.x;
    const int a0 = b % A.nActive[0], a1 = (b / A.nActive[0]) % A.nActive[1],
              a2 = b / (A.nActive[0] * A.nActive[1]) + A.aOff2;
    const int cx = 2 * a0 + A.colour[0], cy = 2 * a1 + A.colour[1], cz = 2 * a2 + A.colour[2];
    const int ownCell = (cz * sg.n[1] + cy) * sg.n[0] + cx;

    // neighbour code c = (dz+1)*9 + (dy+1)*3 + (dx+1).  Staging position p holds the cell with code SW_STAGE_ORDER[p]:
    // the own cell first, then the 6 face, 12 edge and 8 corner neighbours — chunks of candidates are handed out in
    // this order, so the expensive ones (many survivors of the cull) come first and the warps finish together.
    // Warp 0 does the whole set-up in registers (one round trip to the cell table, scan, TMA issue) — the only CTA
    // barrier before the data lands is the one below.
    int ownStartAll, ownCountAll;
    if (warp == 0) {
        int cnt = 0, src = 0, tIdx = 13;
        if (lane < 27) {
            const int code = SW_STAGE_ORDER[lane];
            const int c2 = neighbourCell(sg.n, cx, cy, cz, code % 3 - 1, (code / 3) % 3 - 1, code / 9 - 1, &tIdx);
            src = cellStart[c2];
            cnt = cellStart[c2 + 1] - src;
        }
        if (lane == 0) {
            sStats[0] = 0; sStats[1] = 0; sStats[2] = 0;
            mbarInit(&sBar, 1);
        }
        // exclusive scan of the 27 cell populations -> staging offsets; every lane then issues the 1-D TMA bulk copy
        // of its own cell (lane 27: the trial records of the own cell)
        int incl = cnt;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 26);
        ownStartAll = __shfl_sync(0xffffffffu, src, 0);
        ownCountAll = __shfl_sync(0xffffffffu, cnt, 0);
        const bool ok = ownCountAll > 0 && total <= A.capOld && ownCountAll <= A.capOwn;
        if (lane == 0 && ok) mbarExpectTx(&sBar, static_cast<unsigned int>((total + ownCountAll) * 48));
        __syncwarp();
        if (lane < 27) {
            cellBegin[lane] = incl - cnt;
            cellT[lane] = tIdx;
            if (lane == 26) cellBegin[27] = incl;
            if (ok && cnt > 0)
                bulkCopyG2S(cand + 6 * static_cast<size_t>(incl - cnt), rec + 6 * static_cast<size_t>(src),
                            static_cast<unsigned int>(cnt * 48), &sBar);
        } else if (lane == 27 && ok) {
            bulkCopyG2S(cand + 6 * static_cast<size_t>(total), T.rec + 6 * static_cast<size_t>(ownStartAll),
                        static_cast<unsigned int>(ownCountAll * 48), &sBar);
        }
    } else {
        ownStartAll = cellStart[ownCell];
        ownCountAll = cellStart[ownCell + 1] - ownStartAll;
    }
    // trial meta of the first batch: requested now, used after the staging has landed
    uint2 metaFirst = make_uint2(0u, 0u);
    if (tid < min(ownCountAll, SW_MAXOWN)) metaFirst = T.meta[ownStartAll + tid];
    __syncthreads();
    const int numOld = cellBegin[27], ownBeginAll = cellBegin[0], numOwnAll = cellBegin[1] - cellBegin[0];
    if (numOwnAll == 0) return;
    if (numOld > A.capOld || numOwnAll > A.capOwn) {
        if (tid == 0) atomicOr(&flags[3], 1);
        return;
    }

    // origin of the FP32 local frame: centre of the active cell (image frame)
    const V3 origin = relToAbs(box, v3((cx + 0.5) * sg.relCell[0] + sg.shift[0], (cy + 0.5) * sg.relCell[1] + sg.shift[1],
                                       (cz + 0.5) * sg.relCell[2] + sg.shift[2]));
    if (warp == 0) mbarWait(&sBar, 0);       // the other warps park at the CTA barrier instead of spinning
    __syncthreads();

    // ---- second staging pass (shared memory only): pos2 + lattice translation (HardcodedTranslation) ----
    for (int c = warp; c < 27; c += WARPS) {
        if (cellT[c] == 13) continue;        // translation (0, 0, 0): pos2 + 0 == pos2
        const int begin = cellBegin[c], cnt = cellBegin[c + 1] - begin;
        const V3 t = sg.trans[cellT[c]];
        for (int j = begin + lane; j < begin + cnt; j += 32) {
            double *r = cand + 6 * static_cast<size_t>(j);
            r[0] += t.x; r[1] += t.y; r[2] += t.z;
        }
    }

    // (a spherocylinder of length 0 is admissible, SpherocylinderTraits.cpp:15: its unit axis is irrelevant, not NaN)
    const float invHalfF = (HARD == RP_HARD_SPHEROCYLINDER && A.cullHalf > 0.f) ? __fdividef(1.0f, A.cullHalf) : 0.f;
    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION);

    // A cell with more than SW_MAXOWN particles is processed in batches of SW_MAXOWN trials; the particles of the
    // other batches are frozen candidates meanwhile (their accepted states are patched into shared memory).
    for (int batch0 = 0; batch0 < numOwnAll; batch0 += SW_MAXOWN) {
        const int numOwn = min(SW_MAXOWN, numOwnAll - batch0);
        const int ownBegin = ownBeginAll + batch0, ownStart = ownStartAll + batch0;
        const int newBegin = numOld + batch0;          // new states of this batch's trials inside cand[]
        const int numCand = numOld + numOwn;

        if (tid == 0) { sAccepted = 0ull; sRejMask = 0ull; qCount = 0; sDomain = 0; }
        __syncthreads();       // staging passes done (first batch) / previous batch fully written back
        if (tid < numOwn) {
            const uint2 meta = batch0 == 0 ? metaFirst : T.meta[ownStart + tid];
            const double *r = cand + 6 * static_cast<size_t>(newBegin + tid);
            V3 p = v3(r[0], r[1], r[2]), h = v3(r[3], r[4], r[5]);
            trialF[tid] = make_float4(static_cast<float>(p.x - origin.x), static_cast<float>(p.y - origin.y),
                                      static_cast<float>(p.z - origin.z), 0.f);
            if (HARD == RP_HARD_SPHEROCYLINDER)      // FP32 unit axis of the trial for the capsule cull
                trialU[tid] = make_float4(static_cast<float>(h.x) * invHalfF, static_cast<float>(h.y) * invHalfF,
                                          static_cast<float>(h.z) * invHalfF, 0.f);
            {
                float *tp = reinterpret_cast<float *>(&trialP[tid >> 1][0]) + (tid & 1);
                tp[0] = -trialF[tid].x; tp[2] = -trialF[tid].y; tp[4] = -trialF[tid].z;
                if (HARD == RP_HARD_SPHEROCYLINDER) { tp[6] = trialU[tid].x; tp[8] = trialU[tid].y; tp[10] = trialU[tid].z; }
            }
            rejected[tid] = static_cast<int>(meta.x);
            orderKey[tid] = meta.y;
            ownMask[tid] = 0ull;
            newMask[tid] = 0ull;
            if (meta.x) atomicOr(&sRejMask, 1ull << tid);
            if (meta.x == 2u) atomicAdd(&sDomain, 1);
        }
        if (tid == numOwn && (numOwn & 1)) {        // odd count: the partner of the last trial is out of everyone's reach
            float *tp = reinterpret_cast<float *>(&trialP[tid >> 1][0]) + 1;
            tp[0] = -1e18f; tp[2] = 0.f; tp[4] = 0.f; tp[6] = 0.f; tp[8] = 0.f; tp[10] = 0.f;
        }
        __syncthreads();

        // position of every trial in the cell's random order (needed by phase C; computed here, off the critical path:
        // the other warps are already culling)
        int rank = 0;
        if (tid < numOwn) {
            const unsigned int k = orderKey[tid];
            for (int j = 0; j < numOwn; j++) {
                const unsigned int kj = orderKey[j];
                rank += (kj < k) || (kj == k && j < tid);
            }
        }

        volatile int *vrej = rejected;
        auto record = [&](int ti, int cj) {
            if (cj >= numOld) atomicOr(&newMask[ti], 1ull << (cj - newBegin));
            else {
                const int o = cj - ownBegin;
                if (o >= 0 && o < numOwn) atomicOr(&ownMask[ti], 1ull << o);
                else vrej[ti] = 1;
            }
        };
        // two independent exact tests per lane: their FP64 dependency chains interleave (the pair arithmetic is
        // straight-line code), which is what keeps the FP64 pipe fed at 3-4 warps per scheduler
        auto exactTest2 = [&](unsigned int e0, unsigned int e1, bool v0, bool v1) {
            const int t0 = e0 >> 16, c0 = e0 & 0xffff, t1 = e1 >> 16, c1 = e1 & 0xffff;
            v0 = v0 && !vrej[t0];
            v1 = v1 && !vrej[t1];
            if (!__any_sync(0xffffffffu, v1)) {        // a short tail: one test per lane
                if (v0 && exactOverlap<HARD>(I, cand + 6 * static_cast<size_t>(newBegin + t0),
                                             cand + 6 * static_cast<size_t>(c0)))
                    record(t0, c0);
                return;
            }
            if (!(v0 || v1)) return;
            const bool r0 = exactOverlap<HARD>(I, cand + 6 * static_cast<size_t>(newBegin + t0),
                                               cand + 6 * static_cast<size_t>(c0));
            const bool r1 = exactOverlap<HARD>(I, cand + 6 * static_cast<size_t>(newBegin + t1),
                                               cand + 6 * static_cast<size_t>(c1));
            if (v0 && r0) record(t0, c0);
            if (v1 && r1) record(t1, c1);
        };

        // ---- phases A + B, warp-autonomous (no CTA barrier in between) ----
        // Warps grab chunks of 32 candidates from a shared counter.  A: conservative FP32 cull of the chunk against
        // all trials (spheres: centre distance; spherocylinders: distance of the candidate's centre from the trial's
        // axis segment — a necessary condition for the two segments to come closer than 2r); survivors go to the
        // warp's queue.  B: whenever 64 entries are queued they take (spherocylinders) the FP32 segment-segment cull
        // and its survivors, again in batches of 64, the exact FP64 test with every lane busy.
        {
            const float thrF = A.cullThr2, halfF = A.cullHalf;
            const unsigned long long dead = sRejMask;
            unsigned int *q = queue + warp * SW_WQCAP;       // survivors of the centre-to-axis cull
            unsigned int *q2 = q + SW_WQCAP1;                // survivors of the FP32 segment cull, awaiting FP64
            int qn = 0, q2n = 0;
            const unsigned int ltMask = (1u << lane) - 1u;
            // B1: conservative FP32 segment-segment cull of 64 queued pairs (two per lane), survivors compacted into
            // q2 by ballot; B2: the exact FP64 test on every full batch of 64 survivors
            auto segmentCull = [&](unsigned int e0, unsigned int e1, bool v0, bool v1) {
                if (HARD == RP_HARD_SPHEROCYLINDER) {
                    const int t0 = e0 >> 16, c0 = e0 & 0xffff, t1 = e1 >> 16, c1 = e1 & 0xffff;
                    v0 = v0 && !vrej[t0] && segmentCullPass(cand + 6 * static_cast<size_t>(c0), origin, trialF[t0],
                                                            trialU[t0], halfF, A.cull2Thr2);
                    v1 = v1 && !vrej[t1] && segmentCullPass(cand + 6 * static_cast<size_t>(c1), origin, trialF[t1],
                                                            trialU[t1], halfF, A.cull2Thr2);
                }
                const unsigned int b0 = __ballot_sync(0xffffffffu, v0), b1 = __ballot_sync(0xffffffffu, v1);
                if (v0) q2[q2n + __popc(b0 & ltMask)] = e0;
                if (v1) q2[q2n + __popc(b0) + __popc(b1 & ltMask)] = e1;
                q2n += __popc(b0) + __popc(b1);
                __syncwarp();
                while (q2n >= 64) {
                    q2n -= 64;
                    exactTest2(q2[q2n + lane], q2[q2n + 32 + lane], true, true);
                }
                __syncwarp();
            };
            const int numChunks = (numCand + 31) >> 5;
            for (;;) {
                int chunk = 0;
                if (lane == 0) chunk = atomicAdd(&qCount, 1);
                chunk = __shfl_sync(0xffffffffu, chunk, 0);
                if (chunk >= numChunks) break;
                const int jj = (chunk << 5) + lane;
                const bool valid = jj < numCand;
                const int j = jj < numOwn ? newBegin + jj : jj - numOwn;      // the new states of the trials first
                // a far-away dummy for the invalid slots keeps the loop branch-free
                float cxf = 1e18f, cyf = 0.f, czf = 0.f;
                if (valid) {         // FP32 copy of the position in cell-local coordinates
                    const double *r = cand + 6 * static_cast<size_t>(j);
                    cxf = static_cast<float>(r[0] - origin.x);
                    cyf = static_cast<float>(r[1] - origin.y);
                    czf = static_cast<float>(r[2] - origin.z);
                }
                // two trials per iteration with the packed FP32 instructions of sm_100 (FADD2 / FMUL2 / FFMA2): the
                // same operations in the same order as the scalar form, half the issue slots
                unsigned int mlo = 0, mhi = 0;
                const float2 cx2 = make_float2(cxf, cxf), cy2 = make_float2(cyf, cyf), cz2 = make_float2(czf, czf);
                const int numPairs = (numOwn + 1) >> 1;
                #pragma unroll
                for (int half = 0; half < 2; half++) {
                    const int k0 = half * 16, k1 = min(numPairs, k0 + 16);
                    unsigned int bit = 1, mm = 0;
                    #pragma unroll 4
                    for (int k = k0; k < k1; k++) {
                        const float4 pa = trialP[k][0], pb = trialP[k][1];
                        const float2 dx = __fadd2_rn(cx2, make_float2(pa.x, pa.y)),
                                     dy = __fadd2_rn(cy2, make_float2(pa.z, pa.w)),
                                     dz = __fadd2_rn(cz2, make_float2(pb.x, pb.y));
                        float2 d2;
                        if (HARD == RP_HARD_SPHEROCYLINDER) {
                            const float4 pc = trialP[k][2];
                            const float2 ux = make_float2(pb.z, pb.w), uy = make_float2(pc.x, pc.y),
                                         uz = make_float2(pc.z, pc.w);
                            // |d - s u|^2 = |d|^2 + s (s - 2 t), t = d . u, s = clamp(t), |u| = 1: four packed
                            // instructions fewer than forming d - s u; the rounding error (a few ulp of |d|^2 <= 36
                            // for a pair inside the threshold, ~1e-5) is inside the margin of cullThr2
                            const float2 t = __ffma2_rn(dz, uz, __ffma2_rn(dy, uy, __fmul2_rn(dx, ux)));
                            const float2 dd = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                            const float2 sc = make_float2(fminf(fmaxf(t.x, -halfF), halfF),
                                                          fminf(fmaxf(t.y, -halfF), halfF));
                            d2 = __ffma2_rn(sc, __ffma2_rn(t, make_float2(-2.f, -2.f), sc), dd);
                        } else {
                            d2 = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __fmul2_rn(dx, dx)));
                        }
                        if (d2.x < thrF) mm |= bit;
                        if (d2.y < thrF) mm |= bit << 1;
                        bit <<= 2;
                    }
                    if (half == 0) mlo = mm; else mhi = mm;
                }
                unsigned long long m = (static_cast<unsigned long long>(mhi) << 32) | mlo;
                // never test a particle against itself (old or new state); drop trials that left their cell
                const int o = j - ownBegin, nw = j - newBegin;
                if (o >= 0 && o < numOwn) m &= ~(1ull << o);
                if (nw >= 0 && nw < numOwn) m &= ~(1ull << nw);
                m &= ~dead;

                // push: one entry per lane and round, queue slots from a ballot (no scan); whenever 64 entries are
                // queued they go through the second cull
                #pragma unroll 1
                for (int half = 0; half < 2; half++) {
                    unsigned int mm = half ? static_cast<unsigned int>(m >> 32) : static_cast<unsigned int>(m);
                    unsigned int has;
                    while ((has = __ballot_sync(0xffffffffu, mm != 0u)) != 0u) {
                        if (mm != 0u) {
                            const unsigned int i = static_cast<unsigned int>(__ffs(static_cast<int>(mm)) - 1 + 32 * half);
                            mm &= mm - 1u;
                            q[qn + __popc(has & ltMask)] = (i << 16) | static_cast<unsigned int>(j);
                        }
                        qn += __popc(has);
                        __syncwarp();
                        if (qn >= 64) {
                            qn -= 64;
                            segmentCull(q[qn + lane], q[qn + 32 + lane], true, true);
                        }
                        __syncwarp();
                    }
                }
            }
            if (qn > 0) {
                const bool v0 = lane < qn, v1 = lane + 32 < qn;
                segmentCull(v0 ? q[lane] : 0u, v1 ? q[lane + 32] : 0u, v0, v1);
            }
            if (q2n > 0) {
                const bool v0 = lane < q2n, v1 = lane + 32 < q2n;
                exactTest2(v0 ? q2[lane] : 0u, v1 ? q2[lane + 32] : 0u, v0, v1);
            }
        }

        // ---- phase C: sequential resolution inside the cell, random order, from the bit matrices ----
        __syncthreads();
        if (tid < numOwn) {
            // a rejected trial (left its cell, hit a wall or a frozen particle) is blocked whatever has moved
            const bool alive = rejected[tid] == 0;
            seqOwn[rank] = alive ? ownMask[tid] : ~0ull;
            seqNew[rank] = alive ? newMask[tid] : ~0ull;
            seqBit[rank] = 1ull << tid;
        }
        // global loads of the write-back issued now: their latency passes while thread 0 resolves the cell
        double oNew[9];
        int particleId = 0;
        if (tid < numOwn) {
            particleId = items[ownStart + tid];
            if (rotated) {
                const double *o = T.orient + 9 * static_cast<size_t>(ownStart + tid);
                #pragma unroll
                for (int q = 0; q < 9; q++) oNew[q] = o[q];
            }
        }
        __syncthreads();
        if (tid == 0) {
            // the only loop-carried value is `moved`: select, or, compare, predicated or per step
            unsigned long long moved = 0ull;
            #pragma unroll 4
            for (int step = 0; step < numOwn; step++) {
                const unsigned long long blockers = (seqOwn[step] & ~moved) | (seqNew[step] & moved);
                // hard interactions only: Metropolis u <= exp(-0/T) always accepts (Simulation.cpp:483)
                moved |= blockers == 0ull ? seqBit[step] : 0ull;
            }
            const int nAcc = __popcll(moved), nRejDomain = sDomain;
            sAccepted = moved;
            sStats[0] += nAcc; sStats[1] += nRejDomain; sStats[2] += numOwn - nAcc - nRejDomain;
        }
        __syncthreads();

        // ---- write-back of the accepted moves: sorted record (image frame) and canonical arrays (wrapped) ----
        if (tid < numOwn && ((sAccepted >> tid) & 1ull)) {
            const double *t = cand + 6 * static_cast<size_t>(newBegin + tid);
            double *r = rec + 6 * static_cast<size_t>(ownStart + tid);
            for (int q = 0; q < 6; q++) r[q] = t[q];
            T.meta[ownStart + tid].x = 4u;      // accepted (read back by rp_sweep_trace only)
            if (batch0 + SW_MAXOWN < numOwnAll) {
                // later batches see this particle as a frozen candidate in its new state (own cell: no translation)
                double *cs = cand + 6 * static_cast<size_t>(ownBegin + tid);
                for (int q = 0; q < 6; q++) cs[q] = t[q];
            }
            if (translated) {
                // Shape::translate (core/Shape.cpp:9-12): position += bc.getCorrection(position)
                V3 np = v3(t[0], t[1], t[2]);
                V3 p = np + pbcCorrection(box, np);
                pos[particleId] = p.x; pos[A.cap + particleId] = p.y; pos[2 * A.cap + particleId] = p.z;
            }
            if (rotated) {
                #pragma unroll
                for (int q = 0; q < 9; q++) orient[q * A.cap + particleId] = oNew[q];
            }
            // sharded run over peer memory: the z-neighbour reads my boundary layers in the next colour
            if (peers.n > 0) {
                if (cz == peers.layerLo) {
                    double *pr_ = peers.recLo + 6 * static_cast<size_t>(ownStart + tid);
                    for (int q = 0; q < 6; q++) pr_[q] = t[q];
                }
                if (cz == peers.layerHi) {
                    double *pr_ = peers.recHi + 6 * static_cast<size_t>(ownStart + tid);
                    for (int q = 0; q < 6; q++) pr_[q] = t[q];
                }
            }
        }
        // ... and every other rank learns the verdicts: one accept byte per slot (they hold the same trials)
        if (tid < numOwn) {
            const unsigned char a = static_cast<unsigned char>((sAccepted >> tid) & 1ull);
            #pragma unroll
            for (int pr = 0; pr < SHARD_MAX_PEERS; pr++)      // (unrolled: no dynamic indexing of the argument)
                if (pr < peers.n) peers.acc[pr][ownStart + tid] = a;
        }
    }   // batches

    if (tid == 0) {
        atomicAdd(&counters[0], static_cast<unsigned long long>(numOwnAll));
        atomicAdd(&counters[1], static_cast<unsigned long long>(sStats[0]));
        atomicAdd(&counters[2], static_cast<unsigned long long>(sStats[1]));
        atomicAdd(&counters[3], static_cast<unsigned long long>(sStats[2]));
    }
}

}  // namespace rp
