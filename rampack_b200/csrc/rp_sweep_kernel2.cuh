// rp_sweep_kernel2.cuh — k_sweep, the per-colour sub-sweep kernel (see rp_sweep_kernel.cuh for the overview)
//
// Round-2 layout: the two conservative culls run on 32-byte FP32 cull records (position relative to the cell centre +
// unit axis) that the prepare kernel wrote next to the FP64 records; only those are staged in shared memory (TMA bulk
// copies, one per neighbour cell).  The exact FP64 test — the only stage that decides anything — fetches the FP64
// records of the few surviving pairs (2-3 per trial) from global memory (L2).  Against staging the FP64 records this
// takes a third off the staging traffic and the shared memory per CTA, removes every FP64 -> FP32 conversion from the
// cull loops and lets 5-6 CTAs share an SM instead of 4, so that one CTA's serial sections (set-up, sequential
// resolution, write-back) run under the other CTAs' culls.
#pragma once

namespace rp {

template <int HARD, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
k_sweep(SweepArgs A, SweepGrid sg, SweepCellOffsets off, Box box, const int *__restrict__ cellStart,
        const int *__restrict__ items, double *rec, float4 *rec32, TrialBuffers T, double *__restrict__ pos,
        double *__restrict__ orient, unsigned long long *__restrict__ counters, int *__restrict__ flags,
        PeerView peers)
{
    constexpr int WARPS = THREADS / 32;
    extern __shared__ __align__(16) unsigned char smemRaw[];
    // FP32 trial data for the packed (f32x2) cull loop, two trials per entry: (-x0, -x1, -y0, -y1),
    // (-z0, -z1, ux0, ux1), (uy0, uy1, uz0, uz1)
    __shared__ float4 trialP[SW_MAXOWN / 2 + 1][3];
    __shared__ unsigned long long ownMask[SW_MAXOWN], newMask[SW_MAXOWN], seqOwn[SW_MAXOWN], seqNew[SW_MAXOWN];
    __shared__ unsigned int orderKey[SW_MAXOWN];
    __shared__ unsigned long long seqBit[SW_MAXOWN];
    __shared__ int rejected[SW_MAXOWN];     // 0 = alive, 1 = overlap with a frozen particle, 2 = left the cell
    __shared__ unsigned long long sAccepted, sRejMask, sBar;
    __shared__ int sDomain;
    __shared__ int cellBegin[28], cellT[27], cellSrc[27];
    __shared__ int sStats[3];
    __shared__ int qCount;

    // dynamic shared memory: FP32 cull records (old states of the 27 cells, then the trial = new states of the own
    // cell), the FP64 trial records of the own cell, the warp queues
    const int capCand = A.capOld + A.capOwn;
    float4 *cand32 = reinterpret_cast<float4 *>(smemRaw);
    double *trial64 = reinterpret_cast<double *>(cand32 + 2 * static_cast<size_t>(capCand));
    unsigned int *queue = reinterpret_cast<unsigned int *>(trial64 + 6 * static_cast<size_t>(A.capOwn));

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
        // of its own cell (lanes 27 / 28: the FP32 / FP64 trial records of the own cell)
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
        if (lane == 0 && ok)
            mbarExpectTx(&sBar, static_cast<unsigned int>((total + ownCountAll) * 32 + ownCountAll * 48));
        __syncwarp();
        if (lane < 27) {
            cellBegin[lane] = incl - cnt;
            cellT[lane] = tIdx;
            cellSrc[lane] = src;
            if (lane == 26) cellBegin[27] = incl;
            if (ok && cnt > 0)
                bulkCopyG2S(cand32 + 2 * static_cast<size_t>(incl - cnt), rec32 + 2 * static_cast<size_t>(src),
                            static_cast<unsigned int>(cnt * 32), &sBar);
        } else if (lane == 27 && ok) {
            bulkCopyG2S(cand32 + 2 * static_cast<size_t>(total), T.rec32 + 2 * static_cast<size_t>(ownStartAll),
                        static_cast<unsigned int>(ownCountAll * 32), &sBar);
        } else if (lane == 28 && ok) {
            bulkCopyG2S(trial64, T.rec + 6 * static_cast<size_t>(ownStartAll),
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
        // the staging capacity (sized from the previous sweeps' maxima) is too small for this cell: its trials are not
        // attempted in this sweep; the host learns the count and raises the capacity.  Sharded: the other ranks must
        // not replay stale accept bytes for these slots.
        if (tid == 0) atomicAdd(&flags[3], 1);
        for (int b0 = tid * SW_MAXOWN; b0 < numOwnAll; b0 += THREADS * SW_MAXOWN) {
            #pragma unroll
            for (int pr = 0; pr < SHARD_MAX_PEERS; pr++)
                if (pr < peers.n) peers.acc[pr][cellSrc[0] + b0] = 0ull;
        }
        return;
    }

    if (warp == 0) mbarWait(&sBar, 0);       // the other warps park at the CTA barrier instead of spinning
    __syncthreads();

    // ---- second staging pass (shared memory only): neighbour cells move from their own cell-centred frame into the
    // frame of the active cell (FP32; the periodic image translation is part of the centre-to-centre vector) ----
    for (int c = 1 + warp; c < 27; c += WARPS) {        // c = 0 is the own cell
        const int begin = cellBegin[c], cnt = cellBegin[c + 1] - begin;
        const int code = SW_STAGE_ORDER[c];
        const float ox = off.o[code][0], oy = off.o[code][1], oz = off.o[code][2];
        for (int j = begin + lane; j < begin + cnt; j += 32) {
            float4 r = cand32[2 * j];
            r.x += ox; r.y += oy; r.z += oz;
            cand32[2 * j] = r;
        }
    }

    const bool rotated = (A.kind == RP_MOVE_ROTATION || A.kind == RP_MOVE_ROTOTRANSLATION);
    const bool translated = (A.kind == RP_MOVE_TRANSLATION || A.kind == RP_MOVE_ROTOTRANSLATION);

    // FP64 record of staged candidate c (old states: from the sorted records in global memory + the lattice
    // translation of its cell, exactly the reference's pos2 + translation; new states: the staged trial records)
    auto loadCandidate64 = [&](int c, double *out) {
        if (c >= numOld) {
            const double *s = trial64 + 6 * static_cast<size_t>(c - numOld);
            #pragma unroll
            for (int q = 0; q < 6; q++) out[q] = s[q];
            return;
        }
        int ci = 0;      // staged cell of c: last ci with cellBegin[ci] <= c
        #pragma unroll
        for (int step = 16; step > 0; step >>= 1)
            if (ci + step < 27 && cellBegin[ci + step] <= c) ci += step;
        const double2 *s = reinterpret_cast<const double2 *>(rec + 6 * static_cast<size_t>(cellSrc[ci] + (c - cellBegin[ci])));
        const double2 s0 = s[0], s1 = s[1], s2 = s[2];
        out[0] = s0.x; out[1] = s0.y; out[2] = s1.x; out[3] = s1.y; out[4] = s2.x; out[5] = s2.y;
        const int ti = cellT[ci];
        if (ti != 13) {      // translation (0, 0, 0): pos2 + 0 == pos2
            const V3 t = sg.trans[ti];
            out[0] += t.x; out[1] += t.y; out[2] += t.z;
        }
    };

    // A cell with more than SW_MAXOWN particles is processed in batches of SW_MAXOWN trials; the particles of the
    // other batches are frozen candidates meanwhile (accepted states go to the global records and to the staged cull
    // records at the end of a batch).
    for (int batch0 = 0; batch0 < numOwnAll; batch0 += SW_MAXOWN) {
        const int numOwn = min(SW_MAXOWN, numOwnAll - batch0);
        const int ownBegin = ownBeginAll + batch0, ownStart = ownStartAll + batch0;
        const int newBegin = numOld + batch0;          // new states of this batch's trials inside cand32[]
        const int numCand = numOld + numOwn;

        if (tid == 0) { sAccepted = 0ull; sRejMask = 0ull; qCount = 0; sDomain = 0; }
        __syncthreads();       // staging passes done (first batch) / previous batch fully written back
        if (tid < numOwn) {
            const uint2 meta = batch0 == 0 ? metaFirst : T.meta[ownStart + tid];
            const float4 ra = cand32[2 * (newBegin + tid)], rb = cand32[2 * (newBegin + tid) + 1];
            float *tp = reinterpret_cast<float *>(&trialP[tid >> 1][0]) + (tid & 1);
            tp[0] = -ra.x; tp[2] = -ra.y; tp[4] = -ra.z;
            if (HARD == RP_HARD_SPHEROCYLINDER) { tp[6] = ra.w; tp[8] = rb.x; tp[10] = rb.y; }
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
        // exact FP64 test, one pair per lane (after the two culls the exact stage is a few per cent of the kernel's
        // instructions: register pressure matters more than the FP64 pipe's latency here)
        auto exactTest = [&](unsigned int e, bool v) {
            const int t = e >> 16, c = e & 0xffff;
            v = v && !vrej[t];
            if (!v) return;
            double cr[6];
            loadCandidate64(c, cr);
            if (exactOverlap<HARD>(I, trial64 + 6 * static_cast<size_t>(batch0 + t), cr)) record(t, c);
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
            // q2 by ballot; B2: the exact FP64 test on every full batch of 32 survivors
            auto segmentCull = [&](unsigned int e0, unsigned int e1, bool v0, bool v1) {
                if (HARD == RP_HARD_SPHEROCYLINDER) {
                    const int t0 = e0 >> 16, c0 = e0 & 0xffff, t1 = e1 >> 16, c1 = e1 & 0xffff;
                    v0 = v0 && !vrej[t0]
                         && segmentCullPass32(cand32[2 * c0], cand32[2 * c0 + 1], cand32[2 * (newBegin + t0)],
                                              cand32[2 * (newBegin + t0) + 1], halfF, A.cull2Thr2);
                    v1 = v1 && !vrej[t1]
                         && segmentCullPass32(cand32[2 * c1], cand32[2 * c1 + 1], cand32[2 * (newBegin + t1)],
                                              cand32[2 * (newBegin + t1) + 1], halfF, A.cull2Thr2);
                }
                const unsigned int b0 = __ballot_sync(0xffffffffu, v0), b1 = __ballot_sync(0xffffffffu, v1);
                if (v0) q2[q2n + __popc(b0 & ltMask)] = e0;
                if (v1) q2[q2n + __popc(b0) + __popc(b1 & ltMask)] = e1;
                q2n += __popc(b0) + __popc(b1);
                __syncwarp();
                while (q2n >= 32) {
                    q2n -= 32;
                    exactTest(q2[q2n + lane], true);
                }
                __syncwarp();
            };
            const int numChunks = (numCand + 31) >> 5;
            for (;;) {
                int chunk = 0;
                if (lane == 0)      // (plain atom.shared: the compiler's warp-aggregated atomicAdd costs 16 instructions)
                    asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(chunk) : "r"(smemAddr(&qCount)) : "memory");
                chunk = __shfl_sync(0xffffffffu, chunk, 0);
                if (chunk >= numChunks) break;
                const int jj = (chunk << 5) + lane;
                const bool valid = jj < numCand;
                const int j = jj < numOwn ? newBegin + jj : jj - numOwn;      // the new states of the trials first
                // a far-away dummy for the invalid slots keeps the loop branch-free
                float cxf = 1e18f, cyf = 0.f, czf = 0.f;
                if (valid) {
                    const float4 r = cand32[2 * j];
                    cxf = r.x; cyf = r.y; czf = r.z;
                }
                // two trials per iteration with the packed FP32 instructions of sm_100 (FADD2 / FMUL2 / FFMA2): the
                // same operations in the same order as the scalar form, half the issue slots
                unsigned int mlo = 0, mhi = 0;
                const float2 cx2 = make_float2(cxf, cxf), cy2 = make_float2(cyf, cyf), cz2 = make_float2(czf, czf);
                const float2 negThr = make_float2(-thrF, -thrF);
                const int numPairs = (numOwn + 1) >> 1;
                #pragma unroll
                for (int half = 0; half < 2; half++) {
                    const int k0 = half * 16, k1 = min(numPairs, k0 + 16);
                    // the verdict "squared distance < threshold" is the sign bit of (squared distance - threshold),
                    // shifted into the mask by one funnel shift per trial (first trial ends up in the highest bit)
                    unsigned int mm = 0;
                    #pragma unroll 4
                    for (int k = k0; k < k1; k++) {
                        const float4 pa = trialP[k][0], pb = trialP[k][1];
                        const float2 dx = __fadd2_rn(cx2, make_float2(pa.x, pa.y)),
                                     dy = __fadd2_rn(cy2, make_float2(pa.z, pa.w)),
                                     dz = __fadd2_rn(cz2, make_float2(pb.x, pb.y));
                        // |d|^2 - threshold
                        const float2 dd = __ffma2_rn(dz, dz, __ffma2_rn(dy, dy, __ffma2_rn(dx, dx, negThr)));
                        float2 e;
                        if (HARD == RP_HARD_SPHEROCYLINDER) {
                            const float4 pc = trialP[k][2];
                            const float2 ux = make_float2(pb.z, pb.w), uy = make_float2(pc.x, pc.y),
                                         uz = make_float2(pc.z, pc.w);
                            // |d - s u|^2 = |d|^2 + s (s - 2 t), t = d . u, s = clamp(t), |u| = 1: four packed
                            // instructions fewer than forming d - s u; the rounding error (a few ulp of |d|^2 <= 36
                            // for a pair inside the threshold, ~1e-5) is inside the margin of cullThr2
                            const float2 t = __ffma2_rn(dz, uz, __ffma2_rn(dy, uy, __fmul2_rn(dx, ux)));
                            const float2 sc = make_float2(fminf(fmaxf(t.x, -halfF), halfF),
                                                          fminf(fmaxf(t.y, -halfF), halfF));
                            e = __ffma2_rn(sc, __ffma2_rn(t, make_float2(-2.f, -2.f), sc), dd);
                        } else {
                            e = dd;
                        }
                        mm = __funnelshift_l(__float_as_uint(e.x), mm, 1);
                        mm = __funnelshift_l(__float_as_uint(e.y), mm, 1);
                    }
                    // trial 2 (k0 + i) sits at bit 2 (k1 - k0) - 1 - 2 i: reverse into ascending order
                    const int nb = 2 * (k1 - k0);
                    mm = nb > 0 ? __brev(mm) >> (32 - nb) : 0u;
                    if (half == 0) mlo = mm; else mhi = mm;
                }
                unsigned long long m = (static_cast<unsigned long long>(mhi) << 32) | mlo;
                // never test a particle against itself (old or new state); drop trials that left their cell
                const int o = j - ownBegin, nw = j - newBegin;
                if (o >= 0 && o < numOwn) m &= ~(1ull << o);
                if (nw >= 0 && nw < numOwn) m &= ~(1ull << nw);
                m &= ~dead;

                // push: every lane appends all its entries at the offset a warp scan of the counts gives it (one
                // scan per chunk instead of one ballot round per entry).  A chunk whose survivors would not fit into
                // the queue goes through in rounds of at most SW_PUSH entries per lane.
                const unsigned int lowBits = static_cast<unsigned int>(j);
                #pragma unroll 1
                for (;;) {
                    int cnt = __popcll(m);
                    const int total = __reduce_add_sync(0xffffffffu, cnt);
                    if (total == 0) break;
                    unsigned long long take = m;
                    if (qn + total > SW_WQCAP1) {
                        take = 0ull;
                        unsigned long long rest = m;
                        #pragma unroll
                        for (int e = 0; e < SW_PUSH; e++) {
                            const unsigned long long low = rest & (~rest + 1ull);
                            take |= low;
                            rest ^= low;
                        }
                        cnt = __popcll(take);
                    }
                    m ^= take;
                    int incl = cnt;
                    #pragma unroll
                    for (int d = 1; d < 32; d <<= 1) {
                        const int v = __shfl_up_sync(0xffffffffu, incl, d);
                        if (lane >= d) incl += v;
                    }
                    // (the order of a lane's entries is immaterial: highest trial first, one 32-bit word at a time)
                    unsigned int *dst = q + (qn + incl - cnt);
                    #pragma unroll
                    for (int w = 0; w < 2; w++) {
                        unsigned int x = w == 0 ? static_cast<unsigned int>(take) : static_cast<unsigned int>(take >> 32);
                        const unsigned int tag = lowBits | (static_cast<unsigned int>(32 * w) << 16);
                        while (x) {
                            const int i = 31 - __clz(x);
                            x &= ~(1u << i);
                            *dst++ = tag + (static_cast<unsigned int>(i) << 16);
                        }
                    }
                    qn += __shfl_sync(0xffffffffu, incl, 31);
                    __syncwarp();
                    while (qn >= 64) {
                        qn -= 64;
                        segmentCull(q[qn + lane], q[qn + 32 + lane], true, true);
                    }
                    __syncwarp();
                }
            }
            if (qn > 0) {
                const bool v0 = lane < qn, v1 = lane + 32 < qn;
                segmentCull(v0 ? q[lane] : 0u, v1 ? q[lane + 32] : 0u, v0, v1);
            }
            if (q2n > 0) exactTest(lane < q2n ? q2[lane] : 0u, lane < q2n);
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

        // ---- write-back of the accepted moves: sorted records (image frame) and canonical arrays (wrapped) ----
        if (tid < numOwn && ((sAccepted >> tid) & 1ull)) {
            const double *t = trial64 + 6 * static_cast<size_t>(batch0 + tid);
            const float4 ra = cand32[2 * (newBegin + tid)], rb = cand32[2 * (newBegin + tid) + 1];
            double *r = rec + 6 * static_cast<size_t>(ownStart + tid);
            for (int q = 0; q < 6; q++) r[q] = t[q];
            rec32[2 * static_cast<size_t>(ownStart + tid)] = ra;
            rec32[2 * static_cast<size_t>(ownStart + tid) + 1] = rb;
            T.meta[ownStart + tid].x = 4u;      // accepted (read back by rp_sweep_trace only)
            if (batch0 + SW_MAXOWN < numOwnAll) {
                // later batches see this particle as a frozen candidate in its new state (own cell: no offset); their
                // exact tests read the global record written above (ordered by the barrier that ends the batch)
                cand32[2 * (ownBegin + tid)] = ra;
                cand32[2 * (ownBegin + tid) + 1] = rb;
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
                    peers.rec32Lo[2 * static_cast<size_t>(ownStart + tid)] = ra;
                    peers.rec32Lo[2 * static_cast<size_t>(ownStart + tid) + 1] = rb;
                }
                if (cz == peers.layerHi) {
                    double *pr_ = peers.recHi + 6 * static_cast<size_t>(ownStart + tid);
                    for (int q = 0; q < 6; q++) pr_[q] = t[q];
                    peers.rec32Hi[2 * static_cast<size_t>(ownStart + tid)] = ra;
                    peers.rec32Hi[2 * static_cast<size_t>(ownStart + tid) + 1] = rb;
                }
            }
        }
        // ... and every other rank learns the verdicts: the batch's 64-bit accept mask, ONE 8-byte store per peer
        // (they hold the same trials)
        if (tid < SHARD_MAX_PEERS && tid < peers.n) {
            #pragma unroll
            for (int pr = 0; pr < SHARD_MAX_PEERS; pr++)      // (unrolled: no dynamic indexing of the argument)
                if (pr == tid) peers.acc[pr][ownStart] = sAccepted;
        }
        // every read of sAccepted (write-back, accept bytes) precedes the next batch's reset of it, and the records
        // written above are visible to the next batch's exact tests
        if (batch0 + SW_MAXOWN < numOwnAll) __syncthreads();
    }   // batches

    if (tid == 0) {
        atomicAdd(&counters[0], static_cast<unsigned long long>(numOwnAll));
        atomicAdd(&counters[1], static_cast<unsigned long long>(sStats[0]));
        atomicAdd(&counters[2], static_cast<unsigned long long>(sStats[1]));
        atomicAdd(&counters[3], static_cast<unsigned long long>(sStats[2]));
    }
}

}  // namespace rp
