// rp_shard.cu — see rp_shard.cuh.  NCCL is bound at run time (dlopen) so that librampack_b200.so has no link-time
// dependency on it: a single-GPU user never loads NCCL, and inside a PyTorch process the already loaded
// libnccl.so.2 is reused (RTLD_NOLOAD first) instead of a second copy.
#include "rp_shard.cuh"
#include "rp_sweep_kernel.cuh"
#include <dlfcn.h>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <nccl.h>

namespace rp {

namespace {
struct NcclApi {
    void *lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                              cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi nccl;

template <typename F>
void bind(F &fn, const char *name) {
    fn = reinterpret_cast<F>(dlsym(nccl.lib, name));
    if (!fn) throw rp_error(RP_ERR_STATE, std::string("NCCL symbol missing: ") + name);
}

void loadNccl() {
    if (nccl.lib) return;
    void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) throw rp_error(RP_ERR_STATE, std::string("cannot load NCCL (libnccl.so.2): ") + dlerror());
    nccl.lib = lib;
    bind(nccl.GetUniqueId, "ncclGetUniqueId");
    bind(nccl.CommInitRank, "ncclCommInitRank");
    bind(nccl.CommDestroy, "ncclCommDestroy");
    bind(nccl.Send, "ncclSend");
    bind(nccl.Recv, "ncclRecv");
    bind(nccl.GroupStart, "ncclGroupStart");
    bind(nccl.GroupEnd, "ncclGroupEnd");
    bind(nccl.AllReduce, "ncclAllReduce");
    bind(nccl.AllGather, "ncclAllGather");
    bind(nccl.Broadcast, "ncclBroadcast");
    bind(nccl.GetErrorString, "ncclGetErrorString");
}

#define RP_NCCL(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
    throw rp::rp_error(RP_ERR_CUDA, std::string(#call) + ": " + nccl.GetErrorString(r_)); } while (0)

ncclComm_t commOf(const Handle *h) { return static_cast<ncclComm_t>(h->shard->comm); }
}  // namespace

bool shardActive(const Handle *h) { return h->shard && h->shard->nranks > 1; }

void shardUniqueId(unsigned char id[128]) {
    loadNccl();
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId u;
    RP_NCCL(nccl.GetUniqueId(&u));
    std::memcpy(id, &u, 128);
}

void shardInit(Handle *h, const unsigned char id[128], int rank, int nranks) {
    RP_EXPECT(nranks >= 1 && rank >= 0 && rank < nranks, "0 <= rank < nranks");
    shardDestroy(h);
    h->shard = new Shard;
    h->shard->rank = rank;
    h->shard->nranks = nranks;
    if (nranks == 1) return;
    loadNccl();
    ncclUniqueId u;
    std::memcpy(&u, id, 128);
    ncclComm_t comm;
    RP_NCCL(nccl.CommInitRank(&comm, nranks, u, rank));
    h->shard->comm = comm;
    RP_CUDA(cudaMalloc(&h->shard->dCounts, 64 * sizeof(unsigned long long)));
}

void shardDestroy(Handle *h) {
    if (!h->shard) return;
    Shard *s = h->shard;
    for (int q = 0; q < 64; q++) {
        for (int b = 0; b < 2; b++) if (s->peerBase[q][b]) cudaIpcCloseMemHandle(s->peerBase[q][b]);
        if (s->peerFlags[q]) cudaIpcCloseMemHandle(s->peerFlags[q]);
    }
    if (s->flags) cudaFree(s->flags);
    if (s->acc) cudaFree(s->acc);
    if (s->dHandles) cudaFree(s->dHandles);
    if (s->comm) nccl.CommDestroy(static_cast<ncclComm_t>(s->comm));
    for (void *p : {(void *) s->sendBuf, (void *) s->recvBuf, (void *) s->gatherBuf, (void *) s->dLayerStart,
                    (void *) s->dCounts})
        if (p) cudaFree(p);
    delete s;
    h->shard = nullptr;
}

void shardSlab(int nActiveZ, int nranks, int r, int *aLo, int *aHi) {
    *aLo = static_cast<int>(static_cast<long long>(r) * nActiveZ / nranks);
    *aHi = static_cast<int>(static_cast<long long>(r + 1) * nActiveZ / nranks);
}

HaloPlan shardHaloPlan(int nz, int P, int r, int colour) {
    HaloPlan p;
    shardSlab(nz / 2, P, r, &p.aLo, &p.aHi);
    const int pz = (colour >> 2) & 1;
    // even-z colours changed my bottom layer, which rank r-1 reads next; odd-z colours my top layer (rank r+1)
    p.sendLayer = pz == 0 ? 2 * p.aLo : 2 * p.aHi - 1;
    p.peerSend = pz == 0 ? (r + P - 1) % P : (r + 1) % P;
    p.peerRecv = pz == 0 ? (r + 1) % P : (r + P - 1) % P;
    p.recvLayer = pz == 0 ? (2 * p.aHi) % nz : (2 * p.aLo - 1 + nz) % nz;
    return p;
}

__global__ void k_layer_starts(int nz, int layerCells, const int *__restrict__ cellStart, int *__restrict__ out) {
    int z = blockIdx.x * blockDim.x + threadIdx.x;
    if (z <= nz) out[z] = cellStart[z * layerCells];
}

void shardQueueLayerStarts(Handle *h, int nx, int ny, int nz, bool toHost) {
    Shard *s = h->shard;
    devEnsure(s->dLayerStart, s->layerCap, static_cast<size_t>(nz) + 1);
    k_layer_starts<<<(nz + 1 + 127) / 128, 128, 0, h->stream>>>(nz, nx * ny, h->sgrid.cellStart, s->dLayerStart);
    h->launches++;
    if (!toHost) return;
    s->layerStart.resize(nz + 1);
    RP_CUDA(cudaMemcpyAsync(s->layerStart.data(), s->dLayerStart, (nz + 1) * sizeof(int), cudaMemcpyDeviceToHost,
                            h->stream));
}

// accepted moves (meta.x == 4) of the slots [s0, s1); xyColour >= 0 keeps only cells of that x/y parity
__global__ void k_pack_moves(int s0, int s1, int xyColour, int nx, int ny, const int *__restrict__ items,
                             const int *__restrict__ keys, const uint2 *__restrict__ meta,
                             const double *__restrict__ rec, const double *__restrict__ trialOrient, int rotated,
                             MoveMsg *__restrict__ out)
{
    int slot = s0 + blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= s1) return;
    if (meta[slot].x != 4u) return;
    if (xyColour >= 0) {
        int cell = keys[items[slot]] & 0x0FFFFFFF;
        int cx = cell % nx, cy = (cell / nx) % ny;
        if (((cx & 1) | ((cy & 1) << 1)) != xyColour) return;
    }
    unsigned long long at = atomicAdd(reinterpret_cast<unsigned long long *>(&out[0].slot), 1ull);
    MoveMsg &m = out[1 + at];
    m.slot = slot;
    for (int q = 0; q < 6; q++) m.rec[q] = rec[6 * static_cast<size_t>(slot) + q];
    if (rotated) for (int q = 0; q < 9; q++) m.orient[q] = trialOrient[9 * static_cast<size_t>(slot) + q];
}

// the receiving side of the sweep kernel's write-back (same arithmetic: Shape::translate, core/Shape.cpp:9-12)
__global__ void k_apply_moves(const MoveMsg *__restrict__ in, int translated, int rotated, Box box, size_t cap,
                              const int *__restrict__ items, const int *__restrict__ keys, SweepGrid sg,
                              float invHalf, double *__restrict__ rec, float4 *__restrict__ rec32,
                              double *__restrict__ pos, double *__restrict__ orient, uint2 *__restrict__ meta)
{
    long long count = in[0].slot;
    long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const MoveMsg &m = in[1 + i];
    const int slot = static_cast<int>(m.slot), id = items[slot];
    for (int q = 0; q < 6; q++) rec[6 * static_cast<size_t>(slot) + q] = m.rec[q];
    {   // the FP32 cull record of the new state, by the expression its producer used (k_sweep_prepare)
        const int cell = keys[id] & 0x0FFFFFFF;
        const V3 centre = sweepCellCentre(sg, box, cell % sg.n[0], (cell / sg.n[0]) % sg.n[1],
                                          cell / (sg.n[0] * sg.n[1]));
        makeRec32(v3(m.rec[0], m.rec[1], m.rec[2]), v3(m.rec[3], m.rec[4], m.rec[5]), centre, invHalf,
                  rec32 + 2 * static_cast<size_t>(slot));
    }
    if (translated) {
        V3 np = v3(m.rec[0], m.rec[1], m.rec[2]);
        V3 p = np + pbcCorrection(box, np);
        pos[id] = p.x; pos[cap + id] = p.y; pos[2 * cap + id] = p.z;
    }
    if (rotated) for (int q = 0; q < 9; q++) orient[q * cap + id] = m.orient[q];
    meta[slot].x = 4u;
}

static void ensureMsgBuf(MoveMsg *&buf, size_t &cap, size_t need) {
    if (need <= cap && buf) return;
    if (buf) cudaFree(buf);
    buf = nullptr;
    size_t newCap = need + need / 4 + 64;
    RP_CUDA(cudaMalloc(&buf, (newCap + 1) * sizeof(MoveMsg)));
    cap = newCap;
}

static void pack(Handle *h, int s0, int s1, int xyColour, int nx, int ny, bool rotated, MoveMsg *out) {
    RP_CUDA(cudaMemsetAsync(out, 0, sizeof(MoveMsg), h->stream));
    if (s1 > s0) {
        k_pack_moves<<<(s1 - s0 + 255) / 256, 256, 0, h->stream>>>(s0, s1, xyColour, nx, ny, h->sgrid.items,
                                                                  h->sgrid.keys, static_cast<uint2 *>(h->trialMeta),
                                                                  h->sgrid.rec, h->trialOrient, rotated ? 1 : 0, out);
        h->launches++;
    }
}

static void apply(Handle *h, const MoveMsg *in, size_t capacity, bool translated, bool rotated) {
    if (capacity == 0) return;
    k_apply_moves<<<static_cast<int>((capacity + 255) / 256), 256, 0, h->stream>>>(
        in, translated ? 1 : 0, rotated ? 1 : 0, h->box, h->cap, h->sgrid.items, h->sgrid.keys, h->lastSg,
        h->lastInvHalf, h->sgrid.rec, sweepRec32(h->sgrid), h->pos, h->orient, static_cast<uint2 *>(h->trialMeta));
    h->launches++;
}

void shardHaloExchange(Handle *h, int colour, int nx, int ny, int nz, bool translated, bool rotated) {
    Shard *s = h->shard;
    const int P = s->nranks, r = s->rank;
    const HaloPlan plan = shardHaloPlan(nz, P, r, colour);
    const int sendLayer = plan.sendLayer, recvLayer = plan.recvLayer, peerSend = plan.peerSend,
              peerRecv = plan.peerRecv;
    const std::vector<int> &L = s->layerStart;
    const size_t sendCapNow = static_cast<size_t>(L[sendLayer + 1] - L[sendLayer]);
    const size_t recvCapNow = static_cast<size_t>(L[recvLayer + 1] - L[recvLayer]);
    ensureMsgBuf(s->sendBuf, s->sendCap, sendCapNow);
    ensureMsgBuf(s->recvBuf, s->recvCap, recvCapNow);
    pack(h, L[sendLayer], L[sendLayer + 1], colour & 3, nx, ny, rotated, s->sendBuf);
    const size_t dblPerMsg = sizeof(MoveMsg) / sizeof(double);
    RP_NCCL(nccl.GroupStart());
    RP_NCCL(nccl.Send(s->sendBuf, (sendCapNow + 1) * dblPerMsg, ncclDouble, peerSend, commOf(h), h->stream));
    RP_NCCL(nccl.Recv(s->recvBuf, (recvCapNow + 1) * dblPerMsg, ncclDouble, peerRecv, commOf(h), h->stream));
    RP_NCCL(nccl.GroupEnd());
    apply(h, s->recvBuf, recvCapNow, translated, rotated);
    s->lastHaloBytes += (sendCapNow + 1) * sizeof(MoveMsg);
}

void shardGatherCycle(Handle *h, int nz, bool translated, bool rotated) {
    Shard *s = h->shard;
    const int P = s->nranks, r = s->rank;
    const std::vector<int> &L = s->layerStart;
    std::vector<int> lo(P), hi(P);
    size_t total = 0;
    std::vector<size_t> offset(P);
    for (int q = 0; q < P; q++) {
        int aLo, aHi;
        shardSlab(nz / 2, P, q, &aLo, &aHi);
        lo[q] = L[2 * aLo];
        hi[q] = L[2 * aHi];
        offset[q] = total;
        total += static_cast<size_t>(hi[q] - lo[q]) + 1;      // + header
    }
    ensureMsgBuf(s->gatherBuf, s->gatherCap, total);
    MoveMsg *mine = s->gatherBuf + offset[r];
    pack(h, lo[r], hi[r], -1, 0, 0, rotated, mine);
    // message counts of all ranks -> host (exact broadcast sizes)
    RP_NCCL(nccl.AllGather(&mine[0].slot, s->dCounts, 1, ncclUint64, commOf(h), h->stream));
    unsigned long long counts[64];
    RP_CUDA(cudaMemcpyAsync(counts, s->dCounts, P * sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    const size_t dblPerMsg = sizeof(MoveMsg) / sizeof(double);
    RP_NCCL(nccl.GroupStart());
    for (int q = 0; q < P; q++) {
        MoveMsg *buf = s->gatherBuf + offset[q];
        RP_NCCL(nccl.Broadcast(buf, buf, (counts[q] + 1) * dblPerMsg, ncclDouble, q, commOf(h), h->stream));
        if (q != r) s->lastGatherBytes += (counts[q] + 1) * sizeof(MoveMsg);
    }
    RP_NCCL(nccl.GroupEnd());
    for (int q = 0; q < P; q++)
        if (q != r) apply(h, s->gatherBuf + offset[q], counts[q], translated, rotated);
}

// ---------------------------------------------------------------------------------------------------------------
// peer-memory path
// ---------------------------------------------------------------------------------------------------------------
// Flag block of a rank (one IPC-mapped allocation): [3][64] uint32 epochs (grid-ready / colour / sweep-done, indexed
// by SENDER rank), then [64][4] uint64 sweep counters of every rank.
constexpr size_t FLAG_WORDS = 4 * 64, FLAG_BLOCK_BYTES = FLAG_WORDS * sizeof(unsigned int) + 64 * 4 * sizeof(unsigned long long);
__host__ __device__ inline unsigned long long *counterBlock(unsigned int *flags) {
    return reinterpret_cast<unsigned long long *>(flags + FLAG_WORDS);
}

struct IpcBundle { cudaIpcMemHandle_t mem[4]; cudaIpcMemHandle_t flags; };

bool shardEnsurePeerMapping(Handle *h) {
    Shard *s = h->shard;
    const int P = s->nranks, r = s->rank;
    if (P > SHARD_MAX_PEERS + 1 || std::getenv("RP_SHARD_NCCL")) return false;
    if (s->p2pTried && !s->p2p) return false;
    if (s->accCap < h->sgrid.itemCap || !s->acc) {
        if (s->acc) cudaFree(s->acc);       // (peers close their mapping below, in this same collective call)
        s->acc = nullptr;
        RP_CUDA(cudaMalloc(&s->acc, h->sgrid.itemCap * sizeof(unsigned long long)));
        s->accCap = h->sgrid.itemCap;
    }
    void *local[4] = {h->sgrid.rec, s->acc, h->trialRec, h->trialOrient};
    if (s->p2p && std::memcmp(local, s->mappedLocal, sizeof(local)) == 0) return true;
    // (all ranks take this branch in the same call: the replicas make identical allocation decisions)
    s->p2pTried = true;
    s->p2p = false;
    if (!s->flags) {
        RP_CUDA(cudaMalloc(&s->flags, FLAG_BLOCK_BYTES));
        RP_CUDA(cudaMemset(s->flags, 0, FLAG_BLOCK_BYTES));
        RP_CUDA(cudaMalloc(&s->dHandles, 64 * sizeof(IpcBundle)));
    }
    for (int q = 0; q < P; q++)
        for (int b = 0; b < 4; b++)
            if (s->peerBase[q][b]) { cudaIpcCloseMemHandle(s->peerBase[q][b]); s->peerBase[q][b] = nullptr; }
    IpcBundle mine;
    bool ok = true;
    for (int b = 0; b < 4; b++) ok = ok && cudaIpcGetMemHandle(&mine.mem[b], local[b]) == cudaSuccess;
    ok = ok && cudaIpcGetMemHandle(&mine.flags, s->flags) == cudaSuccess;
    cudaGetLastError();
    std::vector<IpcBundle> all(P);
    unsigned long long okAll = ok ? 1 : 0;
    RP_CUDA(cudaMemcpyAsync(s->dHandles + r * sizeof(IpcBundle), &mine, sizeof(mine), cudaMemcpyHostToDevice,
                            h->stream));
    RP_NCCL(nccl.AllGather(s->dHandles + r * sizeof(IpcBundle), s->dHandles, sizeof(IpcBundle), ncclUint8,
                           commOf(h), h->stream));
    RP_CUDA(cudaMemcpyAsync(all.data(), s->dHandles, P * sizeof(IpcBundle), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    for (int q = 0; q < P && ok; q++) {
        if (q == r) continue;
        for (int b = 0; b < 4 && ok; b++)
            ok = cudaIpcOpenMemHandle(&s->peerBase[q][b], all[q].mem[b], cudaIpcMemLazyEnablePeerAccess)
                 == cudaSuccess;
        if (ok && !s->peerFlags[q])
            ok = cudaIpcOpenMemHandle(reinterpret_cast<void **>(&s->peerFlags[q]), all[q].flags,
                                      cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
    }
    cudaGetLastError();
    // everyone or no one: a rank that cannot map must not be written to through stale mappings
    okAll = ok ? 1 : 0;
    RP_CUDA(cudaMemcpyAsync(s->dCounts, &okAll, sizeof(okAll), cudaMemcpyHostToDevice, h->stream));
    RP_NCCL(nccl.AllReduce(s->dCounts, s->dCounts, 1, ncclUint64, ncclMin, commOf(h), h->stream));
    RP_CUDA(cudaMemcpyAsync(&okAll, s->dCounts, sizeof(okAll), cudaMemcpyDeviceToHost, h->stream));
    RP_CUDA(cudaStreamSynchronize(h->stream));
    s->p2p = okAll != 0;
    std::memcpy(s->mappedLocal, local, sizeof(local));
    return s->p2p;
}

PeerView shardPeerView(const Handle *h, int nLayersZ) {
    PeerView v;
    const Shard *s = h->shard;
    const int P = s->nranks, r = s->rank;
    for (int q = 0; q < P; q++) {
        if (q == r) continue;
        v.acc[v.n++] = static_cast<unsigned long long *>(s->peerBase[q][1]);
    }
    int aLo, aHi;
    shardSlab(nLayersZ / 2, P, r, &aLo, &aHi);
    v.layerLo = 2 * aLo;
    v.layerHi = 2 * aHi - 1;
    v.recLo = static_cast<double *>(s->peerBase[(r + P - 1) % P][0]);
    v.recHi = static_cast<double *>(s->peerBase[(r + 1) % P][0]);
    // the cull records live behind the FP64 records of the same allocation (sweepRec32); every replica makes the
    // same allocation decisions, so the offset is the local one
    v.rec32Lo = reinterpret_cast<float4 *>(v.recLo + 6 * h->sgrid.itemCap);
    v.rec32Hi = reinterpret_cast<float4 *>(v.recHi + 6 * h->sgrid.itemCap);
    return v;
}

// the write-back of k_sweep (rp_sweep_kernel2.cuh) for the accepted trials of the OTHER slabs.  A rank generates the
// trials of its own slab only (k_sweep_prepare), so the accepted states are PULLED from the owner's trial buffers
// (IPC-mapped peer memory: 48 + 72 contiguous bytes per accepted slot over NVLink).
struct ApplySources {
    int n;                                   // ranks
    int layerLo[SHARD_MAX_PEERS + 1], layerHi[SHARD_MAX_PEERS + 1];     // cell layers [lo, hi) of rank q's slab
    const double *rec[SHARD_MAX_PEERS + 1], *orient[SHARD_MAX_PEERS + 1];
};
__global__ void k_apply_accepts(int n, const int *__restrict__ layerStart, int layerLo, int layerHi,
                                const unsigned long long *__restrict__ acc, const int *__restrict__ keys,
                                const int *__restrict__ cellStart, ApplySources src,
                                int translated, int rotated, Box box, size_t cap, const int *__restrict__ items,
                                double *__restrict__ rec, double *__restrict__ pos, double *__restrict__ orient,
                                uint2 *__restrict__ meta)
{
    int slot = blockIdx.x * blockDim.x + threadIdx.x;
    const int slotBegin = layerStart[layerLo], slotEnd = layerStart[layerHi];      // my own slab: already written
    if (slot >= n || (slot >= slotBegin && slot < slotEnd)) return;
    const int id = items[slot];
    // the accept word of the slot's batch sits at the batch's first slot (k_sweep: batches of SW_MAXOWN = 64 trials)
    const int cs = cellStart[keys[id] & 0x0FFFFFFF];
    const int first = cs + ((slot - cs) & ~63);
    if (!((acc[first] >> (slot - first)) & 1ull)) return;
    // the owner of the slot
    const double *trialRec = nullptr, *trialOrient = nullptr;
    #pragma unroll
    for (int q = 0; q < SHARD_MAX_PEERS + 1; q++)
        if (q < src.n && slot >= layerStart[src.layerLo[q]] && slot < layerStart[src.layerHi[q]]) {
            trialRec = src.rec[q];
            trialOrient = src.orient[q];
        }
    if (!trialRec) return;
    const double *t = trialRec + 6 * static_cast<size_t>(slot);
    double tr[6];
    for (int q = 0; q < 6; q++) tr[q] = t[q];
    for (int q = 0; q < 6; q++) rec[6 * static_cast<size_t>(slot) + q] = tr[q];
    if (translated) {
        V3 np = v3(tr[0], tr[1], tr[2]);
        V3 p = np + pbcCorrection(box, np);
        pos[id] = p.x; pos[cap + id] = p.y; pos[2 * cap + id] = p.z;
    }
    if (rotated) for (int q = 0; q < 9; q++) orient[q * cap + id] = trialOrient[9 * static_cast<size_t>(slot) + q];
    meta[slot].x = 4u;
}

void shardApplyAccepts(Handle *h, int aLo, int aHi, int n, bool translated, bool rotated, int nActiveZ,
                       unsigned int epoch) {
    const Shard *s = h->shard;
    ApplySources src;
    src.n = s->nranks;
    for (int q = 0; q < s->nranks; q++) {
        int lo, hi;
        shardSlab(nActiveZ, s->nranks, q, &lo, &hi);
        src.layerLo[q] = 2 * lo;
        src.layerHi[q] = 2 * hi;
        src.rec[q] = q == s->rank ? h->trialRec : static_cast<const double *>(s->peerBase[q][2]);
        src.orient[q] = q == s->rank ? h->trialOrient : static_cast<const double *>(s->peerBase[q][3]);
    }
    k_apply_accepts<<<(n + 255) / 256, 256, 0, h->stream>>>(n, h->shard->dLayerStart, 2 * aLo, 2 * aHi,
                                                           h->shard->acc, h->sgrid.keys, h->sgrid.cellStart, src,
                                                           translated ? 1 : 0,
                                                           rotated ? 1 : 0, h->box, h->cap, h->sgrid.items,
                                                           h->sgrid.rec, h->pos,
                                                           h->orient, static_cast<uint2 *>(h->trialMeta));
    h->launches++;
    // nobody overwrites its trial buffers (the next sweep's k_sweep_prepare) before every rank has pulled from them
    shardSignalWait(h, SHARD_FLAG_APPLIED, epoch, true, false, nullptr);
}

struct FlagTargets {
    int n;
    unsigned int *to[SHARD_MAX_PEERS];     // flag word on the peer that I write
    unsigned int *from[SHARD_MAX_PEERS];   // local flag word the peer writes
    unsigned long long *cntTo[SHARD_MAX_PEERS];   // (sweep-done only) where my 4 counters go on the peer
};

// One launch per synchronisation point: tell the peers "everything I queued before this kernel is done" (the peer
// writes of earlier kernels are complete at their kernel end; the system-scope fence keeps the flag behind them),
// then poll until every peer of the set said the same.
// The poll is bounded (SHARD_WAIT_TIMEOUT_NS of wall clock, ~20 s): a peer that died or left the call sequence must
// not leave this GPU spinning for ever.  A timeout raises *abortFlag, which the host checks at its next
// synchronisation point (RP_ERR_STATE "peer timeout"); the sweep's results are then invalid.
constexpr unsigned long long SHARD_WAIT_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;
__global__ void k_signal_wait(FlagTargets t, unsigned int value, const unsigned long long *__restrict__ counters,
                              int *__restrict__ abortFlag) {
    if (threadIdx.x < t.n) {
        if (counters)
            for (int j = 0; j < 4; j++) t.cntTo[threadIdx.x][j] = counters[j];
        __threadfence_system();
        *reinterpret_cast<volatile unsigned int *>(t.to[threadIdx.x]) = value;
        volatile unsigned int *f = t.from[threadIdx.x];
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        unsigned int spins = 0;
        while (static_cast<int>(*f - value) < 0) {
            __nanosleep(100);
            if ((++spins & 0x3ffu) == 0u) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (now - t0 > SHARD_WAIT_TIMEOUT_NS || *reinterpret_cast<volatile int *>(abortFlag)) {
                    atomicExch(abortFlag, 1);
                    break;
                }
            }
        }
    }
    __threadfence_system();
}

// dCounters[j] = sum over ranks (mine + what the peers stored with their sweep-done signal)
__global__ void k_sum_counters(int nranks, int rank, const unsigned long long *__restrict__ block,
                               unsigned long long *__restrict__ counters)
{
    int j = threadIdx.x;
    if (j >= 4) return;
    unsigned long long sum = counters[j];
    for (int q = 0; q < nranks; q++) if (q != rank) sum += block[4 * q + j];
    counters[j] = sum;
}

static int peerSet(const Shard *s, bool all, int out[SHARD_MAX_PEERS]) {
    int n = 0;
    const int P = s->nranks, r = s->rank;
    if (all) { for (int q = 0; q < P; q++) if (q != r) out[n++] = q; return n; }
    int lo = (r + P - 1) % P, hi = (r + 1) % P;
    out[n++] = lo;
    if (hi != lo) out[n++] = hi;
    return n;
}

void shardSignalWait(Handle *h, ShardFlag kind, unsigned int value, bool all, bool withCounters,
                     unsigned long long *dCounters) {
    if (!dCounters) dCounters = h->dCounters;
    Shard *s = h->shard;
    int q[SHARD_MAX_PEERS];
    FlagTargets t;
    t.n = peerSet(s, all, q);
    for (int i = 0; i < t.n; i++) {
        t.to[i] = s->peerFlags[q[i]] + kind * 64 + s->rank;
        t.from[i] = s->flags + kind * 64 + q[i];
        t.cntTo[i] = counterBlock(s->peerFlags[q[i]]) + 4 * s->rank;
    }
    k_signal_wait<<<1, 32, 0, h->stream>>>(t, value, withCounters ? dCounters : nullptr, &h->dFlags[SHARD_ABORT_FLAG]);
    h->launches++;
    if (withCounters) {
        k_sum_counters<<<1, 32, 0, h->stream>>>(s->nranks, s->rank, counterBlock(s->flags), dCounters);
        h->launches++;
    }
}

// slice of rank q in the standard partition of n particles over P ranks
void shardSlice(unsigned long long n, int P, int q, unsigned long long *first, unsigned long long *count) {
    const unsigned long long a = n * static_cast<unsigned long long>(q) / P, b = n * static_cast<unsigned long long>(q + 1) / P;
    *first = a;
    *count = b - a;
}

// every rank holds its own slice of the canonical SoA arrays: broadcast the 12 rows of every slice from its owner
void shardAllGatherState(Handle *h, unsigned long long n) {
    Shard *s = h->shard;
    RP_NCCL(nccl.GroupStart());
    for (int q = 0; q < s->nranks; q++) {
        unsigned long long first, count;
        shardSlice(n, s->nranks, q, &first, &count);
        if (count == 0) continue;
        for (int row = 0; row < 12; row++) {
            double *base = row < 3 ? h->pos + row * h->cap : h->orient + (row - 3) * h->cap;
            RP_NCCL(nccl.Broadcast(base + first, base + first, count, ncclDouble, q, commOf(h), h->stream));
        }
    }
    RP_NCCL(nccl.GroupEnd());
}

void shardAllReduceMaxInt(Handle *h, int *dValue) {
    RP_NCCL(nccl.AllReduce(dValue, dValue, 1, ncclInt32, ncclMax, commOf(h), h->stream));
}

void shardAllReduceU64(Handle *h, unsigned long long *dValues, int count) {
    RP_NCCL(nccl.AllReduce(dValues, dValues, count, ncclUint64, ncclSum, commOf(h), h->stream));
}

}  // namespace rp
