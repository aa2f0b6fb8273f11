// rp_shard.cuh — spatial sharding of the sweep over the GPUs of one node (north_star: "per-sweep halo exchange over
// NCCL/NVLink and a device-side all-reduce for the volume-move overlap count").  The reference has no analogue (one
// process, OpenMP); its semantic model is the domain decomposition of Simulation::performMovesWithDomainDivision
// (core/Simulation.cpp:389-441): domains that are far enough apart move concurrently.
//
// Round-1 design: the STATE is replicated (every rank holds all particles and builds the same sweep grid from the same
// Philox key — bit-identical on every rank), the COMPUTE is sharded: the cell layers along z are cut into one slab per
// rank (an even number of layers, so both parities are present), rank r runs the active cells of its slab only.
//   * after every colour: the accepted moves of the ONE boundary layer a neighbour will read next are sent to that
//     neighbour (even-z colours: bottom layer -> rank r-1, odd-z colours: top layer -> rank r+1; ncclSend/ncclRecv,
//     fixed-capacity buffers, no host synchronisation);
//   * after the eighth colour: every rank broadcasts the accepted moves of its slab (exact sizes, one count exchange)
//     so that the replicated state is identical again before the next grid build / box move / download.
// Result: an N-GPU run is bit-identical to the 1-GPU run with the same seed (tests/test_gpu_shard.py).
#pragma once

#include "rp_internal.cuh"

namespace rp {

struct MoveMsg {                 // 128 bytes; element 0 of every buffer is the header: slot = number of messages
    long long slot;
    double rec[6];               // new sorted record {P'(3), half axis(3)}
    double orient[9];            // new orientation (rotating samplers only)
};

struct Shard {
    void *comm = nullptr;        // ncclComm_t
    int rank = 0, nranks = 1;
    MoveMsg *sendBuf = nullptr, *recvBuf = nullptr, *gatherBuf = nullptr;
    size_t sendCap = 0, recvCap = 0, gatherCap = 0;     // in messages (without header)
    int *dLayerStart = nullptr;  // slot at which cell layer z begins, z = 0..nz
    size_t layerCap = 0;
    std::vector<int> layerStart; // host copy
    unsigned long long *dCounts = nullptr;    // nranks message counts (all-gathered)
    // ---- peer-memory path (default when the buffers of all ranks can be IPC-mapped; RP_SHARD_NCCL=1 disables it) ----
    // A rank generates the trials of its own slab.  k_sweep stores the accept word of every batch straight into every
    // peer's `acc` array over NVLink; the records of accepted moves in the slab's two boundary layers also go
    // straight into the z-neighbour's sorted records (read there in the next colour).  Per-colour / per-sweep flags
    // (k_signal_wait) order the kernels of different GPUs; after the last colour k_apply_accepts pulls the accepted
    // states of the other slabs from their owners' trial buffers (mapped like the records).
    bool p2p = false, p2pTried = false;
    // accept words: the 64-bit accept mask of a batch of <= 64 trials of one cell, stored at the sorted slot of the
    // batch's first trial by the cell's owner rank (one 8-byte NVLink store per cell and peer; one BYTE per slot — 23
    // one-byte remote stores per cell and peer — was the round-1 form)
    unsigned long long *acc = nullptr;
    size_t accCap = 0;
    void *mappedLocal[4] = {nullptr, nullptr, nullptr, nullptr};    // local rec / acc / trialRec / trialOrient at mapping time
    void *peerBase[64][4] = {};      // IPC-mapped rec / acc / trial records / trial orientations of rank q (nullptr for q == rank)
    unsigned int *flags = nullptr;   // local flag block: [3][64] = grid-ready / colour / sweep-done epochs per sender
    unsigned int *peerFlags[64] = {};
    unsigned char *dHandles = nullptr;    // all-gather staging for the IPC handles
    unsigned int sweepEpoch = 0;
    // timing of the last sharded sweep (ms, CUDA events on the handle's stream)
    double lastHaloMs = 0, lastGatherMs = 0;
    unsigned long long lastHaloBytes = 0, lastGatherBytes = 0;
};

bool shardActive(const Handle *h);
void shardUniqueId(unsigned char id[128]);
void shardInit(Handle *h, const unsigned char id[128], int rank, int nranks);
void shardDestroy(Handle *h);
// slab of rank r in active-cell layers: [aLo, aHi) of nActiveZ
void shardSlab(int nActiveZ, int nranks, int r, int *aLo, int *aHi);
// downloads the layer boundaries of the current sweep grid (call after the grid build, before the host sync)
// who sends which cell layer to whom after a sub-sweep of `colour` (bit 2 = z parity); nz = cell layers along z
struct HaloPlan { int aLo, aHi, sendLayer, recvLayer, peerSend, peerRecv; };
HaloPlan shardHaloPlan(int nz, int nranks, int r, int colour);
// layer boundaries of the current sweep grid -> device array (always) and, toHost, the host copy (valid after the
// next stream synchronisation; the NCCL path sizes its messages from it)
void shardQueueLayerStarts(Handle *h, int nx, int ny, int nz, bool toHost);
void shardHaloExchange(Handle *h, int colour, int nx, int ny, int nz, bool translated, bool rotated);
void shardGatherCycle(Handle *h, int nz, bool translated, bool rotated);
void shardAllReduceU64(Handle *h, unsigned long long *dValues, int count);
// sharded upload: standard partition of n particles over the ranks, all-gather of the canonical arrays over NCCL
void shardSlice(unsigned long long n, int nranks, int rank, unsigned long long *first, unsigned long long *count);
void shardAllGatherState(Handle *h, unsigned long long n);
void shardAllReduceMaxInt(Handle *h, int *dValue);

// peer-memory path
constexpr int SHARD_MAX_PEERS = 7;
struct PeerView {                 // by-value kernel argument of k_sweep
    int n = 0;                    // number of peers
    unsigned long long *acc[SHARD_MAX_PEERS]; // their accept-word arrays (indexed by the first slot of a batch)
    int layerLo = -1, layerHi = -1;           // my slab's bottom / top cell layer
    double *recLo = nullptr, *recHi = nullptr;    // sorted records of the rank below / above (read my boundary layers)
    float4 *rec32Lo = nullptr, *rec32Hi = nullptr;    // ... and their FP32 cull records (same allocation, behind the FP64 records)
};
enum ShardFlag { SHARD_FLAG_GRID = 0, SHARD_FLAG_COLOUR = 1, SHARD_FLAG_DONE = 2, SHARD_FLAG_APPLIED = 3 };
constexpr int SHARD_ABORT_FLAG = 12;     // index into Handle::dFlags: a peer-flag wait timed out (never reset by a sweep)
// (re)maps the peers' buffers when a local buffer moved since the last mapping; collective; returns whether the
// peer-memory path is usable
bool shardEnsurePeerMapping(Handle *h);
PeerView shardPeerView(const Handle *h, int nLayersZ);
// accepted trials of the slots outside [slotBegin, slotEnd) -> sorted records and canonical arrays
// (the slab's slot range is read from the device copy of the layer boundaries: layers [2 aLo, 2 aHi))
void shardApplyAccepts(Handle *h, int aLo, int aHi, int n, bool translated, bool rotated, int nActiveZ,
                       unsigned int epoch);
// One launch: flag[kind][my rank] = value on the peers (everyone, or the two z-neighbours), ordered after all prior
// work of the stream, then poll the local flags of the same peers until they reach `value`.  withCounters (sweep
// done): my four sweep counters travel with the signal and h->dCounters becomes the sum over ranks.
void shardSignalWait(Handle *h, ShardFlag kind, unsigned int value, bool all, bool withCounters,
                     unsigned long long *dCounters = nullptr);

}  // namespace rp
