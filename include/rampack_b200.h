/*
 * rampack_b200.h — C ABI of the B200-native Monte Carlo sweep (librampack_b200.so).
 *
 * The reference (PKua007/rampack) has no FFI layer: its boundary for this path is the C++ contract
 * Simulation -> Packing -> Interaction (src/core/Packing.h:143-390, src/core/Interaction.h:28-130).  This header is
 * the C ABI a maintainer binds *underneath* the reference's `Packing`.  Every entry point names
 * the reference method it replaces.  Plain pointers and sizes only; no C++/torch types cross the boundary.
 *
 * Conventions
 *   - all functions return an rp_status; rp_last_error(h) gives the message (thread-local when h == NULL)
 *   - one host thread per handle; all device work of a handle runs on the handle's CUDA stream
 *   - shapes cross the boundary in the reference's `Shape` layout (src/core/Shape.h:20-24):
 *     12 doubles per particle = position[3] followed by the row-major orientation matrix[9]
 *   - box matrices are row-major 3x3, COLUMNS are the box vectors (src/core/TriclinicBox.h:20-23)
 *   - "+inf" / "-inf" return conventions of tryMove / tryScaling are kept (src/core/Packing.h:306-352)
 *   - there is no CPU fallback: every compute entry point fails with RP_ERR_CUDA when no device is usable
 */
#ifndef RAMPACK_B200_H
#define RAMPACK_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rp_handle rp_handle;

typedef enum rp_status {
    RP_OK = 0,
    RP_ERR_PRECONDITION = 1,   /* reference: PreconditionException from Expects() (src/utils/Exceptions.h:19-66) */
    RP_ERR_CUDA = 2,           /* CUDA runtime failure or no usable device */
    RP_ERR_STATE = 3,          /* reference: std::runtime_error (e.g. un-cached overlap count, Packing.cpp:1059) */
    RP_ERR_UNSUPPORTED = 4,    /* interaction / mode not implemented on the device yet */
    RP_ERR_CAPACITY = 5        /* a device-side buffer bound was exceeded (reported, never silently truncated) */
} rp_status;

/* Hard part of the interaction: replaces Interaction::overlapBetween of the named ShapeTraits */
typedef enum rp_hard_kind {
    RP_HARD_NONE = 0,
    RP_HARD_SPHERE = 1,              /* SphereTraits::HardInteraction        (src/core/shapes/SphereTraits.cpp:56-75) */
    RP_HARD_SPHEROCYLINDER = 2,      /* SpherocylinderTraits                 (src/core/shapes/SpherocylinderTraits.cpp:28-73) */
    RP_HARD_POLYSPHERE = 3,          /* PolysphereTraits::HardInteraction    (src/core/shapes/PolysphereTraits.cpp:97-136) */
    RP_HARD_POLYSPHEROCYLINDER = 4,  /* PolyspherocylinderTraits             (src/core/shapes/PolyspherocylinderTraits.cpp:109-170) */
    RP_HARD_XENOCOLLIDE = 5          /* XenoCollideTraits<...> (generic_convex, smooth_wedge, polyhedral_wedge:
                                        src/core/shapes/XenoCollideTraits.h:147-171, geometry/xenocollide/XenoCollide.h) */
} rp_hard_kind;

/* Soft part: replaces CentralInteraction::calculateEnergyBetween (src/core/interactions/CentralInteraction.h:58-67) */
typedef enum rp_soft_kind {
    RP_SOFT_NONE = 0,
    RP_SOFT_LJ = 1,    /* LennardJonesInteraction.cpp:8-13           range 3*sigma            */
    RP_SOFT_WCA = 2,   /* RepulsiveLennardJonesInteraction.cpp:8-16  range 2^(1/6)*sigma      */
    RP_SOFT_SIC = 3    /* SquareInverseCoreInteraction.cpp:13-15     range sigma              */
} rp_soft_kind;

#define RP_MAX_CENTRES 16

/*
 * XenoCollide geometry as a flat program.  The reference builds a tree of AbstractXCGeometry objects
 * (src/geometry/xenocollide/XCPrimitives.h, XCOperations.h, XCBodyBuilder.cpp); a `describe()` walk over that tree
 * emits one rp_xc_node per object (children before parents) and one rp_xc_entry per child slot of XCSum / XCDiff /
 * XCMax ({geometry, pos, rot} in the order of the object's `entries` vector; XCDiff: geom1 then geom2).
 * params are the constructor arguments of the reference class, in its order:
 *   POINT x y z | SEGMENT halfLength | RECTANGLE halfSideX halfSideY | CUBOID halfSides[3] | DISK radius |
 *   SPHERE radius | ELLIPSE semiAxisX semiAxisY | ELLIPSOID semiAxes[3] | FOOTBALL halfLength radius |
 *   SAUCER radius halfThickness | SMOOTH_WEDGE R r l (SmoothWedgeTraits::CollideGeometry) |
 *   POLYHEDRAL_WEDGE axBottom ayBottom axTop ayTop length (PolyhedralWedgeTraits::CollideGeometry)
 * Circumsphere / insphere radii and centres are derived by the library with the reference's expressions.
 */
typedef enum rp_xc_kind {
    RP_XC_POINT = 0, RP_XC_SEGMENT = 1, RP_XC_RECTANGLE = 2, RP_XC_CUBOID = 3, RP_XC_DISK = 4, RP_XC_SPHERE = 5,
    RP_XC_ELLIPSE = 6, RP_XC_ELLIPSOID = 7, RP_XC_FOOTBALL = 8, RP_XC_SAUCER = 9, RP_XC_SMOOTH_WEDGE = 10,
    RP_XC_POLYHEDRAL_WEDGE = 11,
    RP_XC_SUM = 16,   /* XCSum  (XCOperations.cpp:38-73)  */
    RP_XC_DIFF = 17,  /* XCDiff (XCOperations.cpp:75-99)  */
    RP_XC_MAX = 18    /* XCMax  (XCOperations.cpp:101-185), the `wrap` of XCBodyBuilder */
} rp_xc_kind;
#define RP_XC_MAX_NODES 32
#define RP_XC_MAX_ENTRIES 32
typedef struct rp_xc_node {
    int32_t kind;          /* rp_xc_kind */
    int32_t n_entries;     /* operations: number of child entries */
    int32_t first_entry;   /* operations: index of the first child entry */
    int32_t reserved;
    double params[6];
} rp_xc_node;
typedef struct rp_xc_entry {
    int32_t node;          /* child geometry: index of an EARLIER node */
    int32_t reserved;
    double pos[3];
    double rot[9];         /* row-major */
} rp_xc_entry;
typedef struct rp_xc_program {
    int32_t n_nodes, n_entries;
    rp_xc_node nodes[RP_XC_MAX_NODES];
    rp_xc_entry entries[RP_XC_MAX_ENTRIES];
    int32_t root[RP_MAX_CENTRES];   /* getCollideGeometry(i): node of interaction centre i (root[0] when n_centres == 0) */
} rp_xc_program;

/*
 * POD description of an Interaction (the reference hides these parameters behind a vtable; a `describe()` on the
 * host ShapeTraits emits this — INTEGRATION.md).  Semantics of CompoundInteraction (CompoundInteraction.cpp:5-66)
 * are expressed by setting both a hard and a soft kind.
 */
typedef struct rp_interaction_desc {
    int32_t hard_kind;                 /* rp_hard_kind */
    int32_t soft_kind;                 /* rp_soft_kind */
    double radius;                     /* sphere / spherocylinder radius */
    double length;                     /* spherocylinder axis length (cap centre to cap centre) */
    double epsilon;                    /* soft: energy scale */
    double sigma;                      /* soft: length scale */
    int32_t n_centres;                 /* Interaction::getInteractionCentres().size(); 0 = one centre in the origin */
    int32_t reserved;
    double centres[RP_MAX_CENTRES][3];         /* body-frame centre positions */
    double centre_radius[RP_MAX_CENTRES];      /* polysphere bead radius / polyspherocylinder radius */
    double centre_half_axis[RP_MAX_CENTRES][3];/* polyspherocylinder body-frame half axis */
    const rp_xc_program *xc;           /* RP_HARD_XENOCOLLIDE: the collide geometries (copied by rp_set_interaction) */
} rp_interaction_desc;

typedef struct rp_config {
    int32_t device;        /* CUDA device ordinal */
    int32_t reserved;
    uint64_t seed;         /* Philox key for rp_sweep (Simulation's `seed`, src/core/Simulation.cpp:66-68) */
    void *cuda_stream;     /* cudaStream_t to run on, or NULL for a private stream */
} rp_config;

/* Geometry of the neighbour grid — NeighbourGrid::getCellDivisions (src/core/NeighbourGrid.cpp:372-374) */
typedef struct rp_grid_info {
    int32_t has_grid;            /* 0: reference would run without NG (Packing.cpp:868-881) */
    int32_t cell_divisions[3];   /* real cells per axis (without the 2 ghost layers) */
    double cell_size;            /* the requested cell size (Packing.cpp:862-866) */
    uint64_t num_centres;        /* entries binned */
    uint32_t max_cell_population;
    uint32_t reserved;
} rp_grid_info;

/* One explicit trial move — the arguments of Packing::tryTranslation / tryRotation / tryMove (Packing.cpp:84-164) */
typedef enum rp_move_kind {
    RP_MOVE_TRANSLATION = 0, RP_MOVE_ROTATION = 1, RP_MOVE_ROTOTRANSLATION = 2,
    RP_MOVE_FLIP = 3     /* rp_sweep only: FlipSampler (src/core/move_samplers/FlipSampler.cpp:18-41) */
} rp_move_kind;

typedef struct rp_move {
    uint64_t particle;       /* particleIdx */
    int32_t kind;            /* rp_move_kind */
    int32_t reserved;
    double translation[3];
    double rotation[9];      /* row-major rotation matrix, applied as R * orientation (Shape.cpp:14-16) */
} rp_move;

/* Sampler description for a bulk sweep — MoveSampler semantics (src/core/move_samplers/*.cpp) */
typedef struct rp_sweep_params {
    int32_t kind;                 /* rp_move_kind: translation / rotation / rototranslation sampler */
    int32_t pass;                 /* index of the sampler inside the cycle: part of the Philox counter */
    double translation_step;      /* already clamped by the caller to min box height / 2 (TranslationSampler.cpp:20-37) */
    double rotation_step;         /* already clamped to pi (RototranslationSampler.cpp:33) */
    double temperature;           /* Metropolis: u <= exp(-dE/T) (Simulation.cpp:483) */
    uint64_t cycle;               /* cycle number: part of the Philox counter, makes runs reproducible */
    /* RP_MOVE_FLIP: half-turn about the shape's flip axis (ShapeGeometry::findFlipAxis, body frame), followed by the
     * translation that puts the geometric origin (ShapeGeometry::getGeometricOrigin, body frame) back in place.  The
     * reference attempts N / flip_every flips of randomly drawn particles per cycle; here every particle attempts one
     * with probability 1 / flip_every (same expected number, no particle twice). */
    double flip_axis[3];
    double flip_origin[3];
    uint64_t flip_every;
} rp_sweep_params;

typedef struct rp_sweep_counters {
    uint64_t attempted;           /* trial moves attempted in this call (== N, Simulation.cpp:431-437) */
    uint64_t accepted;
    uint64_t rejected_domain;     /* rejected because the mass centre left the active region (Packing.cpp:95-96) */
    uint64_t rejected_overlap;
    uint64_t rejected_energy;     /* rejected by the Metropolis test on a finite energy difference */
    uint64_t sub_sweeps;          /* checkerboard colours run */
    uint64_t kernel_launches;     /* device kernels launched by this call */
} rp_sweep_counters;

/* One explicit pair for the verdict/energy parity harness — arguments of Interaction::overlapBetween */
typedef struct rp_pair {
    double pos1[3]; double orientation1[9]; uint64_t idx1;
    double pos2[3]; double orientation2[9]; uint64_t idx2;
    double translation[3];        /* BoundaryConditions::getTranslation result (HardcodedTranslation, Packing.cpp:23-43) */
} rp_pair;

/* ---- lifetime: replaces Packing ctor / dtor (src/core/Packing.cpp:46-65) ---- */
rp_status rp_create(const rp_config *config, rp_handle **out);
void rp_destroy(rp_handle *h);
const char *rp_last_error(const rp_handle *h);
const char *rp_version(void);

/* ---- state: Packing::reset / setupForInteraction / begin..end (Packing.cpp:67-82, 931-950) ---- */
rp_status rp_set_interaction(rp_handle *h, const rp_interaction_desc *desc);
rp_status rp_upload(rp_handle *h, const double box[9], const double *shapes /* N x 12 */, uint64_t n);
rp_status rp_download(rp_handle *h, double box_out[9], double *shapes_out /* N x 12 */, uint64_t n);
/* Sharded upload (after rp_comm_init): every rank passes its slice [n * rank / n_ranks, n * (rank + 1) / n_ranks) of
 * ONE packing of n particles; the slices travel host -> device once (1 / n_ranks of the bytes per rank) and are
 * all-gathered over NVLink (NCCL).  Collective: every rank of the communicator must call it.  Without a communicator
 * the slice is the whole packing.  The matching read-back of a slice is rp_download_range. */
rp_status rp_upload_slice(rp_handle *h, const double box[9], const double *shapes /* count x 12 */, uint64_t first,
                          uint64_t count, uint64_t n);
rp_status rp_size(const rp_handle *h, uint64_t *n_out);
/* shapes [first, first + count) only — Packing::operator[] / front / back after a few explicit moves (Packing.h:193-215) */
rp_status rp_download_range(rp_handle *h, uint64_t first, uint64_t count, double *shapes_out /* count x 12 */);
/* Snapshot without stalling the sweep — the producer side of SimulationRecorder / SnapshotWriter
 * (src/core/Simulation.cpp:141-152, src/core/io/RamsnapWriter.cpp:35-47, RamtrjIO.cpp writeShape): the state of this
 * moment is copied on the device (ordered with the work queued so far), then moved to `shapes_out` (pinned host
 * memory for a truly asynchronous copy) on a second stream while later cycles run.  `box_out` is valid on return;
 * `shapes_out` after rp_snapshot_wait, or once rp_snapshot_done reports 1.  One snapshot in flight per handle: a
 * second request waits (on the device) for the first copy to finish. */
rp_status rp_snapshot_async(rp_handle *h, double box_out[9], double *shapes_out /* N x 12 */, uint64_t n);
rp_status rp_snapshot_wait(rp_handle *h);
rp_status rp_snapshot_done(rp_handle *h, int32_t *done_out);

/* ---- neighbour grid: Packing::rebuildNeighbourGrid (Packing.cpp:858-929), NeighbourGrid::positionToCellNo ---- */
rp_status rp_rebuild_grid(rp_handle *h);
rp_status rp_grid_info_get(rp_handle *h, rp_grid_info *out);
/* cell number of every interaction centre, in the reference's numbering incl. ghost layers (NeighbourGrid.cpp:13-34) */
rp_status rp_cell_numbers(rp_handle *h, uint64_t *cell_no_out /* num_centres */);

/* ---- full-system recount: Packing::countTotalOverlaps / getTotalEnergy (Packing.cpp:519-559, 690-709) ---- */
rp_status rp_count_total_overlaps(rp_handle *h, int early_exit, uint64_t *count_out);
rp_status rp_total_energy(rp_handle *h, double *energy_out);
/* Packing::countWallOverlaps (Packing.cpp:561-580); early_exit != 0: 1 when any wall overlap exists */
rp_status rp_count_wall_overlaps(rp_handle *h, int early_exit, uint64_t *count_out);
/* Packing::getParticleEnergyFluctuations (Packing.cpp:841-856): sqrt(var(E_i)) / 2 over the per-particle energies */
rp_status rp_particle_energy_fluctuations(rp_handle *h, double *fluctuations_out);

/* ---- explicit moves (compatibility path): Packing::try* / accept* (Packing.cpp:84-164, 234-302) ---- */
rp_status rp_try_move(rp_handle *h, const rp_move *move, double *dE_out);
rp_status rp_accept_move(rp_handle *h);

/* ---- box move: Packing::tryScaling / revertScaling (Packing.cpp:172-208, 363-371) ---- */
rp_status rp_try_scaling(rp_handle *h, const double new_box[9], double *dE_out);
rp_status rp_revert_scaling(rp_handle *h);

/* ---- bulk sweep: Simulation::performMoves (Simulation.cpp:365-491) as one call per sampler per cycle ---- */
rp_status rp_sweep(rp_handle *h, const rp_sweep_params *params, rp_sweep_counters *counters_out);

/* Parity harness: what the last rp_sweep call did, one entry per particle.  Replaying the entries in the order
 * (colour_step, cell, batch, order_key, slot) through Packing::tryMove / acceptMove of the reference must reproduce every
 * accept / reject decision (tests/test_gpu_sweep_replay.py). */
typedef struct rp_trial_trace {
    uint64_t particle;
    int32_t colour_step;      /* 0..7: the sub-sweep that processed the particle */
    int32_t cell;             /* sweep-grid cell (the active domain of the trial) */
    uint32_t order_key;       /* processing order inside the cell: ascending (batch, order_key, slot) */
    int32_t slot;
    int32_t status;           /* 0 rejected (overlap or Metropolis), 1 accepted, 2 rejected: left the active cell,
                                 3 not attempted (RP_MOVE_FLIP: the particle was not drawn) */
    int32_t batch;            /* cells with more than 64 trials are resolved in batches of 64 slots (hard spheres /
                                 spherocylinders): ascending batch first */
    double position[3];       /* trial mass centre in the image frame of the sweep grid (not wrapped) */
    double orientation[9];    /* trial orientation */
    double image_shift[3];    /* image-frame position of the particle = canonical position + image_shift */
    double u;                 /* Metropolis variate (0 on the hard-only fast path, which draws none) */
    double dE;                /* what Packing::tryMove would have returned for the trial (undefined when status == 2) */
} rp_trial_trace;
rp_status rp_sweep_trace(rp_handle *h, rp_trial_trace *out, uint64_t n);

/* ---- lattice population on the device: Lattice(UnitCell(cell_box, molecules), n_cells) populated
 * by SerialPopulator(axis_order) with n_shapes shapes (core/lattice/Lattice.cpp:43-81, SerialPopulator.cpp:9-40).
 * cell_molecules: m x 12 doubles (RELATIVE position inside the cell, orientation); axis_order: a permutation of
 * {0, 1, 2}, outermost loop first ("xyz" = {0, 1, 2}).  The packing's box becomes Lattice::getLatticeBox(); positions
 * equal the reference's bit for bit.  Replaces the host loop + rp_upload for large N (nothing crosses PCIe but the
 * unit cell). ---- */
rp_status rp_populate_lattice(rp_handle *h, const double cell_box[9], const double *cell_molecules, uint64_t m,
                              const uint32_t n_cells[3], const int axis_order[3], uint64_t n_shapes);
/* RotationRandomizingTransformer (core/lattice/RotationRandomizingTransformer.cpp:21-56) applied to the packing:
 * mode 0 = uniformly random rotations (Tait-Bryan angles), 1 = about the fixed `axis` by U(0, 2 pi), 2 = about the
 * particle's own axis orientation * `axis`.  Per-particle Philox streams keyed by (seed, particle) replace the
 * reference's sequential std::mt19937: same distribution, different variates. */
rp_status rp_randomize_rotations(rp_handle *h, uint64_t seed, int mode, const double axis[3]);

/* ---- Packing::renormalizeOrientations / fixRotationMatrix (Packing.cpp:1178-1219), called by
 * Simulation::fixRotationMatrices every `rotation_matrix_fix_every` cycles.  allow_overlaps != 0 is the counting-mode
 * variant.  rejected_out: fixes not applied because they would create an overlap (per particle: every particle that
 * takes part in an overlap after the bulk fix gets its old matrix back; the check is repeated). ---- */
rp_status rp_renormalize_orientations(rp_handle *h, int allow_overlaps, uint64_t *rejected_out);

/* ---- counting mode + walls: Packing::toggleOverlapCounting / getCachedNumberOfOverlaps (Packing.cpp:1049-1060) ---- */
rp_status rp_toggle_overlap_counting(rp_handle *h, int count_overlaps);
/* Packing::toggleWalls (Packing.cpp:1081-1084): a hard wall on the two box faces normal to axis i when walls[i] != 0;
 * wall overlaps (Interaction::overlapWithWall, Packing::countParticleWallOverlaps, Packing.cpp:1086-1142) then count
 * like pair overlaps in rp_count_total_overlaps, rp_try_move, rp_try_scaling and rp_sweep.  (The reference does not
 * refresh a cached overlap count here either; call it before rp_toggle_overlap_counting.) */
rp_status rp_toggle_walls(rp_handle *h, const int32_t walls[3]);
rp_status rp_cached_overlaps(rp_handle *h, uint64_t *count_out);

/* ---- parity harness: Interaction::overlapBetween / calculateEnergyBetween on explicit pairs ---- */
rp_status rp_pair_test_batch(rp_handle *h, const rp_pair *pairs, uint64_t n, uint8_t *verdict_out, double *energy_out);

/* ---- spatial sharding over the GPUs of one node (no reference analogue: the reference is one OpenMP process;
 * semantic model: Simulation::performMovesWithDomainDivision, src/core/Simulation.cpp:389-441).
 * One process per GPU.  Rank 0 calls rp_comm_unique_id and the host distributes the 128 bytes (MPI, torch.distributed,
 * a file ...); every rank then calls rp_comm_init on its handle.  Afterwards all ranks must make the SAME sequence of
 * calls with the same arguments (rp_upload of the same state included): the state is replicated, rp_sweep runs only
 * the cell layers of this rank's slab and exchanges accepted moves over NCCL, rp_count_total_overlaps /
 * rp_try_scaling split the recount and all-reduce the count on the device.  An N-GPU run is bit-identical to the
 * 1-GPU run with the same seed.  Hard spheres / spherocylinders only in this round. */
rp_status rp_comm_unique_id(uint8_t id_out[128]);
rp_status rp_comm_init(rp_handle *h, const uint8_t id[128], int32_t rank, int32_t n_ranks);
rp_status rp_comm_destroy(rp_handle *h);
/* bytes this rank sent in the per-colour halo exchanges / received in the end-of-cycle gather of the last rp_sweep */
rp_status rp_comm_stats(rp_handle *h, uint64_t *halo_bytes, uint64_t *gather_bytes);
/* host logic of the sharded sweep, callable without a device: rank `rank` of `n_ranks` owns the cell layers
 * [2 * out[0], 2 * out[1]) of `n_layers_z`; after a sub-sweep of `colour` (bit 2 = z parity of the active cells) it
 * sends the accepted moves of layer out[2] to rank out[4] and receives those of layer out[3] from rank out[5] */
rp_status rp_shard_plan(int32_t n_layers_z, int32_t n_ranks, int32_t rank, int32_t colour, int32_t out[6]);

/* ---- device-side order parameters (the callers of a host-readable Packing in the reference; here they read the
 * resident state, so sampling them does not download 96 bytes per particle).
 * rp_nematic_order: NematicOrder::calculate (src/core/observables/NematicOrder.cpp:9-55) — Q = (3 <a a> - 1) / 2 with
 *   a = orientation * primary_axis (ShapeGeometry::getPrimaryAxis in the body frame), row-major Q in q_out, P2 = the
 *   eigenvalue of the largest magnitude.
 * rp_smectic_order: SmecticOrder::calculate (src/core/observables/SmecticOrder.cpp:14-62) with the constant shape
 *   function — tau = max over (h, k, l) in the reference's half-space scan of |sum_j exp(i k . r_j)| / V, r_j = position
 *   + orientation * focal_point (named point in the body frame; NULL or zeros = the position), k = 2 pi M^-T (h, k, l). */
rp_status rp_nematic_order(rp_handle *h, const double primary_axis[3], double q_out[9], double *p2_out);
/* rp_pair_density_correlation: one snapshot of PairDensityCorrelation with RadialEnumerator
 *   (src/core/observables/correlation/PairDensityCorrelation.cpp:8-41, RadialEnumerator.cpp:12-50): every pair i < j of
 *   focal points (position + orientation * focal_point; NULL or zeros = the position) at the reference's minimum-image
 *   distance d <= max_r lands in bin floor(d / (max_r / n_bins)) (the last bin also takes d == max_r); counts_out[b] =
 *   pairs in bin b (equal to the reference's counts), g_out[b] = counts * 2 / (N * expected molecules in the shell at
 *   the packing's number density).  Either output may be NULL.  n_bins <= 4096; max_r <= half of the smallest box
 *   height (RP_ERR_PRECONDITION otherwise: beyond it the reference's relative-coordinate minimum image is not the
 *   nearest image and the caller keeps the host path). */
rp_status rp_pair_density_correlation(rp_handle *h, double max_r, int32_t n_bins, const double focal_point[3],
                                      uint64_t *counts_out, double *g_out);
rp_status rp_smectic_order(rp_handle *h, const int32_t hkl_range[3], const double focal_point[3], double *tau_out,
                           int32_t hkl_out[3], double k_out[3]);
/* rp_nematic_order_static: no counterpart in the reference; added for callers
 *   that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs,
 *   outputs and errors as rp_nematic_order, but it allocates nothing: the per-block partial sums go to a file-scope
 *   __device__ array of 49 200 bytes (held by every GPU that loads the library until exit) instead of the handle's
 *   growable scratch buffer.  Packings of more than RP_NEMATIC_STATIC_MAX_PARTICLES particles are rejected with
 *   RP_ERR_PRECONDITION.  All calls on one device share the array: calls that overlap on the GPU (other streams,
 *   other host threads) race and must be serialised by the caller.  The summation order differs from
 *   rp_nematic_order's, so the two agree to rounding (1e-12), not bit for bit (rampack_b200/csrc/rp_noalloc.cu). */
#define RP_NEMATIC_STATIC_MAX_PARTICLES 4194304ull
rp_status rp_nematic_order_static(rp_handle *h, const double primary_axis[3], double q_out[9], double *p2_out);

/* ---- host cycle loop in C++: Simulation::performCycle / evaluateCounters / tryScaling (src/core/Simulation.cpp:
 * 325-363, 493-607) with the particle moves delegated to rp_sweep and the box move to rp_try_scaling.  One host thread;
 * every particle-level operation runs on the device.  Observables stay with the caller: run a block of cycles, then
 * rp_download (or a device-side observable) — that is Simulation::integrate's averaging loop turned inside out. ---- */
typedef enum rp_scaler_kind {
    RP_SCALER_NONE = 0,             /* Environment::disableBoxScaling: NVT */
    RP_SCALER_DELTA_V = 1,          /* TriclinicAdapter(DeltaVolumeScaler)          (volume_scalers/DeltaVolumeScaler.cpp) */
    RP_SCALER_LINEAR = 2,           /* AnisotropicVolumeScaler + LinearScalingFactorSampler (…/LinearScalingFactorSampler.cpp) */
    RP_SCALER_LOG = 3,              /* AnisotropicVolumeScaler + LogScalingFactorSampler    (…/LogScalingFactorSampler.cpp) */
    RP_SCALER_TRICLINIC_DELTA = 4   /* TriclinicDeltaScaler                         (volume_scalers/TriclinicDeltaScaler.cpp) */
} rp_scaler_kind;

#define RP_SIM_MAX_SAMPLERS 4
typedef struct rp_sim_config {
    uint64_t seed;                  /* host generator for box proposals and their Metropolis draws (mts[0] in the reference) */
    double temperature, pressure;
    int32_t n_samplers;             /* MoveSampler list; every sampler attempts N moves per cycle */
    int32_t scaler_kind;            /* rp_scaler_kind */
    struct {
        int32_t kind; int32_t reserved;
        double translation_step, rotation_step;
        double flip_axis[3], flip_origin[3];     /* RP_MOVE_FLIP */
        uint64_t flip_every;
    } samplers[RP_SIM_MAX_SAMPLERS];
    double scaler_step;
    /* LINEAR / LOG: groups of axes scaled by ONE sampled factor (ScalingDirection: bit 0 = x, 1 = y, 2 = z per group,
     * e.g. isotropic {7}, anisotropic xyz {1, 2, 4}, "xy|z" {3, 4}); independent != 0 samples one random group per
     * move (AnisotropicVolumeScaler.cpp:18-33).  TRICLINIC_DELTA: independent != 0 perturbs one random box vector,
     * 0 all nine entries. */
    int32_t n_direction_groups;
    int32_t direction_groups[3];
    int32_t independent;
    int32_t rotation_matrix_fix_every;   /* Simulation::fixRotationMatrices period in cycles (Simulation.cpp:144-145,
                                            750-769); 0 = never */
} rp_sim_config;

typedef struct rp_sim_stats {
    uint64_t cycles;                          /* performed by this rp_sim */
    uint64_t moves_attempted[RP_SIM_MAX_SAMPLERS], moves_accepted[RP_SIM_MAX_SAMPLERS];
    uint64_t scalings_attempted, scalings_evaluated /* reached rp_try_scaling */, scalings_accepted;
    double translation_step[RP_SIM_MAX_SAMPLERS], rotation_step[RP_SIM_MAX_SAMPLERS];   /* current (adapted) values */
    double scaler_step;
    double box[9];
    double sweep_kernel_ms, sweep_grid_ms, scaling_ms;   /* device time inside the kernels, accumulated */
    uint64_t cells_skipped;                   /* active cells a sweep left out because the shared-memory staging capacity
                                                 (sized from the previous sweeps, no host round trip) was too small; the
                                                 capacity grows at once, the moves of such a cell are simply not attempted */
} rp_sim_stats;

typedef struct rp_sim rp_sim;
rp_status rp_sim_create(rp_handle *h, const rp_sim_config *config, rp_sim **out);
void rp_sim_destroy(rp_sim *s);
/* `cycles` x Simulation::performCycle; adjust_step_sizes != 0 runs evaluateCounters after every cycle
 * (thermalisation phase, Simulation.cpp:128-135), 0 leaves the step sizes alone (averaging phase) */
rp_status rp_sim_run(rp_sim *s, uint64_t cycles, int adjust_step_sizes);
rp_status rp_sim_stats_get(const rp_sim *s, rp_sim_stats *out);
/* box proposal of the configured scaler for a given box (host logic, no device needed; tests) */
rp_status rp_sim_propose_box(rp_sim *s, const double old_box[9], double new_box[9]);

/* ---- timing hook for bench.py: device milliseconds spent inside the kernels of the last call ---- */
rp_status rp_last_kernel_ms(rp_handle *h, double *sweep_ms, double *sweep_grid_ms, double *rebuild_ms,
                            double *scaling_ms);
/* measurement helpers (bench.py's roofline): candidates per interaction centre in the reference-geometry grid —
 * Packing::getAverageNumberOfNeighbours (src/core/Packing.cpp:1000-1033) — and how many of them lie within
 * getRangeRadius(); FP64 vector-pipe peak of the device in GFLOP/s, fused (DFMA) and non-fused (DMUL + DADD) */
rp_status rp_neighbour_stats(rp_handle *h, double *candidates_per_centre, double *within_range_per_centre);
rp_status rp_measure_fp64_peak(rp_handle *h, double *dfma_gflops, double *dmul_dadd_gflops);
/* kernels launched by this handle so far (bench.py's gpu_launches) */
rp_status rp_launch_count(rp_handle *h, uint64_t *launches_out);
rp_status rp_synchronize(rp_handle *h);

#ifdef __cplusplus
}
#endif

#endif /* RAMPACK_B200_H */
