/*
 * rp_oracle.h — TEST INFRASTRUCTURE ONLY.  Plain-C CPU restatement ("port") of the reference's Monte Carlo hot path.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load liboracle_port.so.
 * Nothing under rampack_b200/ links, loads or calls it.  Parity of this restatement is PINNED: tests/test_oracle.py
 * checks it against the reference's own known-answer vectors (tests/golden/, transcribed from the reference's unit
 * tests with the generating script committed) and, where /root/reference is mounted, against the unmodified
 * reference compiled into oracle/_ref/.
 *
 * Every function cites the reference file:line (relative to /root/reference/src) it restates.
 */
#ifndef RP_ORACLE_H
#define RP_ORACLE_H

#include <stddef.h>
#include <stdint.h>
#include "../include/rampack_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rpo_packing rpo_packing;

/* Interaction::overlapBetween / calculateEnergyBetween on explicit pairs (core/Interaction.h:60-85) with the
 * HardcodedTranslation boundary conditions (core/Packing.cpp:23-43) */
int rpo_pair_test_batch(const rp_interaction_desc *desc, const rp_pair *pairs, uint64_t n, uint8_t *verdict,
                        double *energy);
/* the same with the true PeriodicBoundaryConditions of `box` (core/PeriodicBoundaryConditions.cpp:11-50);
 * pair.translation is ignored */
int rpo_pair_test_batch_pbc(const rp_interaction_desc *desc, const double box[9], const rp_pair *pairs, uint64_t n,
                            uint8_t *verdict, double *energy);

/* TriclinicBox (core/TriclinicBox.cpp:27-51) */
void rpo_box_info(const double box[9], double heights[3], double *volume, double inverse[9]);
/* PeriodicBoundaryConditions::getCorrection (core/PeriodicBoundaryConditions.cpp:11-22) */
void rpo_pbc_correction(const double box[9], const double pos[3], double out[3]);
/* NeighbourGrid::positionToCellNo on a grid sized like NeighbourGrid::setupSizes (core/NeighbourGrid.cpp:13-34,
 * 178-201); returns 0 on success */
int rpo_cell_numbers(const double box[9], double cell_size, const double *positions, uint64_t n, uint64_t *out,
                     uint64_t divisions_out[3]);
/* Matrix<3,3>::rotation (geometry/Matrix.tpp:437-449) */
void rpo_rotation_matrix(const double axis[3], double angle, double out[9]);
/* Interaction::getRangeRadius / getTotalRangeRadius (core/Interaction.cpp:55-70) */
double rpo_range_radius(const rp_interaction_desc *desc);
double rpo_total_range_radius(const rp_interaction_desc *desc);

/* Packing (core/Packing.cpp) with PeriodicBoundaryConditions, one thread */
rpo_packing *rpo_packing_create(const rp_interaction_desc *desc, const double box[9], const double *shapes,
                                uint64_t n, char *err, size_t err_len);
void rpo_packing_destroy(rpo_packing *p);
void rpo_packing_get(const rpo_packing *p, double box[9], double *shapes);
int rpo_grid_divisions(const rpo_packing *p, uint64_t out[3]);       /* returns 0 if the packing has no grid */
double rpo_average_neighbours(const rpo_packing *p);                 /* Packing.cpp:1000-1033 */
uint64_t rpo_count_total_overlaps(const rpo_packing *p, int early_exit);   /* Packing.cpp:519-559 */
double rpo_total_energy(const rpo_packing *p);                              /* Packing.cpp:690-709 */
int rpo_try_move(rpo_packing *p, const rp_move *m, double *dE);            /* Packing.cpp:84-164 */
void rpo_accept_move(rpo_packing *p);                                       /* Packing.cpp:278-302 */
int rpo_try_scaling(rpo_packing *p, const double new_box[9], double *dE);  /* Packing.cpp:172-208 */
void rpo_revert_scaling(rpo_packing *p);                                    /* Packing.cpp:363-371 */
void rpo_toggle_overlap_counting(rpo_packing *p, int on);                  /* Packing.cpp:1049-1053 */
uint64_t rpo_cached_overlaps(const rpo_packing *p);

/* Reference-semantics cycle loop (Simulation::performCycle, core/Simulation.cpp:325-531) with one domain:
 * `cycles` x (N rototranslation trial moves picked uniformly at random + optionally one delta-triclinic box move),
 * driven by a 64-bit SplitMix/xoshiro generator (the reference uses std::mt19937: trajectories differ, the
 * distributions are the same).  Used only as the "port" CPU baseline when oracle/_ref is not available. */
typedef struct rpo_run_result {
    uint64_t moves_attempted, moves_accepted, scalings_attempted, scalings_accepted;
    double move_seconds, scaling_seconds;
} rpo_run_result;
int rpo_run_cycles(rpo_packing *p, uint64_t seed, uint64_t cycles, double trans_step, double rot_step,
                   double temperature, double pressure, int box_moves, double box_step, rpo_run_result *res);

#ifdef __cplusplus
}
#endif
#endif
